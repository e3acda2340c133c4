// Context, error handling, memory and timing helpers of libmfv2d_cuda.so.
#include "internal.cuh"

static thread_local std::string g_error;

void mfv2d_set_global_error(const char *msg) { g_error = msg; }

int mfv2d_fail(mfv2d_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx)
        ctx->error = buf;
    g_error = buf;
    return code;
}

extern "C" {

const char *mfv2d_cuda_result_str(int code)
{
    static const char *names[] = {
        "MFV2D_SUCCESS",          "MFV2D_FAILED_ALLOC",    "MFV2D_BAD_ENUM",          "MFV2D_DIMS_MISMATCH",
        "MFV2D_DOUBLE_INCIDENCE", "MFV2D_OUT_OF_INSTRUCTIONS", "MFV2D_STACK_OVERFLOW", "MFV2D_STACK_UNDERFLOW",
        "MFV2D_WRONG_MAT_TYPES",  "MFV2D_STACK_NOT_EMPTY", "MFV2D_NOT_SQUARE",        "MFV2D_NOT_IN_SURFACE",
        "MFV2D_ORDER_MISMATCH",   "MFV2D_NOT_CONVERGED",   "MFV2D_INDEX_OUT_OF_RANGE", "MFV2D_NOT_A_LEAF",
        "MFV2D_FAILED_CALLBACK",  "MFV2D_UNSPECIFIED_ERROR", "MFV2D_BAD_ARGUMENT",    "MFV2D_PYTHON_EXCEPTION",
        "MFV2D_PIVOT_FAILED",     "MFV2D_BAD_VECTOR_PARENT", "MFV2D_MISMATCHED_TRACE"};
    if (code >= 0 && code < (int)(sizeof names / sizeof *names))
        return names[code];
    switch (code)
    {
    case MFV2D_CUDA_ERROR:
        return "MFV2D_CUDA_ERROR";
    case MFV2D_CUDA_NO_DEVICE:
        return "MFV2D_CUDA_NO_DEVICE";
    case MFV2D_CUDA_UNSUPPORTED:
        return "MFV2D_CUDA_UNSUPPORTED";
    case MFV2D_CUDA_NCCL_ERROR:
        return "MFV2D_CUDA_NCCL_ERROR";
    default:
        return "MFV2D_UNKNOWN";
    }
}

int mfv2d_cuda_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess)
    {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int mfv2d_cuda_init(mfv2d_ctx **out, int device)
{
    if (!out)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "ctx pointer is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
    {
        cudaGetLastError();
        return mfv2d_fail(nullptr, MFV2D_CUDA_NO_DEVICE, "no CUDA device available (%s); libmfv2d_cuda has no CPU fallback",
                          e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    }
    if (device < 0 || device >= n)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "device %d out of range [0, %d)", device, n);
    cudaDeviceProp prop;
    MFV2D_CUDA_CHECK(nullptr, cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return mfv2d_fail(nullptr, MFV2D_CUDA_NO_DEVICE, "device %d is sm_%d%d; this library is built for sm_100a only",
                          device, prop.major, prop.minor);
    MFV2D_CUDA_CHECK(nullptr, cudaSetDevice(device));
    mfv2d_ctx *ctx = new mfv2d_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&ctx->ev0) != cudaSuccess || cudaEventCreate(&ctx->ev1) != cudaSuccess ||
        cudaMalloc(&ctx->d_scalars, 64 * sizeof(double)) != cudaSuccess ||
        cudaMallocHost(&ctx->h_scalars, 64 * sizeof(double)) != cudaSuccess)
    {
        int r = mfv2d_fail(nullptr, MFV2D_CUDA_ERROR, "context resources: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return r;
    }
    *out = ctx;
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_comm_destroy(mfv2d_ctx *ctx);

void mfv2d_cuda_destroy(mfv2d_ctx *ctx)
{
    if (!ctx)
        return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    mfv2d_cuda_comm_destroy(ctx);
    if (ctx->flush_buf)
        cudaFree(ctx->flush_buf);
    cudaFree(ctx->d_scalars);
    cudaFreeHost(ctx->h_scalars);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *mfv2d_cuda_last_error(const mfv2d_ctx *ctx) { return ctx ? ctx->error.c_str() : g_error.c_str(); }

int mfv2d_cuda_synchronize(mfv2d_ctx *ctx)
{
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_malloc(mfv2d_ctx *ctx, size_t bytes, void **dptr)
{
    MFV2D_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    *dptr = nullptr;
    if (bytes == 0)
        bytes = 8;
    cudaError_t e = cudaMalloc(dptr, bytes);
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        return mfv2d_fail(ctx, MFV2D_FAILED_ALLOC, "cudaMalloc of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_free(mfv2d_ctx *ctx, void *dptr)
{
    if (!dptr)
        return MFV2D_SUCCESS;
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaFree(dptr));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_memset(mfv2d_ctx *ctx, void *dptr, int value, size_t bytes)
{
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(dptr, value, bytes, ctx->stream));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_upload(mfv2d_ctx *ctx, void *dptr, const void *hptr, size_t bytes)
{
    MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(dptr, hptr, bytes, cudaMemcpyHostToDevice, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_download(mfv2d_ctx *ctx, void *hptr, const void *dptr, size_t bytes)
{
    MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(hptr, dptr, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_host_alloc(mfv2d_ctx *ctx, size_t bytes, void **hptr)
{
    *hptr = nullptr;
    cudaError_t e = cudaMallocHost(hptr, bytes ? bytes : 8);
    if (e != cudaSuccess)
    {
        cudaGetLastError();
        return mfv2d_fail(ctx, MFV2D_FAILED_ALLOC, "cudaMallocHost of %zu bytes failed: %s", bytes, cudaGetErrorString(e));
    }
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_host_free(mfv2d_ctx *ctx, void *hptr)
{
    if (hptr)
        MFV2D_CUDA_CHECK(ctx, cudaFreeHost(hptr));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_timer_start(mfv2d_ctx *ctx)
{
    MFV2D_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_timer_stop(mfv2d_ctx *ctx, double *ms)
{
    MFV2D_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaEventSynchronize(ctx->ev1));
    float t = 0;
    MFV2D_CUDA_CHECK(ctx, cudaEventElapsedTime(&t, ctx->ev0, ctx->ev1));
    *ms = (double)t;
    return MFV2D_SUCCESS;
}

uint64_t mfv2d_cuda_launch_count(const mfv2d_ctx *ctx) { return ctx->launches; }

__global__ void flush_kernel(double *buf, size_t n, double v)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride)
        buf[i] = v;
}

int mfv2d_cuda_flush_l2(mfv2d_ctx *ctx)
{
    const size_t bytes = (size_t)256 << 20; // 256 MiB > 126 MB L2
    if (!ctx->flush_buf)
    {
        MFV2D_TRY(mfv2d_cuda_malloc(ctx, bytes, &ctx->flush_buf));
        ctx->flush_bytes = bytes;
    }
    MFV2D_LAUNCH(flush_kernel, ctx->sm_count * 8, 256, 0, ctx->stream, (double *)ctx->flush_buf, bytes / 8, 1.0);
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

} // extern "C"

// ---- staging helpers -----------------------------------------------------------------------
int stage_in(mfv2d_ctx *ctx, const void *ptr, size_t bytes, int space, Staged *s)
{
    *s = Staged();
    s->ctx = ctx;
    s->bytes = bytes;
    if (!ptr || bytes == 0)
        return MFV2D_SUCCESS;
    if (space == MFV2D_DEVICE)
    {
        s->dev = const_cast<void *>(ptr);
        return MFV2D_SUCCESS;
    }
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, bytes, &s->dev));
    s->owned = true;
    cudaError_t e = cudaMemcpyAsync(s->dev, ptr, bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e != cudaSuccess)
    {
        cudaFree(s->dev);
        s->dev = nullptr;
        s->owned = false;
        return mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "H2D copy failed: %s", cudaGetErrorString(e));
    }
    return MFV2D_SUCCESS;
}

int stage_out_alloc(mfv2d_ctx *ctx, void *ptr, size_t bytes, int space, Staged *s)
{
    *s = Staged();
    s->ctx = ctx;
    s->bytes = bytes;
    s->is_out = true;
    if (!ptr || bytes == 0)
        return MFV2D_SUCCESS;
    if (space == MFV2D_DEVICE)
    {
        s->dev = ptr;
        return MFV2D_SUCCESS;
    }
    s->host = ptr;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, bytes, &s->dev));
    s->owned = true;
    return MFV2D_SUCCESS;
}

int stage_finish(Staged *s)
{
    int r = MFV2D_SUCCESS;
    if (s->is_out && s->host && s->dev)
    {
        cudaError_t e = cudaMemcpyAsync(s->host, s->dev, s->bytes, cudaMemcpyDeviceToHost, s->ctx->stream);
        if (e == cudaSuccess)
            e = cudaStreamSynchronize(s->ctx->stream);
        if (e != cudaSuccess)
            r = mfv2d_fail(s->ctx, MFV2D_CUDA_ERROR, "D2H copy failed: %s", cudaGetErrorString(e));
    }
    if (s->owned && s->dev)
    {
        cudaStreamSynchronize(s->ctx->stream);
        cudaFree(s->dev);
    }
    s->dev = nullptr;
    s->owned = false;
    return r;
}

void stage_abort(Staged *s)
{
    if (s->owned && s->dev)
    {
        cudaStreamSynchronize(s->ctx->stream);
        cudaFree(s->dev);
    }
    s->dev = nullptr;
    s->owned = false;
}
