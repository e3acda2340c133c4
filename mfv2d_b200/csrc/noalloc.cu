// Allocation-free entry point (no counterpart in the reference).
//
// Callers that cannot call cudaMalloc - code under CUDA graph capture, tight serving loops, harnesses that forbid it - get
// ONE variant of a kernel that needs scratch memory: the device-wide in-place Gauss-Jordan inverse of one definite
// matrix (launch_inverse_definite_large of dense.cu, the routine that inverts the coarse operator of the multilevel
// preconditioner and cudaMalloc()s 2n doubles for the pivot row and column of every elimination step).  Here that
// scratch is a file-scope device array of a fixed, modest size:
//
//     static __device__ __align__(256) double noalloc_scratch[2 * MFV2D_NOALLOC_INVERSE_MAX_N]     (64 KiB)
//
// which every GPU that loads the library holds until the process exits.  In HOST code the name of the array is only a
// stand-in for the symbol, not a device address, so the address is resolved once per device - with that device current
// - by cudaGetSymbolAddress() and the resulting pointer is what the kernels get.
//
// The array is shared by every call and keeps the last call's contents.  Nothing in it has to start at zero: each
// elimination step overwrites entries [0, n) and [MAX_N, MAX_N + n) before it reads them, and no call reads beyond its
// own n.  The caller's status word is cleared by the entry point.  Calls that overlap on the GPU (two streams, two
// contexts on one device, two host threads) RACE on the array: serialise them.
#include <mutex>

#include "internal.cuh"

#define NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) double noalloc_scratch[2 * MFV2D_NOALLOC_INVERSE_MAX_N];

// device addresses of noalloc_scratch, one per device, resolved on the first call made for that device
static double *noalloc_scratch_ptr[NOALLOC_MAX_DEVICES];
static std::mutex noalloc_mutex;

// pivot column and pivot row of step k into the scratch (fully rewritten on [0, n) every step)
__global__ void noalloc_gj_snapshot_kernel(int n, int k, const double *__restrict__ M, double *__restrict__ colk, double *__restrict__ rowk,
                                           int32_t *__restrict__ status, double tiny)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        colk[i] = M[(size_t)i * n + k];
        rowk[i] = M[(size_t)k * n + i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && !(fabs(M[(size_t)k * n + k]) > tiny))
        *status = MFV2D_PIVOT_FAILED;
}

// Gauss-Jordan step k on the whole matrix (no pivoting: definite matrices only)
__global__ void __launch_bounds__(256) noalloc_gj_update_kernel(int n, int k, double *__restrict__ M, const double *__restrict__ colk,
                                                                const double *__restrict__ rowk)
{
    const double ip = 1.0 / colk[k];
    const size_t nn = (size_t)n * n;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < nn; idx += (size_t)gridDim.x * blockDim.x)
    {
        const int i = (int)(idx / n), j = (int)(idx - (size_t)i * n);
        double v;
        if (i == k)
            v = (j == k) ? ip : rowk[j] * ip;
        else if (j == k)
            v = -colk[i] * ip;
        else
            v = fma(-colk[i] * ip, rowk[j], M[idx]);
        M[idx] = v;
    }
}

static int noalloc_resolve(mfv2d_ctx *ctx, double **out)
{
    if (ctx->device < 0 || ctx->device >= NOALLOC_MAX_DEVICES)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "inverse_definite_noalloc: device %d beyond the %d supported", ctx->device,
                          NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(noalloc_mutex);
    MFV2D_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    if (!noalloc_scratch_ptr[ctx->device])
    {
        double *ptr = nullptr;
        MFV2D_CUDA_CHECK(ctx, cudaGetSymbolAddress((void **)&ptr, noalloc_scratch));
        if (!ptr)
            return mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "inverse_definite_noalloc: cudaGetSymbolAddress returned a null address");
        noalloc_scratch_ptr[ctx->device] = ptr;
    }
    *out = noalloc_scratch_ptr[ctx->device];
    return MFV2D_SUCCESS;
}

extern "C" {

int mfv2d_cuda_inverse_definite_noalloc(mfv2d_ctx *ctx, int n, const double *in, double *out, int32_t *status)
{
    if (!ctx)
        return MFV2D_BAD_ARGUMENT;
    if (n < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "inverse_definite_noalloc: n must be positive (got %d)", n);
    if (n > MFV2D_NOALLOC_INVERSE_MAX_N)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED,
                          "inverse_definite_noalloc: n = %d exceeds MFV2D_NOALLOC_INVERSE_MAX_N = %d (the static scratch holds "
                          "2 x %d doubles); use mfv2d_cuda_inverse_batch, which allocates",
                          n, MFV2D_NOALLOC_INVERSE_MAX_N, MFV2D_NOALLOC_INVERSE_MAX_N);
    if (!in || !out || !status)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "inverse_definite_noalloc: in, out and status must be device pointers");
    double *scratch = nullptr;
    MFV2D_TRY(noalloc_resolve(ctx, &scratch));
    double *colk = scratch, *rowk = scratch + MFV2D_NOALLOC_INVERSE_MAX_N;
    const size_t nn = (size_t)n * n;
    if (in != out)
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(out, in, nn * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(status, 0, sizeof(int32_t), ctx->stream));
    unsigned grid = (unsigned)((nn + 255) / 256);
    if (grid > (unsigned)ctx->sm_count * 8)
        grid = (unsigned)ctx->sm_count * 8;
    for (int k = 0; k < n; ++k)
    {
        MFV2D_LAUNCH(noalloc_gj_snapshot_kernel, (unsigned)((n + 255) / 256), 256, 0, ctx->stream, n, k, (const double *)out, colk, rowk,
                     status, 1e-300);
        MFV2D_LAUNCH(noalloc_gj_update_kernel, grid, 256, 0, ctx->stream, n, k, out, (const double *)colk, (const double *)rowk);
    }
    ctx->launches += 2 * (uint64_t)n;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

} // extern "C"
