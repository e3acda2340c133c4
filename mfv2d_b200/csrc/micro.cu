// Measured fp64 peaks of THIS device, in the process that reports a roofline against them (bench.py): the issue rate of
// mma.sync.aligned.m8n8k4.f64 (DMMA, fp64 tensor cores) and of plain DFMA chains.  On B200 both end at 64 FMA / clock / SM
// (profiles/micro/dmma_rate_b200.txt): the denominator of the matrix-free apply's roofline.
#include "internal.cuh"

#ifndef MFV2D_EMU
__device__ __forceinline__ void micro_dmma(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int CHAINS> __global__ void micro_dmma_kernel(double *out, int iters, double a, double b)
{
    double c0[CHAINS], c1[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        c0[i] = c1[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
            micro_dmma(c0[i], c1[i], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS> __global__ void micro_dfma_kernel(double *out, int iters, double a, double b)
{
    double c[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        c[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
            c[i] = fma(c[i], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
#endif

extern "C" int mfv2d_cuda_measure_fp64_peak(mfv2d_ctx *ctx, double *tflops_dmma, double *tflops_dfma)
{
#ifdef MFV2D_EMU
    return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "fp64 peak measurement needs the device");
#else
    const int threads = 256, blocks = ctx->sm_count * 4, iters = 8000;
    double *out = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, sizeof(double) * blocks * threads, (void **)&out));
    double best[2] = {0.0, 0.0};
    for (int rep = 0; rep < 3; ++rep)
        for (int which = 0; which < 2; ++which)
        {
            MFV2D_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
            if (which == 0)
                micro_dmma_kernel<8><<<blocks, threads, 0, ctx->stream>>>(out, iters, 1.0000001, 0.9999999);
            else
                micro_dfma_kernel<8><<<blocks, threads, 0, ctx->stream>>>(out, iters, 1.0000001, 0.9999999);
            MFV2D_CUDA_CHECK(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
            MFV2D_CUDA_CHECK(ctx, cudaEventSynchronize(ctx->ev1));
            float ms = 0.f;
            MFV2D_CUDA_CHECK(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
            // one DMMA = 8 x 8 x 4 FMAs = 512 flop per warp instruction; one DFMA = 2 flop per thread
            const double flops = which == 0 ? (double)blocks * (threads / 32) * iters * 8 * 512.0 : (double)blocks * threads * iters * 8 * 2.0;
            const double tf = flops / (ms * 1e-3) / 1e12;
            if (tf > best[which])
                best[which] = tf;
            ctx->launches++;
        }
    MFV2D_TRY(mfv2d_cuda_free(ctx, out));
    if (tflops_dmma)
        *tflops_dmma = best[0];
    if (tflops_dfma)
        *tflops_dfma = best[1];
    return MFV2D_SUCCESS;
#endif
}
