// Micro-benchmark: issue rate of mma.sync.aligned.m8n8k4.f64 (fp64 tensor-core DMMA) and of plain DFMA on one B200.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_rate dmma_rate.cu && ./dmma_rate
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <int CHAINS> __global__ void dmma_kernel(double *out, int iters, double a, double b)
{
    double c0[CHAINS], c1[CHAINS];
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        c0[i] = c1[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; ++it)
#pragma unroll
        for (int i = 0; i < CHAINS; ++i)
            dmma(c0[i], c1[i], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < CHAINS; ++i)
        s += c0[i] + c1[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int CHAINS> __global__ void dfma_kernel(double *out, int iters, double a, double b)
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

int main()
{
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int sms = prop.multiProcessorCount, threads = 256, blocks = sms * 4, iters = 20000;
    double *out;
    cudaMalloc(&out, sizeof(double) * blocks * threads);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float ms;
    const double clk = prop.clockRate * 1e3;
    for (int rep = 0; rep < 2; ++rep)
    {
        cudaEventRecord(e0);
        dmma_kernel<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        const double n_dmma = (double)blocks * (threads / 32) * iters * 8;
        if (rep)
            printf("DMMA m8n8k4: %.3f ms, %.2f TFLOP/s, %.3f DMMA/clk/SM (nominal clock %.0f MHz)\n", ms, n_dmma * 512 / (ms * 1e-3) / 1e12,
                   n_dmma / (ms * 1e-3) / clk / sms, clk / 1e6);
        cudaEventRecord(e0);
        dfma_kernel<8><<<blocks, threads>>>(out, iters, 1.0000001, 0.9999999);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        const double n_fma = (double)blocks * threads * iters * 8;
        if (rep)
            printf("DFMA:        %.3f ms, %.2f TFLOP/s, %.2f thread-DFMA/clk/SM\n", ms, n_fma * 2 / (ms * 1e-3) / 1e12, n_fma / (ms * 1e-3) / clk / sms);
    }
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
