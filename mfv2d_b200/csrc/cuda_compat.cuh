// Thin macros over the CUDA launch / shared-memory syntax.
//
// The product build (nvcc, sm_100a) maps them straight onto CUDA.  tests/emu/ builds the same sources
// with a plain C++ compiler and MFV2D_EMU defined, where every CTA is executed by ONE sequential
// "thread" (blockDim.x == 1, __syncthreads() a no-op) so that the host logic (bytecode planner,
// layouts, index maps) and the kernels' index arithmetic can be unit-tested in a container without a
// GPU.  That build is test infrastructure only: it is never built, shipped or loaded by the package.
#pragma once

#ifndef MFV2D_EMU
#include <cuda_runtime.h>

#define MFV2D_LAUNCH(kernel, grid, block, smem, stream, ...) kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__)
#define MFV2D_DYN_SMEM(type, name) extern __shared__ __align__(16) type name[]
#define MFV2D_LANE0 (threadIdx.x & 31)
#define MFV2D_LANE_STEP 32
#define MFV2D_NWARPS (blockDim.x >> 5)

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double ld_ro(const double *p) { return __ldg(p); }
__device__ __forceinline__ int4 ld_ro(const int4 *p) { return __ldg(p); }
__device__ __forceinline__ int ld_ro(const int *p) { return __ldg(p); }
__device__ __forceinline__ float4 ld_ro(const float4 *p) { return __ldg(p); }

// 8-byte asynchronous global -> shared copy (LDGSTS): no registers are held while the data is in flight
__device__ __forceinline__ void cp_async8(double *smem_dst, const double *gmem_src)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;\n" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }
// 1 / x, IEEE rounded, without the slow paths of the full division
__device__ __forceinline__ double fast_rcp(double x) { return __drcp_rn(x); }

#else // ---------------------------------------------------------------------------------------
#include "../../tests/emu/emu_cuda.h"
#endif
