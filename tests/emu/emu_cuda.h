// TEST INFRASTRUCTURE ONLY - never part of the product build.
//
// A minimal stand-in for the CUDA runtime and device built-ins so that mfv2d_b200/csrc/*.cu can be
// compiled with a plain C++ compiler (g++ -x c++ -DMFV2D_EMU) into tests/emu/_build/libmfv2d_emu.so.
// Every CTA of a launch is executed by ONE sequential thread (blockDim.x == 1): all kernels of the
// library are written as barrier-separated phases of thread-strided loops, for which that execution
// is valid.  It checks the host logic (bytecode planner, shared-memory layouts, index maps) and the
// kernels' index arithmetic in the GPU-less build container; it says nothing about races, timing or
// anything else that needs the real device - the `-m gpu` parity tests do that.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <vector>

#define __global__
#define __device__
#define __host__
#define __forceinline__ inline
#define __noinline__
#define __restrict__
#define __shared__ static thread_local
#define __launch_bounds__(...)
#define __align__(x)

struct emu_dim
{
    unsigned x = 1, y = 1, z = 1;
};
extern thread_local emu_dim threadIdx, blockIdx, blockDim, gridDim;
extern thread_local std::vector<double> emu_smem;

struct double2
{
    double x, y;
};
struct int4
{
    int x, y, z, w;
};
struct int2
{
    int x, y;
};
struct float4
{
    float x, y, z, w;
};
static inline float4 make_float4(float a, float b, float c, float d) { return float4{a, b, c, d}; }
static inline int4 make_int4(int a, int b, int c, int d) { return int4{a, b, c, d}; }
static inline double2 make_double2(double a, double b) { return double2{a, b}; }

static inline double __hiloint2double(int hi, int lo)
{
    const uint64_t u = ((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo;
    double d;
    memcpy(&d, &u, 8);
    return d;
}
static inline void __syncthreads() {}
static inline int __syncthreads_or(int v) { return v; } // one thread per CTA
static inline void __syncwarp() {}
static inline void __threadfence() {}
static inline void __threadfence_system() {}
template <typename T> static inline T __ldg(const T *p) { return *p; }
template <typename T> static inline T ld_ro(const T *p) { return *p; }
static inline double warp_sum(double v) { return v; }
static inline double atomicAdd(double *p, double v)
{
    double o = *p;
    *p += v;
    return o;
}
static inline int atomicAdd(int *p, int v)
{
    int o = *p;
    *p += v;
    return o;
}
static inline unsigned long long atomicAdd(unsigned long long *p, unsigned long long v)
{
    unsigned long long o = *p;
    *p += v;
    return o;
}
static inline int atomicMax(int *p, int v)
{
    int o = *p;
    if (v > o)
        *p = v;
    return o;
}
static inline int atomicExch(int *p, int v)
{
    int o = *p;
    *p = v;
    return o;
}
static inline double __shfl_xor_sync(unsigned, double v, int) { return v; }
static inline double __shfl_sync(unsigned, double v, int) { return v; }
static inline double __shfl_down_sync(unsigned, double v, int) { return v; }
static inline int __shfl_sync(unsigned, int v, int) { return v; }

#define MFV2D_DYN_SMEM(type, name) type *name = reinterpret_cast<type *>(emu_smem.data())
#define MFV2D_LANE0 0
#define MFV2D_LANE_STEP 1
#define MFV2D_NWARPS 1

#define MFV2D_LAUNCH(kernel, grid, block, smem, stream, ...)                                                   \
    do                                                                                                         \
    {                                                                                                          \
        emu_smem.assign((size_t)(smem) / sizeof(double) + 64, std::numeric_limits<double>::quiet_NaN());      \
        gridDim.x = (unsigned)(grid);                                                                          \
        blockDim.x = 1;                                                                                        \
        threadIdx.x = 0;                                                                                       \
        for (unsigned emu_b_ = 0; emu_b_ < (unsigned)(grid); ++emu_b_)                                         \
        {                                                                                                      \
            blockIdx.x = emu_b_;                                                                               \
            kernel(__VA_ARGS__);                                                                               \
        }                                                                                                      \
    } while (0)

// ---- runtime API subset -------------------------------------------------------------------
typedef int cudaError_t;
typedef void *cudaStream_t;
typedef void *cudaEvent_t;
enum
{
    cudaSuccess = 0,
    cudaErrorMemoryAllocation = 2
};
enum cudaMemcpyKind
{
    cudaMemcpyHostToHost,
    cudaMemcpyHostToDevice,
    cudaMemcpyDeviceToHost,
    cudaMemcpyDeviceToDevice,
    cudaMemcpyDefault
};
enum
{
    cudaStreamNonBlocking = 1,
    cudaFuncAttributeMaxDynamicSharedMemorySize = 8
};
struct cudaDeviceProp
{
    int major = 10, minor = 0, multiProcessorCount = 4;
    size_t sharedMemPerBlockOptin = 232448;
};
static inline const char *cudaGetErrorString(cudaError_t) { return "emulated CUDA error"; }
static inline cudaError_t cudaGetLastError() { return cudaSuccess; }
static inline cudaError_t cudaGetDeviceCount(int *n)
{
    *n = 1;
    return cudaSuccess;
}
static inline cudaError_t cudaGetDeviceProperties(cudaDeviceProp *p, int)
{
    *p = cudaDeviceProp();
    return cudaSuccess;
}
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaStreamCreateWithFlags(cudaStream_t *s, unsigned)
{
    *s = nullptr;
    return cudaSuccess;
}
static inline cudaError_t cudaStreamDestroy(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaStreamSynchronize(cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline cudaError_t cudaEventCreate(cudaEvent_t *e)
{
    *e = nullptr;
    return cudaSuccess;
}
static inline cudaError_t cudaEventDestroy(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventRecord(cudaEvent_t, cudaStream_t) { return cudaSuccess; }
static inline cudaError_t cudaEventSynchronize(cudaEvent_t) { return cudaSuccess; }
static inline cudaError_t cudaEventElapsedTime(float *t, cudaEvent_t, cudaEvent_t)
{
    *t = 0.f;
    return cudaSuccess;
}
static inline cudaError_t cudaMalloc(void **p, size_t n)
{
    *p = malloc(n ? n : 8);
    if (!*p)
        return cudaErrorMemoryAllocation;
    memset(*p, 0xFF, n); // NaN pattern: reads of unwritten memory show up
    return cudaSuccess;
}
template <typename T> static inline cudaError_t cudaMalloc(T **p, size_t n) { return cudaMalloc((void **)p, n); }
static inline cudaError_t cudaFree(void *p)
{
    free(p);
    return cudaSuccess;
}
static inline cudaError_t cudaMallocHost(void **p, size_t n)
{
    *p = malloc(n ? n : 8);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
template <typename T> static inline cudaError_t cudaMallocHost(T **p, size_t n) { return cudaMallocHost((void **)p, n); }
static inline cudaError_t cudaFreeHost(void *p)
{
    free(p);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpyAsync(void *d, const void *s, size_t n, cudaMemcpyKind, cudaStream_t)
{
    memmove(d, s, n);
    return cudaSuccess;
}
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind)
{
    memmove(d, s, n);
    return cudaSuccess;
}
static inline cudaError_t cudaMemsetAsync(void *d, int v, size_t n, cudaStream_t)
{
    memset(d, v, n);
    return cudaSuccess;
}
template <typename F> static inline cudaError_t cudaFuncSetAttribute(F, int, int) { return cudaSuccess; }
template <typename F> static inline cudaError_t cudaOccupancyMaxActiveBlocksPerMultiprocessor(int *n, F, int, size_t)
{
    *n = 1;
    return cudaSuccess;
}
// IPC: every "rank" of an emulated multi-rank test is a THREAD of one process, a handle is the pointer itself
struct cudaIpcMemHandle_t
{
    char reserved[64];
};
enum
{
    cudaIpcMemLazyEnablePeerAccess = 1
};
// separately rounded product and sum (the emulation is built without -mfma: no contraction)
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __dadd_rn(double a, double b) { return a + b; }
template <typename T> static inline cudaError_t cudaGetSymbolAddress(void **p, T &symbol)
{
    *p = (void *)&symbol;
    return cudaSuccess;
}
static inline cudaError_t cudaIpcGetMemHandle(cudaIpcMemHandle_t *h, void *p)
{
    memset(h, 0, sizeof *h);
    memcpy(h->reserved, &p, sizeof p);
    return cudaSuccess;
}
static inline cudaError_t cudaIpcOpenMemHandle(void **p, cudaIpcMemHandle_t h, unsigned)
{
    memcpy(p, h.reserved, sizeof *p);
    return cudaSuccess;
}
static inline cudaError_t cudaIpcCloseMemHandle(void *) { return cudaSuccess; }
static inline void cp_async8(double *smem_dst, const double *gmem_src) { *smem_dst = *gmem_src; }
static inline void cp_async_wait_all() {}
static inline double fast_rcp(double x) { return 1.0 / x; }
