// Multi-GPU plumbing: one process per GPU, NCCL over NVLink for the two exchange points of the solver
// half (interface-multiplier sums and Krylov dot products).  NCCL is dlopen()ed on first
// use so that single-GPU use has no dependency on it.
#include "internal.cuh"

#ifndef MFV2D_EMU
#include <dlfcn.h>
#endif

namespace
{
// the handful of NCCL entry points used, with their public signatures (nccl.h)
typedef struct
{
    char internal[128];
} nccl_unique_id;
typedef int (*fn_get_unique_id)(nccl_unique_id *);
typedef int (*fn_comm_init_rank)(void **, int, nccl_unique_id, int);
typedef int (*fn_comm_destroy)(void *);
typedef int (*fn_all_reduce)(const void *, void *, size_t, int, int, void *, cudaStream_t);
typedef const char *(*fn_get_error_string)(int);

struct NcclApi
{
    void *lib = nullptr;
    fn_get_unique_id get_unique_id = nullptr;
    fn_comm_init_rank comm_init_rank = nullptr;
    fn_comm_destroy comm_destroy = nullptr;
    fn_all_reduce all_reduce = nullptr;
    fn_get_error_string get_error_string = nullptr;
    bool tried = false;
} g_nccl;

const int NCCL_FLOAT64 = 8; // ncclDouble
const int NCCL_SUM = 0;     // ncclSum

bool load_nccl()
{
#ifdef MFV2D_EMU
    return false;
#else
    if (g_nccl.tried)
        return g_nccl.lib != nullptr;
    g_nccl.tried = true;
    const char *names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char *n : names)
    {
        g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (g_nccl.lib)
            break;
    }
    if (!g_nccl.lib)
        return false;
    g_nccl.get_unique_id = (fn_get_unique_id)dlsym(g_nccl.lib, "ncclGetUniqueId");
    g_nccl.comm_init_rank = (fn_comm_init_rank)dlsym(g_nccl.lib, "ncclCommInitRank");
    g_nccl.comm_destroy = (fn_comm_destroy)dlsym(g_nccl.lib, "ncclCommDestroy");
    g_nccl.all_reduce = (fn_all_reduce)dlsym(g_nccl.lib, "ncclAllReduce");
    g_nccl.get_error_string = (fn_get_error_string)dlsym(g_nccl.lib, "ncclGetErrorString");
    if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.comm_destroy || !g_nccl.all_reduce)
    {
        dlclose(g_nccl.lib);
        g_nccl.lib = nullptr;
        return false;
    }
    return true;
#endif
}
} // namespace

int comm_allreduce(mfv2d_ctx *ctx, double *dptr, size_t n)
{
    if (ctx->n_ranks <= 1 || n == 0)
        return MFV2D_SUCCESS;
    if (!ctx->nccl_comm)
        return mfv2d_fail(ctx, MFV2D_CUDA_NCCL_ERROR, "all-reduce requested without an NCCL communicator (and the peer exchange of this system "
                                                      "is not connected)");
    const int rc = g_nccl.all_reduce(dptr, dptr, n, NCCL_FLOAT64, NCCL_SUM, ctx->nccl_comm, ctx->stream);
    if (rc != 0)
        return mfv2d_fail(ctx, MFV2D_CUDA_NCCL_ERROR, "ncclAllReduce failed: %s",
                          g_nccl.get_error_string ? g_nccl.get_error_string(rc) : "?");
    return MFV2D_SUCCESS;
}

extern "C" {

int mfv2d_cuda_comm_id_size(void) { return (int)sizeof(nccl_unique_id); }

int mfv2d_cuda_comm_get_id(void *id_bytes)
{
    if (!load_nccl())
        return mfv2d_fail(nullptr, MFV2D_CUDA_NCCL_ERROR, "libnccl.so.2 could not be loaded");
    nccl_unique_id id;
    const int rc = g_nccl.get_unique_id(&id);
    if (rc != 0)
        return mfv2d_fail(nullptr, MFV2D_CUDA_NCCL_ERROR, "ncclGetUniqueId failed (%d)", rc);
    memcpy(id_bytes, &id, sizeof id);
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_comm_init(mfv2d_ctx *ctx, int n_ranks, int rank, const void *id_bytes)
{
    if (n_ranks < 1 || rank < 0 || rank >= n_ranks)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "comm_init: bad rank %d of %d", rank, n_ranks);
    ctx->n_ranks = n_ranks;
    ctx->rank = rank;
    if (n_ranks == 1 || !id_bytes) // id_bytes == NULL: ranks only, every exchange goes through the mapped peer blocks (peer.cu)
        return MFV2D_SUCCESS;
    if (!load_nccl())
        return mfv2d_fail(ctx, MFV2D_CUDA_NCCL_ERROR, "libnccl.so.2 could not be loaded");
    nccl_unique_id id;
    memcpy(&id, id_bytes, sizeof id);
    MFV2D_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    const int rc = g_nccl.comm_init_rank(&ctx->nccl_comm, n_ranks, id, rank);
    if (rc != 0)
    {
        ctx->nccl_comm = nullptr;
        ctx->n_ranks = 1;
        return mfv2d_fail(ctx, MFV2D_CUDA_NCCL_ERROR, "ncclCommInitRank failed: %s",
                          g_nccl.get_error_string ? g_nccl.get_error_string(rc) : "?");
    }
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_comm_destroy(mfv2d_ctx *ctx)
{
    if (ctx->nccl_comm && g_nccl.comm_destroy)
    {
        g_nccl.comm_destroy(ctx->nccl_comm);
        ctx->nccl_comm = nullptr;
    }
    ctx->n_ranks = 1;
    ctx->rank = 0;
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_comm_allreduce_sum(mfv2d_ctx *ctx, double *dptr, size_t n) { return comm_allreduce(ctx, dptr, n); }

} // extern "C"
