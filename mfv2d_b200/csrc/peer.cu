// Peer-to-peer exchange over NVLink / NVSwitch for partitioned systems (one process per GPU).
//
// Every rank owns ONE exchange block (cudaMalloc, exported with cudaIpcGetMemHandle); the host side passes the handles
// around by any means (like the NCCL unique id) and every rank maps the blocks of the others
// (cudaIpcOpenMemHandle, peer access enabled lazily).  From then on the kernels of the solver WRITE their interface
// contributions and partial scalars straight into the blocks of the ranks that need them, followed by a release store of
// an epoch flag; the consumer kernel acquires the flags and sums the contributions in rank order (deterministic, and
// bit-identical on every rank, so all ranks take the same branches).  No NCCL call, memset, pack / unpack kernel or
// device-to-device copy is left in a Krylov iteration: the collective is part of the compute kernels
// (cg1_gather_kernel / cg1_update_kernel in schur_cg.cu).  The generic kernels below give the same all-reduce for the
// operator entry points (apply_trace, apply_schur, dot products) outside the iteration.
//
// Why double buffering by epoch parity is enough: every exchange is all-to-all in its scalars, so a rank can start
// writing epoch e + 2 only after it has received epoch e + 1 from every peer, which each peer sends after it has consumed
// epoch e.
#include <algorithm>

#include "system.cuh"

#ifdef MFV2D_EMU
#include <atomic>
#include <thread>
void emu_peer_fence() { std::atomic_thread_fence(std::memory_order_seq_cst); }
unsigned long long emu_peer_load_flag(const unsigned long long *p)
{
    return reinterpret_cast<const std::atomic<unsigned long long> *>(p)->load(std::memory_order_acquire);
}
void emu_peer_store_flag(unsigned long long *p, unsigned long long v)
{
    reinterpret_cast<std::atomic<unsigned long long> *>(p)->store(v, std::memory_order_release);
}
void emu_peer_backoff() { std::this_thread::yield(); }
#endif

// ---- generic all-reduce of the shared multipliers of a trace vector (+ n_extra scalars) ----------------------------
// one CTA: the interface of a strip is a few thousand doubles; these calls are outside the Krylov iteration
__global__ void __launch_bounds__(1024) peer_send_kernel(PeerView pv, int type, long long n_shared, const int64_t *__restrict__ index,
                                                         const int64_t *__restrict__ slot, const int64_t *__restrict__ holder_ptr,
                                                         const int32_t *__restrict__ holder_rank, const double *__restrict__ t,
                                                         int n_extra, const double *__restrict__ extra, long long extra_at)
{
    if (*(volatile int *)pv.error)
        return;
    const unsigned long long e = pv.epoch[type] + 1;
    const int parity = (int)(e & 1);
    for (long long j = threadIdx.x; j < n_shared; j += blockDim.x)
    {
        const double v = t[index[j]];
        const long long sl = slot[j];
        for (int64_t h = holder_ptr[j]; h < holder_ptr[j + 1]; ++h)
            peer_region(pv, holder_rank[h], type, parity, pv.rank)[sl] = v;
    }
    for (int k = threadIdx.x; k < n_extra * pv.n_ranks; k += blockDim.x)
    {
        const int r = k / n_extra, i = k - r * n_extra;
        if (r != pv.rank)
            peer_region(pv, r, type, parity, pv.rank)[extra_at + i] = extra[i];
    }
    peer_fence();
    __syncthreads();
    if (threadIdx.x == 0)
    {
        peer_fence();
        for (int r = 0; r < pv.n_ranks; ++r)
            if (r != pv.rank)
                peer_store_flag(peer_flag(pv, r, type, pv.rank), e);
        pv.epoch[type] = e;
    }
}

__global__ void __launch_bounds__(1024) peer_recv_kernel(PeerView pv, int type, long long n_shared, const int64_t *__restrict__ index,
                                                         const int64_t *__restrict__ slot, const int64_t *__restrict__ holder_ptr,
                                                         const int32_t *__restrict__ holder_rank, double *__restrict__ t, int n_extra,
                                                         double *__restrict__ extra, long long extra_at)
{
    const unsigned long long e = pv.epoch[type];
    const int parity = (int)(e & 1);
    peer_wait_all(pv, type, e);
    if (*(volatile int *)pv.error)
        return;
    for (long long j = threadIdx.x; j < n_shared; j += blockDim.x)
    {
        // contributions in increasing rank order, this rank's own value at its place: the same sum on every holder
        const double own = t[index[j]];
        const long long sl = slot[j];
        double tot = 0.0;
        bool added = false;
        for (int64_t h = holder_ptr[j]; h < holder_ptr[j + 1]; ++h)
        {
            const int r = holder_rank[h];
            if (!added && r > pv.rank)
            {
                tot += own;
                added = true;
            }
            tot += peer_load(peer_region(pv, pv.rank, type, parity, r) + sl);
        }
        if (!added)
            tot += own;
        t[index[j]] = tot;
    }
    for (int i = threadIdx.x; i < n_extra; i += blockDim.x)
    {
        double tot = 0.0;
        for (int r = 0; r < pv.n_ranks; ++r)
            tot += r == pv.rank ? extra[i] : peer_load(peer_region(pv, pv.rank, type, parity, r) + extra_at + i);
        extra[i] = tot;
    }
}

int peer_interface_sum(const mfv2d_system *s, double *t, int n_extra, double *extra)
{
    mfv2d_ctx *ctx = s->ctx;
    if (n_extra > PEER_XSCALARS)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer exchange carries at most %d scalars", PEER_XSCALARS);
    const long long ns = t ? (long long)s->n_shared : 0;
    MFV2D_LAUNCH(peer_send_kernel, 1, 1024, 0, ctx->stream, s->peer, 0, ns, s->d_shared_index, s->d_shared_slot, s->d_holder_ptr,
                 s->d_holder_rank, (const double *)t, n_extra, (const double *)extra, (long long)s->n_iface);
    MFV2D_LAUNCH(peer_recv_kernel, 1, 1024, 0, ctx->stream, s->peer, 0, ns, s->d_shared_index, s->d_shared_slot, s->d_holder_ptr,
                 s->d_holder_rank, t, n_extra, extra, (long long)s->n_iface);
    ctx->launches += 2;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

void sys_release_peers(mfv2d_system *s)
{
#ifndef MFV2D_EMU
    if (s->cg_graph_exec)
    {
        cudaGraphExecDestroy((cudaGraphExec_t)s->cg_graph_exec);
        s->cg_graph_exec = nullptr;
    }
#endif
    for (void *p : s->peer_mapped)
        if (p)
            cudaIpcCloseMemHandle(p);
    s->peer_mapped.clear();
    if (s->peer_block)
    {
        cudaFree(s->peer_block);
        s->peer_block = nullptr;
    }
    s->peers_ready = false;
}

extern "C" {

int mfv2d_cuda_peer_handle_size(void) { return (int)sizeof(cudaIpcMemHandle_t); }

// Step 1 (after set_interface): allocate this rank's exchange block and export it.  n_coarse = capacity of the
// coarse-residual area (0 when no two-level preconditioner will be used).
int mfv2d_cuda_system_peer_export(mfv2d_system *s, size_t n_coarse, void *handle_bytes)
{
    mfv2d_ctx *ctx = s->ctx;
    if (ctx->n_ranks < 2)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_export: the context has one rank (call mfv2d_cuda_comm_init first)");
    if (ctx->n_ranks > MFV2D_MAX_PEERS)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "peer exchange supports at most %d ranks", MFV2D_MAX_PEERS);
    if (!s->interface_set)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_export: call mfv2d_cuda_system_set_interface first");
    sys_release_peers(s);
    PeerView &pv = s->peer;
    pv = PeerView();
    pv.n_ranks = ctx->n_ranks;
    pv.rank = ctx->rank;
    pv.n_iface = s->n_iface;
    pv.stride[0] = (s->n_iface + PEER_XSCALARS + 1) & ~1ll;
    pv.stride[1] = ((long long)n_coarse + 1) & ~1ll;
    pv.area_off[0] = 0;
    pv.area_off[1] = 2 * pv.n_ranks * pv.stride[0];
    s->n_coarse_cap = (int64_t)n_coarse;
    const size_t bytes = PEER_FLAG_BYTES + (size_t)(pv.area_off[1] + 2 * pv.n_ranks * pv.stride[1]) * sizeof(double);
    MFV2D_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    MFV2D_CUDA_CHECK(ctx, cudaMalloc(&s->peer_block, bytes));
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(s->peer_block, 0, bytes, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    cudaIpcMemHandle_t h;
    MFV2D_CUDA_CHECK(ctx, cudaIpcGetMemHandle(&h, s->peer_block));
    memcpy(handle_bytes, &h, sizeof h);
    return MFV2D_SUCCESS;
}

// Step 2 (after every rank has exported and the handles have been gathered, rank-major): map the peers' blocks and
// upload, for each shared multiplier of this rank (in the order of set_interface), the OTHER ranks that hold it
// (ascending).  Every rank must have called peer_export before any rank calls this (the host's gather is the barrier).
int mfv2d_cuda_system_peer_connect(mfv2d_system *s, const void *handles_all, size_t n_shared, const int64_t *holder_ptr,
                                   const int32_t *holder_rank)
{
    mfv2d_ctx *ctx = s->ctx;
    if (!s->peer_block)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_connect: call mfv2d_cuda_system_peer_export first");
    if ((int64_t)n_shared != s->n_shared)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_connect: %zu shared multipliers, set_interface declared %lld", n_shared,
                          (long long)s->n_shared);
    PeerView &pv = s->peer;
    for (size_t j = 0; j < n_shared; ++j)
        for (int64_t k = holder_ptr[j]; k < holder_ptr[j + 1]; ++k)
            if (holder_rank[k] < 0 || holder_rank[k] >= pv.n_ranks || holder_rank[k] == pv.rank ||
                (k > holder_ptr[j] && holder_rank[k] <= holder_rank[k - 1]))
                return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_connect: holder ranks of shared multiplier %zu must be other ranks in ascending order", j);
    MFV2D_CUDA_CHECK(ctx, cudaSetDevice(ctx->device));
    s->peer_mapped.assign(pv.n_ranks, nullptr);
    for (int r = 0; r < pv.n_ranks; ++r)
    {
        if (r == pv.rank)
        {
            pv.base[r] = (unsigned char *)s->peer_block;
            continue;
        }
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *)handles_all + (size_t)r * sizeof h, sizeof h);
        void *p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess)
        {
            cudaGetLastError();
            sys_release_peers(s);
            return mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "cudaIpcOpenMemHandle of rank %d's exchange block failed: %s", r, cudaGetErrorString(e));
        }
        s->peer_mapped[r] = p;
        pv.base[r] = (unsigned char *)p;
    }
    std::vector<int64_t> hp(holder_ptr, holder_ptr + n_shared + 1);
    std::vector<int32_t> hr(holder_rank, holder_rank + hp[n_shared]);
    MFV2D_TRY(sys_upload(s, hp, &s->d_holder_ptr));
    MFV2D_TRY(sys_upload(s, hr, &s->d_holder_rank));
    // position of every multiplier in the shared list (-1: not shared)
    {
        std::vector<int64_t> idx((size_t)s->n_shared);
        if (s->n_shared)
            MFV2D_TRY(mfv2d_cuda_download(ctx, idx.data(), s->d_shared_index, idx.size() * sizeof(int64_t)));
        std::vector<int32_t> of((size_t)s->n_trace, -1);
        for (int64_t j = 0; j < s->n_shared; ++j)
            of[idx[j]] = (int32_t)j;
        MFV2D_TRY(sys_upload(s, of, &s->d_shared_of));
    }
    void *p = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, 64 + 32 * sizeof(unsigned long long), &p));
    s->allocs.push_back(p);
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(p, 0, 64 + 32 * sizeof(unsigned long long), ctx->stream));
    pv.epoch = (unsigned long long *)p;
    pv.error = (int *)((char *)p + 32);
    pv.stats = (unsigned long long *)((char *)p + 64);
    {
        const char *dbg = getenv("MFV2D_PEER_DEBUG");
        pv.dbg = dbg ? atoi(dbg) : 0;
    }
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    s->peers_ready = true;
    return MFV2D_SUCCESS;
}

// non-zero after a peer wait timed out (a rank died or never joined): the results of that call are invalid
int mfv2d_cuda_system_peer_error(mfv2d_system *s)
{
    if (!s->peers_ready)
        return 0;
    int h = 0;
    cudaMemcpyAsync(&h, s->peer.error, sizeof h, cudaMemcpyDeviceToHost, s->ctx->stream);
    cudaStreamSynchronize(s->ctx->stream);
    return h;
}

// diagnostics (MFV2D_PEER_DEBUG bit 8): the timeline stamps of this rank's stream, nanoseconds accumulated before each of
// the 16 stamp ids (schur_cg.cu / precond.cu say where they sit); the counters are reset by the call
int mfv2d_cuda_system_peer_stats(mfv2d_system *s, uint64_t *out16)
{
    memset(out16, 0, 16 * sizeof(uint64_t));
    if (!s->peers_ready || !s->peer.stats)
        return MFV2D_SUCCESS;
    mfv2d_ctx *ctx = s->ctx;
    MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(out16, s->peer.stats, 16 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(s->peer.stats, 0, 17 * sizeof(uint64_t), ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return MFV2D_SUCCESS;
}

} // extern "C"

// all-reduce (sum, rank order) of n <= capacity doubles through the type-1 area: set-up paths only
int peer_dense_sum(const mfv2d_system *s, double *v, long long n)
{
    mfv2d_ctx *ctx = s->ctx;
    if (n > s->n_coarse_cap)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "peer_dense_sum: %lld doubles exceed the reserved %lld", n, (long long)s->n_coarse_cap);
    MFV2D_LAUNCH(peer_send_kernel, 1, 1024, 0, ctx->stream, s->peer, 1, 0ll, (const int64_t *)nullptr, (const int64_t *)nullptr,
                 (const int64_t *)nullptr, (const int32_t *)nullptr, (const double *)nullptr, (int)n, (const double *)v, 0ll);
    MFV2D_LAUNCH(peer_recv_kernel, 1, 1024, 0, ctx->stream, s->peer, 1, 0ll, (const int64_t *)nullptr, (const int64_t *)nullptr,
                 (const int64_t *)nullptr, (const int32_t *)nullptr, (double *)nullptr, (int)n, v, 0ll);
    ctx->launches += 2;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}
