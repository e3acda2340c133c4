// Two-level preconditioner for the Schur-complement CG on the trace space ("includes a trace-space
// preconditioner"; the reference has only block-Jacobi on the saddle system, python/mfv2d/solving.py:656-661).
//
//   M^-1 = D^-1 + P Ac^-1 P^T,   D = diag(S),   Ac = P^T S P
//
// P has ONE column per aggregate of multipliers: the near-null-space vector z of S (for the mixed Poisson / Darcy systems
// the multipliers of u = const, +-1 per multiplier) restricted to the aggregate.  The CG iteration count of the plain
// method grows like 1/h (12 201 iterations on the 10^6-element mesh of BASELINE.json configs[4]); with aggregates of
// H x H elements it is bounded independently of the mesh size for fixed H (measured: tests/test_precond.py).
// Ac (n_c <= 4096) is assembled element by element from the condensed blocks, summed over the ranks, inverted ONCE on
// the device (dense.cu) and kept as a dense inverse: per iteration the coarse solve is one n_c x n_c mat-vec that every
// rank performs redundantly on the all-reduced restriction - which travels through the same NVLink peer exchange as the
// interface multipliers (type 1 area), or through ncclAllReduce when the peers are not mapped.
#include <algorithm>

#include "system.cuh"

int launch_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *d_in, double *d_out, int32_t *d_status);

#define COARSE_MAX 4096

// Ac[a(r)][a(s)] += z_r z_s S_e[r][s] over the slots of every element (atomics: set-up only; the order of the sums, and
// with it the last bits of the PRECONDITIONER, may differ from run to run - the solution does not depend on it)
__global__ void coarse_assemble_kernel(unsigned long long n_elem, const int64_t *__restrict__ slot_off, const int64_t *__restrict__ slot_trace,
                                       const int64_t *__restrict__ s_off, const double *__restrict__ S, int packed,
                                       const int32_t *__restrict__ agg, const double *__restrict__ cz, long long nc, double *__restrict__ Ac)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        const double *Se = S + s_off[e];
        const int half = m >> 1, kfull = (m + 1) >> 1;
        for (int idx = threadIdx.x; idx < m * m; idx += blockDim.x)
        {
            const int s = idx / m, r = idx - s * m;
            const int64_t ir = slot_trace[s0 + r], is = slot_trace[s0 + s];
            const int ar = agg[ir], as = agg[is];
            if (ar < 0 || as < 0)
                continue;
            double v;
            if (packed)
            {
                // entry (k, row) with (row + k) % m = col, k <= m / 2
                int k = s - r;
                if (k < 0)
                    k += m;
                int row = r;
                if (k > half || (k == half && !(m & 1) && r >= half))
                {
                    k = m - k;
                    row = s;
                }
                (void)kfull;
                v = Se[(size_t)k * m + row];
            }
            else
                v = Se[idx];
            atomicAdd(&Ac[(size_t)ar * nc + as], cz[ir] * cz[is] * v);
        }
    }
}

struct CoarseArgs
{
    long long nc;
    const int64_t *cg_ptr, *cg_idx;
    const double *cz, *own, *r;
    double *rc;   // [nc] this rank's restriction
    int *tickets; // CTA ticket counter (last-block pattern)
    const int *done;
    PeerView pv;
};

// rc[I] = sum over the multipliers i of aggregate I of own_i z_i r_i (one warp per aggregate, fixed order); on a partitioned
// system the values are written into the type-1 area of every peer and the last CTA releases the epoch flag
__global__ void __launch_bounds__(SYS_THREADS) coarse_restrict_kernel(const CoarseArgs a)
{
    if (a.done && *a.done)
        return;
    __shared__ int s_last;
    const PeerView &pv = a.pv;
    const bool peers = pv.n_ranks > 1;
    unsigned long long e = 0;
    int parity = 0;
    if (peers)
    {
        if (*(volatile int *)pv.error)
            return;
        e = pv.epoch[1] + 1;
        parity = (int)(e & 1);
    }
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    for (long long I = (long long)blockIdx.x * nwarps + warp; I < a.nc; I += (long long)gridDim.x * nwarps)
    {
        double acc = 0.0;
        for (int64_t k = a.cg_ptr[I] + MFV2D_LANE0; k < a.cg_ptr[I + 1]; k += MFV2D_LANE_STEP)
        {
            const int64_t i = a.cg_idx[k];
            const double ri = a.own ? a.r[i] * a.own[i] : a.r[i];
            acc = fma(a.cz[i], ri, acc);
        }
        acc = warp_sum(acc);
        if (MFV2D_LANE0 == 0)
        {
            a.rc[I] = acc;
            if (peers)
                for (int r = 0; r < pv.n_ranks; ++r)
                    if (r != pv.rank)
                        peer_region(pv, r, 1, parity, pv.rank)[I] = acc;
        }
    }
    if (!peers)
        return;
    peer_fence(); // every writer orders its own remote stores before the CTA's ticket
    __syncthreads();
    if (threadIdx.x == 0)
    {
        peer_fence();
        s_last = atomicAdd(a.tickets, 1) == (int)gridDim.x - 1;
        if (s_last)
        {
            *a.tickets = 0;
            peer_fence();
            for (int r = 0; r < pv.n_ranks; ++r)
                if (r != pv.rank)
                    peer_store_flag(peer_flag(pv, r, 1, pv.rank), e);
            pv.epoch[1] = e;
        }
    }
}

// yc = Ac^-1 (sum over the ranks of rc), one warp per row; every CTA forms the rank-ordered sum in shared memory
__global__ void __launch_bounds__(SYS_THREADS) coarse_solve_kernel(long long nc, const double *__restrict__ Ainv, const double *__restrict__ rc,
                                                                  double *__restrict__ yc, const int *__restrict__ done, PeerView pv)
{
    MFV2D_DYN_SMEM(double, rs);
    if (done && *done)
        return;
    const bool peers = pv.n_ranks > 1;
    int parity = 0;
    if (peers)
    {
        const unsigned long long e = pv.epoch[1];
        parity = (int)(e & 1);
        peer_wait_all(pv, 1, e);
        if (*(volatile int *)pv.error)
            return;
    }
    for (long long i = threadIdx.x; i < nc; i += blockDim.x)
    {
        double tot = 0.0;
        for (int r = 0; r < pv.n_ranks; ++r)
            tot += r == pv.rank ? rc[i] : peer_load(peer_region(pv, pv.rank, 1, parity, r) + i);
        rs[i] = tot;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    for (long long row = (long long)blockIdx.x * nwarps + warp; row < nc; row += (long long)gridDim.x * nwarps)
    {
        const double *a = Ainv + (size_t)row * nc;
        double acc = 0.0;
        for (long long c = MFV2D_LANE0; c < nc; c += MFV2D_LANE_STEP)
            acc = fma(a[c], rs[c], acc);
        acc = warp_sum(acc);
        if (MFV2D_LANE0 == 0)
            yc[row] = acc;
    }
}

// z[i] += cz[i] yc[agg[i]]
__global__ void coarse_prolong_kernel(long long n, const int32_t *__restrict__ agg, const double *__restrict__ cz, const double *__restrict__ yc,
                                      double *__restrict__ z, const int *__restrict__ done)
{
    if (done && *done)
        return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const int a = agg[i];
        if (a >= 0)
            z[i] = fma(cz[i], yc[a], z[i]);
    }
}

// z += P Ac^-1 P^T r  (three launches; `done` = the CG's device flag so that a captured iteration stops by itself)
int coarse_correction(mfv2d_system *s, const double *r, double *z, const int *done, bool use_peers)
{
    mfv2d_ctx *ctx = s->ctx;
    const long long nc = s->n_coarse;
    if (!s->coarse_ready || nc == 0 || s->n_trace == 0)
        return MFV2D_SUCCESS;
    const int warps = SYS_THREADS / 32;
    PeerView one;
    const unsigned grid = (unsigned)std::min<long long>((nc + warps - 1) / warps, (long long)ctx->sm_count * 4);
    CoarseArgs ca{nc, s->d_cg_ptr, s->d_cg_idx, s->d_cz, s->d_owned, r, s->d_rc, reinterpret_cast<int *>(s->d_cg + 40), done,
                  use_peers ? s->peer : one};
    MFV2D_LAUNCH(coarse_restrict_kernel, grid, SYS_THREADS, 0, ctx->stream, ca);
    if (ctx->n_ranks > 1 && !use_peers)
        MFV2D_TRY(comm_allreduce(ctx, s->d_rc, (size_t)nc));
    const size_t smem = (size_t)nc * sizeof(double);
    MFV2D_LAUNCH(coarse_solve_kernel, grid, SYS_THREADS, smem, ctx->stream, nc, (const double *)s->d_Ac_inv, (const double *)s->d_rc, s->d_yc,
                 done, use_peers ? s->peer : one);
    MFV2D_LAUNCH(coarse_prolong_kernel, grid_for(ctx, s->n_trace, 256), 256, 0, ctx->stream, (long long)s->n_trace,
                 (const int32_t *)s->d_agg, (const double *)s->d_cz, (const double *)s->d_yc, z, done);
    ctx->launches += 3;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

extern "C" {

// Declare the coarse space of the two-level preconditioner: agg[n_trace] = aggregate (0 .. n_coarse-1, or -1) of every
// multiplier of THIS rank, z[n_trace] = the near-null-space vector of S on them (both must agree between the ranks that
// share a multiplier).  Assembles Ac = P^T S P from the condensed blocks, sums it over the ranks and inverts it.
// Collective on a partitioned system (after set_interface / peer_connect).
int mfv2d_cuda_system_set_coarse_space(mfv2d_system *s, size_t n_coarse, const int32_t *agg, const double *z)
{
    mfv2d_ctx *ctx = s->ctx;
    s->coarse_ready = false;
    if (n_coarse == 0)
        return MFV2D_SUCCESS;
    if (n_coarse > COARSE_MAX)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "set_coarse_space: at most %d aggregates (the coarse operator is kept as a dense inverse)", COARSE_MAX);
    MFV2D_TRY(sys_require_interface(s));
    const bool use_peers = sys_use_peers(s);
    if (use_peers && (int64_t)n_coarse > s->n_coarse_cap)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "set_coarse_space: %zu aggregates, but peer_export reserved room for %lld", n_coarse,
                          (long long)s->n_coarse_cap);
    const int64_t nt = s->n_trace;
    std::vector<int32_t> h_agg(agg, agg + nt);
    std::vector<double> h_z(z, z + nt);
    std::vector<int64_t> ptr(n_coarse + 1, 0), idx;
    for (int64_t i = 0; i < nt; ++i)
    {
        if (h_agg[i] >= (int64_t)n_coarse)
            return mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "set_coarse_space: aggregate %d of multiplier %lld is out of range", h_agg[i],
                              (long long)i);
        if (h_agg[i] >= 0)
            ptr[h_agg[i] + 1]++;
    }
    for (size_t a = 0; a < n_coarse; ++a)
        ptr[a + 1] += ptr[a];
    idx.resize((size_t)ptr[n_coarse]);
    {
        std::vector<int64_t> fill(ptr.begin(), ptr.end() - 1);
        for (int64_t i = 0; i < nt; ++i)
            if (h_agg[i] >= 0)
                idx[fill[h_agg[i]]++] = i;
    }
    s->n_coarse = (int64_t)n_coarse;
    MFV2D_TRY(sys_upload(s, h_agg, &s->d_agg));
    MFV2D_TRY(sys_upload(s, h_z, &s->d_cz));
    MFV2D_TRY(sys_upload(s, ptr, &s->d_cg_ptr));
    MFV2D_TRY(sys_upload(s, idx, &s->d_cg_idx));
    const size_t nn = n_coarse * n_coarse;
    double *d_Ac = nullptr;
    int32_t *d_status = nullptr;
    MFV2D_TRY(sys_alloc(s, nn, &d_Ac));
    MFV2D_TRY(sys_alloc(s, nn, &s->d_Ac_inv));
    MFV2D_TRY(sys_alloc(s, n_coarse, &s->d_rc));
    MFV2D_TRY(sys_alloc(s, n_coarse, &s->d_yc));
    MFV2D_TRY(sys_alloc(s, (size_t)2, &d_status));
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(d_Ac, 0, nn * sizeof(double), ctx->stream));
    if (s->n_elem)
    {
        long long grid = (long long)ctx->sm_count * 8;
        if (grid > s->n_elem)
            grid = s->n_elem;
        MFV2D_LAUNCH(coarse_assemble_kernel, (unsigned)grid, 128, 0, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off, s->d_slot_trace,
                     s->d_s_off, (const double *)s->d_S, (int)s->symmetric, (const int32_t *)s->d_agg, (const double *)s->d_cz,
                     (long long)n_coarse, d_Ac);
        ctx->launches++;
        MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    }
    if (ctx->n_ranks > 1)
    {
        if (use_peers)
        {
            // row by row through the type-1 area of the exchange blocks (n_coarse doubles per exchange)
            for (size_t row = 0; row < n_coarse; ++row)
                MFV2D_TRY(peer_dense_sum(s, d_Ac + row * n_coarse, (long long)n_coarse));
        }
        else
            MFV2D_TRY(comm_allreduce(ctx, d_Ac, nn));
    }
    MFV2D_TRY(launch_inverse_batch(ctx, (int)n_coarse, 1, d_Ac, s->d_Ac_inv, d_status));
    int32_t st = 0;
    MFV2D_TRY(mfv2d_cuda_download(ctx, &st, d_status, sizeof st));
    if (st != MFV2D_SUCCESS)
        return mfv2d_fail(ctx, MFV2D_PIVOT_FAILED, "set_coarse_space: the coarse operator P^T S P is singular (an aggregate without multipliers, "
                                                   "or a system whose Schur complement has a null space along z)");
    const size_t smem = n_coarse * sizeof(double);
    if (smem > 48 * 1024)
        MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(coarse_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    s->coarse_ready = true;
#ifndef MFV2D_EMU
    if (s->cg_graph_exec)
    {
        cudaGraphExecDestroy((cudaGraphExec_t)s->cg_graph_exec);
        s->cg_graph_exec = nullptr;
    }
#endif
    return MFV2D_SUCCESS;
}

} // extern "C"
