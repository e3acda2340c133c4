// Two-level preconditioner for the Schur-complement CG on the trace space ("includes a trace-space
// preconditioner"; the reference has only block-Jacobi on the saddle system, python/mfv2d/solving.py:656-661).
//
//   M^-1 = D^-1 + sum_l P_l D_l^-1 P_l^T + P_L A_L^-1 P_L^T,   D = diag(S),  D_l = diag(P_l^T S P_l),  A_L = P_L^T S P_L
//
// an additive multilevel (BPX-type) preconditioner.  Every level l has ONE vector per aggregate of multipliers: the
// near-null-space vector z of S (for the mixed Poisson / Darcy systems the multipliers of u = const, +-1 per multiplier)
// restricted to the aggregate.  The intermediate levels (aggregates of a few elements, local to a rank: shared
// multipliers are left out of them) only need the DIAGONAL of their coarse operator; the last level (<= 4096 aggregates,
// global) is solved exactly: A_L is assembled element by element from the condensed blocks, summed over the ranks,
// inverted ONCE on the device (dense.cu) and kept as a dense inverse, so that per iteration the coarse solve is one
// mat-vec that every rank performs redundantly on the all-reduced restriction - which travels through the same NVLink
// peer exchange as the interface multipliers (type 1 area), or through ncclAllReduce when the peers are not mapped.
// The CG iteration count of the plain method grows like 1/h (12 201 iterations on the 10^6-element mesh of
// BASELINE.json configs[4]); with the hierarchy it is bounded for a fixed size of the last level's aggregates.
// Four launches per application: chunk partial sums of all levels | per-aggregate sums (scaled by D_l^-1, or sent to the
// peers) | coarse mat-vec | prolongation of all levels.
#include <algorithm>

#include "system.cuh"

int launch_inverse_definite_large(mfv2d_ctx *ctx, int n, double *d_M, int *d_flag);

#define COARSE_MAX 4096
#define COARSE_MAX_LEVELS 6
#define COARSE_CHUNK 256

// coarse operator entries from the condensed blocks: z_r z_s S_e[r][s] is added to out[a(r) * nc + a(s)] (full = 1) or, only
// where a(r) == a(s), to out[a(r)] (full = 0: the diagonal of P^T S P).  Atomics: set-up only; the order of the sums, and
// with it the last bits of the PRECONDITIONER, may differ from run to run - the solution does not depend on it.
__global__ void coarse_assemble_kernel(unsigned long long n_elem, const int64_t *__restrict__ slot_off, const int64_t *__restrict__ slot_trace,
                                       const int64_t *__restrict__ s_off, const double *__restrict__ S, int packed,
                                       const int32_t *__restrict__ agg, const double *__restrict__ cz, long long nc, int full,
                                       double *__restrict__ out)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        const double *Se = S + s_off[e];
        const int half = m >> 1;
        for (int idx = threadIdx.x; idx < m * m; idx += blockDim.x)
        {
            const int s = idx / m, r = idx - s * m;
            const int64_t ir = slot_trace[s0 + r], is = slot_trace[s0 + s];
            const int ar = agg[ir], as = agg[is];
            if (ar < 0 || as < 0 || (!full && ar != as))
                continue;
            double v;
            if (packed)
            {
                // S[r][s] is stored as entry (k, r), k = (s - r) mod m, if k is in the stored half, else as its mirror image
                int k = s - r;
                if (k < 0)
                    k += m;
                int row = r;
                if (k > half || (k == half && !(m & 1) && r >= half))
                {
                    k = m - k;
                    row = s;
                }
                v = Se[(size_t)k * m + row];
            }
            else
                v = Se[idx];
            atomicAdd(full ? &out[(size_t)ar * nc + as] : &out[ar], cz[ir] * cz[is] * v);
        }
    }
}

// d[i] = 1 / d[i] (1 where it vanishes)
__global__ void reciprocal_kernel(long long n, double *__restrict__ d)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        d[i] = d[i] != 0.0 ? 1.0 / d[i] : 1.0;
}

// part[c] = sum over the multipliers i of chunk c of own_i z_i r_i (one warp per chunk, fixed order)
__global__ void __launch_bounds__(SYS_THREADS) coarse_chunk_kernel(long long n_chunks, const int64_t *__restrict__ chunk_ptr,
                                                                  const int64_t *__restrict__ chunk_idx, const double *__restrict__ cz,
                                                                  const double *__restrict__ own, const double *__restrict__ r,
                                                                  double *__restrict__ part, const int *__restrict__ done)
{
    if (done && *done)
        return;
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    for (long long c = (long long)blockIdx.x * nwarps + warp; c < n_chunks; c += (long long)gridDim.x * nwarps)
    {
        double acc = 0.0;
        for (int64_t k = chunk_ptr[c] + MFV2D_LANE0; k < chunk_ptr[c + 1]; k += MFV2D_LANE_STEP)
        {
            const int64_t i = chunk_idx[k];
            acc = fma(cz[i], own ? r[i] * own[i] : r[i], acc);
        }
        acc = warp_sum(acc);
        if (MFV2D_LANE0 == 0)
            part[c] = acc;
    }
}

struct CoarseArgs
{
    long long n_agg_total, exact_begin, n_exact; // aggregates of all levels; the exact level is the tail [exact_begin, +n_exact)
    const int64_t *aggchunk_ptr;                 // [n_agg_total + 1] chunks of each aggregate (contiguous)
    const double *part, *ldinv;                  // chunk partial sums; 1 / D_l per aggregate of the diagonal levels
    double *y;                                   // [n_agg_total] scaled restrictions (diagonal levels) | this rank's rc (exact level)
    int *tickets;                                // CTA ticket counter (last-block pattern)
    const int *done;
    PeerView pv;
};

// y[J] = (sum of the chunks of aggregate J) / D_l for the diagonal levels; for the exact level the sums are this rank's part
// of the restricted residual: written into the type-1 area of every peer, and the last CTA releases the epoch flag
__global__ void __launch_bounds__(SYS_THREADS) coarse_aggregate_kernel(const CoarseArgs a)
{
    if (a.done && *a.done)
        return;
    __shared__ int s_last;
    const PeerView &pv = a.pv;
    const bool peers = pv.n_ranks > 1 && a.n_exact > 0;
    unsigned long long e = 0;
    int parity = 0;
    if (peers)
    {
        if (*(volatile int *)pv.error)
            return;
        e = pv.epoch[1] + 1;
        parity = (int)(e & 1);
        if (blockIdx.x == 0 && threadIdx.x == 0)
            peer_stamp(pv, 5); // since the CG update: launch gaps + the chunk sums
    }
    bool wrote = false;
    for (long long J = blockIdx.x * (long long)blockDim.x + threadIdx.x; J < a.n_agg_total; J += (long long)gridDim.x * blockDim.x)
    {
        double acc = 0.0;
        for (int64_t c = a.aggchunk_ptr[J]; c < a.aggchunk_ptr[J + 1]; ++c)
            acc += a.part[c];
        if (J < a.exact_begin)
            a.y[J] = acc * a.ldinv[J];
        else
        {
            a.y[J] = acc;
            if (peers)
            {
                for (int r = 0; r < pv.n_ranks; ++r)
                    if (r != pv.rank)
                        peer_region(pv, r, 1, parity, pv.rank)[J - a.exact_begin] = acc;
                wrote = true;
            }
        }
    }
    if (!peers)
        return;
    if (wrote)
        peer_fence(); // every writer orders its own remote stores before the CTA's ticket
    const int cta_wrote = __syncthreads_or(wrote ? 1 : 0);
    if (threadIdx.x == 0)
    {
        if (cta_wrote)
            peer_fence();
        else
            __threadfence();
        s_last = atomicAdd(a.tickets, 1) == (int)gridDim.x - 1;
        if (s_last)
        {
            *a.tickets = 0;
            peer_fence();
            for (int r = 0; r < pv.n_ranks; ++r)
                if (r != pv.rank)
                    peer_store_flag(peer_flag(pv, r, 1, pv.rank), e);
            pv.epoch[1] = e;
            peer_stamp(pv, 6); // aggregate sums + publication of the coarse residual
        }
    }
}

// rsum = sum over the ranks (in rank order) of the restricted residuals: this rank's rc + the copies the peers have written
// into the type-1 area.  The only kernel of the coarse solve that waits for the peers.
__global__ void __launch_bounds__(SYS_THREADS) coarse_sum_kernel(long long nc, const double *__restrict__ rc, double *__restrict__ rsum,
                                                                const int *__restrict__ done, PeerView pv)
{
    if (done && *done)
        return;
    const unsigned long long e = pv.epoch[1];
    const int parity = (int)(e & 1);
    if (blockIdx.x == 0 && threadIdx.x == 0)
        peer_stamp(pv, 7); // launch gap
    peer_wait_all(pv, 1, e);
    if (blockIdx.x == 0 && threadIdx.x == 0)
        peer_stamp(pv, 8); // waiting for the peers' coarse residuals
    if (*(volatile int *)pv.error)
        return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nc; i += (long long)gridDim.x * blockDim.x)
    {
        double tot = 0.0;
        for (int r = 0; r < pv.n_ranks; ++r)
            tot += r == pv.rank ? rc[i] : peer_load(peer_region(pv, pv.rank, 1, parity, r) + i);
        rsum[i] = tot;
    }
}

// yc = A_L^-1 rs, one warp per row (16-byte loads); every CTA stages rs in shared memory.  rs and yc may not alias.
__global__ void __launch_bounds__(SYS_THREADS) coarse_solve_kernel(long long nc, const double *__restrict__ Ainv, const double *__restrict__ rsum,
                                                                  double *__restrict__ yc, const int *__restrict__ done, PeerView pv)
{
    MFV2D_DYN_SMEM(double, rs);
    if (done && *done)
        return;
    if (pv.n_ranks > 1 && *(volatile int *)pv.error)
        return;
    for (long long i = threadIdx.x; i < nc; i += blockDim.x)
        rs[i] = rsum[i];
    __syncthreads();
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    const long long pairs = nc >> 1;
    for (long long row = (long long)blockIdx.x * nwarps + warp; row < nc; row += (long long)gridDim.x * nwarps)
    {
        const double *a = Ainv + (size_t)row * nc;
        double acc0 = 0.0, acc1 = 0.0;
        if ((((size_t)row * nc) & 1) == 0)
        {
            const double2 *a2 = reinterpret_cast<const double2 *>(a);
            const double2 *r2 = reinterpret_cast<const double2 *>(rs);
            for (long long c = MFV2D_LANE0; c < pairs; c += MFV2D_LANE_STEP)
            {
                const double2 v = a2[c], w = r2[c];
                acc0 = fma(v.x, w.x, acc0);
                acc1 = fma(v.y, w.y, acc1);
            }
            if ((nc & 1) && MFV2D_LANE0 == 0)
                acc0 = fma(a[nc - 1], rs[nc - 1], acc0);
        }
        else
            for (long long c = MFV2D_LANE0; c < nc; c += MFV2D_LANE_STEP)
                acc0 = fma(a[c], rs[c], acc0);
        const double acc = warp_sum(acc0 + acc1);
        if (MFV2D_LANE0 == 0)
            yc[row] = acc;
    }
    if (pv.n_ranks > 1 && blockIdx.x == 0 && threadIdx.x == 0)
        peer_stamp(pv, 9); // CTA 0's share of the coarse mat-vec
}

// z[i] += cz[i] * sum over the levels of y_l[agg_l[i]]  (the exact level reads yc)
__global__ void coarse_prolong_kernel(long long n, int n_levels, const int32_t *__restrict__ lagg, const long long *__restrict__ agg_off,
                                      long long exact_begin, const double *__restrict__ cz, const double *__restrict__ y,
                                      const double *__restrict__ yc, double *__restrict__ z, const int *__restrict__ done)
{
    if (done && *done)
        return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        double acc = 0.0;
        for (int l = 0; l < n_levels; ++l)
        {
            const int a = lagg[(size_t)l * n + i];
            if (a >= 0)
            {
                const long long J = agg_off[l] + a;
                acc += J < exact_begin ? y[J] : yc[J - exact_begin];
            }
        }
        z[i] = fma(cz[i], acc, z[i]);
    }
}

// z += sum_l P_l D_l^-1 P_l^T r + P_L A_L^-1 P_L^T r  (`done` = the CG's device flag: a captured iteration stops by itself)
int coarse_correction(mfv2d_system *s, const double *r, double *z, const int *done, bool use_peers)
{
    mfv2d_ctx *ctx = s->ctx;
    if (!s->coarse_ready || s->n_trace == 0)
        return MFV2D_SUCCESS;
    const int warps = SYS_THREADS / 32;
    const long long cap = (long long)ctx->sm_count * 8;
    PeerView one;
    if (s->n_chunks)
    {
        MFV2D_LAUNCH(coarse_chunk_kernel, (unsigned)std::min<long long>((s->n_chunks + warps - 1) / warps, cap), SYS_THREADS, 0, ctx->stream,
                     s->n_chunks, (const int64_t *)s->d_chunk_ptr, (const int64_t *)s->d_chunk_idx, (const double *)s->d_cz,
                     (const double *)s->d_owned, r, s->d_cpart, done);
        ctx->launches++;
    }
    CoarseArgs ca{s->n_agg_total, s->exact_begin, s->n_exact, s->d_aggchunk_ptr, s->d_cpart, s->d_ldinv, s->d_cy,
                  reinterpret_cast<int *>(s->d_cg + 40), done, use_peers ? s->peer : one};
    MFV2D_LAUNCH(coarse_aggregate_kernel, grid_for(ctx, s->n_agg_total, SYS_THREADS), SYS_THREADS, 0, ctx->stream, ca);
    ctx->launches++;
    if (s->n_exact)
    {
        if (ctx->n_ranks > 1 && !use_peers)
            MFV2D_TRY(comm_allreduce(ctx, s->d_cy + s->exact_begin, (size_t)s->n_exact));
        const double *rsum = s->d_cy + s->exact_begin;
        if (ctx->n_ranks > 1 && use_peers)
        {
            MFV2D_LAUNCH(coarse_sum_kernel, grid_for(ctx, s->n_exact, SYS_THREADS), SYS_THREADS, 0, ctx->stream, s->n_exact, rsum, s->d_crs, done,
                         s->peer);
            ctx->launches++;
            rsum = s->d_crs;
        }
        const unsigned grid = (unsigned)std::min<long long>((s->n_exact + warps - 1) / warps, (long long)ctx->sm_count * 4);
        MFV2D_LAUNCH(coarse_solve_kernel, grid, SYS_THREADS, (size_t)s->n_exact * sizeof(double), ctx->stream, s->n_exact,
                     (const double *)s->d_Ac_inv, rsum, s->d_yc, done, use_peers ? s->peer : one);
        ctx->launches++;
    }
    MFV2D_LAUNCH(coarse_prolong_kernel, grid_for(ctx, s->n_trace, 256), 256, 0, ctx->stream, (long long)s->n_trace, s->n_clevels,
                 (const int32_t *)s->d_lagg, (const long long *)s->d_agg_off, s->exact_begin, (const double *)s->d_cz,
                 (const double *)s->d_cy, (const double *)s->d_yc, z, done);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

extern "C" {

// Declare the aggregate hierarchy of the multilevel preconditioner (precondition = 2 of solve_schur_cg).
//   n_levels, n_agg[n_levels]     aggregates per level
//   agg[n_levels][n_trace]        aggregate (0 .. n_agg[l] - 1, or -1) of every multiplier of THIS rank on level l
//   z[n_trace]                    near-null-space vector of S on them (identical on the ranks sharing a multiplier)
//   exact_last                    != 0: the last level is solved exactly; its aggregates are numbered GLOBALLY (the same ids
//                                 on every rank, n_agg <= 4096); all other levels are rank-local and must not contain
//                                 shared multipliers
// Collective on a partitioned system (after set_interface / peer_connect).
int mfv2d_cuda_system_set_coarse_space(mfv2d_system *s, int n_levels, const int64_t *n_agg, const int32_t *agg, const double *z,
                                       int exact_last)
{
    mfv2d_ctx *ctx = s->ctx;
    s->coarse_ready = false;
#ifndef MFV2D_EMU
    if (s->cg_graph_exec)
    {
        cudaGraphExecDestroy((cudaGraphExec_t)s->cg_graph_exec);
        s->cg_graph_exec = nullptr;
    }
#endif
    if (n_levels == 0)
        return MFV2D_SUCCESS;
    if (n_levels < 0 || n_levels > COARSE_MAX_LEVELS)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "set_coarse_space: 1 to %d levels", COARSE_MAX_LEVELS);
    MFV2D_TRY(sys_require_interface(s));
    const bool use_peers = sys_use_peers(s);
    const int64_t nt = s->n_trace;
    const int64_t n_exact = exact_last ? n_agg[n_levels - 1] : 0;
    if (n_exact > COARSE_MAX)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "set_coarse_space: the exact level has %lld aggregates, at most %d (its operator is kept as a "
                                                       "dense inverse)", (long long)n_exact, COARSE_MAX);
    if (use_peers && n_exact > s->n_coarse_cap)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "set_coarse_space: %lld global aggregates, but peer_export reserved room for %lld",
                          (long long)n_exact, (long long)s->n_coarse_cap);
    std::vector<long long> agg_off(n_levels, 0);
    long long total = 0;
    for (int l = 0; l < n_levels; ++l)
    {
        agg_off[l] = total;
        total += n_agg[l];
    }
    const long long exact_begin = total - n_exact;
    // chunk lists: the multipliers of every aggregate in increasing order, cut into chunks of COARSE_CHUNK
    std::vector<int64_t> count((size_t)total + 1, 0);
    for (int l = 0; l < n_levels; ++l)
        for (int64_t i = 0; i < nt; ++i)
        {
            const int32_t a = agg[(size_t)l * nt + i];
            if (a >= n_agg[l])
                return mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "set_coarse_space: aggregate %d of multiplier %lld on level %d is out of range", a,
                                  (long long)i, l);
            if (a >= 0)
                count[agg_off[l] + a + 1]++;
        }
    for (long long J = 0; J < total; ++J)
        count[J + 1] += count[J];
    std::vector<int64_t> idx((size_t)count[total]);
    {
        std::vector<int64_t> fill(count.begin(), count.end() - 1);
        for (int l = 0; l < n_levels; ++l)
            for (int64_t i = 0; i < nt; ++i)
            {
                const int32_t a = agg[(size_t)l * nt + i];
                if (a >= 0)
                    idx[fill[agg_off[l] + a]++] = i;
            }
    }
    std::vector<int64_t> chunk_ptr(1, 0), aggchunk_ptr((size_t)total + 1, 0);
    for (long long J = 0; J < total; ++J)
    {
        for (int64_t k = count[J]; k < count[J + 1]; k += COARSE_CHUNK)
            chunk_ptr.push_back(std::min<int64_t>(k + COARSE_CHUNK, count[J + 1]));
        aggchunk_ptr[J + 1] = (int64_t)chunk_ptr.size() - 1;
    }
    s->n_clevels = n_levels;
    s->n_agg_total = total;
    s->exact_begin = exact_begin;
    s->n_exact = n_exact;
    s->n_chunks = (long long)chunk_ptr.size() - 1;
    std::vector<int32_t> h_agg(agg, agg + (size_t)n_levels * nt);
    std::vector<double> h_z(z, z + nt);
    if (s->scaled)
    {
        // the blocks are those of D S D: its near-null-space vector is D^-1 z
        std::vector<double> rs((size_t)nt);
        MFV2D_TRY(mfv2d_cuda_download(ctx, rs.data(), s->d_rs, (size_t)nt * sizeof(double)));
        for (int64_t i = 0; i < nt; ++i)
            h_z[i] /= rs[i];
    }
    MFV2D_TRY(sys_upload(s, h_agg, &s->d_lagg));
    MFV2D_TRY(sys_upload(s, agg_off, &s->d_agg_off));
    MFV2D_TRY(sys_upload(s, h_z, &s->d_cz));
    MFV2D_TRY(sys_upload(s, chunk_ptr, &s->d_chunk_ptr));
    MFV2D_TRY(sys_upload(s, idx, &s->d_chunk_idx));
    MFV2D_TRY(sys_upload(s, aggchunk_ptr, &s->d_aggchunk_ptr));
    MFV2D_TRY(sys_alloc(s, (size_t)s->n_chunks, &s->d_cpart));
    MFV2D_TRY(sys_alloc(s, (size_t)total, &s->d_cy));
    MFV2D_TRY(sys_alloc(s, (size_t)exact_begin, &s->d_ldinv));
    MFV2D_TRY(sys_alloc(s, (size_t)n_exact, &s->d_yc));
    MFV2D_TRY(sys_alloc(s, (size_t)n_exact, &s->d_crs));
    long long grid = (long long)ctx->sm_count * 8;
    if (grid > s->n_elem)
        grid = s->n_elem;
    // diagonal levels: D_l = diag(P_l^T S P_l) -> 1 / D_l (1 where an aggregate's diagonal vanishes)
    if (exact_begin)
        MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(s->d_ldinv, 0, (size_t)exact_begin * sizeof(double), ctx->stream));
    for (int l = 0; l < n_levels - (exact_last ? 1 : 0) && s->n_elem; ++l)
    {
        MFV2D_LAUNCH(coarse_assemble_kernel, (unsigned)grid, 128, 0, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off, s->d_slot_trace,
                     s->d_s_off, (const double *)s->d_S, (int)s->symmetric, (const int32_t *)(s->d_lagg + (size_t)l * nt), (const double *)s->d_cz,
                     (long long)n_agg[l], 0, s->d_ldinv + agg_off[l]);
        ctx->launches++;
    }
    if (exact_begin)
    {
        MFV2D_LAUNCH(reciprocal_kernel, grid_for(ctx, exact_begin, 256), 256, 0, ctx->stream, exact_begin, s->d_ldinv);
        ctx->launches++;
    }
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    if (n_exact)
    {
        const size_t nn = (size_t)n_exact * n_exact;
        double *d_Ac = nullptr;
        int32_t *d_status = nullptr;
        MFV2D_TRY(sys_alloc(s, nn, &s->d_Ac_inv));
        MFV2D_TRY(sys_alloc(s, (size_t)2, &d_status));
        d_Ac = s->d_Ac_inv; // assembled, summed and inverted in place
        MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(d_Ac, 0, nn * sizeof(double), ctx->stream));
        if (s->n_elem)
        {
            MFV2D_LAUNCH(coarse_assemble_kernel, (unsigned)grid, 128, 0, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off,
                         s->d_slot_trace, s->d_s_off, (const double *)s->d_S, (int)s->symmetric,
                         (const int32_t *)(s->d_lagg + (size_t)(n_levels - 1) * nt), (const double *)s->d_cz, (long long)n_exact, 1, d_Ac);
            ctx->launches++;
            MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
        }
        if (ctx->n_ranks > 1)
        {
            if (use_peers) // row by row through the type-1 area of the exchange blocks
                for (int64_t row = 0; row < n_exact; ++row)
                    MFV2D_TRY(peer_dense_sum(s, d_Ac + (size_t)row * n_exact, (long long)n_exact));
            else
                MFV2D_TRY(comm_allreduce(ctx, d_Ac, nn));
        }
        // P^T S P is definite like S: no pivoting, every elimination step runs on the whole device
        MFV2D_TRY(launch_inverse_definite_large(ctx, (int)n_exact, d_Ac, d_status));
        int32_t st = 0;
        MFV2D_TRY(mfv2d_cuda_download(ctx, &st, d_status, sizeof st));
        if (st != MFV2D_SUCCESS)
            return mfv2d_fail(ctx, MFV2D_PIVOT_FAILED, "set_coarse_space: the coarse operator P^T S P is singular (an aggregate without multipliers, "
                                                       "or a Schur complement with a null space along z)");
        const size_t smem = (size_t)n_exact * sizeof(double);
        if (smem > 48 * 1024)
            MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(coarse_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    MFV2D_TRY(mfv2d_cuda_synchronize(ctx));
    s->coarse_ready = true;
    return MFV2D_SUCCESS;
}

} // extern "C"
