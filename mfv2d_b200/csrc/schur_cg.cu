// solve_schur_iterative (python/mfv2d/solving.py:439-500) with cg_general (:350-436) on the device:
// conjugate gradients on the Schur complement S = N A^-1 N^T of the multipliers, then the back substitution
// x = A^-1 (y - N^T lambda).
//
// One iteration is THREE kernels, on one GPU and on a partitioned system alike (Chronopoulos & Gear 1989: with
// u = M^-1 r and w = S u the scalars gamma = r.u, delta = u.w and rho = r.r are needed at the same point):
//   schur_local (system.cu)   per-element products S_e u_e (packed symmetric blocks: the HBM-bound part)
//   cg1_gather_kernel         w = sum of the element contributions per multiplier, partial sums of (delta, gamma, rho);
//                             on a partitioned system every thread WRITES the contributions of shared multipliers
//                             straight into the exchange blocks of the ranks that hold them (NVLink peer stores) and the
//                             last CTA to finish reduces the partials, sends the three scalars and releases the epoch flag
//   cg1_update_kernel         acquires the peers' flags, sums interface contributions and scalars in rank order
//                             (bit-identical on all ranks), then beta = gamma / gamma_old,
//                             alpha = gamma / (delta - beta gamma / alpha_old), p = u + beta p, s = w + beta s,
//                             x += alpha p, r -= alpha s, u = M^-1 r (Jacobi fused; block-Jacobi / two-level: precond.cu)
// The exchange is part of the compute kernels: no NCCL call, memset, pack / unpack or device-to-device copy per
// iteration (round 1: 12 graph nodes per iteration on N GPUs, 7 on one).  MFV2D_CG_EXCHANGE=nccl keeps ncclAllReduce
// between the two kernels for comparison.  16 iterations are captured ONCE per system in a CUDA graph (all pointers
// are the system's own buffers) and replayed; the host polls the device flag once per replay.
#include <algorithm>
#include <cmath>

#include "system.cuh"

int coarse_correction(mfv2d_system *s, const double *r, double *z, const int *done, bool use_peers); // precond.cu
extern "C" int mfv2d_cuda_system_peer_error(mfv2d_system *s);
__global__ void jacobi_setup_kernel(long long n, double *__restrict__ d, const double *__restrict__ r, double *__restrict__ z);

// device state of the solver in s->d_cg (64 doubles):
//   sc[0] rho = r.r   sc[1] delta   sc[2] alpha   sc[3] beta   sc[4] tol^2   sc[8..10] this rank's (delta, gamma, rho)
//   sc[12 + 2 (k & 1)] = (gamma, alpha) of update k - 1 (double buffered: update k reads one pair, writes the other)
//   flags = (int *)(sc + 16): [0] done  [1] the reference's iter_cnt  [2] breakdown  [3] updates started  [4] CTA tickets
enum
{
    CG_RHO = 0,
    CG_DELTA = 1,
    CG_ALPHA = 2,
    CG_BETA = 3,
    CG_TOL2 = 4,
    CG_LOCAL = 8,
    CG_STATE = 12,
    CG_FLAGS = 16
};

struct CgGatherArgs
{
    long long n_trace;
    const int64_t *tr_ptr, *tr_slot;
    const int32_t *tr2; // fixed-width (two slots per multiplier) form of the same lists, or nullptr
    const double *slot_val, *u, *r, *own;
    double *w, *partials, *sc;
    const int32_t *shared_of;
    const int64_t *shared_slot, *holder_ptr;
    const int32_t *holder_rank;
    PeerView pv; // n_ranks == 1: nothing is sent (one GPU, or the NCCL all-reduce follows)
};

__global__ void __launch_bounds__(SYS_THREADS) cg1_gather_kernel(const CgGatherArgs a)
{
    int *flags = reinterpret_cast<int *>(a.sc + CG_FLAGS);
    if (flags[0])
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
        e = pv.epoch[0] + 1;
        parity = (int)(e & 1);
        if (blockIdx.x == 0 && threadIdx.x == 0)
            peer_stamp(pv, 0); // since the previous stamp: the tail of the last iteration + the condensed-block product
    }
    double d = 0.0, g = 0.0, rho = 0.0;
    bool wrote = false;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < a.n_trace; i += (long long)gridDim.x * blockDim.x)
    {
        double acc = 0.0;
        if (a.tr2)
        {
            const int2 sl = reinterpret_cast<const int2 *>(a.tr2)[i];
            if (sl.x >= 0)
                acc += a.slot_val[sl.x];
            if (sl.y >= 0)
                acc += a.slot_val[sl.y];
        }
        else
            for (int64_t k = a.tr_ptr[i]; k < a.tr_ptr[i + 1]; ++k)
                acc += a.slot_val[a.tr_slot[k]];
        a.w[i] = acc;
        const double ui = a.u[i], ri = a.r[i];
        const double rw = a.own ? ri * a.own[i] : ri;
        d = fma(ui, acc, d); // local, un-summed product: summed over the ranks it is u . S u
        g = fma(rw, ui, g);
        rho = fma(rw, ri, rho);
        if (peers)
        {
            const int j = a.shared_of[i];
            if (j >= 0)
            {
                const long long sl = a.shared_slot[j];
                for (int64_t h = a.holder_ptr[j]; h < a.holder_ptr[j + 1]; ++h)
                    peer_region(pv, a.holder_rank[h], 0, parity, pv.rank)[sl] = acc;
                wrote = true;
            }
        }
    }
    if (wrote && !(pv.dbg & 1))
        peer_fence(); // every writer orders its own remote stores before the CTA's ticket
    const int cta_wrote = peers ? __syncthreads_or(wrote ? 1 : 0) : 0;
    const int np = gridDim.x;
    block_reduce_store(d, a.partials);
    block_reduce_store(g, a.partials + np);
    block_reduce_store(rho, a.partials + 2 * np);
    // the last CTA to arrive reduces the partials in a fixed order and publishes
    if (threadIdx.x == 0)
    {
        if (cta_wrote)
            peer_fence(); // a CTA without remote stores only publishes its partials: device scope is enough
        else
            __threadfence();
        s_last = atomicAdd(&flags[4], 1) == np - 1;
    }
    __syncthreads();
    if (!s_last)
        return;
    __threadfence();
    if (threadIdx.x < 32)
    {
        const double sd = warp_sum_partials(np, a.partials);
        const double sg = warp_sum_partials(np, a.partials + np);
        const double sr = warp_sum_partials(np, a.partials + 2 * np);
        if (threadIdx.x == 0)
        {
            a.sc[CG_LOCAL + 0] = sd;
            a.sc[CG_LOCAL + 1] = sg;
            a.sc[CG_LOCAL + 2] = sr;
            flags[4] = 0;
            flags[3] += 1;
            if (peers)
            {
                for (int r = 0; r < pv.n_ranks; ++r)
                    if (r != pv.rank)
                    {
                        double *dst = peer_region(pv, r, 0, parity, pv.rank) + pv.n_iface;
                        dst[0] = sd;
                        dst[1] = sg;
                        dst[2] = sr;
                    }
                peer_fence();
                for (int r = 0; r < pv.n_ranks; ++r)
                    if (r != pv.rank)
                        peer_store_flag(peer_flag(pv, r, 0, pv.rank), e);
                pv.epoch[0] = e;
                peer_stamp(pv, 1); // gather + reduction + publication
            }
        }
    }
}

struct CgUpdateArgs
{
    long long n_trace;
    double *sc;
    const double *u; // may alias r (no preconditioner): every entry is read before it is written by the same thread
    const double *w;
    double *p, *sv, *x, *r;
    const double *dinv; // Jacobi: z = dinv r fused here; nullptr otherwise
    double *z;
    const int32_t *shared_of;
    const int64_t *shared_slot, *holder_ptr;
    const int32_t *holder_rank;
    PeerView pv;
};

__global__ void __launch_bounds__(SYS_THREADS) cg1_update_kernel(const CgUpdateArgs a)
{
    int *flags = reinterpret_cast<int *>(a.sc + CG_FLAGS);
    if (flags[0])
        return;
    const PeerView &pv = a.pv;
    const bool peers = pv.n_ranks > 1;
    const int k = flags[3] - 1; // updates carried out before this one
    int parity = 0;
    if (peers)
    {
        const unsigned long long e = pv.epoch[0];
        parity = (int)(e & 1);
        if (blockIdx.x == 0 && threadIdx.x == 0)
            peer_stamp(pv, 2); // launch gap between the two kernels
        peer_wait_all(pv, 0, e);
        if (blockIdx.x == 0 && threadIdx.x == 0)
            peer_stamp(pv, 3); // waiting for the peers' flags
        if (*(volatile int *)pv.error)
            return;
    }
    double delta = 0.0, gamma = 0.0, rho = 0.0;
    for (int r = 0; r < pv.n_ranks; ++r)
    {
        if (r == pv.rank)
        {
            delta += a.sc[CG_LOCAL + 0];
            gamma += a.sc[CG_LOCAL + 1];
            rho += a.sc[CG_LOCAL + 2];
        }
        else
        {
            const double *src = peer_region(pv, pv.rank, 0, parity, r) + pv.n_iface;
            delta += peer_load(src + 0);
            gamma += peer_load(src + 1);
            rho += peer_load(src + 2);
        }
    }
    // rho belongs to the residual after update k - 1
    const bool converged = k > 0 && rho < a.sc[CG_TOL2];
    const double *st = a.sc + CG_STATE + 2 * (k & 1);
    const double beta = k > 0 ? gamma / st[0] : 0.0;
    const double alpha = k > 0 ? gamma / (delta - beta * gamma / st[1]) : gamma / delta;
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        a.sc[CG_RHO] = rho;
        if (k > 0)
            flags[1] = k - 1;
        if (converged)
            flags[0] = 1;
        else
        {
            double *sn = a.sc + CG_STATE + 2 * ((k + 1) & 1);
            a.sc[CG_DELTA] = delta;
            a.sc[CG_ALPHA] = alpha;
            a.sc[CG_BETA] = beta;
            sn[0] = gamma;
            sn[1] = alpha;
            if (!(alpha == alpha) || !(fabs(alpha) <= 1.7976931348623157e308))
                flags[2] = 1; // breakdown (delta = 0: S is singular along u, or not definite)
        }
    }
    if (converged)
        return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < a.n_trace; i += (long long)gridDim.x * blockDim.x)
    {
        double wi = a.w[i];
        if (peers)
        {
            const int j = a.shared_of[i];
            if (j >= 0)
            {
                const long long sl = a.shared_slot[j];
                double tot = 0.0;
                bool added = false;
                for (int64_t h = a.holder_ptr[j]; h < a.holder_ptr[j + 1]; ++h)
                {
                    const int r = a.holder_rank[h];
                    if (!added && r > pv.rank)
                    {
                        tot += wi;
                        added = true;
                    }
                    tot += peer_load(peer_region(pv, pv.rank, 0, parity, r) + sl);
                }
                if (!added)
                    tot += wi;
                wi = tot;
            }
        }
        const double pi = fma(beta, a.p[i], a.u[i]);
        const double si = fma(beta, a.sv[i], wi);
        a.p[i] = pi;
        a.sv[i] = si;
        a.x[i] = fma(alpha, pi, a.x[i]);
        const double ri = fma(-alpha, si, a.r[i]);
        a.r[i] = ri;
        if (a.dinv)
            a.z[i] = ri * a.dinv[i];
    }
    if (peers && blockIdx.x == 0 && threadIdx.x == 0)
        peer_stamp(pv, 4); // CTA 0's share of the update
}

// d[i] = 1 / d[i] (1 where the diagonal vanishes); z[i] = d[i] r[i]
__global__ void jacobi_setup_kernel(long long n, double *__restrict__ d, const double *__restrict__ r, double *__restrict__ z)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double v = d[i];
        const double inv = v != 0.0 ? 1.0 / v : 1.0;
        d[i] = inv;
        z[i] = inv * r[i];
    }
}

extern "C" {

// precondition: 0 none (the reference's cg_general), 1 Jacobi (diag S), 2 Jacobi + coarse space (additive two-level;
// needs mfv2d_cuda_system_set_coarse_space)
int mfv2d_cuda_solve_schur_cg(mfv2d_system *s, const double *rhs, const double *phi, int max_iter, double atol, double rtol,
                              int precondition, double *x, double *lambda, double *residual, int *iterations, int space)
{
    mfv2d_ctx *ctx = s->ctx;
    const long long nt = s->n_trace, nd = s->n_dense;
    if ((nd && (!rhs || !x)) || (nt && (!phi || !lambda)))
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_cg: rhs, phi, x and lambda must not be NULL");
    if (precondition < 0 || precondition > 2)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_cg: unknown preconditioner %d", precondition);
    MFV2D_TRY(sys_require_interface(s));
    Staged srhs, sphi, sx, slam;
    int r = stage_in(ctx, rhs, nd * sizeof(double), space, &srhs);
    if (!r)
        r = stage_in(ctx, phi, nt * sizeof(double), space, &sphi);
    if (!r)
        r = stage_out_alloc(ctx, x, nd * sizeof(double), space, &sx);
    if (!r)
        r = stage_out_alloc(ctx, lambda, nt * sizeof(double), space, &slam);
    double *inv_a_y = s->d_work_dense[0], *tmp_d = s->d_work_dense[1], *tmp_d2 = s->d_work_dense[2];
    double *res = s->d_work_trace[0], *p = s->d_work_trace[1], *ap = s->d_work_trace[2], *z = s->d_work_trace[3];
    double *sv = s->d_work_trace[4], *lam = s->d_work_trace[5];
    double *sc = s->d_cg;
    int *flags = reinterpret_cast<int *>(sc + CG_FLAGS);
    const double *own = s->d_owned;
    const bool use_peers = sys_use_peers(s);
    double h[24];
    int it_done = 0;
    double res_mag2 = 0.0;
    do
    {
        if (r)
            break;
        // A^-1 y ; N A^-1 y - phi
        if ((r = sys_gemv(s, s->d_Ainv, (const double *)srhs.dev, inv_a_y)))
            break;
        if ((r = sys_apply_trace(s, inv_a_y, res)))
            break;
        if (nt)
        {
            MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nt, 256), 256, 0, ctx->stream, nt, 1.0, (const double *)res, -1.0,
                         (const double *)sphi.dev, res);
            ctx->launches++;
        }
        if ((r = sys_scale_rows(s, res, false))) // the CG runs on the equilibrated system D S D (D lambda') = D b
            break;
        if (cudaMemsetAsync(lam, 0, (nt ? nt : 1) * sizeof(double), ctx->stream) != cudaSuccess ||
            cudaMemsetAsync(sc, 0, 32 * sizeof(double), ctx->stream) != cudaSuccess)
        {
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_cg: memset failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        // res_mag2 = rhs . rhs ; tolerance (solving.py:405-412)
        if (nt)
        {
            if ((r = sys_dot_device(s, nt, res, res, own, sc + CG_RHO)))
                break;
            if ((r = read_scalars(ctx, sc, 1, h)))
                break;
            res_mag2 = h[0];
        }
        const double mag = sqrt(res_mag2);
        const double tol = (mag * rtol > atol) ? atol : mag * rtol;
        h[4] = tol * tol;
        if (nt && max_iter > 0 && res_mag2 > 0.0)
        {
            if (cudaMemcpyAsync(sc + CG_TOL2, &h[4], sizeof(double), cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess ||
                cudaStreamSynchronize(ctx->stream) != cudaSuccess)
            {
                r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_cg: scalar upload failed: %s", cudaGetErrorString(cudaGetLastError()));
                break;
            }
            const double *dinv = nullptr;
            if (precondition >= 1)
            {
                // Jacobi on the trace space (not in the reference): diag(S) from the diagonals of the condensed blocks,
                // summed per multiplier and over the ranks; z0 = D^-1 r0
                if ((r = sys_schur_diagonal(s, s->d_diag_inv)))
                    break;
                MFV2D_LAUNCH(jacobi_setup_kernel, grid_for(ctx, nt, 256), 256, 0, ctx->stream, nt, s->d_diag_inv, (const double *)res, z);
                ctx->launches++;
                dinv = s->d_diag_inv;
            }
            if (precondition >= 2)
            {
                if (!s->coarse_ready)
                {
                    r = mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_cg: the two-level preconditioner needs mfv2d_cuda_system_set_coarse_space");
                    break;
                }
                if ((r = coarse_correction(s, res, z, nullptr, use_peers))) // z0 = D^-1 r0 + P Ac^-1 P^T r0
                    break;
            }
            const double *u = precondition ? z : res;
            if (cudaMemsetAsync(p, 0, nt * sizeof(double), ctx->stream) != cudaSuccess ||
                cudaMemsetAsync(sv, 0, nt * sizeof(double), ctx->stream) != cudaSuccess)
            {
                r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_cg: memset failed: %s", cudaGetErrorString(cudaGetLastError()));
                break;
            }
            PeerView one; // a view with one rank: nothing is sent or awaited inside the kernels
            CgGatherArgs ga{nt, s->d_tr_ptr, s->d_tr_slot, s->d_tr2, s->d_slot_val, u, res, own, ap, s->d_partials, sc, s->d_shared_of,
                            s->d_shared_slot, s->d_holder_ptr, s->d_holder_rank, use_peers ? s->peer : one};
            CgUpdateArgs ua{nt, sc, u, ap, p, sv, lam, res, dinv, z, s->d_shared_of, s->d_shared_slot, s->d_holder_ptr, s->d_holder_rank,
                            use_peers ? s->peer : one};
            const int check_every = 16;
            auto enqueue_iteration = [&]() -> int {
                MFV2D_TRY(sys_schur_local(s, u, flags));
                MFV2D_LAUNCH(cg1_gather_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, ga);
                if (ctx->n_ranks > 1 && !use_peers)
                    MFV2D_TRY(sys_interface_sum(s, ap, 3, sc + CG_LOCAL)); // ncclAllReduce of interface values + 3 scalars
                MFV2D_LAUNCH(cg1_update_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, ua);
                ctx->launches += 2;
                if (precondition >= 2)
                    MFV2D_TRY(coarse_correction(s, res, z, flags, use_peers));
                return MFV2D_SUCCESS;
            };
#ifndef MFV2D_EMU
            // the graph of check_every iterations is built once per (system, preconditioner, exchange) and kept
            cudaGraphExec_t graph_exec = nullptr;
            const int key = precondition * 4 + (use_peers ? 2 : 0) + (ctx->n_ranks > 1 ? 1 : 0);
            const char *graph_env = getenv("MFV2D_CG_GRAPH");
            uint64_t launches_per_block = 0;
            if (!(graph_env && graph_env[0] == '0') && max_iter >= 2 * check_every)
            {
                if (s->cg_graph_exec && s->cg_graph_key != key)
                {
                    cudaGraphExecDestroy((cudaGraphExec_t)s->cg_graph_exec);
                    s->cg_graph_exec = nullptr;
                }
                const uint64_t launches_before = ctx->launches;
                if (!s->cg_graph_exec)
                {
                    cudaGraph_t graph = nullptr;
                    if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess)
                    {
                        int rc = MFV2D_SUCCESS;
                        for (int k = 0; k < check_every && rc == MFV2D_SUCCESS; ++k)
                            rc = enqueue_iteration();
                        const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
                        if (rc != MFV2D_SUCCESS || ce != cudaSuccess || cudaGraphInstantiate(&graph_exec, graph, 0) != cudaSuccess)
                            graph_exec = nullptr;
                        if (graph)
                            cudaGraphDestroy(graph);
                    }
                    cudaGetLastError();
                    s->cg_graph_exec = graph_exec;
                    s->cg_graph_key = key;
                    // launches of one replay, counted from the capture (capturing is not launching)
                    s->cg_graph_launches = ctx->launches - launches_before;
                    ctx->launches = launches_before;
                }
                graph_exec = (cudaGraphExec_t)s->cg_graph_exec;
                launches_per_block = s->cg_graph_launches;
            }
#endif
            int it = 0;
            bool finished = false;
            while (it < max_iter && !finished)
            {
                const int stop = (it + check_every < max_iter) ? it + check_every : max_iter;
#ifndef MFV2D_EMU
                if (graph_exec && stop - it == check_every)
                {
                    if (cudaGraphLaunch(graph_exec, ctx->stream) != cudaSuccess)
                    {
                        r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "CG graph launch failed: %s", cudaGetErrorString(cudaGetLastError()));
                        break;
                    }
                    ctx->launches += launches_per_block;
                    it = stop;
                }
#endif
                for (; it < stop && !r; ++it)
                    r = enqueue_iteration();
                if (r)
                    break;
                if (cudaGetLastError() != cudaSuccess)
                {
                    r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "CG kernel launch failed");
                    break;
                }
                if ((r = read_scalars(ctx, sc, 20, h)))
                    break;
                int hf[8];
                memcpy(hf, &h[CG_FLAGS], sizeof(int) * 4);
                res_mag2 = h[CG_RHO];
                it_done = hf[1];
                finished = hf[0] != 0;
                if (hf[2])
                {
                    r = mfv2d_fail(ctx, MFV2D_UNSPECIFIED_ERROR, "Schur CG broke down (u . S u = 0 or not finite): the condensed operator is singular or indefinite");
                    break;
                }
                if (!finished)
                    it_done = it - 1; // the one-synchronisation recurrence tests the residual one product later
            }
            if (r)
                break;
            if (use_peers && mfv2d_cuda_system_peer_error(s))
            {
                r = mfv2d_fail(ctx, MFV2D_CUDA_NCCL_ERROR, "peer exchange timed out: a rank did not reach the same iteration");
                break;
            }
            if (!finished)
            {
                // iterations exhausted: the residual of the LAST update has not been formed yet - do it now, so that
                // (lambda, residual, iterations) are consistent
                if ((r = sys_dot_device(s, nt, res, res, own, sc + CG_RHO)))
                    break;
                if ((r = read_scalars(ctx, sc, 1, h)))
                    break;
                res_mag2 = h[0];
            }
        }
        if ((r = sys_scale_rows(s, lam, false))) // lambda = D lambda'
            break;
        if (nt && cudaMemcpyAsync(slam.dev, lam, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess)
        {
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_cg: copy of lambda failed: %s", cudaGetErrorString(cudaGetLastError()));
            break;
        }
        // x = A^-1 y - A^-1 N^T lambda
        if ((r = sys_spmv(s, nd, s->d_t_ptr, s->d_t_col, s->d_t_val, lam, tmp_d)))
            break;
        if ((r = sys_gemv(s, s->d_Ainv, tmp_d, tmp_d2)))
            break;
        if (nd)
        {
            MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)inv_a_y, -1.0,
                         (const double *)tmp_d2, (double *)sx.dev);
            ctx->launches++;
            // one step of iterative refinement of the element-local solves, x += A^-1 (y - N^T lambda - A x): the blocks
            // are applied as explicit inverses, whose residual grows with cond(A_e) (the 1-form and 2-form masses of a
            // small element differ by 1/h^2); MFV2D_SCHUR_REFINE=0 turns it off
            const char *rf_env = getenv("MFV2D_SCHUR_REFINE");
            if (!(rf_env && rf_env[0] == '0'))
            {
                if ((r = sys_gemv(s, s->d_A, (const double *)sx.dev, tmp_d2)))
                    break;
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)srhs.dev, -1.0,
                             (const double *)tmp_d2, tmp_d2);
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)tmp_d2, -1.0,
                             (const double *)tmp_d, tmp_d2);
                if ((r = sys_gemv(s, s->d_Ainv, tmp_d2, inv_a_y)))
                    break;
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)sx.dev, 1.0,
                             (const double *)inv_a_y, (double *)sx.dev);
                ctx->launches += 3;
            }
        }
    } while (0);
    if (!r)
    {
        r = stage_finish(&sx);
        const int r2 = stage_finish(&slam);
        if (!r)
            r = r2;
        if (!r)
            r = mfv2d_cuda_synchronize(ctx);
    }
    else
    {
        stage_abort(&sx);
        stage_abort(&slam);
    }
    stage_abort(&srhs);
    stage_abort(&sphi);
    if (residual)
        *residual = sqrt(res_mag2);
    if (iterations)
        *iterations = it_done;
    return r;
}

} // extern "C"
