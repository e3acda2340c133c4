// BiCGStab on the Schur complement S = N A^-1 N^T of the multipliers for systems whose element blocks are NOT symmetric
// (advection terms: linear advection-diffusion, linearised Navier-Stokes), then the back substitution
// x = A^-1 (y - N^T lambda).  The reference has no iterative solver for such systems (solve_schur_iterative uses CG,
// python/mfv2d/solving.py:439-500; its production path is SuperLU, solve_system_2d.py:258-282); this is the device
// counterpart that lets the implicit time march of BASELINE.json configs[2] run without a global factorisation.
// All scalars (rho, alpha, omega, beta, the residual and the stopping flag) live on the device; the host reads them back
// once per block of iterations.  Right-preconditioned with the Jacobi diagonal of S (precondition = 1) or with the additive
// multilevel preconditioner of precond.cu (2: Jacobi + aggregate levels, after mfv2d_cuda_system_set_coarse_space).
#include <cmath>

#include "system.cuh"

// sc: [0] rr  [1] rho  [2] rho_old  [3] alpha  [4] omega  [5] beta  [6] tol^2  [7] r0.v  [8] t.s  [9] t.t ; flags at sc + 16
enum
{
    B_RR = 0,
    B_RHO = 1,
    B_RHO_OLD = 2,
    B_ALPHA = 3,
    B_OMEGA = 4,
    B_BETA = 5,
    B_TOL2 = 6,
    B_R0V = 7,
    B_TS = 8,
    B_TT = 9,
    B_FLAGS = 16
};

// partials of up to two ownership-weighted dot products
__global__ void __launch_bounds__(SYS_THREADS) bicg_dot2_kernel(long long n, const double *__restrict__ x1, const double *__restrict__ y1,
                                                               const double *__restrict__ x2, const double *__restrict__ y2,
                                                               const double *__restrict__ w, double *__restrict__ partials,
                                                               const int *__restrict__ done)
{
    if (*done)
        return;
    double a = 0.0, b = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double wi = w ? w[i] : 1.0;
        a = fma(x1[i] * wi, y1[i], a);
        if (x2)
            b = fma(x2[i] * wi, y2[i], b);
    }
    block_reduce_store(a, partials);
    block_reduce_store(b, partials + gridDim.x);
}

__global__ void bicg_sum2_kernel(int np, const double *__restrict__ partials, double *__restrict__ out1, double *__restrict__ out2,
                                 const int *__restrict__ done)
{
    if (*done)
        return;
    const double a = warp_sum_partials(np, partials);
    const double b = warp_sum_partials(np, partials + np);
    if (threadIdx.x == 0)
    {
        *out1 = a;
        if (out2)
            *out2 = b;
    }
}

// the scalar recurrences, one thread; step 0: beta from rho, 1: alpha from r0.v, 2: omega from t.s / t.t, 3: convergence test
__global__ void bicg_scalar_kernel(double *__restrict__ sc, int step, int first)
{
    int *flags = reinterpret_cast<int *>(sc + B_FLAGS);
    if (threadIdx.x != 0 || flags[0])
        return;
    if (step == 0)
    {
        sc[B_BETA] = first ? 0.0 : (sc[B_RHO] / sc[B_RHO_OLD]) * (sc[B_ALPHA] / sc[B_OMEGA]);
        sc[B_RHO_OLD] = sc[B_RHO];
    }
    else if (step == 1)
        sc[B_ALPHA] = sc[B_RHO] / sc[B_R0V];
    else if (step == 2)
        sc[B_OMEGA] = sc[B_TT] != 0.0 ? sc[B_TS] / sc[B_TT] : 0.0;
    else
    {
        flags[1] += 1; // iterations completed
        if (sc[B_RR] < sc[B_TOL2])
            flags[0] = 1;
        else if (!(sc[B_RR] == sc[B_RR]) || sc[B_RHO] == 0.0 || sc[B_OMEGA] == 0.0)
            flags[0] = flags[2] = 1; // breakdown
    }
}

// out = x + (sy * *cy) y + (sz * *cz) z   (cz == nullptr: two terms); dinv != nullptr additionally writes pre = dinv * out
__global__ void bicg_combine_kernel(long long n, double *out, const double *x, const double *y, const double *z, const double *cy, double sy,
                                    const double *cz, double sz, const double *__restrict__ dinv, double *__restrict__ pre,
                                    const int *__restrict__ done)
{
    if (*done)
        return;
    const double a = sy * cy[0], b = cz ? sz * cz[0] : 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        double v = fma(a, y[i], x[i]);
        if (cz)
            v = fma(b, z[i], v);
        out[i] = v;
        if (pre)
            pre[i] = dinv ? v * dinv[i] : v;
    }
}

// p = r + beta (p - omega v)
__global__ void bicg_direction_kernel(long long n, const double *__restrict__ sc, const double *__restrict__ r, const double *__restrict__ v,
                                      double *__restrict__ p, const double *__restrict__ dinv, double *__restrict__ pre,
                                      const int *__restrict__ done)
{
    if (*done)
        return;
    const double beta = sc[B_BETA], omega = sc[B_OMEGA];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double pi = fma(beta, p[i] - omega * v[i], r[i]);
        p[i] = pi;
        pre[i] = dinv ? pi * dinv[i] : pi;
    }
}

__global__ void jacobi_setup_kernel(long long n, double *__restrict__ d, const double *__restrict__ r, double *__restrict__ z);
int coarse_correction(mfv2d_system *s, const double *r, double *z, const int *done, bool use_peers); // precond.cu

extern "C" int mfv2d_cuda_solve_schur_bicgstab(mfv2d_system *s, const double *rhs, const double *phi, int max_iter, double atol, double rtol,
                                               int precondition, double *x, double *lambda, double *residual, int *iterations, int space)
{
    mfv2d_ctx *ctx = s->ctx;
    const long long nt = s->n_trace, nd = s->n_dense;
    if ((nd && (!rhs || !x)) || (nt && (!phi || !lambda)))
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_bicgstab: rhs, phi, x and lambda must not be NULL");
    if (precondition < 0 || precondition > 2)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_bicgstab: unknown preconditioner %d", precondition);
    if (precondition == 2 && !s->coarse_ready)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_schur_bicgstab: the multilevel preconditioner needs mfv2d_cuda_system_set_coarse_space");
    const bool use_peers = sys_use_peers(s);
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
    double *res = s->d_work_trace[0], *r0 = s->d_work_trace[1], *p = s->d_work_trace[2], *v = s->d_work_trace[3], *sv = s->d_work_trace[4],
           *lam = s->d_work_trace[5], *t = s->d_work_trace[6], *hat = s->d_work_trace[7];
    double *sc = s->d_cg;
    int *flags = reinterpret_cast<int *>(sc + B_FLAGS);
    const double *own = s->d_owned;
    const int np = s->n_partials;
    double h[24];
    int it_done = 0;
    double res_mag2 = 0.0;
    auto dot2 = [&](const double *x1, const double *y1, const double *x2, const double *y2, double *o1, double *o2) -> int {
        MFV2D_LAUNCH(bicg_dot2_kernel, (unsigned)np, SYS_THREADS, 0, ctx->stream, nt, x1, y1, x2, y2, own, s->d_partials, (const int *)flags);
        MFV2D_LAUNCH(bicg_sum2_kernel, 1, 32, 0, ctx->stream, np, (const double *)s->d_partials, o1, o2, (const int *)flags);
        ctx->launches += 2;
        if (ctx->n_ranks > 1)
        {
            // the two partial sums of all ranks (peer exchange or NCCL); harmless once the flag is set
            if (sys_use_peers(s))
                MFV2D_TRY(peer_interface_sum(s, nullptr, 1, o1));
            else
                MFV2D_TRY(comm_allreduce(ctx, o1, 1));
            if (o2)
            {
                if (sys_use_peers(s))
                    MFV2D_TRY(peer_interface_sum(s, nullptr, 1, o2));
                else
                    MFV2D_TRY(comm_allreduce(ctx, o2, 1));
            }
        }
        return MFV2D_SUCCESS;
    };
    const unsigned gt = grid_for(ctx, nt, 256);
    do
    {
        if (r)
            break;
        if ((r = sys_gemv(s, s->d_Ainv, (const double *)srhs.dev, inv_a_y)))
            break;
        if ((r = sys_apply_trace(s, inv_a_y, res)))
            break;
        if (nt)
        {
            MFV2D_LAUNCH(axpby_kernel, gt, 256, 0, ctx->stream, nt, 1.0, (const double *)res, -1.0, (const double *)sphi.dev, res);
            ctx->launches++;
        }
        if ((r = sys_scale_rows(s, res, false)))
            break;
        if (cudaMemsetAsync(lam, 0, (nt ? nt : 1) * sizeof(double), ctx->stream) != cudaSuccess ||
            cudaMemsetAsync(sc, 0, 32 * sizeof(double), ctx->stream) != cudaSuccess)
        {
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_bicgstab: memset failed");
            break;
        }
        if (nt)
        {
            if ((r = sys_dot_device(s, nt, res, res, own, sc + B_RR)))
                break;
            if ((r = read_scalars(ctx, sc, 1, h)))
                break;
            res_mag2 = h[0];
        }
        const double mag = sqrt(res_mag2);
        const double tol = (mag * rtol > atol) ? atol : mag * rtol;
        if (nt && max_iter > 0 && res_mag2 > 0.0)
        {
            const double tol2 = tol * tol;
            cudaMemcpyAsync(sc + B_TOL2, &tol2, sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
            cudaMemcpyAsync(r0, res, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            cudaMemsetAsync(p, 0, nt * sizeof(double), ctx->stream);
            cudaMemsetAsync(v, 0, nt * sizeof(double), ctx->stream);
            cudaStreamSynchronize(ctx->stream);
            const double *dinv = nullptr;
            if (precondition)
            {
                if ((r = sys_schur_diagonal(s, s->d_diag_inv)))
                    break;
                MFV2D_LAUNCH(jacobi_setup_kernel, gt, 256, 0, ctx->stream, nt, s->d_diag_inv, (const double *)res, hat);
                ctx->launches++;
                dinv = s->d_diag_inv;
            }
            int it = 0;
            bool finished = false;
            while (it < max_iter && !finished && !r)
            {
                const int stop = (it + 8 < max_iter) ? it + 8 : max_iter;
                for (; it < stop && !r; ++it)
                {
                    if ((r = dot2(r0, res, nullptr, nullptr, sc + B_RHO, nullptr)))
                        break;
                    MFV2D_LAUNCH(bicg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 0, it == 0 ? 1 : 0);
                    MFV2D_LAUNCH(bicg_direction_kernel, gt, 256, 0, ctx->stream, nt, (const double *)sc, (const double *)res, (const double *)v, p,
                                 dinv, hat, (const int *)flags);
                    if (precondition >= 2 && (r = coarse_correction(s, p, hat, flags, use_peers))) // p^ = (D^-1 + sum_l P_l .. P_l^T) p
                        break;
                    if ((r = sys_apply_schur(s, hat, v, flags)))
                        break;
                    if ((r = dot2(r0, v, nullptr, nullptr, sc + B_R0V, nullptr)))
                        break;
                    MFV2D_LAUNCH(bicg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 1, 0);
                    // lambda += alpha p^ ; s = r - alpha v ; s^ = M^-1 s
                    MFV2D_LAUNCH(bicg_combine_kernel, gt, 256, 0, ctx->stream, nt, lam, (const double *)lam, (const double *)hat,
                                 (const double *)nullptr, (const double *)(sc + B_ALPHA), 1.0, (const double *)nullptr, 0.0,
                                 (const double *)nullptr, (double *)nullptr, (const int *)flags);
                    MFV2D_LAUNCH(bicg_combine_kernel, gt, 256, 0, ctx->stream, nt, sv, (const double *)res, (const double *)v,
                                 (const double *)nullptr, (const double *)(sc + B_ALPHA), -1.0, (const double *)nullptr, 0.0, dinv, hat,
                                 (const int *)flags);
                    if (precondition >= 2 && (r = coarse_correction(s, sv, hat, flags, use_peers)))
                        break;
                    if ((r = sys_apply_schur(s, hat, t, flags)))
                        break;
                    if ((r = dot2(t, sv, t, t, sc + B_TS, sc + B_TT)))
                        break;
                    MFV2D_LAUNCH(bicg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 2, 0);
                    // lambda += omega s^ ; r = s - omega t ; rr = r . r
                    MFV2D_LAUNCH(bicg_combine_kernel, gt, 256, 0, ctx->stream, nt, lam, (const double *)lam, (const double *)hat,
                                 (const double *)nullptr, (const double *)(sc + B_OMEGA), 1.0, (const double *)nullptr, 0.0,
                                 (const double *)nullptr, (double *)nullptr, (const int *)flags);
                    MFV2D_LAUNCH(bicg_combine_kernel, gt, 256, 0, ctx->stream, nt, res, (const double *)sv, (const double *)t,
                                 (const double *)nullptr, (const double *)(sc + B_OMEGA), -1.0, (const double *)nullptr, 0.0,
                                 (const double *)nullptr, (double *)nullptr, (const int *)flags);
                    if ((r = dot2(res, res, nullptr, nullptr, sc + B_RR, nullptr)))
                        break;
                    MFV2D_LAUNCH(bicg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 3, 0);
                    ctx->launches += 9;
                }
                if (r)
                    break;
                if (cudaGetLastError() != cudaSuccess)
                {
                    r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "BiCGStab kernel launch failed");
                    break;
                }
                if ((r = read_scalars(ctx, sc, 20, h)))
                    break;
                int hf[4];
                memcpy(hf, &h[B_FLAGS], sizeof hf);
                res_mag2 = h[B_RR];
                it_done = hf[1];
                finished = hf[0] != 0;
                if (hf[2])
                {
                    r = mfv2d_fail(ctx, MFV2D_UNSPECIFIED_ERROR, "Schur BiCGStab broke down (rho or omega vanished)");
                    break;
                }
            }
            if (r)
                break;
        }
        if ((r = sys_scale_rows(s, lam, false)))
            break;
        if (nt && cudaMemcpyAsync(slam.dev, lam, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess)
        {
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve_schur_bicgstab: copy of lambda failed");
            break;
        }
        // x = A^-1 y - A^-1 N^T lambda, then one step of iterative refinement with the stored A (see schur_cg.cu)
        if ((r = sys_spmv(s, nd, s->d_t_ptr, s->d_t_col, s->d_t_val, lam, tmp_d)))
            break;
        if ((r = sys_gemv(s, s->d_Ainv, tmp_d, tmp_d2)))
            break;
        if (nd)
        {
            const unsigned gd = grid_for(ctx, nd, 256);
            MFV2D_LAUNCH(axpby_kernel, gd, 256, 0, ctx->stream, nd, 1.0, (const double *)inv_a_y, -1.0, (const double *)tmp_d2, (double *)sx.dev);
            if ((r = sys_gemv(s, s->d_A, (const double *)sx.dev, tmp_d2)))
                break;
            MFV2D_LAUNCH(axpby_kernel, gd, 256, 0, ctx->stream, nd, 1.0, (const double *)srhs.dev, -1.0, (const double *)tmp_d2, tmp_d2);
            MFV2D_LAUNCH(axpby_kernel, gd, 256, 0, ctx->stream, nd, 1.0, (const double *)tmp_d2, -1.0, (const double *)tmp_d, tmp_d2);
            if ((r = sys_gemv(s, s->d_Ainv, tmp_d2, inv_a_y)))
                break;
            MFV2D_LAUNCH(axpby_kernel, gd, 256, 0, ctx->stream, nd, 1.0, (const double *)sx.dev, 1.0, (const double *)inv_a_y, (double *)sx.dev);
            ctx->launches += 4;
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
