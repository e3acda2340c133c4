// solve_pcg_iterative (python/mfv2d/solving.py:634-684) with pcg_general (:280-347): preconditioned CG on the saddle-point
// system K = [[A, N^T], [N, 0]] with the block-Jacobi preconditioner A^-1 (+) I.  Full vectors are stored as
// [dense | trace]; every scalar of the recurrence (r.z, p.Kp, alpha, beta, |r|^2), the reference's degeneracy test
// log|p.Kp| - log|r|^2 < log 1e-12 (:331-332) and the stopping flag live on the DEVICE - the host reads them back once per
// block of iterations - and the driver works on partitioned systems: the dense part is rank-local, the shared
// multipliers of the trace part are summed over the interface (NVLink peer exchange or NCCL) and counted once in the dot
// products through the ownership weights.
#include <cmath>

#include "system.cuh"

// sc: [0] |r|^2  [1] r.z  [2] p.Kp  [3] alpha  [4] beta  [5] tol^2  [6] new r.z ; flags at sc + 16: [0] done [1] iter_cnt [2] degenerate
enum
{
    P_RR = 0,
    P_RZ = 1,
    P_APA = 2,
    P_ALPHA = 3,
    P_BETA = 4,
    P_TOL2 = 5,
    P_NRZ = 6,
    P_FLAGS = 16
};

// partial sums of x . y over [dense | trace]; trace entries weighted by the ownership mask
__global__ void __launch_bounds__(SYS_THREADS) pcg_dot_kernel(long long nd, long long nf, const double *__restrict__ x, const double *__restrict__ y,
                                                             const double *__restrict__ own, double *__restrict__ partials,
                                                             const int *__restrict__ done)
{
    if (*done)
        return;
    double a = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nf; i += (long long)gridDim.x * blockDim.x)
    {
        const double w = (own && i >= nd) ? own[i - nd] : 1.0;
        a = fma(x[i] * w, y[i], a);
    }
    block_reduce_store(a, partials);
}

__global__ void pcg_sum_kernel(int np, const double *__restrict__ partials, double *__restrict__ out, const int *__restrict__ done)
{
    if (*done)
        return;
    const double a = warp_sum_partials(np, partials);
    if (threadIdx.x == 0)
        *out = a;
}

// step 0: alpha = r.z / p.Kp with the degeneracy test; 1: convergence test after the update; 2: beta = new r.z / r.z
__global__ void pcg_scalar_kernel(double *__restrict__ sc, int step)
{
    int *flags = reinterpret_cast<int *>(sc + P_FLAGS);
    if (threadIdx.x != 0 || flags[0])
        return;
    if (step == 0)
    {
        if (log(fabs(sc[P_APA])) - log(sc[P_RR]) < log(1e-12)) // solving.py:331-332
        {
            flags[0] = flags[2] = 1;
            return;
        }
        sc[P_ALPHA] = sc[P_RZ] / sc[P_APA];
    }
    else if (step == 1)
    {
        if (sc[P_RR] < sc[P_TOL2])
            flags[0] = 1; // iter_cnt stays at the index of this iteration
        // otherwise the iteration goes on to the preconditioner
    }
    else
    {
        sc[P_BETA] = sc[P_NRZ] / sc[P_RZ];
        sc[P_RZ] = sc[P_NRZ];
        flags[1] += 1;
    }
}

// out = x + (s * *c) y.  Product and sum are rounded SEPARATELY, as in the reference's x86 build of the vector updates
// (system_objects.c: no fused multiply-add without -mfma): this recurrence on the indefinite saddle-point system amplifies
// rounding differences by about 100x every ten iterations, and with the same arithmetic as the reference it follows the
// reference's trajectory for longer.
__global__ void pcg_axpy_kernel(long long n, double *out, const double *x, const double *y, const double *c, double s, const int *__restrict__ done)
{
    if (*done)
        return;
    const double a = s * c[0];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = __dadd_rn(x[i], __dmul_rn(a, y[i]));
}

extern "C" int mfv2d_cuda_solve_pcg(mfv2d_system *s, const double *rhs, const double *phi, int max_iter, double atol, double rtol, double *x,
                                    double *lambda, double *residual, int *iterations, int space)
{
    mfv2d_ctx *ctx = s->ctx;
    const long long nd = s->n_dense, nt = s->n_trace, nf = nd + nt;
    if ((nd && (!rhs || !x)) || (nt && (!phi || !lambda)))
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_pcg: rhs, phi, x and lambda must not be NULL");
    MFV2D_TRY(sys_require_interface(s));
    Staged srhs, sphi, sx, slam;
    int r = stage_in(ctx, rhs, nd * sizeof(double), space, &srhs);
    if (!r)
        r = stage_in(ctx, phi, nt * sizeof(double), space, &sphi);
    if (!r)
        r = stage_out_alloc(ctx, x, nd * sizeof(double), space, &sx);
    if (!r)
        r = stage_out_alloc(ctx, lambda, nt * sizeof(double), space, &slam);
    double *buf = nullptr;
    if (!r)
        r = mfv2d_cuda_malloc(ctx, (size_t)(6 * nf + 2 * nd + 8) * sizeof(double), (void **)&buf);
    double res_mag2 = 0.0;
    int iter_cnt = 0;
    if (!r)
    {
        double *b = buf, *xs = b + nf, *res = xs + nf, *pv = res + nf, *z = pv + nf, *ap = z + nf, *t1 = ap + nf, *t2 = t1 + nd;
        double *sc = s->d_cg;
        int *flags = reinterpret_cast<int *>(sc + P_FLAGS);
        const double *own = s->d_owned;
        const int np = s->n_partials;
        const unsigned gf = grid_for(ctx, nf, 256);
        double h[24];
        auto apply_system = [&](const double *v, double *out) -> int { // out = K v (interface sum on the trace part)
            MFV2D_TRY(sys_gemv(s, s->d_A, v, t1));
            MFV2D_TRY(sys_spmv(s, nd, s->d_t_ptr, s->d_t_col, s->d_t_val, v + nd, t2));
            if (nd)
            {
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)t1, 1.0, (const double *)t2, out);
                ctx->launches++;
            }
            return sys_apply_trace(s, v, out + nd);
        };
        auto precondition = [&](const double *v, double *out) -> int { // out = (A^-1 (+) I) v
            MFV2D_TRY(sys_gemv(s, s->d_Ainv, v, out));
            if (nt)
                MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(out + nd, v + nd, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
            return MFV2D_SUCCESS;
        };
        auto dot = [&](const double *u, const double *v, double *out) -> int {
            MFV2D_LAUNCH(pcg_dot_kernel, (unsigned)np, SYS_THREADS, 0, ctx->stream, nd, nf, u, v, own, s->d_partials, (const int *)flags);
            MFV2D_LAUNCH(pcg_sum_kernel, 1, 32, 0, ctx->stream, np, (const double *)s->d_partials, out, (const int *)flags);
            ctx->launches += 2;
            if (ctx->n_ranks > 1)
            {
                if (sys_use_peers(s))
                    MFV2D_TRY(peer_interface_sum(s, nullptr, 1, out));
                else
                    MFV2D_TRY(comm_allreduce(ctx, out, 1));
            }
            return MFV2D_SUCCESS;
        };
        do
        {
            cudaMemsetAsync(sc, 0, 32 * sizeof(double), ctx->stream);
            cudaMemcpyAsync(b, srhs.dev, nd * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if (nt)
                cudaMemcpyAsync(b + nd, sphi.dev, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            cudaMemsetAsync(xs, 0, nf * sizeof(double), ctx->stream);
            cudaMemcpyAsync(res, b, nf * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream); // initial guess zero: res = rhs
            if ((r = precondition(res, z)))
                break;
            cudaMemcpyAsync(pv, z, nf * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if ((r = dot(b, b, sc + P_RR)))
                break;
            if ((r = dot(res, z, sc + P_RZ)))
                break;
            if ((r = read_scalars(ctx, sc, 2, h)))
                break;
            res_mag2 = h[P_RR];
            const double mag = sqrt(res_mag2);
            const double tol = (mag * rtol > atol) ? atol : mag * rtol;
            const double tol2 = tol * tol;
            cudaMemcpyAsync(sc + P_TOL2, &tol2, sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
            cudaStreamSynchronize(ctx->stream);
            int it = 0;
            bool finished = false;
            while (it < max_iter && !finished && !r)
            {
                const int stop = (it + 8 < max_iter) ? it + 8 : max_iter;
                for (; it < stop && !r; ++it)
                {
                    if ((r = apply_system(pv, ap)))
                        break;
                    if ((r = dot(ap, pv, sc + P_APA)))
                        break;
                    MFV2D_LAUNCH(pcg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 0);
                    MFV2D_LAUNCH(pcg_axpy_kernel, gf, 256, 0, ctx->stream, nf, xs, (const double *)xs, (const double *)pv, (const double *)(sc + P_ALPHA),
                                 1.0, (const int *)flags);
                    MFV2D_LAUNCH(pcg_axpy_kernel, gf, 256, 0, ctx->stream, nf, res, (const double *)res, (const double *)ap,
                                 (const double *)(sc + P_ALPHA), -1.0, (const int *)flags);
                    if ((r = dot(res, res, sc + P_RR)))
                        break;
                    MFV2D_LAUNCH(pcg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 1);
                    if ((r = precondition(res, z)))
                        break;
                    if ((r = dot(res, z, sc + P_NRZ)))
                        break;
                    MFV2D_LAUNCH(pcg_scalar_kernel, 1, 32, 0, ctx->stream, sc, 2);
                    // p = z + beta p
                    MFV2D_LAUNCH(pcg_axpy_kernel, gf, 256, 0, ctx->stream, nf, pv, (const double *)z, (const double *)pv, (const double *)(sc + P_BETA),
                                 1.0, (const int *)flags);
                    ctx->launches += 6;
                }
                if (r)
                    break;
                if (cudaGetLastError() != cudaSuccess)
                {
                    r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "PCG kernel launch failed");
                    break;
                }
                if ((r = read_scalars(ctx, sc, 20, h)))
                    break;
                int hf[4];
                memcpy(hf, &h[P_FLAGS], sizeof hf);
                res_mag2 = h[P_RR];
                iter_cnt = hf[1];
                finished = hf[0] != 0;
                if (hf[2])
                {
                    r = mfv2d_fail(ctx, MFV2D_UNSPECIFIED_ERROR, "System degenerated (matrix was probably not SPD).");
                    break;
                }
            }
            if (r)
                break;
            if (!finished && iter_cnt == max_iter && max_iter > 0)
                iter_cnt = max_iter - 1; // Python's loop variable after an exhausted range
            cudaMemcpyAsync(sx.dev, xs, nd * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if (nt)
                cudaMemcpyAsync(slam.dev, xs + nd, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if (cudaGetLastError() != cudaSuccess)
                r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "PCG kernel launch failed");
        } while (0);
    }
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
    if (buf)
    {
        mfv2d_cuda_synchronize(ctx);
        mfv2d_cuda_free(ctx, buf);
    }
    if (residual)
        *residual = sqrt(res_mag2);
    if (iterations)
        *iterations = iter_cnt;
    return r;
}
