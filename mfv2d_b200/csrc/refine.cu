// Batched kernels of the error-estimation helpers (python/mfv2d/refinement.py).
//
// The reference's estimators are per-element Python loops over primitives this library already has batched (projectors,
// element matrices and vectors, local solves, reconstruction); what they add is
//   * compute_legendre_coefficients / compute_legendre_error_estimates (refinement.py:40-152): the Legendre spectrum of
//     the reconstructed solution and of the reconstructed error on an element -> (|err|^2_L2, h-refinement cost);
//   * small dense glue: per-element matrix-vector products with the projectors (:1158-1168, :880-886) and the bordered
//     local systems that pin boundary DoFs / the mean of a form with Lagrange multipliers (:1001-1050).
// Here each of them is one launch over all elements of a batch.
#include "internal.cuh"

#define REFINE_THREADS 128
#define REFINE_MAX_ORDER 15
#define REFINE_MAX_Q 20

// ---- y[e] = A[e] x[e] for n row-major (rows x cols) matrices: one warp per row --------------------------------------
__global__ void __launch_bounds__(REFINE_THREADS) batched_gemv_kernel(unsigned long long n_rows_total, int rows, int cols,
                                                                     const double *__restrict__ A, const double *__restrict__ x,
                                                                     double *__restrict__ y)
{
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    for (unsigned long long row = (unsigned long long)blockIdx.x * nwarps + warp; row < n_rows_total; row += (unsigned long long)gridDim.x * nwarps)
    {
        const unsigned long long e = row / rows;
        const double *a = A + row * cols, *xe = x + e * cols;
        double acc = 0.0;
        for (int c = MFV2D_LANE0; c < cols; c += MFV2D_LANE_STEP)
            acc = fma(a[c], xe[c], acc);
        acc = warp_sum(acc);
        if (MFV2D_LANE0 == 0)
            y[row] = acc;
    }
}

// ---- C[e] = op(A[e]) B[e] for n row-major matrices: a 4 x 4 register tile per thread, operands through L1 / L2 ---------
// op(A) is (m x k): A itself, or the transpose of a stored (k x m) matrix.  Set-up work of the VMS operator (coarse blocks
// P^T A_f P), not a hot kernel: two loads per four FMAs, no shared-memory staging.
__global__ void __launch_bounds__(REFINE_THREADS) batched_gemm_kernel(unsigned long long n, int m, int k, int l, int trans_a,
                                                                     const double *__restrict__ A, const double *__restrict__ B,
                                                                     double *__restrict__ Cm)
{
    const int tm = (m + 3) / 4, tl = (l + 3) / 4;
    const unsigned long long per = (unsigned long long)tm * tl, total = n * per;
    for (unsigned long long g = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; g < total; g += (unsigned long long)gridDim.x * blockDim.x)
    {
        const unsigned long long e = g / per;
        const int t = (int)(g - e * per);
        const int i0 = (t / tl) * 4, j0 = (t - (t / tl) * tl) * 4;
        const double *a = A + e * (unsigned long long)m * k, *b = B + e * (unsigned long long)k * l;
        double acc[4][4] = {};
        for (int kk = 0; kk < k; ++kk)
        {
            double av[4], bv[4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
            {
                const int i = i0 + r < m ? i0 + r : m - 1, j = j0 + r < l ? j0 + r : l - 1;
                av[r] = trans_a ? a[(size_t)kk * m + i] : a[(size_t)i * k + kk];
                bv[r] = b[(size_t)kk * l + j];
            }
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c)
                    acc[r][c] = fma(av[r], bv[c], acc[r][c]);
        }
        double *cm = Cm + e * (unsigned long long)m * l;
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c)
                if (i0 + r < m && j0 + c < l)
                    cm[(size_t)(i0 + r) * l + j0 + c] = acc[r][c];
    }
}

// ---- out[e] = [[A[e], L^T], [L, 0]] : the same `pad` Lagrange rows (COO, entries lval) for every element --------------
__global__ void border_batch_kernel(unsigned long long n_elem, int N, int pad, const double *__restrict__ A, int nnz,
                                    const int32_t *__restrict__ lrow, const int32_t *__restrict__ lcol, const double *__restrict__ lval,
                                    double *__restrict__ out)
{
    const int M = N + pad;
    const unsigned long long per = (unsigned long long)M * M, total = n_elem * per;
    for (unsigned long long g = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; g < total; g += (unsigned long long)gridDim.x * blockDim.x)
    {
        const unsigned long long e = g / per;
        const int idx = (int)(g - e * per);
        const int i = idx / M, j = idx - i * M;
        double v = 0.0;
        if (i < N && j < N)
            v = A[e * (unsigned long long)N * N + (unsigned long long)i * N + j];
        else if (i >= N && j >= N)
            v = 0.0;
        else
        {
            // (N + r, c) and its mirror image (c, N + r)
            const int r = (i >= N ? i : j) - N, c = i >= N ? j : i;
            for (int k = 0; k < nnz; ++k)
                if (lrow[k] == r && lcol[k] == c)
                    v += lval[k];
        }
        out[g] = v;
    }
}

// ---- Legendre error estimates of one element per CTA pass ----------------------------------------------------------------
// u, err: values at the quadrature points [n][q(eta)][q(xi)] (comps == 1) or [n][q][q][2] (1-forms: the Euclidean norm is
// taken, refinement.py:138-140).  out[e] = { sum err^2 w det , | sum of the measure outside the low-order block | }.
struct EstimateArgs
{
    const double *corners, *pnts, *wgts, *u, *err;
    double *out;
    unsigned long long n_elem;
    int q, o1, o2, comps;
};

__global__ void __launch_bounds__(REFINE_THREADS) legendre_estimate_kernel(const EstimateArgs a)
{
    MFV2D_DYN_SMEM(double, smem);
    const int q = a.q, q2 = q * q, n1 = a.o1 + 1, n2 = a.o2 + 1;
    double *leg1 = smem;               // [n1][q] over xi
    double *leg2 = leg1 + n1 * q;      // [n2][q] over eta
    double *wu = leg2 + n2 * q;        // [q2] u w det
    double *we = wu + q2;              // [q2] err w det
    double *rs = we + q2;              // [q2] sqrt(det)
    double *l2 = rs + q2;              // [q2] err^2 w det
    double *meas = l2 + q2;            // [n2][n1]
    for (int i = threadIdx.x; i < q; i += blockDim.x)
    {
        // Bonnet recurrence, the operation order of compute_legendre (src/polynomials/legendre.c / tables.cu legendre_kernel)
        const double x = a.pnts[i];
        double p0 = 1.0, p1 = x;
        leg1[i] = leg2[i] = 1.0;
        if (n1 > 1)
            leg1[q + i] = x;
        if (n2 > 1)
            leg2[q + i] = x;
        const int top = n1 > n2 ? n1 : n2;
        for (int k = 2; k < top; ++k)
        {
            const double c = ((2 * k - 1) * x * p1 - (k - 1) * p0) / (double)k;
            if (k < n1)
                leg1[k * q + i] = c;
            if (k < n2)
                leg2[k * q + i] = c;
            p0 = p1;
            p1 = c;
        }
    }
    __syncthreads();
    for (unsigned long long e = blockIdx.x; e < a.n_elem; e += gridDim.x)
    {
        const double *c = a.corners + e * 8;
        const double x0 = c[0], y0 = c[1], x1 = c[2], y1 = c[3], x2 = c[4], y2 = c[5], x3 = c[6], y3 = c[7];
        for (int idx = threadIdx.x; idx < q2; idx += blockDim.x)
        {
            const int ia = idx / q, ib = idx - ia * q;
            const double eta = a.pnts[ia], xi = a.pnts[ib];
            const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
            const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
            const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
            const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
            const double det = j00 * j11 - j10 * j01;
            const double w = a.wgts[ib] * a.wgts[ia];
            double uv, ev;
            if (a.comps == 2)
            {
                const double *pu = a.u + (e * q2 + idx) * 2, *pe = a.err + (e * q2 + idx) * 2;
                uv = sqrt(pu[0] * pu[0] + pu[1] * pu[1]);
                ev = sqrt(pe[0] * pe[0] + pe[1] * pe[1]);
            }
            else
            {
                uv = a.u[e * q2 + idx];
                ev = a.err[e * q2 + idx];
            }
            wu[idx] = uv * w * det;
            we[idx] = ev * w * det;
            rs[idx] = sqrt(det);
            l2[idx] = ev * ev * w * det;
        }
        __syncthreads();
        for (int k = threadIdx.x; k < n1 * n2; k += blockDim.x)
        {
            const int j2 = k / n1, j1 = k - j2 * n1;
            double cu = 0.0, ce = 0.0;
            for (int ia = 0; ia < q; ++ia)
            {
                const double lb = leg2[j2 * q + ia];
                for (int ib = 0; ib < q; ++ib)
                {
                    const double f = leg1[j1 * q + ib] * lb / rs[ia * q + ib];
                    cu = fma(wu[ia * q + ib], f, cu);
                    ce = fma(we[ia * q + ib], f, ce);
                }
            }
            // coefficients: divide by the norms 2 / (2 j + 1) of both directions; measure: times 4 / ((2 j1 + 1)(2 j2 + 1))
            const double nrm = (2.0 / (2 * j1 + 1)) * (2.0 / (2 * j2 + 1));
            cu /= nrm;
            ce /= nrm;
            const bool low = j2 < a.o2 / 2 && j1 < a.o1 / 2;
            meas[k] = low ? 0.0 : cu * (cu + 2 * ce) * (4.0 / ((2 * j1 + 1) * (double)(2 * j2 + 1)));
        }
        __syncthreads();
        if (threadIdx.x == 0)
        {
            double s = 0.0, m = 0.0;
            for (int idx = 0; idx < q2; ++idx)
                s += l2[idx];
            for (int k = 0; k < n1 * n2; ++k)
                m += meas[k];
            a.out[2 * e] = s;
            a.out[2 * e + 1] = fabs(m);
        }
        __syncthreads();
    }
}

extern "C" {

int mfv2d_cuda_batched_gemv(mfv2d_ctx *ctx, size_t n, int rows, int cols, const double *mats, const double *x, double *y, int space)
{
    if (rows < 1 || cols < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "batched_gemv: rows and cols must be positive");
    if (n == 0)
        return MFV2D_SUCCESS;
    Staged sa, sx, sy;
    int r = stage_in(ctx, mats, n * (size_t)rows * cols * sizeof(double), space, &sa);
    if (!r)
        r = stage_in(ctx, x, n * (size_t)cols * sizeof(double), space, &sx);
    if (!r)
        r = stage_out_alloc(ctx, y, n * (size_t)rows * sizeof(double), space, &sy);
    if (!r)
    {
        const unsigned long long total = (unsigned long long)n * rows;
        const int warps = REFINE_THREADS / 32;
        unsigned long long grid = (total + warps - 1) / warps;
        if (grid > (unsigned long long)ctx->sm_count * 16)
            grid = (unsigned long long)ctx->sm_count * 16;
        MFV2D_LAUNCH(batched_gemv_kernel, (unsigned)grid, REFINE_THREADS, 0, ctx->stream, total, rows, cols, (const double *)sa.dev,
                     (const double *)sx.dev, (double *)sy.dev);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "batched_gemv kernel launch failed");
    }
    if (!r)
        r = stage_finish(&sy);
    else
        stage_abort(&sy);
    stage_abort(&sa);
    stage_abort(&sx);
    return r;
}

int mfv2d_cuda_batched_gemm(mfv2d_ctx *ctx, size_t n, int m, int k, int l, int trans_a, const double *a, const double *b, double *c, int space)
{
    if (m < 1 || k < 1 || l < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "batched_gemm: m, k and l must be positive");
    if (n == 0)
        return MFV2D_SUCCESS;
    Staged sa, sb, sc;
    int r = stage_in(ctx, a, n * (size_t)m * k * sizeof(double), space, &sa);
    if (!r)
        r = stage_in(ctx, b, n * (size_t)k * l * sizeof(double), space, &sb);
    if (!r)
        r = stage_out_alloc(ctx, c, n * (size_t)m * l * sizeof(double), space, &sc);
    if (!r)
    {
        const unsigned long long total = (unsigned long long)n * ((m + 3) / 4) * ((l + 3) / 4);
        unsigned long long grid = (total + REFINE_THREADS - 1) / REFINE_THREADS;
        if (grid > (unsigned long long)ctx->sm_count * 16)
            grid = (unsigned long long)ctx->sm_count * 16;
        MFV2D_LAUNCH(batched_gemm_kernel, (unsigned)grid, REFINE_THREADS, 0, ctx->stream, (unsigned long long)n, m, k, l, trans_a ? 1 : 0,
                     (const double *)sa.dev, (const double *)sb.dev, (double *)sc.dev);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "batched_gemm kernel launch failed");
    }
    if (!r)
        r = stage_finish(&sc);
    else
        stage_abort(&sc);
    stage_abort(&sa);
    stage_abort(&sb);
    return r;
}

int mfv2d_cuda_border_batch(mfv2d_ctx *ctx, size_t n, int size, int pad, const double *mats, int nnz, const int32_t *lrow,
                            const int32_t *lcol, const double *lval, double *out, int space)
{
    if (size < 1 || pad < 0 || nnz < 0)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "border_batch: bad sizes");
    for (int k = 0; k < nnz; ++k)
        if (lrow[k] < 0 || lrow[k] >= pad || lcol[k] < 0 || lcol[k] >= size)
            return mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "border_batch: entry %d (%d, %d) outside the %d x %d border", k, lrow[k],
                              lcol[k], pad, size);
    if (n == 0)
        return MFV2D_SUCCESS;
    const int M = size + pad;
    const int32_t none_i = 0;
    const double none_d = 0.0;
    if (nnz == 0)
    {
        lrow = lcol = &none_i;
        lval = &none_d;
    }
    Staged sa, so, s1, s2, s3;
    int r = stage_in(ctx, mats, n * (size_t)size * size * sizeof(double), space, &sa);
    if (!r)
        r = stage_in(ctx, lrow, (size_t)(nnz ? nnz : 1) * sizeof(int32_t), MFV2D_HOST, &s1);
    if (!r)
        r = stage_in(ctx, lcol, (size_t)(nnz ? nnz : 1) * sizeof(int32_t), MFV2D_HOST, &s2);
    if (!r)
        r = stage_in(ctx, lval, (size_t)(nnz ? nnz : 1) * sizeof(double), MFV2D_HOST, &s3);
    if (!r)
        r = stage_out_alloc(ctx, out, n * (size_t)M * M * sizeof(double), space, &so);
    if (!r)
    {
        unsigned long long grid = ((unsigned long long)n * M * M + 255) / 256;
        if (grid > (unsigned long long)ctx->sm_count * 16)
            grid = (unsigned long long)ctx->sm_count * 16;
        MFV2D_LAUNCH(border_batch_kernel, (unsigned)grid, 256, 0, ctx->stream, (unsigned long long)n, size, pad, (const double *)sa.dev, nnz,
                     (const int32_t *)s1.dev, (const int32_t *)s2.dev, (const double *)s3.dev, (double *)so.dev);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "border_batch kernel launch failed");
    }
    if (!r)
        r = stage_finish(&so);
    else
        stage_abort(&so);
    stage_abort(&sa);
    stage_abort(&s1);
    stage_abort(&s2);
    stage_abort(&s3);
    return r;
}

int mfv2d_cuda_legendre_estimates_batch(const mfv2d_elements *elems, int order_1, int order_2, int form, const double *u,
                                        const double *err, double *out, int space)
{
    mfv2d_ctx *ctx = elems->ctx;
    const int q = elems->basis->q;
    if (order_1 < 0 || order_2 < 0 || order_1 > REFINE_MAX_ORDER || order_2 > REFINE_MAX_ORDER)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "legendre_estimates: orders must lie in [0, %d]", REFINE_MAX_ORDER);
    if (form < MFV2D_FORM_ORDER_0 || form > MFV2D_FORM_ORDER_2)
        return mfv2d_fail(ctx, MFV2D_BAD_ENUM, "legendre_estimates: invalid form order %d", form);
    if (elems->n == 0)
        return MFV2D_SUCCESS;
    const int comps = form == MFV2D_FORM_ORDER_1 ? 2 : 1;
    const size_t vals = elems->n * (size_t)q * q * comps * sizeof(double);
    Staged su, se, so;
    int r = stage_in(ctx, u, vals, space, &su);
    if (!r)
        r = stage_in(ctx, err, vals, space, &se);
    if (!r)
        r = stage_out_alloc(ctx, out, elems->n * 2 * sizeof(double), space, &so);
    if (!r)
    {
        EstimateArgs a;
        a.corners = elems->d_corners;
        a.pnts = elems->basis->d_tab + (size_t)(2 * elems->basis->p + 1) * q;
        a.wgts = a.pnts + q;
        a.u = (const double *)su.dev;
        a.err = (const double *)se.dev;
        a.out = (double *)so.dev;
        a.n_elem = elems->n;
        a.q = q;
        a.o1 = order_1;
        a.o2 = order_2;
        a.comps = comps;
        const size_t smem = ((size_t)(order_1 + order_2 + 2) * q + 4 * (size_t)q * q + (size_t)(order_1 + 1) * (order_2 + 1)) * sizeof(double);
        if (smem > ctx->smem_optin)
            r = mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "legendre_estimates: %zu bytes of shared memory needed", smem);
        else
        {
            if (smem > 48 * 1024)
                cudaFuncSetAttribute(legendre_estimate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            size_t grid = (size_t)ctx->sm_count * 8;
            if (grid > elems->n)
                grid = elems->n;
            MFV2D_LAUNCH(legendre_estimate_kernel, (unsigned)grid, REFINE_THREADS, smem, ctx->stream, a);
            ctx->launches++;
            if (cudaGetLastError() != cudaSuccess)
                r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "legendre_estimates kernel launch failed");
        }
    }
    if (!r)
        r = stage_finish(&so);
    else
        stage_abort(&so);
    stage_abort(&su);
    stage_abort(&se);
    return r;
}

} // extern "C"
