// Two-basis ("projector") mass matrices and element projectors
// (compute_mass_matrix_{node,edge,surf}_double, src/fem_space/fem_space.c:847-1055;
//  compute_element_projector, src/evaluation/element_system.c:480-643).
//
//   M_mixed[w][v] = sum_ab  phi_out_w(a,b) . K(a,b) . phi_in_v(a,b)      rows: output basis, columns: input basis
//   dual != 0: out = M_mixed          dual == 0: out = M_out^-1 M_mixed   (M_out^-1 from the assembly kernel)
//
// Used by the VMS / refinement helpers only (a handful of elements per call), so the kernel is the plain
// quadrature sum - one CTA per element, the metric weights in shared memory - not a sum-factorised one.
#include "internal.cuh"

int launch_mass_inverse_device(const mfv2d_elements *e, int form, double *d_out, int32_t *d_status); // element_api.cu

#define PROJ_THREADS 128

// value of basis function `idx` of component `comp` (0: 0-form, 1: 1-form h, 2: 1-form v, 3: 2-form) of order p at
// the point (a: eta, b: xi); tab = node (p+1) x q | edge p x q  (src/fem_space/fem_space.c:121-217)
__device__ __forceinline__ double basis_value(const double *__restrict__ tab, int p, int q, int comp, int idx, int a, int b)
{
    const double *node = tab, *edge = tab + (p + 1) * q;
    switch (comp)
    {
    case 0: {
        const int i = idx / (p + 1), j = idx - i * (p + 1);
        return node[j * q + b] * node[i * q + a];
    }
    case 1: {
        const int i = idx / p, j = idx - i * p;
        return edge[j * q + b] * node[i * q + a];
    }
    case 2: {
        const int i = idx / (p + 1), j = idx - i * (p + 1);
        return node[j * q + b] * edge[i * q + a];
    }
    default: {
        const int i = idx / p, j = idx - i * p;
        return edge[j * q + b] * edge[i * q + a];
    }
    }
}

__global__ void __launch_bounds__(PROJ_THREADS) mass_double_kernel(const double *__restrict__ tab_out, int p_out,
                                                                   const double *__restrict__ tab_in, int p_in, int q,
                                                                   const double *__restrict__ corners, unsigned long long n_elem,
                                                                   int form, double *__restrict__ out)
{
    MFV2D_DYN_SMEM(double, K); // up to three weight fields of q*q doubles: (hh | vv | hv) or one scalar field
    const int q2 = q * q;
    const double *pnts = tab_out + (2 * p_out + 1) * q, *wgts = pnts + q;
    const int nh_o = p_out * (p_out + 1), nh_i = p_in * (p_in + 1);
    const int n_out = form == MFV2D_FORM_ORDER_0 ? (p_out + 1) * (p_out + 1) : (form == MFV2D_FORM_ORDER_1 ? 2 * nh_o : p_out * p_out);
    const int n_in = form == MFV2D_FORM_ORDER_0 ? (p_in + 1) * (p_in + 1) : (form == MFV2D_FORM_ORDER_1 ? 2 * nh_i : p_in * p_in);
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const double *cs = corners + e * 8;
        __syncthreads();
        for (int pt = threadIdx.x; pt < q2; pt += blockDim.x)
        {
            const int a = pt / q, b = pt - a * q;
            const double eta = pnts[a], xi = pnts[b], w = wgts[a] * wgts[b];
            const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
            const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
            const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
            const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
            const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
            const double det = j00 * j11 - j10 * j01;
            if (form == MFV2D_FORM_ORDER_0)
                K[pt] = det * w;
            else if (form == MFV2D_FORM_ORDER_2)
                K[pt] = w / det;
            else
            {
                K[pt] = (j11 * j11 + j10 * j10) / det * w;          // hh
                K[q2 + pt] = (j00 * j00 + j01 * j01) / det * w;     // vv
                K[2 * q2 + pt] = (j01 * j11 + j00 * j10) / det * w; // hv = vh
            }
        }
        __syncthreads();
        double *o = out + e * (unsigned long long)n_out * n_in;
        for (int idx = threadIdx.x; idx < n_out * n_in; idx += blockDim.x)
        {
            const int wr = idx / n_in, vc = idx - wr * n_in;
            int comp_o, comp_i, io = wr, ii = vc;
            const double *Kf = K;
            if (form == MFV2D_FORM_ORDER_0)
                comp_o = comp_i = 0;
            else if (form == MFV2D_FORM_ORDER_2)
                comp_o = comp_i = 3;
            else
            {
                comp_o = wr < nh_o ? 1 : 2;
                comp_i = vc < nh_i ? 1 : 2;
                io = wr < nh_o ? wr : wr - nh_o;
                ii = vc < nh_i ? vc : vc - nh_i;
                Kf = K + (comp_o != comp_i ? 2 * q2 : (comp_o == 1 ? 0 : q2));
            }
            double acc = 0.0;
            for (int a = 0; a < q; ++a)
                for (int b = 0; b < q; ++b)
                    acc = fma(basis_value(tab_out, p_out, q, comp_o, io, a, b) * basis_value(tab_in, p_in, q, comp_i, ii, a, b),
                              Kf[a * q + b], acc);
            o[idx] = acc;
        }
    }
}

// C[b] = A[b] (n x k) * B[b] (k x m), row-major, one CTA per matrix
__global__ void __launch_bounds__(PROJ_THREADS) batched_gemm_kernel(int n, int k, int m, unsigned long long batch,
                                                                    const double *__restrict__ A, const double *__restrict__ B,
                                                                    double *__restrict__ C)
{
    for (unsigned long long e = blockIdx.x; e < batch; e += gridDim.x)
    {
        const double *a = A + e * (unsigned long long)n * k, *b = B + e * (unsigned long long)k * m;
        double *c = C + e * (unsigned long long)n * m;
        for (int idx = threadIdx.x; idx < n * m; idx += blockDim.x)
        {
            const int r = idx / m, col = idx - r * m;
            double acc = 0.0;
            for (int t = 0; t < k; ++t)
                acc = fma(a[r * k + t], b[t * m + col], acc);
            c[idx] = acc;
        }
    }
}

extern "C" {

int mfv2d_cuda_projector_batch(const mfv2d_elements *e, const mfv2d_basis *basis_in, int form, int dual, double *out,
                               int32_t *status, int space)
{
    mfv2d_ctx *ctx = e->ctx;
    if (!basis_in || basis_in->ctx != ctx)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "projector_batch: input basis does not belong to this context");
    if (form < MFV2D_FORM_ORDER_0 || form > MFV2D_FORM_ORDER_2)
        return mfv2d_fail(ctx, MFV2D_BAD_ENUM, "projector_batch: invalid form order %d", form);
    const mfv2d_basis *bo = e->basis;
    if (basis_in->q != bo->q) // element_system.c:514-523
        return mfv2d_fail(ctx, MFV2D_ORDER_MISMATCH, "projector_batch: input and output bases must share the integration rule (%d vs %d points)",
                          basis_in->q, bo->q);
    if (e->n == 0)
        return MFV2D_SUCCESS;
    const int q = bo->q;
    const size_t n_out = (size_t)form_size(form, bo->p), n_in = (size_t)form_size(form, basis_in->p);
    Staged o, st;
    MFV2D_TRY(stage_out_alloc(ctx, out, e->n * n_out * n_in * sizeof(double), space, &o));
    int r = stage_out_alloc(ctx, status, status ? e->n * sizeof(int32_t) : 0, space, &st);
    void *mixed = nullptr, *minv = nullptr;
    if (!r && !dual)
    {
        r = mfv2d_cuda_malloc(ctx, e->n * n_out * n_in * sizeof(double), &mixed);
        if (!r)
            r = mfv2d_cuda_malloc(ctx, e->n * n_out * n_out * sizeof(double), &minv);
    }
    if (!r)
    {
        size_t grid = e->n < (size_t)ctx->sm_count * 8 ? e->n : (size_t)ctx->sm_count * 8;
        MFV2D_LAUNCH(mass_double_kernel, (unsigned)grid, PROJ_THREADS, (size_t)3 * q * q * sizeof(double), ctx->stream, bo->d_tab, bo->p,
                     basis_in->d_tab, basis_in->p, q, e->d_corners, (unsigned long long)e->n, form,
                     dual ? (double *)o.dev : (double *)mixed);
        ctx->launches++;
        if (!dual)
        {
            r = launch_mass_inverse_device(e, form, (double *)minv, (int32_t *)st.dev);
            if (!r)
            {
                MFV2D_LAUNCH(batched_gemm_kernel, (unsigned)grid, PROJ_THREADS, 0, ctx->stream, (int)n_out, (int)n_out, (int)n_in,
                             (unsigned long long)e->n, (const double *)minv, (const double *)mixed, (double *)o.dev);
                ctx->launches++;
            }
        }
        else if (st.dev)
            cudaMemsetAsync(st.dev, 0, e->n * sizeof(int32_t), ctx->stream);
        if (!r && cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "projector kernels failed to launch");
    }
    if (!r)
    {
        r = stage_finish(&o);
        const int r2 = stage_finish(&st);
        if (!r)
            r = r2;
    }
    else
    {
        stage_abort(&o);
        stage_abort(&st);
    }
    if (!r && (mixed || minv))
        r = mfv2d_cuda_synchronize(ctx); // the temporaries are freed below
    if (mixed)
        mfv2d_cuda_free(ctx, mixed);
    if (minv)
        mfv2d_cuda_free(ctx, minv);
    return r;
}

} // extern "C"
