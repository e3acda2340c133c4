// Batched element-matrix assembly: Jacobian, sum-factorised mass / interior-product matrices in
// shared memory, incidence application as signed gathers and coalesced row-major stores of the
// N_e x N_e element matrices.  One CTA works on one element at a time (persistent over the batch).
//
// Replaces, per element: fem_space_2d_create (src/fem_space/fem_space.c:16-57),
// compute_mass_matrix_{node,edge,surf} (:235-411), compute_mass_matrix_{node_edge,edge_surf}
// (:546-846), matrix_full_invert (src/algebra/matrix.c:361-411), apply_incidence_to_full_{left,right}
// (src/evaluation/incidence.c:16-479) and the block loop of compute_element_matrix
// (src/evaluation/element_system.c:135-181).
//
// Roofline: the kernel is bound by the HBM WRITE of the element matrices, 64 + 8 N_e^2 algorithmic
// bytes per element; the sum-factorised contraction needs O(p^4 q) flops per element.
#include "dense_device.cuh"

#define ASM_THREADS 256

struct AsmArgs
{
    const double *tab;
    const int32_t *inc;
    const double *corners;
    const char *blob;
    const double *fields[16]; // per field: [n_elem][q*q][2]
    double *out;
    int32_t *status;
    unsigned long long n_elem;
    int tab_len, p, q;
    int inc_off[4];
    int n_stages, off_dense, off_terms, off_blocks, off_units;
    int n_total;                   // leading dimension of the output rows
    unsigned long long out_stride; // doubles between consecutive elements of the output
    int arena_off; // offset of the dense arena inside shared memory (doubles)
};


__device__ __forceinline__ double gather_entry(const double *__restrict__ D, int ld, const int4 lm, const int4 rm)
{
    const int lv[4] = {lm.x, lm.y, lm.z, lm.w};
    const int rv[4] = {rm.x, rm.y, rm.z, rm.w};
    double s = 0.0;
#pragma unroll
    for (int a = 0; a < 4; ++a)
    {
        if (lv[a] == 0)
            break;
        const double *row = D + (abs(lv[a]) - 1) * ld;
#pragma unroll
        for (int b = 0; b < 4; ++b)
        {
            if (rv[b] == 0)
                break;
            const double v = row[abs(rv[b]) - 1];
            s += ((lv[a] ^ rv[b]) < 0) ? -v : v;
        }
    }
    return s;
}

__global__ void __launch_bounds__(ASM_THREADS) element_matrix_kernel(const AsmArgs args)
{
    MFV2D_DYN_SMEM(double, smem);
    const int p = args.p, q = args.q;
    const AsmLayout lay = asm_layout(p, q, args.tab_len);
    SmemView s;
    s.tab = smem;
    s.node = s.tab;
    s.edge = s.node + (p + 1) * q;
    s.pnts = s.edge + p * q;
    s.wgts = s.pnts + q;
    s.geo = smem + lay.geo;
    s.G = smem + lay.G;
    s.ldT = lay.ldT;
    s.T = smem + lay.T;
    s.arena = smem + lay.arena;
    __shared__ double cs[8];

    for (int i = threadIdx.x; i < args.tab_len; i += blockDim.x)
        s.tab[i] = args.tab[i];

    const DevStage *stages = reinterpret_cast<const DevStage *>(args.blob);
    const DevDense *dense = reinterpret_cast<const DevDense *>(args.blob + args.off_dense);
    const DevTerm *terms = reinterpret_cast<const DevTerm *>(args.blob + args.off_terms);
    const DevBlock *blocks = reinterpret_cast<const DevBlock *>(args.blob + args.off_blocks);
    const DevUnit *units = reinterpret_cast<const DevUnit *>(args.blob + args.off_units);
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    const int N = args.n_total;

    for (unsigned long long elem = blockIdx.x; elem < args.n_elem; elem += gridDim.x)
    {
        __syncthreads(); // the previous element's emission has finished reading shared memory
        for (int i = threadIdx.x; i < 8; i += blockDim.x)
            cs[i] = args.corners[elem * 8 + i];
        __syncthreads();
        fill_geometry(s.geo, cs, s.pnts, q);
        int bad = 0;
        double *const out = args.out + elem * args.out_stride;
        for (int st = 0; st < args.n_stages; ++st)
        {
            const DevStage sg = stages[st];
            __syncthreads(); // geometry ready / previous stage's emission done
            for (int d = sg.dense_begin; d < sg.dense_end; ++d)
                compute_dense(s, dense[d], p, q, args.fields, elem, &bad);
            __syncthreads();
            // emission: each warp takes units (8 rows x 32 columns of one block)
            for (int u = sg.unit_begin + warp; u < sg.unit_end; u += nwarps)
            {
                const DevUnit un = units[u];
                const DevBlock db = blocks[un.block];
                for (int lane = MFV2D_LANE0; lane < 32; lane += MFV2D_LANE_STEP)
                {
                const int c = un.c0 + lane;
                if (c >= db.ncols)
                    continue;
                double acc[8];
#pragma unroll
                for (int r = 0; r < 8; ++r)
                    acc[r] = 0.0;
                for (int t = db.term_begin; t < db.term_end; ++t)
                {
                    const DevTerm tm = terms[t];
                    if (tm.dense < 0)
                    {
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (r < un.nr && un.r0 + r == c)
                                acc[r] += tm.coef;
                        continue;
                    }
                    const DevDense dd = dense[tm.dense];
                    const double *D = s.arena + dd.smem_off;
                    const int ld = dd.cols;
                    int4 rm = make_int4(c + 1, 0, 0, 0);
                    if (tm.R >= 0) // column c of E_R = row c of its transpose
                        rm = ld_ro(reinterpret_cast<const int4 *>(args.inc) + args.inc_off[tm.R ^ 2] + c);
                    if (tm.L < 0)
                    {
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (r < un.nr)
                                acc[r] = fma(tm.coef, gather_entry(D, ld, make_int4(un.r0 + r + 1, 0, 0, 0), rm), acc[r]);
                    }
                    else
                    {
                        const int4 *lmap = reinterpret_cast<const int4 *>(args.inc) + args.inc_off[tm.L] + un.r0;
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            if (r < un.nr)
                                acc[r] = fma(tm.coef, gather_entry(D, ld, ld_ro(lmap + r), rm), acc[r]);
                    }
                }
                double *o = out + (unsigned long long)(db.row0 + un.r0) * N + db.col0 + c;
                if (db.accumulate)
                {
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < un.nr)
                            o[(unsigned long long)r * N] += acc[r];
                }
                else
                {
#pragma unroll
                    for (int r = 0; r < 8; ++r)
                        if (r < un.nr)
                            o[(unsigned long long)r * N] = acc[r];
                }
                } // lanes
            }
        }
        if (args.status && threadIdx.x == 0)
            args.status[elem] = bad ? MFV2D_PIVOT_FAILED : MFV2D_SUCCESS;
    }
}

// ---- geometry-only kernels ------------------------------------------------------------------
__global__ void jacobian_kernel(const double *__restrict__ corners, const double *__restrict__ pnts, int q,
                                unsigned long long n_elem, double *__restrict__ out, int coords)
{
    const int q2 = q * q;
    const unsigned long long total = n_elem * q2;
    for (unsigned long long g = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; g < total;
         g += (unsigned long long)gridDim.x * blockDim.x)
    {
        const unsigned long long e = g / q2;
        const int idx = (int)(g - e * q2);
        const int a = idx / q, b = idx - a * q;
        const double *c = corners + e * 8;
        const double eta = pnts[a], xi = pnts[b];
        const double x0 = c[0], y0 = c[1], x1 = c[2], y1 = c[3], x2 = c[4], y2 = c[5], x3 = c[6], y3 = c[7];
        if (coords)
        {
            // bilinear blend (integrating_fields.c:92-111)
            const double n0 = (1 - xi) / 2 * (1 - eta) / 2, n1 = (1 + xi) / 2 * (1 - eta) / 2,
                         n2 = (1 + xi) / 2 * (1 + eta) / 2, n3 = (1 - xi) / 2 * (1 + eta) / 2;
            double *o = out + e * 2 * q2;
            o[idx] = x0 * n0 + x1 * n1 + x2 * n2 + x3 * n3;
            o[q2 + idx] = y0 * n0 + y1 * n1 + y2 * n2 + y3 * n3;
        }
        else
        {
            const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
            const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
            const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
            const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
            double *o = out + e * 5 * q2;
            o[idx] = j00;
            o[q2 + idx] = j01;
            o[2 * q2 + idx] = j10;
            o[3 * q2 + idx] = j11;
            o[4 * q2 + idx] = j00 * j11 - j10 * j01;
        }
    }
}

static int geometry_batch(const mfv2d_elements *e, double *out, int space, int coords)
{
    mfv2d_ctx *ctx = e->ctx;
    const int q = e->basis->q;
    const size_t per = (size_t)(coords ? 2 : 5) * q * q;
    if (e->n == 0)
        return MFV2D_SUCCESS;
    Staged o;
    MFV2D_TRY(stage_out_alloc(ctx, out, e->n * per * sizeof(double), space, &o));
    const double *pnts = e->basis->d_tab + (size_t)(2 * e->basis->p + 1) * q;
    size_t blocks = (e->n * q * q + 255) / 256;
    if (blocks > (size_t)ctx->sm_count * 16)
        blocks = (size_t)ctx->sm_count * 16;
    MFV2D_LAUNCH(jacobian_kernel, (unsigned)blocks, 256, 0, ctx->stream, e->d_corners, pnts, q, e->n, (double *)o.dev, coords);
    ctx->launches++;
    cudaError_t err = cudaGetLastError();
    int r = stage_finish(&o);
    if (err != cudaSuccess)
        return mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "jacobian kernel: %s", cudaGetErrorString(err));
    return r;
}

// launch of the fused matrix kernel for a lowered plan; all pointers are device pointers
int launch_element_matrix(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, double *d_out,
                          int32_t *d_status)
{
    mfv2d_ctx *ctx = plan->ctx;
    AsmArgs a{};
    a.tab = plan->basis->d_tab;
    a.inc = plan->basis->d_inc;
    a.corners = e->d_corners;
    a.blob = (const char *)plan->d_blob;
    for (int i = 0; i < 16; ++i)
        a.fields[i] = (fields && i < plan->n_fields) ? fields[i] : nullptr;
    a.out = d_out;
    a.status = d_status;
    a.n_elem = e->n;
    a.tab_len = plan->basis->tab_len;
    a.p = plan->p;
    a.q = plan->q;
    for (int i = 0; i < 4; ++i)
        a.inc_off[i] = plan->basis->inc_off[i];
    a.n_stages = (int)plan->stages.size();
    a.off_dense = plan->off_dense;
    a.off_terms = plan->off_terms;
    a.off_blocks = plan->off_blocks;
    a.off_units = plan->off_units;
    a.n_total = plan->out_ld ? plan->out_ld : plan->n_total;
    a.out_stride = plan->out_stride ? plan->out_stride : (unsigned long long)plan->n_total * plan->n_total;
    a.arena_off = asm_layout(plan->p, plan->q, a.tab_len).arena;
    int max_arena = 0;
    for (const DevStage &sg : plan->stages)
    {
        int end = 0;
        for (int d = sg.dense_begin; d < sg.dense_end; ++d)
        {
            const DevDense &dd = plan->dense[d];
            const int e2 = dd.smem_off + ((dd.rows * dd.cols + 1) & ~1);
            if (e2 > end)
                end = e2;
        }
        if (end > max_arena)
            max_arena = end;
    }
    const size_t smem = (size_t)(a.arena_off + max_arena) * sizeof(double);
    if (smem > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "element matrices of order %d need %zu bytes of shared memory", plan->p,
                          smem);
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_matrix_kernel, ASM_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    if (grid > e->n)
        grid = e->n;
    if (grid == 0)
        return MFV2D_SUCCESS;
    MFV2D_LAUNCH(element_matrix_kernel, (unsigned)grid, ASM_THREADS, smem, ctx->stream, a);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

extern "C" {

int mfv2d_cuda_jacobian_batch(const mfv2d_elements *e, double *out, int space) { return geometry_batch(e, out, space, 0); }

int mfv2d_cuda_coordinates_batch(const mfv2d_elements *e, double *out, int space)
{
    return geometry_batch(e, out, space, 1);
}

} // extern "C"
