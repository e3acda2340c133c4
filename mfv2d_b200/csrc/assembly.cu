// Batched element-matrix assembly: Jacobian, sum-factorised mass / interior-product matrices in
// shared memory, incidence application as signed gathers and coalesced row-major stores of the
// N_e x N_e element matrices.
//
// Replaces, per element: fem_space_2d_create (src/fem_space/fem_space.c:16-57),
// compute_mass_matrix_{node,edge,surf} (:235-411), compute_mass_matrix_{node_edge,edge_surf}
// (:546-846), matrix_full_invert (src/algebra/matrix.c:361-411), apply_incidence_to_full_{left,right}
// (src/evaluation/incidence.c:16-479) and the block loop of compute_element_matrix
// (src/evaluation/element_system.c:135-181).
//
// Organisation: a persistent CTA works on EPB elements at a time ("slots").  Every phase is a
// thread-strided loop over (item, slot) pairs with the slot index fastest, so that the basis tables
// are read as broadcasts and per-slot data sits in distinct banks (slot stride = 4 mod 16 doubles):
//   geometry -> [weight field -> T = sum_b B B' G -> M = sum_a A A' T] per tensor block -> emission.
// Symmetric blocks (M0, M2, the hh / vv blocks of M1) are computed for i <= k only and mirrored.
// The emission applies the incidence maps branch-free from (index, sign) tables: NL x NR gathers per
// output entry, with NL, NR in {1, 2, 4} compiled in.
//
// Roofline: bound by the HBM WRITE of the element matrices, 64 + 8 N_e^2 algorithmic bytes per element
// (SURVEY.md 8d); the sum-factorised contraction needs O(p^4 q) flops per element.
#include "dense_device.cuh"

#define ASM_THREADS 256
#define ASM_ROWS_PER_UNIT 16

struct AsmArgs
{
    const double *tab;
    const int32_t *inc_idx; // [rows][4]
    const float *inc_sgn;   // [rows][4]
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
    int epb_shift;                 // log2(elements per CTA pass)
    int slot_stride;               // doubles between slots
    int slot_off;                  // offset of slot 0 inside shared memory (doubles)
};

// x / n for 0 <= x < 2^20 and 1 <= n < 2^11 without an integer division
__device__ __forceinline__ int fast_div(int x, float inv_n) { return (int)(((float)x + 0.5f) * inv_n); }

struct AsmShared
{
    const double *node, *edge, *pnts, *wgts;
    const int *pairs_n, *pairs_e; // (i << 8) | k for i <= k, for n = p + 1 and n = p
    double *cs;                   // [epb][8]
    double *slot0;
    int slot_stride, epb, epb_shift, n_valid;
    SlotLayout sl;
    int p, q;
};

__device__ __forceinline__ double *slot_base(const AsmShared &s, int e) { return s.slot0 + (size_t)e * s.slot_stride; }

__device__ __forceinline__ void phase_geometry(const AsmShared &s)
{
    const int q = s.q, q2 = q * q;
    const float inv_q = 1.0f / (float)q;
    for (int idx = threadIdx.x; idx < (q2 << s.epb_shift); idx += blockDim.x)
    {
        const int e = idx & (s.epb - 1), pt = idx >> s.epb_shift;
        if (e >= s.n_valid)
            continue;
        const int a = fast_div(pt, inv_q), b = pt - a * q;
        const double *cs = s.cs + 8 * e;
        double *geo = slot_base(s, e);
        const double eta = s.pnts[a], xi = s.pnts[b];
        const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
        const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
        const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
        const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
        const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
        geo[pt] = j00;
        geo[q2 + pt] = j01;
        geo[2 * q2 + pt] = j10;
        geo[3 * q2 + pt] = j11;
        geo[4 * q2 + pt] = j00 * j11 - j10 * j01;
    }
}

__device__ __forceinline__ void phase_weight(const AsmShared &s, int kind, const double *field, unsigned long long elem0)
{
    const int q = s.q, q2 = q * q;
    const float inv_q = 1.0f / (float)q;
    for (int idx = threadIdx.x; idx < (q2 << s.epb_shift); idx += blockDim.x)
    {
        const int e = idx & (s.epb - 1), pt = idx >> s.epb_shift;
        if (e >= s.n_valid)
            continue;
        const int a = fast_div(pt, inv_q), b = pt - a * q;
        double *geo = slot_base(s, e);
        const double w = s.wgts[a] * s.wgts[b];
        const double det = geo[4 * q2 + pt];
        double v;
        switch (kind)
        {
        case W_M0:
            v = det * w;
            break;
        case W_M2:
            v = w / det;
            break;
        case W_HH: {
            const double j10 = geo[2 * q2 + pt], j11 = geo[3 * q2 + pt];
            v = (j11 * j11 + j10 * j10) / det * w;
            break;
        }
        case W_VV: {
            const double j00 = geo[pt], j01 = geo[q2 + pt];
            v = (j00 * j00 + j01 * j01) / det * w;
            break;
        }
        case W_HV: {
            const double j00 = geo[pt], j01 = geo[q2 + pt], j10 = geo[2 * q2 + pt], j11 = geo[3 * q2 + pt];
            v = (j01 * j11 + j00 * j10) / det * w;
            break;
        }
        default: {
            const double j00 = geo[pt], j01 = geo[q2 + pt], j10 = geo[2 * q2 + pt], j11 = geo[3 * q2 + pt];
            const double *fld = field + (elem0 + e) * (unsigned long long)(2 * q2);
            const double fx = fld[2 * pt], fy = fld[2 * pt + 1];
            if (kind == W_P1H)
                v = (fx * j11 - fy * j10) * w;
            else if (kind == W_P1V)
                v = -(fy * j00 - fx * j01) * w;
            else if (kind == W_P2H)
                v = -(fx * j10 + fy * j11) / det * w;
            else
                v = -(fx * j00 + fy * j01) / det * w;
            break;
        }
        }
        geo[s.sl.G + pt] = v;
    }
}

// T[a][j * nB2 + l] = sum_b B[j][b] B2[l][b] G[a][b]      (sym: B == B2, j <= l computed and mirrored)
__device__ __forceinline__ void phase_stage1(const AsmShared &s, const double *__restrict__ B, int nB,
                                             const double *__restrict__ B2, int nB2, bool sym, const int *pairs)
{
    const int q = s.q, ldT = s.sl.ldT;
    const int per_a = sym ? nB * (nB + 1) / 2 : nB * nB2;
    const float inv_per_a = 1.0f / (float)per_a, inv_nb2 = 1.0f / (float)nB2;
    for (int idx = threadIdx.x; idx < ((q * per_a) << s.epb_shift); idx += blockDim.x)
    {
        const int e = idx & (s.epb - 1), item = idx >> s.epb_shift;
        if (e >= s.n_valid)
            continue;
        const int a = fast_div(item, inv_per_a), rem = item - a * per_a;
        int j, l;
        if (sym)
        {
            const int pr = pairs[rem];
            j = pr >> 8;
            l = pr & 255;
        }
        else
        {
            j = fast_div(rem, inv_nb2);
            l = rem - j * nB2;
        }
        double *base = slot_base(s, e);
        const double *g = base + s.sl.G + a * q;
        const double *bj = B + j * q, *bl = B2 + l * q;
        double acc = 0.0;
        for (int b = 0; b < q; ++b)
            acc = fma(bj[b] * bl[b], g[b], acc);
        double *T = base + s.sl.T + a * ldT;
        T[j * nB2 + l] = acc;
        if (sym)
            T[l * nB2 + j] = acc;
    }
}

// dst[(r0 + i nB + j), (c0 + k nB2 + l)] = sum_a A[i][a] A2[k][a] T[a][j nB2 + l]
//   sym:    A == A2, r0 == c0: only i <= k, mirrored into the lower triangle
//   mirror: the block is also written transposed (symmetric off-diagonal block of M1)
//   trans:  the block is written transposed only (transposed interior products)
__device__ __forceinline__ void phase_stage2(const AsmShared &s, int dst_off, int ld, int r0, int c0, bool trans, bool mirror,
                                             bool sym, const double *__restrict__ A, int nA, const double *__restrict__ A2,
                                             int nA2, int nB, int nB2, const int *pairs)
{
    const int q = s.q, ldT = s.sl.ldT;
    const int nJL = nB * nB2;
    const int ntile = (nJL + 3) >> 2;
    const int n_ik = sym ? nA * (nA + 1) / 2 : nA * nA2;
    const float inv_ntile = 1.0f / (float)ntile, inv_na2 = 1.0f / (float)nA2, inv_nb2 = 1.0f / (float)nB2;
    for (int idx = threadIdx.x; idx < ((n_ik * ntile) << s.epb_shift); idx += blockDim.x)
    {
        const int e = idx & (s.epb - 1), item = idx >> s.epb_shift;
        if (e >= s.n_valid)
            continue;
        const int ik = fast_div(item, inv_ntile);
        const int jl0 = (item - ik * ntile) << 2;
        int i, k;
        if (sym)
        {
            const int pr = pairs[ik];
            i = pr >> 8;
            k = pr & 255;
        }
        else
        {
            i = fast_div(ik, inv_na2);
            k = ik - i * nA2;
        }
        double *base = slot_base(s, e);
        const double *T = base + s.sl.T + jl0;
        const double *ai = A + i * q, *ak = A2 + k * q;
        double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
        for (int a = 0; a < q; ++a)
        {
            const double pa = ai[a] * ak[a];
            const double2 t01 = *reinterpret_cast<const double2 *>(T + a * ldT);
            const double2 t23 = *reinterpret_cast<const double2 *>(T + a * ldT + 2);
            acc0 = fma(pa, t01.x, acc0);
            acc1 = fma(pa, t01.y, acc1);
            acc2 = fma(pa, t23.x, acc2);
            acc3 = fma(pa, t23.y, acc3);
        }
        double *dst = base + s.sl.arena + dst_off;
        const double acc[4] = {acc0, acc1, acc2, acc3};
        int j = fast_div(jl0, inv_nb2), l = jl0 - j * nB2;
#pragma unroll
        for (int t = 0; t < 4; ++t)
        {
            if (jl0 + t < nJL)
            {
                const int row = r0 + i * nB + j, col = c0 + k * nB2 + l;
                if (!trans)
                    dst[row * ld + col] = acc[t];
                if (trans || mirror || (sym && i != k))
                    dst[col * ld + row] = acc[t];
            }
            if (++l == nB2)
            {
                l = 0;
                ++j;
            }
        }
    }
}

__device__ __noinline__ void tensor_block_multi(const AsmShared &s, int wkind, const double *field, unsigned long long elem0,
                                                   int dst_off, int ld, int r0, int c0, bool trans, bool mirror, bool sym,
                                                   const double *A, int nA, const double *A2, int nA2, const double *B,
                                                   int nB, const double *B2, int nB2)
{
    const int np1 = s.p + 1;
    phase_weight(s, wkind, field, elem0);
    __syncthreads();
    const bool sym1 = (B == B2);
    phase_stage1(s, B, nB, B2, nB2, sym1, nB == np1 ? s.pairs_n : s.pairs_e);
    __syncthreads();
    phase_stage2(s, dst_off, ld, r0, c0, trans, mirror, sym, A, nA, A2, nA2, nB, nB2, nA == np1 ? s.pairs_n : s.pairs_e);
    // the next phase_weight only writes G (not read by stage 2); stage 1 of the next block is behind a barrier
}

// in-place SPD inverse of the n x n matrices at arena + off of every slot (Gauss-Jordan, no pivoting)
__device__ void invert_spd_multi(const AsmShared &s, int off, int n, int *bad)
{
    const float inv_n = 1.0f / (float)n;
    for (int k = 0; k < n; ++k)
    {
        __syncthreads();
        for (int idx = threadIdx.x; idx < (n << s.epb_shift); idx += blockDim.x)
        {
            const int e = idx & (s.epb - 1), i = idx >> s.epb_shift;
            double *base = slot_base(s, e);
            const double *D = base + s.sl.arena + off;
            double *colk = base + s.sl.T, *rowk = colk + n;
            colk[i] = D[i * n + k];
            rowk[i] = D[k * n + i];
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < ((n * n) << s.epb_shift); idx += blockDim.x)
        {
            const int e = idx & (s.epb - 1), ij = idx >> s.epb_shift;
            if (e >= s.n_valid)
                continue;
            const int i = fast_div(ij, inv_n), j = ij - i * n;
            double *base = slot_base(s, e);
            double *D = base + s.sl.arena + off;
            const double *colk = base + s.sl.T, *rowk = colk + n;
            const double pv = colk[k];
            if (ij == 0 && !(fabs(pv) > 1e-300))
                bad[e] = 1;
            const double ip = 1.0 / pv;
            double v;
            if (i == k)
                v = (j == k) ? ip : rowk[j] * ip;
            else if (j == k)
                v = -colk[i] * ip;
            else
                v = fma(-colk[i] * ip, rowk[j], D[ij]);
            D[ij] = v;
        }
    }
    __syncthreads();
}

__device__ void compute_dense_multi(const AsmShared &s, const DevDense &dd, const double *const *fields,
                                    unsigned long long elem0, int *bad)
{
    const int p = s.p, np1 = p + 1, nh = p * (p + 1);
    const int off = dd.smem_off;
    switch (dd.dk)
    {
    case DK_MASS0:
        tensor_block_multi(s, W_M0, nullptr, elem0, off, np1 * np1, 0, 0, false, false, true, s.node, np1, s.node, np1, s.node,
                           np1, s.node, np1);
        break;
    case DK_MASS2:
        tensor_block_multi(s, W_M2, nullptr, elem0, off, p * p, 0, 0, false, false, true, s.edge, p, s.edge, p, s.edge, p,
                           s.edge, p);
        break;
    case DK_MASS1: {
        const int n1 = 2 * nh;
        tensor_block_multi(s, W_HH, nullptr, elem0, off, n1, 0, 0, false, false, true, s.node, np1, s.node, np1, s.edge, p,
                           s.edge, p);
        tensor_block_multi(s, W_VV, nullptr, elem0, off, n1, nh, nh, false, false, true, s.edge, p, s.edge, p, s.node, np1,
                           s.node, np1);
        tensor_block_multi(s, W_HV, nullptr, elem0, off, n1, 0, nh, false, true, false, s.node, np1, s.edge, p, s.edge, p,
                           s.node, np1);
        break;
    }
    case DK_IPROD1: {
        const double *fld = fields[dd.field];
        const int ld = dd.transpose ? np1 * np1 : 2 * nh;
        tensor_block_multi(s, W_P1H, fld, elem0, off, ld, 0, 0, dd.transpose, false, false, s.node, np1, s.node, np1, s.node,
                           np1, s.edge, p);
        tensor_block_multi(s, W_P1V, fld, elem0, off, ld, 0, nh, dd.transpose, false, false, s.node, np1, s.edge, p, s.node,
                           np1, s.node, np1);
        break;
    }
    case DK_IPROD2: {
        const double *fld = fields[dd.field];
        const int ld = dd.transpose ? 2 * nh : p * p;
        tensor_block_multi(s, W_P2H, fld, elem0, off, ld, 0, 0, dd.transpose, false, false, s.node, np1, s.edge, p, s.edge, p,
                           s.edge, p);
        tensor_block_multi(s, W_P2V, fld, elem0, off, ld, nh, 0, dd.transpose, false, false, s.edge, p, s.edge, p, s.node, np1,
                           s.edge, p);
        break;
    }
    default:
        break;
    }
    if (dd.inverse)
        invert_spd_multi(s, off, dd.rows, bad);
}

// ---- emission ------------------------------------------------------------------------------------
struct EmitCtx
{
    const double *D; // dense matrix of the term in this slot
    int ld;
    double coef;
    const int4 *lidx, *ridx; // incidence tables of the term (nullptr: none), already offset to the kind
    const float4 *lsgn, *rsgn;
};

// one block strip: rows [r0, r0 + nr) x 32 columns starting at c0; single-term blocks, NL x NR gathers
template <int NL, int NR>
__device__ __forceinline__ void emit_strip(const EmitCtx &t, int r0, int nr, int c, bool active, double *__restrict__ o, int N,
                                           bool accumulate)
{
    int rc[NR];
    double rs[NR];
    if (NR == 1)
    {
        rc[0] = active ? c : 0;
        rs[0] = t.coef;
    }
    else
    {
        const int4 iv = active ? ld_ro(t.ridx + c) : make_int4(0, 0, 0, 0);
        const float4 sv = active ? ld_ro(t.rsgn + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        const int ia[4] = {iv.x, iv.y, iv.z, iv.w};
        const float sa[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
        for (int b = 0; b < NR; ++b)
        {
            rc[b] = ia[b];
            rs[b] = t.coef * (double)sa[b];
        }
    }
    for (int r = 0; r < nr; ++r)
    {
        int lr[NL];
        double ls[NL];
        if (NL == 1)
        {
            lr[0] = (r0 + r) * t.ld;
            ls[0] = 1.0;
        }
        else
        {
            const int4 iv = ld_ro(t.lidx + r0 + r);
            const float4 sv = ld_ro(t.lsgn + r0 + r);
            const int ia[4] = {iv.x, iv.y, iv.z, iv.w};
            const float sa[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
            for (int a = 0; a < NL; ++a)
            {
                lr[a] = ia[a] * t.ld;
                ls[a] = (double)sa[a];
            }
        }
        double v = 0.0;
#pragma unroll
        for (int a = 0; a < NL; ++a)
#pragma unroll
            for (int b = 0; b < NR; ++b)
                v = fma(NL == 1 ? rs[b] : ls[a] * rs[b], t.D[lr[a] + rc[b]], v);
        if (active)
        {
            if (accumulate)
                o[(size_t)r * N] += v;
            else
                o[(size_t)r * N] = v;
        }
    }
}

__device__ __forceinline__ int inc_width(int kind) // non-zeros per row of E (left) / per column of E (right, via kind ^ 2)
{
    return kind < 0 ? 1 : ((kind == MFV2D_INC_10 || kind == MFV2D_INC_21_T) ? 2 : 4);
}

__global__ void __launch_bounds__(ASM_THREADS, 2) element_matrix_kernel(const AsmArgs args)
{
    MFV2D_DYN_SMEM(double, smem);
    const int p = args.p, q = args.q, np1 = p + 1;
    AsmShared s;
    s.p = p;
    s.q = q;
    s.node = smem;
    s.edge = s.node + np1 * q;
    s.pnts = s.edge + p * q;
    s.wgts = s.pnts + q;
    int *pairs = reinterpret_cast<int *>(smem + ((args.tab_len + 1) & ~1));
    const int n_pairs_n = np1 * (np1 + 1) / 2, n_pairs_e = p * (p + 1) / 2;
    s.pairs_n = pairs;
    s.pairs_e = pairs + n_pairs_n;
    s.epb_shift = args.epb_shift;
    s.epb = 1 << args.epb_shift;
    s.cs = smem + args.slot_off - 8 * s.epb;
    s.slot0 = smem + args.slot_off;
    s.slot_stride = args.slot_stride;
    s.sl = slot_layout(p, q);
    __shared__ int bad[8];

    for (int i = threadIdx.x; i < args.tab_len; i += blockDim.x)
        smem[i] = args.tab[i];
    for (int n = threadIdx.x; n < np1 * np1; n += blockDim.x)
    {
        const int i = n / np1, k = n - i * np1;
        if (i <= k)
        {
            // position of (i, k) in the row-wise enumeration of the upper triangle
            pairs[i * np1 - i * (i - 1) / 2 + (k - i)] = (i << 8) | k;
            if (k < p)
                pairs[n_pairs_n + i * p - i * (i - 1) / 2 + (k - i)] = (i << 8) | k;
        }
    }

    const DevStage *stages = reinterpret_cast<const DevStage *>(args.blob);
    const DevDense *dense = reinterpret_cast<const DevDense *>(args.blob + args.off_dense);
    const DevTerm *terms = reinterpret_cast<const DevTerm *>(args.blob + args.off_terms);
    const DevBlock *blocks = reinterpret_cast<const DevBlock *>(args.blob + args.off_blocks);
    const DevUnit *units = reinterpret_cast<const DevUnit *>(args.blob + args.off_units);
    const int4 *inc_idx = reinterpret_cast<const int4 *>(args.inc_idx);
    const float4 *inc_sgn = reinterpret_cast<const float4 *>(args.inc_sgn);
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    const int N = args.n_total;
    const unsigned long long n_groups = (args.n_elem + s.epb - 1) >> s.epb_shift;

    for (unsigned long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x)
    {
        const unsigned long long elem0 = grp << s.epb_shift;
        s.n_valid = (int)((args.n_elem - elem0) < (unsigned long long)s.epb ? (args.n_elem - elem0) : s.epb);
        __syncthreads(); // the previous group's emission has finished reading shared memory
        for (int i = threadIdx.x; i < 8 * s.n_valid; i += blockDim.x)
            s.cs[i] = args.corners[elem0 * 8 + i];
        for (int i = threadIdx.x; i < 8; i += blockDim.x)
            bad[i] = 0;
        __syncthreads();
        phase_geometry(s);
        for (int st = 0; st < args.n_stages; ++st)
        {
            const DevStage sg = stages[st];
            __syncthreads(); // geometry ready / previous stage's emission done
            for (int d = sg.dense_begin; d < sg.dense_end; ++d)
                compute_dense_multi(s, dense[d], args.fields, elem0, bad);
            __syncthreads();
            // emission: each warp takes (strip, slot) pairs; a strip is <= 16 rows x 32 columns of one block
            const int n_units = sg.unit_end - sg.unit_begin;
            for (int w = warp; w < (n_units << s.epb_shift); w += nwarps)
            {
                const int e = w & (s.epb - 1);
                if (e >= s.n_valid)
                    continue;
                const DevUnit un = units[sg.unit_begin + (w >> s.epb_shift)];
                const DevBlock db = blocks[un.block];
                const double *arena = slot_base(s, e) + s.sl.arena;
                double *const out = args.out + (elem0 + e) * args.out_stride + (size_t)(db.row0 + un.r0) * N + db.col0;
                const int n_terms = db.term_end - db.term_begin;
                for (int lane = MFV2D_LANE0; lane < 32; lane += MFV2D_LANE_STEP)
                {
                    const int c = un.c0 + lane;
                    const bool active = c < db.ncols;
                    double *o = out + c;
                    if (n_terms == 0)
                    {
                        if (active && !db.accumulate)
                            for (int r = 0; r < un.nr; ++r)
                                o[(size_t)r * N] = 0.0;
                        continue;
                    }
                    for (int ti = 0; ti < n_terms; ++ti)
                    {
                        const DevTerm tm = terms[db.term_begin + ti];
                        const bool acc = db.accumulate || ti > 0;
                        if (tm.dense < 0)
                        {
                            // coef * identity
                            if (active)
                                for (int r = 0; r < un.nr; ++r)
                                {
                                    const double v = (un.r0 + r == c) ? tm.coef : 0.0;
                                    if (acc)
                                        o[(size_t)r * N] += v;
                                    else
                                        o[(size_t)r * N] = v;
                                }
                            continue;
                        }
                        const DevDense dd = dense[tm.dense];
                        EmitCtx t;
                        t.D = arena + dd.smem_off;
                        t.ld = dd.cols;
                        t.coef = tm.coef;
                        t.lidx = tm.L >= 0 ? inc_idx + args.inc_off[tm.L] : nullptr;
                        t.lsgn = tm.L >= 0 ? inc_sgn + args.inc_off[tm.L] : nullptr;
                        // column c of E_R = row c of its transpose
                        t.ridx = tm.R >= 0 ? inc_idx + args.inc_off[tm.R ^ 2] : nullptr;
                        t.rsgn = tm.R >= 0 ? inc_sgn + args.inc_off[tm.R ^ 2] : nullptr;
                        const int nl = inc_width(tm.L), nr_ = inc_width(tm.R < 0 ? -1 : (tm.R ^ 2));
                        switch (nl * 8 + nr_)
                        {
                        case 1 * 8 + 1:
                            emit_strip<1, 1>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 1 * 8 + 2:
                            emit_strip<1, 2>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 1 * 8 + 4:
                            emit_strip<1, 4>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 2 * 8 + 1:
                            emit_strip<2, 1>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 4 * 8 + 1:
                            emit_strip<4, 1>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 2 * 8 + 2:
                            emit_strip<2, 2>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 2 * 8 + 4:
                            emit_strip<2, 4>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        case 4 * 8 + 2:
                            emit_strip<4, 2>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        default:
                            emit_strip<4, 4>(t, un.r0, un.nr, c, active, o, N, acc);
                            break;
                        }
                    }
                }
            }
        }
        __syncthreads();
        if (args.status)
            for (int e = threadIdx.x; e < s.n_valid; e += blockDim.x)
                args.status[elem0 + e] = bad[e] ? MFV2D_PIVOT_FAILED : MFV2D_SUCCESS;
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

// the dense arena one element needs (maximum over the stages), in doubles
static int plan_arena_doubles(const mfv2d_plan *plan)
{
    int max_arena = 0;
    for (const DevStage &sg : plan->stages)
        for (int d = sg.dense_begin; d < sg.dense_end; ++d)
        {
            const DevDense &dd = plan->dense[d];
            const int end = dd.smem_off + ((dd.rows * dd.cols + 1) & ~1);
            if (end > max_arena)
                max_arena = end;
        }
    return max_arena;
}

// launch of the fused matrix kernel for a lowered plan; all pointers are device pointers
int launch_element_matrix(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, double *d_out,
                          int32_t *d_status)
{
    mfv2d_ctx *ctx = plan->ctx;
    if (e->n == 0)
        return MFV2D_SUCCESS;
    AsmArgs a{};
    a.tab = plan->basis->d_tab;
    a.inc_idx = plan->basis->d_inc_idx;
    a.inc_sgn = plan->basis->d_inc_sgn;
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
    // slot size; the stride is padded to 4 (mod 16) doubles so that the slots of a pass sit in distinct banks
    const SlotLayout sl = slot_layout(plan->p, plan->q);
    int slot = sl.arena + plan_arena_doubles(plan);
    slot = ((slot + 15) & ~15) + 4;
    a.slot_stride = slot;
    // elements per pass: as many as keep two CTAs per SM resident (<= ~110 KB each), at most 8
    const size_t budget = (size_t)110 * 1024 / sizeof(double);
    int shift = 0;
    while (shift < 3 && (size_t)asm_fixed_doubles(plan->p, a.tab_len, 2 << shift) + (size_t)(2 << shift) * slot <= budget &&
           (size_t)(2 << shift) <= e->n)
        ++shift;
    a.epb_shift = shift;
    const int epb = 1 << shift;
    a.slot_off = asm_fixed_doubles(plan->p, a.tab_len, epb);
    const size_t smem = ((size_t)a.slot_off + (size_t)epb * slot + 8) * sizeof(double);
    if (smem > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "element matrices of order %d need %zu bytes of shared memory", plan->p,
                          smem);
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_matrix_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_matrix_kernel, ASM_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    const unsigned long long n_groups = (e->n + epb - 1) / epb;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    if (grid > n_groups)
        grid = n_groups;
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
