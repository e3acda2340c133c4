// Batched element-matrix assembly: Jacobian and weight fields, sum-factorised mass / interior-product
// matrices in shared memory, incidence application as signed gathers and row-major stores of the
// N_e x N_e element matrices.
//
// Replaces, per element: fem_space_2d_create (src/fem_space/fem_space.c:16-57),
// compute_mass_matrix_{node,edge,surf} (:235-411), compute_mass_matrix_{node_edge,edge_surf}
// (:546-846), matrix_full_invert (src/algebra/matrix.c:361-411), apply_incidence_to_full_{left,right}
// (src/evaluation/incidence.c:16-479) and the block loop of compute_element_matrix
// (src/evaluation/element_system.c:135-181).
//
// Organisation.  A persistent CTA keeps the plan's descriptor blob (stages, tensor blocks, emission
// units, incidence rows, the basis tables and their pairwise products) in shared memory and works on
// `epb` elements at a time ("slots").  Per pass and plan stage:
//   W   every thread takes quadrature points (e, a, b): Jacobian in registers, all weight fields of the
//       stage written to the slot                                                      (1 barrier)
//   S1  T[(j,l)][a] = sum_b (B_j B2_l)(b) W[a][b]: a thread holds one row W[a][:] in registers and walks a
//       chunk of the (j,l) pair list, reading the product table rows as 16-byte broadcasts (1 barrier)
//   S2  D[(i,j),(k,l)] = sum_a (A_i A2_k)(a) T[(j,l)][a]: a thread holds two columns T[(j,l)][:] in
//       registers and walks a chunk of the (i,k) pair list; i <-> k and transposed images are stored
//       from the same accumulator                                                      (1 barrier)
//   E   emission: a warp takes a tile of one block, lanes = columns (x rows for narrow blocks); the
//       right incidence row is fetched once per lane, the left one once per row; coalesced stores.
// S1 and S2 are templated on the number of quadrature points so that the register tiles are real
// registers.  Symmetric index pairs are computed once.
//
// Roofline: bound by the HBM WRITE of the element matrices, 64 + 8 N_e^2 algorithmic bytes per element
// (SURVEY.md 8d); the sum-factorised contraction needs O(p^4 q) flops per element.
#include "plan.cuh"

#define ASM_THREADS 256

struct AsmArgs
{
    const char *blob;
    const double *corners;
    const double *fields[16]; // per field: [n_elem][q*q][2]
    double *out;
    int32_t *status;
    unsigned long long n_elem;
    int p, q;
    int n_total;                   // leading dimension of the output rows
    unsigned long long out_stride; // doubles between consecutive elements of the output
    AsmLayout lay;
};

// x / n for 0 <= x < 2^20 and 1 <= n < 2^11 without an integer division
__device__ __forceinline__ int fast_div(int x, float inv_n) { return (int)(((float)x + 0.5f) * inv_n); }

template <int Q> __device__ __forceinline__ void load_row(double (&r)[Q], const double *__restrict__ src)
{
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(src + b);
        r[b] = v.x;
        r[b + 1] = v.y;
    }
    if (Q & 1)
        r[Q - 1] = src[Q - 1];
}

template <int Q> __device__ __forceinline__ void dot_row2(const double *__restrict__ row, const double (&t0)[Q],
                                                          const double (&t1)[Q], double &r0, double &r1)
{
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(row + b);
        a0 = fma(v.x, t0[b], a0);
        a1 = fma(v.x, t1[b], a1);
        a0 = fma(v.y, t0[b + 1], a0);
        a1 = fma(v.y, t1[b + 1], a1);
    }
    if (Q & 1)
    {
        const double v = row[Q - 1];
        a0 = fma(v, t0[Q - 1], a0);
        a1 = fma(v, t1[Q - 1], a1);
    }
    r0 = a0;
    r1 = a1;
}

template <int Q> __device__ __forceinline__ double dot_row(const double *__restrict__ row, const double (&g)[Q])
{
    double a0 = 0.0, a1 = 0.0;
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(row + b);
        a0 = fma(v.x, g[b], a0);
        a1 = fma(v.y, g[b + 1], a1);
    }
    if (Q & 1)
        a0 = fma(row[Q - 1], g[Q - 1], a0);
    return a0 + a1;
}

// ---- W: Jacobian (src/fem_space/fem_space.c:39-53) and the weight fields of a stage -------------------
template <int Q>
__device__ __forceinline__ void phase_weights(const AsmArgs &args, double *smem, const DevStage &sg, const DevW *ws,
                                              const double *pnts, const double *wgts, unsigned long long elem0, int n_valid)
{
    const AsmLayout &L = args.lay;
    constexpr int q2 = Q * Q;
    const int ldq = L.ldq;
    for (int idx = threadIdx.x; idx < n_valid * q2; idx += blockDim.x)
    {
        const int e = idx / q2, pt = idx - e * q2;
        const int a = pt / Q, b = pt - a * Q;
        const double *cs = smem + L.off_cs + 8 * e;
        const double eta = pnts[a], xi = pnts[b];
        const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
        const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
        const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
        const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
        const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
        const double det = j00 * j11 - j10 * j01;
        const double w = wgts[a] * wgts[b];
        double *W = smem + L.off_slots + e * L.slot_stride + L.slot_w + a * ldq + b;
        for (int wi = sg.w_begin; wi < sg.w_end; ++wi, W += Q * ldq)
        {
            const DevW dw = ws[wi];
            double v;
            switch (dw.kind)
            {
            case W_M0:
                v = det * w;
                break;
            case W_M2:
                v = w / det;
                break;
            case W_HH:
                v = (j11 * j11 + j10 * j10) / det * w;
                break;
            case W_VV:
                v = (j00 * j00 + j01 * j01) / det * w;
                break;
            case W_HV:
                v = (j01 * j11 + j00 * j10) / det * w;
                break;
            default: {
                const double *fld = args.fields[dw.field] + (elem0 + e) * (unsigned long long)(2 * q2);
                const double fx = fld[2 * pt], fy = fld[2 * pt + 1];
                if (dw.kind == W_P1H)
                    v = (fx * j11 - fy * j10) * w;
                else if (dw.kind == W_P1V)
                    v = -(fy * j00 - fx * j01) * w;
                else if (dw.kind == W_P2H)
                    v = -(fx * j10 + fy * j11) / det * w;
                else
                    v = -(fx * j00 + fy * j01) / det * w;
                break;
            }
            }
            *W = v;
        }
    }
}

// ---- S1: T[(j,l)][a] = sum_b B_j(b) B2_l(b) W[a][b] -----------------------------------------------------
// work item (per element): tensor block, quadrature row a, a chunk of the block's (j,l) pair table
template <int Q>
__device__ __forceinline__ void phase_s1(const AsmArgs &args, double *smem, const DevRound &rd, const DevTB *tbs, int n_valid)
{
    const AsmLayout &L = args.lay;
    const int ldq = L.ldq;
    const char *sb = reinterpret_cast<const char *>(smem);
    const int *xt = reinterpret_cast<const int *>(sb + L.off_x) + rd.x1_begin;
    const DevS1Pair *s1 = reinterpret_cast<const DevS1Pair *>(sb + L.off_s1);
    const int per = rd.x1_count;
    const float inv_per = 1.0f / (float)per;
    for (int x = threadIdx.x; x < per * n_valid; x += blockDim.x)
    {
        const int e = fast_div(x, inv_per);
        const int item = xt[x - e * per];
        const DevTB &tb = tbs[rd.tb_begin + (item & 15)];
        const int a = (item >> 4) & 31, pb = (item >> 9) & 255, len = item >> 17;
        double *slot = smem + L.off_slots + e * L.slot_stride;
        double g[Q];
        load_row<Q>(g, slot + L.slot_w + tb.w * (Q * ldq) + a * ldq);
        double *T = slot + L.slot_t + tb.t_off + a;
        const DevS1Pair *pr = s1 + tb.s1_tab + pb;
#pragma unroll 2
        for (int n = 0; n < len; ++n)
        {
            const int4 d = *reinterpret_cast<const int4 *>(pr + n);
            const double acc = dot_row<Q>(smem + d.x, g);
            T[d.y] = acc;
            if (d.z >= 0)
                T[d.z] = acc;
        }
    }
}

// ---- S2: D[(i,j),(k,l)] = sum_a A_i(a) A2_k(a) T[(j,l)][a] ------------------------------------------------
// work item (per element): tensor block, the two columns (j0,l0) and the following one, a chunk of the (i,k) table
template <int Q>
__device__ __forceinline__ void phase_s2(const AsmArgs &args, double *smem, const DevRound &rd, const DevTB *tbs, int n_valid)
{
    const AsmLayout &L = args.lay;
    const int ldq = L.ldq;
    const char *sb = reinterpret_cast<const char *>(smem);
    const int *xt = reinterpret_cast<const int *>(sb + L.off_x) + rd.x2_begin;
    const DevS2Pair *s2 = reinterpret_cast<const DevS2Pair *>(sb + L.off_s2);
    const int per = rd.x2_count;
    const float inv_per = 1.0f / (float)per;
    for (int x = threadIdx.x; x < per * n_valid; x += blockDim.x)
    {
        const int e = fast_div(x, inv_per);
        const int item = xt[x - e * per];
        const DevTB &tb = tbs[rd.tb_begin + (item & 15)];
        const int j0 = (item >> 4) & 31, l0 = (item >> 9) & 31, pb = (item >> 14) & 255, len = item >> 22;
        const int nB2 = tb.nB2, ld = tb.ld;
        const int jl0 = j0 * nB2 + l0;
        // the item's second column is half the block away, so that the lanes of a warp walk consecutive columns
        const int nJL = tb.nB * nB2, jl1 = jl0 + ((nJL + 1) >> 1);
        const bool has2 = jl1 < nJL;
        double *slot = smem + L.off_slots + e * L.slot_stride;
        const double *T = slot + L.slot_t + tb.t_off;
        double t0[Q], t1[Q];
        load_row<Q>(t0, T + jl0 * ldq);
        load_row<Q>(t1, T + (has2 ? jl1 : jl0) * ldq);
        const int j1 = fast_div(jl1, 1.0f / (float)nB2), l1 = jl1 - j1 * nB2;
        double *D = slot + L.slot_arena + tb.dst_off;
        double *Dd0 = D + (tb.r0 + j0) * ld + tb.c0 + l0, *Df0 = D + (tb.c0 + l0) * ld + tb.r0 + j0;
        double *Dd1 = D + (tb.r0 + j1) * ld + tb.c0 + l1, *Df1 = D + (tb.c0 + l1) * ld + tb.r0 + j1;
        const DevS2Pair *pr = s2 + tb.s2_tab + pb;
#pragma unroll 2
        for (int n = 0; n < len; ++n)
        {
            const int4 d = *reinterpret_cast<const int4 *>(pr + n);
            const int f1 = pr[n].f1;
            double v0, v1;
            dot_row2<Q>(smem + d.x, t0, t1, v0, v1);
            if (d.y >= 0)
                Dd0[d.y] = v0;
            if (d.z >= 0)
                Df0[d.z] = v0;
            if (d.w >= 0)
                Dd0[d.w] = v0;
            if (f1 >= 0)
                Df0[f1] = v0;
            if (has2)
            {
                if (d.y >= 0)
                    Dd1[d.y] = v1;
                if (d.z >= 0)
                    Df1[d.z] = v1;
                if (d.w >= 0)
                    Dd1[d.w] = v1;
                if (f1 >= 0)
                    Df1[f1] = v1;
            }
        }
    }
}

#ifndef MFV2D_EMU
// ---- S1 + S2 on the tensor cores --------------------------------------------------------------
// mma.sync m8n8k4 f64: lane (g = lane / 4, t = lane % 4) holds A[m = g][k = t], B[k = t][n = g] and
// D[m = g][n = 2t, 2t + 1].  The two k-steps of a contraction over 8 quadrature points take the points 2t and
// 2t + 1, so that a lane's pair of D values of stage 1 (T[(j,l) = g][a = 2t, 2t + 1]) is exactly its pair of B
// operands of stage 2: T never leaves the registers and the two stages need no barrier between them.  More than
// 8 points: NQ = 2 tiles of 8 (zero-padded by predicates), the same correspondence per tile.
__device__ __forceinline__ void asm_dmma(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0, %1}, {%2}, {%3}, {%0, %1};"
                 : "+d"(d0), "+d"(d1)
                 : "d"(a), "d"(b));
}

// work item (per element): tensor block, tile of 8 (j,l) pairs.  Stage 1: T = PT_B (8 pairs x Q) W^T (Q x Q rows a);
// stage 2, per tile of 8 (i,k) pairs: PT_A (8 pairs x Q) T^T, scattered to all images of the entries in the arena.
template <int Q>
__device__ __forceinline__ void phase_tc(const AsmArgs &args, double *smem, const DevStage &sg, const DevTB *tbs, int n_valid)
{
    const AsmLayout &L = args.lay;
    const int ldq = L.ldq;
    const char *sb = reinterpret_cast<const char *>(smem);
    const int *xt = reinterpret_cast<const int *>(sb + L.off_tc) + sg.tc_begin;
    const DevS1Pair *s1 = reinterpret_cast<const DevS1Pair *>(sb + L.off_s1);
    const DevS2Pair *s2 = reinterpret_cast<const DevS2Pair *>(sb + L.off_s2);
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    constexpr int NQ = (Q + 7) / 8; // tiles of 8 quadrature points
    const int per = sg.tc_count;
    const float inv_per = 1.0f / (float)per;
    for (int x = threadIdx.x >> 5; x < per * n_valid; x += blockDim.x >> 5)
    {
        const int e = fast_div(x, inv_per);
        const int item = xt[x - e * per];
        const DevTB &tb = tbs[item & 0xffff];
        const int nt = (item >> 16) & 255, mt = item >> 24;
        double *slot = smem + L.off_slots + e * L.slot_stride;
        const int npA = tb.npA, npB = tb.npB, ld = tb.ld;
        // stage 1: NQ tiles of 8 rows a, NQ pairs of k-steps over b
        double T[NQ][2];
        {
            const DevS1Pair *pjl0 = s1 + tb.s1_tab;
            const int njl = 8 * nt + g;
            const bool vjl = njl < npB;
            const double *ptb = smem + pjl0[vjl ? njl : 0].pt + 2 * t;
            const double *wf = slot + L.slot_w + tb.w * (Q * ldq) + 2 * t;
            double2 pb[NQ];
#pragma unroll
            for (int kb = 0; kb < NQ; ++kb)
            {
                pb[kb] = *reinterpret_cast<const double2 *>(ptb + 8 * kb);
                pb[kb].x = (vjl && 8 * kb + 2 * t < Q) ? pb[kb].x : 0.0;
                pb[kb].y = (vjl && 8 * kb + 2 * t + 1 < Q) ? pb[kb].y : 0.0;
            }
#pragma unroll
            for (int na = 0; na < NQ; ++na)
            {
                const int a = 8 * na + g;
                T[na][0] = T[na][1] = 0.0;
#pragma unroll
                for (int kb = 0; kb < NQ; ++kb)
                {
                    double2 w = *reinterpret_cast<const double2 *>(wf + (a < Q ? a : 0) * ldq + 8 * kb);
                    w.x = (a < Q && 8 * kb + 2 * t < Q) ? w.x : 0.0;
                    w.y = (a < Q && 8 * kb + 2 * t + 1 < Q) ? w.y : 0.0;
                    asm_dmma(T[na][0], T[na][1], pb[kb].x, w.x);
                    asm_dmma(T[na][0], T[na][1], pb[kb].y, w.y);
                }
            }
        }
        const DevS1Pair *pjl = s1 + tb.s1_tab;
        // the lane's two columns of stage 2: pairs 8 nt + 2t and the next one
        const int c = 8 * nt + 2 * t;
        const bool vc0 = c < npB, vc1 = c + 1 < npB;
        const int4 e0 = *reinterpret_cast<const int4 *>(pjl + (vc0 ? c : 0));
        const int4 e1 = *reinterpret_cast<const int4 *>(pjl + (vc1 ? c + 1 : 0));
        const int j0 = e0.w & 255, l0 = e0.w >> 8, j1 = e1.w & 255, l1 = e1.w >> 8;
        const bool m0 = vc0 && e0.z >= 0, m1 = vc1 && e1.z >= 0; // the (l,j) image exists
        double *D = slot + L.slot_arena + tb.dst_off;
        // direct entries live at row r0 + (i,j), column c0 + (k,l); flipped ones at the transposed place
        double *Dd0 = D + (tb.r0 + j0) * ld + tb.c0 + l0, *Df0 = D + (tb.c0 + l0) * ld + tb.r0 + j0;
        double *Dm0 = D + (tb.r0 + l0) * ld + tb.c0 + j0, *Dn0 = D + (tb.c0 + j0) * ld + tb.r0 + l0;
        double *Dd1 = D + (tb.r0 + j1) * ld + tb.c0 + l1, *Df1 = D + (tb.c0 + l1) * ld + tb.r0 + j1;
        double *Dm1 = D + (tb.r0 + l1) * ld + tb.c0 + j1, *Dn1 = D + (tb.c0 + j1) * ld + tb.r0 + l1;
        const DevS2Pair *pik = s2 + tb.s2_tab;
        {
            const int nik = 8 * mt + g;
            const bool vik = nik < npA;
            const DevS2Pair *pr = pik + (vik ? nik : 0);
            const int4 d = *reinterpret_cast<const int4 *>(pr);
            const int f1 = pr->f1;
            double v0 = 0.0, v1 = 0.0;
#pragma unroll
            for (int ka = 0; ka < NQ; ++ka)
            {
                double2 pa = *reinterpret_cast<const double2 *>(smem + d.x + 8 * ka + 2 * t);
                pa.x = (vik && 8 * ka + 2 * t < Q) ? pa.x : 0.0;
                pa.y = (vik && 8 * ka + 2 * t + 1 < Q) ? pa.y : 0.0;
                asm_dmma(v0, v1, pa.x, T[ka][0]);
                asm_dmma(v0, v1, pa.y, T[ka][1]);
            }
            if (vik)
            {
                if (d.y >= 0)
                {
                    if (vc0)
                        Dd0[d.y] = v0;
                    if (m0)
                        Dm0[d.y] = v0;
                    if (vc1)
                        Dd1[d.y] = v1;
                    if (m1)
                        Dm1[d.y] = v1;
                }
                if (d.z >= 0)
                {
                    if (vc0)
                        Df0[d.z] = v0;
                    if (m0)
                        Dn0[d.z] = v0;
                    if (vc1)
                        Df1[d.z] = v1;
                    if (m1)
                        Dn1[d.z] = v1;
                }
                if (d.w >= 0)
                {
                    if (vc0)
                        Dd0[d.w] = v0;
                    if (m0)
                        Dm0[d.w] = v0;
                    if (vc1)
                        Dd1[d.w] = v1;
                    if (m1)
                        Dm1[d.w] = v1;
                }
                if (f1 >= 0)
                {
                    if (vc0)
                        Df0[f1] = v0;
                    if (m0)
                        Dn0[f1] = v0;
                    if (vc1)
                        Df1[f1] = v1;
                    if (m1)
                        Dn1[f1] = v1;
                }
            }
        }
    }
}
#endif

// in-place SPD inverse of the n x n matrix (leading dimension ld) at arena + off of every slot (Gauss-Jordan,
// no pivoting needed); the column / row of the pivot are staged in the slot's T region
__device__ __forceinline__ void invert_spd_slots(const AsmArgs &args, int off, int n, int ld, int n_valid, int *bad)
{
    MFV2D_DYN_SMEM(double, smem);
    const AsmLayout &L = args.lay;
    const float inv_n = 1.0f / (float)n, inv_nn = 1.0f / (float)(n * n);
    for (int k = 0; k < n; ++k)
    {
        __syncthreads();
        for (int idx = threadIdx.x; idx < n_valid * n; idx += blockDim.x)
        {
            const int e = fast_div(idx, inv_n), i = idx - e * n;
            double *slot = smem + L.off_slots + e * L.slot_stride;
            const double *D = slot + L.slot_arena + off;
            double *colk = slot + L.slot_t, *rowk = colk + n;
            colk[i] = D[i * ld + k];
            rowk[i] = D[k * ld + i];
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < n_valid * n * n; idx += blockDim.x)
        {
            const int e = fast_div(idx, inv_nn), ij = idx - e * n * n;
            const int i = fast_div(ij, inv_n), j = ij - i * n;
            double *slot = smem + L.off_slots + e * L.slot_stride;
            double *D = slot + L.slot_arena + off;
            const double *colk = slot + L.slot_t, *rowk = colk + n;
            const double pv = colk[k];
            if (ij == 0 && !(fabs(pv) > 1e-12)) // decompose_pivoted_lu's tolerance (src/algebra/matrix.c:454-457)
                bad[e] = 1;
            const double ip = 1.0 / pv;
            double v;
            if (i == k)
                v = (j == k) ? ip : rowk[j] * ip;
            else if (j == k)
                v = -colk[i] * ip;
            else
                v = fma(-colk[i] * ip, rowk[j], D[i * ld + j]);
            D[i * ld + j] = v;
        }
    }
    __syncthreads();
}

// ---- emission ------------------------------------------------------------------------------------
struct EmitTerm
{
    const double *D; // dense matrix of the term in this slot
    int ld;
    double coef;
    const DevIncRow *ltab, *rtab; // incidence rows of the term (nullptr: none)
};

__device__ __forceinline__ void load_inc(const DevIncRow *row, int (&idx)[4], double (&sgn)[4])
{
    const int4 a = *reinterpret_cast<const int4 *>(row->e);
    const int4 b = *reinterpret_cast<const int4 *>(row->e + 4);
    idx[0] = a.x;
    idx[1] = a.z;
    idx[2] = b.x;
    idx[3] = b.z;
    sgn[0] = __hiloint2double(a.y, 0);
    sgn[1] = __hiloint2double(a.w, 0);
    sgn[2] = __hiloint2double(b.y, 0);
    sgn[3] = __hiloint2double(b.w, 0);
}

// one lane's share of a tile: column c, rows r0 + sub, r0 + sub + rstep, ... ; NL x NR gathers per entry.
// D rows and output rows are walked with running offsets (no per-row multiplications).
template <int NL, int NR>
__device__ __forceinline__ void emit_tile(const EmitTerm &t, int r0, int nr, int sub, int rstep, int c, double *__restrict__ o, int N)
{
    int rc[4];
    double rs[4];
    if (NR == 1)
    {
        rc[0] = c;
        rs[0] = t.coef;
    }
    else
    {
        load_inc(t.rtab + c, rc, rs);
#pragma unroll
        for (int b = 0; b < NR; ++b)
            rs[b] *= t.coef;
    }
    o += (size_t)sub * N;
    const size_t ostep = (size_t)rstep * N;
    if (NL == 1)
    {
        const double *Dr = t.D + (r0 + sub) * t.ld;
        const int dstep = rstep * t.ld;
#pragma unroll 4
        for (int r = sub; r < nr; r += rstep, Dr += dstep, o += ostep)
        {
            double v = rs[0] * Dr[rc[0]];
#pragma unroll
            for (int b = 1; b < NR; ++b)
                v = fma(rs[b], Dr[rc[b]], v);
            *o = v;
        }
    }
    else
    {
        const DevIncRow *lt = t.ltab + r0 + sub;
        const double *Dc = t.D + rc[0];
#pragma unroll 2
        for (int r = sub; r < nr; r += rstep, lt += rstep, o += ostep)
        {
            int li[4];
            double ls[4];
            load_inc(lt, li, ls);
            double v = 0.0;
#pragma unroll
            for (int a = 0; a < NL; ++a)
            {
                if (NR == 1)
                    v = fma(ls[a], Dc[li[a] * t.ld], v);
                else
                {
                    const double *Dr = t.D + li[a] * t.ld;
#pragma unroll
                    for (int b = 0; b < NR; ++b)
                        v = fma(ls[a] * rs[b], Dr[rc[b]], v);
                }
            }
            if (NR == 1)
                v *= rs[0];
            *o = v;
        }
    }
}

__device__ __forceinline__ int inc_width(int kind) // non-zeros per row of E (left) / per column of E (right, via kind ^ 2)
{
    return kind < 0 ? 1 : ((kind == MFV2D_INC_10 || kind == MFV2D_INC_21_T) ? 2 : 4);
}

__device__ __forceinline__ EmitTerm make_emit_term(const AsmArgs &args, const DevTerm &tm, const DevDense *dense,
                                                   const DevIncRow *inc, const double *arena)
{
    const DevDense &dd = dense[tm.dense];
    EmitTerm t;
    t.D = arena + dd.smem_off;
    t.ld = dd.ld;
    t.coef = tm.coef;
    t.ltab = tm.L >= 0 ? inc + args.lay.inc_row0[tm.L] : nullptr;
    t.rtab = tm.R >= 0 ? inc + args.lay.inc_row0[tm.R ^ 2] : nullptr; // column c of E_R = row c of its transpose
    return t;
}

__device__ __forceinline__ void emit_stage(const AsmArgs &args, const DevStage &sg, unsigned long long elem0, int n_valid,
                                           int *counter)
{
    MFV2D_DYN_SMEM(double, smem); // re-declared here so that the compiler keeps every access in the shared window
    const AsmLayout &L = args.lay;
    const char *sb = reinterpret_cast<const char *>(smem);
    const DevDense *dense = reinterpret_cast<const DevDense *>(sb + L.off_dense);
    const DevTerm *terms = reinterpret_cast<const DevTerm *>(sb + L.off_terms);
    const DevBlock *blocks = reinterpret_cast<const DevBlock *>(sb + L.off_blocks);
    const DevUnit *units = reinterpret_cast<const DevUnit *>(sb + L.off_units);
    const DevIncRow *inc = reinterpret_cast<const DevIncRow *>(sb + L.off_inc);
    const int N = args.n_total;
    const int n_units = sg.unit_end - sg.unit_begin;
    const float inv_valid = 1.0f / (float)n_valid;
    // units are sorted by decreasing cost; the warps fetch (unit, slot) pairs from a shared counter
    for (;;)
    {
        int w = 0;
        if (MFV2D_LANE0 == 0)
            w = atomicAdd(counter, 1);
        w = __shfl_sync(0xffffffffu, w, 0);
        if (w >= n_units * n_valid)
            break;
        const int u = fast_div(w, inv_valid), e = w - u * n_valid;
        const DevUnit un = units[sg.unit_begin + u];
        const DevBlock db = blocks[un.block];
        const double *arena = smem + L.off_slots + e * L.slot_stride + L.slot_arena;
        double *const out = args.out + (elem0 + e) * args.out_stride + (size_t)(db.row0 + un.r0) * N + db.col0;
        const int n_terms = db.term_end - db.term_begin;
        const int rstep = 32 >> un.cl;
        for (int lane = MFV2D_LANE0; lane < 32; lane += MFV2D_LANE_STEP)
        {
            const int sub = lane >> un.cl;
            const int c = un.c0 + (lane & ((1 << un.cl) - 1));
            const bool active = c < db.ncols;
            double *o = out + c;
            if (n_terms == 0)
            {
                if (active && !db.accumulate)
                {
                    double *oz = o + (size_t)sub * N;
                    const size_t ostep = (size_t)rstep * N;
#pragma unroll 4
                    for (int r = sub; r < un.nr; r += rstep, oz += ostep)
                        *oz = 0.0;
                }
                continue;
            }
            const DevTerm tm0 = terms[db.term_begin];
            if (n_terms == 1 && tm0.dense >= 0 && !db.accumulate)
            {
                const EmitTerm t = make_emit_term(args, tm0, dense, inc, arena);
                const int nl = inc_width(tm0.L), nr_ = inc_width(tm0.R < 0 ? -1 : (tm0.R ^ 2));
                if (!active)
                    continue; // nothing below is warp-collective
                switch (nl * 8 + nr_)
                    {
                    case 1 * 8 + 1:
                        emit_tile<1, 1>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    case 1 * 8 + 2:
                        emit_tile<1, 2>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    case 1 * 8 + 4:
                        emit_tile<1, 4>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    case 2 * 8 + 1:
                        emit_tile<2, 1>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    case 4 * 8 + 1:
                        emit_tile<4, 1>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    default:
                        emit_tile<4, 4>(t, un.r0, un.nr, sub, rstep, c, o, N);
                        break;
                    }
                continue;
            }
            // several terms (or an identity): accumulate the terms of an entry in a register
            for (int r = sub; r < un.nr; r += rstep)
            {
                double v = 0.0;
                for (int ti = 0; ti < n_terms; ++ti)
                {
                    const DevTerm tm = terms[db.term_begin + ti];
                    if (tm.dense < 0)
                    {
                        v += (un.r0 + r == c) ? tm.coef : 0.0;
                        continue;
                    }
                    const EmitTerm t = make_emit_term(args, tm, dense, inc, arena);
                    int li[4] = {un.r0 + r, 0, 0, 0}, rc[4] = {active ? c : 0, 0, 0, 0};
                    double ls[4] = {1.0, 0.0, 0.0, 0.0}, rs[4] = {1.0, 0.0, 0.0, 0.0};
                    int nl = 1, nr_ = 1;
                    if (t.ltab)
                    {
                        load_inc(t.ltab + un.r0 + r, li, ls);
                        nl = 4;
                    }
                    if (t.rtab)
                    {
                        load_inc(t.rtab + (active ? c : 0), rc, rs);
                        nr_ = 4;
                    }
                    double s = 0.0;
                    for (int a = 0; a < nl; ++a)
                        for (int b = 0; b < nr_; ++b)
                            s = fma(ls[a] * rs[b], t.D[li[a] * t.ld + rc[b]], s);
                    v = fma(tm.coef, s, v);
                }
                if (active)
                {
                    if (db.accumulate)
                        o[(size_t)r * N] += v;
                    else
                        o[(size_t)r * N] = v;
                }
            }
        }
    }
}

template <int Q> __global__ void __launch_bounds__(ASM_THREADS, 2) element_matrix_kernel(const AsmArgs args)
{
    MFV2D_DYN_SMEM(double, smem);
    const AsmLayout &L = args.lay;
    __shared__ int bad[8];
    __shared__ int emit_counter;
    {
        // descriptor blob -> shared memory, once per CTA
        const double2 *src = reinterpret_cast<const double2 *>(args.blob);
        double2 *dst = reinterpret_cast<double2 *>(smem);
        for (int i = threadIdx.x; i < (L.blob_bytes >> 4); i += blockDim.x)
            dst[i] = src[i];
    }
    const char *sb = reinterpret_cast<const char *>(smem);
    const DevStage *stages = reinterpret_cast<const DevStage *>(sb + L.off_stages);
    const DevDense *dense = reinterpret_cast<const DevDense *>(sb + L.off_dense);
    const DevW *ws = reinterpret_cast<const DevW *>(sb + L.off_ws);
    const DevTB *tbs = reinterpret_cast<const DevTB *>(sb + L.off_tbs);
    const DevRound *rounds = reinterpret_cast<const DevRound *>(sb + L.off_rounds);
    const double *tab = reinterpret_cast<const double *>(sb + L.off_tab);
    const double *pnts = tab + (2 * args.p + 1) * Q, *wgts = pnts + Q;
    const int epb = L.epb;
    const unsigned long long n_groups = (args.n_elem + epb - 1) / epb;

    for (unsigned long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x)
    {
        const unsigned long long elem0 = grp * epb;
        const int n_valid = (int)((args.n_elem - elem0) < (unsigned long long)epb ? (args.n_elem - elem0) : epb);
        __syncthreads(); // blob in place / the previous pass has finished reading shared memory
        for (int i = threadIdx.x; i < 8 * n_valid; i += blockDim.x)
            smem[L.off_cs + i] = args.corners[elem0 * 8 + i];
        for (int i = threadIdx.x; i < 8; i += blockDim.x)
            bad[i] = 0;
        for (int st = 0; st < L.n_stages; ++st)
        {
            const DevStage sg = stages[st];
            __syncthreads(); // corners ready / previous stage's emission done
            if (threadIdx.x == 0)
                emit_counter = 0;
            phase_weights<Q>(args, smem, sg, ws, pnts, wgts, elem0, n_valid);
#ifndef MFV2D_EMU
            if (L.use_tc)
            {
                __syncthreads();
                phase_tc<Q>(args, smem, sg, tbs, n_valid);
            }
            else
#endif
                for (int r = sg.round_begin; r < sg.round_end; ++r)
                {
                    const DevRound rd = rounds[r];
                    __syncthreads();
                    phase_s1<Q>(args, smem, rd, tbs, n_valid);
                    __syncthreads();
                    phase_s2<Q>(args, smem, rd, tbs, n_valid);
                }
            for (int d = sg.dense_begin; d < sg.dense_end; ++d)
                if (dense[d].inverse)
                    invert_spd_slots(args, dense[d].smem_off, dense[d].rows, dense[d].ld, n_valid, bad);
            __syncthreads();
            emit_stage(args, sg, elem0, n_valid, &emit_counter);
        }
        __syncthreads();
        if (args.status)
            for (int e = threadIdx.x; e < n_valid; e += blockDim.x)
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

// launch of the fused matrix kernel for a lowered plan; all pointers are device pointers
template <int Q> static int launch_q(mfv2d_ctx *ctx, const AsmArgs &a, size_t n_elem)
{
    const size_t smem = (size_t)a.lay.smem_bytes;
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_matrix_kernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_matrix_kernel<Q>, ASM_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    const unsigned long long n_groups = (n_elem + a.lay.epb - 1) / a.lay.epb;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    if (grid > n_groups)
        grid = n_groups;
    MFV2D_LAUNCH(element_matrix_kernel<Q>, (unsigned)grid, ASM_THREADS, smem, ctx->stream, a);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int launch_element_matrix(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, double *d_out,
                          int32_t *d_status)
{
    mfv2d_ctx *ctx = plan->ctx;
    if (e->n == 0)
        return MFV2D_SUCCESS;
    AsmArgs a{};
    a.blob = (const char *)plan->d_blob;
    a.corners = e->d_corners;
    for (int i = 0; i < 16; ++i)
        a.fields[i] = (fields && i < plan->n_fields) ? fields[i] : nullptr;
    a.out = d_out;
    a.status = d_status;
    a.n_elem = e->n;
    a.p = plan->p;
    a.q = plan->q;
    a.n_total = plan->out_ld ? plan->out_ld : plan->n_total;
    a.out_stride = plan->out_stride ? plan->out_stride : (unsigned long long)plan->n_total * plan->n_total;
    a.lay = plan->lay;
    switch (plan->q)
    {
#define MFV2D_Q_CASE(Q)                                                                                        \
    case Q:                                                                                                    \
        return launch_q<Q>(ctx, a, e->n);
        MFV2D_Q_CASE(2)
        MFV2D_Q_CASE(3)
        MFV2D_Q_CASE(4)
        MFV2D_Q_CASE(5)
        MFV2D_Q_CASE(6)
        MFV2D_Q_CASE(7)
        MFV2D_Q_CASE(8)
        MFV2D_Q_CASE(9)
        MFV2D_Q_CASE(10)
        MFV2D_Q_CASE(11)
        MFV2D_Q_CASE(12)
        MFV2D_Q_CASE(13)
        MFV2D_Q_CASE(14)
        MFV2D_Q_CASE(15)
        MFV2D_Q_CASE(16)
#undef MFV2D_Q_CASE
    default:
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "element matrices with %d quadrature points per direction (supported: 2..16)",
                          plan->q);
    }
}

extern "C" {

int mfv2d_cuda_jacobian_batch(const mfv2d_elements *e, double *out, int space) { return geometry_batch(e, out, space, 0); }

int mfv2d_cuda_coordinates_batch(const mfv2d_elements *e, double *out, int space)
{
    return geometry_batch(e, out, space, 1);
}

} // extern "C"
