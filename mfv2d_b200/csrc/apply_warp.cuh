// Warp-autonomous matrix-free apply for mass-type tensor terms (included by apply.cu after FusedArgs).
//
// Same plan as element_vector_fused_kernel (the descriptor blob of plan_lower_vector_fused: row lists of the xi
// contractions, term groups of the eta contractions + metric, output contributions), different execution engine:
//
//   * a WARP owns a group of G = 32 / Q elements from load to store; the only synchronisation is __syncwarp() between
//     the four stages - no CTA barrier, no phase imbalance across 256 threads, warps of a CTA are independent;
//   * the element order P and the number of quadrature points Q are template parameters and every contraction loop is
//     fully unrolled, so the 1-D basis tables - which live in __constant__ memory - are consumed as IMMEDIATE constant-bank
//     operands of the DFMAs (SASS: DFMA R, R, c[0x3][imm], R): no table loads at all, where the CTA-phase kernel spends one
//     shared-memory wavefront per four DFMAs and is bound by exactly that pipe (ncu: l1tex data-pipe 65 %, fp64 27 %);
//   * stage A (xi interpolation) and stage C (xi projection): lane = (row, element), the row's coefficients in registers;
//     stage B (eta interpolation, metric, eta projection): lane = (element, quadrature column); the two layouts meet in
//     the warp's own shared-memory slots (one transpose each way).
//
// Terms: M0 / M1 / M2 (modes 0-2 of tile_op) with optional incidences on either side - every block of the mixed
// Poisson, Darcy, Stokes and direct Poisson systems.  Interior-product terms stay on the CTA-phase kernel.
#pragma once

#ifndef MFV2D_EMU

#define AW_THREADS 128
#define AW_MAXQ 12

// [kind][i * Q + b]: node (kind 0) and edge (kind 1) tables of the (P, Q) in use; points and weights
__constant__ double cw_tab[2][9 * AW_MAXQ];
__constant__ double cw_pnts[AW_MAXQ];
__constant__ double cw_wgts[AW_MAXQ];

template <int NB, int KIND, int Q>
__device__ __noinline__ void aw_row_a(double *__restrict__ slot, int src, int gi, int dst, const int4 *__restrict__ gtab)
{
    double c[NB];
    if (gi < 0)
    {
#pragma unroll
        for (int j = 0; j < NB; ++j)
            c[j] = slot[src + j];
    }
    else
    {
#pragma unroll
        for (int j = 0; j < NB; ++j)
        {
            const int4 m = gtab[gi + j]; // (p0 + p1) - (n0 + n1); unused entries point at the slot's zero cell
            c[j] = (slot[m.x] + slot[m.y]) - (slot[m.z] + slot[m.w]);
        }
    }
    double acc[Q];
#pragma unroll
    for (int b = 0; b < Q; ++b)
        acc[b] = c[0] * cw_tab[KIND][b];
#pragma unroll
    for (int j = 1; j < NB; ++j)
#pragma unroll
        for (int b = 0; b < Q; ++b)
            acc[b] = fma(c[j], cw_tab[KIND][j * Q + b], acc[b]);
#pragma unroll
    for (int b = 0; b < Q; ++b)
        slot[dst + b] = acc[b];
}

template <int NB, int KIND, int Q> __device__ __noinline__ void aw_row_c(double *__restrict__ slot, int src)
{
    double tr[Q];
#pragma unroll
    for (int b = 0; b < Q; ++b)
        tr[b] = slot[src + b];
#pragma unroll
    for (int j = 0; j < NB; ++j)
    {
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int b = 0; b + 1 < Q; b += 2)
        {
            a0 = fma(cw_tab[KIND][j * Q + b], tr[b], a0);
            a1 = fma(cw_tab[KIND][j * Q + b + 1], tr[b + 1], a1);
        }
        if (Q & 1)
            a0 = fma(cw_tab[KIND][j * Q + Q - 1], tr[Q - 1], a0);
        slot[src + j] = a0 + a1;
    }
}

// u[a] = sum_i t[i * LDT] T[i][a]  (eta interpolation of one component of one term on this lane's quadrature column)
template <int NA, int KIND, int Q, int LDT> __device__ __forceinline__ void aw_interp(const double *__restrict__ tc, double (&u)[Q])
{
    {
        const double c = tc[0];
#pragma unroll
        for (int a = 0; a < Q; ++a)
            u[a] = c * cw_tab[KIND][a];
    }
#pragma unroll
    for (int i = 1; i < NA; ++i)
    {
        const double c = tc[i * LDT];
#pragma unroll
        for (int a = 0; a < Q; ++a)
            u[a] = fma(c, cw_tab[KIND][i * Q + a], u[a]);
    }
}

// t[i * LDT] = sum_a T[i][a] u[a]
template <int NA, int KIND, int Q, int LDT> __device__ __forceinline__ void aw_project(double *__restrict__ tc, const double (&u)[Q])
{
#pragma unroll
    for (int i = 0; i < NA; ++i)
    {
        double a0 = 0.0, a1 = 0.0;
#pragma unroll
        for (int a = 0; a + 1 < Q; a += 2)
        {
            a0 = fma(cw_tab[KIND][i * Q + a], u[a], a0);
            a1 = fma(cw_tab[KIND][i * Q + a + 1], u[a + 1], a1);
        }
        if (Q & 1)
            a0 = fma(cw_tab[KIND][i * Q + Q - 1], u[Q - 1], a0);
        tc[i * LDT] = a0 + a1;
    }
}

template <int Q> struct AwLdt
{
    static constexpr int even = (Q + 1) & ~1;
    static constexpr int value = (((even >> 1) & 1) == 0) ? even + 2 : even; // asm_ldq(Q)
};

// one term (NT = 1) or two terms with the same dense factor (NT = 2) on quadrature column b of one element
template <int P, int Q, int MODE, int NT>
__device__ __noinline__ void aw_term_b(double *__restrict__ slot, const int *__restrict__ tb, int b, const ColumnGeo &g)
{
    constexpr int LDT = AwLdt<Q>::value;
    double *tc[NT];
    tc[0] = slot + tb[2] + b;
    if (NT > 1)
        tc[NT - 1] = slot + tb[11] + b;
    if (MODE == 2)
    {
        // M1: component 0 = h edges (eta table: node, P + 1 rows), component 1 = v edges (eta table: edge, P rows)
#pragma unroll
        for (int n = 0; n < NT; ++n)
        {
            double u0[Q], u1[Q];
            aw_interp<P + 1, TK_NODE, Q, LDT>(tc[n], u0);
            aw_interp<P, TK_EDGE, Q, LDT>(tc[n] + (P + 1) * LDT, u1);
#pragma unroll
            for (int a = 0; a < Q; ++a)
            {
                const double eta = cw_pnts[a];
                const double j00 = 0.25 * (g.dx10 * (1 - eta) + g.dx23 * (1 + eta));
                const double j01 = 0.25 * (g.dy10 * (1 - eta) + g.dy23 * (1 + eta));
                const double det = j00 * g.j11 - g.j10 * j01;
                const double wod = cw_wgts[a] * g.wb * fast_rcp(det);
                const double khh = (g.j11 * g.j11 + g.j10 * g.j10) * wod, kvv = (j00 * j00 + j01 * j01) * wod,
                             khv = (j01 * g.j11 + j00 * g.j10) * wod;
                const double a0 = u0[a], a1 = u1[a];
                u0[a] = khh * a0 + khv * a1;
                u1[a] = khv * a0 + kvv * a1;
            }
            aw_project<P + 1, TK_NODE, Q, LDT>(tc[n], u0);
            aw_project<P, TK_EDGE, Q, LDT>(tc[n] + (P + 1) * LDT, u1);
        }
    }
    else
    {
        constexpr int NA = MODE == 0 ? P + 1 : P;
        constexpr int KIND = MODE == 0 ? TK_NODE : TK_EDGE;
        double k[Q];
#pragma unroll
        for (int a = 0; a < Q; ++a)
        {
            const double eta = cw_pnts[a];
            const double j00 = 0.25 * (g.dx10 * (1 - eta) + g.dx23 * (1 + eta));
            const double j01 = 0.25 * (g.dy10 * (1 - eta) + g.dy23 * (1 + eta));
            const double det = j00 * g.j11 - g.j10 * j01;
            k[a] = MODE == 0 ? det * (cw_wgts[a] * g.wb) : cw_wgts[a] * g.wb * fast_rcp(det);
        }
#pragma unroll
        for (int n = 0; n < NT; ++n)
        {
            double u0[Q];
            aw_interp<NA, KIND, Q, LDT>(tc[n], u0);
#pragma unroll
            for (int a = 0; a < Q; ++a)
                u0[a] *= k[a];
            aw_project<NA, KIND, Q, LDT>(tc[n], u0);
        }
    }
}

template <int P, int Q> __global__ void __launch_bounds__(AW_THREADS, 3) element_vector_warp_kernel(const FusedArgs fa)
{
    MFV2D_DYN_SMEM(double, smem);
    const ApplyArgs &args = fa.a;
    const FusedVecInfo &fv = fa.fv;
    constexpr int G = 32 / Q;
    const int N = args.n_total, SD = fv.slot_doubles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    // CTA-shared: coefficients | descriptor blob ; then per warp G element slots
    double *coef = smem;
    int *blob = reinterpret_cast<int *>(coef + ((fa.n_coef + 1) & ~1));
    double *slots = reinterpret_cast<double *>(blob + fv.blob_ints) + (size_t)warp * G * SD;
    for (int i = threadIdx.x; i < fa.n_coef; i += blockDim.x)
        coef[i] = args.coef[i];
    for (int i = threadIdx.x; i < fv.blob_ints; i += blockDim.x)
        blob[i] = fa.blob[i];
    for (int g = lane; g < G; g += 32)
        slots[g * SD + fv.off_vec + N] = 0.0; // the zero cell of every slot (never written again)
    __syncthreads();
    const int4 *gtab = reinterpret_cast<const int4 *>(blob + fv.off_pre);
    const int4 *rowA = reinterpret_cast<const int4 *>(blob + fv.off_rowA);
    const int *termB = blob + fv.off_termB;
    const int4 *rowC = reinterpret_cast<const int4 *>(blob + fv.off_rowC);
    const int2 *yrow = reinterpret_cast<const int2 *>(blob + fv.off_yrow);
    const int4 *contrib = reinterpret_cast<const int4 *>(blob + fv.off_contrib);
    const unsigned long long n_groups = (args.n_elem + G - 1) / G;
    const unsigned long long stride = (unsigned long long)gridDim.x * nwarps;
    for (unsigned long long grp = (unsigned long long)blockIdx.x * nwarps + warp; grp < n_groups; grp += stride)
    {
        const unsigned long long elem0 = grp * G;
        const int nv = (int)((args.n_elem - elem0) < (unsigned long long)G ? (args.n_elem - elem0) : G);
        __syncwarp();
        // ---- load: x and the corners of the group's elements (coalesced over the warp) ----
        for (int i = lane; i < nv * N; i += 32)
        {
            const int g = i / N, k = i - g * N;
            slots[g * SD + fv.off_vec + k] = args.x[elem0 * N + i];
        }
        for (int i = lane; i < nv * 8; i += 32)
            slots[(i >> 3) * SD + (i & 7)] = args.corners[elem0 * 8 + i];
        __syncwarp();
        // ---- stage A: t[row][:] = sum_j c[j] B_j[:]; lane = (row, element) ----
        {
            const int n_items = fv.n_rowA * 2 * G;
            for (int it = lane; it < n_items; it += 32)
            {
                const int g = it % G, ew = it / G, w = ew & 1, e = ew >> 1;
                const int4 d0 = rowA[2 * e], d1 = rowA[2 * e + 1];
                const int src = w ? d1.x : d0.x;
                if (src < 0 || g >= nv)
                    continue;
                double *slot = slots + g * SD;
                const int dst = w ? d1.y : d0.w, gi = w ? d1.w : d1.z;
                if (d0.z == TK_NODE)
                    aw_row_a<P + 1, TK_NODE, Q>(slot, src, gi, dst, gtab);
                else
                    aw_row_a<P, TK_EDGE, Q>(slot, src, gi, dst, gtab);
            }
        }
        __syncwarp();
        // ---- stage B: lane = (element, quadrature column): eta interpolation, metric, eta projection of every term ----
        if (lane < nv * Q)
        {
            const int g = lane / Q, b = lane - g * Q;
            double *slot = slots + g * SD;
            const double x0 = slot[0], y0 = slot[1], x1 = slot[2], y1 = slot[3], x2 = slot[4], y2 = slot[5], x3 = slot[6], y3 = slot[7];
            const double xi = cw_pnts[b];
            ColumnGeo cg;
            cg.wb = cw_wgts[b];
            cg.j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
            cg.j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
            cg.dx10 = x1 - x0;
            cg.dx23 = x2 - x3;
            cg.dy10 = y1 - y0;
            cg.dy23 = y2 - y3;
            for (int t = 0; t < fv.n_terms; ++t)
            {
                const int *tb = termB + 12 * t;
                const int mode = tb[0];
                const bool two = tb[11] >= 0;
                if (mode == 0)
                {
                    if (two)
                        aw_term_b<P, Q, 0, 2>(slot, tb, b, cg);
                    else
                        aw_term_b<P, Q, 0, 1>(slot, tb, b, cg);
                }
                else if (mode == 1)
                {
                    if (two)
                        aw_term_b<P, Q, 1, 2>(slot, tb, b, cg);
                    else
                        aw_term_b<P, Q, 1, 1>(slot, tb, b, cg);
                }
                else
                {
                    if (two)
                        aw_term_b<P, Q, 2, 2>(slot, tb, b, cg);
                    else
                        aw_term_b<P, Q, 2, 1>(slot, tb, b, cg);
                }
            }
        }
        __syncwarp();
        // ---- stage C: out[row][j] = B_j[:] . t'[row][:], written over the head of the row; lane = (row, element) ----
        {
            const int n_items = fv.n_rowC * 2 * G;
            for (int it = lane; it < n_items; it += 32)
            {
                const int g = it % G, ew = it / G, w = ew & 1, e = ew >> 1;
                const int4 d0 = rowC[2 * e], d1 = rowC[2 * e + 1];
                const int src = w ? d1.x : d0.x;
                if (src < 0 || g >= nv)
                    continue;
                double *slot = slots + g * SD;
                if (d0.z == TK_NODE)
                    aw_row_c<P + 1, TK_NODE, Q>(slot, src);
                else
                    aw_row_c<P, TK_EDGE, Q>(slot, src);
            }
        }
        __syncwarp();
        // ---- post: y[k] = sum over the terms of coef * (signed sum of up to four outputs), coalesced store ----
        for (int i = lane; i < nv * N; i += 32)
        {
            const int g = i / N, k = i - g * N;
            const double *slot = slots + g * SD;
            const int2 yr = yrow[k];
            double acc = 0.0;
            for (int c = yr.x; c < yr.x + yr.y; ++c)
            {
                const int4 c0 = contrib[c];
                double v = slot[c0.y & 0xffff];
                if (c0.w) // left incidence: (p0 + p1) - (n0 + n1)
                    v = (v + slot[(unsigned)c0.y >> 16]) - (slot[c0.z & 0xffff] + slot[(unsigned)c0.z >> 16]);
                acc = fma(coef[c0.x], v, acc);
            }
            args.y[elem0 * N + i] = acc;
        }
    }
}

template <int P, int Q> static int launch_warp_pq(mfv2d_ctx *ctx, const FusedArgs &fa, size_t n_elem)
{
    constexpr int G = 32 / Q;
    const int nwarps = AW_THREADS / 32;
    const size_t fixed = (((size_t)fa.n_coef + 1) & ~(size_t)1) * 8 + (size_t)fa.fv.blob_ints * 4;
    const size_t smem = fixed + (size_t)nwarps * G * fa.fv.slot_doubles * 8;
    if (smem > ctx->smem_optin)
        return -1;
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_vector_warp_kernel<P, Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_vector_warp_kernel<P, Q>, AW_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    const unsigned long long need = (n_elem + (size_t)G * nwarps - 1) / ((size_t)G * nwarps);
    if (grid > need)
        grid = need;
    element_vector_warp_kernel<P, Q><<<(unsigned)grid, AW_THREADS, smem, ctx->stream>>>(fa);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

// -1: not applicable (the caller goes on to the CTA-phase kernel)
static int launch_warp(mfv2d_ctx *ctx, const mfv2d_plan *plan, const FusedArgs &fa, size_t n_elem)
{
    const int p = plan->p, q = plan->q;
    if (q > AW_MAXQ || p > 8 || !plan->warp_ok)
        return -1;
    // the tables of this basis as constant-bank data (stream ordered; re-uploaded when another basis is used)
    static const mfv2d_basis *uploaded = nullptr;
    if (uploaded != plan->basis)
    {
        double tab[2][9 * AW_MAXQ] = {};
        for (int i = 0; i <= p; ++i)
            for (int b = 0; b < q; ++b)
                tab[0][i * q + b] = plan->basis->node[i * q + b];
        for (int i = 0; i < p; ++i)
            for (int b = 0; b < q; ++b)
                tab[1][i * q + b] = plan->basis->edge[i * q + b];
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyToSymbolAsync(cw_tab, tab, sizeof tab, 0, cudaMemcpyHostToDevice, ctx->stream));
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyToSymbolAsync(cw_pnts, plan->basis->pnts.data(), q * sizeof(double), 0, cudaMemcpyHostToDevice, ctx->stream));
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyToSymbolAsync(cw_wgts, plan->basis->wgts.data(), q * sizeof(double), 0, cudaMemcpyHostToDevice, ctx->stream));
        MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream)); // `tab` is a stack buffer
        uploaded = plan->basis;
    }
#define AW_CASE(Pv, Qv)                                                                                        \
    if (p == Pv && q == Qv)                                                                                    \
        return launch_warp_pq<Pv, Qv>(ctx, fa, n_elem);
    // integration order p + 3 (the reference's default, SystemSettings.over_integration_order) and p + 2 (FemCache(2) of its tests)
    AW_CASE(2, 6) AW_CASE(3, 7) AW_CASE(4, 8) AW_CASE(5, 9) AW_CASE(6, 10) AW_CASE(7, 11) AW_CASE(8, 12)
    AW_CASE(2, 5) AW_CASE(3, 6) AW_CASE(4, 7) AW_CASE(5, 8) AW_CASE(6, 9)
#undef AW_CASE
    return -1;
}

#endif // MFV2D_EMU
