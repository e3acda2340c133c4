// Device routines shared by the assembly (assembly.cu) and matrix-free apply (apply.cu) kernels:
// Jacobian, weight fields, sum-factorised dense mass / interior-product matrices in shared memory and
// the in-place SPD inverse.  All routines are CTA-cooperative: every thread of the CTA must call them.
#pragma once
#include "plan.cuh"
// T[a][jl] = sum_b B[j][b] * B2[l][b] * G[a][b];   jl = j * nB2 + l, row stride ldT
__device__ __forceinline__ void tm_stage1(const double *__restrict__ G, double *__restrict__ T, int ldT, int q,
                                          const double *__restrict__ B, int nB, const double *__restrict__ B2, int nB2)
{
    const int nJL = nB * nB2;
    for (int idx = threadIdx.x; idx < q * nJL; idx += blockDim.x)
    {
        const int a = idx / nJL;
        const int jl = idx - a * nJL;
        const int j = jl / nB2;
        const int l = jl - j * nB2;
        const double *bj = B + j * q, *bl = B2 + l * q, *g = G + a * q;
        double s = 0.0;
        for (int b = 0; b < q; ++b)
            s = fma(bj[b] * bl[b], g[b], s);
        T[a * ldT + jl] = s;
    }
}

// dst[(r0 + i*nB + j), (c0 + k*nB2 + l)] = sum_a A[i][a] * A2[k][a] * T[a][j*nB2 + l]
// trans: write dst[col][row] instead; mirror: also write the transposed entry (symmetric off-diagonal block)
__device__ __forceinline__ void tm_stage2(double *__restrict__ dst, int ld, int r0, int c0, bool trans, bool mirror,
                                          const double *__restrict__ T, int ldT, int q, const double *__restrict__ A,
                                          int nA, const double *__restrict__ A2, int nA2, int nB, int nB2)
{
    const int nJL = nB * nB2;
    const int ntile = (nJL + 3) >> 2;
    const int n_items = nA * nA2 * ntile;
    for (int item = threadIdx.x; item < n_items; item += blockDim.x)
    {
        const int ik = item / ntile;
        const int jl0 = (item - ik * ntile) << 2;
        const int i = ik / nA2;
        const int k = ik - i * nA2;
        const double *ai = A + i * q, *ak = A2 + k * q;
        double acc0 = 0.0, acc1 = 0.0, acc2 = 0.0, acc3 = 0.0;
        for (int a = 0; a < q; ++a)
        {
            const double pa = ai[a] * ak[a];
            const double2 t01 = *reinterpret_cast<const double2 *>(T + a * ldT + jl0);
            const double2 t23 = *reinterpret_cast<const double2 *>(T + a * ldT + jl0 + 2);
            acc0 = fma(pa, t01.x, acc0);
            acc1 = fma(pa, t01.y, acc1);
            acc2 = fma(pa, t23.x, acc2);
            acc3 = fma(pa, t23.y, acc3);
        }
        const double acc[4] = {acc0, acc1, acc2, acc3};
#pragma unroll
        for (int t = 0; t < 4; ++t)
        {
            const int jl = jl0 + t;
            if (jl < nJL)
            {
                const int j = jl / nB2;
                const int l = jl - j * nB2;
                const int row = r0 + i * nB + j;
                const int col = c0 + k * nB2 + l;
                if (trans)
                    dst[col * ld + row] = acc[t];
                else
                    dst[row * ld + col] = acc[t];
                if (mirror)
                    dst[col * ld + row] = acc[t];
            }
        }
    }
}

// weight fields at the quadrature points (geo = j00, j01, j10, j11, det each q*q); kinds W_* are in plan.cuh
__device__ __forceinline__ void fill_weight(double *__restrict__ G, const double *__restrict__ geo, const double *__restrict__ wg,
                                            int q, int kind, const double *__restrict__ fld)
{
    const int q2 = q * q;
    for (int idx = threadIdx.x; idx < q2; idx += blockDim.x)
    {
        const int a = idx / q, b = idx - a * q;
        const double w = wg[a] * wg[b];
        const double j00 = geo[idx], j01 = geo[q2 + idx], j10 = geo[2 * q2 + idx], j11 = geo[3 * q2 + idx],
                     det = geo[4 * q2 + idx];
        double v;
        switch (kind)
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
        case W_P1H:
            v = (fld[2 * idx] * j11 - fld[2 * idx + 1] * j10) * w;
            break;
        case W_P1V:
            v = -(fld[2 * idx + 1] * j00 - fld[2 * idx] * j01) * w;
            break;
        case W_P2H:
            v = -(fld[2 * idx] * j10 + fld[2 * idx + 1] * j11) / det * w;
            break;
        default: // W_P2V
            v = -(fld[2 * idx] * j00 + fld[2 * idx + 1] * j01) / det * w;
            break;
        }
        G[idx] = v;
    }
}

struct SmemView
{
    double *tab, *node, *edge, *pnts, *wgts, *geo, *G, *T, *arena;
    int ldT;
};

__device__ __forceinline__ void tensor_block(const SmemView &s, int q, int wkind, const double *fld, double *dst, int ld,
                                             int r0, int c0, bool trans, bool mirror, const double *A, int nA,
                                             const double *A2, int nA2, const double *B, int nB, const double *B2, int nB2)
{
    fill_weight(s.G, s.geo, s.wgts, q, wkind, fld);
    __syncthreads();
    tm_stage1(s.G, s.T, s.ldT, q, B, nB, B2, nB2);
    __syncthreads();
    tm_stage2(dst, ld, r0, c0, trans, mirror, s.T, s.ldT, q, A, nA, A2, nA2, nB, nB2);
}

// in-place inverse of a symmetric positive definite n x n matrix (Gauss-Jordan, no pivoting needed);
// scratch holds 2n doubles.  Returns (to every thread) whether a pivot vanished.
static __device__ int invert_spd_inplace(double *D, int n, double *scratch)
{
    double *colk = scratch, *rowk = scratch + n;
    __shared__ int bad;
    if (threadIdx.x == 0)
        bad = 0;
    for (int k = 0; k < n; ++k)
    {
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x)
        {
            colk[i] = D[i * n + k];
            rowk[i] = D[k * n + i];
        }
        __syncthreads();
        const double pv = colk[k];
        if (threadIdx.x == 0 && !(fabs(pv) > 1e-300))
            bad = 1;
        const double ip = 1.0 / pv;
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x)
        {
            const int i = idx / n, j = idx - i * n;
            double v;
            if (i == k)
                v = (j == k) ? ip : rowk[j] * ip;
            else if (j == k)
                v = -colk[i] * ip;
            else
                v = fma(-colk[i] * ip, rowk[j], D[idx]);
            D[idx] = v;
        }
    }
    __syncthreads();
    return bad;
}

static __device__ void compute_dense(const SmemView &s, const DevDense &dd, int p, int q, const double *const *fields,
                              unsigned long long elem, int *bad)
{
    double *dst = s.arena + dd.smem_off;
    const int np1 = p + 1, nh = p * (p + 1);
    const int q2 = q * q;
    switch (dd.dk)
    {
    case DK_MASS0:
        tensor_block(s, q, W_M0, nullptr, dst, np1 * np1, 0, 0, false, false, s.node, np1, s.node, np1, s.node, np1,
                     s.node, np1);
        break;
    case DK_MASS2:
        tensor_block(s, q, W_M2, nullptr, dst, p * p, 0, 0, false, false, s.edge, p, s.edge, p, s.edge, p, s.edge, p);
        break;
    case DK_MASS1: {
        const int n1 = 2 * nh;
        // hh: rows/cols (node_eta i, edge_xi j); vv: (edge_eta i, node_xi j); hv and its mirror
        tensor_block(s, q, W_HH, nullptr, dst, n1, 0, 0, false, false, s.node, np1, s.node, np1, s.edge, p, s.edge, p);
        tensor_block(s, q, W_VV, nullptr, dst, n1, nh, nh, false, false, s.edge, p, s.edge, p, s.node, np1, s.node, np1);
        tensor_block(s, q, W_HV, nullptr, dst, n1, 0, nh, false, true, s.node, np1, s.edge, p, s.edge, p, s.node, np1);
        break;
    }
    case DK_IPROD1: {
        // (n0 x n1): rows (node_eta, node_xi); cols h (node_eta, edge_xi), v (edge_eta, node_xi)
        const double *fld = fields[dd.field] + elem * (unsigned long long)(2 * q2);
        const int ld = dd.transpose ? np1 * np1 : 2 * nh;
        tensor_block(s, q, W_P1H, fld, dst, ld, 0, 0, dd.transpose, false, s.node, np1, s.node, np1, s.node, np1, s.edge, p);
        tensor_block(s, q, W_P1V, fld, dst, ld, 0, nh, dd.transpose, false, s.node, np1, s.edge, p, s.node, np1, s.node,
                     np1);
        break;
    }
    case DK_IPROD2: {
        // (n1 x n2): rows h (node_eta, edge_xi), v (edge_eta, node_xi); cols (edge_eta, edge_xi)
        const double *fld = fields[dd.field] + elem * (unsigned long long)(2 * q2);
        const int ld = dd.transpose ? 2 * nh : p * p;
        tensor_block(s, q, W_P2H, fld, dst, ld, 0, 0, dd.transpose, false, s.node, np1, s.edge, p, s.edge, p, s.edge, p);
        tensor_block(s, q, W_P2V, fld, dst, ld, nh, 0, dd.transpose, false, s.edge, p, s.edge, p, s.node, np1, s.edge, p);
        break;
    }
    default:
        break;
    }
    if (dd.inverse)
    {
        __syncthreads();
        if (invert_spd_inplace(dst, dd.rows, s.T))
            *bad = 1;
    }
}
// Jacobian of the bilinear map at the q x q points (src/fem_space/fem_space.c:39-53):
// geo = j00, j01, j10, j11, det, each q*q with index a*q + b (a: eta row, b: xi column)
__device__ __forceinline__ void fill_geometry(double *__restrict__ geo, const double *__restrict__ cs,
                                              const double *__restrict__ pnts, int q)
{
    const int q2 = q * q;
    for (int idx = threadIdx.x; idx < q2; idx += blockDim.x)
    {
        const int a = idx / q, b = idx - a * q;
        const double eta = pnts[a], xi = pnts[b];
        const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
        const double j00 = 0.25 * ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta));
        const double j01 = 0.25 * ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta));
        const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
        const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
        geo[idx] = j00;
        geo[q2 + idx] = j01;
        geo[2 * q2 + idx] = j10;
        geo[3 * q2 + idx] = j11;
        geo[4 * q2 + idx] = j00 * j11 - j10 * j01;
    }
}
