// Element-block linear system on the device (LinearSystem / DenseVector / TraceVector of
// src/algebra/sparse_system.{h,c} and system_objects.c) and the Krylov drivers of
// python/mfv2d/solving.py running without per-operation host round trips.
//
// Layout (all device resident, SoA):
//   A     concatenated dense element blocks, row-major           (apply_diagonal: HBM read bound)
//   Ainv  their explicit inverses (partially pivoted Gauss-Jordan, dense.cu)
//   N     the constraint matrix in CSR over global dense columns, and N^T in CSR: apply_trace and
//         apply_trace_transpose are deterministic sparse mat-vecs (no atomics).  Trace vectors are kept
//         merged - one value per multiplier - instead of one copy per sharing element.
//   S_e   per-element condensed blocks N_e A_e^-1 N_e^T (column-major, m_e x m_e) with the gather list
//         of the multipliers an element touches and, per multiplier, the list of (element, local row)
//         slots that contribute to it (the reference's trace_offsets/trace_values, sparse_system.h:27-32):
//         the Schur operator costs sum m_e^2 reads per application instead of sum N_e^2.
#include <algorithm>
#include <chrono>
#include <cmath>

#include "system.cuh"

int launch_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *d_in, double *d_out, int32_t *d_status);

// ---- kernels -------------------------------------------------------------------------------------
// y_e = M_e x_e for dense row-major blocks (LinearSystem.apply_diagonal / apply_diagonal_inverse with the stored
// inverse, src/algebra/sparse_system.c:3-77).  HBM-read bound: 8 N_e^2 + 16 N_e bytes per element.
// One CTA per element; a quarter warp (8 lanes) owns a row and reads it with 16-byte loads (VEC: even n and a
// 16-byte aligned block) - four rows per warp and up to four loads per lane are in flight before the
// 3-step shuffle reduction, which is what keeps enough bytes in flight per SM.
template <bool VEC>
__device__ __forceinline__ void block_gemv_rows(const double *__restrict__ A, const double *__restrict__ xs, double *__restrict__ y,
                                                int n)
{
#ifndef MFV2D_EMU
    const int lane = threadIdx.x & 31, sub = lane >> 3, l8 = lane & 7;
    // eight rows per trip (two independent quarter-warp rows per lane): ~2 n / 16 16-byte loads in flight per lane
    for (int r0 = 0; r0 < n; r0 += 8)
    {
        const int ra = r0 + sub, rb = r0 + 4 + sub;
        double a0 = 0.0, a1 = 0.0, b0 = 0.0, b1 = 0.0;
        const double *rowa = A + (size_t)(ra < n ? ra : 0) * n, *rowb = A + (size_t)(rb < n ? rb : 0) * n;
        if (VEC)
        {
#pragma unroll 4
            for (int c = 2 * l8; c < n; c += 16)
            {
                const double2 va = __ldcs(reinterpret_cast<const double2 *>(rowa + c));
                const double2 vb = __ldcs(reinterpret_cast<const double2 *>(rowb + c));
                const double2 x = *reinterpret_cast<const double2 *>(xs + c);
                a0 = fma(va.x, x.x, a0);
                a1 = fma(va.y, x.y, a1);
                b0 = fma(vb.x, x.x, b0);
                b1 = fma(vb.y, x.y, b1);
            }
        }
        else
        {
#pragma unroll 4
            for (int c = l8; c < n; c += 8)
            {
                const double x = xs[c];
                a0 = fma(__ldcs(rowa + c), x, a0);
                b0 = fma(__ldcs(rowb + c), x, b0);
            }
        }
        double acc = a0 + a1, bcc = b0 + b1;
        acc += __shfl_xor_sync(0xffffffffu, acc, 4);
        bcc += __shfl_xor_sync(0xffffffffu, bcc, 4);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        bcc += __shfl_xor_sync(0xffffffffu, bcc, 2);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        bcc += __shfl_xor_sync(0xffffffffu, bcc, 1);
        if (l8 == 0 && ra < n)
            y[ra] = acc;
        if (l8 == 0 && rb < n)
            y[rb] = bcc;
    }
#else
    for (int r = 0; r < n; ++r)
    {
        double acc = 0.0;
        for (int c = 0; c < n; ++c)
            acc = fma(A[(size_t)r * n + c], xs[c], acc);
        y[r] = acc;
    }
#endif
}

// one WARP per element block, no CTA-wide barrier: the warp stages x_e in its own strip of shared memory and
// streams the rows of A_e; the other warps of the SM cover its latencies
__global__ void __launch_bounds__(SYS_THREADS) block_gemv_kernel(unsigned long long n_elem, const int64_t *__restrict__ boff,
                                                                const int64_t *__restrict__ moff,
                                                                const double *__restrict__ M, const double *__restrict__ x,
                                                                double *__restrict__ y, int xs_stride)
{
    MFV2D_DYN_SMEM(double, xs_all);
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    double *xs = xs_all + (size_t)warp * xs_stride;
    for (unsigned long long e = (unsigned long long)blockIdx.x * nwarps + warp; e < n_elem; e += (unsigned long long)gridDim.x * nwarps)
    {
        const int64_t o = boff[e];
        const int n = (int)(boff[e + 1] - o);
        const double *A = M + moff[e];
#ifndef MFV2D_EMU
        __syncwarp();
#endif
        for (int i = MFV2D_LANE0; i < n; i += MFV2D_LANE_STEP)
            xs[i] = x[o + i];
#ifndef MFV2D_EMU
        __syncwarp();
#endif
        if ((n & 1) == 0 && (reinterpret_cast<size_t>(A) & 15) == 0)
            block_gemv_rows<true>(A, xs, y + o, n);
        else
            block_gemv_rows<false>(A, xs, y + o, n);
    }
}

// y[i] = sum_k val[k] x[col[k]] over CSR row i
__global__ void csr_spmv_kernel(long long n_rows, const int64_t *__restrict__ ptr, const int64_t *__restrict__ col,
                                const double *__restrict__ val, const double *__restrict__ x, double *__restrict__ y)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_rows; i += (long long)gridDim.x * blockDim.x)
    {
        double acc = 0.0;
        for (int64_t k = ptr[i]; k < ptr[i + 1]; ++k)
            acc = fma(val[k], x[col[k]], acc);
        y[i] = acc;
    }
}

// S_e[r][s] = sum_{(c,v) in row r} sum_{(d,w) in row s} v w Ainv_e[c][d]
__device__ __forceinline__ double schur_entry(const double *__restrict__ B, int n, const int64_t *__restrict__ lrow_ptr,
                                              const int32_t *__restrict__ lrow_col, const double *__restrict__ lrow_val, int64_t ra,
                                              int64_t rb)
{
    double acc = 0.0;
    for (int64_t a = lrow_ptr[ra]; a < lrow_ptr[ra + 1]; ++a)
    {
        const double *brow = B + (size_t)lrow_col[a] * n;
        double t = 0.0;
        for (int64_t b = lrow_ptr[rb]; b < lrow_ptr[rb + 1]; ++b)
            t = fma(lrow_val[b], brow[lrow_col[b]], t);
        acc = fma(lrow_val[a], t, acc);
    }
    return acc;
}

// full storage: column-major m x m (entry idx = s * m + r).  Packed storage (symmetric systems): entry (k, r) at k * m + r holds
// (S[r][c] + S[c][r]) / 2 with c = (r + k) % m - see packed_size() in system.cuh
__global__ void schur_build_kernel(unsigned long long n_elem, const int64_t *__restrict__ boff, const int64_t *__restrict__ moff,
                                   const double *__restrict__ Ainv, const int64_t *__restrict__ slot_off,
                                   const int64_t *__restrict__ lrow_ptr, const int32_t *__restrict__ lrow_col,
                                   const double *__restrict__ lrow_val, const int64_t *__restrict__ s_off, double *__restrict__ S,
                                   int packed)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int n = (int)(boff[e + 1] - boff[e]);
        const double *B = Ainv + moff[e];
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        double *Se = S + s_off[e];
        if (packed)
        {
            const int total = m * (m + 1) / 2;
            for (int idx = threadIdx.x; idx < total; idx += blockDim.x)
            {
                const int k = idx / m, r = idx - k * m;
                int c = r + k;
                if (c >= m)
                    c -= m;
                const double v = schur_entry(B, n, lrow_ptr, lrow_col, lrow_val, s0 + r, s0 + c);
                Se[idx] = k == 0 ? v : 0.5 * (v + schur_entry(B, n, lrow_ptr, lrow_col, lrow_val, s0 + c, s0 + r));
            }
            continue;
        }
        for (int idx = threadIdx.x; idx < m * m; idx += blockDim.x)
        {
            const int s = idx / m, r = idx - s * m; // column s, row r (column-major storage)
            Se[idx] = schur_entry(B, n, lrow_ptr, lrow_col, lrow_val, s0 + r, s0 + s);
        }
    }
}

// largest |A_e[i][j] - A_e[j][i]| against 1e-12 max|A_e|: flag[0] = 1 if some block is not symmetric to rounding
__global__ void block_asymmetry_kernel(unsigned long long n_elem, const int64_t *__restrict__ boff, const int64_t *__restrict__ moff,
                                       const double *__restrict__ A, int *__restrict__ flag)
{
    __shared__ double red_d[32], red_m[32];
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int n = (int)(boff[e + 1] - boff[e]);
        const double *M = A + moff[e];
        double dmax = 0.0, amax = 0.0;
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x)
        {
            const int i = idx / n, j = idx - i * n;
            const double v = M[idx];
            amax = fmax(amax, fabs(v));
            if (j > i)
                dmax = fmax(dmax, fabs(v - M[(size_t)j * n + i]));
        }
#ifndef MFV2D_EMU
        for (int o = 16; o > 0; o >>= 1)
        {
            dmax = fmax(dmax, __shfl_xor_sync(0xffffffffu, dmax, o));
            amax = fmax(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        }
#endif
        const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
        if ((threadIdx.x & 31) == 0)
        {
            red_d[warp] = dmax;
            red_m[warp] = amax;
        }
        __syncthreads();
        if (threadIdx.x == 0)
        {
            for (int w = 1; w < nwarps; ++w)
            {
                dmax = fmax(dmax, red_d[w]);
                amax = fmax(amax, red_m[w]);
            }
            if (!(dmax <= 1e-12 * amax))
                atomicExch(flag, 1);
        }
        __syncthreads();
    }
}

// slot_val[s0 + r] = sum_s S_e[r][s] t[trace(s0 + s)], full column-major blocks - one warp per element
__global__ void __launch_bounds__(SYS_THREADS) schur_local_kernel(unsigned long long n_elem, const int64_t *__restrict__ slot_off,
                                                                 const int64_t *__restrict__ slot_trace,
                                                                 const int64_t *__restrict__ s_off, const double *__restrict__ S,
                                                                 const double *__restrict__ t, double *__restrict__ slot_val,
                                                                 int max_m, const int *__restrict__ done)
{
    MFV2D_DYN_SMEM(double, lam_all);
    if (done && *done)
        return;
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    double *lam = lam_all + (size_t)warp * (max_m + 32);
    double *part = lam + max_m; // 32 doubles: partial sums of the column groups
    for (unsigned long long e = (unsigned long long)blockIdx.x * nwarps + warp; e < n_elem; e += (unsigned long long)gridDim.x * nwarps)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        if (m == 0)
            continue;
        const double *Se = S + s_off[e];
        __syncwarp();
        for (int i = MFV2D_LANE0; i < m; i += MFV2D_LANE_STEP)
            lam[i] = t[slot_trace[s0 + i]];
        __syncwarp();
#ifndef MFV2D_EMU
        if (m <= 16)
        {
            // 32 / m column groups work side by side on the (coalesced) columns of S_e, then fold
            const int lane = threadIdx.x & 31;
            const int G = 32 / m, g = lane / m, r = lane - g * m;
            double acc = 0.0;
            if (g < G)
            {
#pragma unroll 4
                for (int s = g; s < m; s += G)
                    acc = fma(__ldcs(Se + (size_t)s * m + r), lam[s], acc);
            }
            __syncwarp();
            part[lane] = acc;
            __syncwarp();
            if (lane < m)
            {
                double v = 0.0;
                for (int k = 0; k < G; ++k)
                    v += part[k * m + lane];
                slot_val[s0 + lane] = v;
            }
            continue;
        }
#endif
        for (int r = MFV2D_LANE0; r < m; r += MFV2D_LANE_STEP)
        {
            double acc = 0.0;
#pragma unroll 4
            for (int s = 0; s < m; ++s)
                acc = fma(Se[(size_t)s * m + r], lam[s], acc);
            slot_val[s0 + r] = acc;
        }
    }
}

// the same product from PACKED symmetric blocks, any m: a warp per element, lane r gathers row r from the stored half
// (direct entries S[r][(r + k) % m] and mirrored entries S[(r - k) % m][r]); used when a block has more than 32 rows
__global__ void __launch_bounds__(SYS_THREADS) schur_local_packed_generic_kernel(
    unsigned long long n_elem, const int64_t *__restrict__ slot_off, const int64_t *__restrict__ slot_trace,
    const int64_t *__restrict__ s_off, const double *__restrict__ S, const double *__restrict__ t, double *__restrict__ slot_val, int max_m,
    const int *__restrict__ done)
{
    MFV2D_DYN_SMEM(double, lam_all);
    if (done && *done)
        return;
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    double *lam = lam_all + (size_t)warp * max_m;
    for (unsigned long long e = (unsigned long long)blockIdx.x * nwarps + warp; e < n_elem; e += (unsigned long long)gridDim.x * nwarps)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        if (m == 0)
            continue;
        const double *Se = S + s_off[e];
        const int half = m >> 1, kfull = (m + 1) >> 1;
        const bool even = (m & 1) == 0;
        __syncwarp();
        for (int i = MFV2D_LANE0; i < m; i += MFV2D_LANE_STEP)
            lam[i] = t[slot_trace[s0 + i]];
        __syncwarp();
        for (int r = MFV2D_LANE0; r < m; r += MFV2D_LANE_STEP)
        {
            double acc = 0.0;
            for (int k = 0; k <= half; ++k)
            {
                int c = r + k, sr = r - k;
                if (c >= m)
                    c -= m;
                if (sr < 0)
                    sr += m;
                if (k < kfull || (even && r < half))
                    acc = fma(Se[(size_t)k * m + r], lam[c], acc);
                if (k > 0 && (k < kfull || (even && sr < half)))
                    acc = fma(Se[(size_t)k * m + sr], lam[sr], acc);
            }
            slot_val[s0 + r] = acc;
        }
    }
}

#ifndef MFV2D_EMU
// row r of a packed block times lambda for a block of ANY size m <= W (boundary elements, mixed warps): per-entry index arithmetic
__device__ __forceinline__ double packed_row_general(const double *__restrict__ Se, int m, int W, int g, int r, int lane, double lam_r)
{
    constexpr int UNR = 4;
    const bool row = r < m;
    const int half = m >> 1, kfull = (m + 1) >> 1;
    const int kmax = (W >> 1) + 1;
    const bool low = !(m & 1) && r < half; // holds an entry of the last (half-length) diagonal of an even block
    double acc = 0.0;
    for (int k0 = 0; k0 < kmax; k0 += UNR)
    {
        double a[UNR];
#pragma unroll
        for (int u = 0; u < UNR; ++u)
        {
            const int k = k0 + u;
            a[u] = (row && (k < kfull || (k == half && low))) ? __ldcs(Se + (size_t)k * m + r) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < UNR; ++u)
        {
            const int k = k0 + u;
            int lc = lane, ls = lane;
            if (row && k <= half)
            {
                int c = r + k, sr = r - k;
                if (c >= m)
                    c -= m;
                if (sr < 0)
                    sr += m;
                lc = g * W + c;
                ls = g * W + sr;
            }
            const double lam_c = __shfl_sync(0xffffffffu, lam_r, lc);
            const double recv = __shfl_sync(0xffffffffu, a[u] * lam_r, ls);
            acc = fma(a[u], lam_c, acc);
            if (k > 0)
                acc += recv;
        }
    }
    return acc;
}

// PACKED symmetric blocks of at most W <= 32 rows: 32 / W elements per warp, lane (g, r) owns row r of its element.  Every stored
// entry is read ONCE (coalesced over r for a fixed diagonal k) and ALL diagonals of a block are requested before the first is
// used (W is a template parameter: the walk is fully unrolled, one memory latency per element).  The mirrored use of an entry and
// the lambda values travel between the lanes by shuffles with ONE fixed source lane (r - 1 mod W): walking the diagonals from the
// last to the first, T <- rotate(T + a_k lambda_r) delivers a_k lambda_r to lane r + k after k rotations, and L <- rotate(L)
// turns lambda[r + k] into lambda[r + k - 1].  No shared memory, no barrier, no per-entry index arithmetic, half the bytes
// of the full product.  The 16-byte element descriptor of the NEXT element is requested before the current block is streamed.
template <int W>
__global__ void __launch_bounds__(SYS_THREADS, 4) schur_local_packed_kernel(unsigned long long n_elem, const SchurElemDesc *__restrict__ desc,
                                                                           const int32_t *__restrict__ slot_trace32,
                                                                           const double *__restrict__ S, const double *__restrict__ t,
                                                                           double *__restrict__ slot_val, const int *__restrict__ done)
{
    if (done && *done)
        return;
    constexpr int G = 32 / W, KF = (W + 1) / 2, K = W / 2; // full diagonals 0 .. KF-1; an even block adds the half diagonal K
    constexpr bool EVEN = (W & 1) == 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int g = lane / W, r = lane - g * W;
    const unsigned long long stride = (unsigned long long)gridDim.x * nwarps * G;
    unsigned long long base = ((unsigned long long)blockIdx.x * nwarps + warp) * G;
    if (base >= n_elem)
        return;
    auto load_desc = [&](unsigned long long b) {
        SchurElemDesc d{0, 0, 0};
        const unsigned long long e = b + g;
        if (g < G && e < n_elem)
            d = desc[e];
        return d;
    };
    SchurElemDesc cur = load_desc(base);
    double lam_r = r < cur.m ? t[slot_trace32[cur.s0 + r]] : 0.0;
    for (; base < n_elem; base += stride)
    {
        const bool more = base + stride < n_elem;
        SchurElemDesc nxt{0, 0, 0};
        if (more)
            nxt = load_desc(base + stride);
        const int m = cur.m;
        const double *Se = S + cur.s_off;
        const bool act = g < G && base + g < n_elem;
        double acc;
        if (__all_sync(0xffffffffu, !act || m == W))
        {
            double a[KF], ah = 0.0;
#pragma unroll
            for (int k = 0; k < KF; ++k)
                a[k] = act ? __ldcs(Se + k * W + r) : 0.0;
            if (EVEN)
                ah = (act && r < K) ? __ldcs(Se + K * W + r) : 0.0;
            const int src = act ? g * W + (r == 0 ? W - 1 : r - 1) : lane; // idle lanes shuffle with themselves
            double L = __shfl_sync(0xffffffffu, lam_r, act ? g * W + (r + K >= W ? r + K - W : r + K) : lane);
            double T = 0.0;
            acc = 0.0;
#pragma unroll
            for (int k = K; k >= 1; --k)
            {
                const double ak = (EVEN && k == K) ? ah : a[k < KF ? k : 0];
                acc = fma(ak, L, acc);
                T = fma(ak, lam_r, T);
                T = __shfl_sync(0xffffffffu, T, src);
                if (k > 1)
                    L = __shfl_sync(0xffffffffu, L, src);
            }
            acc += fma(a[0], lam_r, T);
        }
        else
            acc = packed_row_general(Se, m, W, g, r, lane, lam_r);
        if (r < m)
            slot_val[cur.s0 + r] = acc;
        cur = nxt;
        lam_r = (more && r < cur.m) ? t[slot_trace32[cur.s0 + r]] : 0.0;
    }
}

typedef void (*schur_packed_fn)(unsigned long long, const SchurElemDesc *, const int32_t *, const double *, const double *, double *,
                                const int *);
static schur_packed_fn schur_packed_for(int max_m, int *width)
{
    // widths of the conforming meshes (4 p multipliers per element, p = 1 .. 8) and the sizes in between
#define SP_CASE(Wv)                                                                                            \
    if (max_m <= Wv)                                                                                           \
    {                                                                                                          \
        *width = Wv;                                                                                           \
        return schur_local_packed_kernel<Wv>;                                                                  \
    }
    SP_CASE(4) SP_CASE(8) SP_CASE(12) SP_CASE(16) SP_CASE(20) SP_CASE(24) SP_CASE(28) SP_CASE(32)
#undef SP_CASE
    return nullptr;
}
#endif

// slot_val[s0 + r] = S_e[r][r]: the elements' contributions to diag(S)
__global__ void schur_diag_kernel(unsigned long long n_elem, const int64_t *__restrict__ slot_off, const int64_t *__restrict__ s_off,
                                  const double *__restrict__ S, double *__restrict__ slot_val, int packed)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        const double *Se = S + s_off[e];
        for (int r = threadIdx.x; r < m; r += blockDim.x)
            slot_val[s0 + r] = packed ? Se[r] : Se[(size_t)r * m + r];
    }
}

// out[i] = sum of the slots of multiplier i; optionally partial sums of out[i] * w[i] (for p.Ap)
__global__ void __launch_bounds__(SYS_THREADS) trace_gather_kernel(long long n_trace, const int64_t *__restrict__ tr_ptr,
                                                                  const int64_t *__restrict__ tr_slot,
                                                                  const double *__restrict__ slot_val, double *__restrict__ out,
                                                                  const int *__restrict__ done)
{
    if (done && *done)
        return;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_trace; i += (long long)gridDim.x * blockDim.x)
    {
        double acc = 0.0;
        for (int64_t k = tr_ptr[i]; k < tr_ptr[i + 1]; ++k)
            acc += slot_val[tr_slot[k]];
        out[i] = acc;
    }
}

// deterministic two-stage dot product: partials[blockIdx] = sum over the CTA's strided range of x*y(*w)
__global__ void __launch_bounds__(SYS_THREADS) dot_partial_kernel(long long n, const double *__restrict__ x,
                                                                 const double *__restrict__ y, const double *__restrict__ w,
                                                                 double *__restrict__ partials, const int *__restrict__ done)
{
    if (done && *done)
        return;
    double acc = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        acc = fma(w ? x[i] * w[i] : x[i], y[i], acc);
    block_reduce_store(acc, partials);
}

__global__ void sum_partials_kernel(int n_partials, const double *__restrict__ partials, double *__restrict__ result)
{
    const double s = warp_sum_partials(n_partials, partials);
    if (threadIdx.x == 0)
        *result = s;
}

__global__ void axpby_kernel(long long n, double a, const double *__restrict__ x, double b, const double *__restrict__ y,
                             double *__restrict__ out)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        out[i] = a * x[i] + b * y[i];
}

// out[e * os + j] = a x[e * xs + j] + b y[e * ys + j] (j < len): the DoFs of one form of every element at once
__global__ void strided_axpby_kernel(unsigned long long n_blocks, unsigned long long len, double a, const double *x, unsigned long long xs,
                                     double b, const double *y, unsigned long long ys, double *out, unsigned long long os)
{
    const unsigned long long total = n_blocks * len;
    for (unsigned long long g = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x; g < total; g += (unsigned long long)gridDim.x * blockDim.x)
    {
        const unsigned long long e = g / len, j = g - e * len;
        out[e * os + j] = a * x[e * xs + j] + b * y[e * ys + j];
    }
}

// partials[blockIdx] = max |x_i| over the CTA's strided range; a NaN anywhere is reported as NaN (numpy's max does the same)
__global__ void __launch_bounds__(SYS_THREADS) absmax_partial_kernel(long long n, const double *__restrict__ x, double *__restrict__ partials)
{
    __shared__ double red[32];
    double m = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double v = fabs(x[i]);
        m = (v > m || v != v) ? v : m; // once NaN, stays NaN (NaN > m and m > NaN are both false, v != v keeps it)
        if (m != m)
            break;
    }
#ifndef MFV2D_EMU
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
    {
        const double v = __shfl_xor_sync(0xffffffffu, m, o);
        m = (v > m || v != v) ? v : m;
    }
#endif
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    if ((threadIdx.x & 31) == 0)
        red[warp] = m;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double r = 0.0;
        for (int w = 0; w < nwarps; ++w)
            r = (red[w] > r || red[w] != red[w]) ? red[w] : r;
        partials[blockIdx.x] = r;
    }
}

__global__ void absmax_final_kernel(int np, const double *__restrict__ partials, double *__restrict__ result)
{
    if (threadIdx.x == 0 && blockIdx.x == 0)
    {
        double r = 0.0;
        for (int i = 0; i < np; ++i)
            r = (partials[i] > r || partials[i] != partials[i]) ? partials[i] : r;
        *result = r;
    }
}

// v[i] *= d[i] (inverse = 0) or v[i] /= d[i] (inverse = 1)
__global__ void scale_by_kernel(long long n, const double *__restrict__ d, double *__restrict__ v, int inverse)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        v[i] = inverse ? v[i] / d[i] : v[i] * d[i];
}

int sys_scale_rows(const mfv2d_system *s, double *v, bool inverse)
{
    if (!s->scaled || !s->n_trace)
        return MFV2D_SUCCESS;
    MFV2D_LAUNCH(scale_by_kernel, grid_for(s->ctx, s->n_trace, 256), 256, 0, s->ctx->stream, (long long)s->n_trace, (const double *)s->d_rs, v,
                 (int)inverse);
    s->ctx->launches++;
    MFV2D_CUDA_CHECK(s->ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

// multi-GPU: pack the shared multipliers of `t` into the interface buffer / scatter the reduced sums back
__global__ void iface_pack_kernel(long long n_shared, const int64_t *__restrict__ index, const int64_t *__restrict__ slot,
                                  const double *__restrict__ t, double *__restrict__ buf)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_shared; i += (long long)gridDim.x * blockDim.x)
        buf[slot[i]] = t[index[i]];
}
__global__ void iface_unpack_kernel(long long n_shared, const int64_t *__restrict__ index, const int64_t *__restrict__ slot,
                                    const double *__restrict__ buf, double *__restrict__ t)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n_shared; i += (long long)gridDim.x * blockDim.x)
        t[index[i]] = buf[slot[i]];
}

// ---- host helpers ------------------------------------------------------------------------------
int sys_gemv(const mfv2d_system *s, const double *M, const double *x, double *y)
{
    mfv2d_ctx *ctx = s->ctx;
    if (s->n_elem == 0)
        return MFV2D_SUCCESS;
    const int nwarps = SYS_THREADS / 32;
    const int xs_stride = (s->max_n + 3) & ~1;
    const size_t smem = (size_t)nwarps * xs_stride * sizeof(double);
    if (smem > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "block_gemv: blocks of %d rows need %zu bytes of shared memory", s->max_n, smem);
    if (smem > 48 * 1024)
        MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(block_gemv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    long long grid = (long long)ctx->sm_count * 8;
    const long long need = (s->n_elem + nwarps - 1) / nwarps;
    if (grid > need)
        grid = need;
    MFV2D_LAUNCH(block_gemv_kernel, (unsigned)grid, SYS_THREADS, smem, ctx->stream, (unsigned long long)s->n_elem, s->d_block_off,
                 s->d_mat_off, M, x, y, xs_stride);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int sys_spmv(const mfv2d_system *s, long long rows, const int64_t *ptr, const int64_t *col, const double *val,
                    const double *x, double *y)
{
    mfv2d_ctx *ctx = s->ctx;
    if (rows == 0)
        return MFV2D_SUCCESS;
    MFV2D_LAUNCH(csr_spmv_kernel, grid_for(ctx, rows, 256), 256, 0, ctx->stream, rows, ptr, col, val, x, y);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

// a partitioned system whose interface was never declared would hang the other ranks in the collective (or silently
// double count shared multipliers in the dot products): refuse instead
int sys_require_interface(const mfv2d_system *s)
{
    if (s->ctx->n_ranks > 1 && !s->interface_set)
        return mfv2d_fail(s->ctx, MFV2D_CUDA_NCCL_ERROR,
                          "the context has %d ranks but mfv2d_cuda_system_set_interface was not called on this system", s->ctx->n_ranks);
    return MFV2D_SUCCESS;
}

// the mapped peer blocks carry every exchange unless MFV2D_CG_EXCHANGE=nccl asks for the NCCL path (A/B comparison)
bool sys_use_peers(const mfv2d_system *s)
{
    if (s->ctx->n_ranks <= 1 || !s->peers_ready)
        return false;
    const char *ex_env = getenv("MFV2D_CG_EXCHANGE");
    return !(ex_env && ex_env[0] == 'n');
}

// sum the shared multipliers of a trace vector over the ranks (no-op on one GPU)
int sys_interface_sum(const mfv2d_system *s, double *t, int extra_scalars, double *scalars)
{
    mfv2d_ctx *ctx = s->ctx;
    if (ctx->n_ranks <= 1)
        return MFV2D_SUCCESS;
    MFV2D_TRY(sys_require_interface(s));
    if (sys_use_peers(s))
        return peer_interface_sum(s, t, extra_scalars, scalars);
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(s->d_iface, 0, (size_t)(s->n_iface + 8) * sizeof(double), ctx->stream));
    if (s->n_shared)
    {
        MFV2D_LAUNCH(iface_pack_kernel, grid_for(ctx, s->n_shared, 256), 256, 0, ctx->stream, (long long)s->n_shared,
                     s->d_shared_index, s->d_shared_slot, t, s->d_iface);
        ctx->launches++;
    }
    if (extra_scalars)
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(s->d_iface + s->n_iface, scalars, extra_scalars * sizeof(double),
                                              cudaMemcpyDeviceToDevice, ctx->stream));
    MFV2D_TRY(comm_allreduce(ctx, s->d_iface, (size_t)s->n_iface + extra_scalars));
    if (s->n_shared)
    {
        MFV2D_LAUNCH(iface_unpack_kernel, grid_for(ctx, s->n_shared, 256), 256, 0, ctx->stream, (long long)s->n_shared,
                     s->d_shared_index, s->d_shared_slot, s->d_iface, t);
        ctx->launches++;
    }
    if (extra_scalars)
        MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(scalars, s->d_iface + s->n_iface, extra_scalars * sizeof(double),
                                              cudaMemcpyDeviceToDevice, ctx->stream));
    return MFV2D_SUCCESS;
}

int sys_apply_trace(const mfv2d_system *s, const double *x, double *t)
{
    MFV2D_TRY(sys_spmv(s, s->n_trace, s->d_n_ptr, s->d_n_col, s->d_n_val, x, t));
    return sys_interface_sum(s, t, 0, nullptr);
}

// CSR of N^T built on the device from the per-element local rows (set-up): the rows of N^T that belong to element e are its
// dense columns; their entries, ordered by multiplier, are the element's local-row entries bucketed by local column.  One
// CTA per element; the histogram lives in shared memory.  (Building these arrays on the host and uploading them - 8 bytes
// per dense DoF for t_ptr alone - was a quarter of the set-up time.)
__global__ void __launch_bounds__(128) build_transpose_kernel(unsigned long long n_elem, const int64_t *__restrict__ block_off,
                                                             const int64_t *__restrict__ slot_off, const int64_t *__restrict__ slot_trace,
                                                             const int64_t *__restrict__ lrow_ptr, const int32_t *__restrict__ lrow_col,
                                                             const double *__restrict__ lrow_val, const double *__restrict__ rs,
                                                             int64_t *__restrict__ t_ptr, int64_t *__restrict__ t_col, double *__restrict__ t_val)
{
    MFV2D_DYN_SMEM(int, cnt);
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int64_t b0 = block_off[e];
        const int ne = (int)(block_off[e + 1] - b0);
        const int64_t s0 = slot_off[e], s1 = slot_off[e + 1];
        const int64_t a0 = lrow_ptr[s0], a1 = lrow_ptr[s1]; // the element's entries are contiguous; a0 = entries of the elements before it
        for (int c = threadIdx.x; c <= ne; c += blockDim.x)
            cnt[c] = 0;
        __syncthreads();
        for (int64_t a = a0 + threadIdx.x; a < a1; a += blockDim.x)
            atomicAdd(&cnt[lrow_col[a] + 1], 1);
        __syncthreads();
        if (threadIdx.x == 0)
        {
            for (int c = 0; c < ne; ++c)
                cnt[c + 1] += cnt[c];
            if (e == 0)
                t_ptr[0] = 0;
        }
        __syncthreads();
        for (int c = threadIdx.x; c < ne; c += blockDim.x)
            t_ptr[b0 + c + 1] = a0 + cnt[c + 1];
        __syncthreads();
        if (threadIdx.x == 0) // in slot (= multiplier) order, so that the entries of one column keep it
            for (int64_t sl = s0; sl < s1; ++sl)
            {
                const int64_t row = slot_trace[sl];
                const double unscale = rs ? 1.0 / rs[row] : 1.0; // rs is a power of two: exact
                for (int64_t a = lrow_ptr[sl]; a < lrow_ptr[sl + 1]; ++a)
                {
                    const int64_t pos = a0 + cnt[lrow_col[a]]++;
                    t_col[pos] = row;
                    t_val[pos] = lrow_val[a] * unscale;
                }
            }
        __syncthreads();
    }
}

// per-element products S_e t_e -> slot_val (the big kernel of a Schur application)
int sys_schur_local(const mfv2d_system *s, const double *t_in, const int *done)
{
    mfv2d_ctx *ctx = s->ctx;
    if (!s->n_elem || !s->max_m)
        return MFV2D_SUCCESS;
    const int warps = SYS_THREADS / 32;
    const long long cap = (long long)ctx->sm_count * 8;
#ifndef MFV2D_EMU
    if (s->symmetric && s->max_m <= 32 && s->d_edesc)
    {
        int W = 32;
        const schur_packed_fn fn = schur_packed_for(s->max_m, &W);
        const int G = 32 / W;
        long long grid = (s->n_elem + (long long)warps * G - 1) / ((long long)warps * G);
        int per_sm = 0; // one wave of resident CTAs: a second, partly filled wave would cost up to a whole pass
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, SYS_THREADS, 0) != cudaSuccess || per_sm < 1)
            per_sm = 4;
        if (grid > (long long)ctx->sm_count * per_sm)
            grid = (long long)ctx->sm_count * per_sm;
        fn<<<(unsigned)grid, SYS_THREADS, 0, ctx->stream>>>((unsigned long long)s->n_elem, (const SchurElemDesc *)s->d_edesc,
                                                            (const int32_t *)s->d_slot_trace32, (const double *)s->d_S, t_in, s->d_slot_val, done);
        ctx->launches++;
        MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
        return MFV2D_SUCCESS;
    }
#endif
    long long grid = (s->n_elem + warps - 1) / warps;
    if (grid > cap)
        grid = cap;
    const size_t smem = (size_t)warps * (s->max_m + 1 + 32) * sizeof(double);
    if (smem > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "schur product: elements with %d multipliers need %zu bytes of shared memory",
                          s->max_m, smem);
    if (s->symmetric)
    {
        if (smem > 48 * 1024)
            MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(schur_local_packed_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MFV2D_LAUNCH(schur_local_packed_generic_kernel, (unsigned)grid, SYS_THREADS, smem, ctx->stream, (unsigned long long)s->n_elem,
                     s->d_slot_off, s->d_slot_trace, s->d_s_off, s->d_S, t_in, s->d_slot_val, s->max_m + 1 + 32, done);
    }
    else
    {
        if (smem > 48 * 1024)
            MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(schur_local_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        MFV2D_LAUNCH(schur_local_kernel, (unsigned)grid, SYS_THREADS, smem, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off,
                     s->d_slot_trace, s->d_s_off, s->d_S, t_in, s->d_slot_val, s->max_m + 1, done);
    }
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int sys_apply_schur(const mfv2d_system *s, const double *t_in, double *t_out, const int *done, bool interface_sum)
{
    mfv2d_ctx *ctx = s->ctx;
    MFV2D_TRY(sys_schur_local(s, t_in, done));
    if (s->n_trace)
    {
        MFV2D_LAUNCH(trace_gather_kernel, grid_for(ctx, s->n_trace, SYS_THREADS), SYS_THREADS, 0, ctx->stream,
                     (long long)s->n_trace, s->d_tr_ptr, s->d_tr_slot, s->d_slot_val, t_out, done);
        ctx->launches++;
    }
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return interface_sum ? sys_interface_sum(s, t_out, 0, nullptr) : MFV2D_SUCCESS;
}

// diag(S) -> out[n_trace] (summed over the elements and, on a partitioned system, over the ranks)
int sys_schur_diagonal(const mfv2d_system *s, double *out)
{
    mfv2d_ctx *ctx = s->ctx;
    long long grid = (long long)ctx->sm_count * 8;
    if (grid > s->n_elem)
        grid = s->n_elem;
    if (s->n_elem)
    {
        MFV2D_LAUNCH(schur_diag_kernel, (unsigned)grid, 32, 0, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off, s->d_s_off,
                     s->d_S, s->d_slot_val, (int)s->symmetric);
        ctx->launches++;
    }
    if (s->n_trace)
    {
        MFV2D_LAUNCH(trace_gather_kernel, grid_for(ctx, s->n_trace, SYS_THREADS), SYS_THREADS, 0, ctx->stream, (long long)s->n_trace,
                     s->d_tr_ptr, s->d_tr_slot, s->d_slot_val, out, (const int *)nullptr);
        ctx->launches++;
    }
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return sys_interface_sum(s, out, 0, nullptr);
}

// dot product of two trace (w = ownership weights) or dense (w = nullptr) vectors -> device scalar
int sys_dot_device(const mfv2d_system *s, long long n, const double *x, const double *y, const double *w, double *d_result)
{
    mfv2d_ctx *ctx = s->ctx;
    MFV2D_LAUNCH(dot_partial_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, n, x, y, w, s->d_partials,
                 (const int *)nullptr);
    MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, s->n_partials, s->d_partials, d_result);
    ctx->launches += 2;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    if (ctx->n_ranks > 1)
    {
        if (sys_use_peers(s))
            return peer_interface_sum(s, nullptr, 1, d_result);
        MFV2D_TRY(comm_allreduce(ctx, d_result, 1));
    }
    return MFV2D_SUCCESS;
}

int read_scalars(mfv2d_ctx *ctx, const double *d, int n, double *h)
{
    MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(ctx->h_scalars, d, n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    memcpy(h, ctx->h_scalars, n * sizeof(double));
    return MFV2D_SUCCESS;
}

extern "C" {

void mfv2d_cuda_system_destroy(mfv2d_system *s)
{
    if (!s)
        return;
    cudaStreamSynchronize(s->ctx->stream);
    sys_release_peers(s);
    for (void *p : s->allocs)
        cudaFree(p);
    delete s;
}

int64_t mfv2d_cuda_system_dense_size(const mfv2d_system *s) { return s->n_dense; }
int64_t mfv2d_cuda_system_trace_size(const mfv2d_system *s) { return s->n_trace; }

int mfv2d_cuda_system_create(mfv2d_ctx *ctx, size_t n_elem, const int64_t *block_offsets, const double *blocks,
                             int blocks_space, size_t n_trace, size_t n_con, const int64_t *con_row, const int64_t *con_col,
                             const double *con_val, mfv2d_system **out)
{
    *out = nullptr;
    mfv2d_system *s = new mfv2d_system();
    s->ctx = ctx;
    s->n_elem = (int64_t)n_elem;
    s->n_trace = (int64_t)n_trace;
    s->n_con = (int64_t)n_con;
    s->n_dense = n_elem ? block_offsets[n_elem] : 0;
    int r = MFV2D_SUCCESS;
#define SYS_TRY(expr)                                                                                          \
    do                                                                                                         \
    {                                                                                                          \
        r = (expr);                                                                                            \
        if (r != MFV2D_SUCCESS)                                                                                \
        {                                                                                                      \
            mfv2d_cuda_system_destroy(s);                                                                      \
            return r;                                                                                          \
        }                                                                                                      \
    } while (0)

    // MFV2D_SYSTEM_TIMING=1: wall-clock milliseconds of the set-up phases on stderr (the device is synchronised at each mark)
    const bool timing = getenv("MFV2D_SYSTEM_TIMING") != nullptr;
    auto t_last = std::chrono::steady_clock::now();
    auto tick = [&](const char *what) {
        if (!timing)
            return;
        cudaStreamSynchronize(ctx->stream);
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[system_create rank %d] %-28s %8.2f ms\n", ctx->rank, what, std::chrono::duration<double, std::milli>(now - t_last).count());
        t_last = now;
    };
    std::vector<int64_t> boff(block_offsets, block_offsets + n_elem + 1), moff(n_elem + 1, 0);
    if (n_elem == 0)
        boff.assign(1, 0);
    s->uniform = true;
    for (size_t e = 0; e < n_elem; ++e)
    {
        const int64_t n = boff[e + 1] - boff[e];
        if (n <= 0)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "system_create: element %zu has a non-positive block size", e));
        moff[e + 1] = moff[e] + n * n;
        if (n > s->max_n)
            s->max_n = (int)n;
        if (n != boff[1] - boff[0])
            s->uniform = false;
    }
    // ---- constraint structure (host) ----
    // validate ordering and bounds; build CSR of N, CSR of N^T, and the per-element local rows.  Every pass below is
    // sequential or works on cache-resident pieces: the entries are first grouped by element (one counting sort), then N^T and
    // the local rows are built element by element - a histogram over all n_dense columns with random increments costs 5x
    // the time (15 M columns for 125 k order-6 elements).
    std::vector<int64_t> n_ptr(n_trace + 1, 0);
    const int64_t ne0 = n_elem ? boff[1] - boff[0] : 1;
    std::vector<int32_t> elem_of(n_con);
    std::vector<int64_t> ecount(n_elem + 1, 0);
    {
        size_t e = 0;
        for (size_t k = 0; k < n_con; ++k)
        {
            const int64_t c = con_col[k];
            if (con_row[k] < 0 || (size_t)con_row[k] >= n_trace || c < 0 || c >= s->n_dense)
                SYS_TRY(mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "system_create: constraint entry %zu is out of range", k));
            if (k && (con_row[k] < con_row[k - 1] || (con_row[k] == con_row[k - 1] && c <= con_col[k - 1])))
                SYS_TRY(mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "system_create: constraint entries must be sorted by (row, column) without duplicates"));
            n_ptr[con_row[k] + 1]++;
            if (s->uniform)
                e = (size_t)(c / ne0);
            else if (!(e < n_elem && boff[e] <= c && c < boff[e + 1])) // columns are not monotone over k, so search
                e = (size_t)(std::upper_bound(boff.begin(), boff.end(), c) - boff.begin()) - 1;
            elem_of[k] = (int32_t)e;
            ecount[e + 1]++;
        }
    }
    for (size_t i = 0; i < n_trace; ++i)
        n_ptr[i + 1] += n_ptr[i];
    for (size_t e = 0; e < n_elem; ++e)
        ecount[e + 1] += ecount[e];
    std::vector<int64_t> n_col(con_col, con_col + n_con);
    std::vector<double> n_val(con_val, con_val + n_con);
    // entries grouped by element in (row, col) order (stable counting sort); the grouped copies make every later pass sequential
    std::vector<int64_t> g_row(n_con);
    std::vector<int32_t> g_lcol(n_con);
    std::vector<double> g_val(n_con);
    {
        std::vector<int64_t> fill(ecount.begin(), ecount.end() - 1);
        for (size_t k = 0; k < n_con; ++k)
        {
            const int32_t e = elem_of[k];
            const int64_t pos = fill[e]++;
            g_row[pos] = con_row[k];
            g_lcol[pos] = (int32_t)(con_col[k] - boff[e]);
            g_val[pos] = con_val[k];
        }
    }
    std::vector<int32_t>().swap(elem_of);
    // Row equilibration of the CONDENSED system only: S' = D S D with D = 2^-round(log2 max|row of N|) (exact scaling).  The
    // hanging-side rows of non-conforming meshes carry coefficients up to 1e8 (inverse interpolation matrices,
    // python/mfv2d/continuity.py:256-265), which would put 16 orders of magnitude between the entries of S and make the
    // residual test of the CG blind to every other row.  Conforming rows (+-1) get D = 1: nothing changes for them, bit for
    // bit.  N, A and the operators keep the caller's scaling; partitioned systems are not equilibrated (shared rows would
    // need the same D on every rank; conforming cuts have +-1 anyway).
    std::vector<double> rs(n_trace, 1.0);
    if (ctx->n_ranks <= 1)
    {
        std::vector<double> mx(n_trace, 0.0);
        for (size_t k = 0; k < n_con; ++k)
            mx[con_row[k]] = std::max(mx[con_row[k]], std::fabs(con_val[k]));
        for (size_t i = 0; i < n_trace; ++i)
            if (mx[i] > 0.0 && std::isfinite(mx[i]) && mx[i] != 1.0)
            {
                rs[i] = std::ldexp(1.0, -(int)std::lround(std::log2(mx[i])));
                if (rs[i] != 1.0)
                    s->scaled = true;
            }
    }
    // element by element: the local rows (slots) and the element's rows of N^T (its dense columns, entries by multiplier)
    std::vector<int64_t> slot_off(n_elem + 1, 0), slot_trace, lrow_ptr(1, 0);
    std::vector<int32_t> lrow_col(n_con);
    std::vector<double> lrow_val(n_con);
    slot_trace.reserve(n_con);
    lrow_ptr.reserve(n_con + 1);
    {
        for (size_t e = 0; e < n_elem; ++e)
        {
            const int64_t lo = ecount[e], hi = ecount[e + 1];
            int64_t prev_row = -1;
            for (int64_t a = lo; a < hi; ++a)
            {
                if (g_row[a] != prev_row)
                {
                    if (prev_row >= 0)
                        lrow_ptr.push_back(a);
                    slot_trace.push_back(g_row[a]);
                    prev_row = g_row[a];
                }
                lrow_col[a] = g_lcol[a];
                lrow_val[a] = g_val[a] * rs[g_row[a]];
            }
            if (prev_row >= 0)
                lrow_ptr.push_back(hi);
            slot_off[e + 1] = (int64_t)slot_trace.size();
            const int m = (int)(slot_off[e + 1] - slot_off[e]);
            if (m > s->max_m)
                s->max_m = m;
        }
    }
    s->n_slots = (int64_t)slot_trace.size();
    std::vector<int64_t> tr_ptr(n_trace + 1, 0), tr_slot(s->n_slots);
    for (int64_t k = 0; k < s->n_slots; ++k)
        tr_ptr[slot_trace[k] + 1]++;
    for (size_t i = 0; i < n_trace; ++i)
        tr_ptr[i + 1] += tr_ptr[i];
    {
        std::vector<int64_t> fill(tr_ptr.begin(), tr_ptr.end() - 1);
        for (int64_t k = 0; k < s->n_slots; ++k)
            tr_slot[fill[slot_trace[k]]++] = k;
    }
    tick("host constraint structure");
    // ---- upload ----
    SYS_TRY(sys_upload(s, boff, &s->d_block_off));
    SYS_TRY(sys_upload(s, moff, &s->d_mat_off));
    SYS_TRY(sys_upload(s, n_ptr, &s->d_n_ptr));
    SYS_TRY(sys_upload(s, n_col, &s->d_n_col));
    SYS_TRY(sys_upload(s, n_val, &s->d_n_val));
    SYS_TRY(sys_upload(s, slot_off, &s->d_slot_off));
    SYS_TRY(sys_upload(s, slot_trace, &s->d_slot_trace));
    SYS_TRY(sys_upload(s, lrow_ptr, &s->d_lrow_ptr));
    SYS_TRY(sys_upload(s, lrow_col, &s->d_lrow_col));
    SYS_TRY(sys_upload(s, lrow_val, &s->d_lrow_val));
    SYS_TRY(sys_upload(s, tr_ptr, &s->d_tr_ptr));
    SYS_TRY(sys_upload(s, tr_slot, &s->d_tr_slot));
    if (s->scaled)
        SYS_TRY(sys_upload(s, rs, &s->d_rs));
    // N^T (CSR over the dense DoFs) on the device, from the local rows
    SYS_TRY(sys_alloc(s, (size_t)s->n_dense + 1, &s->d_t_ptr));
    SYS_TRY(sys_alloc(s, n_con ? n_con : 1, &s->d_t_col));
    SYS_TRY(sys_alloc(s, n_con ? n_con : 1, &s->d_t_val));
    if (n_elem)
    {
        long long grid = (long long)ctx->sm_count * 16;
        if (grid > (long long)n_elem)
            grid = (long long)n_elem;
        const size_t smem = (size_t)(s->max_n + 2) * sizeof(int);
        if (smem > 48 * 1024)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "system_create: blocks of %d rows exceed the shared memory of the transpose kernel", s->max_n));
        MFV2D_LAUNCH(build_transpose_kernel, (unsigned)grid, 128, smem, ctx->stream, (unsigned long long)n_elem, s->d_block_off, s->d_slot_off,
                     s->d_slot_trace, s->d_lrow_ptr, s->d_lrow_col, s->d_lrow_val, s->scaled ? (const double *)s->d_rs : (const double *)nullptr,
                     s->d_t_ptr, s->d_t_col, s->d_t_val);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "system_create: transpose kernel launch failed"));
    }
    else
        cudaMemsetAsync(s->d_t_ptr, 0, sizeof(int64_t), ctx->stream);
    {
        // conforming meshes: at most two elements per multiplier -> a fixed-width 32-bit table for the CG's gather kernel
        bool two = s->n_slots < (int64_t)1 << 31;
        for (size_t i = 0; i < n_trace && two; ++i)
            two = tr_ptr[i + 1] - tr_ptr[i] <= 2;
        if (two && n_trace)
        {
            std::vector<int32_t> tr2(2 * n_trace, -1);
            for (size_t i = 0; i < n_trace; ++i)
                for (int64_t k = tr_ptr[i]; k < tr_ptr[i + 1]; ++k)
                    tr2[2 * i + (k - tr_ptr[i])] = (int32_t)tr_slot[k];
            SYS_TRY(sys_upload(s, tr2, &s->d_tr2));
        }
    }
    tick("uploads");
    const size_t mat_total = (size_t)moff[n_elem];
    if (blocks_space == MFV2D_DEVICE)
    {
        s->d_A = const_cast<double *>(blocks); // borrowed: the caller keeps it alive
        s->owns_A = false;
    }
    else
    {
        SYS_TRY(sys_alloc(s, mat_total, &s->d_A));
        if (mat_total)
            SYS_TRY(mfv2d_cuda_upload(ctx, s->d_A, blocks, mat_total * sizeof(double)));
    }
    // symmetric blocks (every system of the reference's examples without an advection term) -> packed condensed blocks
    {
        const char *pk = getenv("MFV2D_SCHUR_PACKED");
        s->symmetric = n_elem > 0;
        if (pk && pk[0] == '0')
            s->symmetric = false;
        else if (n_elem)
        {
            int *d_flag = reinterpret_cast<int *>(ctx->d_scalars + 32);
            cudaMemsetAsync(d_flag, 0, sizeof(int), ctx->stream);
            long long grid = (long long)ctx->sm_count * 8;
            if (grid > (long long)n_elem)
                grid = (long long)n_elem;
            MFV2D_LAUNCH(block_asymmetry_kernel, (unsigned)grid, 256, 0, ctx->stream, (unsigned long long)n_elem, s->d_block_off,
                         s->d_mat_off, (const double *)s->d_A, d_flag);
            ctx->launches++;
            int h_flag = 1;
            if (cudaMemcpyAsync(&h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
                cudaStreamSynchronize(ctx->stream) != cudaSuccess)
                SYS_TRY(mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "system_create: symmetry test failed: %s", cudaGetErrorString(cudaGetLastError())));
            s->symmetric = h_flag == 0;
        }
    }
    tick("symmetry test");
    std::vector<int64_t> s_off(n_elem + 1, 0);
    for (size_t e = 0; e < n_elem; ++e)
    {
        const int64_t m = slot_off[e + 1] - slot_off[e];
        s_off[e + 1] = s_off[e] + (s->symmetric ? packed_size(m) : m * m);
    }
    SYS_TRY(sys_upload(s, s_off, &s->d_s_off));
    if (s->n_slots < (int64_t)1 << 31 && (int64_t)n_trace < (int64_t)1 << 31)
    {
        // compact per-element descriptors and 32-bit multiplier indices for the packed Schur product
        std::vector<SchurElemDesc> edesc(n_elem);
        for (size_t e = 0; e < n_elem; ++e)
            edesc[e] = SchurElemDesc{(long long)s_off[e], (int)slot_off[e], (int)(slot_off[e + 1] - slot_off[e])};
        std::vector<int32_t> st32(slot_trace.begin(), slot_trace.end());
        SYS_TRY(sys_upload(s, edesc, &s->d_edesc));
        SYS_TRY(sys_upload(s, st32, &s->d_slot_trace32));
    }
    SYS_TRY(sys_alloc(s, mat_total, &s->d_Ainv));
    SYS_TRY(sys_alloc(s, (size_t)s_off[n_elem], &s->d_S));
    SYS_TRY(sys_alloc(s, (size_t)s->n_slots, &s->d_slot_val));
    SYS_TRY(sys_alloc(s, (size_t)s->n_trace, &s->d_diag_inv));
    for (int i = 0; i < 3; ++i)
        SYS_TRY(sys_alloc(s, (size_t)s->n_dense, &s->d_work_dense[i]));
    for (int i = 0; i < 8; ++i)
        SYS_TRY(sys_alloc(s, (size_t)s->n_trace, &s->d_work_trace[i]));
    SYS_TRY(sys_alloc(s, (size_t)64, &s->d_cg));
    cudaMemsetAsync(s->d_cg, 0, 64 * sizeof(double), ctx->stream);
    {
        // CTAs of the vector kernels of the Krylov solvers (= partial sums of their dot products) per SM
        const char *env = getenv("MFV2D_CG_CTAS_PER_SM");
        const int per_sm = env ? atoi(env) : 4;
        s->n_partials = ctx->sm_count * (per_sm >= 1 && per_sm <= 16 ? per_sm : 4);
    }
    SYS_TRY(sys_alloc(s, (size_t)3 * s->n_partials, &s->d_partials));
    tick("allocations");
    // ---- factorise: explicit inverses of the dense blocks ----
    int32_t *d_status = nullptr;
    SYS_TRY(sys_alloc(s, n_elem, &d_status));
    if (s->uniform && n_elem)
        SYS_TRY(launch_inverse_batch(ctx, s->max_n, n_elem, s->d_A, s->d_Ainv, d_status));
    else
        for (size_t e = 0; e < n_elem;)
        {
            // runs of equal block size
            size_t e2 = e + 1;
            const int64_t n = boff[e + 1] - boff[e];
            while (e2 < n_elem && boff[e2 + 1] - boff[e2] == n)
                ++e2;
            SYS_TRY(launch_inverse_batch(ctx, (int)n, e2 - e, s->d_A + moff[e], s->d_Ainv + moff[e], d_status + e));
            e = e2;
        }
    tick("block inverses");
    if (n_elem)
    {
        std::vector<int32_t> status(n_elem);
        SYS_TRY(mfv2d_cuda_download(ctx, status.data(), d_status, n_elem * sizeof(int32_t)));
        for (size_t e = 0; e < n_elem; ++e)
            if (status[e] != MFV2D_SUCCESS)
                SYS_TRY(mfv2d_fail(ctx, MFV2D_PIVOT_FAILED, "system_create: the dense block of element %zu is singular", e));
        // condensed blocks
        long long grid = (long long)ctx->sm_count * 8;
        if (grid > (long long)n_elem)
            grid = (long long)n_elem;
        MFV2D_LAUNCH(schur_build_kernel, (unsigned)grid, 128, 0, ctx->stream, (unsigned long long)n_elem, s->d_block_off,
                     s->d_mat_off, s->d_Ainv, s->d_slot_off, s->d_lrow_ptr, s->d_lrow_col, s->d_lrow_val, s->d_s_off, s->d_S,
                     (int)s->symmetric);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "schur_build kernel launch failed"));
    }
    SYS_TRY(mfv2d_cuda_synchronize(ctx));
    tick("condensed blocks");
#undef SYS_TRY
    *out = s;
    return MFV2D_SUCCESS;
}

} // extern "C"

// generic wrapper: stage host vectors, run `fn` on device pointers
template <typename F> static int sys_unary(const mfv2d_system *s, const double *x, size_t nx, double *y, size_t ny, int space, F fn)
{
    mfv2d_ctx *ctx = s->ctx;
    Staged i, o;
    int r = stage_in(ctx, x, nx * sizeof(double), space, &i);
    if (!r)
        r = stage_out_alloc(ctx, y, ny * sizeof(double), space, &o);
    if (!r)
        r = fn((const double *)i.dev, (double *)o.dev);
    if (!r)
        r = stage_finish(&o);
    else
        stage_abort(&o);
    stage_abort(&i);
    return r;
}

extern "C" {

int mfv2d_cuda_system_apply_diagonal(const mfv2d_system *s, const double *x, double *y, int space)
{
    return sys_unary(s, x, s->n_dense, y, s->n_dense, space, [&](const double *dx, double *dy) { return sys_gemv(s, s->d_A, dx, dy); });
}

int mfv2d_cuda_system_apply_diagonal_inverse(const mfv2d_system *s, const double *x, double *y, int space)
{
    return sys_unary(s, x, s->n_dense, y, s->n_dense, space,
                     [&](const double *dx, double *dy) { return sys_gemv(s, s->d_Ainv, dx, dy); });
}

int mfv2d_cuda_system_apply_trace(const mfv2d_system *s, const double *x, double *t, int space)
{
    return sys_unary(s, x, s->n_dense, t, s->n_trace, space, [&](const double *dx, double *dt) { return sys_apply_trace(s, dx, dt); });
}

int mfv2d_cuda_system_apply_trace_transpose(const mfv2d_system *s, const double *t, double *y, int space)
{
    return sys_unary(s, t, s->n_trace, y, s->n_dense, space, [&](const double *dt, double *dy) {
        return sys_spmv(s, s->n_dense, s->d_t_ptr, s->d_t_col, s->d_t_val, dt, dy);
    });
}

int mfv2d_cuda_system_apply_schur(const mfv2d_system *s, const double *t_in, double *t_out, int space)
{
    return sys_unary(s, t_in, s->n_trace, t_out, s->n_trace, space, [&](const double *di, double *dout) {
        if (!s->scaled)
            return sys_apply_schur(s, di, dout, nullptr);
        // the stored blocks are those of S' = D S D: S t = D^-1 S' (D^-1 t)
        double *tmp = s->d_work_trace[7];
        MFV2D_CUDA_CHECK(s->ctx, cudaMemcpyAsync(tmp, di, (size_t)s->n_trace * sizeof(double), cudaMemcpyDeviceToDevice, s->ctx->stream));
        MFV2D_TRY(sys_scale_rows(s, tmp, true));
        MFV2D_TRY(sys_apply_schur(s, tmp, dout, nullptr));
        return sys_scale_rows(s, dout, true);
    });
}

int mfv2d_cuda_vec_axpby(mfv2d_ctx *ctx, size_t n, double a, const double *x, double b, const double *y, double *out,
                         int space)
{
    if (n == 0)
        return MFV2D_SUCCESS;
    Staged sx, sy, so;
    int r = stage_in(ctx, x, n * sizeof(double), space, &sx);
    if (!r)
        r = stage_in(ctx, y, n * sizeof(double), space, &sy);
    if (!r)
        r = stage_out_alloc(ctx, out, n * sizeof(double), space, &so);
    if (!r)
    {
        MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, (long long)n, 256), 256, 0, ctx->stream, (long long)n, a, (const double *)sx.dev,
                     b, (const double *)sy.dev, (double *)so.dev);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "axpby kernel launch failed");
    }
    if (!r)
        r = stage_finish(&so);
    else
        stage_abort(&so);
    stage_abort(&sx);
    stage_abort(&sy);
    return r;
}

int mfv2d_cuda_vec_absmax(mfv2d_ctx *ctx, size_t n, const double *x, double *result, int space)
{
    *result = 0.0;
    if (n == 0)
        return MFV2D_SUCCESS;
    Staged sx;
    const int np = ctx->sm_count * 4; // partial maxima live in the context's scalar buffer neighbourhood: allocate per call
    void *partials = nullptr;
    int r = stage_in(ctx, x, n * sizeof(double), space, &sx);
    if (!r)
        r = mfv2d_cuda_malloc(ctx, (size_t)np * sizeof(double), &partials);
    if (!r)
    {
        MFV2D_LAUNCH(absmax_partial_kernel, (unsigned)np, SYS_THREADS, 0, ctx->stream, (long long)n, (const double *)sx.dev, (double *)partials);
        MFV2D_LAUNCH(absmax_final_kernel, 1, 32, 0, ctx->stream, np, (const double *)partials, ctx->d_scalars);
        ctx->launches += 2;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "absmax kernel launch failed");
    }
    if (!r)
        r = read_scalars(ctx, ctx->d_scalars, 1, result);
    stage_abort(&sx);
    if (partials)
        mfv2d_cuda_free(ctx, partials);
    return r;
}

int mfv2d_cuda_strided_axpby(mfv2d_ctx *ctx, size_t n_blocks, size_t len, double a, const double *x, size_t x_stride, double b, const double *y,
                             size_t y_stride, double *out, size_t out_stride)
{
    if (n_blocks == 0 || len == 0)
        return MFV2D_SUCCESS;
    if (!x || !y || !out || x_stride < len || y_stride < len || out_stride < len)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "strided_axpby: device pointers and strides of at least `len` are required");
    MFV2D_LAUNCH(strided_axpby_kernel, grid_for(ctx, (long long)(n_blocks * len), 256), 256, 0, ctx->stream, (unsigned long long)n_blocks,
                 (unsigned long long)len, a, x, (unsigned long long)x_stride, b, y, (unsigned long long)y_stride, out, (unsigned long long)out_stride);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_vec_dot(mfv2d_ctx *ctx, size_t n, const double *x, const double *y, double *result, int space)
{
    *result = 0.0;
    if (n == 0)
        return MFV2D_SUCCESS;
    Staged sx, sy;
    void *partials = nullptr;
    const int np = ctx->sm_count * 4;
    int r = stage_in(ctx, x, n * sizeof(double), space, &sx);
    if (!r)
        r = stage_in(ctx, y, n * sizeof(double), space, &sy);
    if (!r)
        r = mfv2d_cuda_malloc(ctx, (size_t)np * sizeof(double), &partials);
    if (!r)
    {
        MFV2D_LAUNCH(dot_partial_kernel, (unsigned)np, SYS_THREADS, 0, ctx->stream, (long long)n, (const double *)sx.dev,
                     (const double *)sy.dev, (const double *)nullptr, (double *)partials, (const int *)nullptr);
        MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, np, (const double *)partials, ctx->d_scalars);
        ctx->launches += 2;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "dot kernel launch failed");
    }
    if (!r)
        r = read_scalars(ctx, ctx->d_scalars, 1, result);
    stage_abort(&sx);
    stage_abort(&sy);
    if (partials)
        mfv2d_cuda_free(ctx, partials);
    return r;
}

int mfv2d_cuda_system_set_interface(mfv2d_system *s, size_t n_shared, const int64_t *shared_index, const int64_t *slot,
                                    size_t n_slots, const uint8_t *owned)
{
    std::vector<int64_t> hi(shared_index, shared_index + n_shared), hs(slot, slot + n_shared);
    std::vector<double> w((size_t)s->n_trace);
    for (int64_t i = 0; i < s->n_trace; ++i)
        w[i] = owned[i] ? 1.0 : 0.0;
    s->n_shared = (int64_t)n_shared;
    s->n_iface = (int64_t)n_slots;
    s->interface_set = true;
    MFV2D_TRY(sys_upload(s, hi, &s->d_shared_index));
    MFV2D_TRY(sys_upload(s, hs, &s->d_shared_slot));
    MFV2D_TRY(sys_upload(s, w, &s->d_owned));
    MFV2D_TRY(sys_alloc(s, n_slots + 8, &s->d_iface));
    return MFV2D_SUCCESS;
}

} // extern "C"
