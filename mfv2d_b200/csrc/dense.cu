// Batched dense algebra on small matrices: inverse and linear solve with partial pivoting
// (replaces decompose_pivoted_lu / solve_lu / matrix_full_invert, src/algebra/matrix.c:361-534, and the
// Python-facing _compute_matrix_inverse / _solve_linear_system, :546-660).
//
// One CTA per matrix (persistent over the batch).  The working copy lives in shared memory when
// n*(n+nrhs) doubles fit (n <= ~165 for an inverse), otherwise in the output buffer in global memory
// (L2 resident).  The reference picks a static row order from the ORIGINAL column magnitudes and fails
// with MFV2D_PIVOT_FAILED when a column has no unused entry above 1e-12 (matrix.c:424-466); here true
// partial pivoting is used (same result to rounding for regular matrices) and the same code is reported
// when the best available pivot is below 1e-12.
#include "internal.cuh"

#define DENSE_THREADS 256
#define PIVOT_TOL 1e-12

// CTA-wide argmax of |col[r]| for r in [k, n): returns the row index to every thread.
__device__ __forceinline__ int block_argmax(const double *__restrict__ M, int ld, int k, int n, double *red_v, int *red_i)
{
    double best = -1.0;
    int bi = k;
    for (int r = k + threadIdx.x; r < n; r += blockDim.x)
    {
        const double v = fabs(M[(size_t)r * ld + k]);
        if (v > best)
        {
            best = v;
            bi = r;
        }
    }
#ifndef MFV2D_EMU
    for (int o = 16; o > 0; o >>= 1)
    {
        const double ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi))
        {
            best = ov;
            bi = oi;
        }
    }
#endif
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    if ((threadIdx.x & 31) == 0)
    {
        red_v[warp] = best;
        red_i[warp] = bi;
    }
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double b = red_v[0];
        int i = red_i[0];
        for (int w = 1; w < nwarps; ++w)
            if (red_v[w] > b || (red_v[w] == b && red_i[w] < i))
            {
                b = red_v[w];
                i = red_i[w];
            }
        red_v[0] = b;
        red_i[0] = i;
    }
    __syncthreads();
    const int res = red_i[0];
    return res;
}

// In-place Gauss-Jordan inverse with partial pivoting of the n x n matrix M (leading dimension n).
// scratch: 2n doubles (column and row of the pivot), piv: n ints.  Returns non-zero on pivot failure.
__device__ int gauss_jordan_inverse(double *M, int n, double *scratch, int *piv, double *red_v, int *red_i)
{
    double *colk = scratch, *rowk = scratch + n;
    int bad = 0;
    for (int k = 0; k < n; ++k)
    {
        __syncthreads();
        const int r = block_argmax(M, n, k, n, red_v, red_i);
        if (!(red_v[0] >= PIVOT_TOL))
            bad = 1;
        if (threadIdx.x == 0)
            piv[k] = r;
        if (r != k)
            for (int j = threadIdx.x; j < n; j += blockDim.x)
            {
                const double t = M[(size_t)k * n + j];
                M[(size_t)k * n + j] = M[(size_t)r * n + j];
                M[(size_t)r * n + j] = t;
            }
        __syncthreads();
        for (int i = threadIdx.x; i < n; i += blockDim.x)
        {
            colk[i] = M[(size_t)i * n + k];
            rowk[i] = M[(size_t)k * n + i];
        }
        __syncthreads();
        const double ip = 1.0 / colk[k];
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x)
        {
            const int i = idx / n, j = idx - i * n;
            double v;
            if (i == k)
                v = (j == k) ? ip : rowk[j] * ip;
            else if (j == k)
                v = -colk[i] * ip;
            else
                v = fma(-colk[i] * ip, rowk[j], M[idx]);
            M[idx] = v;
        }
    }
    __syncthreads();
    // (P A)^-1 = A^-1 P^-1  ->  undo the row swaps as column swaps in reverse order
    for (int k = n - 1; k >= 0; --k)
    {
        const int r = piv[k];
        if (r != k)
        {
            for (int i = threadIdx.x; i < n; i += blockDim.x)
            {
                const double t = M[(size_t)i * n + k];
                M[(size_t)i * n + k] = M[(size_t)i * n + r];
                M[(size_t)i * n + r] = t;
            }
            __syncthreads();
        }
    }
    return bad;
}

__global__ void __launch_bounds__(DENSE_THREADS) inverse_kernel(int n, unsigned long long batch, const double *__restrict__ in,
                                                               double *__restrict__ out, int32_t *__restrict__ status,
                                                               int use_smem)
{
    MFV2D_DYN_SMEM(double, smem);
    __shared__ double red_v[32];
    __shared__ int red_i[32];
    const size_t nn = (size_t)n * n;
    double *scratch = smem;                                   // 2n
    int *piv = reinterpret_cast<int *>(smem + 2 * n);         // n ints
    double *work_s = smem + 2 * n + ((n + 1) / 2) + 1;        // n*n when use_smem
    work_s = smem + (((int)(work_s - smem) + 1) & ~1);
    for (unsigned long long b = blockIdx.x; b < batch; b += gridDim.x)
    {
        __syncthreads();
        double *M = use_smem ? work_s : out + b * nn;
        const double *src = in + b * nn;
        if (use_smem || src != M)
            for (size_t i = threadIdx.x; i < nn; i += blockDim.x)
                M[i] = src[i];
        __syncthreads();
        const int bad = gauss_jordan_inverse(M, n, scratch, piv, red_v, red_i);
        __syncthreads();
        if (use_smem)
            for (size_t i = threadIdx.x; i < nn; i += blockDim.x)
                out[b * nn + i] = M[i];
        if (status && threadIdx.x == 0)
            status[b] = bad ? MFV2D_PIVOT_FAILED : MFV2D_SUCCESS;
    }
}

#ifndef MFV2D_EMU
// ---- register-resident batched inverse (n <= 128) ------------------------------------------------------------------
// The same partially pivoted Gauss-Jordan sweep, but the working copy of the block never touches shared memory: the CTA
// is a TI x 16 grid of threads and thread (ti, tj) keeps the entries (ti + TI a, tj + 16 b) of the block in REGISTERS for
// the whole elimination (cyclic distribution: every thread stays busy until the last column).  Per pivot only the pivot
// column and the two swapped rows cross the CTA through shared memory (2 n + 16 doubles), so a step is RA + CB
// shared-memory reads for RA x CB fused multiply-adds per thread, where inverse_kernel pays three reads and a write per
// FMA.  Row swaps are undone as a column permutation when the inverse is written out.  Pivot test of
// decompose_pivoted_lu (src/algebra/matrix.c:454-457): |pivot| < 1e-12 -> MFV2D_PIVOT_FAILED.
// CTA-uniform dispatch on a run-time index in [0, N), N <= 8: BODY sees the index as the compile-time constant I, so a row
// or column of the register tile is addressed directly.  (A predicated loop over all N candidates costs N times the
// instructions: with it a Gauss-Jordan step was 566 warp instructions for 32 FMAs per thread.)
#define UNIFORM_CASE(idx, N, C, BODY)                                                                                            \
    else if ((C) < (N) && (idx) == (C))                                                                                          \
    {                                                                                                                            \
        constexpr int I = (C) < (N) ? (C) : 0;                                                                                   \
        BODY                                                                                                                     \
    }
#define UNIFORM_SWITCH(idx, N, BODY)                                                                                             \
    do                                                                                                                           \
    {                                                                                                                            \
        if (false)                                                                                                               \
        {                                                                                                                        \
        }                                                                                                                        \
        UNIFORM_CASE(idx, N, 0, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 1, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 2, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 3, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 4, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 5, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 6, BODY)                                                                                            \
        UNIFORM_CASE(idx, N, 7, BODY)                                                                                            \
    } while (0)

template <int TI, int RA, int CB>
__global__ void __launch_bounds__(TI * 16) inverse_reg_kernel(int n, unsigned long long batch, const double *__restrict__ in,
                                                             double *__restrict__ out, int32_t *__restrict__ status)
{
    constexpr int TJ = 16, NT = TI * TJ;
    __shared__ double colk[TI * RA], row_a[TJ * CB], row_b[TJ * CB], cand_v[TI];
    __shared__ int cand_i[TI], piv[TI * RA], cperm[TJ * CB], cinv[TJ * CB];
    __shared__ double ip_s;
    const int tj = threadIdx.x & 15, ti = threadIdx.x >> 4;
    const size_t nn = (size_t)n * n;
    for (unsigned long long blk = blockIdx.x; blk < batch; blk += gridDim.x)
    {
        const double *src = in + blk * nn;
        double e[RA][CB];
#pragma unroll
        for (int a = 0; a < RA; ++a)
#pragma unroll
            for (int b = 0; b < CB; ++b)
            {
                const int i = ti + TI * a, j = tj + TJ * b;
                e[a][b] = (i < n && j < n) ? src[(size_t)i * n + j] : (i == j ? 1.0 : 0.0); // identity padding
            }
        int bad = 0;
        // the 16 pivots of one column group share the register column kb (and, TI being 16 or 32, the register row ka): the
        // outer loop is unrolled so that both are compile-time constants and the pivot column / pivot row of the tile are
        // addressed directly - only the row of the pivot ELEMENT (ra) is a run-time index
        static_assert(TI % 16 == 0, "ka must be constant over a column group");
#pragma unroll
        for (int kb = 0; kb < CB; ++kb)
        for (int ktj = 0; ktj < 16; ++ktj)
        {
            const int k = 16 * kb + ktj;
            if (k >= n)
                break;
            const int ka = (16 * kb) / TI, kti = k - ka * TI;
            // 1. pivot search in column k, rows >= k: the 16-lane column owners propose, everybody picks the winner
            if (tj == ktj)
            {
                double best = -1.0;
                int bi = k;
#pragma unroll
                for (int a = 0; a < RA; ++a)
                {
                    const int i = ti + TI * a;
                    const double v = fabs(e[a][kb]);
                    if (i >= k && i < n && v > best)
                    {
                        best = v;
                        bi = i;
                    }
                }
                cand_v[ti] = best;
                cand_i[ti] = bi;
            }
            __syncthreads();
            // every warp reduces the TI candidates by itself (no second barrier, no serial scan): non-negative doubles order
            // like their bit patterns, so the arg-max is three hardware warp reductions (REDUX) - largest high word, largest
            // low word among those, smallest row index among the winners (the tie rule of the shuffle version it replaces)
            int r;
            {
                const int lane = threadIdx.x & 31;
                const double cv = lane < TI ? cand_v[lane] : -1.0;
                const unsigned ci = lane < TI ? (unsigned)cand_i[lane] : 0x7fffffffu;
                const bool valid = cv >= 0.0; // false for the "no row" sentinel and for NaN
                const unsigned hi = valid ? (unsigned)__double2hiint(cv) : 0u, lo = valid ? (unsigned)__double2loint(cv) : 0u;
                const unsigned mh = __reduce_max_sync(0xffffffffu, hi);
                const unsigned ml = __reduce_max_sync(0xffffffffu, hi == mh ? lo : 0u);
                r = (int)__reduce_min_sync(0xffffffffu, (hi == mh && lo == ml) ? ci : 0x7fffffffu);
                const double best = __hiloint2double((int)mh, (int)ml);
                if (!(best >= PIVOT_TOL))
                    bad = 1;
                if (threadIdx.x == 0)
                    piv[k] = r;
            }
            const int ra = r / TI, rti = r - ra * TI;
            // 2. rows k and r and the (pre-swap) pivot column through shared memory: row_a = old row r (the new pivot row),
            //    row_b = old row k
            if (ti == rti)
                UNIFORM_SWITCH(ra, RA, {
#pragma unroll
                    for (int b = 0; b < CB; ++b)
                        row_a[tj + TJ * b] = e[I][b];
                    if (tj == ktj)
                        ip_s = 1.0 / e[I][kb]; // one division per pivot instead of one per thread
                });
            if (ti == kti && r != k)
            {
#pragma unroll
                for (int b = 0; b < CB; ++b)
                    row_b[tj + TJ * b] = e[ka][b];
            }
            if (tj == ktj)
            {
#pragma unroll
                for (int a = 0; a < RA; ++a)
                    colk[ti + TI * a] = e[a][kb];
            }
            __syncthreads();
            if (r != k)
            {
                if (ti == kti)
                {
#pragma unroll
                    for (int b = 0; b < CB; ++b)
                        e[ka][b] = row_a[tj + TJ * b];
                }
                if (ti == rti)
                    UNIFORM_SWITCH(ra, RA, {
#pragma unroll
                        for (int b = 0; b < CB; ++b)
                            e[I][b] = row_b[tj + TJ * b];
                    });
            }
            // 4. Gauss-Jordan update of the whole block in registers: the plain rank-one update everywhere, then the pivot
            //    column (f) and the pivot row (row * 1 / pivot, 1 / pivot on the diagonal) are overwritten
            const double ip = ip_s;
            double rk[CB], f[RA];
#pragma unroll
            for (int b = 0; b < CB; ++b)
                rk[b] = row_a[tj + TJ * b];
#pragma unroll
            for (int a = 0; a < RA; ++a)
            {
                const int i = ti + TI * a;
                f[a] = -colk[i == k ? r : (i == r ? k : i)] * ip; // the column was published before the swap
#pragma unroll
                for (int b = 0; b < CB; ++b)
                    e[a][b] = fma(f[a], rk[b], e[a][b]);
            }
            if (tj == ktj)
            {
#pragma unroll
                for (int a = 0; a < RA; ++a)
                    e[a][kb] = f[a];
            }
            if (ti == kti)
            {
#pragma unroll
                for (int b = 0; b < CB; ++b)
                    e[ka][b] = (tj + TJ * b == k) ? ip : rk[b] * ip;
            }
            // colk / row_a / row_b are read here and rewritten after the NEXT step's first barrier; cand_* are rewritten before
            // it, but were last read before this step's second barrier: two barriers per pivot are enough
        }
        // (P A)^-1 = A^-1 P^-1: the row swaps come back as column swaps in reverse order -> one permutation of the columns
        __syncthreads();
        for (int j = threadIdx.x; j < TJ * CB; j += NT)
            cperm[j] = cinv[j] = j;
        __syncthreads();
        if (threadIdx.x == 0)
        {
            // cperm[c] = output position of working column c, cinv = its inverse; the swaps (k, piv[k]), k = n-1 .. 0, exchange
            // the columns that CURRENTLY sit at positions k and piv[k]
            for (int k = n - 1; k >= 0; --k)
            {
                const int r = piv[k];
                if (r != k)
                {
                    const int ca = cinv[k], cb = cinv[r];
                    cperm[ca] = r;
                    cperm[cb] = k;
                    cinv[k] = cb;
                    cinv[r] = ca;
                }
            }
        }
        __syncthreads();
        double *dst = out + blk * nn;
#pragma unroll
        for (int a = 0; a < RA; ++a)
#pragma unroll
            for (int b = 0; b < CB; ++b)
            {
                const int i = ti + TI * a, j = tj + TJ * b;
                if (i < n && j < n)
                    dst[(size_t)i * n + cperm[j]] = e[a][b];
            }
        if (status && threadIdx.x == 0)
            status[blk] = bad ? MFV2D_PIVOT_FAILED : MFV2D_SUCCESS;
        __syncthreads();
    }
}

template <int TI, int RA, int CB>
static int launch_inverse_reg(mfv2d_ctx *ctx, int n, size_t batch, const double *d_in, double *d_out, int32_t *d_status)
{
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, inverse_reg_kernel<TI, RA, CB>, TI * 16, 0));
    if (per_sm < 1)
        per_sm = 1;
    size_t grid = (size_t)ctx->sm_count * per_sm;
    if (grid > batch)
        grid = batch;
    inverse_reg_kernel<TI, RA, CB><<<(unsigned)grid, TI * 16, 0, ctx->stream>>>(n, (unsigned long long)batch, d_in, d_out, d_status);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}
#endif

// Gaussian elimination with partial pivoting on the augmented system [A | B] (n x (n + nrhs)), then
// back substitution; the working copy W has leading dimension ld = n + nrhs.
__global__ void __launch_bounds__(DENSE_THREADS) solve_kernel(int n, int nrhs, unsigned long long batch,
                                                             const double *__restrict__ a, const double *__restrict__ bmat,
                                                             double *__restrict__ x, double *__restrict__ work_g,
                                                             int32_t *__restrict__ status, int use_smem)
{
    MFV2D_DYN_SMEM(double, smem);
    __shared__ double red_v[32];
    __shared__ int red_i[32];
    const int ld = n + nrhs;
    const size_t per = (size_t)n * ld;
    for (unsigned long long b = blockIdx.x; b < batch; b += gridDim.x)
    {
        __syncthreads();
        double *W = use_smem ? smem : work_g + (size_t)blockIdx.x * per;
        for (size_t idx = threadIdx.x; idx < per; idx += blockDim.x)
        {
            const int i = (int)(idx / ld), j = (int)(idx - (size_t)i * ld);
            W[idx] = j < n ? a[b * (size_t)n * n + (size_t)i * n + j] : bmat[b * (size_t)n * nrhs + (size_t)i * nrhs + (j - n)];
        }
        __syncthreads();
        int bad = 0;
        for (int k = 0; k < n; ++k)
        {
            const int r = block_argmax(W, ld, k, n, red_v, red_i);
            if (!(red_v[0] >= PIVOT_TOL))
                bad = 1;
            if (r != k)
                for (int j = threadIdx.x; j < ld; j += blockDim.x)
                {
                    const double t = W[(size_t)k * ld + j];
                    W[(size_t)k * ld + j] = W[(size_t)r * ld + j];
                    W[(size_t)r * ld + j] = t;
                }
            __syncthreads();
            const double ip = 1.0 / W[(size_t)k * ld + k];
            // eliminate below: rows i > k, columns j > k
            const int nr = n - k - 1, ncw = ld - k - 1;
            for (int idx = threadIdx.x; idx < nr * ncw; idx += blockDim.x)
            {
                const int i = k + 1 + idx / ncw, j = k + 1 + idx % ncw;
                const double f = W[(size_t)i * ld + k] * ip;
                W[(size_t)i * ld + j] = fma(-f, W[(size_t)k * ld + j], W[(size_t)i * ld + j]);
            }
            __syncthreads();
        }
        // back substitution, one right-hand side per thread
        for (int c = threadIdx.x; c < nrhs; c += blockDim.x)
        {
            for (int i = n - 1; i >= 0; --i)
            {
                double s = W[(size_t)i * ld + n + c];
                for (int j = i + 1; j < n; ++j)
                    s = fma(-W[(size_t)i * ld + j], W[(size_t)j * ld + n + c], s);
                W[(size_t)i * ld + n + c] = s / W[(size_t)i * ld + i];
            }
        }
        __syncthreads();
        for (int idx = threadIdx.x; idx < n * nrhs; idx += blockDim.x)
        {
            const int i = idx / nrhs, c = idx - i * nrhs;
            x[b * (size_t)n * nrhs + idx] = W[(size_t)i * ld + n + c];
        }
        if (status && threadIdx.x == 0)
            status[b] = bad ? MFV2D_PIVOT_FAILED : MFV2D_SUCCESS;
    }
}

// ---- one LARGE definite matrix (the coarse operator of the multilevel preconditioner, precond.cu): in-place Gauss-Jordan
// without pivoting, the whole device working on every elimination step (two launches per pivot)
__global__ void gj_snapshot_kernel(int n, int k, const double *__restrict__ M, double *__restrict__ colk, double *__restrict__ rowk,
                                   int *__restrict__ flag, double tiny)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
    {
        colk[i] = M[(size_t)i * n + k];
        rowk[i] = M[(size_t)k * n + i];
    }
    if (blockIdx.x == 0 && threadIdx.x == 0 && !(fabs(M[(size_t)k * n + k]) > tiny))
        *flag = 1;
}

__global__ void __launch_bounds__(256) gj_update_kernel(int n, int k, double *__restrict__ M, const double *__restrict__ colk,
                                                        const double *__restrict__ rowk)
{
    const double ip = 1.0 / colk[k];
    const size_t nn = (size_t)n * n;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < nn; idx += (size_t)gridDim.x * blockDim.x)
    {
        const int i = (int)(idx / n), j = (int)(idx - (size_t)i * n);
        double v;
        if (i == k)
            v = (j == k) ? ip : rowk[j] * ip;
        else if (j == k)
            v = -colk[i] * ip;
        else
            v = fma(-colk[i] * ip, rowk[j], M[idx]);
        M[idx] = v;
    }
}

// ---- the same inverse, BLOCKED: GJB pivots per pass.  The unblocked version reads and writes the whole matrix once per pivot
// (n = 4096: 268 MB per pivot, 52 us at 5.2 TB/s, 0.22 s in all - HBM bound); here one pass inverts the GJB x GJB diagonal block
// P = D^-1 in shared memory, forms the new row panel R = P M[K, :] and keeps a copy C of the old column panel M[:, K], and one
// tiled rank-GJB update rewrites the matrix:  M[i, j] <- M[i, j] - C[i, :] R[:, j]  (i, j outside K),  M[K, :] <- R,
// M[i, K] <- -C[i, :] P,  M[K, K] <- P.  Traffic drops by the block size; the update is a register-tiled fp64 GEMM.
#define GJB 32
#define GJT 64 // tile of the update kernel: 256 (virtual) threads x 4 x 4 entries

__global__ void __launch_bounds__(256) gjb_diag_kernel(int n, int k0, int bs, const double *__restrict__ M, double *__restrict__ P,
                                                       int *__restrict__ flag, double tiny)
{
    __shared__ double D[GJB][GJB + 1];
    __shared__ double ck[GJB], rk[GJB];
    for (int idx = threadIdx.x; idx < bs * bs; idx += blockDim.x)
        D[idx / bs][idx % bs] = M[(size_t)(k0 + idx / bs) * n + k0 + idx % bs];
    __syncthreads();
    for (int k = 0; k < bs; ++k)
    {
        for (int i = threadIdx.x; i < bs; i += blockDim.x)
        {
            ck[i] = D[i][k];
            rk[i] = D[k][i];
        }
        __syncthreads();
        if (threadIdx.x == 0 && !(fabs(ck[k]) > tiny))
            *flag = 1;
        const double ip = 1.0 / ck[k];
        for (int idx = threadIdx.x; idx < bs * bs; idx += blockDim.x)
        {
            const int i = idx / bs, j = idx - i * bs;
            double v;
            if (i == k)
                v = (j == k) ? ip : rk[j] * ip;
            else if (j == k)
                v = -ck[i] * ip;
            else
                v = fma(-ck[i] * ip, rk[j], D[i][j]);
            D[i][j] = v;
        }
        __syncthreads();
    }
    for (int idx = threadIdx.x; idx < GJB * GJB; idx += blockDim.x)
    {
        const int i = idx / GJB, j = idx - i * GJB;
        P[idx] = (i < bs && j < bs) ? D[i][j] : 0.0;
    }
}

// R[r, j] = sum_s P[r, s] M[k0 + s, j] (columns of K: P itself);  C[i, s] = M[i, k0 + s]
__global__ void gjb_panels_kernel(int n, int k0, int bs, const double *__restrict__ M, const double *__restrict__ P, double *__restrict__ R,
                                  double *__restrict__ Cp)
{
    const size_t total = (size_t)bs * n;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x)
    {
        const int r = (int)(idx / n), j = (int)(idx - (size_t)r * n);
        double v = 0.0;
        if (j >= k0 && j < k0 + bs)
            v = P[r * GJB + (j - k0)];
        else
            for (int s = 0; s < bs; ++s)
                v = fma(P[r * GJB + s], M[(size_t)(k0 + s) * n + j], v);
        R[idx] = v;
    }
    const size_t totc = (size_t)n * GJB;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < totc; idx += (size_t)gridDim.x * blockDim.x)
    {
        const int i = (int)(idx / GJB), s = (int)(idx - (size_t)i * GJB);
        Cp[idx] = s < bs ? M[(size_t)i * n + k0 + s] : 0.0;
    }
}

__global__ void __launch_bounds__(256) gjb_update_kernel(int n, int k0, int bs, double *__restrict__ M, const double *__restrict__ P,
                                                         const double *__restrict__ R, const double *__restrict__ Cp)
{
    __shared__ double Cs[GJT][GJB + 1];
    __shared__ double Ys[GJB][GJT];
    const int tn = (n + GJT - 1) / GJT;
    for (int tile = blockIdx.x; tile < tn * tn; tile += gridDim.x)
    {
        const int i0 = (tile / tn) * GJT, j0 = (tile % tn) * GJT;
        for (int idx = threadIdx.x; idx < GJT * GJB; idx += blockDim.x)
        {
            const int ii = idx / GJB, s = idx - ii * GJB;
            Cs[ii][s] = (i0 + ii < n) ? Cp[(size_t)(i0 + ii) * GJB + s] : 0.0;
        }
        for (int idx = threadIdx.x; idx < GJB * GJT; idx += blockDim.x)
        {
            const int s = idx / GJT, jj = idx - s * GJT, j = j0 + jj;
            double v = 0.0;
            if (j < n && s < bs)
                v = (j >= k0 && j < k0 + bs) ? P[s * GJB + (j - k0)] : R[(size_t)s * n + j];
            Ys[s][jj] = v;
        }
        __syncthreads();
        for (int vt = threadIdx.x; vt < 256; vt += blockDim.x) // one trip on the device; the host emulation walks the 256 tiles
        {
            const int ty = vt >> 4, tx = vt & 15;
            double acc[4][4] = {};
#pragma unroll 8
            for (int s = 0; s < GJB; ++s)
            {
                double av[4], bv[4];
#pragma unroll
                for (int a = 0; a < 4; ++a)
                {
                    av[a] = Cs[ty * 4 + a][s];
                    bv[a] = Ys[s][tx + 16 * a];
                }
#pragma unroll
                for (int a = 0; a < 4; ++a)
#pragma unroll
                    for (int b = 0; b < 4; ++b)
                        acc[a][b] = fma(av[a], bv[b], acc[a][b]);
            }
#pragma unroll
            for (int a = 0; a < 4; ++a)
            {
                const int i = i0 + ty * 4 + a;
                if (i >= n)
                    continue;
                const bool row_in = i >= k0 && i < k0 + bs;
#pragma unroll
                for (int b = 0; b < 4; ++b)
                {
                    const int j = j0 + tx + 16 * b;
                    if (j >= n)
                        continue;
                    double v;
                    if (row_in)
                        v = R[(size_t)(i - k0) * n + j];
                    else if (j >= k0 && j < k0 + bs)
                        v = -acc[a][b];
                    else
                        v = M[(size_t)i * n + j] - acc[a][b];
                    M[(size_t)i * n + j] = v;
                }
            }
        }
        __syncthreads();
    }
}

static int launch_inverse_definite_blocked(mfv2d_ctx *ctx, int n, double *d_M, int *d_flag)
{
    double *scratch = nullptr;
    const size_t np = GJB * GJB, nr = (size_t)GJB * n, nc = (size_t)n * GJB;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, (np + nr + nc) * sizeof(double), (void **)&scratch));
    double *P = scratch, *R = scratch + np, *Cp = R + nr;
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(d_flag, 0, sizeof(int), ctx->stream));
    const int tn = (n + GJT - 1) / GJT;
    unsigned grid_u = (unsigned)(tn * tn), grid_p = (unsigned)((nr + 255) / 256);
    if (grid_u > (unsigned)ctx->sm_count * 8)
        grid_u = (unsigned)ctx->sm_count * 8;
    if (grid_p > (unsigned)ctx->sm_count * 8)
        grid_p = (unsigned)ctx->sm_count * 8;
    for (int k0 = 0; k0 < n; k0 += GJB)
    {
        const int bs = n - k0 < GJB ? n - k0 : GJB;
        MFV2D_LAUNCH(gjb_diag_kernel, 1, 256, 0, ctx->stream, n, k0, bs, (const double *)d_M, P, d_flag, 1e-300);
        MFV2D_LAUNCH(gjb_panels_kernel, grid_p, 256, 0, ctx->stream, n, k0, bs, (const double *)d_M, (const double *)P, R, Cp);
        MFV2D_LAUNCH(gjb_update_kernel, grid_u, 256, 0, ctx->stream, n, k0, bs, d_M, (const double *)P, (const double *)R, (const double *)Cp);
        ctx->launches += 3;
    }
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return mfv2d_cuda_free(ctx, scratch);
}

// M (n x n, definite) <- M^-1 in place; *d_flag (device int, zeroed here) is set when a pivot vanishes
int launch_inverse_definite_large(mfv2d_ctx *ctx, int n, double *d_M, int *d_flag)
{
    {
        const char *env = getenv("MFV2D_INVERSE_BLOCKED"); // =0: one pivot per pass (the first version, kept for comparison)
        if (n >= 2 * GJB && !(env && env[0] == '0'))
            return launch_inverse_definite_blocked(ctx, n, d_M, d_flag);
    }
    double *scratch = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, (size_t)2 * n * sizeof(double), (void **)&scratch));
    MFV2D_CUDA_CHECK(ctx, cudaMemsetAsync(d_flag, 0, sizeof(int), ctx->stream));
    const size_t nn = (size_t)n * n;
    unsigned grid = (unsigned)((nn + 255) / 256);
    if (grid > (unsigned)ctx->sm_count * 8)
        grid = (unsigned)ctx->sm_count * 8;
    for (int k = 0; k < n; ++k)
    {
        MFV2D_LAUNCH(gj_snapshot_kernel, (unsigned)((n + 255) / 256), 256, 0, ctx->stream, n, k, (const double *)d_M, scratch, scratch + n,
                     d_flag, 1e-300);
        MFV2D_LAUNCH(gj_update_kernel, grid, 256, 0, ctx->stream, n, k, d_M, (const double *)scratch, (const double *)(scratch + n));
    }
    ctx->launches += 2 * (uint64_t)n;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return mfv2d_cuda_free(ctx, scratch);
}

// device-pointer entry used by system.cu as well
int launch_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *d_in, double *d_out, int32_t *d_status)
{
    if (batch == 0 || n == 0)
        return MFV2D_SUCCESS;
#ifndef MFV2D_EMU
    {
        // register-resident kernel for blocks up to 128 x 128 (MFV2D_INVERSE_REG=0: the shared-memory kernel)
        const char *env = getenv("MFV2D_INVERSE_REG");
        if (!(env && env[0] == '0') && d_in != d_out)
        {
            if (n <= 16)
                return launch_inverse_reg<16, 1, 1>(ctx, n, batch, d_in, d_out, d_status);
            if (n <= 32)
                return launch_inverse_reg<16, 2, 2>(ctx, n, batch, d_in, d_out, d_status);
            if (n <= 48)
                return launch_inverse_reg<16, 3, 3>(ctx, n, batch, d_in, d_out, d_status);
            if (n <= 64)
                return launch_inverse_reg<16, 4, 4>(ctx, n, batch, d_in, d_out, d_status);
            if (n <= 96)
                return launch_inverse_reg<16, 6, 6>(ctx, n, batch, d_in, d_out, d_status);
            if (n <= 128)
                return launch_inverse_reg<32, 4, 8>(ctx, n, batch, d_in, d_out, d_status);
        }
    }
#endif
    const size_t head = (size_t)2 * n + (n + 1) / 2 + 4;
    const size_t smem_full = (head + (size_t)n * n) * sizeof(double);
    const int use_smem = smem_full <= ctx->smem_optin;
    const size_t smem = use_smem ? smem_full : head * sizeof(double);
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(inverse_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, inverse_kernel, DENSE_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    size_t grid = (size_t)ctx->sm_count * per_sm;
    if (grid > batch)
        grid = batch;
    MFV2D_LAUNCH(inverse_kernel, (unsigned)grid, DENSE_THREADS, smem, ctx->stream, n, (unsigned long long)batch, d_in, d_out,
                 d_status, use_smem);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

extern "C" {

int mfv2d_cuda_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *in, double *out, int32_t *status,
                             int space)
{
    if (n < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "inverse_batch: n must be positive");
    if (batch == 0)
        return MFV2D_SUCCESS;
    const size_t bytes = batch * (size_t)n * n * sizeof(double);
    Staged i, o, st;
    int r = stage_in(ctx, in, bytes, space, &i);
    if (!r)
        r = stage_out_alloc(ctx, out, bytes, space, &o);
    if (!r)
        r = stage_out_alloc(ctx, status, status ? batch * sizeof(int32_t) : 0, space, &st);
    if (!r)
        r = launch_inverse_batch(ctx, n, batch, (const double *)i.dev, (double *)o.dev, (int32_t *)st.dev);
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
    stage_abort(&i);
    return r;
}

int mfv2d_cuda_inverse_definite(mfv2d_ctx *ctx, int n, const double *in, double *out, int32_t *status, int space)
{
    if (status)
        *status = MFV2D_SUCCESS;
    if (n < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "inverse_definite: n must be positive");
    const size_t bytes = (size_t)n * n * sizeof(double);
    Staged i, o;
    int *d_flag = nullptr;
    int r = stage_in(ctx, in, bytes, space, &i);
    if (!r)
        r = stage_out_alloc(ctx, out, bytes, space, &o);
    if (!r)
        r = mfv2d_cuda_malloc(ctx, sizeof(int), (void **)&d_flag);
    if (!r && o.dev != i.dev)
        r = cudaMemcpyAsync(o.dev, i.dev, bytes, cudaMemcpyDeviceToDevice, ctx->stream) == cudaSuccess
                ? MFV2D_SUCCESS
                : mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "inverse_definite: device copy failed");
    if (!r)
        r = launch_inverse_definite_large(ctx, n, (double *)o.dev, d_flag);
    int flag = 0;
    if (!r)
        r = mfv2d_cuda_download(ctx, &flag, d_flag, sizeof flag);
    if (!r)
        r = stage_finish(&o);
    else
        stage_abort(&o);
    stage_abort(&i);
    if (d_flag)
        mfv2d_cuda_free(ctx, d_flag);
    if (!r && flag)
    {
        if (status)
            *status = MFV2D_PIVOT_FAILED;
        return mfv2d_fail(ctx, MFV2D_PIVOT_FAILED, "inverse_definite: a pivot vanished (the matrix is singular or not definite)");
    }
    return r;
}

int mfv2d_cuda_solve_batch(mfv2d_ctx *ctx, int n, int nrhs, size_t batch, const double *a, const double *b, double *x,
                           int32_t *status, int space)
{
    if (n < 1 || nrhs < 1)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "solve_batch: n and nrhs must be positive");
    if (batch == 0)
        return MFV2D_SUCCESS;
    const size_t per = (size_t)n * (n + nrhs);
    const size_t smem_full = per * sizeof(double);
    const int use_smem = smem_full <= ctx->smem_optin;
    const size_t smem = use_smem ? smem_full : 0;
    Staged sa, sb, sx, st;
    void *work = nullptr;
    int r = stage_in(ctx, a, batch * (size_t)n * n * sizeof(double), space, &sa);
    if (!r)
        r = stage_in(ctx, b, batch * (size_t)n * nrhs * sizeof(double), space, &sb);
    if (!r)
        r = stage_out_alloc(ctx, x, batch * (size_t)n * nrhs * sizeof(double), space, &sx);
    if (!r)
        r = stage_out_alloc(ctx, status, status ? batch * sizeof(int32_t) : 0, space, &st);
    size_t grid = 0;
    if (!r)
    {
        cudaFuncSetAttribute(solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        int per_sm = 1;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, solve_kernel, DENSE_THREADS, smem);
        if (per_sm < 1)
            per_sm = 1;
        grid = (size_t)ctx->sm_count * per_sm;
        if (grid > batch)
            grid = batch;
        if (!use_smem)
            r = mfv2d_cuda_malloc(ctx, grid * per * sizeof(double), &work);
    }
    if (!r)
    {
        MFV2D_LAUNCH(solve_kernel, (unsigned)grid, DENSE_THREADS, smem, ctx->stream, n, nrhs, (unsigned long long)batch,
                     (const double *)sa.dev, (const double *)sb.dev, (double *)sx.dev, (double *)work, (int32_t *)st.dev,
                     use_smem);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "solve kernel launch failed");
    }
    if (!r)
    {
        r = stage_finish(&sx);
        const int r2 = stage_finish(&st);
        if (!r)
            r = r2;
    }
    else
    {
        stage_abort(&sx);
        stage_abort(&st);
    }
    stage_abort(&sa);
    stage_abort(&sb);
    if (work)
        mfv2d_cuda_free(ctx, work);
    return r;
}

} // extern "C"
