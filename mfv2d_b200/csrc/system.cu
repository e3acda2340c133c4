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
#include <cmath>

#include "internal.cuh"

int launch_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *d_in, double *d_out, int32_t *d_status);
int comm_allreduce(mfv2d_ctx *ctx, double *dptr, size_t n);

struct mfv2d_system
{
    mfv2d_ctx *ctx;
    int64_t n_elem = 0, n_dense = 0, n_trace = 0, n_con = 0, n_slots = 0;
    int max_n = 0, max_m = 0;
    bool uniform = false; // all blocks have the same size
    // device arrays
    int64_t *d_block_off = nullptr; // [n_elem+1] dense offsets
    int64_t *d_mat_off = nullptr;   // [n_elem+1] offsets into A / Ainv
    double *d_A = nullptr, *d_Ainv = nullptr;
    bool owns_A = true;
    int64_t *d_n_ptr = nullptr, *d_n_col = nullptr; // N   CSR  [n_trace+1], [n_con]
    double *d_n_val = nullptr;
    int64_t *d_t_ptr = nullptr, *d_t_col = nullptr; // N^T CSR  [n_dense+1], [n_con]
    double *d_t_val = nullptr;
    // per element local constraint rows
    int64_t *d_slot_off = nullptr;  // [n_elem+1] first slot of element e (m_e = difference)
    int64_t *d_slot_trace = nullptr; // [n_slots] multiplier of each slot
    int64_t *d_lrow_ptr = nullptr;  // [n_slots+1] entries of each local row
    int32_t *d_lrow_col = nullptr;  // local column index
    double *d_lrow_val = nullptr;
    int64_t *d_s_off = nullptr;     // [n_elem+1] offsets into S
    double *d_S = nullptr;          // condensed blocks, column-major
    int64_t *d_tr_ptr = nullptr, *d_tr_slot = nullptr; // per multiplier: contributing slots
    double *d_slot_val = nullptr;   // [n_slots] scratch: local results
    double *d_diag_inv = nullptr;   // [n_trace] 1 / diag(S) (Jacobi preconditioner on the trace space)
    // work vectors
    double *d_work_dense[3] = {nullptr, nullptr, nullptr};
    double *d_work_trace[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double *d_partials = nullptr; // reduction partials
    int n_partials = 0;
    // multi-GPU interface (comm.cu)
    int64_t n_shared = 0, n_iface = 0;
    int64_t *d_shared_index = nullptr, *d_shared_slot = nullptr;
    double *d_iface = nullptr;   // [n_iface + 8]
    double *d_owned = nullptr;   // [n_trace] 1.0 / 0.0 ownership weights for dot products (nullptr: all owned)
    std::vector<void *> allocs;
};

#define SYS_THREADS 256

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

// S_e[r][s] = sum_{(c,v) in row r} sum_{(d,w) in row s} v w Ainv_e[c][d]; stored column-major (s fastest... r fastest)
__global__ void schur_build_kernel(unsigned long long n_elem, const int64_t *__restrict__ boff, const int64_t *__restrict__ moff,
                                   const double *__restrict__ Ainv, const int64_t *__restrict__ slot_off,
                                   const int64_t *__restrict__ lrow_ptr, const int32_t *__restrict__ lrow_col,
                                   const double *__restrict__ lrow_val, const int64_t *__restrict__ s_off, double *__restrict__ S)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int n = (int)(boff[e + 1] - boff[e]);
        const double *B = Ainv + moff[e];
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        double *Se = S + s_off[e];
        for (int idx = threadIdx.x; idx < m * m; idx += blockDim.x)
        {
            const int s = idx / m, r = idx - s * m; // column s, row r (column-major storage)
            double acc = 0.0;
            for (int64_t a = lrow_ptr[s0 + r]; a < lrow_ptr[s0 + r + 1]; ++a)
            {
                const double *brow = B + (size_t)lrow_col[a] * n;
                double t = 0.0;
                for (int64_t b = lrow_ptr[s0 + s]; b < lrow_ptr[s0 + s + 1]; ++b)
                    t = fma(lrow_val[b], brow[lrow_col[b]], t);
                acc = fma(lrow_val[a], t, acc);
            }
            Se[idx] = acc;
        }
    }
}

// slot_val[s0 + r] = sum_s S_e[r][s] t[trace(s0 + s)]   - one warp per element
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

// slot_val[s0 + r] = S_e[r][r]: the elements' contributions to diag(S)
__global__ void schur_diag_kernel(unsigned long long n_elem, const int64_t *__restrict__ slot_off, const int64_t *__restrict__ s_off,
                                  const double *__restrict__ S, double *__restrict__ slot_val)
{
    for (unsigned long long e = blockIdx.x; e < n_elem; e += gridDim.x)
    {
        const int64_t s0 = slot_off[e];
        const int m = (int)(slot_off[e + 1] - s0);
        const double *Se = S + s_off[e];
        for (int r = threadIdx.x; r < m; r += blockDim.x)
            slot_val[s0 + r] = Se[(size_t)r * m + r];
    }
}

// d[i] = 1 / d[i] (1 where the diagonal vanishes); z[i] = d[i] r[i]
__global__ void jacobi_setup_kernel(long long n, double *__restrict__ d, const double *__restrict__ r, double *__restrict__ z)
{
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double v = d[i];
        const double inv = v != 0.0 ? 1.0 / v : 1.0;
        d[i] = inv;
        z[i] = inv * r[i];
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
__device__ __forceinline__ void block_reduce_store(double acc, double *partials)
{
    __shared__ double red[32];
    acc = warp_sum(acc);
    const int warp = threadIdx.x >> 5, nwarps = MFV2D_NWARPS;
    if ((threadIdx.x & 31) == 0)
        red[warp] = acc;
    __syncthreads();
    if (threadIdx.x == 0)
    {
        double s = 0.0;
        for (int w = 0; w < nwarps; ++w)
            s += red[w];
        partials[blockIdx.x] = s;
    }
    __syncthreads();
}

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

// deterministic sum of the per-CTA partials by ONE warp (launch <<<1, 32>>>): lane-strided, then a butterfly
__device__ __forceinline__ double warp_sum_partials(int n_partials, const double *__restrict__ partials)
{
    double s = 0.0;
    for (int i = MFV2D_LANE0; i < n_partials; i += MFV2D_LANE_STEP)
        s += partials[i];
    return warp_sum(s);
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

// ---- CG on the trace space, scalars on the device --------------------------------------------------
// sc[0] = rr, sc[1] = pAp, sc[2] = alpha, sc[3] = beta, sc[4] = tol^2, sc[5] = rz (preconditioned)
// flags[0] = done, flags[1] = iterations completed (the reference's iter_cnt), flags[2] = breakdown
__global__ void cg_alpha_kernel(int n_partials, const double *__restrict__ partials, double *__restrict__ sc, int *flags,
                                int precond)
{
    const double s = warp_sum_partials(n_partials, partials);
    if (blockIdx.x == 0 && threadIdx.x == 0 && !flags[0])
    {
        sc[1] = s;
        sc[2] = (precond ? sc[5] : sc[0]) / s;
    }
}

// x += alpha p ; r -= alpha Ap ; z = Minv r ; partials of r.r (and r.z)
__global__ void __launch_bounds__(SYS_THREADS) cg_update_kernel(long long n, const double *__restrict__ sc, const int *flags,
                                                               const double *__restrict__ p, const double *__restrict__ ap,
                                                               double *__restrict__ x, double *__restrict__ r,
                                                               const double *__restrict__ w, const double *__restrict__ dinv,
                                                               double *__restrict__ z, double *__restrict__ partials,
                                                               double *__restrict__ partials2)
{
    if (flags[0])
        return;
    const double alpha = sc[2];
    double acc = 0.0, acc2 = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        x[i] = fma(alpha, p[i], x[i]);
        const double ri = fma(-alpha, ap[i], r[i]);
        r[i] = ri;
        const double wi = w ? w[i] : 1.0;
        acc = fma(ri * wi, ri, acc);
        if (dinv)
        {
            const double zi = ri * dinv[i];
            z[i] = zi;
            acc2 = fma(ri * wi, zi, acc2);
        }
    }
    block_reduce_store(acc, partials);
    if (dinv)
        block_reduce_store(acc2, partials2);
}

__global__ void cg_beta_kernel(int n_partials, const double *__restrict__ partials, const double *__restrict__ partials2,
                               double *__restrict__ sc, int *flags, int precond, int iter)
{
    const double s = warp_sum_partials(n_partials, partials);
    const double s2 = precond ? warp_sum_partials(n_partials, partials2) : 0.0;
    if (blockIdx.x == 0 && threadIdx.x == 0 && !flags[0])
    {
        // flags[1] = index of the last iteration carried out (the reference's iter_cnt), counted on the device so
        // that the launches of an iteration do not depend on it (CUDA graph replay); iter < 0 selects the counter
        flags[1] = iter >= 0 ? iter : flags[3];
        flags[3] += 1;
        if (s < sc[4])
        {
            sc[0] = s;
            flags[0] = 1;
            return;
        }
        if (precond)
        {
            sc[3] = s2 / sc[5];
            sc[5] = s2;
        }
        else
            sc[3] = s / sc[0];
        sc[0] = s;
    }
}

// p = (z or r) + beta p
__global__ void cg_direction_kernel(long long n, const double *__restrict__ sc, const int *flags, const double *__restrict__ r,
                                    double *__restrict__ p)
{
    if (flags[0])
        return;
    const double beta = sc[3];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        p[i] = fma(beta, p[i], r[i]);
}

// ---- CG with ONE synchronisation per iteration (Chronopoulos & Gear 1989), used when the system is partitioned -----
// u = Minv r, w = S u;  gamma = r.u, delta = u.w, rho = r.r are needed at the same point, so they travel in the ONE
// all-reduce that also sums the interface multipliers of w (the two-reduction recurrence above pays a second
// all-reduce for r.r after the update):
//   beta = gamma / gamma_old ; alpha = gamma / (delta - beta gamma / alpha_old)
//   p = u + beta p ; s = w + beta s ; x += alpha p ; r -= alpha s ; u = Minv r
// sc[6] = gamma_old, sc[7] = alpha_old, sc[8..10] = (delta, gamma, rho); the other scalars and the flags as above.
__global__ void __launch_bounds__(SYS_THREADS) cg1_dots_kernel(long long n, const double *__restrict__ u, const double *__restrict__ wv,
                                                              const double *__restrict__ r, const double *__restrict__ own,
                                                              double *__restrict__ partials, int n_partials,
                                                              const int *__restrict__ done)
{
    if (*done)
        return;
    double d = 0.0, g = 0.0, rho = 0.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double ui = u[i], ri = r[i];
        const double rw = own ? ri * own[i] : ri;
        d = fma(ui, wv[i], d); // local, un-summed product: summed over the ranks it is u . S u
        g = fma(rw, ui, g);
        rho = fma(rw, ri, rho);
    }
    block_reduce_store(d, partials);
    block_reduce_store(g, partials + n_partials);
    block_reduce_store(rho, partials + 2 * n_partials);
}

__global__ void cg1_sum3_kernel(int n_partials, const double *__restrict__ partials, double *__restrict__ out, const int *done)
{
    if (*done)
        return;
    const double a = warp_sum_partials(n_partials, partials);
    const double b = warp_sum_partials(n_partials, partials + n_partials);
    const double c = warp_sum_partials(n_partials, partials + 2 * n_partials);
    if (threadIdx.x == 0)
    {
        out[0] = a;
        out[1] = b;
        out[2] = c;
    }
}

__global__ void cg1_scalar_kernel(double *__restrict__ sc, int *flags)
{
    if (blockIdx.x != 0 || threadIdx.x != 0 || flags[0])
        return;
    const double delta = sc[8], gamma = sc[9], rho = sc[10];
    const int k = flags[3]; // updates carried out so far: rho belongs to the residual after update k - 1
    flags[3] = k + 1;
    sc[0] = rho;
    if (k > 0)
    {
        flags[1] = k - 1;
        if (rho < sc[4])
        {
            flags[0] = 1;
            return;
        }
    }
    const double beta = k > 0 ? gamma / sc[6] : 0.0;
    const double alpha = k > 0 ? gamma / (delta - beta * gamma / sc[7]) : gamma / delta;
    sc[1] = delta;
    sc[2] = alpha;
    sc[3] = beta;
    sc[6] = gamma;
    sc[7] = alpha;
}

// u may alias r (no preconditioner): every entry is read before it is written by the same thread
__global__ void __launch_bounds__(SYS_THREADS) cg1_update_kernel(long long n, const double *__restrict__ sc, const int *flags,
                                                                const double *u, const double *__restrict__ wv,
                                                                double *__restrict__ p, double *__restrict__ sv,
                                                                double *__restrict__ x, double *r, const double *__restrict__ dinv,
                                                                double *z)
{
    if (flags[0])
        return;
    const double alpha = sc[2], beta = sc[3];
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    {
        const double pi = fma(beta, p[i], u[i]);
        const double si = fma(beta, sv[i], wv[i]);
        p[i] = pi;
        sv[i] = si;
        x[i] = fma(alpha, pi, x[i]);
        const double ri = fma(-alpha, si, r[i]);
        r[i] = ri;
        if (dinv)
            z[i] = ri * dinv[i];
    }
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
template <typename T> static int sys_upload(mfv2d_system *s, const std::vector<T> &h, T **d)
{
    void *p = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(s->ctx, (h.size() ? h.size() : 1) * sizeof(T), &p));
    s->allocs.push_back(p);
    *d = (T *)p;
    if (!h.empty())
        MFV2D_TRY(mfv2d_cuda_upload(s->ctx, p, h.data(), h.size() * sizeof(T)));
    return MFV2D_SUCCESS;
}

template <typename T> static int sys_alloc(mfv2d_system *s, size_t n, T **d)
{
    void *p = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(s->ctx, (n ? n : 1) * sizeof(T), &p));
    s->allocs.push_back(p);
    *d = (T *)p;
    return MFV2D_SUCCESS;
}

static unsigned grid_for(const mfv2d_ctx *ctx, long long n, int threads)
{
    long long g = (n + threads - 1) / threads;
    const long long cap = (long long)ctx->sm_count * 8;
    if (g > cap)
        g = cap;
    if (g < 1)
        g = 1;
    return (unsigned)g;
}

static int sys_gemv(const mfv2d_system *s, const double *M, const double *x, double *y)
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

static int sys_spmv(const mfv2d_system *s, long long rows, const int64_t *ptr, const int64_t *col, const double *val,
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

// sum the shared multipliers of a trace vector over the ranks (no-op on one GPU)
static int sys_interface_sum(const mfv2d_system *s, double *t, int extra_scalars, double *scalars)
{
    mfv2d_ctx *ctx = s->ctx;
    if (ctx->n_ranks <= 1 || !s->d_iface)
        return MFV2D_SUCCESS;
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

static int sys_apply_trace(const mfv2d_system *s, const double *x, double *t)
{
    MFV2D_TRY(sys_spmv(s, s->n_trace, s->d_n_ptr, s->d_n_col, s->d_n_val, x, t));
    return sys_interface_sum(s, t, 0, nullptr);
}

static int sys_apply_schur(const mfv2d_system *s, const double *t_in, double *t_out, const int *done, bool interface_sum = true)
{
    mfv2d_ctx *ctx = s->ctx;
    if (s->n_elem)
    {
        const int warps = SYS_THREADS / 32;
        long long grid = (s->n_elem + warps - 1) / warps;
        const long long cap = (long long)ctx->sm_count * 8;
        if (grid > cap)
            grid = cap;
        MFV2D_LAUNCH(schur_local_kernel, (unsigned)grid, SYS_THREADS, (size_t)warps * (s->max_m + 1 + 32) * sizeof(double), ctx->stream,
                     (unsigned long long)s->n_elem, s->d_slot_off, s->d_slot_trace, s->d_s_off, s->d_S, t_in, s->d_slot_val,
                     s->max_m + 1, done);
        ctx->launches++;
    }
    if (s->n_trace)
    {
        MFV2D_LAUNCH(trace_gather_kernel, grid_for(ctx, s->n_trace, SYS_THREADS), SYS_THREADS, 0, ctx->stream,
                     (long long)s->n_trace, s->d_tr_ptr, s->d_tr_slot, s->d_slot_val, t_out, done);
        ctx->launches++;
    }
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return interface_sum ? sys_interface_sum(s, t_out, 0, nullptr) : MFV2D_SUCCESS;
}

// dot product of two trace (w = ownership weights) or dense (w = nullptr) vectors -> device scalar
static int sys_dot_device(const mfv2d_system *s, long long n, const double *x, const double *y, const double *w, double *d_result)
{
    mfv2d_ctx *ctx = s->ctx;
    MFV2D_LAUNCH(dot_partial_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, n, x, y, w, s->d_partials,
                 (const int *)nullptr);
    MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, s->n_partials, s->d_partials, d_result);
    ctx->launches += 2;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    if (ctx->n_ranks > 1)
        MFV2D_TRY(comm_allreduce(ctx, d_result, 1));
    return MFV2D_SUCCESS;
}

static int read_scalars(mfv2d_ctx *ctx, const double *d, int n, double *h)
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
    // validate ordering and bounds; build CSR of N, CSR of N^T, and the per-element local rows
    std::vector<int64_t> n_ptr(n_trace + 1, 0), t_ptr(s->n_dense + 1, 0);
    for (size_t k = 0; k < n_con; ++k)
    {
        if (con_row[k] < 0 || (size_t)con_row[k] >= n_trace || con_col[k] < 0 || con_col[k] >= s->n_dense)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "system_create: constraint entry %zu is out of range", k));
        if (k && (con_row[k] < con_row[k - 1] || (con_row[k] == con_row[k - 1] && con_col[k] <= con_col[k - 1])))
            SYS_TRY(mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "system_create: constraint entries must be sorted by (row, column) without duplicates"));
        n_ptr[con_row[k] + 1]++;
        t_ptr[con_col[k] + 1]++;
    }
    for (size_t i = 0; i < n_trace; ++i)
        n_ptr[i + 1] += n_ptr[i];
    for (int64_t i = 0; i < s->n_dense; ++i)
        t_ptr[i + 1] += t_ptr[i];
    std::vector<int64_t> n_col(con_col, con_col + n_con), t_col(n_con);
    std::vector<double> n_val(con_val, con_val + n_con), t_val(n_con);
    {
        std::vector<int64_t> fill(t_ptr.begin(), t_ptr.end() - 1);
        for (size_t k = 0; k < n_con; ++k)
        {
            const int64_t pos = fill[con_col[k]]++;
            t_col[pos] = con_row[k];
            t_val[pos] = con_val[k];
        }
    }
    // element of each entry; slots = (element, multiplier) pairs in (element, multiplier) order
    std::vector<int64_t> elem_of(n_con);
    {
        size_t e = 0;
        // columns are not monotone over k, so search
        for (size_t k = 0; k < n_con; ++k)
        {
            const int64_t c = con_col[k];
            if (!(e < n_elem && boff[e] <= c && c < boff[e + 1]))
                e = (size_t)(std::upper_bound(boff.begin(), boff.end(), c) - boff.begin()) - 1;
            elem_of[k] = (int64_t)e;
        }
    }
    // order entries by (element, row, col): stable counting sort by element keeps (row, col) order
    std::vector<int64_t> ecount(n_elem + 1, 0);
    for (size_t k = 0; k < n_con; ++k)
        ecount[elem_of[k] + 1]++;
    for (size_t e = 0; e < n_elem; ++e)
        ecount[e + 1] += ecount[e];
    std::vector<int64_t> order(n_con);
    {
        std::vector<int64_t> fill(ecount.begin(), ecount.end() - 1);
        for (size_t k = 0; k < n_con; ++k)
            order[fill[elem_of[k]]++] = (int64_t)k;
    }
    std::vector<int64_t> slot_off(n_elem + 1, 0), slot_trace, lrow_ptr(1, 0);
    std::vector<int32_t> lrow_col(n_con);
    std::vector<double> lrow_val(n_con);
    for (size_t e = 0; e < n_elem; ++e)
    {
        int64_t prev_row = -1;
        for (int64_t a = ecount[e]; a < ecount[e + 1]; ++a)
        {
            const int64_t k = order[a];
            if (con_row[k] != prev_row)
            {
                if (prev_row >= 0)
                    lrow_ptr.push_back(a);
                slot_trace.push_back(con_row[k]);
                prev_row = con_row[k];
            }
            lrow_col[a] = (int32_t)(con_col[k] - boff[e]);
            lrow_val[a] = con_val[k];
        }
        if (prev_row >= 0)
            lrow_ptr.push_back(ecount[e + 1]);
        slot_off[e + 1] = (int64_t)slot_trace.size();
        const int m = (int)(slot_off[e + 1] - slot_off[e]);
        if (m > s->max_m)
            s->max_m = m;
    }
    s->n_slots = (int64_t)slot_trace.size();
    std::vector<int64_t> s_off(n_elem + 1, 0);
    for (size_t e = 0; e < n_elem; ++e)
    {
        const int64_t m = slot_off[e + 1] - slot_off[e];
        s_off[e + 1] = s_off[e] + m * m;
    }
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
    // ---- upload ----
    SYS_TRY(sys_upload(s, boff, &s->d_block_off));
    SYS_TRY(sys_upload(s, moff, &s->d_mat_off));
    SYS_TRY(sys_upload(s, n_ptr, &s->d_n_ptr));
    SYS_TRY(sys_upload(s, n_col, &s->d_n_col));
    SYS_TRY(sys_upload(s, n_val, &s->d_n_val));
    SYS_TRY(sys_upload(s, t_ptr, &s->d_t_ptr));
    SYS_TRY(sys_upload(s, t_col, &s->d_t_col));
    SYS_TRY(sys_upload(s, t_val, &s->d_t_val));
    SYS_TRY(sys_upload(s, slot_off, &s->d_slot_off));
    SYS_TRY(sys_upload(s, slot_trace, &s->d_slot_trace));
    SYS_TRY(sys_upload(s, lrow_ptr, &s->d_lrow_ptr));
    SYS_TRY(sys_upload(s, lrow_col, &s->d_lrow_col));
    SYS_TRY(sys_upload(s, lrow_val, &s->d_lrow_val));
    SYS_TRY(sys_upload(s, s_off, &s->d_s_off));
    SYS_TRY(sys_upload(s, tr_ptr, &s->d_tr_ptr));
    SYS_TRY(sys_upload(s, tr_slot, &s->d_tr_slot));
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
    SYS_TRY(sys_alloc(s, mat_total, &s->d_Ainv));
    SYS_TRY(sys_alloc(s, (size_t)s_off[n_elem], &s->d_S));
    SYS_TRY(sys_alloc(s, (size_t)s->n_slots, &s->d_slot_val));
    SYS_TRY(sys_alloc(s, (size_t)s->n_trace, &s->d_diag_inv));
    for (int i = 0; i < 3; ++i)
        SYS_TRY(sys_alloc(s, (size_t)s->n_dense, &s->d_work_dense[i]));
    for (int i = 0; i < 6; ++i)
        SYS_TRY(sys_alloc(s, (size_t)s->n_trace, &s->d_work_trace[i]));
    s->n_partials = ctx->sm_count * 4;
    SYS_TRY(sys_alloc(s, (size_t)3 * s->n_partials, &s->d_partials));
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
                     s->d_mat_off, s->d_Ainv, s->d_slot_off, s->d_lrow_ptr, s->d_lrow_col, s->d_lrow_val, s->d_s_off, s->d_S);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess)
            SYS_TRY(mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "schur_build kernel launch failed"));
    }
    SYS_TRY(mfv2d_cuda_synchronize(ctx));
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
    return sys_unary(s, t_in, s->n_trace, t_out, s->n_trace, space,
                     [&](const double *di, double *dout) { return sys_apply_schur(s, di, dout, nullptr); });
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

// solve_schur_iterative (python/mfv2d/solving.py:439-500) with cg_general (:350-436)
int mfv2d_cuda_solve_schur_cg(mfv2d_system *s, const double *rhs, const double *phi, int max_iter, double atol, double rtol,
                              int precondition, double *x, double *lambda, double *residual, int *iterations, int space)
{
    mfv2d_ctx *ctx = s->ctx;
    const long long nt = s->n_trace, nd = s->n_dense;
    Staged srhs, sphi, sx, slam;
    int r = stage_in(ctx, rhs, nd * sizeof(double), space, &srhs);
    if (!r)
        r = stage_in(ctx, phi, nt * sizeof(double), space, &sphi);
    if (!r)
        r = stage_out_alloc(ctx, x, nd * sizeof(double), space, &sx);
    if (!r)
        r = stage_out_alloc(ctx, lambda, nt * sizeof(double), space, &slam);
    double *inv_a_y = s->d_work_dense[0], *tmp_d = s->d_work_dense[1], *tmp_d2 = s->d_work_dense[2];
    double *res = s->d_work_trace[0], *p = s->d_work_trace[1], *ap = s->d_work_trace[2], *z = s->d_work_trace[3];
    double *lam = (double *)slam.dev;
    double *sc = ctx->d_scalars;
    int *flags = reinterpret_cast<int *>(ctx->d_scalars + 16);
    const double *w = s->d_owned;
    double h[24];
    int it_done = 0;
    double res_mag2 = 0.0;
    do
    {
        if (r)
            break;
        // A^-1 y ; N A^-1 y - phi
        if ((r = sys_gemv(s, s->d_Ainv, (const double *)srhs.dev, inv_a_y)))
            break;
        if ((r = sys_apply_trace(s, inv_a_y, res)))
            break;
        if (nt)
        {
            MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nt, 256), 256, 0, ctx->stream, nt, 1.0, (const double *)res, -1.0,
                         (const double *)sphi.dev, res);
            ctx->launches++;
        }
        cudaMemsetAsync(lam, 0, nt * sizeof(double), ctx->stream);
        cudaMemsetAsync(sc, 0, 32 * sizeof(double), ctx->stream);
        // res_mag2 = rhs . rhs ; tolerance (solving.py:405-412)
        if (nt)
        {
            if ((r = sys_dot_device(s, nt, res, res, w, sc + 0)))
                break;
            if ((r = read_scalars(ctx, sc, 1, h)))
                break;
            res_mag2 = h[0];
        }
        const double mag = sqrt(res_mag2);
        const double tol = (mag * rtol > atol) ? atol : mag * rtol;
        h[4] = tol * tol;
        if (nt && max_iter > 0)
        {
            cudaMemcpyAsync(sc + 4, &h[4], sizeof(double), cudaMemcpyHostToDevice, ctx->stream);
            cudaStreamSynchronize(ctx->stream);
            const double *dinv = nullptr;
            if (precondition)
            {
                // Jacobi on the trace space (not in the reference): diag(S) from the diagonals of the condensed blocks,
                // summed per multiplier and over the ranks; z0 = D^-1 r0, p0 = z0, rz = r0 . z0
                long long grid = (long long)ctx->sm_count * 8;
                if (grid > s->n_elem)
                    grid = s->n_elem;
                if (s->n_elem)
                {
                    MFV2D_LAUNCH(schur_diag_kernel, (unsigned)grid, 32, 0, ctx->stream, (unsigned long long)s->n_elem, s->d_slot_off,
                                 s->d_s_off, s->d_S, s->d_slot_val);
                    ctx->launches++;
                }
                MFV2D_LAUNCH(trace_gather_kernel, grid_for(ctx, nt, SYS_THREADS), SYS_THREADS, 0, ctx->stream, nt, s->d_tr_ptr,
                             s->d_tr_slot, s->d_slot_val, s->d_diag_inv, (const int *)nullptr);
                ctx->launches++;
                if ((r = sys_interface_sum(s, s->d_diag_inv, 0, nullptr)))
                    break;
                MFV2D_LAUNCH(jacobi_setup_kernel, grid_for(ctx, nt, 256), 256, 0, ctx->stream, nt, s->d_diag_inv, (const double *)res, z);
                ctx->launches++;
                if ((r = sys_dot_device(s, nt, res, z, w, sc + 5)))
                    break;
                dinv = s->d_diag_inv;
            }
            // partitioned systems take the one-synchronisation recurrence (MFV2D_CG_SINGLE_REDUCE=0 / 1 forces either)
            const char *sr_env = getenv("MFV2D_CG_SINGLE_REDUCE");
            const bool single_reduce = sr_env ? sr_env[0] == '1' : ctx->n_ranks > 1;
            double *sv = s->d_work_trace[4];
            if (single_reduce)
            {
                cudaMemsetAsync(p, 0, nt * sizeof(double), ctx->stream);
                cudaMemsetAsync(sv, 0, nt * sizeof(double), ctx->stream);
            }
            else
                cudaMemcpyAsync(p, precondition ? z : res, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            const int check_every = 16;
            // one CG iteration = 7 small launches; on one GPU a block of check_every iterations is captured ONCE
            // into a CUDA graph and replayed (the kernels stop by themselves once the device flag is set)
            auto enqueue_iteration = [&]() -> int {
                if (single_reduce)
                {
                    const double *u = precondition ? z : res;
                    MFV2D_TRY(sys_apply_schur(s, u, ap, flags, false));
                    MFV2D_LAUNCH(cg1_dots_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, nt, u, (const double *)ap,
                                 (const double *)res, w, s->d_partials, s->n_partials, (const int *)flags);
                    MFV2D_LAUNCH(cg1_sum3_kernel, 1, 32, 0, ctx->stream, s->n_partials, (const double *)s->d_partials, sc + 8,
                                 (const int *)flags);
                    MFV2D_TRY(sys_interface_sum(s, ap, 3, sc + 8));
                    MFV2D_LAUNCH(cg1_scalar_kernel, 1, 32, 0, ctx->stream, sc, flags);
                    MFV2D_LAUNCH(cg1_update_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, nt, (const double *)sc,
                                 (const int *)flags, u, (const double *)ap, p, sv, lam, res, dinv, z);
                    ctx->launches += 4;
                    return MFV2D_SUCCESS;
                }
                if (ctx->n_ranks > 1)
                {
                    // p . S p = sum over the ranks of p_r . (S_r p_r) (local, un-summed products, no ownership weights):
                    // the scalar rides in the SAME all-reduce that sums the interface multipliers of S p
                    MFV2D_TRY(sys_apply_schur(s, p, ap, flags, false));
                    MFV2D_LAUNCH(dot_partial_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, nt, (const double *)ap,
                                 (const double *)p, (const double *)nullptr, s->d_partials, (const int *)flags);
                    MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, s->n_partials, (const double *)s->d_partials,
                                 s->d_partials);
                    MFV2D_TRY(sys_interface_sum(s, ap, 1, s->d_partials));
                    MFV2D_LAUNCH(cg_alpha_kernel, 1, 32, 0, ctx->stream, 1, (const double *)s->d_partials, sc, flags, precondition);
                    ctx->launches++;
                }
                else
                {
                    MFV2D_TRY(sys_apply_schur(s, p, ap, flags));
                    MFV2D_LAUNCH(dot_partial_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, nt, (const double *)ap,
                                 (const double *)p, w, s->d_partials, (const int *)flags);
                    MFV2D_LAUNCH(cg_alpha_kernel, 1, 32, 0, ctx->stream, s->n_partials, (const double *)s->d_partials, sc, flags,
                                 precondition);
                }
                MFV2D_LAUNCH(cg_update_kernel, (unsigned)s->n_partials, SYS_THREADS, 0, ctx->stream, nt, (const double *)sc,
                             (const int *)flags, (const double *)p, (const double *)ap, lam, res, w, dinv, z, s->d_partials,
                             s->d_partials + s->n_partials);
                if (ctx->n_ranks > 1)
                {
                    // r.r (and r.z) of all ranks: two scalars in one all-reduce
                    MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, s->n_partials, (const double *)s->d_partials, sc + 8);
                    MFV2D_LAUNCH(sum_partials_kernel, 1, 32, 0, ctx->stream, precondition ? s->n_partials : 0,
                                 (const double *)(s->d_partials + s->n_partials), sc + 9);
                    MFV2D_TRY(comm_allreduce(ctx, sc + 8, 2));
                    MFV2D_LAUNCH(cg_beta_kernel, 1, 32, 0, ctx->stream, 1, (const double *)(sc + 8), (const double *)(sc + 9), sc, flags,
                                 precondition, -1);
                    ctx->launches += 2;
                }
                else
                    MFV2D_LAUNCH(cg_beta_kernel, 1, 32, 0, ctx->stream, s->n_partials, (const double *)s->d_partials,
                                 (const double *)(s->d_partials + s->n_partials), sc, flags, precondition, -1);
                MFV2D_LAUNCH(cg_direction_kernel, grid_for(ctx, nt, 256), 256, 0, ctx->stream, nt, (const double *)sc,
                             (const int *)flags, (const double *)(precondition ? z : res), p);
                ctx->launches += 5;
                return MFV2D_SUCCESS;
            };
#ifndef MFV2D_EMU
            cudaGraphExec_t graph_exec = nullptr;
            // (with more than one rank the NCCL all-reduces are captured too; MFV2D_CG_GRAPH=0 turns the graph off)
            const char *graph_env = getenv("MFV2D_CG_GRAPH");
            if (!(graph_env && graph_env[0] == '0') && max_iter >= 2 * check_every)
            {
                cudaGraph_t graph = nullptr;
                const uint64_t launches_before = ctx->launches;
                if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess)
                {
                    int rc = MFV2D_SUCCESS;
                    for (int k = 0; k < check_every && rc == MFV2D_SUCCESS; ++k)
                        rc = enqueue_iteration();
                    const cudaError_t ce = cudaStreamEndCapture(ctx->stream, &graph);
                    if (rc != MFV2D_SUCCESS || ce != cudaSuccess || cudaGraphInstantiate(&graph_exec, graph, 0) != cudaSuccess)
                        graph_exec = nullptr;
                    if (graph)
                        cudaGraphDestroy(graph);
                }
                cudaGetLastError();
                ctx->launches = launches_before; // capturing is not launching
            }
#endif
            int it = 0;
            bool finished = false;
            while (it < max_iter && !finished)
            {
                const int stop = (it + check_every < max_iter) ? it + check_every : max_iter;
#ifndef MFV2D_EMU
                if (graph_exec && stop - it == check_every)
                {
                    if (cudaGraphLaunch(graph_exec, ctx->stream) != cudaSuccess)
                    {
                        r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "CG graph launch failed: %s", cudaGetErrorString(cudaGetLastError()));
                        break;
                    }
                    ctx->launches += (single_reduce ? (ctx->n_ranks > 1 ? 8 : 6) : (ctx->n_ranks > 1 ? 13 : 7)) * check_every;
                    it = stop;
                }
#endif
                for (; it < stop && !r; ++it)
                    r = enqueue_iteration();
                if (r)
                    break;
                if (cudaGetLastError() != cudaSuccess)
                {
                    r = mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "CG kernel launch failed");
                    break;
                }
                if ((r = read_scalars(ctx, sc, 20, h)))
                    break;
                int hf[4];
                memcpy(hf, &h[16], sizeof hf);
                res_mag2 = h[0];
                it_done = hf[1];
                finished = hf[0] != 0;
                if (single_reduce && !finished)
                    it_done = it - 1; // the one-synchronisation recurrence tests the residual one product later
            }
#ifndef MFV2D_EMU
            if (graph_exec)
                cudaGraphExecDestroy(graph_exec);
#endif
            if (r)
                break;
        }
        // x = A^-1 y - A^-1 N^T lambda
        if ((r = sys_spmv(s, nd, s->d_t_ptr, s->d_t_col, s->d_t_val, lam, tmp_d)))
            break;
        if ((r = sys_gemv(s, s->d_Ainv, tmp_d, tmp_d2)))
            break;
        if (nd)
        {
            MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)inv_a_y, -1.0,
                         (const double *)tmp_d2, (double *)sx.dev);
            ctx->launches++;
            // one step of iterative refinement of the element-local solves, x += A^-1 (y - N^T lambda - A x): the blocks
            // are applied as explicit inverses, whose residual grows with cond(A_e) (the 1-form and 2-form masses of a
            // small element differ by 1/h^2); MFV2D_SCHUR_REFINE=0 turns it off
            const char *rf_env = getenv("MFV2D_SCHUR_REFINE");
            if (!(rf_env && rf_env[0] == '0'))
            {
                if ((r = sys_gemv(s, s->d_A, (const double *)sx.dev, tmp_d2)))
                    break;
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)srhs.dev, -1.0,
                             (const double *)tmp_d2, tmp_d2);
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)tmp_d2, -1.0,
                             (const double *)tmp_d, tmp_d2);
                if ((r = sys_gemv(s, s->d_Ainv, tmp_d2, inv_a_y)))
                    break;
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)sx.dev, 1.0,
                             (const double *)inv_a_y, (double *)sx.dev);
                ctx->launches += 3;
            }
        }
    } while (0);
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
    if (residual)
        *residual = sqrt(res_mag2);
    if (iterations)
        *iterations = it_done;
    return r;
}

// solve_pcg_iterative (python/mfv2d/solving.py:634-684) with pcg_general (:280-347): preconditioned CG on the
// saddle-point system K = [[A, N^T], [N, 0]] with the block-Jacobi preconditioner A^-1 (+) I.  Full vectors are
// stored as [dense | trace] so that every BLAS-1 step and dot product is one launch; the scalars are read back
// once per dot product (this driver follows the reference's control flow, including its degeneracy test; the
// production path is the Schur-complement CG above).
int mfv2d_cuda_solve_pcg(mfv2d_system *s, const double *rhs, const double *phi, int max_iter, double atol, double rtol, double *x,
                         double *lambda, double *residual, int *iterations, int space)
{
    mfv2d_ctx *ctx = s->ctx;
    if (ctx->n_ranks > 1)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "solve_pcg runs on one GPU (use solve_schur_cg for partitioned systems)");
    const long long nd = s->n_dense, nt = s->n_trace, nf = nd + nt;
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
        double *sc = ctx->d_scalars;
        double h[4];
        const unsigned gf = grid_for(ctx, nf, 256);
        auto apply_system = [&](const double *v, double *out) -> int { // out = K v
            MFV2D_TRY(sys_gemv(s, s->d_A, v, t1));
            MFV2D_TRY(sys_spmv(s, nd, s->d_t_ptr, s->d_t_col, s->d_t_val, v + nd, t2));
            if (nd)
            {
                MFV2D_LAUNCH(axpby_kernel, grid_for(ctx, nd, 256), 256, 0, ctx->stream, nd, 1.0, (const double *)t1, 1.0,
                             (const double *)t2, out);
                ctx->launches++;
            }
            return sys_spmv(s, nt, s->d_n_ptr, s->d_n_col, s->d_n_val, v, out + nd);
        };
        auto precondition = [&](const double *v, double *out) -> int { // out = (A^-1 (+) I) v
            MFV2D_TRY(sys_gemv(s, s->d_Ainv, v, out));
            if (nt)
                MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(out + nd, v + nd, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
            return MFV2D_SUCCESS;
        };
        auto dot = [&](const double *u, const double *v, double *result) -> int {
            MFV2D_TRY(sys_dot_device(s, nf, u, v, nullptr, sc));
            MFV2D_TRY(read_scalars(ctx, sc, 1, h));
            *result = h[0];
            return MFV2D_SUCCESS;
        };
        auto axpby = [&](double a, const double *u, double c, const double *v, double *out) {
            MFV2D_LAUNCH(axpby_kernel, gf, 256, 0, ctx->stream, nf, a, u, c, v, out);
            ctx->launches++;
        };
        do
        {
            cudaMemcpyAsync(b, srhs.dev, nd * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if (nt)
                cudaMemcpyAsync(b + nd, sphi.dev, nt * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            cudaMemsetAsync(xs, 0, nf * sizeof(double), ctx->stream);
            // the initial guess is zero: res = rhs - K 0
            cudaMemcpyAsync(res, b, nf * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if ((r = precondition(res, z)))
                break;
            cudaMemcpyAsync(pv, z, nf * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream);
            if ((r = dot(b, b, &res_mag2)))
                break;
            const double mag = sqrt(res_mag2);
            const double tol = (mag * rtol > atol) ? atol : mag * rtol;
            double rz = 0.0;
            if ((r = dot(res, z, &rz)))
                break;
            for (iter_cnt = 0; iter_cnt < max_iter; ++iter_cnt)
            {
                if ((r = apply_system(pv, ap)))
                    break;
                double apa = 0.0;
                if ((r = dot(ap, pv, &apa)))
                    break;
                if (log(fabs(apa)) - log(res_mag2) < log(1e-12)) // solving.py:331-332
                {
                    r = mfv2d_fail(ctx, MFV2D_UNSPECIFIED_ERROR, "System degenerated (matrix was probably not SPD).");
                    break;
                }
                const double alpha = rz / apa;
                axpby(1.0, xs, alpha, pv, xs);
                axpby(1.0, res, -alpha, ap, res);
                if ((r = dot(res, res, &res_mag2)))
                    break;
                if (res_mag2 < tol * tol)
                    break;
                if ((r = precondition(res, z)))
                    break;
                double new_rz = 0.0;
                if ((r = dot(res, z, &new_rz)))
                    break;
                const double beta = new_rz / rz;
                rz = new_rz;
                axpby(1.0, z, beta, pv, pv);
            }
            if (r)
                break;
            if (iter_cnt == max_iter && max_iter > 0)
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

int mfv2d_cuda_system_set_interface(mfv2d_system *s, size_t n_shared, const int64_t *shared_index, const int64_t *slot,
                                    size_t n_slots, const uint8_t *owned)
{
    std::vector<int64_t> hi(shared_index, shared_index + n_shared), hs(slot, slot + n_shared);
    std::vector<double> w((size_t)s->n_trace);
    for (int64_t i = 0; i < s->n_trace; ++i)
        w[i] = owned[i] ? 1.0 : 0.0;
    s->n_shared = (int64_t)n_shared;
    s->n_iface = (int64_t)n_slots;
    MFV2D_TRY(sys_upload(s, hi, &s->d_shared_index));
    MFV2D_TRY(sys_upload(s, hs, &s->d_shared_slot));
    MFV2D_TRY(sys_upload(s, w, &s->d_owned));
    MFV2D_TRY(sys_alloc(s, n_slots + 8, &s->d_iface));
    return MFV2D_SUCCESS;
}

} // extern "C"
