// Matrix-free evaluation of bytecode programs applied to element DoF vectors
// (compute_element_vector, src/evaluation/element_system.c:245-411) - the "apply_recompute" path of
// y_e = sum_c block(r, c) x_c from the 8 corner doubles and the shared basis tables,
// without ever forming an element matrix.  Mass and interior-product factors are applied by sum
// factorisation (interpolate to the q x q points, scale by the metric weight, project back),
// incidences as signed gathers; only an inverse mass needs a dense matrix, which is built and
// inverted in shared memory.
//
// The same evaluator, fed with unit vectors, is the generic (non-fused) fallback of the matrix path for
// programs whose blocks are not of the form  coef * [E] * D * [E].
#include "dense_device.cuh"

#include <algorithm>
#include <array>

#define APPLY_THREADS 128

// flattened program: [n_terms] then per term: row_form, col_form, coef_index, n_factors, 4 ints per factor
// factor: (type, a, b, c)  type 0: incidence (a = kind) | 1: mass (a = DK_MASS*, b = inverse) |
//                          2: interior product (a = DK_IPROD*, b = field, c = transpose)
static void flatten(const mfv2d_plan *plan, const std::vector<BlockProgram> &blocks, std::vector<int32_t> &prog,
                    std::vector<double> &coef, bool *needs_dense)
{
    const int nf = plan->n_forms;
    prog.clear();
    coef.clear();
    prog.push_back(0);
    int n_terms = 0;
    for (int r = 0; r < nf; ++r)
        for (int c = 0; c < nf; ++c)
        {
            const BlockProgram &bp = blocks[r * nf + c];
            if (!bp.present)
                continue;
            if (bp.identity)
            {
                prog.insert(prog.end(), {r, c, (int32_t)coef.size(), 0});
                coef.push_back(bp.identity_coef);
                ++n_terms;
                continue;
            }
            for (const Term &t : bp.terms)
            {
                prog.insert(prog.end(), {r, c, (int32_t)coef.size(), (int32_t)t.f.size()});
                coef.push_back(t.coef);
                for (const Factor &f : t.f)
                {
                    if (!f.is_dense)
                        prog.insert(prog.end(), {0, f.inc_kind, 0, 0});
                    else if (f.dk <= DK_MASS2)
                    {
                        prog.insert(prog.end(), {1, f.dk, f.inverse, 0});
                        if (f.inverse)
                            *needs_dense = true;
                    }
                    else
                        prog.insert(prog.end(), {2, f.dk, f.field, f.transpose});
                }
                ++n_terms;
            }
        }
    prog[0] = n_terms;
}

static int plan_lower_vector_fused(mfv2d_plan *plan);
static int plan_lower_vector_dmma(mfv2d_plan *plan);

int plan_lower_vector(mfv2d_plan *plan)
{
    std::vector<int32_t> vprog, mprog;
    std::vector<double> vcoef, mcoef;
    bool nd = false;
    flatten(plan, plan->vec_blocks, vprog, vcoef, &nd);
    flatten(plan, plan->mat_blocks, mprog, mcoef, &nd);
    plan->vector_needs_dense = nd;
    // device layout: [vprog | mprog], [vcoef | mcoef]
    plan->vprog = vprog;
    plan->vcoef = vcoef;
    plan->mprog_off = (int)vprog.size();
    plan->mcoef_off = (int)vcoef.size();
    plan->vprog.insert(plan->vprog.end(), mprog.begin(), mprog.end());
    plan->vcoef.insert(plan->vcoef.end(), mcoef.begin(), mcoef.end());
    if (plan->vcoef.empty())
        plan->vcoef.push_back(0.0);
    MFV2D_TRY(mfv2d_cuda_malloc(plan->ctx, plan->vprog.size() * sizeof(int32_t), (void **)&plan->d_vprog));
    MFV2D_TRY(mfv2d_cuda_malloc(plan->ctx, plan->vcoef.size() * sizeof(double), (void **)&plan->d_vcoef));
    MFV2D_TRY(mfv2d_cuda_upload(plan->ctx, plan->d_vprog, plan->vprog.data(), plan->vprog.size() * sizeof(int32_t)));
    MFV2D_TRY(mfv2d_cuda_upload(plan->ctx, plan->d_vcoef, plan->vcoef.data(), plan->vcoef.size() * sizeof(double)));
    MFV2D_TRY(plan_lower_vector_fused(plan));
    return plan_lower_vector_dmma(plan);
}

struct ApplyArgs
{
    const double *tab;
    const int32_t *inc;
    const double *corners;
    const int32_t *prog;
    const double *coef;
    const double *fields[16];
    const double *x; // [n_elem][N]  (unused in column mode)
    double *y;       // [n_elem][N]  or  [n_elem][N][N] in column mode
    unsigned long long n_elem;
    int tab_len, p, q;
    int inc_off[4];
    int n_forms;
    int form_offsets[17];
    int n_total;
    int column_mode; // evaluate the programs on unit vectors and write the columns of the element matrix
    int needs_dense;
    int vmax; // length of the ping-pong vectors
};

struct ApplyView
{
    const double *node, *edge, *pnts, *wgts;
    double *geo;    // 5 q^2
    double *u0, *u1; // q^2 each: values at the quadrature points
    double *tmp;    // 2 (p+1) q: intermediate of the two-stage contractions
    double *v0, *v1; // ping-pong coefficient vectors
    double *xs, *ys; // element input / output vectors (N each)
};

// u[a][b] (+)= sum_i A[i][a] sum_j c[i*nB + j] B[j][b]
__device__ __forceinline__ void interp(const ApplyView &s, int q, const double *__restrict__ c, const double *__restrict__ A,
                                       int nA, const double *__restrict__ B, int nB, double *__restrict__ u)
{
    for (int idx = threadIdx.x; idx < nA * q; idx += blockDim.x)
    {
        const int i = idx / q, b = idx - i * q;
        double r = 0.0;
        for (int j = 0; j < nB; ++j)
            r = fma(c[i * nB + j], B[j * q + b], r);
        s.tmp[idx] = r;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < q * q; idx += blockDim.x)
    {
        const int a = idx / q, b = idx - a * q;
        double r = 0.0;
        for (int i = 0; i < nA; ++i)
            r = fma(A[i * q + a], s.tmp[i * q + b], r);
        u[idx] = r;
    }
    __syncthreads();
}

// out[i*nB + j] = sum_a A[i][a] sum_b B[j][b] g[a][b]
__device__ __forceinline__ void project(const ApplyView &s, int q, const double *__restrict__ g, const double *__restrict__ A,
                                        int nA, const double *__restrict__ B, int nB, double *__restrict__ out)
{
    for (int idx = threadIdx.x; idx < nA * q; idx += blockDim.x)
    {
        const int i = idx / q, b = idx - i * q;
        double r = 0.0;
        for (int a = 0; a < q; ++a)
            r = fma(A[i * q + a], g[a * q + b], r);
        s.tmp[idx] = r;
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < nA * nB; idx += blockDim.x)
    {
        const int i = idx / nB, j = idx - i * nB;
        double r = 0.0;
        for (int b = 0; b < q; ++b)
            r = fma(B[j * q + b], s.tmp[i * q + b], r);
        out[idx] = r;
    }
    __syncthreads();
}

// pointwise metric scaling between interpolation and projection.  mode selects the weights:
//  0: u0 *= det W                     (M0)          1: u0 *= W / det                 (M2)
//  2: (u0,u1) <- (khh u0 + khv u1, khv u0 + kvv u1) W / det            (M1; u0 = h part, u1 = v part)
//  3: u0 <- kh u0 + kv u1  (P1: 1-form in)         4: (u0,u1) <- (kh u0, kv u0)     (P1^T: 0-form in)
//  5: (u0,u1) <- (kh u0, kv u0)  (P2: 2-form in)   6: u0 <- kh u0 + kv u1           (P2^T: 1-form in)
__device__ __forceinline__ void scale_points(const ApplyView &s, int q, int mode, const double *__restrict__ fld)
{
    const int q2 = q * q;
    for (int idx = threadIdx.x; idx < q2; idx += blockDim.x)
    {
        const int a = idx / q, b = idx - a * q;
        const double w = s.wgts[a] * s.wgts[b];
        const double j00 = s.geo[idx], j01 = s.geo[q2 + idx], j10 = s.geo[2 * q2 + idx], j11 = s.geo[3 * q2 + idx],
                     det = s.geo[4 * q2 + idx];
        const double a0 = s.u0[idx];
        switch (mode)
        {
        case 0:
            s.u0[idx] = a0 * det * w;
            break;
        case 1:
            s.u0[idx] = a0 * w / det;
            break;
        case 2: {
            const double a1 = s.u1[idx];
            const double khh = (j11 * j11 + j10 * j10) / det * w, kvv = (j00 * j00 + j01 * j01) / det * w,
                         khv = (j01 * j11 + j00 * j10) / det * w;
            s.u0[idx] = khh * a0 + khv * a1;
            s.u1[idx] = khv * a0 + kvv * a1;
            break;
        }
        case 3:
        case 4: {
            const double fx = fld[2 * idx], fy = fld[2 * idx + 1];
            const double kh = (fx * j11 - fy * j10) * w, kv = -(fy * j00 - fx * j01) * w;
            if (mode == 3)
                s.u0[idx] = kh * a0 + kv * s.u1[idx];
            else
            {
                s.u0[idx] = kh * a0;
                s.u1[idx] = kv * a0;
            }
            break;
        }
        default: {
            const double fx = fld[2 * idx], fy = fld[2 * idx + 1];
            const double kh = -(fx * j10 + fy * j11) / det * w, kv = -(fx * j00 + fy * j01) / det * w;
            if (mode == 6)
                s.u0[idx] = kh * a0 + kv * s.u1[idx];
            else
            {
                s.u0[idx] = kh * a0;
                s.u1[idx] = kv * a0;
            }
            break;
        }
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(APPLY_THREADS) element_vector_kernel(const ApplyArgs args)
{
    MFV2D_DYN_SMEM(double, smem);
    const int p = args.p, q = args.q, q2 = q * q, np1 = p + 1, nh = p * (p + 1);
    const int N = args.n_total;
    ApplyView s;
    s.node = smem;
    s.edge = s.node + np1 * q;
    s.pnts = s.edge + p * q;
    s.wgts = s.pnts + q;
    double *w = smem + ((args.tab_len + 1) & ~1);
    s.geo = w;
    w += 5 * q2;
    s.u0 = w;
    w += q2;
    s.u1 = w;
    w += q2;
    s.tmp = w;
    w += 2 * np1 * q;
    s.v0 = w;
    w += args.vmax;
    s.v1 = w;
    w += args.vmax;
    s.xs = w;
    w += N;
    s.ys = w;
    w += N;
    // dense part (only when an inverse mass appears): the assembly kernel's scratch + one matrix
    SmemView dv;
    dv.tab = smem;
    dv.node = smem;
    dv.edge = smem + np1 * q;
    dv.pnts = dv.edge + p * q;
    dv.wgts = dv.pnts + q;
    dv.geo = s.geo;
    w = smem + (((int)(w - smem) + 1) & ~1);
    dv.G = w;
    w += (q2 + 1) & ~1;
    dv.ldT = (np1 * np1 + 3) & ~3;
    dv.T = w;
    w += q * dv.ldT + 4;
    dv.arena = w;
    __shared__ double cs[8];

    for (int i = threadIdx.x; i < args.tab_len; i += blockDim.x)
        smem[i] = args.tab[i];

    const int n_terms = args.prog[0];
    const int n_cols = args.column_mode ? N : 1;

    for (unsigned long long elem = blockIdx.x; elem < args.n_elem; elem += gridDim.x)
    {
        __syncthreads();
        for (int i = threadIdx.x; i < 8; i += blockDim.x)
            cs[i] = args.corners[elem * 8 + i];
        __syncthreads();
        fill_geometry(s.geo, cs, s.pnts, q);
        for (int col = 0; col < n_cols; ++col)
        {
            __syncthreads();
            for (int i = threadIdx.x; i < N; i += blockDim.x)
            {
                s.xs[i] = args.column_mode ? (i == col ? 1.0 : 0.0) : args.x[elem * N + i];
                s.ys[i] = 0.0;
            }
            __syncthreads();
            int pc = 1;
            for (int t = 0; t < n_terms; ++t)
            {
                const int rf = args.prog[pc], cf = args.prog[pc + 1], ci = args.prog[pc + 2], nfac = args.prog[pc + 3];
                pc += 4;
                const int c0 = args.form_offsets[cf], nc = args.form_offsets[cf + 1] - c0;
                const int r0 = args.form_offsets[rf];
                double *cur = s.v0, *nxt = s.v1;
                int n_cur = nc;
                for (int i = threadIdx.x; i < nc; i += blockDim.x)
                    cur[i] = s.xs[c0 + i];
                __syncthreads();
                for (int f = 0; f < nfac; ++f, pc += 4)
                {
                    const int type = args.prog[pc], fa = args.prog[pc + 1], fb = args.prog[pc + 2], fc = args.prog[pc + 3];
                    int n_out;
                    if (type == 0)
                    {
                        // incidence: signed gather through the row map of E
                        const int rows = (fa == MFV2D_INC_10 || fa == MFV2D_INC_21_T) ? 2 * nh : (fa == MFV2D_INC_21 ? p * p : np1 * np1);
                        const int4 *map = reinterpret_cast<const int4 *>(args.inc) + args.inc_off[fa];
                        for (int r = threadIdx.x; r < rows; r += blockDim.x)
                        {
                            const int4 e = ld_ro(map + r);
                            const int ev[4] = {e.x, e.y, e.z, e.w};
                            double acc = 0.0;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                            {
                                if (ev[k] == 0)
                                    break;
                                const double v = cur[abs(ev[k]) - 1];
                                acc += ev[k] < 0 ? -v : v;
                            }
                            nxt[r] = acc;
                        }
                        __syncthreads();
                        n_out = rows;
                    }
                    else if (type == 1 && fb)
                    {
                        // inverse mass: dense matrix in shared memory, inverted in place, then a mat-vec
                        DevDense dd{fa, 1, -1, 0, 0, 0, 0};
                        dd.rows = dd.cols = fa == DK_MASS0 ? np1 * np1 : (fa == DK_MASS1 ? 2 * nh : p * p);
                        int bad = 0;
                        __syncthreads();
                        compute_dense(dv, dd, p, q, args.fields, elem, &bad);
                        __syncthreads();
                        const int n = dd.rows;
                        for (int r = threadIdx.x; r < n; r += blockDim.x)
                        {
                            double acc = 0.0;
                            for (int k = 0; k < n; ++k)
                                acc = fma(dv.arena[r * n + k], cur[k], acc);
                            nxt[r] = acc;
                        }
                        __syncthreads();
                        n_out = n;
                    }
                    else if (type == 1)
                    {
                        if (fa == DK_MASS0)
                        {
                            interp(s, q, cur, s.node, np1, s.node, np1, s.u0);
                            scale_points(s, q, 0, nullptr);
                            project(s, q, s.u0, s.node, np1, s.node, np1, nxt);
                            n_out = np1 * np1;
                        }
                        else if (fa == DK_MASS2)
                        {
                            interp(s, q, cur, s.edge, p, s.edge, p, s.u0);
                            scale_points(s, q, 1, nullptr);
                            project(s, q, s.u0, s.edge, p, s.edge, p, nxt);
                            n_out = p * p;
                        }
                        else
                        {
                            interp(s, q, cur, s.node, np1, s.edge, p, s.u0);      // h part
                            interp(s, q, cur + nh, s.edge, p, s.node, np1, s.u1); // v part
                            scale_points(s, q, 2, nullptr);
                            project(s, q, s.u0, s.node, np1, s.edge, p, nxt);
                            project(s, q, s.u1, s.edge, p, s.node, np1, nxt + nh);
                            n_out = 2 * nh;
                        }
                    }
                    else
                    {
                        const double *fld = args.fields[fb] + elem * (unsigned long long)(2 * q2);
                        if (fa == DK_IPROD1 && !fc)
                        { // (n0 x n1): 1-form in, 0-form out
                            interp(s, q, cur, s.node, np1, s.edge, p, s.u0);
                            interp(s, q, cur + nh, s.edge, p, s.node, np1, s.u1);
                            scale_points(s, q, 3, fld);
                            project(s, q, s.u0, s.node, np1, s.node, np1, nxt);
                            n_out = np1 * np1;
                        }
                        else if (fa == DK_IPROD1)
                        { // transposed: 0-form in, 1-form out
                            interp(s, q, cur, s.node, np1, s.node, np1, s.u0);
                            scale_points(s, q, 4, fld);
                            project(s, q, s.u0, s.node, np1, s.edge, p, nxt);
                            project(s, q, s.u1, s.edge, p, s.node, np1, nxt + nh);
                            n_out = 2 * nh;
                        }
                        else if (!fc)
                        { // (n1 x n2): 2-form in, 1-form out
                            interp(s, q, cur, s.edge, p, s.edge, p, s.u0);
                            scale_points(s, q, 5, fld);
                            project(s, q, s.u0, s.node, np1, s.edge, p, nxt);
                            project(s, q, s.u1, s.edge, p, s.node, np1, nxt + nh);
                            n_out = 2 * nh;
                        }
                        else
                        { // transposed: 1-form in, 2-form out
                            interp(s, q, cur, s.node, np1, s.edge, p, s.u0);
                            interp(s, q, cur + nh, s.edge, p, s.node, np1, s.u1);
                            scale_points(s, q, 6, fld);
                            project(s, q, s.u0, s.edge, p, s.edge, p, nxt);
                            n_out = p * p;
                        }
                    }
                    double *sw = cur;
                    cur = nxt;
                    nxt = sw;
                    n_cur = n_out;
                }
                const double cf_ = args.coef[ci];
                for (int i = threadIdx.x; i < n_cur; i += blockDim.x)
                    s.ys[r0 + i] = fma(cf_, cur[i], s.ys[r0 + i]);
                __syncthreads();
            }
            if (args.column_mode)
            {
                double *o = args.y + elem * (unsigned long long)N * N;
                for (int i = threadIdx.x; i < N; i += blockDim.x)
                    o[(unsigned long long)i * N + col] = s.ys[i];
            }
            else
            {
                for (int i = threadIdx.x; i < N; i += blockDim.x)
                    args.y[elem * N + i] = s.ys[i];
            }
        }
    }
}

// =====================================================================================================
// Tensor terms.  A dense factor (mass / interior product) acts on a form through its tensor components: the
// coefficient (i, j) of a component multiplies A_i(eta) B_j(xi) with A, B the nodal or the edge 1-D basis.  The
// fast kernels below (fused DFMA kernel, tensor-core kernel) are built on this description; programs that do not
// have the tensor shape (inverse mass inside a vector, identity blocks, two dense factors in one term) and the
// column mode use the CTA-per-element kernel above.
// =====================================================================================================
#define APPLYT_THREADS 256

enum
{
    TK_NODE = 0,
    TK_EDGE = 1
};

// one tensor component of a form: coefficient (i, j) multiplies A_i(eta) B_j(xi); stored at off + i nB + j
struct TileComp
{
    int kA, kB, off;
};

struct TileOp
{
    int n_in, n_out, mode;
    TileComp in[2], out[2];
};

__host__ __device__ __forceinline__ TileOp tile_op(int type, int fa, int fc, int p)
{
    const int nh = p * (p + 1);
    const TileComp c0{TK_NODE, TK_NODE, 0}, c2{TK_EDGE, TK_EDGE, 0}, ch{TK_NODE, TK_EDGE, 0}, cv{TK_EDGE, TK_NODE, nh};
    TileOp op;
    op.in[1] = op.out[1] = c0;
    if (type == 1)
    {
        if (fa == DK_MASS0)
        {
            op.n_in = op.n_out = 1;
            op.mode = 0;
            op.in[0] = op.out[0] = c0;
        }
        else if (fa == DK_MASS2)
        {
            op.n_in = op.n_out = 1;
            op.mode = 1;
            op.in[0] = op.out[0] = c2;
        }
        else
        {
            op.n_in = op.n_out = 2;
            op.mode = 2;
            op.in[0] = op.out[0] = ch;
            op.in[1] = op.out[1] = cv;
        }
    }
    else if (fa == DK_IPROD1 && !fc) // (n0 x n1): 1-form in, 0-form out
    {
        op.n_in = 2;
        op.n_out = 1;
        op.mode = 3;
        op.in[0] = ch;
        op.in[1] = cv;
        op.out[0] = c0;
    }
    else if (fa == DK_IPROD1) // transposed: 0-form in, 1-form out
    {
        op.n_in = 1;
        op.n_out = 2;
        op.mode = 4;
        op.in[0] = c0;
        op.out[0] = ch;
        op.out[1] = cv;
    }
    else if (!fc) // (n1 x n2): 2-form in, 1-form out
    {
        op.n_in = 1;
        op.n_out = 2;
        op.mode = 5;
        op.in[0] = c2;
        op.out[0] = ch;
        op.out[1] = cv;
    }
    else // transposed: 1-form in, 2-form out
    {
        op.n_in = 2;
        op.n_out = 1;
        op.mode = 6;
        op.in[0] = ch;
        op.in[1] = cv;
        op.out[0] = c2;
    }
    return op;
}


// =====================================================================================================
// Fused matrix-free apply.  When every term of a vector program has the shape  coef * [E_L] * D * [E_R]
// (D a mass or an interior-product matrix, at most one incidence on each side - all the reference's example
// systems do), the terms do not depend on each other and are evaluated TOGETHER in five barrier-separated
// phases per pass of `epb` elements instead of one chain after the other:
//   pre   right incidences: signed gathers of x into the terms' input vectors
//   A     thread (e, row pair):      t[i][:] = sum_j c[i, j] B_j[:]                (interpolate along xi)
//   B     thread (e, b):             for every term (pair): interpolate along eta, metric (one division per
//                                    point and pair), project along eta - the values at the points stay in registers
//   C     thread (e, row pair):      out[i, j] = B'_j[:] . t'[i][:]                (project along xi)
//   post  thread (e, k):             y[k] = sum_terms coef * (out[k] | left incidence of out), stored to HBM
// The kernel is bound by shared-memory wavefronts (one 16-byte table read feeds two DFMAs), so rows that use the
// same 1-D table are processed two at a time in A / C and terms of the same kind two at a time in B: every table
// read then feeds four DFMAs, and the metric of a pair is computed once.  x and the corners of the next pass
// are prefetched into registers under phases C / post.  All descriptors are resolved on the host
// (plan_lower_vector_fused) into one int blob staged in shared memory.
// =====================================================================================================
static int inc_rows_of(int kind, int p)
{
    const int nh = p * (p + 1);
    return (kind == MFV2D_INC_10 || kind == MFV2D_INC_21_T) ? 2 * nh : (kind == MFV2D_INC_21 ? p * p : (p + 1) * (p + 1));
}

struct FusedRow // one 1-D contraction row of phase A (src = coefficients, dst = t row) or C (src = dst = t row)
{
    int src, nB, kB, dst;
    int g; // phase A: -1, or first row of the right incidence whose signed gathers of x (at src) are the coefficients
};

// rows with the same table (kB) two at a time; 8 ints per entry: src0, nB, kB, dst0, src1 (-1: single), dst1, g0, g1
static void pair_rows(std::vector<FusedRow> rows, std::vector<int32_t> &out)
{
    std::stable_sort(rows.begin(), rows.end(), [](const FusedRow &a, const FusedRow &b) {
        return (a.g >= 0) != (b.g >= 0) ? (a.g >= 0) < (b.g >= 0) : a.kB < b.kB; // gathered rows together: uniform warps
    });
    std::vector<int32_t> singles;
    for (size_t i = 0; i < rows.size();)
    {
        if (i + 1 < rows.size() && rows[i + 1].kB == rows[i].kB && (rows[i + 1].g >= 0) == (rows[i].g >= 0))
        {
            out.insert(out.end(), {rows[i].src, rows[i].nB, rows[i].kB, rows[i].dst, rows[i + 1].src, rows[i + 1].dst, rows[i].g, rows[i + 1].g});
            i += 2;
        }
        else
        {
            singles.insert(singles.end(), {rows[i].src, rows[i].nB, rows[i].kB, rows[i].dst, -1, -1, rows[i].g, -1});
            i += 1;
        }
    }
    out.insert(out.end(), singles.begin(), singles.end());
}

struct FT // one term  coef * [E_L] * D * [E_R]  of a vector program
{
    int rf, cf, ci, L, R;
    TileOp op;
    int fld, dk, tr;
};

// all terms of the vector program in evaluation order; false if one of them does not have the tensor shape
static bool collect_tensor_terms(const mfv2d_plan *plan, std::vector<FT> &fts)
{
    const int p = plan->p, nf = plan->n_forms;
    int ci = 0;
    for (int r = 0; r < nf; ++r)
        for (int c = 0; c < nf; ++c)
        {
            const BlockProgram &bp = plan->vec_blocks[r * nf + c];
            if (!bp.present)
                continue;
            if (bp.identity)
                return false; // not a tensor term: the term-by-term kernel handles it
            for (const Term &t : bp.terms)
            {
                FT ft{r, c, ci++, -1, -1, TileOp{}, 0, 0, 0};
                int dense_at = -1;
                for (size_t i = 0; i < t.f.size(); ++i)
                    if (t.f[i].is_dense)
                    {
                        if (dense_at >= 0)
                            return false;
                        dense_at = (int)i;
                    }
                if (dense_at < 0 || dense_at > 1 || (int)t.f.size() - dense_at > 2)
                    return false;
                const Factor &d = t.f[dense_at];
                if (d.inverse)
                    return false;
                if (dense_at == 1)
                    ft.R = t.f[0].inc_kind;
                if ((int)t.f.size() == dense_at + 2)
                    ft.L = t.f[dense_at + 1].inc_kind;
                ft.op = tile_op(d.dk <= DK_MASS2 ? 1 : 2, d.dk, d.transpose, p);
                ft.fld = d.dk <= DK_MASS2 ? 0 : d.field;
                ft.dk = d.dk;
                ft.tr = d.dk <= DK_MASS2 ? 0 : d.transpose;
                fts.push_back(ft);
            }
        }
    return !fts.empty() && fts.size() <= 64;
}

static int plan_lower_vector_fused(mfv2d_plan *plan)
{
    plan->vector_fused = false;
    const int p = plan->p, q = plan->q, np1 = p + 1, nf = plan->n_forms, N = plan->n_total;
    if (plan->vector_needs_dense || q < 2 || q > 16)
        return MFV2D_SUCCESS;
    const int ldt = asm_ldq(q); // t rows: 16-byte row accesses of consecutive rows fall into distinct bank groups
    auto ksize = [&](int k) { return k == TK_NODE ? np1 : p; };
    std::vector<FT> fts;
    if (!collect_tensor_terms(plan, fts))
        return MFV2D_SUCCESS;

    // slot layout: corners[2][8] (double buffered) | x[N] | t rows of all terms.  The results of phase C overwrite
    // the head of their own t row, so neither the inputs after a right incidence nor the outputs need vectors
    std::vector<int> t_off(fts.size()), n_in(fts.size()), n_out(fts.size());
    int t_total = 0;
    for (size_t i = 0; i < fts.size(); ++i)
    {
        const TileOp &op = fts[i].op;
        int ni = 0, no = 0, ri = 0, ro = 0;
        for (int k = 0; k < op.n_in; ++k)
        {
            ni += ksize(op.in[k].kA) * ksize(op.in[k].kB);
            ri += ksize(op.in[k].kA);
        }
        for (int k = 0; k < op.n_out; ++k)
        {
            no += ksize(op.out[k].kA) * ksize(op.out[k].kB);
            ro += ksize(op.out[k].kA);
        }
        n_in[i] = ni;
        n_out[i] = no;
        const int c0 = plan->form_offsets[fts[i].cf];
        if (fts[i].R >= 0 ? inc_rows_of(fts[i].R, p) != ni : plan->form_offsets[fts[i].cf + 1] - c0 != ni)
            return MFV2D_SUCCESS;
        t_off[i] = t_total;
        t_total += (ri > ro ? ri : ro) * ldt;
    }
    FusedVecInfo fv{};
    fv.n_terms = (int)fts.size();
    fv.off_vec = 16;
    fv.off_t = fv.off_vec + ((N + 2) & ~1); // at least one pad cell after x: it holds 0.0 (target of unused gather entries)
    const int zero_cell = fv.off_vec + N;
    fv.slot_doubles = fv.off_t + t_total;
    // thread (e, b) of phase B reads slot[e][... + b]: with the slot stride congruent to q modulo 16 doubles the
    // addresses of a warp are congruent to the thread index, i.e. conflict free (even stride keeps 16-byte alignment)
    fv.slot_doubles += 6; // room for the per-element rotation of the slot base (0, 2, 4 or 6 doubles)
    while ((fv.slot_doubles & 15) != ((q + (q & 1)) & 15))
        ++fv.slot_doubles;
    if (fv.slot_doubles > 32000)
        return MFV2D_SUCCESS; // slot addresses are packed into 16 bits

    std::vector<int32_t> rowA, termB, rowC, yrow((size_t)2 * N, 0), contrib, gtab;
    // signed sum of up to four cells as (p0 + p1) - (n0 + n1): positive entries first, the zero cell where there is none
    auto split_signs = [&](const int32_t *ev, const std::vector<int> *addr_of, int base, int32_t out[4]) -> bool {
        int np = 0, nn = 0;
        out[0] = out[1] = out[2] = out[3] = zero_cell;
        for (int z = 0; z < 4 && ev[z] != 0; ++z)
        {
            const int idx = (ev[z] < 0 ? -ev[z] : ev[z]) - 1;
            const int a = addr_of ? (*addr_of)[idx] : base + idx;
            if (ev[z] > 0)
            {
                if (np == 2)
                    return false;
                out[np++] = a;
            }
            else
            {
                if (nn == 2)
                    return false;
                out[2 + nn++] = a;
            }
        }
        return true;
    };
    std::vector<FusedRow> rowsA, rowsC;
    std::vector<std::vector<std::array<int32_t, 8>>> per_row(N);
    const int x_base = fv.off_vec, t_base = fv.off_t;
    for (size_t i = 0; i < fts.size(); ++i)
    {
        const FT &ft = fts[i];
        const TileOp &op = ft.op;
        const int c0 = plan->form_offsets[ft.cf], r0 = plan->form_offsets[ft.rf];
        int row = 0;
        for (int k = 0; k < op.n_in; ++k)
        {
            const int nA = ksize(op.in[k].kA), nB = ksize(op.in[k].kB);
            for (int a = 0; a < nA; ++a, ++row)
            {
                const int first = op.in[k].off + a * nB; // index of the row's first coefficient in the term's input vector
                if (ft.R >= 0)
                {
                    rowsA.push_back({x_base + c0, nB, op.in[k].kB, t_base + t_off[i] + row * ldt, (int)gtab.size() / 4});
                    for (int j = 0; j < nB; ++j)
                    {
                        int32_t e4[4];
                        if (!split_signs(plan->basis->h_inc.data() + 4 * (size_t)(plan->basis->inc_off[ft.R] + first + j), nullptr, x_base + c0, e4))
                            return MFV2D_SUCCESS;
                        gtab.insert(gtab.end(), e4, e4 + 4);
                    }
                }
                else
                    rowsA.push_back({x_base + c0 + first, nB, op.in[k].kB, t_base + t_off[i] + row * ldt, -1});
            }
        }
        // address of entry m of the term's output vector: head of the t row of its (component, row)
        std::vector<int> out_addr(n_out[i]);
        row = 0;
        for (int k = 0; k < op.n_out; ++k)
        {
            const int nA = ksize(op.out[k].kA), nB = ksize(op.out[k].kB);
            for (int a = 0; a < nA; ++a, ++row)
            {
                const int tr = t_base + t_off[i] + row * ldt;
                rowsC.push_back({tr, nB, op.out[k].kB, tr, -1});
                for (int j = 0; j < nB; ++j)
                    out_addr[op.out[k].off + a * nB + j] = tr + j;
            }
        }
        const int n_rows_form = plan->form_offsets[ft.rf + 1] - r0;
        if (ft.L >= 0)
        {
            if (inc_rows_of(ft.L, p) != n_rows_form)
                return MFV2D_SUCCESS;
            const int32_t *map = plan->basis->h_inc.data() + 4 * (size_t)plan->basis->inc_off[ft.L];
            for (int k = 0; k < n_rows_form; ++k)
            {
                int32_t e4[4];
                if (!split_signs(map + 4 * k, &out_addr, 0, e4))
                    return MFV2D_SUCCESS;
                per_row[r0 + k].push_back({ft.ci, e4[0], e4[1], e4[2], e4[3], 1, 0, 0});
            }
        }
        else
        {
            if (n_out[i] != n_rows_form)
                return MFV2D_SUCCESS;
            for (int k = 0; k < n_rows_form; ++k)
                per_row[r0 + k].push_back({ft.ci, out_addr[k], zero_cell, zero_cell, zero_cell, 0, 0, 0});
        }
    }
    // phase B: terms with the same dense factor (kind, field, transposition) two at a time
    {
        std::vector<char> used(fts.size(), 0);
        for (size_t i = 0; i < fts.size(); ++i)
        {
            if (used[i])
                continue;
            used[i] = 1;
            int partner = -1;
            for (size_t j = i + 1; j < fts.size(); ++j)
                if (!used[j] && fts[j].dk == fts[i].dk && fts[j].fld == fts[i].fld && fts[j].tr == fts[i].tr)
                {
                    partner = (int)j;
                    used[j] = 1;
                    break;
                }
            const TileOp &op = fts[i].op;
            termB.insert(termB.end(), {op.mode, fts[i].fld, t_base + t_off[i], ksize(op.in[0].kA), op.in[0].kA,
                                       op.n_in > 1 ? ksize(op.in[1].kA) : 0, op.in[1].kA, ksize(op.out[0].kA), op.out[0].kA,
                                       op.n_out > 1 ? ksize(op.out[1].kA) : 0, op.out[1].kA,
                                       partner >= 0 ? t_base + t_off[partner] : -1});
        }
        fv.n_terms = (int)termB.size() / 12; // term groups
    }
    pair_rows(rowsA, rowA);
    pair_rows(rowsC, rowC);
    for (int k = 0; k < N; ++k)
    {
        yrow[2 * k] = (int32_t)(contrib.size() / 4);
        yrow[2 * k + 1] = (int32_t)per_row[k].size();
        for (const auto &c : per_row[k]) // 16 bytes per contribution: coefficient, 2 + 2 16-bit addresses, flag (0: only the first)
            contrib.insert(contrib.end(), {c[0], (int32_t)(((uint32_t)c[1] & 0xffffu) | ((uint32_t)c[2] << 16)),
                                           (int32_t)(((uint32_t)c[3] & 0xffffu) | ((uint32_t)c[4] << 16)), c[5]});
    }
    fv.n_pre = (int)gtab.size() / 4;
    fv.n_rowA = (int)rowA.size() / 8;
    fv.n_rowC = (int)rowC.size() / 8;
    fv.n_contrib = (int)contrib.size() / 4;
    std::vector<int32_t> &blob = plan->fblob;
    blob.clear();
    auto put = [&](const std::vector<int32_t> &v) {
        while (blob.size() & 3)
            blob.push_back(0);
        const int off = (int)blob.size();
        blob.insert(blob.end(), v.begin(), v.end());
        return off;
    };
    fv.off_pre = put(gtab);
    fv.off_rowA = put(rowA);
    fv.off_termB = put(termB);
    fv.off_rowC = put(rowC);
    fv.off_yrow = put(yrow);
    fv.off_contrib = put(contrib);
    {
        // fast output stage: when no row has more than two contributions, every row gets a fixed record of two
        // (coefficient, four 16-bit addresses); unused addresses point at the slot's zero cell, so the kernel needs neither
        // the loop over the contributions nor the incidence branch
        bool two = true;
        for (int k = 0; k < N && two; ++k)
            two = per_row[k].size() >= 1 && per_row[k].size() <= 2;
        fv.off_yfast = -1;
        fv.yfast_pairs = 0;
        bool pairs = two;
        for (int k = 0; k < N && pairs; ++k)
            for (const auto &c : per_row[k])
                pairs = pairs && c[2] == zero_cell && c[4] == zero_cell;
        if (pairs)
        {
            // ... and when every contribution is (one positive) - (one negative) entry (mass terms, E21^T, E10 / E21 on the right
            // of a mass): one 16-byte record per row, four value reads
            std::vector<int32_t> yfast;
            auto pack = [](int32_t lo, int32_t hi) { return (int32_t)(((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16)); };
            for (int k = 0; k < N; ++k)
            {
                const auto &c0 = per_row[k][0];
                const bool has1 = per_row[k].size() > 1;
                const auto &c1 = per_row[k][has1 ? 1 : 0];
                const int32_t z = (int32_t)zero_cell;
                yfast.insert(yfast.end(), {c0[0], pack(c0[1], c0[3]), c1[0], has1 ? pack(c1[1], c1[3]) : pack(z, z)});
            }
            fv.off_yfast = put(yfast);
            fv.yfast_pairs = 1;
        }
        else if (two)
        {
            std::vector<int32_t> yfast;
            auto pack = [](int32_t lo, int32_t hi) { return (int32_t)(((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16)); };
            for (int k = 0; k < N; ++k)
            {
                const auto &c0 = per_row[k][0];
                const bool has1 = per_row[k].size() > 1;
                const auto &c1 = per_row[k][has1 ? 1 : 0];
                const int32_t z = (int32_t)zero_cell;
                yfast.insert(yfast.end(), {c0[0], pack(c0[1], c0[2]), pack(c0[3], c0[4]), c1[0], has1 ? pack(c1[1], c1[2]) : pack(z, z),
                                           has1 ? pack(c1[3], c1[4]) : pack(z, z), 0, 0});
            }
            fv.off_yfast = put(yfast);
        }
    }
    while (blob.size() & 3)
        blob.push_back(0);
    fv.blob_ints = (int)blob.size();
    MFV2D_TRY(mfv2d_cuda_malloc(plan->ctx, blob.size() * sizeof(int32_t), (void **)&plan->d_fblob));
    MFV2D_TRY(mfv2d_cuda_upload(plan->ctx, plan->d_fblob, blob.data(), blob.size() * sizeof(int32_t)));
    plan->fv = fv;
    plan->vector_fused = true;
    plan->warp_ok = true;
    for (const FT &ft : fts)
        plan->warp_ok = plan->warp_ok && ft.op.mode <= 2;
    if (getenv("MFV2D_PLAN_DEBUG"))
        fprintf(stderr, "[mfv2d] fused vector plan: p %d q %d, %d term groups, %d pre, %d + %d row pairs, slot %d doubles, blob %d ints\n",
                p, q, fv.n_terms, fv.n_pre, fv.n_rowA, fv.n_rowC, fv.slot_doubles, fv.blob_ints);
    return MFV2D_SUCCESS;
}

struct FusedArgs
{
    ApplyArgs a;
    FusedVecInfo fv;
    const int32_t *blob;
    int n_coef, inc_rows;
    int epb;  // elements per pass
    int bulk; // x of a pass arrives by cp.async.bulk (one 1-D bulk copy per element, mbarrier completion) instead of 8-byte cp.async
};

#ifndef MFV2D_EMU
// ---- 1-D bulk copies global -> shared with mbarrier completion (TMA engine, SASS UBLKCP) ----------------------------
__device__ __forceinline__ void mbar_init(unsigned long long *bar, int count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *smem_dst, const void *gmem_src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     (unsigned)__cvta_generic_to_shared(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n"
                 ".reg .pred P1;\n"
                 "LAB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
                 "@P1 bra DONE;\n"
                 "bra LAB_WAIT;\n"
                 "DONE:\n"
                 "}\n" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
                 "r"(parity)
                 : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
#endif


__device__ __forceinline__ double signed_gather(const int4 m, const double *__restrict__ src)
{
    const int ev[4] = {m.x, m.y, m.z, m.w};
    double acc = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k)
    {
        if (ev[k] == 0)
            break;
        const double v = src[abs(ev[k]) - 1];
        acc += ev[k] < 0 ? -v : v;
    }
    return acc;
}

// acc[n][:] += c[n] * row[:] for NT accumulator rows: one 16-byte table read feeds 2 NT DFMAs
template <int Q, int NT> __device__ __forceinline__ void row_axpy_n(double (&acc)[NT][Q], const double (&c)[NT], const double *__restrict__ row)
{
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(row + b);
#pragma unroll
        for (int n = 0; n < NT; ++n)
        {
            acc[n][b] = fma(c[n], v.x, acc[n][b]);
            acc[n][b + 1] = fma(c[n], v.y, acc[n][b + 1]);
        }
    }
    if (Q & 1)
    {
        const double v = row[Q - 1];
#pragma unroll
        for (int n = 0; n < NT; ++n)
            acc[n][Q - 1] = fma(c[n], v, acc[n][Q - 1]);
    }
}

// acc[n][:] = c[n] * row[:]
template <int Q, int NT> __device__ __forceinline__ void row_mul_n(double (&acc)[NT][Q], const double (&c)[NT], const double *__restrict__ row)
{
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(row + b);
#pragma unroll
        for (int n = 0; n < NT; ++n)
        {
            acc[n][b] = c[n] * v.x;
            acc[n][b + 1] = c[n] * v.y;
        }
    }
    if (Q & 1)
    {
        const double v = row[Q - 1];
#pragma unroll
        for (int n = 0; n < NT; ++n)
            acc[n][Q - 1] = c[n] * v;
    }
}

// res[n] = row[:] . g[n][:]
template <int Q, int NT> __device__ __forceinline__ void row_dot_n(const double *__restrict__ row, const double (&g)[NT][Q], double (&res)[NT])
{
    double a0[NT], a1[NT];
#pragma unroll
    for (int n = 0; n < NT; ++n)
        a0[n] = a1[n] = 0.0;
#pragma unroll
    for (int b = 0; b + 1 < Q; b += 2)
    {
        const double2 v = *reinterpret_cast<const double2 *>(row + b);
#pragma unroll
        for (int n = 0; n < NT; ++n)
        {
            a0[n] = fma(v.x, g[n][b], a0[n]);
            a1[n] = fma(v.y, g[n][b + 1], a1[n]);
        }
    }
    if (Q & 1)
    {
        const double v = row[Q - 1];
#pragma unroll
        for (int n = 0; n < NT; ++n)
            a0[n] = fma(v, g[n][Q - 1], a0[n]);
    }
#pragma unroll
    for (int n = 0; n < NT; ++n)
        res[n] = a0[n] + a1[n];
}

// coefficient j of a phase-A row: read in place, or gathered from x through the right incidence
__device__ __forceinline__ double fused_coef(const double *__restrict__ slot, int src, int g, int j, const int4 *__restrict__ gtab)
{
    if (g < 0)
        return slot[src + j];
    const int4 m = gtab[g + j]; // (p0 + p1) - (n0 + n1); unused entries point at the slot's zero cell
    return (slot[m.x] + slot[m.y]) - (slot[m.z] + slot[m.w]);
}

// trip-count-2 unrolling of the row loops pays from 10 quadrature points on (measured: +6 % at Q = 10, -3 % at Q = 8)
template <int Q> struct FusedUnroll
{
    static constexpr int value = Q >= 10 ? 2 : 1;
};

template <int Q, int NT>
__device__ __forceinline__ void fused_rows_a(const int4 d0, const int4 d1, double *__restrict__ slot, const double *__restrict__ smem, int edge_off,
                                             const int4 *__restrict__ gtab)
{
    constexpr int UNR = FusedUnroll<Q>::value;
    constexpr int LDQ = (Q + 1) & ~1;
    const double *B = smem + d0.z * edge_off;
    double acc[NT][Q];
    {
        // first coefficient: plain products instead of zero + fma
        double c[NT];
        c[0] = fused_coef(slot, d0.x, d1.z, 0, gtab);
        if (NT > 1)
            c[NT - 1] = fused_coef(slot, d1.x, d1.w, 0, gtab);
        row_mul_n<Q, NT>(acc, c, B);
    }
#pragma unroll(UNR)
    for (int j = 1; j < d0.y; ++j)
    {
        double c[NT];
        c[0] = fused_coef(slot, d0.x, d1.z, j, gtab);
        if (NT > 1)
            c[NT - 1] = fused_coef(slot, d1.x, d1.w, j, gtab);
        row_axpy_n<Q, NT>(acc, c, B + j * LDQ);
    }
#pragma unroll
    for (int n = 0; n < NT; ++n)
    {
        double *dst = slot + (n == 0 ? d0.w : d1.y);
#pragma unroll
        for (int b = 0; b + 1 < Q; b += 2)
            *reinterpret_cast<double2 *>(dst + b) = make_double2(acc[n][b], acc[n][b + 1]);
        if (Q & 1)
            dst[Q - 1] = acc[n][Q - 1];
    }
}

// phase C: the nB results of a row overwrite the head of the row itself (it has been read into registers)
template <int Q, int NT>
__device__ __forceinline__ void fused_rows_c(const int4 d0, const int4 d1, double *__restrict__ slot, const double *__restrict__ smem, int edge_off)
{
    constexpr int UNR = FusedUnroll<Q>::value;
    constexpr int LDQ = (Q + 1) & ~1;
    double tr[NT][Q];
#pragma unroll
    for (int n = 0; n < NT; ++n)
    {
        const double *src = slot + (n == 0 ? d0.x : d1.x);
#pragma unroll
        for (int b = 0; b + 1 < Q; b += 2)
        {
            const double2 v = *reinterpret_cast<const double2 *>(src + b);
            tr[n][b] = v.x;
            tr[n][b + 1] = v.y;
        }
        if (Q & 1)
            tr[n][Q - 1] = src[Q - 1];
    }
    const double *B = smem + d0.z * edge_off;
    double *dst0 = slot + d0.x, *dst1 = slot + (NT > 1 ? d1.x : d0.x);
#pragma unroll(UNR)
    for (int j = 0; j < d0.y; ++j)
    {
        double res[NT];
        row_dot_n<Q, NT>(B + j * LDQ, tr, res);
        dst0[j] = res[0];
        if (NT > 1)
            dst1[j] = res[NT - 1];
    }
}

// geometry of one quadrature column (fixed xi): what the metric at its Q points is computed from
struct ColumnGeo
{
    double j10, j11, dx10, dx23, dy10, dy23, wb;
};

// one term (NT = 1) or two terms of the same kind (NT = 2) on the quadrature column b of one element
template <int Q, int NT>
__device__ __forceinline__ void fused_terms_b(const int *__restrict__ tb, double *__restrict__ slot, int b, const double *__restrict__ smem,
                                              int edge_off, const double *__restrict__ pnts, const double *__restrict__ wgts,
                                              const ColumnGeo &g, const double *__restrict__ fp)
{
    constexpr int UNR = FusedUnroll<Q>::value;
    constexpr int LDQ = (Q + 1) & ~1;
    const int LDT = asm_ldq(Q);
    const int mode = tb[0];
    double *tc[NT];
    tc[0] = slot + tb[2] + b;
    if (NT > 1)
        tc[NT - 1] = slot + tb[11] + b;
    const int nA0 = tb[3], nA1 = tb[5], nO0 = tb[7], nO1 = tb[9];
    double u0[NT][Q], u1[NT][Q];
#pragma unroll
    for (int n = 0; n < NT; ++n)
#pragma unroll
        for (int a = 0; a < Q; ++a)
            u0[n][a] = u1[n][a] = 0.0;
    {
        const double *A = smem + tb[4] * edge_off;
#pragma unroll(UNR)
        for (int i = 0; i < nA0; ++i)
        {
            double c[NT];
#pragma unroll
            for (int n = 0; n < NT; ++n)
                c[n] = tc[n][i * LDT];
            row_axpy_n<Q, NT>(u0, c, A + i * LDQ);
        }
        const double *A1 = smem + tb[6] * edge_off;
#pragma unroll(UNR)
        for (int i = 0; i < nA1; ++i)
        {
            double c[NT];
#pragma unroll
            for (int n = 0; n < NT; ++n)
                c[n] = tc[n][(nA0 + i) * LDT];
            row_axpy_n<Q, NT>(u1, c, A1 + i * LDQ);
        }
    }
    {
#pragma unroll
        for (int a = 0; a < Q; ++a)
        {
            const double eta = pnts[a];
            const double j00 = 0.25 * (g.dx10 * (1 - eta) + g.dx23 * (1 + eta));
            const double j01 = 0.25 * (g.dy10 * (1 - eta) + g.dy23 * (1 + eta));
            const double det = j00 * g.j11 - g.j10 * j01;
            if (mode == 0)
            {
                const double k = det * (wgts[a] * g.wb);
#pragma unroll
                for (int n = 0; n < NT; ++n)
                    u0[n][a] = u0[n][a] * k;
                continue;
            }
            const double wod = wgts[a] * g.wb * fast_rcp(det); // one reciprocal per point and term group
            if (mode == 1)
            {
#pragma unroll
                for (int n = 0; n < NT; ++n)
                    u0[n][a] = u0[n][a] * wod;
            }
            else if (mode == 2)
            {
                const double khh = (g.j11 * g.j11 + g.j10 * g.j10) * wod, kvv = (j00 * j00 + j01 * j01) * wod,
                             khv = (j01 * g.j11 + j00 * g.j10) * wod;
#pragma unroll
                for (int n = 0; n < NT; ++n)
                {
                    const double a0 = u0[n][a], a1 = u1[n][a];
                    u0[n][a] = khh * a0 + khv * a1;
                    u1[n][a] = khv * a0 + kvv * a1;
                }
            }
            else
            {
                const double fx = fp[2 * a * Q], fy = fp[2 * a * Q + 1];
                double kh, kv;
                if (mode <= 4)
                {
                    const double w = wgts[a] * g.wb;
                    kh = (fx * g.j11 - fy * g.j10) * w;
                    kv = -(fy * j00 - fx * j01) * w;
                }
                else
                {
                    kh = -(fx * g.j10 + fy * g.j11) * wod;
                    kv = -(fx * j00 + fy * j01) * wod;
                }
#pragma unroll
                for (int n = 0; n < NT; ++n)
                {
                    const double a0 = u0[n][a], a1 = u1[n][a];
                    if (mode == 3 || mode == 6)
                        u0[n][a] = kh * a0 + kv * a1;
                    else
                    {
                        u0[n][a] = kh * a0;
                        u1[n][a] = kv * a0;
                    }
                }
            }
        }
    }
    {
        const double *A = smem + tb[8] * edge_off;
#pragma unroll(UNR)
        for (int i = 0; i < nO0; ++i)
        {
            double res[NT];
            row_dot_n<Q, NT>(A + i * LDQ, u0, res);
#pragma unroll
            for (int n = 0; n < NT; ++n)
                tc[n][i * LDT] = res[n];
        }
        const double *A1 = smem + tb[10] * edge_off;
#pragma unroll(UNR)
        for (int i = 0; i < nO1; ++i)
        {
            double res[NT];
            row_dot_n<Q, NT>(A1 + i * LDQ, u1, res);
#pragma unroll
            for (int n = 0; n < NT; ++n)
                tc[n][(nO0 + i) * LDT] = res[n];
        }
    }
}

template <int Q> __global__ void __launch_bounds__(APPLYT_THREADS, 2) element_vector_fused_kernel(const FusedArgs fa)
{
    MFV2D_DYN_SMEM(double, smem);
    constexpr int LDQ = (Q + 1) & ~1;
    const ApplyArgs &args = fa.a;
    const FusedVecInfo &fv = fa.fv;
    const int p = args.p, np1 = p + 1, N = args.n_total;
    // CTA-shared: padded tables | points, weights | coefficients | descriptor blob | incidence maps | element slots
    const int edge_off = np1 * LDQ;
    double *pnts = smem + (2 * p + 1) * LDQ, *wgts = pnts + LDQ;
    double *coef = wgts + LDQ;
    int *blob = reinterpret_cast<int *>(coef + ((fa.n_coef + 1) & ~1));
    int4 *inc = reinterpret_cast<int4 *>(blob + fv.blob_ints);
    double *slots = reinterpret_cast<double *>(inc + fa.inc_rows);
    const int epb = fa.epb, SD = fv.slot_doubles;
    // slot of element e: stride SD (congruent to Q modulo 16 doubles: phase B, lanes = (e, b), is conflict free) plus a
    // small rotation so that 16-byte accesses of phases A / C, lanes = consecutive elements, fall into distinct bank groups
    // The rotation depends on s = (SD / 2) mod 8, the slot stride in 16-byte units: the bases of 8 consecutive elements must fall
    // into 8 distinct 16-byte bank groups.  s odd (Q = 5, 6, 9, 10, 13, 14): already a permutation, no rotation (a rotation
    // costs a factor 2 there: measured x3.9 / x2.0 / x1.6 wavefronts on the 8-byte, 16-byte and phase-B accesses at Q = 10);
    // s = 2, 6: one step every 4 elements; s = 0, 4: one step every 2 elements.
    const int rot_s = (SD >> 1) & 7;
    const int rot_shift = (rot_s & 1) ? 0 : ((rot_s & 2) ? 2 : 1), rot_mask = (rot_s & 1) ? 0 : ((rot_s & 2) ? 1 : 3);
#define SLOT(e) ((e) * SD + 2 * (((e) >> rot_shift) & rot_mask))
    const int off_xs = fv.off_vec;
    const unsigned long long n_groups = (args.n_elem + epb - 1) / epb;
    const float inv_n = 1.0f / (float)N;

    // x of a pass and its corners are copied global -> shared asynchronously (no registers in flight): issued
    // right after phase A of the previous pass (x is dead from there on; the corners are double buffered)
#ifndef MFV2D_EMU
    __shared__ unsigned long long x_bar; // completion of the bulk copies of one pass (phase = pass parity)
    if (fa.bulk && threadIdx.x == 0)
    {
        mbar_init(&x_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
#endif
    auto prefetch = [&](unsigned long long grp, int parity) {
        if (grp >= n_groups)
            return;
        const unsigned long long elem0 = grp * epb;
        const int nv = (int)((args.n_elem - elem0) < (unsigned long long)epb ? (args.n_elem - elem0) : epb);
        for (int i = threadIdx.x; i < nv * 8; i += blockDim.x)
            cp_async8(slots + SLOT(i >> 3) + parity * 8 + (i & 7), args.corners + elem0 * 8 + i);
#ifndef MFV2D_EMU
        if (fa.bulk)
        {
            // one thread arms the barrier with the byte count of the pass and issues one 1-D bulk copy per element (N * 8 bytes,
            // 16-byte aligned on both sides): 25 instructions instead of 3 000 8-byte cp.async, none of them on the LSU
            if (threadIdx.x == 0)
            {
                fence_proxy_async(); // the slots were last touched through the generic proxy (phase A of this pass)
                mbar_expect_tx(&x_bar, (unsigned)(nv * N * 8));
                for (int e = 0; e < nv; ++e)
                    bulk_g2s(slots + SLOT(e) + off_xs, args.x + (elem0 + e) * N, (unsigned)(N * 8), &x_bar);
            }
            return;
        }
#endif
        for (int i = threadIdx.x; i < nv * N; i += blockDim.x)
        {
            const int e = (int)(((float)i + 0.5f) * inv_n);
            cp_async8(slots + SLOT(e) + off_xs + (i - e * N), args.x + elem0 * N + i);
        }
    };
    prefetch(blockIdx.x, 0);

    for (int i = threadIdx.x; i < (2 * p + 1) * LDQ; i += blockDim.x)
    {
        const int r = i / LDQ, b = i - r * LDQ;
        smem[i] = b < Q ? args.tab[r * Q + b] : 0.0; // tab = node (p+1) x q | edge p x q | pnts | wgts
    }
    for (int i = threadIdx.x; i < Q; i += blockDim.x)
    {
        pnts[i] = args.tab[(2 * p + 1) * Q + i];
        wgts[i] = args.tab[(2 * p + 2) * Q + i];
    }
    for (int i = threadIdx.x; i < epb; i += blockDim.x)
        slots[SLOT(i) + off_xs + N] = 0.0; // the zero cell of every slot (never written again)
    for (int i = threadIdx.x; i < fa.n_coef; i += blockDim.x)
        coef[i] = args.coef[i];
    for (int i = threadIdx.x; i < fv.blob_ints; i += blockDim.x)
        blob[i] = fa.blob[i];
    for (int i = threadIdx.x; i < fa.inc_rows; i += blockDim.x)
        inc[i] = reinterpret_cast<const int4 *>(args.inc)[i];
    const int4 *gtab = reinterpret_cast<const int4 *>(blob + fv.off_pre);
    const int4 *rowA = reinterpret_cast<const int4 *>(blob + fv.off_rowA);
    const int *termB = blob + fv.off_termB;
    const int4 *rowC = reinterpret_cast<const int4 *>(blob + fv.off_rowC);
    const int2 *yrow = reinterpret_cast<const int2 *>(blob + fv.off_yrow);
    const int4 *contrib = reinterpret_cast<const int4 *>(blob + fv.off_contrib);

    int parity = 0;
    for (unsigned long long grp = blockIdx.x; grp < n_groups; grp += gridDim.x, parity ^= 1)
    {
        const unsigned long long elem0 = grp * epb;
        const int nv = (int)((args.n_elem - elem0) < (unsigned long long)epb ? (args.n_elem - elem0) : epb);
        cp_async_wait_all();
        __syncthreads(); // corners (and x, 8-byte path) of this pass have landed; the previous pass is done with the t rows
#ifndef MFV2D_EMU
        if (fa.bulk)
            mbar_wait(&x_bar, (unsigned)parity); // x of this pass: the bulk copies have completed their bytes
#endif
        // ---- phase A: t[row][:] = sum_j c[j] B_j[:], rows of the same table two at a time ----
        {
            // lanes = consecutive elements of ONE row pair: descriptors, tables and branches are warp uniform
            const int R = fv.n_rowA;
            const float inv_nv = 1.0f / (float)nv;
            for (int idx = threadIdx.x; idx < nv * R; idx += blockDim.x)
            {
                const int row = (int)(((float)idx + 0.5f) * inv_nv), e = idx - row * nv;
                double *slot = slots + SLOT(e);
                const int4 d0 = rowA[2 * row], d1 = rowA[2 * row + 1];
                if (d1.x >= 0)
                    fused_rows_a<Q, 2>(d0, d1, slot, smem, edge_off, gtab);
                else
                    fused_rows_a<Q, 1>(d0, d1, slot, smem, edge_off, gtab);
            }
        }
        __syncthreads();
        prefetch(grp + gridDim.x, parity ^ 1);
        // ---- phase B: thread (e, b) owns the quadrature column b of element e for all terms ----
        for (int idx = threadIdx.x; idx < nv * Q; idx += blockDim.x)
        {
            const int e = idx / Q, b = idx - e * Q;
            double *slot = slots + SLOT(e);
            const double *cs = slot + parity * 8;
            const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
            const double xi = pnts[b];
            ColumnGeo g;
            g.wb = wgts[b];
            g.j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
            g.j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
            g.dx10 = x1 - x0;
            g.dx23 = x2 - x3;
            g.dy10 = y1 - y0;
            g.dy23 = y2 - y3;
            for (int t = 0; t < fv.n_terms; ++t)
            {
                const int *tb = termB + 12 * t;
                const double *fp = tb[0] >= 3 ? args.fields[tb[1]] + (elem0 + e) * (unsigned long long)(2 * Q * Q) + 2 * b : nullptr;
                if (tb[11] >= 0)
                    fused_terms_b<Q, 2>(tb, slot, b, smem, edge_off, pnts, wgts, g, fp);
                else
                    fused_terms_b<Q, 1>(tb, slot, b, smem, edge_off, pnts, wgts, g, fp);
            }
        }
        __syncthreads();
        // ---- phase C: out[i, j] = B_j[:] . t'[row][:], rows of the same table two at a time ----
        {
            const int R = fv.n_rowC;
            const float inv_nv = 1.0f / (float)nv;
            for (int idx = threadIdx.x; idx < nv * R; idx += blockDim.x)
            {
                const int row = (int)(((float)idx + 0.5f) * inv_nv), e = idx - row * nv;
                double *slot = slots + SLOT(e);
                const int4 d0 = rowC[2 * row], d1 = rowC[2 * row + 1];
                if (d1.x >= 0)
                    fused_rows_c<Q, 2>(d0, d1, slot, smem, edge_off);
                else
                    fused_rows_c<Q, 1>(d0, d1, slot, smem, edge_off);
            }
        }
        __syncthreads();
        // ---- post: y[k] = sum over the terms of coef * (signed sum of up to four outputs), stored to HBM ----
        if (fv.off_yfast >= 0)
        {
            const int4 *yf = reinterpret_cast<const int4 *>(blob + fv.off_yfast);
            if (fv.yfast_pairs)
            {
                for (int i = threadIdx.x; i < nv * N; i += blockDim.x)
                {
                    const int e = (int)(((float)i + 0.5f) * inv_n), k = i - e * N;
                    const double *slot = slots + SLOT(e);
                    const int4 d = yf[k]; // coef0, (p | n) of contribution 0, coef1, (p | n) of contribution 1
                    const double v0 = slot[d.y & 0xffff] - slot[(unsigned)d.y >> 16];
                    const double v1 = slot[d.w & 0xffff] - slot[(unsigned)d.w >> 16];
                    args.y[elem0 * N + i] = fma(coef[d.z], v1, coef[d.x] * v0);
                }
                continue;
            }
            for (int i = threadIdx.x; i < nv * N; i += blockDim.x)
            {
                const int e = (int)(((float)i + 0.5f) * inv_n), k = i - e * N;
                const double *slot = slots + SLOT(e);
                const int4 d0 = yf[2 * k], d1 = yf[2 * k + 1];
                const double v0 = (slot[d0.y & 0xffff] + slot[(unsigned)d0.y >> 16]) - (slot[d0.z & 0xffff] + slot[(unsigned)d0.z >> 16]);
                const double v1 = (slot[d1.x & 0xffff] + slot[(unsigned)d1.x >> 16]) - (slot[d1.y & 0xffff] + slot[(unsigned)d1.y >> 16]);
                args.y[elem0 * N + i] = fma(coef[d0.w], v1, coef[d0.x] * v0);
            }
            continue;
        }
        for (int i = threadIdx.x; i < nv * N; i += blockDim.x)
        {
            const int e = (int)(((float)i + 0.5f) * inv_n), k = i - e * N;
            const double *slot = slots + SLOT(e);
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
#undef SLOT

template <int Q> static int launch_fused_q(mfv2d_ctx *ctx, FusedArgs &fa, size_t fixed_bytes, size_t n_elem)
{
    const size_t slot_bytes = (size_t)fa.fv.slot_doubles * 8;
    const size_t budget = ((size_t)228 * 1024 - 2048) / 2; // two CTAs per SM
    int epb = (int)((budget > fixed_bytes ? budget - fixed_bytes : 0) / slot_bytes);
    const int want = APPLYT_THREADS / Q; // phase B: one thread per (element, quadrature column), one round
    if (epb > want)
        epb = want;
    if (const char *env = getenv("MFV2D_APPLY_EPB")) // experiment: elements per CTA pass
    {
        const int v = atoi(env);
        if (v >= 1 && v < epb)
            epb = v;
    }
    if (epb < 1)
        epb = 1;
    fa.epb = epb;
    {
        // bulk copies need 16-byte alignment of every element's x (N even), of the slot position of x and of the buffer
        const char *env = getenv("MFV2D_APPLY_BULK");
        fa.bulk = !(env && env[0] == '0') && (fa.a.n_total % 2 == 0) && (fa.fv.off_vec % 2 == 0) && (fa.fv.slot_doubles % 2 == 0) &&
                  ((size_t)fa.a.x % 16 == 0) && (fixed_bytes % 16 == 0);
    }
    const size_t smem = fixed_bytes + (size_t)epb * slot_bytes;
    if (smem > ctx->smem_optin)
        return -1; // the caller falls back to the term-by-term kernel
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_vector_fused_kernel<Q>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_vector_fused_kernel<Q>, APPLYT_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    const unsigned long long need = (n_elem + epb - 1) / epb;
    if (grid > need)
        grid = need;
    MFV2D_LAUNCH(element_vector_fused_kernel<Q>, (unsigned)grid, APPLYT_THREADS, smem, ctx->stream, fa);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

#include "apply_warp.cuh"

// =====================================================================================================
// Tensor-core variant of the fused apply for q <= 8 and p <= 7 (every 1-D extent fits one 8 x 8 tile):
// ONE WARP PER ELEMENT, no CTA barrier.  All four contractions of a term are chains of fp64
// mma.sync.m8n8k4 (DMMA): the 1-D basis tables live in REGISTERS as A / B fragments for the whole kernel, the
// intermediate 8 x 8 tiles never leave the accumulator fragments (two shuffles turn a D fragment into the
// next B or A fragment), and every lane owns two fixed quadrature points of the element, so the metric and its
// reciprocal are computed once per point and element.  On B200 a DMMA issues at exactly the DFMA flop rate
// (profiles/micro/dmma_rate_b200.txt): what it buys here is issue slots (256 FMAs per instruction) and all the
// shared-memory traffic of the table reads.
//   fragments (PTX ISA, m8n8k4.f64), g = lane / 4, t = lane % 4:
//     A[m = g][k = t]      B[k = t][n = g]      C/D[m = g][n = 2 t], [m = g][n = 2 t + 1]
// =====================================================================================================
enum
{
    DM_STRIDE = 24 // ints per term descriptor
};

static int plan_lower_vector_dmma(mfv2d_plan *plan)
{
    plan->vector_dmma = false;
    const int p = plan->p, q = plan->q, np1 = p + 1, N = plan->n_total;
    if (plan->vector_needs_dense || q < 2 || q > 8 || np1 > 8)
        return MFV2D_SUCCESS;
    std::vector<FT> fts;
    if (!collect_tensor_terms(plan, fts))
        return MFV2D_SUCCESS;
    auto ksize = [&](int k) { return k == TK_NODE ? np1 : p; };
    // per-warp staging (doubles): x[2][Nx] (double buffered, each with a zero cell at N) | corners[2][8] | y[N] |
    // output vector of the current term.  Gather addresses are relative to the current x buffer (right incidence)
    // or to the strip (left incidence; the zero cell of buffer 0 serves as its zero)
    const int zero_cell = N;
    const int Nx = (N + 2) & ~1;
    const int dump_cell = 2 * Nx + 16; // receives the stores of the tile's padding lanes
    const int off_ys = 2 * Nx + 18;
    const int off_outv = off_ys + ((N + 1) & ~1);
    int max_out = 0;
    std::vector<int32_t> terms, gtab, dtab, ltab, otab;
    for (const FT &ft : fts)
    {
        const TileOp &op = ft.op;
        int ni = 0, no = 0;
        for (int k = 0; k < op.n_in; ++k)
            ni += ksize(op.in[k].kA) * ksize(op.in[k].kB);
        for (int k = 0; k < op.n_out; ++k)
            no += ksize(op.out[k].kA) * ksize(op.out[k].kB);
        max_out = no > max_out ? no : max_out;
        const int c0 = plan->form_offsets[ft.cf], r0 = plan->form_offsets[ft.rf];
        const int n_rows_form = plan->form_offsets[ft.rf + 1] - r0;
        if (ft.R >= 0 ? inc_rows_of(ft.R, p) != ni : plan->form_offsets[ft.cf + 1] - c0 != ni)
            return MFV2D_SUCCESS;
        if (ft.L >= 0 ? inc_rows_of(ft.L, p) != n_rows_form : no != n_rows_form)
            return MFV2D_SUCCESS;
        // signed sum of up to four cells as (p0 + p1) - (n0 + n1); the zero cell where there is no entry
        auto split = [&](const int32_t *ev, int base, std::vector<int32_t> &out) -> bool {
            int32_t e4[4] = {zero_cell, zero_cell, zero_cell, zero_cell};
            int np_ = 0, nn = 0;
            for (int z = 0; z < 4 && ev[z] != 0; ++z)
            {
                const int a = base + (ev[z] < 0 ? -ev[z] : ev[z]) - 1;
                if (ev[z] > 0 ? np_ == 2 : nn == 2)
                    return false;
                if (ev[z] > 0)
                    e4[np_++] = a;
                else
                    e4[2 + nn++] = a;
            }
            out.insert(out.end(), e4, e4 + 4);
            return true;
        };
        int lbase = -1;
        if (ft.L >= 0)
        {
            lbase = (int)ltab.size() / 4;
            for (int k = 0; k < n_rows_form; ++k)
                if (!split(plan->basis->h_inc.data() + 4 * (size_t)(plan->basis->inc_off[ft.L] + k), off_outv, ltab))
                    return MFV2D_SUCCESS;
        }
        // per-lane addresses (g = lane / 4, t = lane % 4): the coefficient c[i = g][j = 2 t + ks] of every input
        // component as (p0 + p1) - (n0 + n1) in the x buffer (zero cell where there is nothing), and the two cells
        // out[i = g][j = 2 t + s] of every output component in the strip (dump cell for the padding of the tile)
        // gathered terms: int4 records, [component][k-step][lane]; direct terms: int2 {address of step 0, of step 1}, [component][lane]
        const int lbase_in = ft.R >= 0 ? (int)gtab.size() / 4 : (int)dtab.size() / 2;
        for (int k = 0; k < 2; ++k)
        {
            const int nA = k < op.n_in ? ksize(op.in[k].kA) : 0, nB = k < op.n_in ? ksize(op.in[k].kB) : 0;
            if (ft.R >= 0)
            {
                for (int ks = 0; ks < 2; ++ks)
                    for (int lane = 0; lane < 32; ++lane)
                    {
                        const int g = lane >> 2, j = 2 * (lane & 3) + ks;
                        if (g < nA && j < nB)
                        {
                            if (!split(plan->basis->h_inc.data() + 4 * (size_t)(plan->basis->inc_off[ft.R] + op.in[k].off + g * nB + j), c0, gtab))
                                return MFV2D_SUCCESS;
                        }
                        else
                            gtab.insert(gtab.end(), {zero_cell, zero_cell, zero_cell, zero_cell});
                    }
            }
            else
                for (int lane = 0; lane < 32; ++lane)
                    for (int ks = 0; ks < 2; ++ks)
                    {
                        const int g = lane >> 2, j = 2 * (lane & 3) + ks;
                        dtab.push_back(g < nA && j < nB ? c0 + op.in[k].off + g * nB + j : zero_cell);
                    }
        }
        const int lbase_out = (int)otab.size() / 2;
        for (int k = 0; k < 2; ++k)
            for (int lane = 0; lane < 32; ++lane)
            {
                const int g = lane >> 2, t = lane & 3;
                const int nA = k < op.n_out ? ksize(op.out[k].kA) : 0, nB = k < op.n_out ? ksize(op.out[k].kB) : 0;
                for (int sfx = 0; sfx < 2; ++sfx)
                {
                    const int j = 2 * t + sfx;
                    if (g < nA && j < nB)
                        otab.push_back((ft.L >= 0 ? off_outv : off_ys + r0) + op.out[k].off + g * nB + j);
                    else
                        otab.push_back(dump_cell);
                }
            }
        const int32_t d[DM_STRIDE] = {op.mode, ft.fld, ft.ci, lbase_in, lbase_out, ft.R >= 0 ? 1 : 0, lbase, n_rows_form, off_ys + r0, 0, 0, 0,
                                      0,       0,      0,     0,        0,         0,                0,     0,           0,           0, 0, 0};
        terms.insert(terms.end(), d, d + DM_STRIDE);
    }
    std::vector<int32_t> blob;
    auto put = [&](const std::vector<int32_t> &v) {
        while (blob.size() & 3)
            blob.push_back(0);
        const int off = (int)blob.size();
        blob.insert(blob.end(), v.begin(), v.end());
        return off;
    };
    for (std::vector<int32_t> *v : {&gtab, &dtab, &ltab, &otab}) // the kernel addresses shared memory by byte offsets
        for (int32_t &a : *v)
            a *= 8;
    plan->dm_terms = (int)fts.size();
    plan->dm_off_terms = put(terms);
    plan->dm_off_gtab = put(gtab);
    plan->dm_off_ltab = put(ltab);
    plan->dm_off_otab = put(otab);
    plan->dm_off_dtab = put(dtab);
    while (blob.size() & 3)
        blob.push_back(0);
    plan->dm_blob_ints = (int)blob.size();
    plan->dm_off_ys = off_ys;
    plan->dm_off_outv = off_outv;
    plan->dm_warp_doubles = (off_outv + max_out + 1) & ~1;
    MFV2D_TRY(mfv2d_cuda_malloc(plan->ctx, blob.size() * sizeof(int32_t), (void **)&plan->d_dmblob));
    MFV2D_TRY(mfv2d_cuda_upload(plan->ctx, plan->d_dmblob, blob.data(), blob.size() * sizeof(int32_t)));
    plan->vector_dmma = true;
    return MFV2D_SUCCESS;
}

#ifndef MFV2D_EMU
struct DmmaArgs
{
    ApplyArgs a;
    const int32_t *blob;
    int n_terms, off_terms, off_gtab, off_ltab, off_otab, off_dtab, blob_ints;
    int warp_doubles, off_ys, off_outv, n_coef;
};

#define DMMA_THREADS 256

// shared-memory accesses through 32-bit window addresses (the address tables hold byte offsets)
__device__ __forceinline__ double lds64(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts64(unsigned a, double v) { asm volatile("st.shared.f64 [%0], %1;\n" ::"r"(a), "d"(v) : "memory"); }

// D (8 x 8, fragment d0 / d1) += A (8 x 4, fragment a) * B (4 x 8, fragment b)
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// Everything is kept TRANSPOSED with the xi index b as the row (m = g) of every tile, and the contraction index of
// k-step ks is k = 2 t + ks: a D fragment X[g][2 t + s] then IS the A fragment of the next product (step ks = s), no
// data movement; only the last product (which contracts b) needs the tile transposed through shuffles.
struct DmmaLane
{
    int g, t, src0; // src0 = 8 t + g / 2: source lane of the transposing shuffle (k-step 0)
    double fa[2][2]; // fa[kind][ks] = tab_kind[basis 2 t + ks][point g]
    double fb[2][2]; // fb[kind][ks] = tab_kind[basis g][point 2 t + ks]
};

// values at the lane's two points (a = 2 t + s, b = g) of one input component (kA along eta, kB along xi);
// `adr` = the lane's two address records {p0, p1, n0, n1} (k-steps 0 and 1) of the coefficient c[i = g][j = 2 t + ks]
template <int KA, int KB>
__device__ __forceinline__ void dmma_in(const DmmaLane &L, unsigned xs, const int4 *__restrict__ gadr, const int2 *__restrict__ dadr,
                                        bool gathered, double &u0, double &u1)
{
    // T^T[b][i] = sum_j B_j(b) c[i][j]:  A = B-table (m = b, k = j), B = c (k = j, n = i = g)
    double c0, c1;
    if (gathered)
    {
        const int4 q0 = gadr[0], q1 = gadr[32];
        c0 = (lds64(xs + q0.x) + lds64(xs + q0.y)) - (lds64(xs + q0.z) + lds64(xs + q0.w));
        c1 = (lds64(xs + q1.x) + lds64(xs + q1.y)) - (lds64(xs + q1.z) + lds64(xs + q1.w));
    }
    else
    {
        const int2 q = dadr[0];
        c0 = lds64(xs + q.x);
        c1 = lds64(xs + q.y);
    }
    double t0 = 0.0, t1 = 0.0;
    dmma884(t0, t1, L.fa[KB][0], c0);
    dmma884(t0, t1, L.fa[KB][1], c1);
    // U^T[b][a] += sum_i T^T[b][i] A_i(a):  A = T^T straight from the D fragment, B = A-table (k = i, n = a = g)
    dmma884(u0, u1, t0, L.fa[KA][0]);
    dmma884(u0, u1, t1, L.fa[KA][1]);
}

// out[i][j] of one output component from the weighted values at the points; o0 / o1 = out[g][2 t], out[g][2 t + 1]
template <int KA, int KB> __device__ __forceinline__ void dmma_out(const DmmaLane &L, double u0, double u1, double &o0, double &o1)
{
    // T'^T[b][i] = sum_a U'^T[b][a] A_i(a):  A = U'^T from the D fragment, B = A-table (k = a, n = i = g)
    double t0 = 0.0, t1 = 0.0;
    dmma884(t0, t1, u0, L.fb[KA][0]);
    dmma884(t0, t1, u1, L.fb[KA][1]);
    // out[i][j] = sum_b T'^T[b][i] B_j(b):  A[m = i = g][k = b = 2 t + ks] = T'^T[2 t + ks][g] lives in lane
    // 4 (2 t + ks) + g / 2, register g % 2;  B = B-table (k = b, n = j = g)
    o0 = o1 = 0.0;
#pragma unroll
    for (int ks = 0; ks < 2; ++ks)
    {
        const int src = L.src0 + 4 * ks;
        const double v0 = __shfl_sync(0xffffffffu, t0, src), v1 = __shfl_sync(0xffffffffu, t1, src);
        dmma884(o0, o1, (L.g & 1) ? v1 : v0, L.fb[KB][ks]);
    }
}

__global__ void __launch_bounds__(DMMA_THREADS, 3) element_vector_dmma_kernel(const DmmaArgs da)
{
    MFV2D_DYN_SMEM(double, smem);
    const ApplyArgs &args = da.a;
    const int p = args.p, Q = args.q, np1 = p + 1, N = args.n_total, nh = p * np1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    DmmaLane L;
    L.g = lane >> 2;
    L.t = lane & 3;
    L.src0 = 8 * L.t + (L.g >> 1);
    const int g = L.g, t = L.t;
    // CTA-shared: coefficients | descriptor blob ; then one staging strip per warp: x[2][Nx] | corners[2][8] | y | term output
    double *coef = smem;
    int *blob = reinterpret_cast<int *>(coef + ((da.n_coef + 1) & ~1));
    double *wbuf = reinterpret_cast<double *>(blob + da.blob_ints) + (size_t)warp * da.warp_doubles;
    for (int i = threadIdx.x; i < da.n_coef; i += blockDim.x)
        coef[i] = args.coef[i];
    for (int i = threadIdx.x; i < da.blob_ints; i += blockDim.x)
        blob[i] = da.blob[i];
    const int *terms = blob + da.off_terms;
    const int4 *gtab = reinterpret_cast<const int4 *>(blob + da.off_gtab);
    const int4 *ltab = reinterpret_cast<const int4 *>(blob + da.off_ltab);
    const int2 *otab = reinterpret_cast<const int2 *>(blob + da.off_otab);
    const int2 *dtab = reinterpret_cast<const int2 *>(blob + da.off_dtab);
#pragma unroll
    for (int k = 0; k < 2; ++k)
    {
        const double *tab = args.tab + (k ? np1 * Q : 0);
        const int nk = k ? p : np1;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks)
        {
            L.fa[k][ks] = (2 * t + ks < nk && g < Q) ? tab[(2 * t + ks) * Q + g] : 0.0;
            L.fb[k][ks] = (g < nk && 2 * t + ks < Q) ? tab[g * Q + 2 * t + ks] : 0.0;
        }
    }
    const double *pnts = args.tab + (2 * p + 1) * Q, *wgts = pnts + Q;
    // the lane's two quadrature points: (eta_{2t+s}, xi_g); padded points carry weight 0
    const bool vb = g < Q;
    const double xi = vb ? pnts[g] : 0.0, wb = vb ? wgts[g] : 0.0;
    double eta[2], wq[2];
#pragma unroll
    for (int s = 0; s < 2; ++s)
    {
        const bool va = 2 * t + s < Q;
        eta[s] = va ? pnts[2 * t + s] : 0.0;
        wq[s] = va ? wgts[2 * t + s] * wb : 0.0;
    }
    const int Nx = (N + 2) & ~1, off_cs = 2 * Nx;
    if (lane == 0)
        wbuf[N] = wbuf[Nx + N] = 0.0; // zero cells of both x buffers
    __syncthreads();
    double *ys = wbuf + da.off_ys;
    const unsigned wsm = (unsigned)__cvta_generic_to_shared(wbuf);
    const unsigned long long total_warps = (unsigned long long)gridDim.x * nwarps;
    auto prefetch = [&](unsigned long long e, int buf) {
        if (e >= args.n_elem)
            return;
        for (int i = lane; i < N; i += 32)
            cp_async8(wbuf + buf * Nx + i, args.x + e * N + i);
        if (lane < 8)
            cp_async8(wbuf + off_cs + buf * 8 + lane, args.corners + e * 8 + lane);
    };
    unsigned long long e = (unsigned long long)blockIdx.x * nwarps + warp;
    prefetch(e, 0);
    for (int buf = 0; e < args.n_elem; e += total_warps, buf ^= 1)
    {
        cp_async_wait_all();
        __syncwarp();
        prefetch(e + total_warps, buf ^ 1);
        const double *cs = wbuf + off_cs + buf * 8;
        const unsigned xs = wsm + (unsigned)(buf * Nx) * 8u;
        for (int i = lane; i < N; i += 32)
            ys[i] = 0.0;
        const double x0 = cs[0], y0 = cs[1], x1 = cs[2], y1 = cs[3], x2 = cs[4], y2 = cs[5], x3 = cs[6], y3 = cs[7];
        const double j10 = 0.25 * ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi));
        const double j11 = 0.25 * ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi));
        double j00[2], j01[2], det[2], wod[2];
#pragma unroll
        for (int s = 0; s < 2; ++s)
        {
            j00[s] = 0.25 * ((x1 - x0) * (1 - eta[s]) + (x2 - x3) * (1 + eta[s]));
            j01[s] = 0.25 * ((y1 - y0) * (1 - eta[s]) + (y2 - y3) * (1 + eta[s]));
            det[s] = j00[s] * j11 - j10 * j01[s];
            wod[s] = wq[s] * fast_rcp(det[s]);
        }
        __syncwarp();
        for (int ti = 0; ti < da.n_terms; ++ti)
        {
            const int *d = terms + ti * DM_STRIDE;
            const int mode = d[0], lbase = d[6];
            const bool gathered = d[5] != 0, direct = lbase < 0;
            const double cf = coef[d[2]];
            const int4 *ga = gtab + d[3] + lane;  // + 64 for the second component
            const int2 *dadr = dtab + d[3] + lane; // + 32 for the second component
            double uh[2] = {0.0, 0.0}, uv[2] = {0.0, 0.0}, o0, o1;
            // mode 0: M0   1: M2   2: M1   3: node x edge   4: its transpose   5: edge x surf   6: its transpose
            if (mode == 0 || mode == 4)
                dmma_in<0, 0>(L, xs, ga, dadr, gathered, uh[0], uh[1]);
            else if (mode == 1 || mode == 5)
                dmma_in<1, 1>(L, xs, ga, dadr, gathered, uh[0], uh[1]);
            else
            {
                dmma_in<0, 1>(L, xs, ga, dadr, gathered, uh[0], uh[1]);
                dmma_in<1, 0>(L, xs, ga + 64, dadr + 32, gathered, uv[0], uv[1]);
            }
#pragma unroll
            for (int s = 0; s < 2; ++s)
            {
                const double a0 = uh[s], a1 = uv[s];
                if (mode == 0)
                    uh[s] = a0 * det[s] * wq[s];
                else if (mode == 1)
                    uh[s] = a0 * wod[s];
                else if (mode == 2)
                {
                    const double khh = (j11 * j11 + j10 * j10) * wod[s], kvv = (j00[s] * j00[s] + j01[s] * j01[s]) * wod[s],
                                 khv = (j01[s] * j11 + j00[s] * j10) * wod[s];
                    uh[s] = khh * a0 + khv * a1;
                    uv[s] = khv * a0 + kvv * a1;
                }
                else
                {
                    const bool ok = vb && 2 * t + s < Q;
                    const double *fp = args.fields[d[1]] + ((e * Q + 2 * t + s) * (unsigned long long)Q + g) * 2;
                    const double fx = ok ? fp[0] : 0.0, fy = ok ? fp[1] : 0.0;
                    double kh, kv;
                    if (mode <= 4)
                    {
                        kh = (fx * j11 - fy * j10) * wq[s];
                        kv = -(fy * j00[s] - fx * j01[s]) * wq[s];
                    }
                    else
                    {
                        kh = -(fx * j10 + fy * j11) * wod[s];
                        kv = -(fx * j00[s] + fy * j01[s]) * wod[s];
                    }
                    if (mode == 3 || mode == 6)
                        uh[s] = kh * a0 + kv * a1;
                    else
                    {
                        uh[s] = kh * a0;
                        uv[s] = kv * a0;
                    }
                }
            }
            const int2 *oa = otab + d[4] + lane; // + 32 for the second component
            const int2 w0 = oa[0];
            if (mode == 0 || mode == 3)
                dmma_out<0, 0>(L, uh[0], uh[1], o0, o1);
            else if (mode == 1 || mode == 6)
                dmma_out<1, 1>(L, uh[0], uh[1], o0, o1);
            else
                dmma_out<0, 1>(L, uh[0], uh[1], o0, o1);
            if (direct)
            {
                sts64(wsm + w0.x, fma(cf, o0, lds64(wsm + w0.x)));
                sts64(wsm + w0.y, fma(cf, o1, lds64(wsm + w0.y)));
            }
            else
            {
                sts64(wsm + w0.x, o0);
                sts64(wsm + w0.y, o1);
            }
            if (mode == 2 || mode == 4 || mode == 5)
            {
                const int2 w1 = oa[32];
                dmma_out<1, 0>(L, uv[0], uv[1], o0, o1);
                if (direct)
                {
                    sts64(wsm + w1.x, fma(cf, o0, lds64(wsm + w1.x)));
                    sts64(wsm + w1.y, fma(cf, o1, lds64(wsm + w1.y)));
                }
                else
                {
                    sts64(wsm + w1.x, o0);
                    sts64(wsm + w1.y, o1);
                }
            }
            __syncwarp();
            if (!direct)
            {
                const int n_rows = d[7];
                double *yr = wbuf + d[8];
                for (int k = lane; k < n_rows; k += 32)
                {
                    const int4 q4 = ltab[lbase + k];
                    const double v = (lds64(wsm + q4.x) + lds64(wsm + q4.y)) - (lds64(wsm + q4.z) + lds64(wsm + q4.w));
                    yr[k] = fma(cf, v, yr[k]);
                }
                __syncwarp();
            }
        }
        for (int i = lane; i < N; i += 32)
            args.y[e * N + i] = ys[i];
        __syncwarp();
    }
}

static int launch_dmma(mfv2d_ctx *ctx, const mfv2d_plan *plan, const ApplyArgs &a, size_t n_elem)
{
    DmmaArgs da{};
    da.a = a;
    da.blob = plan->d_dmblob;
    da.n_terms = plan->dm_terms;
    da.off_terms = plan->dm_off_terms;
    da.off_gtab = plan->dm_off_gtab;
    da.off_ltab = plan->dm_off_ltab;
    da.off_otab = plan->dm_off_otab;
    da.off_dtab = plan->dm_off_dtab;
    da.blob_ints = plan->dm_blob_ints;
    da.warp_doubles = plan->dm_warp_doubles;
    da.off_ys = plan->dm_off_ys;
    da.off_outv = plan->dm_off_outv;
    da.n_coef = plan->mcoef_off ? plan->mcoef_off : (int)plan->vcoef.size();
    const int nwarps = DMMA_THREADS / 32;
    const size_t smem = (((size_t)da.n_coef + 1) & ~(size_t)1) * 8 + (size_t)da.blob_ints * 4 + (size_t)nwarps * da.warp_doubles * 8;
    if (smem > ctx->smem_optin)
        return -1;
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_vector_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_vector_dmma_kernel, DMMA_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    const unsigned long long need = (n_elem + nwarps - 1) / nwarps;
    if (grid > need)
        grid = need;
    MFV2D_LAUNCH(element_vector_dmma_kernel, (unsigned)grid, DMMA_THREADS, smem, ctx->stream, da);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}
#endif // MFV2D_EMU

static int launch_apply(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, const double *d_x,
                        double *d_y, int column_mode)
{
    mfv2d_ctx *ctx = plan->ctx;
    ApplyArgs a{};
    a.tab = plan->basis->d_tab;
    a.inc = plan->basis->d_inc;
    a.corners = e->d_corners;
    a.prog = plan->d_vprog + (column_mode ? plan->mprog_off : 0);
    a.coef = plan->d_vcoef + (column_mode ? plan->mcoef_off : 0);
    for (int i = 0; i < 16; ++i)
        a.fields[i] = (fields && i < plan->n_fields) ? fields[i] : nullptr;
    a.x = d_x;
    a.y = d_y;
    a.n_elem = e->n;
    a.tab_len = plan->basis->tab_len;
    a.p = plan->p;
    a.q = plan->q;
    for (int i = 0; i < 4; ++i)
        a.inc_off[i] = plan->basis->inc_off[i];
    a.n_forms = plan->n_forms;
    for (int i = 0; i <= plan->n_forms; ++i)
        a.form_offsets[i] = plan->form_offsets[i];
    a.n_total = plan->n_total;
    a.column_mode = column_mode;
    a.needs_dense = plan->vector_needs_dense ? 1 : 0;
    const int p = plan->p, q = plan->q, q2 = q * q, np1 = p + 1;
    a.vmax = (2 * p * np1 > np1 * np1 ? 2 * p * np1 : np1 * np1) + 2;
    a.vmax = (a.vmax + 1) & ~1;
    size_t doubles = ((a.tab_len + 1) & ~1) + 5 * q2 + 2 * q2 + 2 * np1 * q + 2 * a.vmax + 2 * (size_t)a.n_total + 2;
    if (plan->vector_needs_dense)
    {
        const int ldT = (np1 * np1 + 3) & ~3;
        const int nmax = 2 * p * np1;
        doubles += ((q2 + 1) & ~1) + q * ldT + 4 + (size_t)nmax * nmax + 8;
    }
#ifndef MFV2D_EMU
    if (!column_mode && plan->vector_fused && plan->warp_ok && e->n > 0)
    {
        // warp-autonomous kernel with the tables as uniform-register operands: an experiment that did NOT pay on B200
        // (28.6 against 39.0 GDoF/s at order 6, profiles/r2/ncu_full_element_vector_warp_poisson6_r02.csv: every DFMA needs
        // its own ULDC, 1 721 warp instructions per element); kept behind MFV2D_APPLY_WARP=1 and covered by a parity test
        const char *env = getenv("MFV2D_APPLY_WARP");
        if (env && env[0] == '1')
        {
            FusedArgs fa{};
            fa.a = a;
            fa.fv = plan->fv;
            fa.blob = plan->d_fblob;
            fa.n_coef = plan->mcoef_off ? plan->mcoef_off : (int)plan->vcoef.size();
            const int r = launch_warp(ctx, plan, fa, e->n);
            if (r >= 0)
                return r;
        }
    }
    if (!column_mode && plan->vector_dmma && e->n > 0)
    {
        // opt-in since round 2: with the bulk-copy prefetch and the fixed-record output stage the DFMA kernel below is faster
        // on every order measured (Stokes p4: 39.7 vs 38.6 GDoF/s, mixed Poisson p4: 46.4 vs 41.1, Stokes p3: 35.5 vs 24.0;
        // profiles/apply_dmma_vs_fused.py)
        const char *env = getenv("MFV2D_APPLY_DMMA");
        if (env && env[0] == '1')
        {
            const int r = launch_dmma(ctx, plan, a, e->n);
            if (r >= 0)
                return r;
        }
    }
#endif
    if (!column_mode && plan->vector_fused && e->n > 0)
    {
        FusedArgs fa{};
        fa.a = a;
        fa.fv = plan->fv;
        fa.blob = plan->d_fblob;
        fa.n_coef = plan->mcoef_off ? plan->mcoef_off : (int)plan->vcoef.size();
        fa.inc_rows = plan->basis->inc_off[4];
        const int ldq = (q + 1) & ~1;
        const size_t fixed_bytes = ((size_t)(2 * p + 1) * ldq + 2 * ldq + (((size_t)fa.n_coef + 1) & ~(size_t)1)) * 8 +
                                   (size_t)fa.fv.blob_ints * 4 + (size_t)fa.inc_rows * 16;
        int r = -1;
        switch (q)
        {
#define MFV2D_Q_CASE(Q)                                                                                        \
    case Q:                                                                                                    \
        r = launch_fused_q<Q>(ctx, fa, fixed_bytes, e->n);                                                     \
        break;
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
            break;
        }
        if (r >= 0)
            return r;
    }
    const size_t smem = doubles * sizeof(double);
    if (smem > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "element vectors of order %d need %zu bytes of shared memory", p, smem);
    MFV2D_CUDA_CHECK(ctx, cudaFuncSetAttribute(element_vector_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 1;
    MFV2D_CUDA_CHECK(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, element_vector_kernel, APPLY_THREADS, smem));
    if (per_sm < 1)
        per_sm = 1;
    unsigned long long grid = (unsigned long long)ctx->sm_count * per_sm;
    if (grid > e->n)
        grid = e->n;
    if (grid == 0)
        return MFV2D_SUCCESS;
    MFV2D_LAUNCH(element_vector_kernel, (unsigned)grid, APPLY_THREADS, smem, ctx->stream, a);
    ctx->launches++;
    MFV2D_CUDA_CHECK(ctx, cudaGetLastError());
    return MFV2D_SUCCESS;
}

int launch_element_vector(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, const double *d_x,
                          double *d_y)
{
    return launch_apply(plan, e, fields, d_x, d_y, 0);
}

int launch_matrix_by_columns(const mfv2d_plan *plan, const mfv2d_elements *e, const double *const *fields, double *d_out)
{
    return launch_apply(plan, e, fields, nullptr, d_out, 1);
}
