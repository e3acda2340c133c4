// Host-side polynomial tables: GLL nodes/weights, Lagrange values and derivatives, Legendre, and the
// Basis1D nodal/edge tables.  O(p*q) doubles computed once per (p, q) bucket and uploaded; they are
// shared by every element of the bucket.
#include <cmath>

#include "internal.cuh"

// P_{n-1}(x), P_n(x) with Bonnet's recursion.  The GLL iteration below follows the published
// scheme of the reference (src/polynomials/gauss_lobatto.c:18-33,52-85) operation for operation so
// that nodes and weights come out bit-identical.
static inline void legendre_pair(int n, double x, double *pm1, double *pn)
{
    double a = 1.0, b = x;
    for (int i = 2; i <= n; ++i)
    {
        const double c = ((2 * i - 1) * x * b - (i - 1) * a) / (double)i;
        a = b;
        b = c;
    }
    *pm1 = a;
    *pn = b;
}

void host_gll(int n, int max_iter, double tol, double *x, double *w, int *not_converged)
{
    *not_converged = 0;
    if (n == 1)
    {
        x[0] = 0.0;
        w[0] = 2.0;
        return;
    }
    x[0] = -1.0;
    x[n - 1] = +1.0;
    w[0] = w[n - 1] = 2.0 / (double)(n * (n - 1));
    const double kx_1 = 1.0 - 3.0 * (n - 2) / (8.0 * (n - 1) * (n - 1) * (n - 1));
    const double kx_2 = M_PI / (4.0 * (n - 1) + 1);
    for (int i = 2; i < n; ++i)
    {
        double xn = kx_1 * cos(kx_2 * (4 * i - 3));
        double err = 1.0;
        int it = 0;
        for (; it < max_iter && err > tol; ++it)
        {
            double p0, p1;
            legendre_pair(n - 1, xn, &p0, &p1);
            const double den = 1 - xn * xn;
            const double dy = (n - 1) * (p0 - xn * p1) / den;
            const double d2y = (2 * xn * dy - (n - 1) * n * p1) / den;
            const double d3y = (4 * xn * d2y - ((n - 1) * n - 2) * dy) / den;
            const double dx = 2 * dy * d2y / (2 * d2y * d2y - dy * d3y);
            xn -= dx;
            err = fabs(dx);
        }
        if (err > tol)
            *not_converged += 1;
        x[n - i] = xn;
        double p0, p1;
        legendre_pair(n - 1, xn, &p0, &p1);
        w[n - i] = 2.0 / (n * (n - 1) * p1 * p1);
    }
}

// l_j(x) = prod_{i != j} (x - r_i) / prod_{i != j} (r_j - r_i).  Both products run over i in increasing order and the
// quotient is formed as numerator * (1 / denominator): the operation order of src/polynomials/lagrange.c:9-101, so that the
// values are the reference's bit for bit (the hanging-side constraint rows invert matrices of these values with condition
// numbers up to 1e8: a last-bit difference here is a 1e-8 difference there).
void host_lagrange(int n, const double *roots, int m, const double *pos, double *out)
{
    std::vector<double> inv_den(n);
    for (int j = 0; j < n; ++j)
    {
        double d = 1.0;
        for (int i = 0; i < n; ++i)
            if (i != j)
                d *= (roots[j] - roots[i]);
        inv_den[j] = 1.0 / d;
    }
    for (int k = 0; k < m; ++k)
        for (int j = 0; j < n; ++j)
        {
            double num = 1.0;
            for (int i = 0; i < n; ++i)
                if (i != j)
                    num *= (pos[k] - roots[i]);
            out[(size_t)k * n + j] = num * inv_den[j];
        }
}

// l_j'(x) = (1/den_j) sum_{m != j} prod_{i != j, m} (x - r_i)
void host_dlagrange(int n, const double *roots, int m, const double *pos, double *out)
{
    std::vector<double> inv_den(n), dif(n);
    for (int j = 0; j < n; ++j)
    {
        double d = 1.0;
        for (int i = 0; i < n; ++i)
            if (i != j)
                d *= (roots[j] - roots[i]);
        inv_den[j] = 1.0 / d;
    }
    for (int k = 0; k < m; ++k)
    {
        for (int i = 0; i < n; ++i)
            dif[i] = pos[k] - roots[i];
        for (int j = 0; j < n; ++j)
        {
            double s = 0.0;
            for (int mm = 0; mm < n; ++mm)
            {
                if (mm == j)
                    continue;
                double t = 1.0;
                for (int i = 0; i < n; ++i)
                    if (i != j && i != mm)
                        t *= dif[i];
                s += t;
            }
            out[(size_t)k * n + j] = s * inv_den[j];
        }
    }
}

void host_basis_tables(int p, int q, double *roots, double *pnts, double *wgts, double *node, double *edge)
{
    int nc;
    std::vector<double> wr(p + 1);
    host_gll(p + 1, 10, 1e-15, roots, wr.data(), &nc);
    host_gll(q, 10, 1e-15, pnts, wgts, &nc);
    std::vector<double> val((size_t)q * (p + 1)), der((size_t)q * (p + 1));
    host_lagrange(p + 1, roots, q, pnts, val.data());
    host_dlagrange(p + 1, roots, q, pnts, der.data());
    for (int i = 0; i <= p; ++i)
        for (int k = 0; k < q; ++k)
            node[i * q + k] = val[(size_t)k * (p + 1) + i];
    // edge_j(xi_k) = -sum_{i <= j} l_i'(xi_k)   (src/fem_space/basis.c:77-86)
    for (int k = 0; k < q; ++k)
    {
        double acc = 0.0;
        for (int j = 0; j < p; ++j)
        {
            acc -= der[(size_t)k * (p + 1) + j];
            edge[j * q + k] = acc;
        }
    }
}

// Row maps of the incidence matrices (src/evaluation/incidence.c:33-56,79-92 and their transposes
// :117-157,182-223): row r of E has at most 4 non-zeros; entry = +(col+1) or -(col+1), 0 = unused.
void host_incidence_rowmap(int p, int kind, std::vector<int32_t> &rows)
{
    const int nh = p * (p + 1);
    auto node = [&](int r, int c) { return (p + 1) * r + c; };
    auto hedge = [&](int r, int c) { return p * r + c; };
    auto vedge = [&](int r, int c) { return nh + (p + 1) * r + c; };
    auto surf = [&](int r, int c) { return p * r + c; };
    auto push = [&](int32_t a, int32_t b, int32_t c, int32_t d) {
        rows.push_back(a);
        rows.push_back(b);
        rows.push_back(c);
        rows.push_back(d);
    };
    switch (kind)
    {
    case MFV2D_INC_10:
        for (int r = 0; r <= p; ++r)
            for (int c = 0; c < p; ++c)
                push(+(node(r, c) + 1), -(node(r, c + 1) + 1), 0, 0);
        for (int r = 0; r < p; ++r)
            for (int c = 0; c <= p; ++c)
                push(+(node(r + 1, c) + 1), -(node(r, c) + 1), 0, 0);
        break;
    case MFV2D_INC_21:
        for (int r = 0; r < p; ++r)
            for (int c = 0; c < p; ++c)
                push(+(hedge(r, c) + 1), -(hedge(r + 1, c) + 1), +(vedge(r, c) + 1), -(vedge(r, c + 1) + 1));
        break;
    case MFV2D_INC_10_T:
        for (int r = 0; r <= p; ++r)
            for (int c = 0; c <= p; ++c)
            {
                int32_t e[4] = {0, 0, 0, 0};
                int k = 0;
                if (c < p)
                    e[k++] = +(hedge(r, c) + 1);
                if (c >= 1)
                    e[k++] = -(hedge(r, c - 1) + 1);
                if (r >= 1)
                    e[k++] = +(vedge(r - 1, c) + 1);
                if (r < p)
                    e[k++] = -(vedge(r, c) + 1);
                push(e[0], e[1], e[2], e[3]);
            }
        break;
    default: // MFV2D_INC_21_T
        for (int r = 0; r <= p; ++r)
            for (int c = 0; c < p; ++c)
            {
                int32_t e[4] = {0, 0, 0, 0};
                int k = 0;
                if (r < p)
                    e[k++] = +(surf(r, c) + 1);
                if (r >= 1)
                    e[k++] = -(surf(r - 1, c) + 1);
                push(e[0], e[1], e[2], e[3]);
            }
        for (int r = 0; r < p; ++r)
            for (int c = 0; c <= p; ++c)
            {
                int32_t e[4] = {0, 0, 0, 0};
                int k = 0;
                if (c < p)
                    e[k++] = +(surf(r, c) + 1);
                if (c >= 1)
                    e[k++] = -(surf(r, c - 1) + 1);
                push(e[0], e[1], e[2], e[3]);
            }
        break;
    }
}

__global__ void legendre_kernel(int order, size_t m, const double *__restrict__ pos, double *__restrict__ out)
{
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < m; i += stride)
    {
        const double x = pos[i];
        double a = 1.0, b = x;
        out[i] = a;
        if (order > 0)
            out[m + i] = b;
        for (int k = 2; k <= order; ++k)
        {
            const double c = ((2 * k - 1) * x * b - (k - 1) * a) / (double)k;
            out[(size_t)k * m + i] = c;
            a = b;
            b = c;
        }
    }
}

extern "C" {

int mfv2d_gll(int n, int max_iter, double tol, double *x, double *w)
{
    if (n < 1 || !x || !w)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "mfv2d_gll: n must be >= 1");
    int nc = 0;
    host_gll(n, max_iter, tol, x, w, &nc);
    return nc ? MFV2D_NOT_CONVERGED : MFV2D_SUCCESS;
}

int mfv2d_lagrange1d(int n, const double *roots, int m, const double *pos, double *out)
{
    if (n < 1 || m < 0)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "mfv2d_lagrange1d: bad sizes");
    host_lagrange(n, roots, m, pos, out);
    return MFV2D_SUCCESS;
}

int mfv2d_dlagrange1d(int n, const double *roots, int m, const double *pos, double *out)
{
    if (n < 1 || m < 0)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "mfv2d_dlagrange1d: bad sizes");
    host_dlagrange(n, roots, m, pos, out);
    return MFV2D_SUCCESS;
}

int mfv2d_legendre(int order, int m, const double *pos, double *out)
{
    if (order < 0 || m < 0)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "mfv2d_legendre: bad sizes");
    for (int i = 0; i < m; ++i)
    {
        const double x = pos[i];
        double a = 1.0, b = x;
        out[i] = a;
        if (order > 0)
            out[(size_t)m + i] = b;
        for (int k = 2; k <= order; ++k)
        {
            const double c = ((2 * k - 1) * x * b - (k - 1) * a) / (double)k;
            out[(size_t)k * m + i] = c;
            a = b;
            b = c;
        }
    }
    return MFV2D_SUCCESS;
}

int mfv2d_basis_tables(int p, int q, double *roots, double *pnts, double *wgts, double *node, double *edge)
{
    if (p < 1 || q < 1)
        return mfv2d_fail(nullptr, MFV2D_BAD_ARGUMENT, "mfv2d_basis_tables: need p >= 1, q >= 1");
    host_basis_tables(p, q, roots, pnts, wgts, node, edge);
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_legendre(mfv2d_ctx *ctx, int order, size_t m, const double *pos, double *out, int space)
{
    if (order < 0)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "mfv2d_cuda_legendre: order < 0");
    if (m == 0)
        return MFV2D_SUCCESS;
    Staged in, o;
    MFV2D_TRY(stage_in(ctx, pos, m * sizeof(double), space, &in));
    int r = stage_out_alloc(ctx, out, (size_t)(order + 1) * m * sizeof(double), space, &o);
    if (r)
    {
        stage_abort(&in);
        return r;
    }
    const int threads = 256;
    size_t blocks = (m + threads - 1) / threads;
    if (blocks > (size_t)ctx->sm_count * 16)
        blocks = (size_t)ctx->sm_count * 16;
    MFV2D_LAUNCH(legendre_kernel, (unsigned)blocks, threads, 0, ctx->stream, order, m, (const double *)in.dev, (double *)o.dev);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    r = stage_finish(&o);
    stage_abort(&in);
    if (e != cudaSuccess)
        return mfv2d_fail(ctx, MFV2D_CUDA_ERROR, "legendre kernel: %s", cudaGetErrorString(e));
    return r;
}

int mfv2d_cuda_basis_create(mfv2d_ctx *ctx, int p, int q, mfv2d_basis **out)
{
    *out = nullptr;
    if (p < 1 || q < 1 || p > 24 || q > 32)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "basis_create: need 1 <= p <= 24 and 1 <= q <= 32 (got %d, %d)", p, q);
    mfv2d_basis *b = new mfv2d_basis();
    b->ctx = ctx;
    b->p = p;
    b->q = q;
    b->roots.resize(p + 1);
    b->pnts.resize(q);
    b->wgts.resize(q);
    b->node.resize((size_t)(p + 1) * q);
    b->edge.resize((size_t)p * q);
    host_basis_tables(p, q, b->roots.data(), b->pnts.data(), b->wgts.data(), b->node.data(), b->edge.data());
    std::vector<double> tab;
    tab.insert(tab.end(), b->node.begin(), b->node.end());
    tab.insert(tab.end(), b->edge.begin(), b->edge.end());
    tab.insert(tab.end(), b->pnts.begin(), b->pnts.end());
    tab.insert(tab.end(), b->wgts.begin(), b->wgts.end());
    b->tab_len = (int)tab.size();
    for (int kind = 0; kind < 4; ++kind)
    {
        b->inc_off[kind] = (int)(b->h_inc.size() / 4);
        host_incidence_rowmap(p, kind, b->h_inc);
    }
    b->inc_off[4] = (int)(b->h_inc.size() / 4);
    int r = mfv2d_cuda_malloc(ctx, tab.size() * sizeof(double), (void **)&b->d_tab);
    if (!r)
        r = mfv2d_cuda_malloc(ctx, b->h_inc.size() * sizeof(int32_t), (void **)&b->d_inc);
    if (!r)
        r = mfv2d_cuda_upload(ctx, b->d_tab, tab.data(), tab.size() * sizeof(double));
    if (!r)
        r = mfv2d_cuda_upload(ctx, b->d_inc, b->h_inc.data(), b->h_inc.size() * sizeof(int32_t));
    std::vector<int32_t> idx(b->h_inc.size());
    std::vector<float> sgn(b->h_inc.size());
    for (size_t i = 0; i < b->h_inc.size(); ++i)
    {
        const int32_t v = b->h_inc[i];
        idx[i] = v == 0 ? 0 : (v < 0 ? -v : v) - 1;
        sgn[i] = v == 0 ? 0.f : (v < 0 ? -1.f : 1.f);
    }
    if (!r)
        r = mfv2d_cuda_malloc(ctx, idx.size() * sizeof(int32_t), (void **)&b->d_inc_idx);
    if (!r)
        r = mfv2d_cuda_malloc(ctx, sgn.size() * sizeof(float), (void **)&b->d_inc_sgn);
    if (!r)
        r = mfv2d_cuda_upload(ctx, b->d_inc_idx, idx.data(), idx.size() * sizeof(int32_t));
    if (!r)
        r = mfv2d_cuda_upload(ctx, b->d_inc_sgn, sgn.data(), sgn.size() * sizeof(float));
    if (r)
    {
        mfv2d_cuda_basis_destroy(b);
        return r;
    }
    *out = b;
    return MFV2D_SUCCESS;
}

void mfv2d_cuda_basis_destroy(mfv2d_basis *b)
{
    if (!b)
        return;
    if (b->d_tab)
        mfv2d_cuda_free(b->ctx, b->d_tab);
    if (b->d_inc)
        mfv2d_cuda_free(b->ctx, b->d_inc);
    if (b->d_inc_idx)
        mfv2d_cuda_free(b->ctx, b->d_inc_idx);
    if (b->d_inc_sgn)
        mfv2d_cuda_free(b->ctx, b->d_inc_sgn);
    delete b;
}

int mfv2d_cuda_elements_create(mfv2d_ctx *ctx, const mfv2d_basis *basis, size_t n_elem, const double *corners,
                               int space, mfv2d_elements **out)
{
    *out = nullptr;
    if (!basis || basis->ctx != ctx)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "elements_create: basis does not belong to this context");
    mfv2d_elements *e = new mfv2d_elements();
    e->ctx = ctx;
    e->basis = basis;
    e->n = n_elem;
    int r = mfv2d_cuda_malloc(ctx, n_elem * 8 * sizeof(double), (void **)&e->d_corners);
    if (r)
    {
        delete e;
        return r;
    }
    if (corners)
    {
        r = mfv2d_cuda_elements_set_corners(e, corners, space);
        if (r)
        {
            mfv2d_cuda_elements_destroy(e);
            return r;
        }
    }
    *out = e;
    return MFV2D_SUCCESS;
}

int mfv2d_cuda_elements_set_corners(mfv2d_elements *e, const double *corners, int space)
{
    mfv2d_ctx *ctx = e->ctx;
    MFV2D_CUDA_CHECK(ctx, cudaMemcpyAsync(e->d_corners, corners, e->n * 8 * sizeof(double),
                                          space == MFV2D_HOST ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice,
                                          ctx->stream));
    if (space == MFV2D_HOST)
        MFV2D_CUDA_CHECK(ctx, cudaStreamSynchronize(ctx->stream));
    return MFV2D_SUCCESS;
}

void mfv2d_cuda_elements_destroy(mfv2d_elements *e)
{
    if (!e)
        return;
    if (e->d_corners)
        mfv2d_cuda_free(e->ctx, e->d_corners);
    delete e;
}

} // extern "C"
