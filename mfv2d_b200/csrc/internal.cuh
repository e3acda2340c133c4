// Internal declarations shared by the translation units of libmfv2d_cuda.so.
#pragma once
#include "cuda_compat.cuh"

#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/mfv2d_cuda.h"

struct mfv2d_ctx
{
    int device = 0;
    int sm_count = 148;
    size_t smem_optin = 0; // max dynamic shared memory per block (opt-in)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    uint64_t launches = 0;
    std::string error;
    // scratch buffer used by the host-buffer paths and the L2 flush
    void *flush_buf = nullptr;
    size_t flush_bytes = 0;
    // NCCL (dlopen()ed lazily, comm.cu)
    void *nccl_comm = nullptr;
    int n_ranks = 1, rank = 0;
    // small device scalars for reductions
    double *d_scalars = nullptr; // 64 doubles
    double *h_scalars = nullptr; // pinned, 64 doubles
};

int mfv2d_fail(mfv2d_ctx *ctx, int code, const char *fmt, ...);
void mfv2d_set_global_error(const char *msg);

#define MFV2D_CUDA_CHECK(ctx, call)                                                                            \
    do                                                                                                         \
    {                                                                                                          \
        cudaError_t e__ = (call);                                                                              \
        if (e__ != cudaSuccess)                                                                                \
            return mfv2d_fail((ctx), MFV2D_CUDA_ERROR, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__,    \
                              cudaGetErrorString(e__));                                                        \
    } while (0)

#define MFV2D_TRY(expr)                                                                                        \
    do                                                                                                         \
    {                                                                                                          \
        int r__ = (expr);                                                                                      \
        if (r__ != MFV2D_SUCCESS)                                                                              \
            return r__;                                                                                        \
    } while (0)

// RAII-free helper: device buffer that is either the caller's device pointer or a staged copy of a
// host pointer.  `stage_in` uploads, `stage_out` downloads + frees.
struct Staged
{
    mfv2d_ctx *ctx = nullptr;
    void *dev = nullptr;
    void *host = nullptr;
    size_t bytes = 0;
    bool owned = false;
    bool is_out = false;
};
int stage_in(mfv2d_ctx *ctx, const void *ptr, size_t bytes, int space, Staged *s);
int stage_out_alloc(mfv2d_ctx *ctx, void *ptr, size_t bytes, int space, Staged *s);
int stage_finish(Staged *s); // download if output + host, free if owned
void stage_abort(Staged *s);

static inline int form_size(int form, int p)
{
    switch (form)
    {
    case MFV2D_FORM_ORDER_0:
        return (p + 1) * (p + 1);
    case MFV2D_FORM_ORDER_1:
        return 2 * p * (p + 1);
    case MFV2D_FORM_ORDER_2:
        return p * p;
    default:
        return -1;
    }
}

struct mfv2d_basis
{
    mfv2d_ctx *ctx;
    int p, q;
    // host copies
    std::vector<double> roots, pnts, wgts, node, edge;
    // device: one packed table [node (p+1)*q | edge p*q | pnts q | wgts q]
    double *d_tab = nullptr;
    int tab_len = 0;
    // incidence row maps for the four kinds, concatenated: for each kind a [rows][4] int32 array of
    // signed (index+1) entries, 0 = unused.  d_inc_off[kind] = offset of the kind (in rows).
    int32_t *d_inc = nullptr;
    int inc_off[5] = {0, 0, 0, 0, 0};
    std::vector<int32_t> h_inc;
    // the same maps in branch-free form: [rows][4] column indices (0 when unused) and [rows][4] signs
    // (+1, -1, or 0 when unused), used by the assembly kernel's emission
    int32_t *d_inc_idx = nullptr;
    float *d_inc_sgn = nullptr;
};

struct mfv2d_elements
{
    mfv2d_ctx *ctx;
    const mfv2d_basis *basis;
    size_t n;
    double *d_corners = nullptr; // [n][8]
};

// host polynomial routines (tables.cpp)
void host_gll(int n, int max_iter, double tol, double *x, double *w, int *not_converged);
void host_lagrange(int n, const double *roots, int m, const double *pos, double *out);
void host_dlagrange(int n, const double *roots, int m, const double *pos, double *out);
void host_basis_tables(int p, int q, double *roots, double *pnts, double *wgts, double *node, double *edge);
void host_incidence_rowmap(int p, int kind, std::vector<int32_t> &rows); // appends rows*4 entries
static inline int incidence_rows(int p, int kind)
{
    switch (kind)
    {
    case MFV2D_INC_10:
        return 2 * p * (p + 1);
    case MFV2D_INC_21:
        return p * p;
    case MFV2D_INC_10_T:
        return (p + 1) * (p + 1);
    default:
        return 2 * p * (p + 1);
    }
}
static inline int incidence_cols(int p, int kind)
{
    switch (kind)
    {
    case MFV2D_INC_10:
        return (p + 1) * (p + 1);
    case MFV2D_INC_21:
        return 2 * p * (p + 1);
    case MFV2D_INC_10_T:
        return 2 * p * (p + 1);
    default:
        return p * p;
    }
}
