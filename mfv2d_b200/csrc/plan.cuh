// Bytecode plan: the result of symbolically executing the reference's stack machine once per
// system (instead of once per element per call) and lowering each block to fused device work.
#pragma once
#include "internal.cuh"

// kinds of dense per-element matrices a plan can need
enum
{
    DK_MASS0 = 0,  // (p+1)^2 x (p+1)^2
    DK_MASS1 = 1,  // 2p(p+1) x 2p(p+1)
    DK_MASS2 = 2,  // p^2 x p^2
    DK_IPROD1 = 3, // node x edge  (n0 x n1), compute_mass_matrix_node_edge
    DK_IPROD2 = 4, // edge x surf  (n1 x n2), compute_mass_matrix_edge_surf
};

// one factor of a product chain, applied right-to-left
struct Factor
{
    int is_dense;  // 0: incidence, 1: dense
    int inc_kind;  // incidence kind when !is_dense
    int dk;        // DK_* when dense
    int inverse;   // mass inverse
    int field;     // field index for interior products
    int transpose; // interior product transposed
    int rows, cols;
};

struct Term
{
    double coef;
    std::vector<Factor> f; // f[0] is applied first (right-most)
    int rows, cols;        // dims of the product (cols == -1 while it is still applied to the seed)
};

struct BlockProgram
{
    bool present = false;
    bool identity = false; // matrix mode: block is coef * I
    double identity_coef = 0.0;
    std::vector<Term> terms;
};

// ---- device-side description of the fused matrix path ---------------------------------------
struct DevDense
{
    int dk, inverse, field, transpose;
    int rows, cols;
    int smem_off; // offset (in doubles) of the matrix inside the stage's shared-memory arena
};

struct DevTerm
{
    double coef;
    int dense; // index into the dense list, -1: identity
    int L, R;  // incidence kinds applied on the left / right, -1: none
};

struct DevUnit // one warp work item: a strip of rows x 32 columns of one block
{
    short block; // index into blocks
    short c0;    // first column (relative to the block)
    short r0;    // first row (relative to the block)
    short nr;    // number of rows
};

struct DevBlock
{
    int row0, col0, nrows, ncols;
    int term_begin, term_end; // terms emitted in this stage
    int accumulate;           // 0: store, 1: add to what is already in the output
};

struct DevStage
{
    int dense_begin, dense_end;
    int block_begin, block_end;
    int unit_begin, unit_end;
};

struct mfv2d_plan
{
    mfv2d_ctx *ctx;
    const mfv2d_basis *basis;
    int p, q;
    int n_forms;
    std::vector<int> form_orders, form_offsets; // offsets has n_forms + 1 entries
    int n_total;
    int n_fields;
    std::vector<int> field_form; // -1: supplied, else unknown index
    std::vector<BlockProgram> mat_blocks, vec_blocks; // n_forms^2, symbolic results of both modes
    bool needs_dofs = false;

    // fused matrix path
    bool matrix_fused = false; // every term is L * D * R with at most one incidence on each side
    std::vector<DevDense> dense;
    std::vector<DevTerm> terms;
    std::vector<DevBlock> blocks;
    std::vector<DevUnit> units;
    std::vector<DevStage> stages;
    int out_ld = 0;                    // 0: n_total (rectangular single-matrix plans override)
    unsigned long long out_stride = 0; // 0: n_total^2
    int smem_doubles = 0; // dynamic shared memory of the matrix kernel (in doubles)
    void *d_blob = nullptr; // device copy: [stages | dense | terms | blocks | units]
    int off_dense = 0, off_terms = 0, off_blocks = 0, off_units = 0; // byte offsets inside the blob
    size_t blob_bytes = 0;

    // vector path (flattened programs), see apply.cu
    std::vector<int32_t> vprog;
    std::vector<double> vcoef;
    int32_t *d_vprog = nullptr;
    double *d_vcoef = nullptr;
    bool vector_needs_dense = false; // an inverse mass appears in a vector program
    int mprog_off = 0, mcoef_off = 0; // matrix-mode programs (column-by-column fallback) follow the vector ones
};

// shared-memory layout of the assembly kernel (in doubles); used by the planner and the kernel.
// One slot per element of a pass: geo 5 q^2 | G | T | arena
struct SlotLayout
{
    int G, T, ldT, arena;
};
static inline __host__ __device__ SlotLayout slot_layout(int p, int q)
{
    SlotLayout l;
    const int q2 = q * q;
    l.G = 5 * q2;
    l.T = (l.G + q2 + 3) & ~3;
    l.ldT = ((p + 1) * (p + 1) + 3) & ~3;
    l.arena = l.T + q * l.ldT + 4;
    return l;
}
// fixed part: tables | pair LUTs (two ints per double) | corners of the slots
static inline __host__ __device__ int asm_fixed_doubles(int p, int tab_len, int epb)
{
    const int n_pairs = (p + 1) * (p + 2) / 2 + p * (p + 1) / 2;
    int n = ((tab_len + 1) & ~1) + ((n_pairs + 1) >> 1) + 8 * epb;
    return (n + 15) & ~15;
}

int plan_lower_matrix(mfv2d_plan *plan);
int plan_lower_vector(mfv2d_plan *plan);
