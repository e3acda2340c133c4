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
// Everything below is packed by the host into ONE blob (plan_finalize_matrix) that the assembly kernel
// copies into shared memory once per CTA; all descriptor reads of the kernel are shared-memory reads.
struct DevDense
{
    int dk, inverse, field, transpose;
    int rows, cols;
    int smem_off; // offset (in doubles) of the matrix inside the slot's arena
    int ld;       // leading dimension inside the arena (cols rounded up to odd: conflict-free column walks)
};

struct DevTerm
{
    double coef;
    int dense; // index into the dense list, -1: identity
    int L, R;  // incidence kinds applied on the left / right, -1: none
};

struct DevUnit // one warp work item of the emission: nr rows x (1 << cl) columns of one block
{
    short block; // index into blocks
    short c0;    // first column (relative to the block)
    short r0;    // first row (relative to the block)
    short nr;    // number of rows
    short cl;    // log2 of the tile width in lanes (3, 4 or 5); 32 >> cl rows are written per warp store
    short pad[3];
};

struct DevBlock
{
    int row0, col0, nrows, ncols;
    int term_begin, term_end; // terms emitted in this stage
    int accumulate;           // 0: store, 1: add to what is already in the output
};

// kinds of weight fields at the quadrature points (Appendix A of SURVEY.md)
enum
{
    W_M0,  // det w                                 (0-form mass)
    W_M2,  // w / det                               (2-form mass)
    W_HH,  // (j11^2 + j10^2) / det w               (1-form mass, h-h block)
    W_VV,  // (j00^2 + j01^2) / det w               (v-v block)
    W_HV,  // (j01 j11 + j00 j10) / det w           (h-v block)
    W_P1H, // (fx j11 - fy j10) w                   (interior product 1 -> 0, h columns)
    W_P1V, // -(fy j00 - fx j01) w                  (v columns)
    W_P2H, // -(fx j10 + fy j11) / det w            (interior product 2 -> 1, h rows)
    W_P2V  // -(fx j00 + fy j01) / det w            (v rows)
};

struct DevW // a weight field at the quadrature points (W_* of assembly.cu), `field` for interior products
{
    int kind, field;
};

// pair kinds: which two 1-D tables a product table row multiplies
enum
{
    PK_NN = 0, // node x node, symmetric: rows enumerate i <= k
    PK_EE = 1, // edge x edge, symmetric
    PK_NE = 2, // node_i x edge_k, all pairs
    PK_EN = 3, // edge_i x node_k, all pairs (rows of the NE table)
};

// One sum-factorised tensor block:  D[(i,j),(k,l)] = sum_a A_i(a) A2_k(a) sum_b B_j(b) B2_l(b) W(a,b)
struct DevTB
{
    int w;                    // weight slot inside the stage
    int pkA, pkB;             // pair kinds of (A, A2) and (B, B2)
    int nA, nA2, nB, nB2;
    int dst_off, ld, r0, c0;  // destination: arena offset, leading dimension, first row / column
    int trans, mirror;        // write transposed only / also write the transposed entry
    int t_off;                // offset of T[(j,l)][a] inside the slot's T region
    int round;                // TB round inside the stage (rounds reuse the T region)
    int n_sub1, n_sub2;       // number of chunks the (j,l) / (i,k) pair lists are cut into
    int s1_tab, s2_tab;       // first entry of the block's DevS1Pair / DevS2Pair tables
    int npA, npB;             // entries of the (i,k) / (j,l) pair tables
};

// stage-1 work of one (j,l) pair: product-table row and where T goes
struct DevS1Pair
{
    int pt;       // offset (doubles from the start of shared memory) of the row (B_j B2_l)(b)
    int t1, t2;   // offsets of T[(j,l)][0] and of its mirror image T[(l,j)][0] (-1: none) inside the block's T
    int jl;       // j | l << 8
};

// stage-2 work of one (i,k) pair: product-table row and the (up to four) images of the entry in the arena,
// relative to a per-thread base: d* are added to (r0 + j) ld + c0 + l, f* to (c0 + l) ld + r0 + j; -1: none
struct DevS2Pair
{
    int pt;
    int d0, f0; // ((i,j),(k,l)) and its transposed image
    int d1, f1; // ((k,j),(i,l)), the i <-> k image of a symmetric (A == A2) block, and its transposed image
    int pad[3];
};

// work items of the two contraction stages for ONE element of a pass, packed into an int:
//   stage 1: tb | a << 4 | first pair << 9 | number of pairs << 17
//   stage 2: tb | j0 << 4 | l0 << 9 | first pair << 14 | number of pairs << 22     (columns (j0,l0) and the next one)
struct DevRound
{
    int tb_begin, tb_end;
    int x1_begin, x1_count; // entries in the x tables (per element)
    int x2_begin, x2_count;
};

struct DevStage
{
    int dense_begin, dense_end;
    int block_begin, block_end;
    int unit_begin, unit_end;
    int w_begin, w_end;
    int round_begin, round_end;
    int tc_begin, tc_count; // tensor-core work items of the stage (per element): tensor block | (j,l) tile << 16
};

// row of an incidence matrix in gather form: up to 4 (column, sign) pairs.  The sign is kept as the HIGH word
// of the double +1.0 / -1.0 / 0.0 (unused entry), so that a row is two 16-byte loads and the sign costs nothing
// to rebuild (__hiloint2double(hi, 0)).
struct DevIncRow
{
    int e[8]; // idx0, sgn0.hi, idx1, sgn1.hi, ...
};

// byte offsets of the sections of the blob and the shared-memory layout of the assembly kernel
struct AsmLayout
{
    int off_stages, off_dense, off_terms, off_blocks, off_units, off_ws, off_tbs, off_rounds;
    int off_s1, off_s2, off_x; // DevS1Pair / DevS2Pair tables and the packed work items
    int off_tc, use_tc;        // work items of the tensor-core contraction (q <= 8), 1: the kernel takes that path
    int off_inc;      // DevIncRow rows of the four incidence kinds, inc_row0[kind] = first row
    int inc_row0[4];
    int off_pt[3];    // product tables NN, EE, NE: rows of ldq doubles
    int off_tab;      // node | edge | pnts | wgts
    int blob_bytes;
    // shared memory (in doubles, after the blob): corners, then epb slots of [W | T | arena]
    int ldq;          // padded quadrature row length
    int epb;          // elements per pass
    int off_cs, off_slots, slot_stride;
    int slot_w, slot_t, slot_arena; // offsets inside a slot
    int n_stages;
    int smem_bytes;
};

// layout of the fused matrix-free apply (element_vector_fused_kernel): int offsets into the descriptor blob,
// double offsets into an element slot
struct FusedVecInfo
{
    int n_terms, n_pre, n_rowA, n_rowC, n_contrib;
    int off_pre, off_rowA, off_termB, off_rowC, off_yrow, off_contrib; // ints, each a multiple of 4
    int blob_ints;
    int off_vec, off_t, slot_doubles; // slot: corners[8] | in / out vectors (aliased) | t rows
    int off_yfast;                    // -1, or 8 ints per output row: two branch-free contributions (every row has at most two)
    int yfast_pairs;                  // 1: every contribution is one positive and one negative address -> 4 ints per output row
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
    std::vector<DevW> ws;
    std::vector<DevTB> tbs;
    std::vector<DevRound> rounds;
    AsmLayout lay{};
    int out_ld = 0;                    // 0: n_total (rectangular single-matrix plans override)
    unsigned long long out_stride = 0; // 0: n_total^2
    void *d_blob = nullptr; // device copy of the descriptor blob (AsmLayout gives the sections)

    // vector path (flattened programs), see apply.cu
    std::vector<int32_t> vprog;
    std::vector<double> vcoef;
    int32_t *d_vprog = nullptr;
    double *d_vcoef = nullptr;
    bool vector_needs_dense = false; // an inverse mass appears in a vector program
    int mprog_off = 0, mcoef_off = 0; // matrix-mode programs (column-by-column fallback) follow the vector ones

    // fused vector path (apply.cu): every term is  coef * [E] * D * [E]  with D a mass / interior product
    bool vector_fused = false;
    FusedVecInfo fv{};
    std::vector<int32_t> fblob;
    int32_t *d_fblob = nullptr;

    // tensor-core (fp64 DMMA) vector path for q <= 8, p <= 7: one warp per element (apply.cu)
    bool warp_ok = false; // every fused term is a mass term (modes 0-2): the warp-autonomous kernel applies (apply_warp.cuh)
    bool vector_dmma = false;
    int dm_terms = 0, dm_off_terms = 0, dm_off_gtab = 0, dm_off_ltab = 0, dm_off_otab = 0, dm_off_dtab = 0, dm_blob_ints = 0;
    int dm_warp_doubles = 0, dm_off_ys = 0, dm_off_outv = 0; // per-warp staging: x | zero cell | y | term output
    int32_t *d_dmblob = nullptr;
};

// padded row length of the quadrature-indexed tables: even, and half of it odd, so that 16-byte
// shared-memory reads of consecutive rows fall into distinct bank groups
static inline __host__ __device__ int asm_ldq(int q)
{
    int l = (q + 1) & ~1;
    if (((l >> 1) & 1) == 0)
        l += 2;
    return l;
}

// lowered matrix blocks -> stages, tensor blocks, emission units, blob and shared-memory layout
struct LoweredTerm
{
    double coef;
    int L, R;
    Factor d;
    bool identity;
};
int plan_finalize_matrix(mfv2d_plan *plan, const std::vector<std::vector<LoweredTerm>> &lowered, int n_row_forms,
                         int n_col_forms, const std::vector<int> &row_offsets, const std::vector<int> &col_offsets);
int plan_lower_matrix(mfv2d_plan *plan);
int plan_lower_vector(mfv2d_plan *plan);
