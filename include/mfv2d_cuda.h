/*
 * mfv2d_cuda.h - C-ABI of libmfv2d_cuda.so: the B200 (sm_100a) implementation of mfv2d's
 * element-level hot path.
 *
 * This is the drop-in boundary.  The reference has no FFI of its own: every C
 * function is visibility("hidden") (src/common/common_defines.h:13-15) and the only exported
 * symbol is PyInit__mfv2d (src/module.c:256-259).  The entry points below are therefore what the
 * reference's CPython extension `_mfv2d` would bind if its per-element C routines were replaced by
 * batched device routines; each one cites the reference routine(s) it replaces.  INTEGRATION.md
 * shows the reference-side stub for every group.
 *
 * Conventions
 *  - plain C linkage, plain pointers and sizes, no C++/torch types.
 *  - every function returns 0 (MFV2D_SUCCESS) or one of the reference's mfv2d_result_t values
 *    (src/common/error.h:10-36, same numeric values) or an MFV2D_CUDA_* code >= 100.  A text
 *    message for the last failure is available from mfv2d_cuda_last_error().
 *  - bulk pointers carry a memory-space flag: MFV2D_HOST buffers are copied by the library
 *    (H2D before / D2H after the kernels, on the context's stream, followed by a stream
 *    synchronise); MFV2D_DEVICE buffers must come from mfv2d_cuda_malloc() of the same context and
 *    the call is asynchronous on the context's stream.
 *  - all floating point data is IEEE fp64, all indices int32/uint32, matrices dense row-major;
 *    index conventions (DoF order inside an element, quadrature point order) are the reference's
 *    (SURVEY.md Appendix A).
 *  - a context is bound to one CUDA device and one stream and is not thread-safe.
 */
#ifndef MFV2D_CUDA_H
#define MFV2D_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define MFV2D_API __attribute__((visibility("default")))
#else
#define MFV2D_API
#endif

/* ---- result codes: numeric values of mfv2d_result_t, src/common/error.h:10-36 ------------- */
enum
{
    MFV2D_SUCCESS = 0,
    MFV2D_FAILED_ALLOC = 1,
    MFV2D_BAD_ENUM = 2,
    MFV2D_DIMS_MISMATCH = 3,
    MFV2D_DOUBLE_INCIDENCE = 4,
    MFV2D_OUT_OF_INSTRUCTIONS = 5,
    MFV2D_STACK_OVERFLOW = 6,
    MFV2D_STACK_UNDERFLOW = 7,
    MFV2D_WRONG_MAT_TYPES = 8,
    MFV2D_STACK_NOT_EMPTY = 9,
    MFV2D_NOT_SQUARE = 10,
    MFV2D_NOT_IN_SURFACE = 11,
    MFV2D_ORDER_MISMATCH = 12,
    MFV2D_NOT_CONVERGED = 13,
    MFV2D_INDEX_OUT_OF_RANGE = 14,
    MFV2D_NOT_A_LEAF = 15,
    MFV2D_FAILED_CALLBACK = 16,
    MFV2D_UNSPECIFIED_ERROR = 17,
    MFV2D_BAD_ARGUMENT = 18,
    MFV2D_PYTHON_EXCEPTION = 19,
    MFV2D_PIVOT_FAILED = 20,
    MFV2D_BAD_VECTOR_PARENT = 21,
    MFV2D_MISMATCHED_TRACE = 22,
    /* codes of this library */
    MFV2D_CUDA_ERROR = 100,       /* a CUDA runtime call failed (message has the CUDA string)   */
    MFV2D_CUDA_NO_DEVICE = 101,   /* no usable sm_100 device                                   */
    MFV2D_CUDA_UNSUPPORTED = 102, /* valid request outside what the kernels implement          */
    MFV2D_CUDA_NCCL_ERROR = 103,  /* NCCL missing or a collective failed                        */
};

/* memory spaces for bulk pointers */
enum
{
    MFV2D_HOST = 0,
    MFV2D_DEVICE = 1
};

/* bytecode operation codes: matrix_op_type_t, src/evaluation/bytecode.h:12-33 */
enum
{
    MFV2D_MATOP_INVALID = 0,
    MFV2D_MATOP_IDENTITY = 1,
    MFV2D_MATOP_MASS = 2,
    MFV2D_MATOP_INCIDENCE = 3,
    MFV2D_MATOP_PUSH = 4,
    MFV2D_MATOP_SCALE = 5,
    MFV2D_MATOP_SUM = 6,
    MFV2D_MATOP_INTERPROD = 7
};

/* form orders: form_order_t, src/evaluation/forms.h:7-13 */
enum
{
    MFV2D_FORM_ORDER_0 = 1,
    MFV2D_FORM_ORDER_1 = 2,
    MFV2D_FORM_ORDER_2 = 3
};

/* incidence kinds: incidence_type_t, src/algebra/matrix.h:33-40 */
enum
{
    MFV2D_INC_10 = 0,
    MFV2D_INC_21 = 1,
    MFV2D_INC_10_T = 2,
    MFV2D_INC_21_T = 3
};

/* One decoded bytecode instruction (the union matrix_op_t of bytecode.h:89-99, flattened).
 *   IDENTITY, PUSH : no operands
 *   MASS           : a = form order (1..3), b = invert (0/1)
 *   INCIDENCE      : a = starting form order (1..2), b = transpose (0/1)
 *   SCALE          : k = factor
 *   SUM            : a = number of stack entries to pop
 *   INTERPROD      : a = starting form order (2..3), b = field index, c = transpose (0/1)     */
typedef struct
{
    int32_t op;
    int32_t a;
    int32_t b;
    int32_t c;
    double k;
} mfv2d_op_t;

typedef struct mfv2d_ctx mfv2d_ctx;
typedef struct mfv2d_basis mfv2d_basis;
typedef struct mfv2d_elements mfv2d_elements;
typedef struct mfv2d_plan mfv2d_plan;
typedef struct mfv2d_system mfv2d_system;

/* =========================================================================================
 * context, memory, errors
 * ========================================================================================= */
/* create a context on CUDA device `device` (cudaSetDevice + one non-blocking stream). */
MFV2D_API int mfv2d_cuda_init(mfv2d_ctx **ctx, int device);
MFV2D_API void mfv2d_cuda_destroy(mfv2d_ctx *ctx);
/* message of the last failure on this context (or of the last failed init when ctx == NULL). */
MFV2D_API const char *mfv2d_cuda_last_error(const mfv2d_ctx *ctx);
MFV2D_API const char *mfv2d_cuda_result_str(int code); /* mfv2d_result_str, src/common/error.c */
MFV2D_API int mfv2d_cuda_synchronize(mfv2d_ctx *ctx);
MFV2D_API int mfv2d_cuda_device_count(void);
MFV2D_API int mfv2d_cuda_malloc(mfv2d_ctx *ctx, size_t bytes, void **dptr);
MFV2D_API int mfv2d_cuda_free(mfv2d_ctx *ctx, void *dptr);
MFV2D_API int mfv2d_cuda_memset(mfv2d_ctx *ctx, void *dptr, int value, size_t bytes);
MFV2D_API int mfv2d_cuda_upload(mfv2d_ctx *ctx, void *dptr, const void *hptr, size_t bytes);
MFV2D_API int mfv2d_cuda_download(mfv2d_ctx *ctx, void *hptr, const void *dptr, size_t bytes);
/* pinned host memory for the host-buffer (e2e) path */
MFV2D_API int mfv2d_cuda_host_alloc(mfv2d_ctx *ctx, size_t bytes, void **hptr);
MFV2D_API int mfv2d_cuda_host_free(mfv2d_ctx *ctx, void *hptr);
/* CUDA-event timing on the context's stream (bench.py times the kernels with these). */
MFV2D_API int mfv2d_cuda_timer_start(mfv2d_ctx *ctx);
MFV2D_API int mfv2d_cuda_timer_stop(mfv2d_ctx *ctx, double *milliseconds);
/* number of kernel launches issued by this library on this context since creation */
MFV2D_API uint64_t mfv2d_cuda_launch_count(const mfv2d_ctx *ctx);
/* write `bytes` of zeros-then-ones through a scratch buffer larger than L2 (bench L2 flush) */
MFV2D_API int mfv2d_cuda_flush_l2(mfv2d_ctx *ctx);

/* measured fp64 peaks of the device (TFLOP/s): mma.sync.m8n8k4.f64 chains and DFMA chains; the denominator of the
 * matrix-free apply's roofline in bench.py (no counterpart in the reference) */
MFV2D_API int mfv2d_cuda_measure_fp64_peak(mfv2d_ctx *ctx, double *tflops_dmma, double *tflops_dfma);

/* =========================================================================================
 * polynomial tables (host code; O(p*q) doubles, computed once per (p, q))
 * ========================================================================================= */
/* n Gauss-Lobatto-Legendre nodes and weights.  Replaces gauss_lobatto_nodes_weights
 * (src/polynomials/gauss_lobatto.c:52-85).  Returns MFV2D_NOT_CONVERGED if a node needed more
 * than max_iter iterations (nodes are still written), as the reference warns (:249-255). */
MFV2D_API int mfv2d_gll(int n, int max_iter, double tol, double *x, double *w);
/* values / first derivatives of the Lagrange basis on roots[n] at pos[m] -> out[m*n], basis index
 * fastest.  Replaces lagrange_polynomial_values / _first_derivative
 * (src/polynomials/lagrange.c:63-107,110-171). */
MFV2D_API int mfv2d_lagrange1d(int n, const double *roots, int m, const double *pos, double *out);
MFV2D_API int mfv2d_dlagrange1d(int n, const double *roots, int m, const double *pos, double *out);
/* Legendre P_0..P_order at pos[m] -> out[(order+1)*m], order index slowest.  Replaces
 * legendre_eval_bonnet_all (src/polynomials/legendre.c:12-36). */
MFV2D_API int mfv2d_legendre(int order, int m, const double *pos, double *out);
/* Basis1D tables for element order p on the GLL rule with q points: roots[p+1], pnts[q], wgts[q],
 * node[(p+1)*q], edge[p*q].  Replaces compute_nodal_and_edge_values (src/fem_space/basis.c:5-94)
 * and IntegrationRule1D (src/fem_space/integration_rule.c:19-67). */
MFV2D_API int mfv2d_basis_tables(int p, int q, double *roots, double *pnts, double *wgts, double *node,
                                 double *edge);
/* bulk Legendre evaluation on the device: pos[m] -> out[(order+1)*m] (a4, refinement estimators) */
MFV2D_API int mfv2d_cuda_legendre(mfv2d_ctx *ctx, int order, size_t m, const double *pos, double *out,
                                  int space);

/* =========================================================================================
 * device-resident basis tables and element geometry
 * ========================================================================================= */
/* Basis2D(Basis1D(p, IntegrationRule1D(q-1)), same) resident on the device. */
MFV2D_API int mfv2d_cuda_basis_create(mfv2d_ctx *ctx, int p, int q, mfv2d_basis **basis);
MFV2D_API void mfv2d_cuda_basis_destroy(mfv2d_basis *basis);

/* A batch of n_elem elements of one (p, q) bucket; corners[n_elem][4][2], counter-clockwise from
 * bottom-left (quad_info_t, src/common/common.h:110-116).  Replaces n_elem calls of
 * fem_space_2d_create (src/fem_space/fem_space.c:16-57): the Jacobian is recomputed inside the
 * kernels from the 8 corner doubles instead of being stored. */
MFV2D_API int mfv2d_cuda_elements_create(mfv2d_ctx *ctx, const mfv2d_basis *basis, size_t n_elem,
                                         const double *corners, int space, mfv2d_elements **elems);
MFV2D_API void mfv2d_cuda_elements_destroy(mfv2d_elements *elems);
/* replace the corners of an existing batch (same n_elem); used by the host-buffer path */
MFV2D_API int mfv2d_cuda_elements_set_corners(mfv2d_elements *elems, const double *corners, int space);

/* Jacobian at the q*q points: out[n_elem][5][q][q] = j00, j01, j10, j11, det, point order
 * (eta row, xi column).  ElementFemSpace2D.jacobian / .jacobian_determinant
 * (src/fem_space/element_fem_space.c:170-219). */
MFV2D_API int mfv2d_cuda_jacobian_batch(const mfv2d_elements *elems, double *out, int space);
/* physical coordinates of the quadrature points: out[n_elem][2][q][q] (x then y); the inputs of
 * user callables (src/evaluation/integrating_fields.c:78-126). */
MFV2D_API int mfv2d_cuda_coordinates_batch(const mfv2d_elements *elems, double *out, int space);

/* =========================================================================================
 * mass matrices, interior products, projectors, field reconstruction
 * ========================================================================================= */
/* Mass matrix (or its inverse) of `form` for every element: out[n_elem][n_k][n_k].
 * Replaces compute_mass_matrix_{node,edge,surf} (src/fem_space/fem_space.c:235-411),
 * element_mass_cache_get_* (src/fem_space/element_fem_space.c:445-524) and, for inverse != 0,
 * matrix_full_invert (src/algebra/matrix.c:361-411).  status[n_elem] (may be NULL, same space as
 * out) receives MFV2D_PIVOT_FAILED for a singular block. */
MFV2D_API int mfv2d_cuda_mass_batch(const mfv2d_elements *elems, int form, int inverse, double *out,
                                    int32_t *status, int space);
/* Interior-product matrices.  start_form = MFV2D_FORM_ORDER_1: (n0 x n1), compute_mass_matrix_node_edge
 * (fem_space.c:546-635); MFV2D_FORM_ORDER_2: (n1 x n2), compute_mass_matrix_edge_surf (:753-846), WITHOUT
 * the VM's -1 (element_eval.c:362-371).  field[n_elem][q][q][2]; transpose writes the transposed
 * matrix. */
MFV2D_API int mfv2d_cuda_interprod_batch(const mfv2d_elements *elems, int start_form, int transpose,
                                         const double *field, double *out, int space);
/* Field reconstruction at the quadrature points from element DoFs of one form
 * (reconstruct_field_{0,1,2}_form, src/evaluation/integrating_fields.c:238-352).
 * dofs[n_elem][n_k] -> out[n_elem][q][q] (0-/2-forms) or out[n_elem][q][q][2] (1-forms). */
MFV2D_API int mfv2d_cuda_reconstruct_batch(const mfv2d_elements *elems, int form, const double *dofs,
                                           double *out, int space);
/* Reconstruction on an arbitrary tensor grid of reference points: out[n_elem][m_eta][m_xi] (0-/2-forms) or [...][2]
 * (1-forms) at (xi[b], eta[a]).  Replaces the per-element NumPy double loops of `reconstruct`
 * (python/mfv2d/mimetic2d.py:1172-1279) behind reconstruct_mesh_from_solution (python/mfv2d/solve_system.py:163-271).
 * 0-forms: plain interpolant; 1-forms: pushed forward with the Jacobian / det; 2-forms: / det. */
MFV2D_API int mfv2d_cuda_reconstruct_grid_batch(const mfv2d_elements *elems, int form, const double *dofs, int m_xi,
                                                const double *xi, int m_eta, const double *eta, double *out, int space);
/* Dual degrees of freedom (L2 projection integrals) of a function given by its values at the quadrature points
 * (element_dual_dofs, python/mfv2d/mimetic2d.py:1003-1123; the per-element NumPy loops behind compute_element_rhs,
 * python/mfv2d/solve_system.py:129-160).  values[n_elem][q][q] (0-/2-forms) or [n_elem][q][q][2] (1-forms)
 * -> out[n_elem][n_k]. */
MFV2D_API int mfv2d_cuda_dual_dofs_batch(const mfv2d_elements *elems, int form, const double *values, double *out,
                                         int space);
/* Two-basis masses / projectors (compute_element_projector, src/evaluation/element_system.c:480-643;
 * compute_mass_matrix_*_double, fem_space.c:847-1055).  elems_out carries the output basis and the
 * geometry, basis_in must share its integration rule.  dual != 0: out = M_mixed (n_k(out) x n_k(in));
 * dual == 0: out = M_out^-1 * M_mixed. */
MFV2D_API int mfv2d_cuda_projector_batch(const mfv2d_elements *elems_out, const mfv2d_basis *basis_in,
                                         int form, int dual, double *out, int32_t *status, int space);

/* =========================================================================================
 * incidence application (check_incidence, src/evaluation/incidence.c:445-479,593-637)
 * ========================================================================================= */
/* out = E * in (right == 0, in is (cols(E) x n) ) or out = in * E (right != 0, in is (n x rows(E))),
 * E = incidence `kind` of order p, batched over `batch` matrices.  No E is formed. */
MFV2D_API int mfv2d_cuda_incidence_apply(mfv2d_ctx *ctx, int p, int kind, int right, size_t batch, int n,
                                         const double *in, double *out, int space);

/* =========================================================================================
 * bytecode plans and element systems
 * ========================================================================================= */
/* Compile the n_forms x n_forms matrix of bytecode programs of a system for order p.
 *   form_orders[n_forms]           form order of each unknown (forms.c:21-72 gives the layout)
 *   block_present[n_forms^2]       0 where the expression is None
 *   op_offsets[n_forms^2 + 1]      range of each block's instructions in ops[]
 *   n_fields, field_form[n_fields] per INTERPROD field: -1 = values supplied by the caller
 *                                  (callable evaluated on the host), else index of the unknown
 *                                  whose DoFs it is reconstructed from (bytecode.c:131-199)
 * The plan symbolically executes the stack machine of src/evaluation/element_eval.c:400-479 once
 * (instead of once per element per call, system_template.c:38-164), reproducing its quirks
 * (SURVEY.md Q1-Q3), and lowers each block to fused device work.  A program the reference would
 * reject returns the reference's error code here. */
MFV2D_API int mfv2d_cuda_plan_create(mfv2d_ctx *ctx, const mfv2d_basis *basis, int n_forms,
                                     const int32_t *form_orders, const uint8_t *block_present,
                                     const int32_t *op_offsets, const mfv2d_op_t *ops, int n_fields,
                                     const int32_t *field_form, mfv2d_plan **plan);
MFV2D_API void mfv2d_cuda_plan_destroy(mfv2d_plan *plan);
/* total element DoF count N_e of the plan (ElementFormSpecification.total_size) */
MFV2D_API int mfv2d_cuda_plan_size(const mfv2d_plan *plan);

/* compute_element_matrix for every element (src/evaluation/element_system.c:13-212):
 * out[n_elem][N_e][N_e].  dofs[n_elem][N_e] may be NULL when no field depends on an unknown;
 * field_values[n_fields][n_elem][q][q][2] supplies the caller-evaluated fields (entries of
 * unknown-based fields are ignored and may be absent: pass NULL when n_fields == 0). */
MFV2D_API int mfv2d_cuda_element_matrix_batch(const mfv2d_plan *plan, const mfv2d_elements *elems,
                                              const double *dofs, const double *field_values, double *out,
                                              int space);
/* compute_element_vector for every element (element_system.c:245-411): y[n_elem][N_e] =
 * sum over column blocks of block(x_col), evaluated matrix-free (no element matrix is formed). */
MFV2D_API int mfv2d_cuda_element_vector_batch(const mfv2d_plan *plan, const mfv2d_elements *elems,
                                              const double *dofs, const double *field_values, double *out,
                                              int space);

/* =========================================================================================
 * dense algebra (src/algebra/matrix.c:361-660)
 * ========================================================================================= */
/* out[b] = in[b]^-1 for `batch` n x n matrices (_compute_matrix_inverse). status[batch] or NULL. */
MFV2D_API int mfv2d_cuda_inverse_batch(mfv2d_ctx *ctx, int n, size_t batch, const double *in, double *out,
                                       int32_t *status, int space);
/* out = in^-1 for ONE large definite matrix (n x n, row-major, in place allowed: out == in), Gauss-Jordan WITHOUT pivoting,
 * blocked 32 pivots per pass (a shared-memory block inverse, two panels, one register-tiled rank-32 update): the routine behind
 * the exact level of the multilevel preconditioner (mfv2d_cuda_system_set_coarse_space).  The reference inverts dense
 * matrices with pivoted LU (`_compute_matrix_inverse`, src/algebra/matrix.c:546-600) - use mfv2d_cuda_inverse_batch for
 * matrices that need pivoting.  *status (HOST int, may be NULL) and the return code are MFV2D_PIVOT_FAILED when a pivot
 * vanishes. */
MFV2D_API int mfv2d_cuda_inverse_definite(mfv2d_ctx *ctx, int n, const double *in, double *out, int32_t *status,
                                          int space);
/* X[b] = A[b]^-1 B[b], B is n x nrhs row-major (_solve_linear_system). */
MFV2D_API int mfv2d_cuda_solve_batch(mfv2d_ctx *ctx, int n, int nrhs, size_t batch, const double *a,
                                     const double *b, double *x, int32_t *status, int space);

/* out = in^-1 for ONE definite n x n matrix (row-major) WITHOUT any device allocation.
 *   NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 *   cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic as the
 *   device-wide Gauss-Jordan that inverts the coarse operator of the multilevel preconditioner (no pivoting:
 *   symmetric definite input), with its 2n doubles of scratch in a static 64 KiB device array of the library.
 *   in, out (n*n doubles, may alias) and status (one int32: 0 or MFV2D_PIVOT_FAILED) are DEVICE pointers of the
 *   caller; the work is queued on the context's stream and the call does not synchronise.
 *   n > MFV2D_NOALLOC_INVERSE_MAX_N returns MFV2D_CUDA_UNSUPPORTED.  The scratch is shared by all calls on a
 *   device: calls that overlap on the GPU race.  The first call on each device resolves the scratch address
 *   (cudaGetSymbolAddress): make it before a graph capture starts. */
#define MFV2D_NOALLOC_INVERSE_MAX_N 4096
MFV2D_API int mfv2d_cuda_inverse_definite_noalloc(mfv2d_ctx *ctx, int n, const double *in, double *out,
                                                  int32_t *status);

/* =========================================================================================
 * error-estimation helpers (python/mfv2d/refinement.py), one launch per batch
 * ========================================================================================= */
/* y[e] = A[e] x[e], n row-major (rows x cols) matrices: the projector products `projector @ dofs` of
 * refinement.py:880-886, 1158-1168 (sp.block_diag(...) @ element_solution) for every element. */
MFV2D_API int mfv2d_cuda_batched_gemv(mfv2d_ctx *ctx, size_t n, int rows, int cols, const double *mats,
                                      const double *x, double *y, int space);
/* c[e] = op(a[e]) b[e], n row-major matrices: op(a) is (m x k) - a itself, or (trans_a != 0) the transpose of a stored
 * (k x m) matrix; b is (k x l), c (m x l).  The element-wise triple products of the VMS operator,
 * `projector_c2f.T @ fine_matrix @ projector_c2f` (python/mfv2d/solve_system.py:751), for every element at once. */
MFV2D_API int mfv2d_cuda_batched_gemm(mfv2d_ctx *ctx, size_t n, int m, int k, int l, int trans_a, const double *a,
                                      const double *b, double *c, int space);
/* out[e] = [[A[e], L^T], [L, 0]] ((size + pad)^2 doubles each): the local systems of refinement.py:1001-1050 whose
 * Lagrange rows L (pad x size, COO lrow/lcol/lval[nnz], HOST arrays, the same for every element) force boundary DoFs of
 * strongly-zeroed forms to zero and pin the sum of constrained forms. */
MFV2D_API int mfv2d_cuda_border_batch(mfv2d_ctx *ctx, size_t n, int size, int pad, const double *mats, int nnz,
                                      const int32_t *lrow, const int32_t *lcol, const double *lval, double *out,
                                      int space);
/* compute_legendre_error_estimates (refinement.py:88-152, with compute_legendre_coefficients :40-85) for every element:
 * u, err = values of the solution and of the error field at the quadrature points of elems' rule, [n][q][q] or
 * [n][q][q][2] for 1-forms (the Euclidean norm is taken); order_1, order_2 = orders of the Legendre spectrum;
 * out[n][2] = { sum err^2 w det , |h-refinement cost| }.  Weights and Jacobian determinants are recomputed from the corners. */
MFV2D_API int mfv2d_cuda_legendre_estimates_batch(const mfv2d_elements *elems, int order_1, int order_2, int form,
                                                  const double *u, const double *err, double *out, int space);

/* =========================================================================================
 * element-block linear system (src/algebra/sparse_system.{h,c}, system_objects.c)
 * ========================================================================================= */
/* system_new / system_fill_out (system_objects.c:73-199,4-71).
 *   n_elem, block_offsets[n_elem+1]   dense DoF offsets (element e owns [off[e], off[e+1]) )
 *   blocks                            concatenated dense A_e (row-major, sizes N_e^2), `blocks_space`
 *                                     may be MFV2D_DEVICE (e.g. the output of element_matrix_batch)
 *   n_trace                           number of multipliers (global constraint rows)
 *   n_con, con_row/con_col/con_val    the constraint matrix N (n_trace x n_dense) as COO with GLOBAL
 *                                     dense column indices, sorted by (row, col); host arrays
 * The per-element CRS blocks, trace_offsets/trace_values (sparse_system.h:27-32) and the element
 * copies of the multipliers are derived inside.  The dense blocks are factorised (explicit inverse by
 * partially pivoted Gauss-Jordan; decompose_pivoted_lu's failure test, matrix.c:454-457, is kept). */
MFV2D_API int mfv2d_cuda_system_create(mfv2d_ctx *ctx, size_t n_elem, const int64_t *block_offsets,
                                       const double *blocks, int blocks_space, size_t n_trace, size_t n_con,
                                       const int64_t *con_row, const int64_t *con_col, const double *con_val,
                                       mfv2d_system **sys);
MFV2D_API void mfv2d_cuda_system_destroy(mfv2d_system *sys);
MFV2D_API int64_t mfv2d_cuda_system_dense_size(const mfv2d_system *sys);
MFV2D_API int64_t mfv2d_cuda_system_trace_size(const mfv2d_system *sys);
/* sparse_system_apply_diagonal / _diagonal_inverse / _trace / _trace_transpose
 * (sparse_system.c:3-40, 42-77, 79-166, 168-207).  Dense vectors are packed [n_dense]; trace
 * vectors are kept MERGED, one value per multiplier [n_trace] (every element copy of the
 * reference holds the same value; TraceVector.as_merged, system_objects.c). */
MFV2D_API int mfv2d_cuda_system_apply_diagonal(const mfv2d_system *sys, const double *x, double *y, int space);
MFV2D_API int mfv2d_cuda_system_apply_diagonal_inverse(const mfv2d_system *sys, const double *x, double *y,
                                                       int space);
MFV2D_API int mfv2d_cuda_system_apply_trace(const mfv2d_system *sys, const double *x, double *t, int space);
MFV2D_API int mfv2d_cuda_system_apply_trace_transpose(const mfv2d_system *sys, const double *t, double *y,
                                                      int space);
/* the Schur operator S = N A^-1 N^T applied through the per-element condensed blocks
 * S_e = N_e A_e^-1 N_e^T (LinearSystem.apply_full_trace_system, python/mfv2d/solving.py:99-123) */
MFV2D_API int mfv2d_cuda_system_apply_schur(const mfv2d_system *sys, const double *t_in, double *t_out,
                                            int space);

/* BLAS-1 of DenseVector / TraceVector (system_objects.c:945-1300, 1634-2060) on device vectors:
 * out = a*x + b*y ; dot(x, y).  n doubles; `space` applies to x, y, out. */
MFV2D_API int mfv2d_cuda_vec_axpby(mfv2d_ctx *ctx, size_t n, double a, const double *x, double b,
                                   const double *y, double *out, int space);
MFV2D_API int mfv2d_cuda_vec_dot(mfv2d_ctx *ctx, size_t n, const double *x, const double *y, double *result,
                                 int space);
/* max_i |x_i| (NaN if any entry is NaN): the residual test of non_linear_solve_run
 * (python/mfv2d/solve_system.py:430-447: np.max(np.abs(residual))) without bringing the vector to the host. */
MFV2D_API int mfv2d_cuda_vec_absmax(mfv2d_ctx *ctx, size_t n, const double *x, double *result, int space);
/* out[e*out_stride + j] = a*x[e*x_stride + j] + b*y[e*y_stride + j], j < len, e < n_blocks, DEVICE pointers (already offset to
 * the first entry): gathers / scatter-adds the DoFs of one form of every element at once - the time-carry terms of the
 * trapezoidal march (find_time_carry_indices, python/mfv2d/solve_system.py:604-617, solve_system_2d.py:325-397). */
MFV2D_API int mfv2d_cuda_strided_axpby(mfv2d_ctx *ctx, size_t n_blocks, size_t len, double a, const double *x,
                                       size_t x_stride, double b, const double *y, size_t y_stride, double *out,
                                       size_t out_stride);

/* solve_schur_iterative (python/mfv2d/solving.py:439-500) with cg_general (:350-436) running on
 * the device: lambda from CG on S, then x = A^-1 y - A^-1 N^T lambda.
 *   rhs[n_dense], phi[n_trace] -> x[n_dense], lambda[n_trace]; residual = sqrt(r.r) at exit,
 *   iterations = the reference's iter_cnt.  precondition: 0 = none (the reference's method), 1 = Jacobi on the
 *   trace space, 2 = Jacobi + coarse space (mfv2d_cuda_system_set_coarse_space); neither is in the reference,
 *   the solution is the same to tolerance. */
MFV2D_API int mfv2d_cuda_solve_schur_cg(mfv2d_system *sys, const double *rhs, const double *phi, int max_iter,
                                        double atol, double rtol, int precondition, double *x, double *lambda,
                                        double *residual, int *iterations, int space);
/* The same for systems whose element blocks are NOT symmetric (advection terms): BiCGStab on S, right-preconditioned with
 * diag(S) when precondition != 0.  The reference has no iterative counterpart (its CG needs symmetry; production path:
 * SuperLU, python/mfv2d/solve_system_2d.py:258-282). */
MFV2D_API int mfv2d_cuda_solve_schur_bicgstab(mfv2d_system *sys, const double *rhs, const double *phi, int max_iter,
                                              double atol, double rtol, int precondition, double *x, double *lambda,
                                              double *residual, int *iterations, int space);
/* solve_pcg_iterative (solving.py:634-684): PCG on [[A, N^T], [N, 0]] with (A^-1 (+) I). */
MFV2D_API int mfv2d_cuda_solve_pcg(mfv2d_system *sys, const double *rhs, const double *phi, int max_iter,
                                   double atol, double rtol, double *x, double *lambda, double *residual,
                                   int *iterations, int space);

/* =========================================================================================
 * multi-GPU (one process per GPU)
 * ========================================================================================= */
/* size of the opaque NCCL unique id and its creation on rank 0 */
MFV2D_API int mfv2d_cuda_comm_id_size(void);
MFV2D_API int mfv2d_cuda_comm_get_id(void *id_bytes);
/* join the communicator (ncclCommInitRank); NCCL is dlopen()ed on first use */
MFV2D_API int mfv2d_cuda_comm_init(mfv2d_ctx *ctx, int n_ranks, int rank, const void *id_bytes);
MFV2D_API int mfv2d_cuda_comm_destroy(mfv2d_ctx *ctx);
/* in-place sum all-reduce of n doubles on the context's stream (device pointer) */
MFV2D_API int mfv2d_cuda_comm_allreduce_sum(mfv2d_ctx *ctx, double *dptr, size_t n);
/* Declare which multipliers of `sys` are shared with other ranks: shared_index[n_shared] are
 * positions in this rank's trace vector, listed in the same global order on every rank that holds
 * them; slot[n_shared] their position in the packed interface buffer of n_slots doubles;
 * owned[n_trace] is 1 where this rank owns the multiplier for dot products.  After this call
 * apply_trace / apply_schur / vec dot on trace vectors include the NCCL interface sum. */
MFV2D_API int mfv2d_cuda_system_set_interface(mfv2d_system *sys, size_t n_shared, const int64_t *shared_index,
                                              const int64_t *slot, size_t n_slots, const uint8_t *owned);

/* ---- NVLink peer exchange: the interface sums and Krylov scalars written by the solver's own kernels straight into
 * the other ranks' memory (replaces the ncclAllReduce inside every iteration) ----
 * 1. every rank: peer_export (after set_interface) allocates its exchange block and returns an IPC handle of
 *    mfv2d_cuda_peer_handle_size() bytes; n_coarse reserves room for the coarse residual of the two-level preconditioner;
 * 2. the host gathers the handles rank-major by any means (as for the NCCL unique id);
 * 3. every rank: peer_connect maps the other blocks; holder_ptr[n_shared + 1] / holder_rank[] list, for each shared
 *    multiplier in the order of set_interface, the OTHER ranks that hold it (ascending).
 * NCCL is then no longer needed by this system (mfv2d_cuda_comm_init may be called with id_bytes == NULL). */
MFV2D_API int mfv2d_cuda_peer_handle_size(void);
MFV2D_API int mfv2d_cuda_system_peer_export(mfv2d_system *sys, size_t n_coarse, void *handle_bytes);
MFV2D_API int mfv2d_cuda_system_peer_connect(mfv2d_system *sys, const void *handles_all, size_t n_shared,
                                             const int64_t *holder_ptr, const int32_t *holder_rank);
/* non-zero after a peer wait timed out (a rank died): results of that call are invalid */
MFV2D_API int mfv2d_cuda_system_peer_error(mfv2d_system *sys);
/* Diagnostics (no counterpart in the reference): with MFV2D_PEER_DEBUG bit 8 set at peer_connect, the kernels of the
 * distributed CG stamp this rank's timeline; out16 receives the nanoseconds accumulated before each stamp
 * (0 product + tail, 1 gather + publish, 2 launch gap, 3 wait for the peers, 4 update, 5 chunk sums, 6 aggregate + publish,
 * 7 gap, 8 wait for the coarse residuals, 9 coarse solve) and the counters are reset. */
MFV2D_API int mfv2d_cuda_system_peer_stats(mfv2d_system *sys, uint64_t *out16);

/* Aggregate hierarchy of the additive multilevel preconditioner of solve_schur_cg (precondition = 2):
 *   M^-1 = diag(S)^-1 + sum_l P_l diag(P_l^T S P_l)^-1 P_l^T + P_L (P_L^T S P_L)^-1 P_L^T,
 * P_l = one column per aggregate of level l holding z (the near-null-space vector of S = N A^-1 N^T) on the aggregate.
 *   n_agg[n_levels]; agg[n_levels][n_trace] = aggregate (or -1) of every multiplier of this rank on each level;
 *   exact_last != 0: the last level is solved exactly (global aggregate ids, <= 4096); the other levels are rank-local
 *   and must leave out the multipliers shared with other ranks.
 * Collective on a partitioned system.  The reference has no counterpart (its only preconditioner is block-Jacobi on the
 * saddle system, python/mfv2d/solving.py:656-661). */
MFV2D_API int mfv2d_cuda_system_set_coarse_space(mfv2d_system *sys, int n_levels, const int64_t *n_agg, const int32_t *agg,
                                                 const double *z, int exact_last);

#ifdef __cplusplus
}
#endif
#endif /* MFV2D_CUDA_H */
