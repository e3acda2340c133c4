// Host-side bytecode planner.
//
// The reference re-parses the Python bytecode tuples and runs its stack machine for every element of
// every call (src/evaluation/system_template.c:38-164, element_eval.c:400-479).  Here the machine is
// executed ONCE, symbolically: the state of the machine (invalid / identity / incidence / dense, plus
// its scalar coefficient) is tracked exactly as the reference does - including the quirks listed in
// SURVEY.md (Q1 dense incidence fails, Q2 a dense*dense product drops the accumulated coefficient, Q3
// SUM only for same-kind operands) - but a dense state is kept as a sum of product chains instead of
// numbers.  The chains are then lowered to fused device work (assembly.cu, apply.cu).
#include <algorithm>
#include <cstdlib>

#include "plan.cuh"

namespace
{
enum StateType
{
    ST_INVALID,
    ST_IDENTITY,
    ST_INCIDENCE,
    ST_FULL
};

struct State
{
    StateType type = ST_INVALID;
    double coef = 0.0;
    int inc = -1;
    std::vector<Term> terms;
    int rows = 0, cols = 0; // dims when FULL
};

struct Sym
{
    int p;
    bool vector_mode;
    int n_seed; // rows of the seed vector (vector mode)
    int n_fields;
};

State seed_state(const Sym &s)
{
    State st;
    if (s.vector_mode)
    {
        st.type = ST_FULL;
        st.coef = 1.0;
        Term t;
        t.coef = 1.0;
        t.rows = s.n_seed;
        t.cols = 1;
        st.terms.push_back(t);
        st.rows = s.n_seed;
        st.cols = 1;
    }
    return st;
}

Factor inc_factor(int p, int kind)
{
    Factor f{};
    f.is_dense = 0;
    f.inc_kind = kind;
    f.rows = incidence_rows(p, kind);
    f.cols = incidence_cols(p, kind);
    return f;
}

// left-multiply the state by a dense factor (operation_mass / operation_interprod,
// element_eval.c:179-275, 311-397)
int left_dense(State &cur, const Factor &d, double extra)
{
    Term t;
    t.coef = 1.0;
    switch (cur.type)
    {
    case ST_INVALID:
    case ST_IDENTITY: {
        const double c = cur.type == ST_INVALID ? 1.0 : cur.coef;
        cur = State();
        cur.type = ST_FULL;
        cur.coef = c * extra;
        t.f.push_back(d);
        t.rows = d.rows;
        t.cols = d.cols;
        cur.terms.push_back(t);
        cur.rows = d.rows;
        cur.cols = d.cols;
        return MFV2D_SUCCESS;
    }
    case ST_INCIDENCE:
        return MFV2D_BAD_ENUM; // handled by apply_dense
    case ST_FULL:
        if (d.cols != cur.rows)
            return MFV2D_DIMS_MISMATCH;
        for (Term &tt : cur.terms)
        {
            tt.f.push_back(d);
            tt.rows = d.rows;
        }
        cur.rows = d.rows;
        cur.coef = 1.0 * extra; // Q2: the accumulated coefficient is dropped (element_eval.c:269,389)
        return MFV2D_SUCCESS;
    }
    return MFV2D_BAD_ENUM;
}

int apply_dense(const Sym &s, State &cur, const Factor &d, double extra)
{
    if (cur.type == ST_INCIDENCE)
    {
        const Factor e = inc_factor(s.p, cur.inc);
        if (d.cols != e.rows)
            return MFV2D_DIMS_MISMATCH;
        const double c = cur.coef;
        cur = State();
        cur.type = ST_FULL;
        cur.coef = c * extra;
        Term t;
        t.coef = 1.0;
        t.f.push_back(e);
        t.f.push_back(d);
        t.rows = d.rows;
        t.cols = e.cols;
        cur.terms.push_back(t);
        cur.rows = d.rows;
        cur.cols = e.cols;
        return MFV2D_SUCCESS;
    }
    return left_dense(cur, d, extra);
}

// matrix_add (src/algebra/matrix.c:188-276): result = right + left
int add_states(State &right, const State &left)
{
    if (left.type == ST_INVALID || right.type == ST_INVALID)
        return MFV2D_BAD_ENUM;
    if (left.type == right.type)
    {
        if (left.type == ST_IDENTITY)
        {
            right.coef += left.coef;
            return MFV2D_SUCCESS;
        }
        if (left.type == ST_INCIDENCE && left.inc == right.inc)
        {
            right.coef += left.coef;
            return MFV2D_SUCCESS;
        }
        if (left.type == ST_FULL)
        {
            if (left.rows != right.rows || left.cols != right.cols)
                return MFV2D_DIMS_MISMATCH;
            for (Term &t : right.terms)
                t.coef *= right.coef;
            for (Term t : left.terms)
            {
                t.coef *= left.coef;
                right.terms.push_back(t);
            }
            right.coef = 1.0;
            return MFV2D_SUCCESS;
        }
    }
    // a dense incidence / identity would be needed: the reference fails here (Q1, incidence.c:504,590)
    return MFV2D_FAILED_ALLOC;
}

int run_block(const Sym &s, const mfv2d_op_t *ops, int n_ops, int form_row, int n_rows, int n_cols, BlockProgram *out,
              std::string *why)
{
    State cur = seed_state(s);
    std::vector<State> stack;
    char buf[160];
    for (int i = 0; i < n_ops; ++i)
    {
        const mfv2d_op_t &op = ops[i];
        int res = MFV2D_SUCCESS;
        switch (op.op)
        {
        case MFV2D_MATOP_IDENTITY:
            if (cur.type == ST_INVALID)
            {
                cur.type = ST_IDENTITY;
                cur.coef = 1.0;
            }
            break;
        case MFV2D_MATOP_SCALE:
            if (cur.type == ST_INVALID)
            {
                cur.type = ST_IDENTITY;
                cur.coef = op.k;
            }
            else
                cur.coef *= op.k;
            break;
        case MFV2D_MATOP_PUSH:
            stack.push_back(cur);
            cur = seed_state(s);
            break;
        case MFV2D_MATOP_INCIDENCE: {
            if (op.a < MFV2D_FORM_ORDER_0 || op.a > MFV2D_FORM_ORDER_1)
            {
                res = MFV2D_BAD_ENUM;
                break;
            }
            const int kind = (op.a - 1) + (op.b ? 2 : 0);
            if (cur.type == ST_INVALID)
            {
                cur.type = ST_INCIDENCE;
                cur.coef = 1.0;
                cur.inc = kind;
            }
            else if (cur.type == ST_IDENTITY)
            {
                cur.type = ST_INCIDENCE;
                cur.inc = kind;
            }
            else if (cur.type == ST_INCIDENCE)
                res = MFV2D_FAILED_ALLOC; // Q1: incidence o incidence needs a dense incidence
            else
            {
                const Factor e = inc_factor(s.p, kind);
                if (e.cols != cur.rows)
                {
                    res = MFV2D_DIMS_MISMATCH;
                    break;
                }
                for (Term &t : cur.terms)
                {
                    t.f.push_back(e);
                    t.rows = e.rows;
                }
                cur.rows = e.rows;
            }
            break;
        }
        case MFV2D_MATOP_MASS: {
            if (op.a < MFV2D_FORM_ORDER_0 || op.a > MFV2D_FORM_ORDER_2)
            {
                res = MFV2D_BAD_ENUM;
                break;
            }
            Factor d{};
            d.is_dense = 1;
            d.dk = DK_MASS0 + (op.a - 1);
            d.inverse = op.b ? 1 : 0;
            d.field = -1;
            d.rows = d.cols = form_size(op.a, s.p);
            res = apply_dense(s, cur, d, 1.0);
            break;
        }
        case MFV2D_MATOP_INTERPROD: {
            Factor d{};
            d.is_dense = 1;
            d.field = op.b;
            d.transpose = op.c ? 1 : 0;
            double k;
            if (op.a == MFV2D_FORM_ORDER_1)
            {
                d.dk = DK_IPROD1;
                d.rows = form_size(MFV2D_FORM_ORDER_0, s.p);
                d.cols = form_size(MFV2D_FORM_ORDER_1, s.p);
                k = 1.0;
            }
            else if (op.a == MFV2D_FORM_ORDER_2)
            {
                d.dk = DK_IPROD2;
                d.rows = form_size(MFV2D_FORM_ORDER_1, s.p);
                d.cols = form_size(MFV2D_FORM_ORDER_2, s.p);
                k = -1.0; // element_eval.c:362-371
            }
            else
            {
                res = MFV2D_BAD_ENUM;
                break;
            }
            if (op.b < 0 || op.b >= s.n_fields)
            {
                res = MFV2D_INDEX_OUT_OF_RANGE;
                break;
            }
            if (d.transpose)
            {
                const int t = d.rows;
                d.rows = d.cols;
                d.cols = t;
            }
            res = apply_dense(s, cur, d, k);
            break;
        }
        case MFV2D_MATOP_SUM: {
            if (op.a < 0 || (size_t)op.a > stack.size())
            {
                res = MFV2D_STACK_UNDERFLOW;
                break;
            }
            for (int k = 0; k < op.a && res == MFV2D_SUCCESS; ++k)
            {
                State left = stack.back();
                stack.pop_back();
                res = add_states(cur, left);
            }
            break;
        }
        default:
            res = MFV2D_BAD_ENUM;
            break;
        }
        if (res != MFV2D_SUCCESS)
        {
            snprintf(buf, sizeof buf, "Failed evaluating matrix operation %d (code %d): %s", i, op.op,
                     mfv2d_cuda_result_str(res));
            *why = buf;
            return res;
        }
    }
    // matrix_as_full (element_eval.c:18-61)
    out->present = true;
    if (cur.type == ST_IDENTITY)
    {
        const int n = n_rows;
        (void)form_row;
        if (n != n_cols)
        {
            *why = "identity block is not square";
            return MFV2D_DIMS_MISMATCH;
        }
        out->identity = true;
        out->identity_coef = cur.coef;
        return MFV2D_SUCCESS;
    }
    if (cur.type == ST_FULL)
    {
        if (cur.rows != n_rows || cur.cols != (s.vector_mode ? 1 : n_cols))
        {
            snprintf(buf, sizeof buf, "block evaluates to %d x %d, expected %d x %d", cur.rows, cur.cols, n_rows,
                     s.vector_mode ? 1 : n_cols);
            *why = buf;
            return MFV2D_DIMS_MISMATCH;
        }
        for (Term &t : cur.terms)
            t.coef *= cur.coef;
        out->terms = cur.terms;
        return MFV2D_SUCCESS;
    }
    if (cur.type == ST_INCIDENCE)
    {
        *why = "block ends in a bare incidence (the reference cannot densify it, quirk Q1)";
        return MFV2D_FAILED_ALLOC;
    }
    *why = "empty program";
    return MFV2D_BAD_ENUM;
}
} // namespace

// ---- lowering of the matrix-mode programs to the fused assembly kernel --------------------------
int plan_lower_matrix(mfv2d_plan *plan)
{
    const int nf = plan->n_forms;
    plan->matrix_fused = true;
    std::vector<std::vector<LoweredTerm>> lowered(nf * nf);
    std::vector<char> present(nf * nf, 0);
    for (int b = 0; b < nf * nf; ++b)
    {
        const BlockProgram &bp = plan->mat_blocks[b];
        if (!bp.present)
            continue;
        if (bp.identity)
        {
            LoweredTerm lt{bp.identity_coef, -1, -1, Factor{}, true};
            lowered[b].push_back(lt);
            continue;
        }
        for (const Term &t : bp.terms)
        {
            LoweredTerm lt{t.coef, -1, -1, Factor{}, false};
            size_t i = 0;
            if (i < t.f.size() && !t.f[i].is_dense)
                lt.R = t.f[i++].inc_kind;
            if (i < t.f.size() && t.f[i].is_dense)
                lt.d = t.f[i++];
            else
            {
                plan->matrix_fused = false;
                return MFV2D_SUCCESS;
            }
            if (i < t.f.size() && !t.f[i].is_dense)
                lt.L = t.f[i++].inc_kind;
            if (i != t.f.size())
            {
                plan->matrix_fused = false; // longer chains go through the generic evaluator
                return MFV2D_SUCCESS;
            }
            lowered[b].push_back(lt);
        }
    }
    const int res = plan_finalize_matrix(plan, lowered, nf, nf, plan->form_offsets, plan->form_offsets);
    if (res == MFV2D_CUDA_UNSUPPORTED)
    {
        // the staged dense matrices of this (system, p, q) do not fit into shared memory (e.g. Navier-Stokes at order 8): the
        // matrix is then evaluated column by column with the matrix-free kernels (launch_matrix_by_columns) - slower, same result
        plan->matrix_fused = false;
        plan->stages.clear();
        plan->units.clear();
        return MFV2D_SUCCESS;
    }
    return res;
}

namespace
{
struct TBSpec
{
    int wkind, pkA, pkB, r0, c0, mirror;
};

// the tensor blocks of one dense matrix kind
int dense_tensor_blocks(int dk, int p, TBSpec out[3])
{
    const int nh = p * (p + 1);
    switch (dk)
    {
    case DK_MASS0:
        out[0] = {W_M0, PK_NN, PK_NN, 0, 0, 0};
        return 1;
    case DK_MASS2:
        out[0] = {W_M2, PK_EE, PK_EE, 0, 0, 0};
        return 1;
    case DK_MASS1: // hh: (node_eta i, edge_xi j); vv: (edge_eta i, node_xi j); hv and its mirror image
        out[0] = {W_HH, PK_NN, PK_EE, 0, 0, 0};
        out[1] = {W_VV, PK_EE, PK_NN, nh, nh, 0};
        out[2] = {W_HV, PK_NE, PK_EN, 0, nh, 1};
        return 3;
    case DK_IPROD1: // (n0 x n1): rows (node_eta, node_xi); cols h (node_eta, edge_xi), v (edge_eta, node_xi)
        out[0] = {W_P1H, PK_NN, PK_NE, 0, 0, 0};
        out[1] = {W_P1V, PK_NE, PK_NN, 0, nh, 0};
        return 2;
    case DK_IPROD2: // (n1 x n2): rows h (node_eta, edge_xi), v (edge_eta, node_xi); cols (edge_eta, edge_xi)
        out[0] = {W_P2H, PK_NE, PK_EE, 0, 0, 0};
        out[1] = {W_P2V, PK_EE, PK_NE, nh, 0, 0};
        return 2;
    default:
        return 0;
    }
}

inline int pk_first(int pk, int p) { return (pk == PK_NN || pk == PK_NE) ? p + 1 : p; }
inline int pk_second(int pk, int p) { return (pk == PK_NN || pk == PK_EN) ? p + 1 : p; }
inline int pk_pairs(int pk, int p)
{
    const int n = p + 1;
    switch (pk)
    {
    case PK_NN:
        return n * (n + 1) / 2;
    case PK_EE:
        return p * (p + 1) / 2;
    default:
        return n * p;
    }
}
} // namespace

int plan_finalize_matrix(mfv2d_plan *plan, const std::vector<std::vector<LoweredTerm>> &lowered, int nrf, int ncf,
                         const std::vector<int> &roff, const std::vector<int> &coff)
{
    mfv2d_ctx *ctx = plan->ctx;
    const int p = plan->p, q = plan->q, np1 = p + 1;
    // q <= 8: the two contraction stages run on the tensor cores and T stays in registers (assembly.cu, phase_tc);
    // its 16-byte fragment loads want rows of 8 doubles, the scalar path's row walks the padded length.  With two
    // tiles of 8 points (q > 8) the path is correct but slower than the scalar one (mixed Poisson p6: 20.9M against
    // 26.3M matrices/s), so it is only taken on request (MFV2D_ASM_DMMA=2; =0 turns it off altogether)
#ifdef MFV2D_EMU
    const bool use_tc = false;
#else
    const char *tc_env = getenv("MFV2D_ASM_DMMA");
    const bool use_tc = (tc_env && tc_env[0] == '2') || (q <= 8 && !(tc_env && tc_env[0] == '0'));
#endif
    const int ldq = (use_tc && q <= 8) ? 8 : asm_ldq(q);
    const int threads = 256; // ASM_THREADS of assembly.cu

    // ---- element-independent tables of the blob --------------------------------------------------
    const mfv2d_basis *bs = plan->basis;
    // incidence rows of the kinds the terms use (left: rows of E_L; right: rows of the transpose of E_R)
    std::vector<DevIncRow> inc_rows;
    int inc_row0[4] = {-1, -1, -1, -1};
    for (const std::vector<LoweredTerm> &lts : lowered)
        for (const LoweredTerm &lt : lts)
            for (int kind : {lt.L, lt.R >= 0 ? (lt.R ^ 2) : -1})
            {
                if (kind < 0 || inc_row0[kind] >= 0)
                    continue;
                inc_row0[kind] = (int)inc_rows.size();
                for (int r = bs->inc_off[kind]; r < bs->inc_off[kind + 1]; ++r)
                {
                    DevIncRow row{};
                    for (int t = 0; t < 4; ++t)
                    {
                        const int32_t v = bs->h_inc[(size_t)r * 4 + t];
                        row.e[2 * t] = v == 0 ? 0 : (v > 0 ? v - 1 : -v - 1);
                        row.e[2 * t + 1] = v == 0 ? 0 : (v > 0 ? 0x3FF00000 : (int)0xBFF00000);
                    }
                    inc_rows.push_back(row);
                }
            }
    std::vector<int32_t> pairs;
    int pair_off[4], pair_n[4];
    for (int pk = 0; pk < 4; ++pk)
    {
        pair_off[pk] = (int)pairs.size();
        const int n1 = pk_first(pk, p), n2 = pk_second(pk, p);
        int row = 0;
        for (int i = 0; i < n1; ++i)
            for (int k = 0; k < n2; ++k)
            {
                if ((pk == PK_NN || pk == PK_EE) && k < i)
                    continue;
                const int prow = pk == PK_EN ? k * p + i : row; // E_i N_k = row (k, i) of the NE table
                pairs.push_back(i | (k << 8) | (prow << 16));
                ++row;
            }
        pair_n[pk] = (int)pairs.size() - pair_off[pk];
    }
    std::vector<double> pt[3];
    {
        const double *node = bs->node.data(), *edge = bs->edge.data();
        for (int pk = 0; pk < 3; ++pk)
        {
            const double *t1 = (pk == PK_EE) ? edge : node, *t2 = (pk == PK_NN) ? node : edge;
            const int n1 = pk_first(pk, p), n2 = pk_second(pk, p);
            for (int i = 0; i < n1; ++i)
                for (int k = 0; k < n2; ++k)
                {
                    if (pk != PK_NE && k < i)
                        continue;
                    for (int a = 0; a < ldq; ++a)
                        pt[pk].push_back(a < q ? t1[i * q + a] * t2[k * q + a] : 0.0);
                }
        }
    }
    std::vector<double> tab;
    tab.insert(tab.end(), bs->node.begin(), bs->node.end());
    tab.insert(tab.end(), bs->edge.begin(), bs->edge.end());
    tab.insert(tab.end(), bs->pnts.begin(), bs->pnts.end());
    tab.insert(tab.end(), bs->wgts.begin(), bs->wgts.end());
    const size_t table_bytes = inc_rows.size() * sizeof(DevIncRow) + pairs.size() * (4 + 16 + 32) +
                               (pt[0].size() + pt[1].size() + pt[2].size() + tab.size()) * 8 + 256;

    // ---- stages: greedy packing of the dense matrices under a per-slot capacity -------------------
    struct DenseNeed
    {
        int arena, t_sum, t_max;
    };
    auto dense_need = [&](const Factor &d, int *ld_out) {
        const int ld = d.cols | 1;
        TBSpec sp[3];
        const int n = dense_tensor_blocks(d.dk, p, sp);
        DenseNeed nd{(d.rows * ld + 1) & ~1, 0, 0};
        for (int i = 0; i < n; ++i)
        {
            const int t = pk_first(sp[i].pkB, p) * pk_second(sp[i].pkB, p) * ldq;
            nd.t_sum += t;
            nd.t_max = t > nd.t_max ? t : nd.t_max;
        }
        if (d.inverse && nd.t_max < 2 * d.rows)
            nd.t_sum = nd.t_max = 2 * d.rows; // scratch of the in-place inverse
        *ld_out = ld;
        return nd;
    };
    auto find_dense = [&](std::vector<DevDense> &list, int begin, const Factor &d) {
        for (int i = begin; i < (int)list.size(); ++i)
            if (list[i].dk == d.dk && list[i].inverse == d.inverse && list[i].field == d.field &&
                list[i].transpose == d.transpose)
                return i;
        return -1;
    };
    const int wq = q * ldq; // doubles of one weight field

    // descriptor bytes are not known before the stages are; 6 KB covers every system the planner accepts
    const size_t fixed_bytes = table_bytes + 6144;
    const long long cap_max = ((long long)ctx->smem_optin - (long long)fixed_bytes - 8 * 64) / 8;
    const long long cap_pref = (((long long)228 * 1024 - 2 * 1024) / 2 - (long long)fixed_bytes - 8 * 64) / 8;

    auto build = [&](long long capacity, bool serial_rounds) -> int {
        plan->dense.clear();
        plan->terms.clear();
        plan->blocks.clear();
        plan->units.clear();
        plan->stages.clear();
        plan->ws.clear();
        plan->tbs.clear();
        plan->rounds.clear();
        DevStage stage{};
        int arena = 0, tsum = 0, tmax = 0;
        std::vector<char> started(nrf * ncf, 0);
        auto stage_slot = [&](int ar, int ts, int tm, int nw) { return (long long)nw * wq + (serial_rounds ? tm : ts) + ar; };
        auto close_stage = [&]() {
            stage.dense_end = (int)plan->dense.size();
            stage.block_end = (int)plan->blocks.size();
            stage.w_end = (int)plan->ws.size();
            plan->stages.push_back(stage);
            stage = DevStage{};
            stage.dense_begin = (int)plan->dense.size();
            stage.block_begin = (int)plan->blocks.size();
            stage.w_begin = (int)plan->ws.size();
            arena = tsum = tmax = 0;
        };
        for (int r = 0; r < nrf; ++r)
            for (int c = 0; c < ncf; ++c)
            {
                const int b = r * ncf + c;
                const std::vector<LoweredTerm> &lts = lowered[b];
                size_t done = 0;
                bool first = true;
                while (first || done < lts.size())
                {
                    first = false;
                    DevBlock db{};
                    db.row0 = roff[r];
                    db.col0 = coff[c];
                    db.nrows = roff[r + 1] - db.row0;
                    db.ncols = coff[c + 1] - db.col0;
                    db.term_begin = (int)plan->terms.size();
                    db.accumulate = started[b];
                    size_t added = 0;
                    for (; done < lts.size(); ++done, ++added)
                    {
                        const LoweredTerm &lt = lts[done];
                        int di = -1;
                        if (!lt.identity)
                        {
                            di = find_dense(plan->dense, stage.dense_begin, lt.d);
                            if (di < 0)
                            {
                                int ld = 0;
                                const DenseNeed nd = dense_need(lt.d, &ld);
                                // weights this matrix adds to the stage
                                TBSpec sp[3];
                                const int ntb = dense_tensor_blocks(lt.d.dk, p, sp);
                                int new_w = 0;
                                for (int i = 0; i < ntb; ++i)
                                {
                                    bool have = false;
                                    for (int w = stage.w_begin; w < (int)plan->ws.size(); ++w)
                                        have |= plan->ws[w].kind == sp[i].wkind && plan->ws[w].field == lt.d.field;
                                    new_w += have ? 0 : 1;
                                }
                                const int nw_now = (int)plan->ws.size() - stage.w_begin;
                                const int tm = nd.t_max > tmax ? nd.t_max : tmax;
                                if (stage_slot(nd.arena, nd.t_sum, nd.t_max, new_w) > capacity)
                                    return MFV2D_CUDA_UNSUPPORTED;
                                if (stage_slot(arena + nd.arena, tsum + nd.t_sum, tm, nw_now + new_w) > capacity)
                                    break; // start a new stage
                                DevDense dd{lt.d.dk, lt.d.inverse, lt.d.field, lt.d.transpose, lt.d.rows, lt.d.cols, arena, ld};
                                for (int i = 0; i < ntb; ++i)
                                {
                                    int w = -1;
                                    for (int k = stage.w_begin; k < (int)plan->ws.size(); ++k)
                                        if (plan->ws[k].kind == sp[i].wkind && plan->ws[k].field == lt.d.field)
                                            w = k - stage.w_begin;
                                    if (w < 0)
                                    {
                                        plan->ws.push_back(DevW{sp[i].wkind, lt.d.field});
                                        w = (int)plan->ws.size() - 1 - stage.w_begin;
                                    }
                                    DevTB tb{};
                                    tb.w = w;
                                    tb.pkA = sp[i].pkA;
                                    tb.pkB = sp[i].pkB;
                                    tb.nA = pk_first(tb.pkA, p);
                                    tb.nA2 = pk_second(tb.pkA, p);
                                    tb.nB = pk_first(tb.pkB, p);
                                    tb.nB2 = pk_second(tb.pkB, p);
                                    tb.dst_off = arena;
                                    tb.ld = ld;
                                    tb.r0 = sp[i].r0;
                                    tb.c0 = sp[i].c0;
                                    tb.trans = lt.d.transpose;
                                    tb.mirror = sp[i].mirror;
                                    tb.round = (int)plan->stages.size(); // stage index for now
                                    plan->tbs.push_back(tb);
                                }
                                arena += nd.arena;
                                tsum += nd.t_sum;
                                tmax = tm;
                                plan->dense.push_back(dd);
                                di = (int)plan->dense.size() - 1;
                            }
                        }
                        plan->terms.push_back(DevTerm{lt.coef, di, lt.L, lt.R});
                    }
                    db.term_end = (int)plan->terms.size();
                    if (added > 0 || lts.empty())
                    {
                        plan->blocks.push_back(db);
                        started[b] = 1;
                    }
                    if (done < lts.size())
                        close_stage();
                    if (lts.empty())
                        break;
                }
            }
        close_stage();
        return MFV2D_SUCCESS;
    };

    bool serial = false;
    long long cap_used = cap_pref;
    int res = cap_pref > 0 ? build(cap_pref, false) : MFV2D_CUDA_UNSUPPORTED;
    if (res != MFV2D_SUCCESS)
    {
        cap_used = cap_max;
        res = build(cap_max, false);
    }
    if (res != MFV2D_SUCCESS)
    {
        serial = true;
        res = build(cap_max, true);
    }
    if (res != MFV2D_SUCCESS)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED,
                          "an element matrix of order %d with %d quadrature points does not fit into shared memory", p, q);

    // ---- emission units -----------------------------------------------------------------------------
    for (DevStage &sg : plan->stages)
    {
        sg.unit_begin = (int)plan->units.size();
        for (int bi = sg.block_begin; bi < sg.block_end; ++bi)
        {
            const DevBlock &db = plan->blocks[bi];
            // tile width: 32 lanes, or 16 (one 128-byte row segment per half warp) if that wastes > 10 % fewer lanes
            int cl = 5;
            double best = (double)((db.ncols + 31) / 32 * 32) / db.ncols;
            for (int c = 4; c >= 4; --c)
            {
                const int w = 1 << c;
                const double waste = (double)((db.ncols + w - 1) / w * w) / db.ncols;
                if (waste < best - 0.10)
                {
                    best = waste;
                    cl = c;
                }
            }
            const int w = 1 << cl, rows_per_store = 32 >> cl;
            const int rg = 32 * rows_per_store; // at most 32 warp stores per unit
            for (int c0 = 0; c0 < db.ncols; c0 += w)
                for (int r0 = 0; r0 < db.nrows; r0 += rg)
                {
                    DevUnit u{};
                    u.block = (short)bi;
                    u.c0 = (short)c0;
                    u.r0 = (short)r0;
                    u.nr = (short)((db.nrows - r0) < rg ? (db.nrows - r0) : rg);
                    u.cl = (short)cl;
                    plan->units.push_back(u);
                }
        }
        // heaviest units first: the warps of a CTA fetch units dynamically (longest-processing-time order)
        auto unit_cost = [&](const DevUnit &u) {
            const DevBlock &db = plan->blocks[u.block];
            int gathers = 1;
            for (int t = db.term_begin; t < db.term_end; ++t)
            {
                const DevTerm &tm = plan->terms[t];
                const int nl = tm.L < 0 ? 1 : ((tm.L == MFV2D_INC_10 || tm.L == MFV2D_INC_21_T) ? 2 : 4);
                const int rk = tm.R < 0 ? -1 : (tm.R ^ 2);
                const int nr = rk < 0 ? 1 : ((rk == MFV2D_INC_10 || rk == MFV2D_INC_21_T) ? 2 : 4);
                gathers += 2 * nl * nr + (nl > 1 ? 6 : 0);
            }
            const int iters = (u.nr + (32 >> u.cl) - 1) / (32 >> u.cl);
            return 40 + iters * (6 + gathers);
        };
        std::stable_sort(plan->units.begin() + sg.unit_begin, plan->units.end(),
                         [&](const DevUnit &a, const DevUnit &b) { return unit_cost(a) > unit_cost(b); });
        sg.unit_end = (int)plan->units.size();
    }

    // ---- slot layout and elements per pass ---------------------------------------------------------
    AsmLayout &L = plan->lay;
    L = AsmLayout{};
    L.ldq = ldq;
    int slot_w = 0, slot_t = 0, slot_a = 0;
    {
        size_t tb_i = 0;
        for (size_t si = 0; si < plan->stages.size(); ++si)
        {
            const DevStage &sg = plan->stages[si];
            const int nw = sg.w_end - sg.w_begin;
            slot_w = nw * wq > slot_w ? nw * wq : slot_w;
            int ar = 0, ts = 0, tm = 0;
            for (int d = sg.dense_begin; d < sg.dense_end; ++d)
            {
                const DevDense &dd = plan->dense[d];
                const int end = dd.smem_off + ((dd.rows * dd.ld + 1) & ~1);
                ar = end > ar ? end : ar;
                if (dd.inverse)
                    tm = 2 * dd.rows > tm ? 2 * dd.rows : tm;
            }
            for (; tb_i < plan->tbs.size() && plan->tbs[tb_i].round == (int)si; ++tb_i)
            {
                const int t = use_tc ? 0 : plan->tbs[tb_i].nB * plan->tbs[tb_i].nB2 * ldq;
                ts += t;
                tm = t > tm ? t : tm;
            }
            const int tt = serial ? tm : (ts > tm ? ts : tm);
            slot_t = tt > slot_t ? tt : slot_t;
            slot_a = ar > slot_a ? ar : slot_a;
        }
    }
    L.slot_w = 0;
    L.slot_t = (slot_w + 1) & ~1;
    L.slot_arena = L.slot_t + ((slot_t + 1) & ~1);
    int slot = L.slot_arena + slot_a + 2;
    slot = ((slot + 15) & ~15) + 4; // slot stride = 4 (mod 16) doubles: the slots of a pass sit in distinct banks
    L.slot_stride = slot;
    // ---- rounds: groups of tensor blocks whose T arrays live at the same time -------------------------
    {
        size_t tb_i = 0;
        for (size_t si = 0; si < plan->stages.size(); ++si)
        {
            DevStage &sg = plan->stages[si];
            sg.round_begin = (int)plan->rounds.size();
            const size_t tb_first = tb_i;
            while (tb_i < plan->tbs.size() && plan->tbs[tb_i].round == (int)si)
                ++tb_i;
            int t_off = 0;
            for (size_t a = tb_first; a < tb_i;)
            {
                const size_t b = serial ? a + 1 : (a + 16 < tb_i ? a + 16 : tb_i); // 4 bits name a block inside a round
                if (serial)
                    t_off = 0;
                for (size_t t = a; t < b; ++t)
                {
                    DevTB &tb = plan->tbs[t];
                    tb.round = (int)plan->rounds.size();
                    tb.t_off = t_off;
                    t_off += tb.nB * tb.nB2 * ldq;
                }
                plan->rounds.push_back(DevRound{(int)a, (int)b, 0, 0, 0, 0});
                a = b;
            }
            sg.round_end = (int)plan->rounds.size();
        }
    }

    // ---- blob, part 1: the element-independent tables (their offsets go into the pair tables) -----------
    std::vector<char> blob;
    auto append = [&](const void *ptr, size_t bytes) {
        size_t off = (blob.size() + 15) & ~size_t(15);
        blob.resize(off + bytes);
        if (bytes)
            memcpy(blob.data() + off, ptr, bytes);
        return (int)off;
    };
    for (int k = 0; k < 3; ++k)
        L.off_pt[k] = append(pt[k].data(), pt[k].size() * sizeof(double));
    L.off_tab = append(tab.data(), tab.size() * sizeof(double));
    L.off_inc = append(inc_rows.data(), inc_rows.size() * sizeof(DevIncRow));
    for (int k = 0; k < 4; ++k)
        L.inc_row0[k] = inc_row0[k];

    // per tensor block: what one (j,l) pair does in stage 1 and one (i,k) pair in stage 2
    std::vector<DevS1Pair> s1tab;
    std::vector<DevS2Pair> s2tab;
    for (DevTB &tb : plan->tbs)
    {
        tb.s1_tab = (int)s1tab.size();
        tb.npA = pair_n[tb.pkA];
        tb.npB = pair_n[tb.pkB];
        for (int n = 0; n < pair_n[tb.pkB]; ++n)
        {
            const int pr = pairs[pair_off[tb.pkB] + n];
            const int j = pr & 255, l = (pr >> 8) & 255, row = pr >> 16;
            DevS1Pair e{};
            e.pt = L.off_pt[tb.pkB == PK_EN ? PK_NE : tb.pkB] / 8 + row * ldq;
            e.t1 = (j * tb.nB2 + l) * ldq;
            e.t2 = (tb.pkB <= PK_EE && j != l) ? (l * tb.nB2 + j) * ldq : -1;
            e.jl = j | (l << 8);
            s1tab.push_back(e);
        }
        tb.s2_tab = (int)s2tab.size();
        for (int n = 0; n < pair_n[tb.pkA]; ++n)
        {
            const int pr = pairs[pair_off[tb.pkA] + n];
            const int i = pr & 255, k = (pr >> 8) & 255, row = pr >> 16;
            const bool direct = !tb.trans, flipped = tb.trans || tb.mirror, image = tb.pkA <= PK_EE && i != k;
            DevS2Pair e{};
            e.pt = L.off_pt[tb.pkA == PK_EN ? PK_NE : tb.pkA] / 8 + row * ldq;
            e.d0 = direct ? i * tb.nB * tb.ld + k * tb.nB2 : -1;
            e.f0 = flipped ? k * tb.nB2 * tb.ld + i * tb.nB : -1;
            e.d1 = (direct && image) ? k * tb.nB * tb.ld + i * tb.nB2 : -1;
            e.f1 = (flipped && image) ? i * tb.nB2 * tb.ld + k * tb.nB : -1;
            s2tab.push_back(e);
        }
    }

    // work items of the tensor-core contraction: one warp takes 8 (j,l) pairs of one tensor block
    std::vector<int32_t> tctab;
    for (DevStage &sg : plan->stages)
    {
        sg.tc_begin = (int)tctab.size();
        if (sg.round_end > sg.round_begin)
            for (int t = plan->rounds[sg.round_begin].tb_begin; t < plan->rounds[sg.round_end - 1].tb_end; ++t)
                for (int nt = 0; 8 * nt < plan->tbs[t].npB; ++nt)
                    for (int mt = 0; 8 * mt < plan->tbs[t].npA; ++mt)
                        tctab.push_back(t | (nt << 16) | (mt << 24));
        sg.tc_count = (int)tctab.size() - sg.tc_begin;
    }
    L.use_tc = use_tc;

    // ---- chunking of the pair lists, elements per pass and the packed work items -------------------------
    // cost model of a round: passes of `threads` work items, each item costs a set-up plus its pairs
    auto choose = [&](int round, bool stage2, int epb, int min_chunk, int *count) {
        const DevRound &rd = plan->rounds[round];
        long long best_cost = -1;
        int best = min_chunk;
        for (int chunk = min_chunk; chunk <= 31; ++chunk)
        {
            long long items = 0;
            for (int t = rd.tb_begin; t < rd.tb_end; ++t)
            {
                const DevTB &tb = plan->tbs[t];
                const int np = pk_pairs(stage2 ? tb.pkA : tb.pkB, p);
                const int per = stage2 ? (tb.nB * tb.nB2 + 1) / 2 : q;
                items += (long long)((np + chunk - 1) / chunk) * per;
            }
            const long long passes = (items * epb + threads - 1) / threads;
            const long long cost = passes * (30 + (stage2 ? 28 : 16) * chunk);
            if (best_cost < 0 || cost < best_cost)
            {
                best_cost = cost;
                best = chunk;
                *count = (int)items;
            }
        }
        return best;
    };
    std::vector<int32_t> xtab;
    bool fits = false;
    // first choice: two CTAs per SM (if the stages were packed for that), any chunking; then one CTA per SM
    const long long cta_budget[2] = {cap_used == cap_pref ? ((long long)228 * 1024 - 2 * 1024) / 2 : (long long)ctx->smem_optin,
                                     (long long)ctx->smem_optin};
    for (int attempt = 0; attempt < 10 && !fits; ++attempt)
    {
        const int min_chunk = (const int[]){1, 4, 8, 16, 31}[attempt % 5];
        const long long per_cta = cta_budget[attempt / 5];
        // upper bound of the work-item table for chunks of at least min_chunk pairs
        size_t x_bound = 0;
        for (const DevTB &tb : plan->tbs)
            x_bound += (size_t)((pk_pairs(tb.pkB, p) + min_chunk - 1) / min_chunk) * q +
                       (size_t)((pk_pairs(tb.pkA, p) + min_chunk - 1) / min_chunk) * ((tb.nB * tb.nB2 + 1) / 2);
        size_t fixed_exact = (blob.size() + 15) & ~size_t(15);
        auto sec = [&](size_t bytes) { fixed_exact = ((fixed_exact + 15) & ~size_t(15)) + bytes; };
        sec(plan->stages.size() * sizeof(DevStage));
        sec(plan->dense.size() * sizeof(DevDense));
        sec(plan->terms.size() * sizeof(DevTerm));
        sec(plan->blocks.size() * sizeof(DevBlock));
        sec(plan->units.size() * sizeof(DevUnit));
        sec(plan->ws.size() * sizeof(DevW));
        sec(plan->tbs.size() * sizeof(DevTB));
        sec(plan->rounds.size() * sizeof(DevRound));
        sec(s1tab.size() * sizeof(DevS1Pair));
        sec(s2tab.size() * sizeof(DevS2Pair));
        sec(x_bound * sizeof(int32_t));
        sec(tctab.size() * sizeof(int32_t));
        fixed_exact += 8 * 64 + 256; // corners of up to 8 slots, alignment slack
        long long epb = (per_cta - (long long)fixed_exact) / 8 / slot;
        if (epb < 1)
            continue;
        if (epb > 8)
            epb = 8;
        L.epb = (int)epb;
        xtab.clear();
        for (size_t r = 0; r < plan->rounds.size(); ++r)
        {
            DevRound &rd = plan->rounds[r];
            int dummy = 0;
            const int chunk1 = choose((int)r, false, L.epb, min_chunk, &dummy);
            const int chunk2 = choose((int)r, true, L.epb, min_chunk, &dummy);
            rd.x1_begin = (int)xtab.size();
            for (int t = rd.tb_begin; t < rd.tb_end; ++t)
            {
                DevTB &tb = plan->tbs[t];
                const int np = pk_pairs(tb.pkB, p);
                tb.n_sub1 = (np + chunk1 - 1) / chunk1;
                for (int sub = 0; sub < tb.n_sub1; ++sub)
                {
                    const int pb = sub * np / tb.n_sub1, pe = (sub + 1) * np / tb.n_sub1;
                    for (int a = 0; a < q; ++a)
                        xtab.push_back((t - rd.tb_begin) | (a << 4) | (pb << 9) | ((pe - pb) << 17));
                }
            }
            rd.x1_count = (int)xtab.size() - rd.x1_begin;
            rd.x2_begin = (int)xtab.size();
            for (int t = rd.tb_begin; t < rd.tb_end; ++t)
            {
                DevTB &tb = plan->tbs[t];
                const int np = pk_pairs(tb.pkA, p);
                tb.n_sub2 = (np + chunk2 - 1) / chunk2;
                const int nJL = tb.nB * tb.nB2;
                for (int sub = 0; sub < tb.n_sub2; ++sub)
                {
                    const int pb = sub * np / tb.n_sub2, pe = (sub + 1) * np / tb.n_sub2;
                    for (int jl = 0; jl < (nJL + 1) / 2; ++jl) // columns jl and jl + ceil(nJL / 2)
                        xtab.push_back((t - rd.tb_begin) | ((jl / tb.nB2) << 4) | ((jl % tb.nB2) << 9) | (pb << 14) |
                                       ((pe - pb) << 22));
                }
            }
            rd.x2_count = (int)xtab.size() - rd.x2_begin;
        }
        fits = true;
        break;
    }
    if (!fits)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED,
                          "an element matrix of order %d with %d quadrature points does not fit into shared memory", p, q);
    const int epb = L.epb;

    // ---- blob, part 2: descriptors -----------------------------------------------------------------------
    L.n_stages = (int)plan->stages.size();
    L.off_stages = append(plan->stages.data(), plan->stages.size() * sizeof(DevStage));
    L.off_dense = append(plan->dense.data(), plan->dense.size() * sizeof(DevDense));
    L.off_terms = append(plan->terms.data(), plan->terms.size() * sizeof(DevTerm));
    L.off_blocks = append(plan->blocks.data(), plan->blocks.size() * sizeof(DevBlock));
    L.off_units = append(plan->units.data(), plan->units.size() * sizeof(DevUnit));
    L.off_ws = append(plan->ws.data(), plan->ws.size() * sizeof(DevW));
    L.off_tbs = append(plan->tbs.data(), plan->tbs.size() * sizeof(DevTB));
    L.off_rounds = append(plan->rounds.data(), plan->rounds.size() * sizeof(DevRound));
    L.off_s1 = append(s1tab.data(), s1tab.size() * sizeof(DevS1Pair));
    L.off_s2 = append(s2tab.data(), s2tab.size() * sizeof(DevS2Pair));
    L.off_x = append(xtab.data(), xtab.size() * sizeof(int32_t));
    L.off_tc = append(tctab.data(), tctab.size() * sizeof(int32_t));
    blob.resize((blob.size() + 15) & ~size_t(15));
    L.blob_bytes = (int)blob.size();
    L.off_cs = L.blob_bytes / 8;
    L.off_slots = (L.off_cs + 8 * epb + 15) & ~15;
    L.smem_bytes = (L.off_slots + epb * L.slot_stride + 8) * 8;
    if ((size_t)L.smem_bytes > ctx->smem_optin)
        return mfv2d_fail(ctx, MFV2D_CUDA_UNSUPPORTED, "assembly kernel needs %d bytes of shared memory (order %d, %d points)",
                          L.smem_bytes, p, q);
    if (getenv("MFV2D_PLAN_DEBUG"))
    {
        fprintf(stderr, "[mfv2d plan] p %d q %d ldq %d: %zu stages, %zu dense, %zu tensor blocks, %zu units; epb %d, slot %d doubles "
                        "(W %d, T %d, arena %d), blob %d B, smem %d B%s\n",
                p, q, ldq, plan->stages.size(), plan->dense.size(), plan->tbs.size(), plan->units.size(), epb, L.slot_stride,
                L.slot_t, L.slot_arena - L.slot_t, slot_a, L.blob_bytes, L.smem_bytes, serial ? ", serial rounds" : "");
        for (size_t r = 0; r < plan->rounds.size(); ++r)
        {
            const DevRound &rd = plan->rounds[r];
            fprintf(stderr, "  round %zu: S1 %d items/element, S2 %d items/element;", r, rd.x1_count, rd.x2_count);
            for (int t = rd.tb_begin; t < rd.tb_end; ++t)
                fprintf(stderr, " tb%d(%dx%d pairs, chunks %d/%d)", t, pk_pairs(plan->tbs[t].pkB, p), pk_pairs(plan->tbs[t].pkA, p),
                        plan->tbs[t].n_sub1, plan->tbs[t].n_sub2);
            fprintf(stderr, "\n");
        }
    }
    if (plan->d_blob)
        mfv2d_cuda_free(ctx, plan->d_blob);
    plan->d_blob = nullptr;
    MFV2D_TRY(mfv2d_cuda_malloc(ctx, blob.size(), &plan->d_blob));
    MFV2D_TRY(mfv2d_cuda_upload(ctx, plan->d_blob, blob.data(), blob.size()));
    return MFV2D_SUCCESS;
}

extern "C" {

int mfv2d_cuda_plan_create(mfv2d_ctx *ctx, const mfv2d_basis *basis, int n_forms, const int32_t *form_orders,
                           const uint8_t *block_present, const int32_t *op_offsets, const mfv2d_op_t *ops, int n_fields,
                           const int32_t *field_form, mfv2d_plan **out)
{
    *out = nullptr;
    if (!basis || basis->ctx != ctx)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "plan_create: basis does not belong to this context");
    if (n_forms < 1 || n_forms > 16)
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "plan_create: n_forms must be in [1, 16]");
    if (n_fields < 0 || n_fields > 16) // INTEGRATING_FIELDS_MAX_COUNT, integrating_fields.h:9-12
        return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "plan_create: at most 16 fields");
    mfv2d_plan *plan = new mfv2d_plan();
    plan->ctx = ctx;
    plan->basis = basis;
    plan->p = basis->p;
    plan->q = basis->q;
    plan->n_forms = n_forms;
    plan->form_offsets.push_back(0);
    for (int i = 0; i < n_forms; ++i)
    {
        const int n = form_size(form_orders[i], plan->p);
        if (n < 0)
        {
            delete plan;
            return mfv2d_fail(ctx, MFV2D_BAD_ENUM, "plan_create: form order %d of unknown %d is invalid", form_orders[i], i);
        }
        plan->form_orders.push_back(form_orders[i]);
        plan->form_offsets.push_back(plan->form_offsets.back() + n);
    }
    plan->n_total = plan->form_offsets.back();
    plan->n_fields = n_fields;
    for (int i = 0; i < n_fields; ++i)
    {
        if (field_form[i] >= n_forms)
        {
            delete plan;
            return mfv2d_fail(ctx, MFV2D_INDEX_OUT_OF_RANGE, "plan_create: field %d refers to unknown %d", i, field_form[i]);
        }
        if (field_form[i] >= 0 && form_orders[field_form[i]] != MFV2D_FORM_ORDER_1)
        {
            // bytecode fields are always two-component (bytecode.c:131-199)
            delete plan;
            return mfv2d_fail(ctx, MFV2D_BAD_ARGUMENT, "plan_create: field %d must come from a 1-form unknown", i);
        }
        plan->field_form.push_back(field_form[i]);
        if (field_form[i] >= 0)
            plan->needs_dofs = true;
    }
    plan->mat_blocks.resize(n_forms * n_forms);
    plan->vec_blocks.resize(n_forms * n_forms);
    for (int r = 0; r < n_forms; ++r)
        for (int c = 0; c < n_forms; ++c)
        {
            const int b = r * n_forms + c;
            if (!block_present[b])
                continue;
            const int n_rows = plan->form_offsets[r + 1] - plan->form_offsets[r];
            const int n_cols = plan->form_offsets[c + 1] - plan->form_offsets[c];
            const mfv2d_op_t *bops = ops + op_offsets[b];
            const int n_ops = op_offsets[b + 1] - op_offsets[b];
            std::string why;
            Sym sm{plan->p, false, n_cols, n_fields};
            int res = run_block(sm, bops, n_ops, form_orders[r], n_rows, n_cols, &plan->mat_blocks[b], &why);
            if (res == MFV2D_SUCCESS)
            {
                Sym sv{plan->p, true, n_cols, n_fields};
                res = run_block(sv, bops, n_ops, form_orders[r], n_rows, n_cols, &plan->vec_blocks[b], &why);
            }
            if (res != MFV2D_SUCCESS)
            {
                delete plan;
                return mfv2d_fail(ctx, res, "block (%d, %d): %s", r, c, why.c_str());
            }
        }
    int res = plan_lower_matrix(plan);
    if (res == MFV2D_SUCCESS)
        res = plan_lower_vector(plan);
    if (res != MFV2D_SUCCESS)
    {
        mfv2d_cuda_plan_destroy(plan);
        return res;
    }
    *out = plan;
    return MFV2D_SUCCESS;
}

void mfv2d_cuda_plan_destroy(mfv2d_plan *plan)
{
    if (!plan)
        return;
    if (plan->d_blob)
        mfv2d_cuda_free(plan->ctx, plan->d_blob);
    if (plan->d_vprog)
        mfv2d_cuda_free(plan->ctx, plan->d_vprog);
    if (plan->d_vcoef)
        mfv2d_cuda_free(plan->ctx, plan->d_vcoef);
    if (plan->d_fblob)
        mfv2d_cuda_free(plan->ctx, plan->d_fblob);
    if (plan->d_dmblob)
        mfv2d_cuda_free(plan->ctx, plan->d_dmblob);
    delete plan;
}

int mfv2d_cuda_plan_size(const mfv2d_plan *plan) { return plan->n_total; }

} // extern "C"
