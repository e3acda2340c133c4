// Host-side bytecode planner.
//
// The reference re-parses the Python bytecode tuples and runs its stack machine for every element of
// every call (src/evaluation/system_template.c:38-164, element_eval.c:400-479).  Here the machine is
// executed ONCE, symbolically: the state of the machine (invalid / identity / incidence / dense, plus
// its scalar coefficient) is tracked exactly as the reference does - including the quirks listed in
// SURVEY.md (Q1 dense incidence fails, Q2 a dense*dense product drops the accumulated coefficient, Q3
// SUM only for same-kind operands) - but a dense state is kept as a sum of product chains instead of
// numbers.  The chains are then lowered to fused device work (assembly.cu, apply.cu).
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
    const int p = plan->p, q = plan->q, nf = plan->n_forms;
    plan->matrix_fused = true;
    plan->dense.clear();
    plan->terms.clear();
    plan->blocks.clear();
    plan->units.clear();
    plan->stages.clear();

    struct LTerm
    {
        double coef;
        int L, R;
        Factor d;
        bool identity;
    };
    std::vector<std::vector<LTerm>> lowered(nf * nf);
    for (int b = 0; b < nf * nf; ++b)
    {
        const BlockProgram &bp = plan->mat_blocks[b];
        if (!bp.present)
            continue;
        if (bp.identity)
        {
            LTerm lt{bp.identity_coef, -1, -1, Factor{}, true};
            lowered[b].push_back(lt);
            continue;
        }
        for (const Term &t : bp.terms)
        {
            LTerm lt{t.coef, -1, -1, Factor{}, false};
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

    // fixed shared memory: tables, geometry (5 q^2), weight field G (q^2), T scratch (q * (p+1)^2, padded)
    const int fixed = asm_fixed_doubles(p, plan->basis->tab_len, 1) + slot_layout(p, q).arena;
    const int capacity = (int)(plan->ctx->smem_optin / sizeof(double)) - fixed - 64;

    auto find_dense = [&](std::vector<DevDense> &list, int begin, const Factor &d) {
        for (int i = begin; i < (int)list.size(); ++i)
            if (list[i].dk == d.dk && list[i].inverse == d.inverse && list[i].field == d.field &&
                list[i].transpose == d.transpose)
                return i;
        return -1;
    };

    DevStage stage{};
    stage.dense_begin = 0;
    stage.block_begin = 0;
    stage.unit_begin = 0;
    int arena = 0, max_arena = 0;
    std::vector<char> started(nf * nf, 0); // block already written in an earlier stage -> accumulate

    auto close_stage = [&]() {
        stage.dense_end = (int)plan->dense.size();
        stage.block_end = (int)plan->blocks.size();
        // units of this stage
        for (int bi = stage.block_begin; bi < stage.block_end; ++bi)
        {
            const DevBlock &db = plan->blocks[bi];
            const int RG = 16; // rows per emission strip (ASM_ROWS_PER_UNIT)
            for (int c0 = 0; c0 < db.ncols; c0 += 32)
                for (int r0 = 0; r0 < db.nrows; r0 += RG)
                {
                    DevUnit u;
                    u.block = (short)bi;
                    u.c0 = (short)c0;
                    u.r0 = (short)r0;
                    u.nr = (short)((db.nrows - r0) < RG ? (db.nrows - r0) : RG);
                    plan->units.push_back(u);
                }
        }
        stage.unit_end = (int)plan->units.size();
        plan->stages.push_back(stage);
        if (arena > max_arena)
            max_arena = arena;
        stage = DevStage{};
        stage.dense_begin = (int)plan->dense.size();
        stage.block_begin = (int)plan->blocks.size();
        stage.unit_begin = (int)plan->units.size();
        arena = 0;
    };

    for (int r = 0; r < nf; ++r)
        for (int c = 0; c < nf; ++c)
        {
            const int b = r * nf + c;
            const std::vector<LTerm> &lts = lowered[b];
            size_t done = 0;
            bool first = true;
            while (first || done < lts.size())
            {
                first = false;
                DevBlock db{};
                db.row0 = plan->form_offsets[r];
                db.col0 = plan->form_offsets[c];
                db.nrows = plan->form_offsets[r + 1] - db.row0;
                db.ncols = plan->form_offsets[c + 1] - db.col0;
                db.term_begin = (int)plan->terms.size();
                db.accumulate = started[b];
                // add as many terms of the block as fit into the current stage
                size_t added = 0;
                for (; done < lts.size(); ++done, ++added)
                {
                    const LTerm &lt = lts[done];
                    int di = -1;
                    if (!lt.identity)
                    {
                        di = find_dense(plan->dense, stage.dense_begin, lt.d);
                        if (di < 0)
                        {
                            const int need = ((lt.d.rows * lt.d.cols) + 1) & ~1;
                            if (need > capacity)
                                return mfv2d_fail(plan->ctx, MFV2D_CUDA_UNSUPPORTED,
                                                  "a %d x %d element matrix does not fit into shared memory (order %d)",
                                                  lt.d.rows, lt.d.cols, p);
                            if (arena + need > capacity)
                                break; // start a new stage
                            DevDense dd{lt.d.dk, lt.d.inverse, lt.d.field, lt.d.transpose, lt.d.rows, lt.d.cols, arena};
                            arena += need;
                            plan->dense.push_back(dd);
                            di = (int)plan->dense.size() - 1;
                        }
                    }
                    DevTerm dt{lt.coef, di, lt.L, lt.R};
                    plan->terms.push_back(dt);
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
    plan->smem_doubles = fixed + max_arena;

    // blob
    std::vector<char> blob;
    auto append = [&](const void *ptr, size_t bytes) {
        size_t off = (blob.size() + 15) & ~size_t(15);
        blob.resize(off + bytes);
        memcpy(blob.data() + off, ptr, bytes);
        return (int)off;
    };
    append(plan->stages.data(), plan->stages.size() * sizeof(DevStage));
    plan->off_dense = append(plan->dense.data(), plan->dense.size() * sizeof(DevDense));
    plan->off_terms = append(plan->terms.data(), plan->terms.size() * sizeof(DevTerm));
    plan->off_blocks = append(plan->blocks.data(), plan->blocks.size() * sizeof(DevBlock));
    plan->off_units = append(plan->units.data(), plan->units.size() * sizeof(DevUnit));
    plan->blob_bytes = blob.size();
    MFV2D_TRY(mfv2d_cuda_malloc(plan->ctx, blob.size(), &plan->d_blob));
    MFV2D_TRY(mfv2d_cuda_upload(plan->ctx, plan->d_blob, blob.data(), blob.size()));
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
    delete plan;
}

int mfv2d_cuda_plan_size(const mfv2d_plan *plan) { return plan->n_total; }

} // extern "C"
