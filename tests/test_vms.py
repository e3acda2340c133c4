"""The batched fine-scale Green's operator (mfv2d_b200/vms.py) against the reference's own ``SuyashGreenOperator``
(tests/golden/vms.npz, written by tests/golden/make_golden_vms.py from python/mfv2d/solve_system.py:653-961).
`-m "not gpu"`: through the host emulation of the kernels; `-m gpu`: through libmfv2d_cuda.so on the B200.

The reference factorises its two saddle systems with SuperLU; here they are solved by CG on the Schur complement to a
relative residual of 1e-13, so the comparison is to 1e-8 of the largest entry (the fixed point amplifies solver error by
its contraction constant), not bit-exact."""
import json
import os
import sys

import numpy as np
import pytest
from conftest import golden

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

import fields as F

TOL = 1e-8


def rel_max(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(np.asarray(b))))


def decode(codes):
    out = []
    for rowc in codes:
        r = []
        for blk in rowc:
            if blk is None:
                r.append(None)
                continue
            ops = []
            for item in blk:
                item = list(item)
                if item[0] == 7 and isinstance(item[2], str) and item[2].startswith("fn:"):
                    item[2] = F.FIELDS[item[2][3:]]
                ops.append(tuple(item))
            r.append(tuple(ops))
        out.append(tuple(r))
    return tuple(out)


def run_checks(M):
    from mfv2d_b200 import solving as SV
    from mfv2d_b200 import vms as V

    g = golden("vms.npz")
    raw = json.loads(str(g["codes"]))
    p, k, inc = int(g["order"]), int(g["k"]), int(g["increase"])
    specs = M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))
    rule = M.IntegrationRule1D(p + k)
    b1 = M.Basis1D(p, rule)
    batch = M.ElementBatch(M.Basis2D(b1, b1), np.ascontiguousarray(g["corners"]))
    n, nc, nf = batch.n, specs.total_size(p, p), specs.total_size(p + inc, p + inc)
    fine = tuple(g[f"fine_{t}"] for t in ("row", "col", "val")) + (int(g["fine_rows"]),)
    coarse = tuple(g[f"coarse_{t}"] for t in ("row", "col", "val")) + (int(g["coarse_rows"]),)
    conv = SV.ConvergenceSettings(60, 1e-13, 1e-12)

    def make(relaxation):
        return V.SuyashGreenOperator(specs, decode(raw["symmetric"]), decode(raw["advection"]), None, batch, inc, g["fine_forcing"], fine, coarse,
                                     conv, relaxation, krylov=SV.ConvergenceSettings(5000, 1e300, 1e-14))

    op = make(1.0)
    assert (op.n_coarse, op.n_fine) == (nc, nf) and op.fine_batch.p == p + inc and op.fine_batch.q == batch.q
    # projectors of element 0 and the batched matrix product behind the coarse blocks
    assert rel_max(op.projector_c2f.get()[0], g["projector_c2f_0"]) < 1e-12
    assert rel_max(op.projector_f2c.get()[0], g["projector_f2c_0"]) < 1e-12
    rng = np.random.default_rng(1)
    a, b = rng.standard_normal((5, 7, 9)), rng.standard_normal((5, 9, 6))
    assert np.allclose(V.batched_matmul(batch.ctx, a, b), a @ b, rtol=1e-13, atol=1e-13)
    at = np.ascontiguousarray(a.transpose(0, 2, 1))
    assert np.allclose(V.batched_matmul(batch.ctx, at, b, trans_a=True), a @ b, rtol=1e-13, atol=1e-13)
    # the methods of the operator
    cs = g["coarse_solution"]
    assert rel_max(op.fine_scale_greens_function(g["x_fine"]), g["green"]) < TOL
    un = op.compute_unresolved_contributions(cs[: n * nc], None)
    assert rel_max(un, g["unresolved"]) < TOL, rel_max(un, g["unresolved"])
    assert rel_max(op.compute_unresolved_contributions(cs[: n * nc], g["guess"]), g["unresolved_guess"]) < TOL
    assert rel_max(op.recover_unresolved(cs, g["unresolved"]), g["recovered"]) < 1e-7
    assert rel_max(op.fine_results_to_coarse_dofs(g["x_fine"], dual=True), g["to_coarse_dual"]) < 1e-12
    assert rel_max(op.fine_results_to_coarse_dofs(g["x_fine"], dual=False), g["to_coarse_primal"]) < 1e-12
    assert rel_max(make(0.7).compute_unresolved_contributions(cs[: n * nc], None), g["unresolved_relaxed"]) < TOL
    op.update_nonlinear_advection(cs)  # linear advection: a no-op, like the reference's
    assert op.fine_advection is op.fine_linear_advection
    # non-linear advection: blocks re-evaluated at the projected coarse solution
    raw2 = json.loads(str(g["codes_nonlin"]))
    op2 = V.SuyashGreenOperator(specs, decode(raw["symmetric"]), decode(raw2["advection"]), decode(raw2["advection_nonlin"]), batch, inc,
                                g["fine_forcing"], fine, coarse, conv, 1.0, krylov=SV.ConvergenceSettings(5000, 1e300, 1e-14))
    small = 0.3 * cs
    op2.update_nonlinear_advection(small)
    assert rel_max(op2.fine_advection.get()[0], g["advection_nonlin_0"]) < 1e-12
    op2.update_nonlinear_advection(small)  # a second update replaces the non-linear part, it does not accumulate
    assert rel_max(op2.fine_advection.get()[0], g["advection_nonlin_0"]) < 1e-12
    assert rel_max(op2.compute_unresolved_contributions(small[: n * nc], None), g["unresolved_nonlin"]) < TOL
    # the steady solve with the closure (solve_system.py:432-440 inside non_linear_solve_run) against the reference's own
    # solve_system_2d(vms_settings=...): same residual history, solution and fine scales; corrections by BiCGStab on the
    # Schur complement of the full (advective) operator instead of SuperLU
    from mfv2d_b200 import callers as CL

    lhs = decode(json.loads(str(g["solve_codes"]))["lhs"])
    con = (g["solve_con_row"], g["solve_con_col"], g["solve_con_val"])
    n_trace = g["solve_con_rhs"].size
    system = SV.compute_linear_system(specs, lhs, batch, con, n_trace)
    kry = SV.ConvergenceSettings(5000, 1e300, 1e-14)

    def solve(sysm, rd, rt):
        return SV.solve_schur_bicgstab(sysm, rd, rt, kry)[:2]

    for tag, sg in (("plain", None), ("vms", make(1.0))):
        hist = []
        out = CL.non_linear_solve_run(specs, lhs, None, batch, system, g["solve_explicit"], g["solve_con_rhs"], np.zeros(n * nc), np.zeros(n_trace),
                                      max_iterations=40, atol=1e-11, rtol=0.0, solve=solve, sg_operator=sg, residual_history=hist)
        ref_hist = g[f"solve_{tag}_residuals"]
        assert out[2] == ref_hist.size - 1, (tag, out[2], hist, ref_hist)
        assert np.allclose(hist[:-1], ref_hist[:-1], rtol=1e-6, atol=1e-11), (tag, hist, ref_hist)
        assert rel_max(out[0], g[f"solve_{tag}_solution"]) < 1e-9, (tag, rel_max(out[0], g[f"solve_{tag}_solution"]))
        if sg is not None:
            assert len(out) == 5 and rel_max(out[4], g["solve_vms_fine_scales"]) < 1e-7
        else:
            assert len(out) == 4
    return op.krylov_iterations


def test_vms_host_emulation():
    from build_emu import build

    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(build())
    try:
        from mfv2d_b200 import _mfv2d as M

        print(run_checks(M))
    finally:
        capi._use_library_for_testing(capi.LIB_PATH)


@pytest.mark.gpu
def test_vms_gpu():
    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(capi.LIB_PATH)
    from mfv2d_b200 import _mfv2d as M
    from mfv2d_b200.runtime import get_context

    assert get_context().handle
    print(run_checks(M))
