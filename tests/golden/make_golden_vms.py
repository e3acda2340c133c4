"""Fixture for the fine-scale Green's operator (mfv2d_b200/vms.py): the reference's own ``SuyashGreenOperator``.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_vms.py
Writes tests/golden/vms.npz.  On the deformed 3 x 2 mesh of order 3 (FemCache(2)), order increase 1:
  symmetric system      p.derivative @ u - p @ q == 0,  v @ q.derivative == v @ source        (mixed diffusion)
  non-symmetric system  0 (p @ q) == 0,                 adv_field * v @ q == 0                 (linear advection)
  full system           their sum
the operator built by python/mfv2d/solve_system.py:653-856 and the outputs of its methods (:858-961) on fixed pseudo-random
vectors: the bytecode of the systems, the constraint rows ``add_system_constraints`` returned for the fine and the coarse
mesh, the fine forcing, the two projectors (element 0), ``fine_scale_greens_function``, ``compute_unresolved_contributions``
(with and without an initial guess, relaxation 1 and 0.7), ``recover_unresolved`` and ``fine_results_to_coarse_dofs``; and
the same operator with the non-linear advection ``0.3 (q * v @ q)`` added: ``update_nonlinear_advection`` at a scaled coarse
solution, the advection block of element 0 and the unresolved contributions after it.  ``solve_*``: the steady solve of the
full system without and with the closure by the reference's ``solve_system_2d`` (see ``solve_fixture``).
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

from oracle.ref_loader import load_ref_python  # noqa: E402

mfv2d = load_ref_python()
import fields as F  # noqa: E402
import mfv2d.solve_system as SS  # noqa: E402
from make_golden import encode_codes  # noqa: E402
from mfv2d._mfv2d import ElementFemSpace2D  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache  # noqa: E402
from mfv2d.solve_system import ConvergenceSettings, SuyashGreenOperator, VMSSettings  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O1, O2 = UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def source(x, y):
    return 3.0 * np.cos(np.pi * x) * np.sin(0.7 * np.pi * y)


def coo(m):
    m = m.tocoo()
    o = np.lexsort((m.col, m.row))
    return m.row[o].astype(np.int64), m.col[o].astype(np.int64), m.data[o].astype(np.float64), int(m.shape[0])


def solve_fixture(mesh, sym, non, full, increase, conv):
    """The steady solve of the full system with the VMS closure through the reference's own driver (solve_system_2d with
    vms_settings, python/mfv2d/solve_system_2d.py:303-316, solve_system.py:432-440): right-hand side and constraint rows of
    compute_linear_system, the residual history of non_linear_solve_run, the solution and the fine scales it ends with."""
    import mfv2d.solve_system_2d  # noqa: F401
    from mfv2d.solve_system import SolverSettings, SystemSettings

    S2 = sys.modules["mfv2d.solve_system_2d"]
    cap = {}
    orig_recon, orig_nl, orig_lin = S2.reconstruct_mesh_from_solution, S2.non_linear_solve_run, S2.compute_linear_system

    def recon(form_specs, recon_order, spaces, solution, fine_scales):
        cap["solution"] = np.array(solution, np.float64)
        cap["fine"] = None if fine_scales is None else np.array(fine_scales, np.float64)
        return orig_recon(form_specs, 1, spaces, solution, None)

    def nl(*a, **kw):
        out = orig_nl(*a, **kw)
        cap["iterations"], cap["residuals"] = int(out[2]), np.array(out[3]).ravel()[: int(out[2]) + 1]
        return out

    def lin(*a, **kw):
        out = orig_lin(*a, **kw)
        cap["explicit"] = np.concatenate([np.array(x) for x in out[0]])
        cap["lag"], cap["lag_vec"] = out[2], np.array(out[3])
        return out

    S2.reconstruct_mesh_from_solution, S2.non_linear_solve_run, S2.compute_linear_system = recon, nl, lin
    out = {}
    try:
        for tag, vms in (("plain", None), ("vms", VMSSettings(sym, non, increase, conv, 1.0))):
            S2.solve_system_2d(mesh, system_settings=SystemSettings(full, over_integration_order=2),
                               solver_settings=SolverSettings(ConvergenceSettings(maximum_iterations=40, relative_tolerance=0, absolute_tolerance=1e-11)),
                               vms_settings=vms, recon_order=1)
            out[f"solve_{tag}_solution"], out[f"solve_{tag}_residuals"] = cap["solution"], cap["residuals"]
            if vms is not None:
                out["solve_vms_fine_scales"] = cap["fine"]
    finally:
        S2.reconstruct_mesh_from_solution, S2.non_linear_solve_run, S2.compute_linear_system = orig_recon, orig_nl, orig_lin
    cfull = CompiledSystem(full)
    assert cfull.rhs_codes is None
    out["solve_codes"] = json.dumps({"lhs": encode_codes(cfull.lhs_codes)})
    out["solve_explicit"] = cap["explicit"]
    r, c, val, rows = coo(cap["lag"])
    out["solve_con_row"], out["solve_con_col"], out["solve_con_val"], out["solve_con_rhs"] = r, c, val, cap["lag_vec"]
    return out


def main(nh=3, nv=2, order=3, increase=1):
    u, q = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v, p = u.weight, q.weight
    sym = KFormSystem(p.derivative @ u - p @ q == 0, (v @ q.derivative) == v @ source)
    non = KFormSystem(0 * (p @ q) == 0, (F.adv_field * v @ q) == 0)
    full = KFormSystem(p.derivative @ u - p @ q == 0, (v @ q.derivative) + (F.adv_field * v @ q) == v @ source)
    mesh = unit_square_mesh(nh, nv, order, F.deform)
    cache = FemCache(order_difference=2)
    leaf = tuple(int(i) for i in mesh.get_leaf_indices())
    corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
    spaces = [ElementFemSpace2D(cache.get_basis2d(*mesh.get_leaf_orders(i)), corners[k]) for k, i in enumerate(leaf)]
    captured = []
    orig = SS.add_system_constraints

    def capture(*args, **kwargs):
        out = orig(*args, **kwargs)
        captured.append(out)
        return out

    SS.add_system_constraints = capture
    try:
        conv = ConvergenceSettings(maximum_iterations=60, absolute_tolerance=1e-13, relative_tolerance=1e-12)
        op = SuyashGreenOperator(full, VMSSettings(sym, non, increase, conv, 1.0), spaces, cache, mesh, leaf, [], [])
        op_relaxed = SuyashGreenOperator(full, VMSSettings(sym, non, increase, conv, 0.7), spaces, cache, mesh, leaf, [], [])
    finally:
        SS.add_system_constraints = orig
    (fine_mat, fine_vec), (coarse_mat, coarse_vec) = captured[0], captured[1]
    assert not np.any(fine_vec) and not np.any(coarse_vec)
    specs = sym.unknown_forms
    csym, cnon = CompiledSystem(sym), CompiledSystem(non)
    assert cnon.nonlin_codes is None
    codes = {"form_names": list(specs.names), "form_orders": [int(x) for x in specs.orders], "symmetric": encode_codes(csym.lhs_codes),
             "advection": encode_codes(cnon.linear_codes)}
    n, nf, nc = len(leaf), specs.total_size(order + increase, order + increase), specs.total_size(order, order)
    rng = np.random.default_rng(5)
    x_fine = rng.standard_normal(n * nf)
    coarse_solution = rng.standard_normal(n * nc + op.coarse_padding)
    guess = 0.1 * rng.standard_normal(n * nf)
    data = dict(corners=corners, order=order, k=2, increase=increase, codes=json.dumps(codes), fine_forcing=op.fine_forcing,
                x_fine=x_fine, coarse_solution=coarse_solution, guess=guess,
                projector_c2f_0=op.projector_c2f.toarray()[:nf, :nc], projector_f2c_0=op.projector_f2c.toarray()[:nc, :nf])
    for tag, m in (("fine", fine_mat), ("coarse", coarse_mat)):
        r, c, val, rows = coo(m)
        data[f"{tag}_row"], data[f"{tag}_col"], data[f"{tag}_val"], data[f"{tag}_rows"] = r, c, val, rows
    data["green"] = op.fine_scale_greens_function(x_fine)
    data["unresolved"] = op.compute_unresolved_contributions(coarse_solution[: n * nc], None)
    data["unresolved_guess"] = op.compute_unresolved_contributions(coarse_solution[: n * nc], guess)
    data["unresolved_relaxed"] = op_relaxed.compute_unresolved_contributions(coarse_solution[: n * nc], None)
    data["recovered"] = op.recover_unresolved(coarse_solution, data["unresolved"])
    data["to_coarse_dual"] = op.fine_results_to_coarse_dofs(x_fine, dual=True)
    data["to_coarse_primal"] = op.fine_results_to_coarse_dofs(x_fine, dual=False)
    # non-linear advection (the unknown 1-form advects itself): update_nonlinear_advection (:922-945) at a scaled coarse solution
    non2 = KFormSystem(0 * (p @ q) == 0, (F.adv_field * v @ q) + 0.3 * (q * v @ q) == 0)
    full2 = KFormSystem(p.derivative @ u - p @ q == 0, (v @ q.derivative) + (F.adv_field * v @ q) + 0.3 * (q * v @ q) == v @ source)
    cnon2 = CompiledSystem(non2)
    assert cnon2.nonlin_codes is not None
    codes2 = {"advection": encode_codes(cnon2.linear_codes), "advection_nonlin": encode_codes(cnon2.nonlin_codes)}
    op2 = SuyashGreenOperator(full2, VMSSettings(sym, non2, increase, conv, 1.0), spaces, cache, mesh, leaf, [], [])
    small = 0.3 * coarse_solution
    op2.update_nonlinear_advection(small)
    data["codes_nonlin"] = json.dumps(codes2)
    data["unresolved_nonlin"] = op2.compute_unresolved_contributions(small[: n * nc], None)
    data["advection_nonlin_0"] = op2.fine_advection_operator.toarray()[:nf, :nf]
    data.update(solve_fixture(mesh, sym, non, full, increase, conv))
    np.savez_compressed(os.path.join(HERE, "vms.npz"), **data)
    for k_, v_ in data.items():
        print(k_, np.asarray(v_).shape, float(np.abs(np.asarray(v_, dtype=float)).max()) if not isinstance(v_, str) else "")


if __name__ == "__main__":
    main()
