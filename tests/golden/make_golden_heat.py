"""Fixture for a time march with a 0-FORM unknown: the unsteady direct heat equation of examples/unsteady/plot_heat_direct.py.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_heat.py
The reference's own `solve_system_2d` marches  alpha (dv, du) == beta (v, u_s) - (beta - alpha pi^2 / 2) (v, u)  (:110-113) with
`TimeSettings(dt, nt, {v: u})` and the strong condition u = u_s on the whole boundary on `unit_square_mesh(4, 3, 3, deform)`.
What is new against timemarch_advdif.npz: the constraint matrix carries 0-form continuity rows (edge rows AND corner chains)
and strong boundary rows with a non-zero right-hand side, the system has right-hand-side bytecode (the term in u on the right:
every step is a fixed-point iteration), and the element blocks are symmetric positive definite (CG on the Schur complement).
Recorded: bytecode of the time-march-updated system, explicit right-hand side, constraint matrix and vector, DoF vectors after
every step, iterations per step.  Writes tests/golden/timemarch_heat.npz.
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
import mfv2d.solve_system_2d  # noqa: E402,F401

S2 = sys.modules["mfv2d.solve_system_2d"]
from make_golden import encode_codes  # noqa: E402
from mfv2d.boundary import BoundaryCondition2DSteady  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.solve_system import ConvergenceSettings, SolverSettings, SystemSettings, TimeSettings  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

ALPHA, BETA = 0.02, 1.0


def steady_u(x, y):
    return np.cos(np.pi * x / 2) * np.cos(np.pi * y / 2)


def main(nh=4, nv=3, order=3, dt=0.1, nt=3):
    u = KFormUnknown("u", UnknownFormOrder.FORM_ORDER_0)
    v = u.weight
    system = KFormSystem(ALPHA * (v.derivative @ u.derivative) == BETA * (v @ steady_u) - (BETA - ALPHA * np.pi**2 / 2) * (v @ u))
    mesh = unit_square_mesh(nh, nv, order, F.deform)
    cap = {"solutions": [], "iters": []}
    orig_recon, orig_lin, orig_nl = S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.non_linear_solve_run

    def recon(form_specs, recon_order, spaces, solution, fine_scales):
        cap["solutions"].append(np.array(solution, np.float64))
        return orig_recon(form_specs, 1, spaces, solution, fine_scales)

    def lin(*a, **kw):
        out = orig_lin(*a, **kw)
        cap["vectors"], cap["mat0"], cap["lag"], cap["lag_vec"] = [np.array(x) for x in out[0]], np.array(out[1][0]), out[2].tocoo(), np.array(out[3])
        return out

    def nl(*a, **kw):
        out = orig_nl(*a, **kw)
        cap["iters"].append(int(out[2]))
        return out

    S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.non_linear_solve_run = recon, lin, nl
    try:
        ts = TimeSettings(dt=dt, nt=nt, time_march_relations={v: u})
        S2.solve_system_2d(mesh, system_settings=SystemSettings(system, boundary_conditions=[BoundaryCondition2DSteady(u, mesh.boundary_indices, steady_u)]),
                           solver_settings=SolverSettings(ConvergenceSettings(maximum_iterations=40, relative_tolerance=0, absolute_tolerance=1e-12)),
                           time_settings=ts, recon_order=1)
    finally:
        S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.non_linear_solve_run = orig_recon, orig_lin, orig_nl
    marched = S2.update_system_for_time_march(ts, system)
    comp = CompiledSystem(marched)
    specs = marched.unknown_forms
    leaf = list(mesh.get_leaf_indices())
    corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
    m = cap["lag"]
    o = np.lexsort((m.col, m.row))
    codes = {"form_names": list(specs.names), "form_orders": [int(x) for x in specs.orders], "lhs": encode_codes(comp.lhs_codes),
             "rhs": encode_codes(comp.rhs_codes)}
    assert comp.rhs_codes is not None and len(cap["solutions"]) == nt + 1
    np.savez_compressed(os.path.join(HERE, "timemarch_heat.npz"), corners=corners, order=order, k=3, dt=dt, nt=nt, nh=nh, nv=nv, codes=json.dumps(codes),
                        mat0=cap["mat0"], explicit=np.concatenate(cap["vectors"]), con_row=m.row[o].astype(np.int64), con_col=m.col[o].astype(np.int64),
                        con_val=m.data[o], con_rhs=cap["lag_vec"], solutions=np.stack(cap["solutions"]), iterations=np.array(cap["iters"]))
    print("timemarch_heat.npz:", corners.shape[0], "elements, order", order, "rows", m.shape[0], "iterations per step", cap["iters"],
          "|u| per step", [float(np.abs(s).max()) for s in cap["solutions"]])


if __name__ == "__main__":
    main()
