"""Fixture for BASELINE.json configs[2] (slice): unsteady linear advection-diffusion with the trapezoidal time march.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_timemarch.py
The reference's own `solve_system_2d` (python/mfv2d/solve_system_2d.py:36-478) marches the system of
examples/steady/plot_linear_adv_dif.py:124-127 with `TimeSettings(dt, nt, {v: u})` on `unit_square_mesh(8, 8, 4, deform)` (SystemSettings default over_integration_order = 3);
the raw DoF vectors it passes to `reconstruct_mesh_from_solution` after every step, the output of its
`compute_linear_system` (explicit right-hand side incl. the weak boundary terms, constraint matrix) and the bytecode of the
time-march-updated system are recorded.  Writes tests/golden/timemarch_advdif.npz.
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

S2 = sys.modules["mfv2d.solve_system_2d"]  # the package attribute of that name is the function
from make_golden import encode_codes  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.solve_system import ConvergenceSettings, SolverSettings, SystemSettings, TimeSettings  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O1, O2 = UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2
NU = 0.05


def u_bc(x, y):
    return np.sin(0.5 * np.pi * x) * np.cos(0.5 * np.pi * y) + 1.0


def source(x, y):
    return 3.0 * np.cos(np.pi * x) * np.sin(0.7 * np.pi * y)


def main(nh=8, nv=8, order=4, dt=0.05, nt=3):
    u, q = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v, p = u.weight, q.weight
    # the second equation of the steady example with its sign flipped: the trapezoidal march adds + 2/dt (v @ u) to the LEFT side
    # (update_system_for_time_march), which is dissipative only for u_t - nu lap(u) - a . grad(u) = f
    system = KFormSystem(p.derivative @ u - p @ q == p ^ u_bc, NU * (v @ q.derivative) + (F.adv_field * v @ q) == v @ source)
    mesh = unit_square_mesh(nh, nv, order, F.deform)
    captured = {"solutions": []}
    orig_recon, orig_lin = S2.reconstruct_mesh_from_solution, S2.compute_linear_system

    def recon(form_specs, recon_order, spaces, solution, fine_scales):
        captured["solutions"].append(np.array(solution, np.float64))
        return orig_recon(form_specs, 1, spaces, solution, fine_scales)

    def lin(*args, **kwargs):
        out = orig_lin(*args, **kwargs)
        captured["vectors"] = [np.array(x) for x in out[0]]
        captured["mat0"] = np.array(out[1][0])
        captured["lagrange_mat"] = out[2].tocoo()
        captured["lagrange_vec"] = np.array(out[3])
        return out

    S2.reconstruct_mesh_from_solution, S2.compute_linear_system = recon, lin
    try:
        time_settings = TimeSettings(dt=dt, nt=nt, time_march_relations={v: u})
        S2.solve_system_2d(mesh, system_settings=SystemSettings(system),
                           solver_settings=SolverSettings(ConvergenceSettings(maximum_iterations=20, relative_tolerance=0, absolute_tolerance=1e-13)),
                           time_settings=time_settings, recon_order=1)
    finally:
        S2.reconstruct_mesh_from_solution, S2.compute_linear_system = orig_recon, orig_lin
    marched = S2.update_system_for_time_march(time_settings, system)
    comp = CompiledSystem(marched)
    specs = marched.unknown_forms
    leaf = list(mesh.get_leaf_indices())
    corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
    m = captured["lagrange_mat"]
    o = np.lexsort((m.col, m.row))
    codes = {"form_names": list(specs.names), "form_orders": [int(x) for x in specs.orders], "lhs": encode_codes(comp.lhs_codes),
             "rhs": encode_codes(comp.rhs_codes)}
    assert len(captured["solutions"]) == nt + 1
    np.savez_compressed(os.path.join(HERE, "timemarch_advdif.npz"), corners=corners, order=order, k=3, dt=dt, nt=nt, nh=nh, nv=nv,
                        codes=json.dumps(codes), mat0=captured["mat0"], explicit=np.concatenate(captured["vectors"]),
                        con_row=m.row[o].astype(np.int64), con_col=m.col[o].astype(np.int64), con_val=m.data[o], con_rhs=captured["lagrange_vec"],
                        march_form=int(specs.names.index("u")) if hasattr(specs, "names") else 1, solutions=np.stack(captured["solutions"]))
    print("timemarch_advdif.npz:", corners.shape[0], "elements, order", order, "steps", nt, "|u| per step",
          [float(np.abs(s).max()) for s in captured["solutions"]], "rhs codes", comp.rhs_codes is not None)


if __name__ == "__main__":
    main()
