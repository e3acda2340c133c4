"""Fixtures for the remaining time-march examples with INITIAL CONDITIONS: examples/unsteady/plot_reaction.py (u0, q1),
plot_reaction_mixed.py (u2, q1, weak boundary term) and plot_vector_reaction.py (u1); and plot_heat_mixed.py (u2, q1: no initial
condition, weak boundary term, right-hand-side bytecode).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_marches.py
For each, the reference's own `solve_system_2d` marches the example's system (:118-122, :111-114, :86-89) with its initial
conditions (L2 projection: `compute_element_dual` -> `compute_element_primal_from_dual`, solve_system_2d.py:196-212) for three
steps on `unit_square_mesh(4, 3, 3, deform)`; recorded are the bytecode of the time-march-updated system, the explicit right-hand
side and constraint matrix of `compute_linear_system`, the initial dual and primal vectors and the DoF vectors after every step.
Writes tests/golden/timemarch_reaction.npz (keys prefixed with the case name).
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
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.solve_system import ConvergenceSettings, SolverSettings, SystemSettings, TimeSettings  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


# the examples' own functions -------------------------------------------------------------------------------------------
def r_initial_u(x, y):
    return 2 * np.cos(np.pi * x / 2) * np.cos(np.pi * y / 2)


def r_final_u(x, y):
    return np.sin(np.pi * x) * np.cos(np.pi * y)


def m_final_u(x, y):
    return (1 - x**4) * (1 - y**4)


def m_initial_u(x, y):
    return (1 - x**2) * (1 - y**2)


def m_initial_q(x, y):
    return np.stack((-2 * x * (1 - y**2), -2 * y * (1 - x**2)), axis=-1)


def v_initial_u(x, y):
    return np.stack((2 * x * y, x**2 * y), axis=-1)


def v_final_u(x, y):
    return np.stack((-(x**2) * y, -x * y), axis=-1)


def cases():
    out = {}
    u0, q1 = KFormUnknown("u", O0), KFormUnknown("q", O1)
    v0, p1 = u0.weight, q1.weight
    out["reaction"] = (KFormSystem(0.25 * (v0 @ u0) == 0.25 * (v0 @ r_final_u), p1 @ q1 - p1 @ u0.derivative == 0, sorting=lambda f: f.order),
                       {v0: u0}, {u0: r_initial_u})
    u2, q = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v2, p = u2.weight, q.weight
    out["reaction_mixed"] = (KFormSystem(1.0 * (v2 @ u2) == 1.0 * (v2 @ m_final_u), p.derivative @ u2 - p @ q == p ^ m_final_u), {v2: u2},
                             {u2: m_initial_u, q: m_initial_q})
    u1 = KFormUnknown("u", O1)
    v1 = u1.weight
    out["vector_reaction"] = (KFormSystem(0.5 * (v1 @ u1) == 0.5 * (v1 @ v_final_u), sorting=lambda f: f.order), {v1: u1}, {u1: v_initial_u})
    # examples/unsteady/plot_heat_mixed.py:64-68: no initial condition, weak boundary term, the term in u on the right (fixed point)
    uh, qh = KFormUnknown("u", O2), KFormUnknown("q", O1)
    vh, ph = uh.weight, qh.weight
    out["heat_mixed"] = (KFormSystem(ph.derivative @ uh - ph @ qh == ph ^ h_steady_u,
                                     0.02 * (vh @ qh.derivative) == 1.0 * (vh @ h_steady_u) - (1.0 - 0.02 * np.pi**2 / 2) * (vh @ uh)), {vh: uh}, {})
    return out


def h_steady_u(x, y):
    return np.cos(np.pi * x / 2) * np.cos(np.pi * y / 2)


def main(nh=4, nv=3, order=3, dt=0.25, nt=3):
    data = dict(order=order, k=3, dt=dt, nt=nt, nh=nh, nv=nv)
    mesh0 = unit_square_mesh(nh, nv, order, F.deform)
    leaf = list(mesh0.get_leaf_indices())
    data["corners"] = np.array([np.array(mesh0.get_leaf_corners(i), np.float64) for i in leaf])
    for name, (system, relations, initial) in cases().items():
        mesh = unit_square_mesh(nh, nv, order, F.deform)
        cap = {"solutions": [], "duals": []}
        orig_recon, orig_lin, orig_dual = S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.compute_element_dual

        def recon(form_specs, recon_order, spaces, solution, fine_scales):
            cap["solutions"].append(np.array(solution, np.float64))
            return orig_recon(form_specs, 1, spaces, solution, fine_scales)

        def lin(*a, **kw):
            out = orig_lin(*a, **kw)
            cap["vectors"], cap["lag"], cap["lag_vec"] = [np.array(x) for x in out[0]], out[2], np.array(out[3])
            return out

        def dual(*a, **kw):
            out = orig_dual(*a, **kw)
            cap["duals"].append(np.array(out, np.float64))
            return out

        S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.compute_element_dual = recon, lin, dual
        try:
            ts = TimeSettings(dt=dt, nt=nt, time_march_relations=relations)
            S2.solve_system_2d(mesh, system_settings=SystemSettings(system, initial_conditions=initial),
                               solver_settings=SolverSettings(ConvergenceSettings(maximum_iterations=20, relative_tolerance=0, absolute_tolerance=1e-12)),
                               time_settings=ts, recon_order=1)
        finally:
            S2.reconstruct_mesh_from_solution, S2.compute_linear_system, S2.compute_element_dual = orig_recon, orig_lin, orig_dual
        marched = S2.update_system_for_time_march(ts, system)
        comp = CompiledSystem(marched)
        specs = marched.unknown_forms
        codes = {"form_names": list(specs.names), "form_orders": [int(x) for x in specs.orders], "lhs": encode_codes(comp.lhs_codes),
                 "rhs": encode_codes(comp.rhs_codes), "march_forms": sorted(list(marched.weight_forms).index(w) for w in relations)}
        assert len(cap["solutions"]) == nt + 1 and len(cap["duals"]) == (len(leaf) if initial else 0)
        data[name + "_codes"] = json.dumps(codes)
        data[name + "_explicit"] = np.concatenate(cap["vectors"])
        data[name + "_initial_dual"] = np.concatenate(cap["duals"]) if initial else np.zeros(0)
        data[name + "_solutions"] = np.stack(cap["solutions"])
        if cap["lag"] is not None:
            m = cap["lag"].tocoo()
            o = np.lexsort((m.col, m.row))
            data[name + "_con_row"], data[name + "_con_col"], data[name + "_con_val"] = m.row[o].astype(np.int64), m.col[o].astype(np.int64), m.data[o]
        else:
            data[name + "_con_row"], data[name + "_con_col"], data[name + "_con_val"] = np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0)
        data[name + "_con_rhs"] = cap["lag_vec"]
        print(name, codes["form_names"], codes["form_orders"], "march", codes["march_forms"], "rhs codes", comp.rhs_codes is not None, "rows",
              cap["lag_vec"].size, "|x| per step", [float(np.abs(s).max()) for s in cap["solutions"]])
    np.savez_compressed(os.path.join(HERE, "timemarch_reaction.npz"), **data)


if __name__ == "__main__":
    main()
