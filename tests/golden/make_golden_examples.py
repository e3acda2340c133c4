"""Element-level fixtures for the example systems that tests/golden/make_golden.py does not cover: the five remaining
examples/unsteady cases, in the form the reference evaluates them (after `update_system_for_time_march`).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_examples.py
  cavity_flow       examples/unsteady/plot_cavity_flow.py:64-69      vor0 / vel1 / pre2, convective term, march on vel (dt = 5)
  heat_direct       examples/unsteady/plot_heat_direct.py:110-113    u0, march on u
  reaction          examples/unsteady/plot_reaction.py:118-122       u0 / q1 sorted by order, march on u
  reaction_mixed    examples/unsteady/plot_reaction_mixed.py:111-114 u2 / q1, march on u
  vector_reaction   examples/unsteady/plot_vector_reaction.py:86-89  u1, march on u
(the steady examples and heat_mixed are in systems.json; the examples/refinement scripts solve the direct-Poisson and
advection-diffusion systems that are already there).  Writes tests/golden/systems_examples.json (bytecode of the
reference's CompiledSystem) and tests/golden/element_<name>.npz (its compute_element_matrix / compute_element_vector outputs on
the test quads, same layout as make_golden.py).
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
from make_golden import encode_codes, fem_space, zero, zero2  # noqa: E402
from mfv2d._mfv2d import compute_element_matrix, compute_element_vector  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.solve_system import TimeSettings  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

S2 = sys.modules["mfv2d.solve_system_2d"]
O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def make_systems():
    out = {}
    pre, vel, vor = KFormUnknown("pre", O2), KFormUnknown("vel", O1), KFormUnknown("vor", O0)
    w_pre, w_vel, w_vor = pre.weight, vel.weight, vor.weight
    re = 100.0
    cavity = KFormSystem(w_vor.derivative @ vel - w_vor @ vor == w_vor ^ zero2,
                         (1 / re) * (w_vel @ vor.derivative) + w_vel.derivative @ pre == -(vel * w_vel @ vor), w_pre @ vel.derivative == 0,
                         sorting=lambda f: f.order)
    out["cavity_flow"] = S2.update_system_for_time_march(TimeSettings(dt=5, nt=20, time_march_relations={w_vel: vel}), cavity)
    alpha, beta = 0.02, 1.0
    u0 = KFormUnknown("u", O0)
    v0 = u0.weight
    heat = KFormSystem(alpha * (v0.derivative @ u0.derivative) == beta * (v0 @ zero) - (beta - alpha * np.pi**2 / 2) * (v0 @ u0))
    out["heat_direct"] = S2.update_system_for_time_march(TimeSettings(dt=0.1, nt=10, time_march_relations={v0: u0}), heat)
    q1 = KFormUnknown("q", O1)
    p1 = q1.weight
    reaction = KFormSystem(alpha * (v0 @ u0) == alpha * (v0 @ zero), p1 @ q1 - p1 @ u0.derivative == 0, sorting=lambda f: f.order)
    out["reaction"] = S2.update_system_for_time_march(TimeSettings(dt=0.05, nt=10, time_march_relations={v0: u0}), reaction)
    u2 = KFormUnknown("u", O2)
    v2 = u2.weight
    reaction_mixed = KFormSystem(alpha * (v2 @ u2) == alpha * (v2 @ zero), p1.derivative @ u2 - p1 @ q1 == p1 ^ zero)
    out["reaction_mixed"] = S2.update_system_for_time_march(TimeSettings(dt=0.05, nt=10, time_march_relations={v2: u2}), reaction_mixed)
    u1 = KFormUnknown("u", O1)
    v1 = u1.weight
    vector_reaction = KFormSystem(alpha * (v1 @ u1) == alpha * (v1 @ zero2), sorting=lambda f: f.order)
    out["vector_reaction"] = S2.update_system_for_time_march(TimeSettings(dt=0.05, nt=10, time_march_relations={v1: u1}), vector_reaction)
    return out


def main():
    sysjson = {}
    rng = np.random.default_rng(4242)
    for name, system in make_systems().items():
        comp = CompiledSystem(system)
        specs = system.unknown_forms
        sysjson[name] = {"form_names": list(specs.names), "form_orders": [int(o) for o in specs.orders], "lhs": encode_codes(comp.lhs_codes),
                         "rhs": encode_codes(comp.rhs_codes), "linear": encode_codes(comp.linear_codes), "nonlin": encode_codes(comp.nonlin_codes)}
        data = {}
        for p, k in ((1, 2), (3, 3), (4, 2), (5, 0)):
            for iq, quad in enumerate(F.QUADS):
                space, _, _ = fem_space(p, k, quad)
                dofs = rng.standard_normal(specs.total_size(p, p))
                tag = f"p{p}k{k}q{iq}"
                data[tag + "_dofs"] = dofs
                data[tag + "_lhs"] = compute_element_matrix(specs, comp.lhs_codes, space, dofs)
                data[tag + "_lhs_vec"] = compute_element_vector(specs, comp.lhs_codes, space, dofs)
                if comp.rhs_codes is not None:
                    data[tag + "_rhs_vec"] = compute_element_vector(specs, comp.rhs_codes, space, dofs)
                if comp.nonlin_codes is not None:
                    data[tag + "_nonlin"] = compute_element_matrix(specs, comp.nonlin_codes, space, dofs)
        np.savez_compressed(os.path.join(HERE, f"element_{name}.npz"), **data)
        print(name, sysjson[name]["form_names"], sysjson[name]["form_orders"], "rhs" if comp.rhs_codes is not None else "-",
              "nonlin" if comp.nonlin_codes is not None else "-")
    with open(os.path.join(HERE, "systems_examples.json"), "w") as f:
        json.dump(sysjson, f, indent=1)


if __name__ == "__main__":
    main()
