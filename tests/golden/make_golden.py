"""Generate the committed golden fixtures from the REFERENCE itself.

Run in the build container only (needs /root/reference and oracle/_ref/_mfv2d_ref.so):

    python tests/golden/make_golden.py

Outputs (all small, committed):
  systems.json        bytecode of the example form systems as produced by the reference's own
                      Python front end (mfv2d.eval.CompiledSystem); callables replaced by "fn:<name>"
  element_<sys>.npz   reference C outputs (basis tables, masses, inverses, element matrices,
                      element vectors) on the quads of tests/fields.py
  solve_<mesh>.npz    mixed-Poisson element-block system of the reference's tests/test_solving.py:
                      corners, constraints (COO), right-hand sides, reference operator outputs and
                      the reference's solve_schur_iterative solution
(constraints.npz - the rows of the reference's connect_elements / add_system_constraints on unit_square_mesh for the
vectorised builder - is written by make_golden_constraints.py, hp_refined*.npz by make_golden_hp.py,
timemarch_advdif.npz by make_golden_timemarch.py)
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

from oracle.ref_loader import load_ref_python  # noqa: E402

mfv2d = load_ref_python()
assert mfv2d is not None, "reference not available"

import fields as F  # noqa: E402
from mfv2d._mfv2d import (  # noqa: E402
    Basis1D,
    Basis2D,
    DenseVector,
    ElementFemSpace2D,
    IntegrationRule1D,
    TraceVector,
    compute_element_matrix,
    compute_element_vector,
)
from mfv2d.continuity import connect_elements  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache  # noqa: E402
from mfv2d.solve_system import compute_element_rhs  # noqa: E402
from mfv2d.solving import ConvergenceSettings, LinearSystem, solve_schur_iterative  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def zero(x, y):
    return 0 * x


def zero2(x, y):
    return np.stack((0 * x, 0 * x), axis=-1)


def make_systems():
    out = {}
    # mixed Poisson (examples/steady/plot_mixed_poisson.py:160-163)
    u = KFormUnknown("u", O2)
    q = KFormUnknown("q", O1)
    v, p = u.weight, q.weight
    out["mixed_poisson"] = KFormSystem(p.derivative @ u - p @ q == p ^ zero, v @ q.derivative == -(v @ zero))
    # test_solving variant (tests/test_solving.py:478-482), sorted by order
    out["mixed_poisson_sorted"] = KFormSystem(
        p @ q + p.derivative @ u == 0, v @ q.derivative == v @ zero, sorting=lambda f: f.order
    )
    # linear advection-diffusion (examples/steady/plot_linear_adv_dif.py:124-127)
    out["adv_dif"] = KFormSystem(
        p.derivative @ u - p @ q == p ^ zero,
        -0.05 * (v @ q.derivative) - (F.adv_field * v @ q) == -(v @ zero),
    )
    # direct Poisson (examples/steady/plot_direct_poisson.py:154-158)
    u0 = KFormUnknown("u", O0)
    q1 = KFormUnknown("q", O1)
    v0, p1 = u0.weight, q1.weight
    out["direct_poisson"] = KFormSystem(
        v0.derivative @ u0.derivative == -(v0 @ zero),
        p1 @ q1 - p1 @ u0.derivative == 0,
    )
    # Stokes (examples/steady/plot_stokes_flow.py:246-252)
    vor = KFormUnknown("vor", O0)
    vel = KFormUnknown("vel", O1)
    prs = KFormUnknown("prs", O2)
    w_vor, w_vel, w_prs = vor.weight, vel.weight, prs.weight
    out["stokes"] = KFormSystem(
        w_vor.derivative @ vel + w_vor @ vor == w_vor ^ zero2,
        w_vel @ vor.derivative + w_vel.derivative @ prs == (w_vel ^ zero) + w_vel @ zero2,
        w_prs @ vel.derivative == 0,
        sorting=lambda f: f.order,
    )
    # Navier-Stokes (examples/steady/plot_navier_stokes.py:136-143), without the div extraction
    out["navier_stokes"] = KFormSystem(
        w_vor.derivative @ vel - w_vor @ vor == w_vor ^ zero2,
        (1 / 10.0) * (w_vel @ vor.derivative) + w_vel.derivative @ prs == w_vel @ zero2 - (vel * w_vel @ vor),
        (w_prs @ vel.derivative) == 0,
        sorting=lambda f: f.order,
    )
    # heat-like system with a reaction (mass) term on the 2-form, as update_system_for_time_march adds
    out["heat_mixed"] = KFormSystem(
        p.derivative @ u - p @ q == p ^ zero,
        v @ q.derivative + 20.0 * (v @ u) == -(v @ zero) + 3.0 * (v @ u),
    )
    return out


def encode_codes(codes):
    if codes is None:
        return None
    rows = []
    for row in codes:
        r = []
        for blk in row:
            if blk is None:
                r.append(None)
                continue
            ops = []
            for item in blk:
                enc = []
                for x in item:
                    if callable(x):
                        assert x.__name__ in F.FIELDS, x
                        enc.append("fn:" + x.__name__)
                    elif isinstance(x, (bool, np.bool_)):
                        enc.append(bool(x))
                    elif isinstance(x, (int, np.integer)) or hasattr(x, "value"):
                        enc.append(int(x))
                    elif isinstance(x, float):
                        enc.append(float(x))
                    elif isinstance(x, str):
                        enc.append(x)
                    else:
                        raise TypeError(type(x))
                ops.append(enc)
            r.append(ops)
        rows.append(r)
    return rows


def fem_space(p, k, corners):
    rule = IntegrationRule1D(p + k)
    b1 = Basis1D(p, rule)
    b2 = Basis2D(b1, b1)
    return ElementFemSpace2D(b2, np.ascontiguousarray(corners, np.float64)), b1, rule


def main():
    systems = make_systems()
    sysjson = {}
    rng = np.random.default_rng(2024)
    for name, system in systems.items():
        comp = CompiledSystem(system)
        specs = system.unknown_forms
        orders = [int(o) for o in specs.orders]
        form_names = list(specs.names)
        sysjson[name] = {
            "form_names": form_names,
            "form_orders": orders,
            "lhs": encode_codes(comp.lhs_codes),
            "rhs": encode_codes(comp.rhs_codes),
            "linear": encode_codes(comp.linear_codes),
            "nonlin": encode_codes(comp.nonlin_codes),
        }
        # reference outputs on the test quads for two (p, k)
        data = {}
        for p, k in ((1, 2), (3, 3), (4, 2), (5, 0)):
            for iq, quad in enumerate(F.QUADS):
                space, b1, rule = fem_space(p, k, quad)
                n = specs.total_size(p, p)
                dofs = rng.standard_normal(n)
                tag = f"p{p}k{k}q{iq}"
                data[tag + "_dofs"] = dofs
                data[tag + "_lhs"] = compute_element_matrix(specs, comp.lhs_codes, space, dofs)
                data[tag + "_lhs_vec"] = compute_element_vector(specs, comp.lhs_codes, space, dofs)
                if comp.rhs_codes is not None:
                    data[tag + "_rhs_vec"] = compute_element_vector(specs, comp.rhs_codes, space, dofs)
                if comp.nonlin_codes is not None:
                    data[tag + "_nonlin"] = compute_element_matrix(specs, comp.nonlin_codes, space, dofs)
                if iq == 4 and name == "stokes":
                    data[tag + "_node"] = np.array(b1.node)
                    data[tag + "_edge"] = np.array(b1.edge)
                    data[tag + "_pnts"] = np.array(rule.nodes)
                    data[tag + "_wgts"] = np.array(rule.weights)
                    for fo in (1, 2, 3):
                        data[tag + f"_m{fo - 1}"] = space.mass_from_order(UnknownFormOrder(fo), False)
                        data[tag + f"_m{fo - 1}i"] = space.mass_from_order(UnknownFormOrder(fo), True)
        np.savez_compressed(os.path.join(HERE, f"element_{name}.npz"), **data)
    with open(os.path.join(HERE, "systems.json"), "w") as f:
        json.dump(sysjson, f, indent=1)

    # ---- element-block systems of tests/test_solving.py:478-538 ---------------------------------
    system = systems["mixed_poisson_sorted"]
    comp = CompiledSystem(system)

    def f_exact(x, y):
        return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)

    u = KFormUnknown("u", O2)
    q = KFormUnknown("q", O1)
    v, p_ = u.weight, q.weight
    system_f = KFormSystem(p_ @ q + p_.derivative @ u == 0, v @ q.derivative == v @ f_exact, sorting=lambda f: f.order)
    comp_f = CompiledSystem(system_f)
    for nh, nv, order, deformed in ((10, 10, 3, False), (3, 4, 4, False), (5, 2, 5, True), (6, 6, 3, True)):
        mesh = unit_square_mesh(nh, nv, order, F.deform if deformed else None)
        cache = FemCache(2)
        leaf = mesh.get_leaf_indices()
        spaces = [
            ElementFemSpace2D(cache.get_basis2d(*mesh.get_leaf_orders(i)), np.astype(mesh.get_leaf_corners(i), np.float64, copy=False))
            for i in leaf
        ]
        corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
        vectors = [compute_element_rhs(system_f, s) for s in spaces]
        mats = [compute_element_matrix(system_f.unknown_forms, comp_f.lhs_codes, s) for s in spaces]
        cons = connect_elements(system_f.unknown_forms, mesh)
        orders = np.array([mesh.get_leaf_orders(i) for i in leaf], np.uint32)
        ls = LinearSystem(mesh.leaf_count, system_f.unknown_forms, orders, mats, cons)
        dv = DenseVector(ls, *vectors)
        tv = TraceVector(ls, np.array([c.rhs for c in cons], np.float64))
        # constraint COO: (row, element, local dof, coeff)
        rows, elems, dofs, vals = [], [], [], []
        for ic, c in enumerate(cons):
            for ec in c.element_constraints:
                for d, cf in zip(ec.dofs, ec.coeffs):
                    rows.append(ic), elems.append(ec.i_e), dofs.append(int(d)), vals.append(float(cf))
        # reference operator outputs on a random vector
        rngv = np.random.default_rng(7)
        n_dense = sum(m.shape[0] for m in mats)
        xs = [rngv.standard_normal(m.shape[0]) for m in mats]
        xv = DenseVector(ls, *xs)
        o1, o2 = DenseVector(ls), DenseVector(ls)
        ls.apply_diagonal(xv, o1)
        ls.apply_diagonal_inverse(xv, o2)
        t1 = TraceVector(ls)
        ls.apply_trace(xv, t1)
        tin = TraceVector(ls, rngv.standard_normal(len(cons)))
        o3 = DenseVector(ls)
        ls.apply_trace_transpose(tin, o3)
        sol_d, sol_t, resid, iters = solve_schur_iterative(ls, dv, tv, ConvergenceSettings(2000, 1e-30, 1e-12))
        np.savez_compressed(
            os.path.join(HERE, f"solve_{nh}x{nv}_p{order}{'_def' if deformed else ''}.npz"),
            corners=corners,
            order=order,
            k=2,
            con_row=np.array(rows, np.uint32),
            con_elem=np.array(elems, np.uint32),
            con_dof=np.array(dofs, np.uint32),
            con_val=np.array(vals),
            con_rhs=np.array([c.rhs for c in cons]),
            rhs=np.concatenate(vectors),
            x=np.concatenate(xs),
            trace_in=np.array(tin.as_merged()),
            apply_diagonal=np.array(o1.as_merged()),
            apply_diagonal_inverse=np.array(o2.as_merged()),
            apply_trace=np.array(t1.as_merged()),
            apply_trace_transpose=np.array(o3.as_merged()),
            sol_dense=np.array(sol_d.as_merged()),
            sol_trace=np.array(sol_t.as_merged()),
            iters=iters,
            resid=resid,
            mat0=mats[0],
        )
        print("solve fixture", nh, nv, order, "iters", iters, "resid", resid)


if __name__ == "__main__":
    main()
