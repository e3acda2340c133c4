"""Iteration counts / residuals of the REFERENCE's solve_pcg_iterative (python/mfv2d/solving.py:634-684; on the small
systems also solve_cg_iterative :590-631 and solve_gmres_iterative :553-587) on the
meshes of the solve_*.npz fixtures.  Build container only (needs /root/reference and oracle/_ref):

    python tests/golden/make_golden_pcg.py        ->  tests/golden/pcg.json
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

import fields as F  # noqa: E402
from oracle.ref_loader import load_ref_python  # noqa: E402


def main():
    mf = load_ref_python()
    assert mf is not None
    from mfv2d._mfv2d import DenseVector, ElementFemSpace2D, TraceVector, compute_element_matrix
    from mfv2d.continuity import connect_elements
    from mfv2d.eval import CompiledSystem
    from mfv2d.examples import unit_square_mesh
    from mfv2d.kform import KFormUnknown, UnknownFormOrder
    from mfv2d.system import KFormSystem
    from mfv2d.mimetic2d import FemCache
    from mfv2d.solve_system import compute_element_rhs
    from mfv2d.solving import (ConvergenceSettings, LinearSystem, solve_cg_iterative, solve_gmres_iterative,
                               solve_pcg_iterative)

    def f_exact(x, y):
        return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)

    u = KFormUnknown("u", UnknownFormOrder.FORM_ORDER_2)
    q = KFormUnknown("q", UnknownFormOrder.FORM_ORDER_1)
    v, p_ = u.weight, q.weight
    system_f = KFormSystem(p_ @ q + p_.derivative @ u == 0, v @ q.derivative == v @ f_exact, sorting=lambda f: f.order)
    comp_f = CompiledSystem(system_f)
    out = {}
    for nh, nv, order, deformed in ((10, 10, 3, False), (3, 4, 4, False), (5, 2, 5, True), (6, 6, 3, True)):
        mesh = unit_square_mesh(nh, nv, order, F.deform if deformed else None)
        cache = FemCache(2)
        leaf = mesh.get_leaf_indices()
        spaces = [ElementFemSpace2D(cache.get_basis2d(*mesh.get_leaf_orders(i)), np.astype(mesh.get_leaf_corners(i), np.float64, copy=False))
                  for i in leaf]
        vectors = [compute_element_rhs(system_f, s) for s in spaces]
        mats = [compute_element_matrix(system_f.unknown_forms, comp_f.lhs_codes, s) for s in spaces]
        cons = connect_elements(system_f.unknown_forms, mesh)
        orders = np.array([mesh.get_leaf_orders(i) for i in leaf], np.uint32)
        ls = LinearSystem(mesh.leaf_count, system_f.unknown_forms, orders, mats, cons)
        dv = DenseVector(ls, *vectors)
        tv = TraceVector(ls, np.array([c.rhs for c in cons], np.float64))
        sd, st, resid, iters = solve_pcg_iterative(ls, dv, tv, ConvergenceSettings(5000, 1e-30, 1e-10))
        name = f"solve_{nh}x{nv}_p{order}{'_def' if deformed else ''}"
        out[name] = {"iterations": int(iters), "residual": float(resid)}
        if mesh.leaf_count <= 12:  # the other two drivers of solving.py on the small systems
            ref_d, ref_t = np.array(sd.as_merged()), np.array(st.as_merged())

            def err(d, t):
                return float(max(np.abs(np.array(d.as_merged()) - ref_d).max() / np.abs(ref_d).max(),
                                 np.abs(np.array(t.as_merged()) - ref_t).max() / np.abs(ref_t).max()))

            for key, fn, settings in (("cg", solve_cg_iterative, ConvergenceSettings(3000, 1e-30, 1e-10)),
                                      ("gmres", solve_gmres_iterative, ConvergenceSettings(400, 1e-30, 1e-10))):
                try:
                    d2, t2, r2, i2 = fn(ls, dv, tv, settings)
                    out[name][key] = {"iterations": int(i2), "residual": float(r2), "err_vs_pcg": err(d2, t2)}
                except Exception as exc:  # recorded: the reference's own driver does not run on its own system
                    out[name][key] = {"reference_raises": f"{type(exc).__name__}: {exc}"}
        print(name, out[name])
    with open(os.path.join(HERE, "pcg.json"), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
