"""Fixtures for the array-based constraint builder (mfv2d_b200/structured.py): the reference's own constraint rows.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_constraints.py
Writes tests/golden/constraints.npz: for a few conforming meshes and the form layouts of the example systems the COO
(row, element, local dof, coefficient) and right-hand sides produced by
  * connect_elements             (python/mfv2d/continuity.py:610-759): 0-form edge + corner rows, 1-form edge rows
  * add_system_constraints       (:762-872): the same followed by gauge rows (constrained_forms) and strong-BC rows
    (mesh_boundary_conditions, python/mfv2d/boundary.py:382-498)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.ref_loader import load_ref_python  # noqa: E402

mfv2d = load_ref_python()
import fields as F  # noqa: E402
from mfv2d.boundary import BoundaryCondition2DSteady  # noqa: E402
from mfv2d.continuity import add_system_constraints, connect_elements  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402
from mfv2d.mimetic2d import FemCache  # noqa: E402

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def coo(cons):
    rows, elems, dofs, vals = [], [], [], []
    for ic, c in enumerate(cons):
        for ec in c.element_constraints:
            for d, cf in zip(ec.dofs, ec.coeffs):
                rows.append(ic), elems.append(ec.i_e), dofs.append(int(d)), vals.append(float(cf))
    return (np.array(rows, np.int64), np.array(elems, np.int64), np.array(dofs, np.int64), np.array(vals, np.float64),
            np.array([c.rhs for c in cons], np.float64))


def bc_u(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y - 0.1)


def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)


def main():
    data = {}
    # Stokes (vor0, vel1, prs2), sorted by order: continuity of the 0-form and of the 1-form
    vor, vel, prs = KFormUnknown("vor", O0), KFormUnknown("vel", O1), KFormUnknown("prs", O2)
    w_vor, w_vel, w_prs = vor.weight, vel.weight, prs.weight
    stokes = KFormSystem(w_vor.derivative @ vel + w_vor @ vor == 0, w_vel @ vor.derivative + w_vel.derivative @ prs == 0,
                         w_prs @ vel.derivative == 0, sorting=lambda f: f.order)
    # direct Poisson (u0, q1)
    u, q = KFormUnknown("u", O0), KFormUnknown("q", O1)
    w_u, w_q = u.weight, q.weight
    direct = KFormSystem(w_q @ q - w_q @ u.derivative == 0, w_u.derivative @ q == 0, sorting=lambda f: f.order)
    for nh, nv, p in ((4, 3, 3), (3, 3, 2), (5, 2, 4), (2, 2, 1)):
        mesh = unit_square_mesh(nh, nv, p, F.deform)
        for name, system in (("stokes", stokes), ("direct", direct)):
            tag = f"{name}_{nh}x{nv}_p{p}"
            r, e, d, v, rhs = coo(connect_elements(system.unknown_forms, mesh))
            data[tag + "_row"], data[tag + "_elem"], data[tag + "_dof"], data[tag + "_val"], data[tag + "_rhs"] = r, e, d, v, rhs
        # full constraint matrix of add_system_constraints: continuity, gauge on prs, strong BC on vel (1-form) / on u (0-form)
        leaf = list(mesh.get_leaf_indices())
        cache = FemCache(2)
        for name, system, gauge, bc in (("stokes", stokes, [(0.25, prs)], BoundaryCondition2DSteady(vel, mesh.boundary_indices, bc_vec)),
                                        ("direct", direct, [], BoundaryCondition2DSteady(u, mesh.boundary_indices, bc_u))):
            sizes = [system.unknown_forms.total_size(*mesh.get_leaf_orders(i)) for i in leaf]
            off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
            mat, vec = add_system_constraints(system, mesh, cache, gauge, [bc], leaf, off, None)
            m = mat.tocoo()
            order = np.lexsort((m.col, m.row))
            tag = f"full_{name}_{nh}x{nv}_p{p}"
            data[tag + "_row"], data[tag + "_col"], data[tag + "_val"], data[tag + "_rhs"] = m.row[order], m.col[order], m.data[order], vec
        data[f"boundary_{nh}x{nv}"] = np.array(mesh.boundary_indices, np.int64)
    np.savez_compressed(os.path.join(HERE, "constraints.npz"), **data)
    print("constraints.npz:", len(data), "arrays")


if __name__ == "__main__":
    main()
