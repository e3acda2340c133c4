"""Fixture for the weak (natural) boundary terms of the right-hand side (mfv2d_b200/structured.py: weak_bc_contributions).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_weak_bc.py
Writes tests/golden/weak_bc.npz: for a few conforming deformed meshes, FemCache(2), the dense ``(n, N_e)`` image of the weak
ElementConstraints returned by the reference's ``mesh_boundary_conditions`` (python/mfv2d/boundary.py:382-498, which calls
``_element_weak_boundary_condition`` :109-214 per boundary line) for
  * Stokes (vor0, vel1, prs2) with ``w_vor ^ bc_vec`` (tangential integral, 0-form weight) and ``w_vel ^ bc_u``
    (normal integral, 1-form weight), no strong conditions;
  * the same with a strong condition on vel on the first half of the boundary lines (those lines get no weak term for vel).
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
from mfv2d.boundary import BoundaryCondition2DSteady, mesh_boundary_conditions  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def bc_u(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y - 0.1)


def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)


def main():
    vor, vel, prs = KFormUnknown("vor", O0), KFormUnknown("vel", O1), KFormUnknown("prs", O2)
    w_vor, w_vel, w_prs = vor.weight, vel.weight, prs.weight
    stokes = KFormSystem(w_vor.derivative @ vel + w_vor @ vor == 0.5 * (w_vor ^ bc_vec),
                         w_vel @ vor.derivative + w_vel.derivative @ prs == -2.0 * (w_vel ^ bc_u), w_prs @ vel.derivative == 0,
                         sorting=lambda f: f.order)
    specs = stokes.unknown_forms
    data = {"form_orders": np.array([int(o) for o in specs.orders])}
    cache = FemCache(2)
    for nh, nv, p in ((4, 3, 3), (3, 3, 2), (2, 2, 1), (5, 2, 4)):
        mesh = unit_square_mesh(nh, nv, p, F.deform)
        leaf = list(mesh.get_leaf_indices())
        data[f"corners_{nh}x{nv}"] = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
        ne = specs.total_size(p, p)
        rights = [eq.right for eq in stokes.equations]
        bidx = np.array(mesh.boundary_indices)
        half = bidx[: bidx.size // 2]
        for tag, strong in (("none", [[] for _ in rights]),
                            ("half", [[BoundaryCondition2DSteady(vel, half, bc_vec)] if eq.weight.base_form == vel else [] for eq in stokes.equations])):
            _, weak = mesh_boundary_conditions(rights, specs, mesh, strong, cache)
            dense = np.zeros((len(leaf), ne))
            for ec in weak:
                np.add.at(dense[ec.i_e], np.asarray(ec.dofs, np.int64), np.asarray(ec.coeffs, np.float64))
            data[f"weak_{tag}_{nh}x{nv}_p{p}"] = dense
        data[f"half_lines_{nh}x{nv}"] = np.arange(bidx.size // 2)
    np.savez_compressed(os.path.join(HERE, "weak_bc.npz"), **data)
    for k, v in data.items():
        print(k, np.asarray(v).shape)


if __name__ == "__main__":
    main()
