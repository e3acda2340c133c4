"""Fixture for boundary conditions on NON-conforming quadtree meshes (mfv2d_b200/nonconforming.py: boundary_terms_forest).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_forest_bc.py
On the three hp-refined meshes of make_golden_forest.py (rebuilt here by the same calls) the strong and weak
ElementConstraints of the reference's ``mesh_boundary_conditions`` (python/mfv2d/boundary.py:382-498; leaves along a refined
boundary side are visited by the recursions of :152-172 and :264-289), FemCache(2):
  * Stokes layout (vor0, vel1, prs2): weak ``0.5 (w_vor ^ bc_vec)`` everywhere, strong ``vel = bc_vec`` on the first half of the
    boundary lines, weak ``-2 (w_vel ^ bc_u)`` on the others;
  * direct Poisson layout (u0, q1): strong ``u = bc_u`` on the whole boundary (first-come ownership of the mesh nodes; the node
    between two children of a refined side is constrained from both sides), weak ``b ^ bc_u`` in the q equation on every line.
Also the (root element, side) pair of every boundary line in ``mesh.boundary_indices`` order.  Writes tests/golden/forest_bc.npz.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, HERE)

from make_golden_forest import predicate  # noqa: E402  (loads the reference and patches compute_gll)
import fields as F  # noqa: E402
import mfv2d.boundary as _bnd  # noqa: E402
from mfv2d.boundary import BoundaryCondition2DSteady, mesh_boundary_conditions  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache, find_surface_boundary_id_line  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def bc_u(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y - 0.1)


def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)


def flat(ecs):
    leaf, dof, val = [], [], []
    for ec in ecs:
        for d, c in zip(ec.dofs, ec.coeffs):
            leaf.append(int(ec.i_e)), dof.append(int(d)), val.append(float(c))
    return np.array(leaf, np.int64), np.array(dof, np.int64), np.array(val, np.float64)


def main():
    vor, vel, prs = KFormUnknown("vor", O0), KFormUnknown("vel", O1), KFormUnknown("prs", O2)
    wv, wu, wp = vor.weight, vel.weight, prs.weight
    stokes = KFormSystem(wv.derivative @ vel + wv @ vor == 0.5 * (wv ^ bc_vec), wu @ vor.derivative + wu.derivative @ prs == -2.0 * (wu ^ bc_u),
                         wp @ vel.derivative == 0, sorting=lambda f: f.order)
    u0, g1 = KFormUnknown("u", O0), KFormUnknown("q", O1)
    a, b = u0.weight, g1.weight
    direct = KFormSystem(b @ g1 - b @ u0.derivative == b ^ bc_u, a.derivative @ g1 == 0, sorting=lambda f: f.order)
    meshes = {
        "a": unit_square_mesh(3, 3, 2, F.deform).split_depth_first(2, predicate, [0, 0, 2, 7, 3]),
        "b": unit_square_mesh(2, 2, 3, F.deform).split_depth_first(1, predicate, [0, 0, 3, 3, 2]),
        "c": unit_square_mesh(2, 3, 2, None).split_depth_first(3, predicate, [1, 0, 2, 4, 2]),
    }
    cache = FemCache(2)
    data = {}
    for name, mesh in meshes.items():
        bidx = np.array(mesh.boundary_indices)
        lines = []
        for i_boundary in bidx:
            dual_line = mesh.dual.get_line(int(i_boundary) + 1)
            id_surf = dual_line.begin if dual_line.begin else dual_line.end
            side = find_surface_boundary_id_line(mesh.primal.get_surface(id_surf), int(i_boundary))
            lines.append((id_surf.index, side.value))
        data[f"{name}_lines"] = np.array(lines, np.int64)
        half = bidx[: bidx.size // 2]
        for sname, system, strong in (
            ("stokes", stokes, lambda eq: [BoundaryCondition2DSteady(vel, half, bc_vec)] if eq.weight.base_form == vel else []),
            ("direct", direct, lambda eq: [BoundaryCondition2DSteady(u0, bidx, bc_u)] if eq.weight.base_form == u0 else []),
        ):
            s_bcs, w_bcs = mesh_boundary_conditions([eq.right for eq in system.equations], system.unknown_forms, mesh,
                                                    [strong(eq) for eq in system.equations], cache)
            tag = f"{name}_{sname}"
            data[tag + "_strong_leaf"], data[tag + "_strong_dof"], data[tag + "_strong_val"] = flat(s_bcs)
            data[tag + "_weak_leaf"], data[tag + "_weak_dof"], data[tag + "_weak_val"] = flat(w_bcs)
            # the whole constraint matrix and the weak additions as add_system_constraints assembles them (continuity.py:762-872):
            # continuity rows, a gauge row on the last form (Stokes: prs), strong rows; linear_vectors start from zero
            from mfv2d.continuity import add_system_constraints

            leaf_ids = list(mesh.get_leaf_indices())
            sizes = [system.unknown_forms.total_size(*mesh.get_leaf_orders(i)) for i in leaf_ids]
            off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint32)
            vectors = [np.zeros(n) for n in sizes]
            gauge = [(0.25, prs)] if sname == "stokes" else []
            bcs = [bc for eq in system.equations for bc in strong(eq)]
            mat, vec = add_system_constraints(system, mesh, cache, gauge, bcs, leaf_ids, off, vectors)
            m = mat.tocoo()
            o = np.lexsort((m.col, m.row))
            data[tag + "_full_row"], data[tag + "_full_col"], data[tag + "_full_val"], data[tag + "_full_rhs"] = (
                m.row[o].astype(np.int64), m.col[o].astype(np.int64), m.data[o], vec)
            data[tag + "_full_weak"] = np.concatenate(vectors)
            print(tag, "lines", bidx.size, "strong rows", data[tag + "_strong_val"].size, "weak entries", data[tag + "_weak_val"].size)
    np.savez_compressed(os.path.join(HERE, "forest_bc.npz"), **data)


if __name__ == "__main__":
    main()
