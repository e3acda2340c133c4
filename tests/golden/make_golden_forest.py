"""Fixture for the array-based constraint builder on NON-conforming quadtree meshes (mfv2d_b200/nonconforming.py).

Run in the build container (needs /root/reference):  python tests/golden/make_golden_forest.py
Three hp-refined meshes made by the reference itself (`unit_square_mesh` + `split_depth_first`, child orders cycling through a
range) and, for the form layouts of the example systems (mixed Poisson: 1-form + 2-form; direct Poisson: 0-form + 1-form;
Stokes: 0-, 1- and 2-form), the rows of the reference's `connect_elements` (python/mfv2d/continuity.py:610-759: intra-element
rows of split elements, edge rows with hanging sides, corner chains).  The tree is stored as arrays: children of every element
(-1 for leaves), the element ids of the leaves in leaf order, the order of every leaf.  Writes tests/golden/forest.npz.
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
import mfv2d.continuity as _cont  # noqa: E402
from mfv2d.continuity import connect_elements  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

# see make_golden_hp.py: the plain-gcc build of the reference needs compute_gll's defaults spelled out
_gll = _cont.compute_gll
_cont.compute_gll = lambda order: _gll(order, 10, 1e-15)

O0, O1, O2 = UnknownFormOrder.FORM_ORDER_0, UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def predicate(m, ie, counter):
    """split `counter[4]` of every four visited elements; the children get orders counter[2] + (running index mod counter[3])"""
    counter[0] += 1
    if (counter[0] & 3) < counter[4]:
        out = []
        for _ in range(4):
            o = counter[2] + counter[1] % counter[3]
            counter[1] += 1
            out.append((o, o))
        return tuple(out)
    return None


def systems():
    u2, q1 = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v, w = u2.weight, q1.weight
    mixed = KFormSystem(w @ q1 + w.derivative @ u2 == 0, v @ q1.derivative == 0, sorting=lambda f: f.order)
    u0, g1 = KFormUnknown("u", O0), KFormUnknown("q", O1)
    a, b = u0.weight, g1.weight
    direct = KFormSystem(b @ g1 - b @ u0.derivative == 0, a.derivative @ g1 == 0, sorting=lambda f: f.order)
    vor, vel, prs = KFormUnknown("vor", O0), KFormUnknown("vel", O1), KFormUnknown("prs", O2)
    wv, wu, wp = vor.weight, vel.weight, prs.weight
    stokes = KFormSystem(wv.derivative @ vel + wv @ vor == 0, wu @ vor.derivative + wu.derivative @ prs == 0, wp @ vel.derivative == 0,
                         sorting=lambda f: f.order)
    return {"mixed": mixed, "direct": direct, "stokes": stokes}


def main():
    data = {}
    meshes = {
        "a": unit_square_mesh(3, 3, 2, F.deform).split_depth_first(2, predicate, [0, 0, 2, 7, 3]),
        "b": unit_square_mesh(2, 2, 3, F.deform).split_depth_first(1, predicate, [0, 0, 3, 3, 2]),
        "c": unit_square_mesh(2, 3, 2, None).split_depth_first(3, predicate, [1, 0, 2, 4, 2]),
    }
    dims = {"a": (3, 3), "b": (2, 2), "c": (2, 3)}
    for name, mesh in meshes.items():
        n_el = mesh.element_count
        children = np.full((n_el, 4), -1, np.int32)
        for e in range(n_el):
            c = mesh.get_element_children(e)
            if c is not None:
                children[e] = c
        leaf = np.array(list(mesh.get_leaf_indices()), np.int32)
        orders = np.array([mesh.get_leaf_orders(int(i))[0] for i in leaf], np.int32)
        assert all(mesh.get_leaf_orders(int(i))[0] == mesh.get_leaf_orders(int(i))[1] for i in leaf)
        data[f"{name}_dims"] = np.array(dims[name], np.int32)
        data[f"{name}_children"], data[f"{name}_leaf"], data[f"{name}_orders"] = children, leaf, orders
        data[f"{name}_corners"] = np.array([np.array(mesh.get_leaf_corners(int(i)), np.float64) for i in leaf])
        for sname, system in systems().items():
            cons = connect_elements(system.unknown_forms, mesh)
            rows, elems, dofs, vals = [], [], [], []
            for ic, c in enumerate(cons):
                for ec in c.element_constraints:
                    for d, cf in zip(ec.dofs, ec.coeffs):
                        rows.append(ic), elems.append(int(ec.i_e)), dofs.append(int(d)), vals.append(float(cf))
            tag = f"{name}_{sname}"
            data[tag + "_row"], data[tag + "_elem"], data[tag + "_dof"], data[tag + "_val"] = (np.array(rows, np.int64), np.array(elems, np.int64),
                                                                                              np.array(dofs, np.int64), np.array(vals, np.float64))
            data[tag + "_n"] = np.array(len(cons))
            per_row = np.bincount(np.array(rows))
            print(tag, "elements", n_el, "leaves", leaf.size, "orders", sorted(set(orders.tolist())), "rows", len(cons), "entries", len(rows),
                  "max entries per row", per_row.max())
    np.savez_compressed(os.path.join(HERE, "forest.npz"), **data)


if __name__ == "__main__":
    main()
