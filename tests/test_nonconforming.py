"""`-m "not gpu"`: the array-based constraint builder for non-conforming quadtree meshes (mfv2d_b200/nonconforming.py) against the
reference's own rows (tests/golden/forest.npz: connect_elements on hp-refined meshes with hanging sides, orders 2..8, systems with
0-, 1- and 2-form unknowns; tests/golden/make_golden_forest.py).  Row count, row order, elements and DoF indices must agree
exactly, and so do the coefficients (the hanging-side rows come out of inverted interpolation matrices of Lagrange values that
the library evaluates in the reference's operation order)."""
import numpy as np
import pytest
from conftest import golden

import fields as F
from mfv2d_b200 import nonconforming as NC


def bc_u(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y - 0.1)


def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)

FORMS = {"mixed": [2, 3], "direct": [1, 2], "stokes": [1, 2, 3]}


@pytest.mark.parametrize("system", sorted(FORMS))
@pytest.mark.parametrize("mesh", ["a", "b", "c"])
def test_forest_rows_match_reference(mesh, system):
    g = golden("forest.npz")
    nh, nv = (int(v) for v in g[f"{mesh}_dims"])
    forest = NC.QuadForest(nh, nv, g[f"{mesh}_children"], g[f"{mesh}_leaf"], g[f"{mesh}_orders"])
    row, col, val, n, offsets = NC.continuity_constraints_forest(forest, FORMS[system])
    tag = f"{mesh}_{system}"
    assert n == int(g[tag + "_n"])
    ref_col = offsets[g[tag + "_elem"]] + g[tag + "_dof"]
    assert np.array_equal(row, g[tag + "_row"]), "row structure"
    assert np.array_equal(col, ref_col), "columns (in the reference's order inside every row)"
    # bit for bit: same nodes (compute_gll), same Lagrange values (operation order of src/polynomials/lagrange.c), same numpy
    # inverse - although the hanging-side coefficients reach 4e8 on the orders-2..8 mesh
    assert np.array_equal(val, g[tag + "_val"])


@pytest.mark.parametrize("mesh,deformed", [("a", True), ("b", True), ("c", False)])
def test_forest_leaf_corners(mesh, deformed):
    import fields as F
    from mfv2d_b200 import structured as ST

    g = golden("forest.npz")
    nh, nv = (int(v) for v in g[f"{mesh}_dims"])
    forest = NC.QuadForest(nh, nv, g[f"{mesh}_children"], g[f"{mesh}_leaf"], g[f"{mesh}_orders"])
    got = forest.leaf_corners(ST.unit_square_corners(nh, nv, F.deform if deformed else None))
    assert np.max(np.abs(got - g[f"{mesh}_corners"])) <= 1e-15


def test_forest_reduces_to_the_structured_builder():
    """without any split element the forest builder gives the rows of structured.continuity_constraints"""
    from mfv2d_b200 import structured as ST

    nh, nv, p = 4, 3, 3
    n = nh * nv
    forest = NC.QuadForest(nh, nv, np.full((n, 4), -1), np.arange(n), np.full(n, p))
    for orders in ([2, 3], [1, 2, 3], [1, 2]):
        row, col, val, nrows, _ = NC.continuity_constraints_forest(forest, orders)
        o = np.lexsort((col, row))
        r2, c2, v2, n2 = ST.continuity_constraints(nh, nv, p, orders)
        assert nrows == n2 and np.array_equal(row[o], r2) and np.array_equal(col[o], c2) and np.array_equal(val[o], v2)


def test_forest_input_checks():
    with pytest.raises(ValueError):
        NC.QuadForest(1, 1, np.full((1, 4), -1), [0], [2, 3])
    with pytest.raises(ValueError):
        NC.QuadForest(1, 1, np.array([[1, 2, 3, 4], [-1] * 4, [-1] * 4, [-1] * 4, [-1] * 4]), [0, 1, 2, 3], [2, 2, 2, 2])


@pytest.mark.parametrize("mesh", ["a", "b", "c"])
def test_boundary_terms_forest(mesh):
    """Strong rows and weak right-hand-side terms on the boundary of the hp-refined meshes against the reference's
    mesh_boundary_conditions (tests/golden/forest_bc.npz): same leaves, same DoFs, in the same order; values to rounding."""
    from mfv2d_b200 import structured as ST

    g, gb = golden("forest.npz"), golden("forest_bc.npz")
    nh, nv = (int(v) for v in g[f"{mesh}_dims"])
    forest = NC.QuadForest(nh, nv, g[f"{mesh}_children"], g[f"{mesh}_leaf"], g[f"{mesh}_orders"])
    assert [tuple(t) for t in gb[f"{mesh}_lines"]] == ST.boundary_lines(nh, nv)
    root_corners = ST.unit_square_corners(nh, nv, None if mesh == "c" else F.deform)
    n_lines = 2 * (nh + nv)
    cases = {
        "stokes": ([1, 2, 3], [(1, bc_vec, range(n_lines // 2))], [(0, [(0.5, bc_vec)]), (1, [(-2.0, bc_u)])]),
        "direct": ([1, 2], [(0, bc_u, None)], [(1, [(1.0, bc_u)])]),
    }
    for name, (orders, strong, weak) in cases.items():
        (leaf, dof, val), weak_vec = NC.boundary_terms_forest(forest, orders, root_corners, 2, strong, weak)
        tag = f"{mesh}_{name}"
        assert np.array_equal(leaf, gb[tag + "_strong_leaf"]) and np.array_equal(dof, gb[tag + "_strong_dof"])
        assert np.max(np.abs(val - gb[tag + "_strong_val"])) <= 1e-14 * max(1.0, np.max(np.abs(val)))
        sizes = np.array([int(ST.form_sizes(orders, int(p)).sum()) for p in forest.leaf_orders])
        offsets = np.concatenate([[0], np.cumsum(sizes)])
        ref = np.zeros(offsets[-1])
        np.add.at(ref, offsets[gb[tag + "_weak_leaf"]] + gb[tag + "_weak_dof"], gb[tag + "_weak_val"])
        assert np.max(np.abs(weak_vec - ref)) <= 1e-14 * np.max(np.abs(ref)), np.max(np.abs(weak_vec - ref))
        # the whole matrix of add_system_constraints (continuity + gauge + strong rows) and its weak additions
        gauge = [(0.25, 2)] if name == "stokes" else []
        row, col, val, rhs, n, offsets2, weak2 = NC.system_constraints_forest(forest, orders, root_corners, 2, gauge, strong, weak)
        assert n == gb[tag + "_full_rhs"].size and np.array_equal(offsets2, offsets)
        keep = gb[tag + "_full_val"] != 0  # scipy keeps explicit zeros of cancelled duplicates; the builder may or may not produce them
        keep2 = val != 0
        assert np.array_equal(row[keep2], gb[tag + "_full_row"][keep]) and np.array_equal(col[keep2], gb[tag + "_full_col"][keep])
        assert np.max(np.abs(val[keep2] - gb[tag + "_full_val"][keep])) <= 1e-12
        assert np.max(np.abs(rhs - gb[tag + "_full_rhs"])) <= 1e-14 * max(1.0, np.max(np.abs(rhs)))
        assert np.max(np.abs(weak2 - gb[tag + "_full_weak"])) <= 1e-14 * np.max(np.abs(ref))
