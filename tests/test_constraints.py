"""`-m "not gpu"`: the array-based constraint builder (mfv2d_b200/structured.py) against the reference's own rows, bit for bit.

Fixtures: tests/golden/constraints.npz (connect_elements and add_system_constraints of the reference on conforming meshes
with 0-, 1- and 2-form unknowns, a gauge row and strong boundary conditions; tests/golden/make_golden_constraints.py) and the
con_row / con_elem / con_dof / con_val arrays of tests/golden/solve_*.npz."""
import numpy as np
import pytest
from conftest import golden

import fields as F
from mfv2d_b200 import structured as ST
from oracle import mfv2d_oracle as O

FORMS = {"stokes": [1, 2, 3], "direct": [1, 2]}
MESHES = ((4, 3, 3), (3, 3, 2), (5, 2, 4), (2, 2, 1))


def bc_u(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y - 0.1)


def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)


@pytest.mark.parametrize("nh,nv,p", MESHES)
@pytest.mark.parametrize("name", sorted(FORMS))
def test_continuity_rows_bit_exact(name, nh, nv, p):
    g = golden("constraints.npz")
    tag = f"{name}_{nh}x{nv}_p{p}"
    fo = FORMS[name]
    ne = int(ST.form_sizes(fo, p).sum())
    col = g[tag + "_elem"] * ne + g[tag + "_dof"]
    o = np.lexsort((col, g[tag + "_row"]))
    row2, col2, val2, n = ST.continuity_constraints(nh, nv, p, fo)
    assert n == g[tag + "_rhs"].size
    assert np.array_equal(row2, g[tag + "_row"][o]) and np.array_equal(col2, col[o]) and np.array_equal(val2, g[tag + "_val"][o])


@pytest.mark.parametrize("nh,nv,p", MESHES)
@pytest.mark.parametrize("name", sorted(FORMS))
def test_system_constraints_bit_exact(name, nh, nv, p):
    """continuity + gauge row + strong boundary conditions = the constraint matrix of add_system_constraints"""
    g = golden("constraints.npz")
    tag = f"full_{name}_{nh}x{nv}_p{p}"
    corners = ST.unit_square_corners(nh, nv, F.deform)
    roots, _ = O.gll(p + 1)
    rn, rw = O.gll(p + 3)  # FemCache(2): integration order p + 2
    gauge, bcs = ([(0.25, 2)], [(1, bc_vec)]) if name == "stokes" else ([], [(0, bc_u)])
    row, col, val, rhs, n = ST.system_constraints(nh, nv, p, FORMS[name], gauge, bcs, corners, roots, rn, rw)
    assert n == g[tag + "_rhs"].size
    assert np.array_equal(row, g[tag + "_row"]) and np.array_equal(col, g[tag + "_col"]) and np.array_equal(val, g[tag + "_val"])
    assert np.max(np.abs(rhs - g[tag + "_rhs"])) <= 1e-15 * max(1.0, np.max(np.abs(rhs)))
    assert np.array_equal(np.array([e * 0 + s for e, s in ST.boundary_lines(nh, nv)]).shape, (2 * (nh + nv),))


@pytest.mark.parametrize("fix,nh,nv", [("solve_10x10_p3", 10, 10), ("solve_3x4_p4", 3, 4), ("solve_5x2_p5_def", 5, 2), ("solve_6x6_p3_def", 6, 6)])
def test_solve_fixture_constraints_bit_exact(fix, nh, nv):
    g = golden(fix + ".npz")
    p = int(g["order"])
    fo = [2, 3]
    ne = int(ST.form_sizes(fo, p).sum())
    col = g["con_elem"].astype(np.int64) * ne + g["con_dof"].astype(np.int64)
    o = np.lexsort((col, g["con_row"]))
    row2, col2, val2, n = ST.continuity_constraints_1form(nh, nv, p, fo)
    assert n == g["con_rhs"].size
    assert np.array_equal(row2, g["con_row"][o].astype(np.int64)) and np.array_equal(col2, col[o]) and np.array_equal(val2, g["con_val"][o])


@pytest.mark.parametrize("nh,nv,p", MESHES)
def test_weak_boundary_terms(nh, nv, p):
    """weak_bc_contributions against the weak ElementConstraints of the reference's mesh_boundary_conditions
    (tests/golden/weak_bc.npz): tangential integrals for the 0-form weight, normal integrals for the 1-form weight, lines
    with a strong condition skipped.  The 1-D tables come from the oracle here; the sums are the reference's, so the
    agreement is to rounding."""
    g = golden("weak_bc.npz")
    fo = [int(o) for o in g["form_orders"]]
    corners = g[f"corners_{nh}x{nv}"]
    assert np.array_equal(corners, ST.unit_square_corners(nh, nv, F.deform))
    b = O.basis(p, p + 3)  # FemCache(2): p + 3 points
    for tag, skip in (("none", ()), ("half", g[f"half_lines_{nh}x{nv}"])):
        ref = g[f"weak_{tag}_{nh}x{nv}_p{p}"]
        out = ST.weak_bc_contributions(nh, nv, p, fo, 0, [(0.5, bc_vec)], corners, b.node, b.edge, b.pnts, b.wgts)
        ST.weak_bc_contributions(nh, nv, p, fo, 1, [(-2.0, bc_u)], corners, b.node, b.edge, b.pnts, b.wgts, skip_lines=skip, out=out)
        assert np.max(np.abs(out - ref)) <= 1e-14 * np.max(np.abs(ref)), (tag, np.max(np.abs(out - ref)))
    with pytest.raises(ValueError):
        ST.weak_bc_contributions(nh, nv, p, fo, 2, [(1.0, bc_u)], corners, b.node, b.edge, b.pnts, b.wgts)
