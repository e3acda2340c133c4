"""Parity checks of the mfv2d_b200 API against the oracle (numpy restatement pinned to the reference) and
against the golden fixtures produced by the reference itself.  The same functions run

* under ``-m gpu`` on the B200 through libmfv2d_cuda.so (tests/test_gpu_parity.py) - the parity tests proper;
* under ``-m "not gpu"`` through the host emulation of the same sources (tests/test_emu_kernels.py), which
  checks the planner / layouts / index arithmetic in the GPU-less container.

Tolerances: fp64 element matrices 1e-12 relative (north_star), inverses 1e-10, solutions 1e-8 against the
reference's own iterative solution (both stop at a residual tolerance), indices bit-exact.
"""
from __future__ import annotations

import json
import os

import numpy as np
import pytest
from conftest import golden, load_systems, relerr

import fields as F
from oracle import mfv2d_oracle as O

SYSTEMS = load_systems()
TOL_MAT = 1e-12
TOL_INV = 1e-10


def make_basis(M, p, k):
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + k))
    return M.Basis2D(b1, b1)


def specs_of(M, s):
    return M._ElementFormSpecification(*zip(s["form_names"], s["form_orders"]))


def check_tables(M, cases=((1, 2), (3, 3), (4, 2), (5, 0))):
    g = golden("element_stokes.npz")
    for p, k in cases:
        tag = f"p{p}k{k}q4"
        rule = M.IntegrationRule1D(p + k)
        b1 = M.Basis1D(p, rule)
        assert np.array_equal(rule.nodes, g[tag + "_pnts"])  # GLL iteration is reproduced bit for bit
        assert np.array_equal(rule.weights, g[tag + "_wgts"])
        assert relerr(b1.node, g[tag + "_node"]) < 1e-14
        assert relerr(b1.edge, g[tag + "_edge"]) < 1e-13
    for n in range(0, 13):
        x, w = M.compute_gll(n)
        xo, wo = O.compute_gll(n)
        assert np.array_equal(x, xo) and np.array_equal(w, wo)
        assert abs(w.sum() - 2.0) < 1e-14
    rng = np.random.default_rng(3)
    roots = np.sort(rng.uniform(-1, 1, 7))
    pos = rng.uniform(-1, 1, (5, 4))
    assert relerr(M.lagrange1d(roots, pos), O.lagrange1d(roots, pos)) < 1e-12
    assert relerr(M.dlagrange1d(roots, pos), O.dlagrange1d(roots, pos)) < 1e-11
    assert relerr(M.compute_legendre(9, pos), O.legendre(9, pos)) < 1e-14
    assert relerr(M.compute_legendre_device(9, pos), O.legendre(9, pos)) < 1e-14


def check_masses_golden(M, cases=((1, 2), (3, 3), (4, 2), (5, 0))):
    """Mass matrices and inverses against the reference's own output on the skewed quad."""
    g = golden("element_stokes.npz")
    for p, k in cases:
        tag = f"p{p}k{k}q4"
        batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(F.QUADS[4:5]))
        for fo in (1, 2, 3):
            assert relerr(batch.mass(fo)[0], g[tag + f"_m{fo - 1}"]) < TOL_MAT
            assert relerr(batch.mass(fo, True)[0], g[tag + f"_m{fo - 1}i"]) < TOL_INV
        jac = batch.jacobian()[0]
        ob = O.basis(p, p + k + 1)
        for a, b in zip(jac, O.jacobian(F.QUADS[4], ob.pnts)):
            assert relerr(a, b) < 1e-15


def check_masses_oracle(M, cases, n_random=64, seed=0):
    """Masses on the test quads + randomly perturbed unit quads (tests/test_basis.py:83-87 distribution)."""
    rng = np.random.default_rng(seed)
    unit = np.array([(-1, -1), (1, -1), (1, 1), (-1, 1)], float)
    corners = np.concatenate([F.QUADS, unit + rng.uniform(-0.1, 0.1, (n_random, 4, 2))])
    for p, k in cases:
        ob = O.basis(p, p + k + 1)
        batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(corners))
        for fo in (1, 2, 3):
            m = batch.mass(fo)
            ref = O.mass(ob, corners, fo)
            assert relerr(m, ref) < TOL_MAT, (p, k, fo)
            mi = batch.mass(fo, True)
            eye = np.einsum("eij,ejk->eik", m, mi)
            assert np.max(np.abs(eye - np.eye(m.shape[1]))) < 1e-9, (p, k, fo)  # tests/test_basis.py:80-100


def check_system_golden(M, name, cases=((1, 2), (3, 3), (4, 2), (5, 0))):
    """Element matrices and vectors of one example system against the reference's outputs."""
    s = SYSTEMS[name]
    g = golden(f"element_{name}.npz")
    specs = specs_of(M, s)
    for p, k in cases:
        batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(F.QUADS))
        n = specs.total_size(p, p)
        dofs = np.stack([g[f"p{p}k{k}q{iq}_dofs"] for iq in range(len(F.QUADS))])
        assert dofs.shape == (len(F.QUADS), n)
        mats = M.compute_element_matrices(specs, s["lhs"], batch, dofs)
        vecs = M.compute_element_vectors(specs, s["lhs"], batch, dofs)
        for iq in range(len(F.QUADS)):
            tag = f"p{p}k{k}q{iq}"
            assert relerr(mats[iq], g[tag + "_lhs"]) < TOL_MAT, (name, tag)
            assert relerr(vecs[iq], g[tag + "_lhs_vec"]) < TOL_MAT, (name, tag)
        if s["rhs"] is not None:
            vecs = M.compute_element_vectors(specs, s["rhs"], batch, dofs)
            for iq in range(len(F.QUADS)):
                assert relerr(vecs[iq], g[f"p{p}k{k}q{iq}_rhs_vec"]) < TOL_MAT, (name, iq)
        if s["nonlin"] is not None:
            mats = M.compute_element_matrices(specs, s["nonlin"], batch, dofs)
            for iq in range(len(F.QUADS)):
                assert relerr(mats[iq], g[f"p{p}k{k}q{iq}_nonlin"]) < TOL_MAT, (name, iq)
        # per-element entry points (batch of one), as the reference's Python loops call them
        sp = M.ElementFemSpace2D(make_basis(M, p, k), np.ascontiguousarray(F.QUADS[4]))
        tag = f"p{p}k{k}q4"
        assert relerr(M.compute_element_matrix(specs, s["lhs"], sp, dofs[4]), g[tag + "_lhs"]) < TOL_MAT
        assert relerr(M.compute_element_vector(specs, s["lhs"], sp, dofs[4]), g[tag + "_lhs_vec"]) < TOL_MAT


def check_matrix_vector_consistency(M, name, p, k, n=16, seed=5):
    """tests/test_explicit.py: ``compute_element_matrix @ x == compute_element_vector(x)`` for linear systems."""
    s = SYSTEMS[name]
    specs = specs_of(M, s)
    rng = np.random.default_rng(seed)
    unit = np.array([(-1, -1), (1, -1), (1, 1), (-1, 1)], float)
    corners = np.ascontiguousarray(unit + rng.uniform(-0.2, 0.2, (n, 4, 2)))
    batch = M.ElementBatch(make_basis(M, p, k), corners)
    ne = specs.total_size(p, p)
    x = rng.standard_normal((n, ne))
    mats = M.compute_element_matrices(specs, s["lhs"], batch, x)
    vecs = M.compute_element_vectors(specs, s["lhs"], batch, x)
    assert relerr(np.einsum("eij,ej->ei", mats, x), vecs) < 1e-12


def check_mixed_order_buckets(M, orders=(2, 3, 5, 8), per_order=5, name="mixed_poisson", k=3, seed=9):
    """BASELINE.json configs[3] (h/p-refined mesh, mixed orders 2-8): the leaf elements are bucketed by order, one
    ElementBatch per (p, q) bucket, and the matrices / matrix-free vectors are scattered back to mesh order; every
    element is compared with the oracle's restatement of the reference at its own order."""
    s = SYSTEMS[name]
    specs = specs_of(M, s)
    rng = np.random.default_rng(seed)
    unit = np.array([(-1, -1), (1, -1), (1, 1), (-1, 1)], float)
    elem_orders = rng.permutation(np.repeat(np.array(orders), per_order))
    corners = unit + rng.uniform(-0.25, 0.25, (elem_orders.size, 4, 2))
    mats, vecs, xs = [None] * elem_orders.size, [None] * elem_orders.size, [None] * elem_orders.size
    for p in sorted(set(int(o) for o in elem_orders)):
        idx = np.flatnonzero(elem_orders == p)
        batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(corners[idx]))
        ne = specs.total_size(p, p)
        x = rng.standard_normal((idx.size, ne))
        m = M.compute_element_matrices(specs, s["lhs"], batch, x)
        v = M.compute_element_vectors(specs, s["lhs"], batch, x)
        assert m.shape == (idx.size, ne, ne)
        for j, e in enumerate(idx):
            mats[e], vecs[e], xs[e] = m[j], v[j], x[j]
    blocks, fs = O.decode_expressions(s["lhs"], s["form_names"])
    for e, p in enumerate(elem_orders):
        ob = O.basis(int(p), int(p) + k + 1)
        ref = O.element_matrix(ob, corners[e], s["form_orders"], blocks, fs)
        assert relerr(mats[e], ref) < TOL_MAT, (e, p)
        assert relerr(vecs[e], ref @ xs[e]) < TOL_MAT, (e, p)


def _dual_f_scalar(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y) + 0.5 * x * y


def _dual_f_vector(x, y):
    return np.stack((np.cos(x) * y + 1.0, np.sin(0.5 * y) - 0.25 * x * x), axis=-1)


def check_dual_dofs(M, cases=((1, 2), (3, 3), (4, 2), (6, 3))):
    """element_dual_dofs (python/mfv2d/mimetic2d.py:1003-1123): against the reference's own output
    (tests/golden/dual_dofs.npz, same functions as tests/golden/make_golden_dual.py) and the oracle on all quads."""
    g = golden("dual_dofs.npz")
    for p, k in cases:
        basis = make_basis(M, p, k)
        batch = M.ElementBatch(basis, np.ascontiguousarray(F.QUADS))
        ob = O.basis(p, p + k + 1)
        for fo, fn in ((1, _dual_f_scalar), (2, _dual_f_vector), (3, _dual_f_scalar)):
            d = batch.dual_dofs(fo, fn)
            assert d.shape == (len(F.QUADS), batch.form_size(fo))
            for iq in (2, 4, 5):
                ref = g[f"p{p}k{k}q{iq}_f{fo}"]
                assert relerr(d[iq], ref) < TOL_MAT, (p, k, iq, fo, relerr(d[iq], ref))
            for iq, quad in enumerate(F.QUADS):
                x, y = O.physical_coordinates(quad, ob.pnts)
                assert relerr(d[iq], O.dual_dofs(ob, quad, fo, fn(x, y))) < TOL_MAT, (p, k, iq, fo)
        sp = M.ElementFemSpace2D(basis, np.ascontiguousarray(F.QUADS[4]))
        assert relerr(M.element_dual_dofs(2, sp, _dual_f_vector), g[f"p{p}k{k}q4_f2"]) < TOL_MAT
    # compute_element_dual_from_primal / _primal_from_dual (solve_system.py:302-351): block-diagonal mass applies
    specs3 = M._ElementFormSpecification(("a", 1), ("b", 2), ("c", 3))
    batch3 = M.ElementBatch(make_basis(M, 3, 3), np.ascontiguousarray(F.QUADS))
    rng3 = np.random.default_rng(8)
    prim = rng3.standard_normal((len(F.QUADS), specs3.total_size(3, 3)))
    ref_dual = np.concatenate([np.einsum("eij,ej->ei", batch3.mass(fo), prim[:, o:o + batch3.form_size(fo)])
                               for fo, o in zip((1, 2, 3), (0, batch3.form_size(1), batch3.form_size(1) + batch3.form_size(2)))], axis=1)
    dual = M.compute_element_dual_from_primal(specs3, prim, batch3)
    assert relerr(dual, ref_dual) < TOL_MAT
    assert relerr(M.compute_element_primal_from_dual(specs3, dual, batch3), prim) < 1e-9
    sp3 = M.ElementFemSpace2D(make_basis(M, 3, 3), np.ascontiguousarray(F.QUADS[4]))
    assert relerr(M.compute_element_dual_from_primal(specs3, prim[4], sp3), ref_dual[4]) < TOL_MAT
    # consistency with the mass matrix: the dual dofs of a reconstructed field are M times its primal dofs (1- and
    # 2-forms; the reference's 0-form reconstruction carries an extra det, integrating_fields.c:238-262)
    p, k = 3, 3
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(F.QUADS))
    rng = np.random.default_rng(4)
    for fo in (2, 3):
        c = rng.standard_normal((len(F.QUADS), batch.form_size(fo)))
        vals = batch.reconstruct(fo, c)
        assert relerr(batch.dual_dofs(fo, vals), np.einsum("eij,ej->ei", batch.mass(fo), c)) < 1e-11, fo


def check_incidence(M, orders=(1, 2, 3, 5)):
    """tests/test_apply_incidence-c.py: all 8 incidence kernels against explicit matrices."""
    rng = np.random.default_rng(11)
    for p in orders:
        for t in range(4):
            e = O.incidence(t, p)
            m = rng.standard_normal((e.shape[1], 5))
            assert np.array_equal(M.check_incidence(m, p, t % 2, t >= 2, False), e @ m) or relerr(
                M.check_incidence(m, p, t % 2, t >= 2, False), e @ m) < 1e-15
            mr = rng.standard_normal((6, e.shape[0]))
            assert relerr(M.check_incidence(mr, p, t % 2, t >= 2, True), mr @ e) < 1e-15
    # structure must be bit-exact: apply to identity and compare the integer matrices
    for p in orders:
        for t in range(4):
            e = O.incidence(t, p)
            got = M.check_incidence(np.eye(e.shape[1]), p, t % 2, t >= 2, False)
            assert np.array_equal(got, e)


def check_dense(M):
    """tests/test_linalg_inverse.py: inverse and solve against numpy.linalg up to 100 x 100."""
    rng = np.random.default_rng(2)
    for n in (1, 2, 7, 30, 100):
        a = rng.standard_normal((n, n)) + 0.1 * n * np.eye(n)
        assert relerr(M._compute_matrix_inverse(a), np.linalg.inv(a)) < 1e-10
        b = rng.standard_normal((n, 3))
        assert relerr(M._solve_linear_system(a, b), np.linalg.solve(a, b)) < 1e-10
        v = rng.standard_normal(n)
        assert relerr(M._solve_linear_system(a, v), np.linalg.solve(a, v)) < 1e-10
    # a matrix that needs pivoting, and a singular one
    a = np.array([[0.0, 2.0], [3.0, 1.0]])
    assert relerr(M._compute_matrix_inverse(a), np.linalg.inv(a)) < 1e-14
    try:
        M._compute_matrix_inverse(np.zeros((3, 3)))
    except RuntimeError as e:
        assert "MFV2D_PIVOT_FAILED" in str(e)
    else:
        raise AssertionError("singular matrix must raise MFV2D_PIVOT_FAILED")


def check_empty_and_unconstrained(M, SV):
    """Edge cases: a batch without elements, a system without constraint rows (every solver reduces to the block solve), a
    system without elements, an empty vector."""
    from mfv2d_b200 import callers as CL

    s = SYSTEMS["mixed_poisson_sorted"]
    specs = specs_of(M, s)
    p, k = 3, 2
    ne = specs.total_size(p, p)
    b0 = M.ElementBatch(make_basis(M, p, k), np.zeros((0, 4, 2)))
    assert M.compute_element_matrices(specs, s["lhs"], b0).shape == (0, ne, ne)
    assert M.compute_element_vectors(specs, s["lhs"], b0, np.zeros((0, ne))).shape == (0, ne)
    assert b0.mass(2).shape == (0, 2 * p * (p + 1), 2 * p * (p + 1))
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(F.QUADS[:3]))
    mats = M.compute_element_matrices(specs, s["lhs"], batch)
    none = (np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0))
    sysm = SV.LinearSystem(mats, none, 0)
    rhs = np.random.default_rng(8).standard_normal(3 * ne)
    ref = np.concatenate([np.linalg.solve(mats[e], rhs[e * ne:(e + 1) * ne]) for e in range(3)])
    conv = SV.ConvergenceSettings(10, 1e-12, 1e-12)
    for solve in (SV.solve_schur_iterative, SV.solve_schur_bicgstab, SV.solve_pcg_iterative):
        x, lam, _res, it = solve(sysm, rhs, np.zeros(0), conv)
        assert lam.shape == (0,) and it == 0 and relerr(x, ref) < 1e-11, solve.__name__
    assert sysm.apply_trace(rhs).shape == (0,) and not np.any(sysm.apply_trace_transpose(np.zeros(0)))
    empty = SV.LinearSystem(np.zeros((0, ne, ne)), none, 0)
    assert empty.n_dense == 0 and empty.apply_diagonal(np.zeros(0)).shape == (0,)
    assert CL.vec_absmax(np.zeros(0), batch.ctx) == 0.0
    assert relerr(M.inverse_definite(np.array([[4.0]])), np.array([[0.25]])) < 1e-15
    # buffers of the wrong size never reach the library (it reads n_dense / n_trace entries)
    one = SV.LinearSystem(mats, (np.array([0, 0]), np.array([0, ne]), np.array([1.0, -1.0])), 1)
    for solve in (SV.solve_schur_iterative, SV.solve_schur_bicgstab, SV.solve_pcg_iterative):
        for bad in ((np.ones(5), np.zeros(1)), (rhs, np.zeros(3))):
            with pytest.raises(ValueError):
                solve(one, bad[0], bad[1], conv)
    with pytest.raises(ValueError):
        SV.vec_axpby(1.0, np.ones(4), 1.0, np.ones(3))
    with pytest.raises(ValueError):
        SV.vec_dot(np.ones(4), np.ones(3))
    with pytest.raises(ValueError):
        one.apply_diagonal(np.ones(3))
    adv = SYSTEMS["adv_dif"]
    with pytest.raises(ValueError):  # caller-evaluated fields of the wrong size
        M.compute_element_matrices(specs_of(M, adv), adv["lhs"], batch, None, np.zeros((1, 3, 2, 2, 2)))


def check_inverse_definite(M, sizes=(5, 64, 100, 333)):
    """The large-matrix inverse behind the preconditioner's exact level (blocked Gauss-Jordan without pivoting from 64 x 64 on,
    one pivot per pass below) against numpy on SPD and on non-symmetric diagonally dominant matrices, sizes that are not
    multiples of the block or of the update tile; a singular matrix must be reported."""
    rng = np.random.default_rng(21)
    for n in sizes:
        g = rng.standard_normal((n, n))
        for a in (g @ g.T + n * np.eye(n), g + 2.0 * n * np.eye(n)):
            inv = M.inverse_definite(a)
            assert np.max(np.abs(inv @ a - np.eye(n))) < 1e-11, (n, np.max(np.abs(inv @ a - np.eye(n))))
            assert relerr(inv, np.linalg.inv(a)) < 1e-11
    z = np.zeros((70, 70))
    with pytest.raises(Exception) as info:
        M.inverse_definite(z)
    assert "PIVOT" in str(info.value).upper()


def check_fields(M, p=3, k=3):
    """tests/test_compute_fields.py: callable and DoF-based fields at the quadrature points."""
    rng = np.random.default_rng(4)
    ob = O.basis(p, p + k + 1)
    specs = M._ElementFormSpecification(("a", 1), ("b", 2), ("c", 3))
    quad = np.ascontiguousarray(F.QUADS[4])
    sp = M.ElementFemSpace2D(make_basis(M, p, k), quad)
    dofs = rng.standard_normal(specs.total_size(p, p))
    offs = specs.form_offsets(p, p)
    out = M.compute_integrating_fields(sp, specs, (1, 2, 3, 2), (0, 1, 2, F.adv_field), dofs)
    for i, fo in enumerate((1, 2, 3)):
        ref = O.reconstruct_field(ob, quad, fo, dofs[offs[i] : offs[i + 1]])
        assert relerr(out[i], ref) < 1e-13
    x, y = O.physical_coordinates(quad, ob.pnts)
    assert relerr(out[3], F.adv_field(x, y)) < 1e-14


def check_interior_products(M, p=4, k=2):
    rng = np.random.default_rng(8)
    ob = O.basis(p, p + k + 1)
    corners = np.ascontiguousarray(F.QUADS)
    batch = M.ElementBatch(make_basis(M, p, k), corners)
    fld = rng.standard_normal((len(corners), ob.q, ob.q, 2))
    for tr in (False, True):
        ref = np.stack([O.interprod_node_edge(ob, c, f, tr) for c, f in zip(corners, fld)])
        assert relerr(batch.interior_product(2, fld, tr), ref) < TOL_MAT
        ref = np.stack([O.interprod_edge_surf(ob, c, f, tr) for c, f in zip(corners, fld)])
        assert relerr(batch.interior_product(3, fld, tr), ref) < TOL_MAT


def check_bytecode_errors(M):
    """Programs the reference rejects must be rejected with the same error code."""
    specs = M._ElementFormSpecification(("u", 1), ("q", 2))
    sp = M.ElementFemSpace2D(make_basis(M, 2, 1), np.ascontiguousarray(F.QUADS[0]))

    def run(block):
        return M.compute_element_matrix(specs, ((block, None), (None, None)), sp)

    for block, code in (
        (((3, 1, False), (3, 2, False)), "MFV2D_FAILED_ALLOC"),  # incidence o incidence (Q1)
        (((3, 1, True),), "MFV2D_FAILED_ALLOC"),  # bare incidence (Q1)
        (((2, 2, False),), "MFV2D_DIMS_MISMATCH"),  # 1-form mass in a 0-form block
        (((6, 1),), "MFV2D_STACK_UNDERFLOW"),
        (((2, 7, False),), "MFV2D_BAD_ENUM"),
        (((2, 1, False), (4,), (1,), (6, 1)), "MFV2D_FAILED_ALLOC"),  # dense + identity (Q3 / Q1)
    ):
        try:
            run(block)
        except RuntimeError as e:
            assert code in str(e), (block, str(e))
        else:
            raise AssertionError(f"{block} must fail with {code}")
    # Q2: a scalar in front of a dense*dense product is dropped in matrix mode
    m = run(((2, 1, False), (5, 3.0), (2, 1, True)))
    n0 = 9
    assert relerr(m[:n0, :n0], np.eye(n0)) < 1e-10
    # identity block and scaled identity
    m = run(((1,), (5, -2.5)))
    assert np.array_equal(m[:n0, :n0], -2.5 * np.eye(n0))
    # check_bytecode round trip (tests/test_bytecode.py)
    prog = ((2, 1, False), (5, 0.5), (4,), (7, 2, "q", True), (6, 1))
    assert M.check_bytecode(specs, prog) == prog


def check_generic_chain(M, p=3, k=2):
    """A block that is not of the fused form (two dense factors): M1^-1 * (E10 M0 ...) style chains go
    through the column-by-column evaluator and must agree with the oracle's stack machine."""
    rng = np.random.default_rng(9)
    specs = M._ElementFormSpecification(("u", 1), ("q", 2))
    names, orders = ["u", "q"], [1, 2]
    prog = ((3, 1, False), (2, 2, False), (5, 0.5), (2, 2, True), (3, 1, True), (2, 1, True))
    exprs = ((prog, None), (((3, 1, False), (2, 2, True)), ((2, 2, False), (5, 2.0), (4,), (2, 2, True), (6, 1))))
    corners = np.ascontiguousarray(F.QUADS[3:6])
    batch = M.ElementBatch(make_basis(M, p, k), corners)
    n = specs.total_size(p, p)
    x = rng.standard_normal((3, n))
    mats = M.compute_element_matrices(specs, exprs, batch)
    vecs = M.compute_element_vectors(specs, exprs, batch, x)
    ob = O.basis(p, p + k + 1)
    blocks, fs = O.decode_expressions(exprs, names)
    for i in range(3):
        assert relerr(mats[i], O.element_matrix(ob, corners[i], orders, blocks, fs)) < 1e-10
        assert relerr(vecs[i], O.element_vector(ob, corners[i], orders, blocks, x[i], fs)) < 1e-10


def check_linear_system(M, SV, ST, fix, nh, nv):
    """tests/test_solving.py: operators against the reference's outputs, Schur solve against its solution."""
    g = golden(fix + ".npz")
    s = SYSTEMS["mixed_poisson_sorted"]
    p, k = int(g["order"]), int(g["k"])
    specs = specs_of(M, s)
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
    mats = M.compute_element_matrices(specs, s["lhs"], batch)
    assert relerr(mats[0], g["mat0"]) < TOL_MAT
    ne = mats.shape[1]
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    assert relerr(sysm.apply_diagonal(g["x"]), g["apply_diagonal"]) < 1e-12
    assert relerr(sysm.apply_diagonal_inverse(g["x"]), g["apply_diagonal_inverse"]) < 1e-9
    assert relerr(sysm.apply_trace(g["x"]), g["apply_trace"]) < 1e-12
    assert relerr(sysm.apply_trace_transpose(g["trace_in"]), g["apply_trace_transpose"]) < 1e-12
    # the extension type's own constructor, LinearSystem(((A_e, MatrixCRS), ...), offsets, indices), fed the way
    # python/mfv2d/solving.py:62-97 feeds it: one build_row per (constraint, element)
    blocks = sysm.get_constraint_blocks()
    dense = sysm.get_dense_blocks()
    assert len(blocks) == len(dense) == sysm.n_elem and np.array_equal(dense[0], mats[0]) and blocks[0].shape == (nt, ne)
    elem = col // ne
    rebuilt = tuple((dense[e], SV.MatrixCRS(nt, ne)) for e in range(sysm.n_elem))
    order_rc = np.lexsort((col, elem, row))
    r_s, e_s, c_s, v_s = row[order_rc], elem[order_rc], col[order_rc] - elem[order_rc] * ne, val[order_rc]
    cuts = np.flatnonzero(np.diff(r_s * sysm.n_elem + e_s)) + 1
    starts, stops = np.concatenate([[0], cuts]), np.concatenate([cuts, [r_s.size]])
    next_row = np.zeros(sysm.n_elem, np.int64)
    t_len, t_elem = np.zeros(nt, np.int64), []
    for lo, hi in zip(starts, stops):
        e, r = int(e_s[lo]), int(r_s[lo])
        for empty in range(next_row[e], r):
            rebuilt[e][1].build_row(empty)
        rebuilt[e][1].build_row(r, SV.SparseVector.from_entries(ne, c_s[lo:hi], v_s[lo:hi]))
        next_row[e] = r + 1
        t_len[r] += 1
        t_elem.append(e)
    for e in range(sysm.n_elem):
        for empty in range(next_row[e], nt):
            rebuilt[e][1].build_row(empty)
        assert np.array_equal(rebuilt[e][1].position_pairs, blocks[e].position_pairs) and np.array_equal(rebuilt[e][1].values, blocks[e].values)
    t_off = np.concatenate([[0], np.cumsum(t_len)]).astype(np.uint32)
    sys2 = SV.LinearSystem(rebuilt, t_off, np.array(t_elem, np.uint32))
    assert np.array_equal(sys2.apply_trace(g["x"]), sysm.apply_trace(g["x"]))
    assert np.array_equal(sys2.apply_diagonal_inverse(g["x"]), sysm.apply_diagonal_inverse(g["x"]))
    assert np.array_equal(sys2.apply_trace_transpose(g["trace_in"]), sysm.apply_trace_transpose(g["trace_in"]))
    try:
        SV.LinearSystem(rebuilt, t_off, np.array(t_elem[::-1], np.uint32))
        raise AssertionError("wrong element lists accepted")
    except ValueError:
        pass
    del sys2
    t = g["trace_in"]
    composed = sysm.apply_trace(sysm.apply_diagonal_inverse(sysm.apply_trace_transpose(t)))
    assert relerr(sysm.apply_full_trace_system(t), composed) < 1e-12
    x, lam, res, it = SV.solve_schur_iterative(sysm, g["rhs"], g["con_rhs"], SV.ConvergenceSettings(2000, 1e-30, 1e-12))
    assert relerr(x, g["sol_dense"]) < 1e-8
    assert relerr(lam, g["sol_trace"]) < 1e-8
    assert abs(it - int(g["iters"])) <= 12
    # residual of the REFERENCE-assembled saddle system (north_star: 1e-10 relative)
    n_dense = x.size
    r1 = g["rhs"] - sysm.apply_diagonal(x) - sysm.apply_trace_transpose(lam)
    r2 = g["con_rhs"] - sysm.apply_trace(x)
    rel = np.sqrt(r1 @ r1 + r2 @ r2) / np.sqrt(g["rhs"] @ g["rhs"] + g["con_rhs"] @ g["con_rhs"])
    assert rel < 1e-10, rel
    # Jacobi-preconditioned Schur CG: same solution, no more iterations than the plain one
    xj, lj, _, itj = SV.solve_schur_iterative(sysm, g["rhs"], g["con_rhs"], SV.ConvergenceSettings(2000, 1e-30, 1e-12), precondition=True)
    assert relerr(xj, g["sol_dense"]) < 1e-8 and relerr(lj, g["sol_trace"]) < 1e-8
    assert itj <= it
    # solve_pcg_iterative (python/mfv2d/solving.py:634-684) with the reference's own settings; its iteration count
    # on these systems is in tests/golden/pcg.json (the stopping test sits at 1e-30 absolute, far below the rounding
    # floor of the recurrence, so the count moves by ten to fifteen percent with the last bits of the matrices and
    # of the dot products)
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pcg.json")) as f:
        ref_pcg = json.load(f)[fix]
    xp, lp, resp, itp = SV.solve_pcg_iterative(sysm, g["rhs"], g["con_rhs"], SV.ConvergenceSettings(5000, 1e-30, 1e-10))
    assert relerr(xp, g["sol_dense"]) < 1e-8 and relerr(lp, g["sol_trace"]) < 1e-8
    assert resp < 1e-30
    # The count itself is not a stable quantity: this recurrence runs on an indefinite matrix with an indefinite preconditioner
    # and amplifies rounding differences ~100x per ten iterations (profiles/r2/dbg_pcg*.py); 200 runs of a numpy restatement
    # with the right-hand side perturbed by 1e-15 give a median of 306 iterations on the 6 x 6 fixture, but also 347 and 1 453.
    # Typical runs land within 20 % of the reference's count; the gate only excludes a different ALGORITHM.
    assert itp <= 5 * ref_pcg["iterations"], (itp, ref_pcg)
    # DenseVector / TraceVector objects (system_objects.c:725-2152) around the same device operations
    dv, dw = SV.DenseVector(sysm, *np.split(g["x"], sysm.offsets[1:-1])), SV.DenseVector(sysm)
    tv = SV.TraceVector(sysm, g["trace_in"])
    assert dv.parent is sysm and np.array_equal(dv.as_merged(), g["x"]) and np.array_equal(dw.as_merged(), np.zeros(sysm.n_dense))
    assert len(dv.as_split()) == sysm.n_elem and np.array_equal(np.concatenate(dv.as_split()), g["x"])
    assert relerr(sysm.apply_diagonal(dv, dw).as_merged(), g["apply_diagonal"]) < 1e-12
    assert relerr(sysm.apply_trace(dv).as_merged(), g["apply_trace"]) < 1e-12
    assert relerr(sysm.apply_trace_transpose(tv).as_merged(), g["apply_trace_transpose"]) < 1e-12
    SV.DenseVector.add(dv, dw, dw, -0.5)
    assert relerr(dw.as_merged(), g["x"] - 0.5 * g["apply_diagonal"]) < 1e-14
    SV.DenseVector.subtract(dv, dv.copy(), dw, 3.0)
    assert relerr(dw.as_merged(), -2.0 * g["x"]) < 1e-15
    SV.TraceVector.scale(tv, 2.0, tv)
    assert relerr(tv.as_merged(), 2.0 * g["trace_in"]) < 1e-15
    assert abs(SV.TraceVector.dot(tv, tv) - 4.0 * g["trace_in"] @ g["trace_in"]) < 1e-10 * abs(4.0 * g["trace_in"] @ g["trace_in"]) + 1e-300
    sv0 = tv.as_split()[0]
    idx0, val0 = sv0.indices.astype(np.int64), sv0.values
    assert np.array_equal(val0, 2.0 * g["trace_in"][idx0]) and idx0.size > 0
    xs_, ls_, _, its_ = SV.solve_schur_iterative(sysm, SV.DenseVector(sysm, *np.split(g["rhs"], sysm.offsets[1:-1])),
                                                SV.TraceVector(sysm, g["con_rhs"]), SV.ConvergenceSettings(2000, 1e-30, 1e-12))
    assert isinstance(xs_, SV.DenseVector) and its_ == it and np.array_equal(xs_.as_merged(), x) and np.array_equal(ls_.as_merged(), lam)
    # compute_element_rhs (solve_system.py:129-160) of the same system: the fixture's rhs IS the reference's output
    from types import SimpleNamespace as NS

    def f_exact(x_, y_):
        return -2 * np.pi**2 * np.cos(np.pi * x_) * np.cos(np.pi * y_)

    duck = NS(equations=[NS(weight=NS(order=2), right=NS(explicit_terms=[])),
                         NS(weight=NS(order=3), right=NS(explicit_terms=[(1.0, NS(func=f_exact, weight=NS(order=3)))]))])
    assert relerr(M.compute_element_rhs(duck, batch).ravel(), g["rhs"]) < 1e-12
    # the assembly loop of the callers collapsed: matrices stay on the device (solve_system.py:988-1008)
    dev_sys = SV.compute_linear_system(specs, s["lhs"], batch, (row, col, val), nt)
    assert relerr(dev_sys.apply_diagonal(g["x"]), g["apply_diagonal"]) < 1e-12
    assert relerr(dev_sys.apply_full_trace_system(t), composed) < 1e-11
    # the reference's constructor call: LinearSystem(n_elem, form_specs, orders, element_matrices, constraints)
    cons = [SV.Constraint(float(g["con_rhs"][i]), *[SV.ElementConstraint(int(e_), col[(row == i) & (el == e_)] - sysm.offsets[e_],
                                                                           val[(row == i) & (el == e_)])
                                                     for e_ in np.unique(el[row == i])])
            for el in [np.searchsorted(sysm.offsets, col, side="right") - 1] for i in range(nt)]
    other = SV.LinearSystem(sysm.n_elem, specs, np.full((sysm.n_elem, 2), p), list(mats), cons)
    assert other.n_trace == nt and np.array_equal(other.constraint_rhs, g["con_rhs"])
    assert relerr(other.apply_trace(g["x"]), g["apply_trace"]) < 1e-12
    with pytest.raises(ValueError):
        SV.DenseVector.add(dv, SV.DenseVector(other), dw, 1.0)
    with pytest.raises(ValueError):
        other.apply_diagonal(dv)
    # BLAS-1
    a, b = g["x"], g["rhs"]
    assert relerr(SV.vec_axpby(2.0, a, -0.5, b), 2.0 * a - 0.5 * b) < 1e-15
    assert abs(SV.vec_dot(a, b) - a @ b) < 1e-10 * abs(a @ b) + 1e-12


def check_projectors(M):
    """compute_element_projector against the reference's own output (tests/golden/projector.npz) and the oracle."""
    g = golden("projector.npz")
    specs = M._ElementFormSpecification(("a", 1), ("b", 2), ("c", 3))
    for po, pi, ro in g["cases"]:
        po, pi, ro = int(po), int(pi), int(ro)
        rule = M.IntegrationRule1D(ro)
        bo1, bi1 = M.Basis1D(po, rule), M.Basis1D(pi, rule)
        bo, bi = M.Basis2D(bo1, bo1), M.Basis2D(bi1, bi1)
        for iq in (4, 5):
            sp = M.ElementFemSpace2D(bo, np.ascontiguousarray(F.QUADS[iq]))
            for dual in (False, True):
                res = M.compute_element_projector(specs, bi, sp, dual)
                assert len(res) == 3
                for fo, m in zip((1, 2, 3), res):
                    ref = g[f"po{po}_pi{pi}_r{ro}_q{iq}_d{int(dual)}_f{fo}"]
                    assert m.shape == ref.shape
                    assert relerr(m, ref) < (TOL_MAT if dual else TOL_INV), (po, pi, iq, dual, fo, relerr(m, ref))
        # batched entry point and the oracle's two-basis mass on all the test quads
        batch = M.ElementBatch(bo, np.ascontiguousarray(F.QUADS))
        ob_o, ob_i = O.basis(po, ro + 1), O.basis(pi, ro + 1)
        for fo in (1, 2, 3):
            mm = M.compute_element_projectors(fo, bi, batch, dual=True)
            assert relerr(mm, O.mass_double(ob_o, ob_i, F.QUADS, fo)) < TOL_MAT
    # projecting onto the same basis is the identity; mismatched rules are rejected like the reference does
    rule = M.IntegrationRule1D(6)
    b1 = M.Basis1D(3, rule)
    b = M.Basis2D(b1, b1)
    sp = M.ElementFemSpace2D(b, np.ascontiguousarray(F.QUADS[6]))
    for m in M.compute_element_projector((1, 2, 3), b, sp):
        assert relerr(m, np.eye(m.shape[0])) < 1e-10
    b2 = M.Basis1D(3, M.IntegrationRule1D(5))
    with pytest.raises(ValueError):
        M.compute_element_projector((1,), M.Basis2D(b2, b2), sp)


def check_saddle_residual_small_elements(M, SV, ST, nh=12, nv=10, p=4, scale=2e-3):
    """north_star: solutions within 1e-10 relative residual.  On small elements (h ~ 1e-3, as on the 1000 x 1000 mesh of
    BASELINE.json configs[4]) the 2-form mass is ~1/h^2 times the 1-form mass and the element blocks, applied as
    explicit inverses, lose cond(A_e) * eps: the solve's refinement step (system.cu) has to bring the residual of
    the saddle-point system [[A, N^T], [N, 0]] back below the gate."""
    s = SYSTEMS["mixed_poisson"]
    specs = specs_of(M, s)

    def deform(xi, eta):
        return (xi + 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta), eta - 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta))

    corners = np.ascontiguousarray(ST.unit_square_corners(nh, nv, deform) * scale)
    batch = M.ElementBatch(make_basis(M, p, 3), corners)
    mats = M.compute_element_matrices(specs, s["lhs"], batch)
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    # right-hand side as in profiles/config4_solve.py: dual DoFs of a smooth source on the 2-form rows
    ne = specs.total_size(p, p)
    rhs = np.zeros((corners.shape[0], ne))
    rhs[:, specs.form_offset(1, p, p):] = batch.dual_dofs(s["form_orders"][1], lambda x, y: np.cos(np.pi * x / scale) * np.cos(np.pi * y / scale))
    rhs, phi = rhs.ravel(), np.zeros(nt)

    def residual(refine):
        os.environ["MFV2D_SCHUR_REFINE"] = refine
        try:
            x, lam, _, it = SV.solve_schur_iterative(sysm, rhs, phi, SV.ConvergenceSettings(20000, 1e300, 1e-13))
        finally:
            del os.environ["MFV2D_SCHUR_REFINE"]
        r1 = rhs - sysm.apply_diagonal(x) - sysm.apply_trace_transpose(lam)
        r2 = phi - sysm.apply_trace(x)
        return np.sqrt(r1 @ r1 + r2 @ r2) / np.sqrt(rhs @ rhs), it

    res_plain, it0 = residual("0")
    res_refined, it1 = residual("1")
    assert abs(it0 - it1) <= 2 and it0 < 20000
    assert res_refined < 1e-10, (res_plain, res_refined)
    assert res_refined <= res_plain
    return res_plain, res_refined


def check_degenerate_element(M):
    """A collapsed element inside a batch: every path that inverts a mass matrix must report MFV2D_PIVOT_FAILED with the
    element's index (matrix_full_invert -> decompose_pivoted_lu, src/algebra/matrix.c:361-460), never return inf / NaN."""
    quads = np.ascontiguousarray(F.QUADS[:3]).copy()
    quads[1] = quads[1][0]  # all four corners at one point: det = 0 everywhere
    batch = M.ElementBatch(make_basis(M, 3, 2), quads)
    for order in (1, 2, 3):
        try:
            batch.mass(order, inverse=True)
            raise AssertionError("a singular mass matrix must raise")
        except RuntimeError as e:
            assert "MFV2D_PIVOT_FAILED" in str(e) and "element 1" in str(e), str(e)
    assert np.all(np.isfinite(batch.mass(1)))  # the forward 0-form mass of a collapsed element is just zero (1-/2-forms divide by det, as the reference does)
    specs = M._ElementFormSpecification(("q", 2), ("u", 3))
    try:
        M.compute_element_matrices(specs, M._mass_expressions(specs, True), batch)
        raise AssertionError("an element matrix with a singular inverse mass must raise")
    except RuntimeError as e:
        assert "MFV2D_PIVOT_FAILED" in str(e), str(e)
    good = M.ElementBatch(make_basis(M, 3, 2), np.ascontiguousarray(F.QUADS[:3]))
    inv = M.compute_element_matrices(specs, M._mass_expressions(specs, True), good)
    fwd = M.compute_element_matrices(specs, M._mass_expressions(specs, False), good)
    assert relerr(inv @ fwd, np.broadcast_to(np.eye(inv.shape[1]), inv.shape)) < 1e-10


def check_partitioned_solve(M, SV, ST, ctx, rank, world, all_gather, fix, nh, nv):
    """One rank of a partitioned Schur solve of a golden mixed-Poisson system through the NVLink peer exchange
    (peer.cu): strips of elements, interface sums and Krylov scalars written into the other ranks' exchange blocks by
    the solver's own kernels.  ``all_gather(obj) -> list`` is the host-side plumbing (gloo between processes on the GPU
    box, a barrier between the threads of the emulated run).  Returns (iterations, error) after asserting parity with
    the reference's solution."""
    g = golden(fix + ".npz")
    p, k = int(g["order"]), int(g["k"])
    s = SYSTEMS["mixed_poisson_sorted"]
    ne = int(ST.form_sizes(s["form_orders"], p).sum())
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    P = ST.partition_constraints(row, col, val, nt, bounds, rank)
    lo, hi = parts[rank]
    rhs_l = np.ascontiguousarray(g["rhs"][bounds[rank]:bounds[rank + 1]])
    phi_l = np.ascontiguousarray(g["con_rhs"][P.global_rows])
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(g["corners"][lo:hi]), ctx)
    mats = M.compute_element_matrices(specs_of(M, s), s["lhs"], batch)
    sysm = SV.LinearSystem(mats, (P.row, P.col, P.val), P.n_trace, ctx=ctx)
    if world > 1:
        sysm.set_interface(P.shared_index, P.slot, P.n_slots, P.owned)
        sysm.connect_peers(P.holder_ptr, P.holder_rank, all_gather)
    xg = np.ascontiguousarray(g["x"][bounds[rank]:bounds[rank + 1]])
    assert relerr(sysm.apply_trace(xg), g["apply_trace"][P.global_rows]) < 1e-12
    conv = SV.ConvergenceSettings(2000, 1e-30, 1e-12)
    x, lam, res, it = SV.solve_schur_iterative(sysm, rhs_l, phi_l, conv)
    assert not sysm.peer_error()
    err = max(relerr(x, g["sol_dense"][bounds[rank]:bounds[rank + 1]]), relerr(lam, g["sol_trace"][P.global_rows]))
    assert err < 1e-8, (rank, err)
    assert abs(it - int(g["iters"])) <= 12, (it, int(g["iters"]))
    xj, lj, _, itj = SV.solve_schur_iterative(sysm, rhs_l, phi_l, conv, precondition="jacobi")
    assert relerr(xj, g["sol_dense"][bounds[rank]:bounds[rank + 1]]) < 1e-8 and itj <= it + 4
    # solve_pcg_iterative on the partitioned saddle-point system: device scalars, interface sums, ownership-weighted dots
    xp, lp, _, it2 = SV.solve_pcg_iterative(sysm, rhs_l, phi_l, SV.ConvergenceSettings(5000, 1e-30, 1e-10))
    assert relerr(xp, g["sol_dense"][bounds[rank]:bounds[rank + 1]]) < 1e-8 and relerr(lp, g["sol_trace"][P.global_rows]) < 1e-8
    # every rank took the same number of iterations (identical scalars on all ranks)
    counts = all_gather((it, itj, it2))
    assert all(c == counts[0] for c in counts), counts
    return it, itj, it2, err


def _timemarch_system(M):
    """bytecode + form layout of the time-march-updated advection-diffusion system of tests/golden/timemarch_advdif.npz"""
    g = golden("timemarch_advdif.npz")
    raw = json.loads(str(g["codes"]))
    lhs = []
    for rowc in raw["lhs"]:
        r = []
        for blk in rowc:
            if blk is None:
                r.append(None)
                continue
            ops = []
            for item in blk:
                item = list(item)
                if item[0] == 7 and isinstance(item[2], str) and item[2].startswith("fn:"):
                    item[2] = F.FIELDS[item[2][3:]]
                ops.append(tuple(item))
            r.append(tuple(ops))
        lhs.append(tuple(r))
    return g, raw, tuple(lhs), M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))


def check_partitioned_bicgstab(M, SV, ST, ctx, rank, world, all_gather, nh=8, nv=8):
    """One rank of a partitioned BiCGStab solve on the Schur complement of a NON-symmetric system (the advection-diffusion
    operator of the time-march fixture on its 8 x 8 order-4 mesh): the strip solution must agree with a direct sparse solve
    of the global saddle-point system (SciPy SuperLU on the host, the reference's production path) and every rank must take
    the same number of iterations."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl

    g, raw, lhs, specs = _timemarch_system(M)
    p, k = int(g["order"]), int(g["k"])
    ne = specs.total_size(p, p)
    corners = np.ascontiguousarray(g["corners"])
    n = corners.shape[0]
    assert n == nh * nv
    row, col, val, nt = ST.continuity_constraints(nh, nv, p, raw["form_orders"])
    rhs = np.asarray(g["explicit"], np.float64)
    phi = np.zeros(nt)
    # global direct solve (host)
    mats_all = M.compute_element_matrices(specs, lhs, M.ElementBatch(make_basis(M, p, k), corners, ctx))
    A = sp.block_diag([sp.csr_matrix(m) for m in mats_all], format="csr")
    N = sp.coo_matrix((val, (row, col)), shape=(nt, n * ne)).tocsr()
    sol = spl.spsolve(sp.bmat([[A, N.T], [N, None]], format="csc"), np.concatenate([rhs, phi]))
    # this rank's strip
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    lo, hi = parts[rank]
    if world > 1:
        P = ST.partition_constraints(row, col, val, nt, bounds, rank)
        prow, pcol, pval, pnt, rows_glob = P.row, P.col, P.val, P.n_trace, P.global_rows
    else:
        prow, pcol, pval, pnt, rows_glob = row, col, val, nt, np.arange(nt)
    sysm = SV.LinearSystem(np.ascontiguousarray(mats_all[lo:hi]), (prow, pcol, pval), pnt, ctx=ctx)
    _, hier, _ = ST.strip_system_1form(nh, nv, p, raw["form_orders"], world, rank, (2,), 4)
    if world > 1:
        sysm.set_interface(P.shared_index, P.slot, P.n_slots, P.owned)
        sysm.connect_peers(P.holder_ptr, P.holder_rank, all_gather, n_coarse=hier[2][-1])
    x, lam, res, it = SV.solve_schur_bicgstab(sysm, np.ascontiguousarray(rhs[bounds[rank]:bounds[rank + 1]]), np.ascontiguousarray(phi[rows_glob]),
                                              SV.ConvergenceSettings(4000, 1e300, 1e-13))
    assert not (world > 1 and sysm.peer_error())
    ex = relerr(x, sol[bounds[rank]:bounds[rank + 1]])
    el = float(np.max(np.abs(lam - sol[n * ne:][rows_glob])) / max(np.max(np.abs(sol[n * ne:])), 1e-300))
    assert ex < 1e-9 and el < 1e-9, (rank, ex, el, it)
    counts = all_gather(int(it))
    assert all(c == counts[0] for c in counts), counts
    # the multilevel preconditioner (aggregates of 2 x 2 elements + exact level of 4 x 4) on the same non-symmetric system
    sysm.set_coarse_space(*hier)
    x2, lam2, _, it2 = SV.solve_schur_bicgstab(sysm, np.ascontiguousarray(rhs[bounds[rank]:bounds[rank + 1]]), np.ascontiguousarray(phi[rows_glob]),
                                               SV.ConvergenceSettings(4000, 1e300, 1e-13), precondition="multilevel")
    assert not (world > 1 and sysm.peer_error())
    assert relerr(x2, sol[bounds[rank]:bounds[rank + 1]]) < 1e-9, (rank, it2)
    counts2 = all_gather(int(it2))
    assert all(c == counts2[0] for c in counts2), counts2
    assert it2 <= it + 4, (it2, it)  # 8 x 8 elements with the 2/dt mass term: well conditioned, the coarse levels cannot gain here
    return int(it), max(ex, el)


def check_partitioned_march(M, SV, ST, ctx, rank, world, all_gather, nh=8, nv=8):
    """One rank of the PARTITIONED time march (configs[2] on several GPUs, in small): the strips of the 8 x 8 order-4
    advection-diffusion fixture march together - interface sums and BiCGStab scalars through the peer blocks, the stopping
    test of every correction through `reduce_max` - with host vectors and with device-resident vectors; every rank's strip of
    every step must equal the reference's own march (tests/golden/timemarch_advdif.npz)."""
    from mfv2d_b200 import callers as CL

    g, raw, lhs, specs = _timemarch_system(M)
    p, k = int(g["order"]), int(g["k"])
    ne = specs.total_size(p, p)
    corners = np.ascontiguousarray(g["corners"])
    row, col, val, nt = g["con_row"], g["con_col"], g["con_val"], g["con_rhs"].size
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    lo, hi = parts[rank]
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(corners[lo:hi]), ctx)
    if world > 1:
        P = ST.partition_constraints(row, col, val, nt, bounds, rank)
        prow, pcol, pval, pnt, rows_glob = P.row, P.col, P.val, P.n_trace, P.global_rows
    else:
        prow, pcol, pval, pnt, rows_glob = row, col, val, nt, np.arange(nt)
    explicit = np.ascontiguousarray(g["explicit"][bounds[rank]:bounds[rank + 1]])
    con_rhs = np.ascontiguousarray(g["con_rhs"][rows_glob])
    reduce_max = (lambda v: max(all_gather(float(v)))) if world > 1 else None
    ref = g["solutions"]
    out = []
    for device in (False, True):
        system = SV.compute_linear_system(specs, lhs, batch, (prow, pcol, pval), pnt)
        if world > 1:
            system.set_interface(P.shared_index, P.slot, P.n_slots, P.owned)
            system.connect_peers(P.holder_ptr, P.holder_rank, all_gather)
        sols, iters = CL.trapezoidal_march(specs, lhs, None, batch, (prow, pcol, pval), pnt, con_rhs, explicit, [raw["form_names"].index("u")],
                                           float(g["dt"]), int(g["nt"]), system=system, reduce_max=reduce_max, device=device)
        assert not (world > 1 and system.peer_error())
        for i, sv in enumerate(sols):
            assert relerr(sv, ref[i + 1][bounds[rank]:bounds[rank + 1]]) < 1e-8, (rank, device, i)
        counts = all_gather(tuple(int(i) for i in iters))
        assert all(c == counts[0] for c in counts), counts
        out.append(tuple(int(i) for i in iters))
    assert out[0] == out[1]
    return out[0]


def check_multilevel(M, SV, ST, ctx, rank, world, all_gather, nh=24, nv=24, p=2, k=2, mid_sizes=(2,), exact_size=6, deform=None):
    """The multilevel preconditioner (precond.cu) on a synthetic deformed mesh, one rank of ``world``: same solution as the
    Jacobi-preconditioned CG, clearly fewer iterations, identical counts on every rank."""
    s = SYSTEMS["mixed_poisson_sorted"]
    ne = int(ST.form_sizes(s["form_orders"], p).sum())
    corners = ST.unit_square_corners(nh, nv, deform or F.deform)
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    P = ST.partition_constraints(row, col, val, nt, bounds, rank)
    lo, hi = parts[rank]
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(corners[lo:hi]), ctx)
    mats = M.compute_element_matrices(specs_of(M, s), s["lhs"], batch)
    sysm = SV.LinearSystem(mats, (P.row, P.col, P.val), P.n_trace, ctx=ctx)
    strip_rows = [parts[r][0] // nh for r in range(world)] + [nv]
    agg, z, n_agg = ST.coarse_hierarchy_1form(nh, nv, p, s["form_orders"], mid_sizes, exact_size, strip_rows)
    if world > 1:
        sysm.set_interface(P.shared_index, P.slot, P.n_slots, P.owned)
        sysm.connect_peers(P.holder_ptr, P.holder_rank, all_gather, n_coarse=n_agg[-1])
    sysm.set_coarse_space(*ST.localize_hierarchy(agg, z, n_agg, P))
    rng = np.random.default_rng(11)
    rhs = rng.standard_normal(nh * nv * ne)[bounds[rank]:bounds[rank + 1]]
    phi = np.zeros(P.n_trace)
    conv = SV.ConvergenceSettings(5000, 1e300, 1e-11)
    xj, lj, _, itj = SV.solve_schur_iterative(sysm, rhs, phi, conv, precondition="jacobi")
    x2, l2, _, it2 = SV.solve_schur_iterative(sysm, rhs, phi, conv, precondition="multilevel")
    assert not sysm.peer_error()
    assert relerr(x2, xj) < 1e-7 and relerr(l2, lj) < 1e-7, (relerr(x2, xj), relerr(l2, lj))
    assert it2 < 0.75 * itj, (it2, itj)
    counts = all_gather((itj, it2))
    assert all(c == counts[0] for c in counts), counts
    return itj, it2


def check_hp_refined(M, SV, ST, fixture="hp_refined", gate=1e-10):
    """BASELINE.json configs[3]: the h/p-refined, non-conforming meshes of tests/golden/hp_refined*.npz (93 leaves, hanging-side
    constraint rows with up to 26 entries - all produced by the reference).  Element matrices per (p, q) bucket, ONE
    LinearSystem with blocks of several different sizes (the non-uniform path of system.cu), its operators against the
    reference-assembled ones, and the Schur-complement solve (constraint groups of hanging sides re-expressed in an
    orthonormal basis, solving.ConstraintConditioner) checked by the relative residual of the REFERENCE-assembled
    saddle-point system.  `hp_refined` (orders 2..8) has hanging-side coefficients up to 4e8: the
    reference's own SuperLU solution leaves 1.7e-10 there, the gate for it is therefore looser; `hp_refined_345`
    (orders 2..5) meets north_star's 1e-10."""
    g = golden(fixture + ".npz")
    s = SYSTEMS["mixed_poisson_sorted"]
    specs = specs_of(M, s)
    k = int(g["k"])
    orders = g["orders"][:, 0].astype(int)
    assert np.array_equal(g["orders"][:, 0], g["orders"][:, 1])
    mats = [None] * orders.size
    for p in np.unique(orders):
        idx = np.flatnonzero(orders == p)
        batch = M.ElementBatch(make_basis(M, int(p), k), np.ascontiguousarray(g["corners"][idx]))
        out = M.compute_element_matrices(specs, s["lhs"], batch)
        for j, e in enumerate(idx):
            mats[e] = out[j]
    assert relerr(mats[0], g["mat_first"]) < TOL_MAT and relerr(mats[-1], g["mat_last"]) < TOL_MAT
    sizes = np.array([m.shape[0] for m in mats])
    off = np.concatenate([[0], np.cumsum(sizes)])
    assert np.array_equal(sizes, [specs.total_size(int(p), int(p)) for p in orders]) and len(set(sizes.tolist())) >= 4
    row = g["con_row"].astype(np.int64)
    col = off[g["con_elem"].astype(np.int64)] + g["con_dof"].astype(np.int64)
    o = np.lexsort((col, row))
    row, col, val = row[o], col[o], g["con_val"][o]
    nt = g["con_rhs"].size
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    assert relerr(sysm.apply_diagonal(g["x"]), g["apply_diagonal"]) < 1e-12
    assert relerr(sysm.apply_diagonal_inverse(g["x"]), g["apply_diagonal_inverse"]) < 1e-8
    assert relerr(sysm.apply_trace(g["x"]), g["apply_trace"]) < 1e-12
    assert relerr(sysm.apply_trace_transpose(g["trace_in"]), g["apply_trace_transpose"]) < 1e-12
    t = g["trace_in"]
    composed = sysm.apply_trace(sysm.apply_diagonal_inverse(sysm.apply_trace_transpose(t)))
    assert relerr(sysm.apply_full_trace_system(t), composed) < 1e-10
    cond = SV.ConstraintConditioner(row, col, val, nt, off)
    assert len(cond.blocks) > 10
    sysc = SV.LinearSystem(mats, (cond.row, cond.col, cond.val), nt)
    x, lam_c, res, it = SV.solve_schur_iterative(sysc, g["rhs"], cond.rhs(g["con_rhs"]), SV.ConvergenceSettings(20000, 1e300, 1e-13), precondition="jacobi")
    lam = cond.multipliers(lam_c)
    # residual of the REFERENCE-assembled saddle-point system (its N, its A) with our solution
    r1 = g["rhs"] - sysm.apply_diagonal(x) - sysm.apply_trace_transpose(lam)
    r2 = g["con_rhs"] - sysm.apply_trace(x)
    rel = np.sqrt(r1 @ r1 + r2 @ r2) / np.sqrt(g["rhs"] @ g["rhs"] + g["con_rhs"] @ g["con_rhs"])
    assert rel < gate, rel
    assert relerr(x, g["sol_dense"]) < 1e-7, relerr(x, g["sol_dense"])
    assert it < 3000, it
    return it, rel


def check_hp_forest(M, SV, ST, mesh="b", gate=1e-10, sol_tol=1e-7, bc=False):
    """BASELINE.json configs[3] without any reference-generated input at run time: the quadtree of tests/golden/forest.npz as
    arrays -> leaf corners by subdivision, constraint rows from mfv2d_b200.nonconforming (pinned bit for bit to the reference in
    tests/test_nonconforming.py), element matrices per order bucket, ONE LinearSystem with mixed block sizes, the hanging-side row
    groups conditioned, Schur CG; checked by the relative residual of the saddle-point system and against a direct sparse solve."""
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl

    from mfv2d_b200 import nonconforming as NC

    g = golden("forest.npz")
    nh, nv = (int(v) for v in g[f"{mesh}_dims"])
    forest = NC.QuadForest(nh, nv, g[f"{mesh}_children"], g[f"{mesh}_leaf"], g[f"{mesh}_orders"])
    s = SYSTEMS["mixed_poisson_sorted"]
    specs = specs_of(M, s)
    k = 2
    corners = forest.leaf_corners(ST.unit_square_corners(nh, nv, F.deform if mesh != "c" else None))
    orders = forest.leaf_orders.astype(int)
    mats, rhs_parts = [None] * orders.size, [None] * orders.size

    def source(x, y):
        return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)

    for p in np.unique(orders):
        idx = np.flatnonzero(orders == p)
        batch = M.ElementBatch(make_basis(M, int(p), k), np.ascontiguousarray(corners[idx]))
        out = M.compute_element_matrices(specs, s["lhs"], batch)
        f = np.asarray(batch.dual_dofs(3, source))
        for j, e in enumerate(idx):
            mats[e] = out[j]
            r = np.zeros(out[j].shape[0])
            r[-int(p) * int(p):] = f[j]
            rhs_parts[e] = r
    if bc:
        # boundary conditions from arrays (nonconforming.system_constraints_forest): the flux q.n prescribed strongly on the first
        # half of the boundary lines, the value of u weakly (natural condition) on the others
        root_corners = ST.unit_square_corners(nh, nv, F.deform if mesh != "c" else None)
        n_lines = 2 * (nh + nv)
        row, col, val, phi, nt, off, weak_vec = NC.system_constraints_forest(
            forest, s["form_orders"], root_corners, k, (),
            [(0, lambda x, y: np.stack((np.pi * np.sin(np.pi * x) * np.cos(np.pi * y), np.pi * np.cos(np.pi * x) * np.sin(np.pi * y)), axis=-1), range(n_lines // 2))],
            [(0, [(1.0, lambda x, y: np.cos(np.pi * x) * np.cos(np.pi * y))])])
        assert np.array_equal(np.diff(off), [m.shape[0] for m in mats]) and np.any(weak_vec) and np.any(phi)
        rhs = np.concatenate(rhs_parts) + weak_vec
    else:
        row, col, val, nt, off = NC.continuity_constraints_forest(forest, s["form_orders"])
        assert np.array_equal(np.diff(off), [m.shape[0] for m in mats])
        o = np.lexsort((col, row))
        row, col, val = row[o], col[o], val[o]
        rhs, phi = np.concatenate(rhs_parts), np.zeros(nt)
    A = sp.block_diag([sp.csr_matrix(m) for m in mats], format="csr")
    N = sp.coo_matrix((val, (row, col)), shape=(nt, int(off[-1]))).tocsr()
    K = sp.bmat([[A, N.T], [N, None]], format="csc")
    sol = spl.spsolve(K, np.concatenate([rhs, phi]))
    direct_rel = np.linalg.norm(K @ sol - np.concatenate([rhs, phi])) / np.linalg.norm(rhs)
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    cond = SV.ConstraintConditioner(row, col, val, nt, off)
    sysc = SV.LinearSystem(mats, (cond.row, cond.col, cond.val), nt)
    x, lam_c, res, it = SV.solve_schur_iterative(sysc, rhs, cond.rhs(phi), SV.ConvergenceSettings(20000, 1e300, 1e-13), precondition="jacobi")
    lam = cond.multipliers(lam_c)
    r1 = rhs - sysm.apply_diagonal(x) - sysm.apply_trace_transpose(lam)
    r2 = phi - sysm.apply_trace(x)
    rel = np.sqrt(r1 @ r1 + r2 @ r2) / np.sqrt(rhs @ rhs)
    assert rel < gate, (rel, direct_rel, it, float(np.abs(val).max()))
    assert relerr(x, sol[:x.size]) < sol_tol, relerr(x, sol[:x.size])
    return it, rel, direct_rel


def check_time_march(M, SV, ST):
    """BASELINE.json configs[2] (slice): three trapezoidal steps of the unsteady advection-diffusion system on 8 x 8 elements
    of order 4 against the reference's own march (tests/golden/timemarch_advdif.npz, produced by solve_system_2d): element
    matrices with the 2/dt mass terms, BiCGStab on the (non-symmetric) Schur complement instead of SuperLU, dual <- primal
    by the matrix-free mass apply."""
    from mfv2d_b200 import callers as CL

    g = golden("timemarch_advdif.npz")
    raw = json.loads(str(g["codes"]))

    def dec(codes):
        out = []
        for rowc in codes:
            r = []
            for blk in rowc:
                if blk is None:
                    r.append(None)
                    continue
                ops = []
                for item in blk:
                    item = list(item)
                    if item[0] == 7 and isinstance(item[2], str) and item[2].startswith("fn:"):
                        item[2] = F.FIELDS[item[2][3:]]
                    ops.append(tuple(item))
                r.append(tuple(ops))
            out.append(tuple(r))
        return tuple(out)

    lhs = dec(raw["lhs"])
    assert raw["rhs"] is None
    p, k = int(g["order"]), int(g["k"])
    specs = M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
    mat0 = M.compute_element_matrices(specs, lhs, batch)[0]
    assert relerr(mat0, g["mat0"]) < TOL_MAT
    assert relerr(mat0, mat0.T) > 1e-3  # the advection term makes the blocks non-symmetric
    # the explicit right-hand side of the reference's compute_linear_system rebuilt from arrays: dual DoFs of the source on
    # the u rows (device) + the weak boundary terms `p ^ u_bc` on the q rows (structured.weak_bc_contributions)
    ne = specs.total_size(p, p)
    iq, iu = raw["form_names"].index("q"), raw["form_names"].index("u")
    rhs = np.zeros((batch.n, ne))
    ou = specs.form_offset(iu, p, p)
    rhs[:, ou:ou + specs.form_size(iu, p, p)] = batch.dual_dofs(3, lambda x, y: 3.0 * np.cos(np.pi * x) * np.sin(0.7 * np.pi * y))
    b1 = batch.basis_2d.basis_xi
    ST.weak_bc_contributions(int(g["nh"]), int(g["nv"]), p, raw["form_orders"], iq,
                             [(1.0, lambda x, y: np.sin(0.5 * np.pi * x) * np.cos(0.5 * np.pi * y) + 1.0)], g["corners"], np.asarray(b1.node),
                             np.asarray(b1.edge), np.asarray(b1.rule.nodes), np.asarray(b1.rule.weights), out=rhs)
    assert np.max(np.abs(rhs.ravel() - g["explicit"])) < 1e-13 * np.max(np.abs(g["explicit"]))
    nt = g["con_rhs"].size
    sols, iters = CL.trapezoidal_march(specs, lhs, None, batch, (g["con_row"], g["con_col"], g["con_val"]), nt, g["con_rhs"], g["explicit"],
                                       [raw["form_names"].index("u")], float(g["dt"]), int(g["nt"]))
    ref = g["solutions"]
    assert len(sols) == ref.shape[0] - 1
    for i, s in enumerate(sols):
        assert relerr(s, ref[i + 1]) < 1e-8, (i, relerr(s, ref[i + 1]))
    # the same march with every vector resident on the device: same iteration counts, same solutions up to the Krylov tolerance;
    # steps that `keep` drops are not downloaded
    dsols, diters = CL.trapezoidal_march(specs, lhs, None, batch, (g["con_row"], g["con_col"], g["con_val"]), nt, g["con_rhs"], g["explicit"],
                                         [raw["form_names"].index("u")], float(g["dt"]), int(g["nt"]), device=True,
                                         keep=lambda step: step != 1)
    assert diters == iters and dsols[1] is None
    for i, s in enumerate(dsols):
        if s is not None:
            assert relerr(s, sols[i]) < 1e-10 and relerr(s, ref[i + 1]) < 1e-8, (i, relerr(s, sols[i]))
    # BLAS-1 helpers of the device march: max |x| (NaN propagates, empty -> 0) on host and device buffers
    ctx = batch.ctx
    rng = np.random.default_rng(3)
    v = rng.standard_normal(100003)
    v[77777] = -9.5
    assert CL.vec_absmax(v, ctx) == 9.5 and CL.vec_absmax(ctx.to_device(v), ctx) == 9.5
    assert CL.vec_absmax(np.zeros(0), ctx) == 0.0
    v[5] = np.nan
    assert np.isnan(CL.vec_absmax(ctx.to_device(v), ctx))
    return iters


def check_time_march_heat(M, SV, ST):
    """A march with a 0-FORM unknown against the reference's own solve_system_2d (tests/golden/timemarch_heat.npz: the unsteady
    direct heat equation of examples/unsteady/plot_heat_direct.py): 0-form continuity rows with corner chains and strong
    boundary rows rebuilt from arrays, right-hand-side bytecode (a fixed-point iteration per step), CG on the Schur complement
    of the symmetric positive definite blocks; host vectors and device-resident vectors."""
    from mfv2d_b200 import callers as CL

    g = golden("timemarch_heat.npz")
    raw = json.loads(str(g["codes"]))

    def dec(codes):
        return tuple(tuple(None if blk is None else tuple(tuple(item) for item in blk) for blk in rowc) for rowc in codes)

    lhs, rhs_codes = dec(raw["lhs"]), dec(raw["rhs"])
    p, k, nh, nv = int(g["order"]), int(g["k"]), int(g["nh"]), int(g["nv"])
    specs = M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
    assert relerr(M.compute_element_matrices(specs, lhs, batch)[0], g["mat0"]) < TOL_MAT

    def steady_u(x, y):
        return np.cos(np.pi * x / 2) * np.cos(np.pi * y / 2)

    # constraint matrix and explicit right-hand side from arrays = those of the reference's compute_linear_system
    b1 = batch.basis_2d.basis_xi
    row, col, val, con_rhs, n_trace = ST.system_constraints(nh, nv, p, raw["form_orders"], (), [(0, steady_u)], g["corners"], np.asarray(b1.roots),
                                                            np.asarray(b1.rule.nodes), np.asarray(b1.rule.weights))
    assert np.array_equal(row, g["con_row"]) and np.array_equal(col, g["con_col"]) and np.array_equal(val, g["con_val"])
    assert np.max(np.abs(con_rhs - g["con_rhs"])) < 1e-15
    explicit = 1.0 * np.asarray(batch.dual_dofs(1, steady_u)).ravel()
    assert np.max(np.abs(explicit - g["explicit"])) < 1e-13 * np.max(np.abs(g["explicit"]))
    ref = g["solutions"]
    for device in (False, True):
        sols, iters = CL.trapezoidal_march(specs, lhs, rhs_codes, batch, (row, col, val), n_trace, con_rhs, explicit, [0], float(g["dt"]), int(g["nt"]),
                                           max_iterations=40, atol=1e-12, rtol=0.0, symmetric=True, device=device)
        assert [int(i) for i in iters] == [int(i) for i in g["iterations"]], (device, iters, g["iterations"])
        for i, sv in enumerate(sols):
            assert relerr(sv, ref[i + 1]) < 1e-8, (device, i, relerr(sv, ref[i + 1]))
    return iters


def check_time_march_reaction(M, SV, ST, name):
    """The time-march examples with INITIAL CONDITIONS against the reference's own solve_system_2d (tests/golden/
    timemarch_reaction.npz: examples/unsteady/plot_reaction.py, plot_reaction_mixed.py, plot_vector_reaction.py): the L2 projection
    of the initial functions (dual DoFs, then the matrix-free inverse-mass apply) and three trapezoidal steps from it; and
    plot_heat_mixed.py (no initial condition; weak boundary term and a fixed-point iteration on the right-hand-side bytecode)."""
    from mfv2d_b200 import callers as CL

    g = golden("timemarch_reaction.npz")
    raw = json.loads(str(g[name + "_codes"]))
    def dec(codes):
        return None if codes is None else tuple(tuple(None if blk is None else tuple(tuple(item) for item in blk) for blk in rowc) for rowc in codes)

    lhs, rhs_codes = dec(raw["lhs"]), dec(raw["rhs"])
    assert (rhs_codes is not None) == (name == "heat_mixed")
    p, k = int(g["order"]), int(g["k"])
    specs = M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
    n, ne = batch.n, specs.total_size(p, p)
    initial = {
        "reaction": {"u": lambda x, y: 2 * np.cos(np.pi * x / 2) * np.cos(np.pi * y / 2)},
        "reaction_mixed": {"u": lambda x, y: (1 - x**2) * (1 - y**2), "q": lambda x, y: np.stack((-2 * x * (1 - y**2), -2 * y * (1 - x**2)), axis=-1)},
        "vector_reaction": {"u": lambda x, y: np.stack((2 * x * y, x**2 * y), axis=-1)},
        "heat_mixed": {},
    }[name]
    # initial conditions: dual DoFs of the given functions (zero for the other forms), primal = M^-1 dual per form
    dual = np.zeros((n, ne))
    for fi, (fname, order) in enumerate(zip(raw["form_names"], raw["form_orders"])):
        if fname in initial:
            o = specs.form_offset(fi, p, p)
            dual[:, o:o + specs.form_size(fi, p, p)] = batch.dual_dofs(int(order), initial[fname])
    if initial:
        assert np.max(np.abs(dual.ravel() - g[name + "_initial_dual"])) < 1e-13 * np.max(np.abs(g[name + "_initial_dual"]))
    primal = np.asarray(M.compute_element_primal_from_dual(specs, dual, batch)).ravel()
    ref = g[name + "_solutions"]
    assert (relerr(primal, ref[0]) < 1e-11) if initial else not np.any(ref[0]), relerr(primal, ref[0])
    con = (g[name + "_con_row"], g[name + "_con_col"], g[name + "_con_val"])
    for device in (False, True):
        sols, iters = CL.trapezoidal_march(specs, lhs, rhs_codes, batch, con, g[name + "_con_rhs"].size, g[name + "_con_rhs"], g[name + "_explicit"],
                                           raw["march_forms"], float(g["dt"]), int(g["nt"]), max_iterations=20, atol=1e-12, rtol=0.0,
                                           symmetric=(name == "vector_reaction"), initial_solution=primal, initial_dual=dual.ravel(), device=device)
        for i, sv in enumerate(sols):
            assert relerr(sv, ref[i + 1]) < 1e-8, (name, device, i, relerr(sv, ref[i + 1]))
    return iters


def check_reconstruct_grid(M, p=4, k=2):
    """reconstruct (python/mfv2d/mimetic2d.py:1172-1279) on tensor grids of arbitrary reference points, restated here with
    numpy straight from the formulas (Lagrange values for nodal factors, minus the running sum of derivatives for edge
    factors, push-forward with the Jacobian), and reconstruct_mesh_from_solution's arrays."""
    rng = np.random.default_rng(12)
    ob = O.basis(p, p + k + 1)
    roots = O.gll(p + 1)[0]
    xi, eta = np.linspace(-1, 1, 7), np.array([-0.9, -0.2, 0.35, 1.0])
    batch = M.ElementBatch(make_basis(M, p, k), np.ascontiguousarray(F.QUADS))
    nx, ny = O.lagrange1d(roots, xi), O.lagrange1d(roots, eta)  # (m, p+1)
    ex, ey = -np.cumsum(O.dlagrange1d(roots, xi), axis=1)[:, :p], -np.cumsum(O.dlagrange1d(roots, eta), axis=1)[:, :p]
    for order, nk in ((1, (p + 1) ** 2), (2, 2 * p * (p + 1)), (3, p * p)):
        dofs = rng.standard_normal((F.QUADS.shape[0], nk))
        got = batch.reconstruct_grid(order, dofs, xi, eta)
        for e, quad in enumerate(F.QUADS):
            c = dofs[e]
            if order == 1:
                ref = np.einsum("ij,ai,bj->ab", c.reshape(p + 1, p + 1), ny, nx)
            else:
                x0, y0, x1, y1, x2, y2, x3, y3 = quad.ravel()
                X, E = xi[None, :], eta[:, None]
                j00 = 0.25 * ((x1 - x0) * (1 - E) + (x2 - x3) * (1 + E)) + 0 * X
                j01 = 0.25 * ((y1 - y0) * (1 - E) + (y2 - y3) * (1 + E)) + 0 * X
                j10 = 0.25 * ((x3 - x0) * (1 - X) + (x2 - x1) * (1 + X)) + 0 * E
                j11 = 0.25 * ((y3 - y0) * (1 - X) + (y2 - y1) * (1 + X)) + 0 * E
                det = j00 * j11 - j10 * j01
                if order == 3:
                    ref = np.einsum("ij,ai,bj->ab", c.reshape(p, p), ey, ex) / det
                else:
                    v_eta = np.einsum("ij,ai,bj->ab", c[: p * (p + 1)].reshape(p + 1, p), ny, ex)
                    v_xi = np.einsum("ij,ai,bj->ab", c[p * (p + 1):].reshape(p, p + 1), ey, nx)
                    ref = np.stack((v_xi * j00 + v_eta * j10, v_xi * j01 + v_eta * j11), axis=-1) / det[..., None]
            assert relerr(got[e], ref) < 1e-12, (order, e, relerr(got[e], ref))
        # the per-element entry point with the argument shapes reconstruct_mesh_from_solution uses
        sp = M.ElementFemSpace2D(make_basis(M, p, k), np.ascontiguousarray(F.QUADS[4]))
        one = M.reconstruct(sp, order, dofs[4], xi[None, :], eta[:, None])
        assert np.array_equal(one, got[4])
    # at the quadrature points the grid variant agrees with the integrating-field reconstruction up to its Jacobian factor
    dofs = rng.standard_normal((F.QUADS.shape[0], p * p))
    assert relerr(batch.reconstruct_grid(3, dofs, ob.pnts, ob.pnts), batch.reconstruct(3, dofs)) < 1e-12
    specs = M._ElementFormSpecification(("q", 2), ("u", 3))
    sol = rng.standard_normal((F.QUADS.shape[0], specs.total_size(p, p)))
    grid = M.reconstruct_mesh_from_solution(specs, 5, batch, sol)
    assert grid["x"].shape == (F.QUADS.shape[0], 6, 6) and grid["q"].shape == (F.QUADS.shape[0], 6, 6, 2) and grid["u"].shape == grid["x"].shape
    assert np.allclose(grid["x"][:, 0, 0], F.QUADS[:, 0, 0]) and np.allclose(grid["y"][:, -1, -1], F.QUADS[:, 2, 1])
    assert relerr(grid["u"], batch.reconstruct_grid(3, np.ascontiguousarray(sol[:, 2 * p * (p + 1):]), np.linspace(-1, 1, 6), np.linspace(-1, 1, 6))) == 0.0
