"""The oracle (numpy restatement) against the committed golden fixtures that were produced by the
reference itself, and - when oracle/_ref is built - against the live reference C extension."""
import numpy as np
import pytest
from conftest import golden, load_systems, relerr

import fields as F
from oracle import mfv2d_oracle as O
from oracle.ref_loader import load_ref

SYSTEMS = load_systems()
CASES = [(1, 2), (3, 3), (4, 2), (5, 0)]


def test_tables_and_masses_vs_golden():
    g = golden("element_stokes.npz")
    for p, k in CASES:
        tag = f"p{p}k{k}q4"
        b = O.basis(p, p + k + 1)
        assert np.array_equal(b.pnts, g[tag + "_pnts"])
        assert np.array_equal(b.wgts, g[tag + "_wgts"])
        assert relerr(b.node, g[tag + "_node"]) < 1e-14
        assert relerr(b.edge, g[tag + "_edge"]) < 1e-13
        for fo in (1, 2, 3):
            assert relerr(O.mass(b, F.QUADS[4], fo), g[tag + f"_m{fo - 1}"]) < 1e-13
            assert relerr(O.mass(b, F.QUADS[4], fo, True), g[tag + f"_m{fo - 1}i"]) < 1e-10


@pytest.mark.parametrize("name", sorted(SYSTEMS))
def test_element_matrix_and_vector_vs_golden(name):
    s = SYSTEMS[name]
    g = golden(f"element_{name}.npz")
    for p, k in CASES:
        b = O.basis(p, p + k + 1)
        for iq, quad in enumerate(F.QUADS):
            tag = f"p{p}k{k}q{iq}"
            dofs = g[tag + "_dofs"]
            blocks, fs = O.decode_expressions(s["lhs"], s["form_names"])
            m = O.element_matrix(b, quad, s["form_orders"], blocks, fs, dofs)
            assert relerr(m, g[tag + "_lhs"]) < 1e-12
            v = O.element_vector(b, quad, s["form_orders"], blocks, dofs, fs)
            assert relerr(v, g[tag + "_lhs_vec"]) < 1e-12
            if s["rhs"] is not None:
                blocks, fs = O.decode_expressions(s["rhs"], s["form_names"])
                v = O.element_vector(b, quad, s["form_orders"], blocks, dofs, fs)
                assert relerr(v, g[tag + "_rhs_vec"]) < 1e-12


@pytest.mark.parametrize("fix", ["solve_10x10_p3", "solve_3x4_p4", "solve_5x2_p5_def", "solve_6x6_p3_def"])
def test_system_operators_and_schur_vs_golden(fix):
    g = golden(fix + ".npz")
    s = SYSTEMS["mixed_poisson_sorted"]
    p, k = int(g["order"]), int(g["k"])
    b = O.basis(p, p + k + 1)
    blocks, fs = O.decode_expressions(s["lhs"], s["form_names"])
    mats = [O.element_matrix(b, c, s["form_orders"], blocks, fs) for c in g["corners"]]
    assert relerr(mats[0], g["mat0"]) < 1e-12
    ne = mats[0].shape[0]
    offs = np.arange(len(mats) + 1) * ne
    sysm = O.System(mats, offs, g["con_row"].astype(int), (g["con_elem"].astype(int) * ne + g["con_dof"]).astype(int),
                    g["con_val"], len(g["con_rhs"]))
    assert relerr(sysm.apply_diagonal(g["x"]), g["apply_diagonal"]) < 1e-12
    assert relerr(sysm.apply_diagonal_inverse(g["x"]), g["apply_diagonal_inverse"]) < 1e-9
    assert relerr(sysm.apply_trace(g["x"]), g["apply_trace"]) < 1e-12
    assert relerr(sysm.apply_trace_transpose(g["trace_in"]), g["apply_trace_transpose"]) < 1e-12
    x, lam, r, it = O.solve_schur(sysm, g["rhs"], g["con_rhs"], 2000, 1e-30, 1e-12)
    assert relerr(x, g["sol_dense"]) < 1e-8
    assert abs(it - int(g["iters"])) <= 5


def test_oracle_vs_live_reference():
    ref = load_ref()
    if ref is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(0)
    for p, k in ((2, 3), (6, 3), (8, 0)):
        b = O.basis(p, p + k + 1)
        rule = ref.IntegrationRule1D(p + k)
        b1 = ref.Basis1D(p, rule)
        assert relerr(b.node, np.array(b1.node)) < 1e-14
        assert relerr(b.edge, np.array(b1.edge)) < 1e-13
        # (compute_gll itself raises RuntimeWarning-as-error for non-converged nodes,
        #  gauss_lobatto.c:249-255; the rule object exposes the same nodes without that path)
        assert np.array_equal(b.pnts, np.array(rule.nodes)) and np.array_equal(b.wgts, np.array(rule.weights))
        quad = F.QUADS[0] + rng.uniform(-0.1, 0.1, (4, 2))
        sp = ref.ElementFemSpace2D(ref.Basis2D(b1, b1), quad)
        for fo in (1, 2, 3):
            mref = ref.compute_element_mass_matrix(fo, quad, ref.Basis2D(b1, b1), False)
            assert relerr(O.mass(b, quad, fo), mref) < 1e-13
        a = rng.standard_normal((30, 30)) + 5 * np.eye(30)
        assert relerr(O.invert(a), ref._compute_matrix_inverse(a)) < 1e-12
        for t in range(4):
            e = O.incidence(t, p)
            m = rng.standard_normal((e.shape[1], 5))
            got = ref.check_incidence(m, p, t % 2, t >= 2, False)
            assert relerr(e @ m, got) < 1e-15
            mr = rng.standard_normal((5, e.shape[0]))
            assert relerr(mr @ e, ref.check_incidence(mr, p, t % 2, t >= 2, True)) < 1e-15
        del sp


def test_dual_dofs_vs_golden():
    """oracle.dual_dofs against the reference's element_dual_dofs (tests/golden/dual_dofs.npz)."""
    import parity_cases as P

    g = golden("dual_dofs.npz")
    for p, k in ((1, 2), (3, 3), (4, 2), (6, 3)):
        ob = O.basis(p, p + k + 1)
        for iq in (2, 4, 5):
            x, y = O.physical_coordinates(F.QUADS[iq], ob.pnts)
            for fo, fn in ((1, P._dual_f_scalar), (2, P._dual_f_vector), (3, P._dual_f_scalar)):
                assert relerr(O.dual_dofs(ob, F.QUADS[iq], fo, fn(x, y)), g[f"p{p}k{k}q{iq}_f{fo}"]) < 1e-13
