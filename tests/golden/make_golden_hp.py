"""Fixture for BASELINE.json configs[3]: an h/p-refined, NON-conforming mesh with mixed orders 2..8.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_hp.py
The reference itself produces everything: `unit_square_mesh(3, 3, 2)` split depth-first twice with the counter predicate
of examples/refinement/plot_direct_poisson_refined_pre.py:129-144 (three of four visited leaves are split), child orders
cycling 2..8; mixed Poisson (q1, u2); element matrices from compute_element_matrix, the constraint rows (including the
hanging-side rows of continuity.py:192-318) from connect_elements, the right-hand side from compute_element_rhs, the
operators of the extension's LinearSystem on a random vector and the solution of solve_schur_iterative.
Writes tests/golden/hp_refined.npz (orders 2..8) and tests/golden/hp_refined_345.npz (orders 3..5).
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
from mfv2d._mfv2d import DenseVector, ElementFemSpace2D, TraceVector, compute_element_matrix  # noqa: E402
from mfv2d.continuity import connect_elements  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache  # noqa: E402
from mfv2d.solve_system import compute_element_rhs  # noqa: E402
from mfv2d.solving import ConvergenceSettings, LinearSystem, solve_schur_iterative  # noqa: E402
from mfv2d.system import KFormSystem  # noqa: E402

import mfv2d.continuity as _cont  # noqa: E402

# compute_gll's optional arguments are left uninitialised by the reference's own argument parser in the plain-gcc build of
# oracle/build_ref.sh (every node is then reported "non-converged" and PyErr_Format(PyExc_RuntimeWarning) leaves an exception
# set, src/polynomials/gauss_lobatto.c:243-249): pass the documented defaults explicitly - same nodes, bit for bit
_gll = _cont.compute_gll
_cont.compute_gll = lambda order: _gll(order, 10, 1e-15)

O1, O2 = UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def f_exact(x, y):
    return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)


def predicate(m, ie, counter):
    counter[0] += 1
    if (counter[0] & 3) != 0:
        out = []
        for _ in range(4):
            o = counter[2] + counter[1] % counter[3]
            counter[1] += 1
            out.append((o, o))
        return tuple(out)
    return None


def main(name="hp_refined", lowest=2, count=7):
    u, q = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v, p_ = u.weight, q.weight
    system = KFormSystem(p_ @ q + p_.derivative @ u == 0, v @ q.derivative == v @ f_exact, sorting=lambda f: f.order)
    comp = CompiledSystem(system)
    mesh = unit_square_mesh(3, 3, 2, F.deform).split_depth_first(2, predicate, [0, 0, lowest, count])
    leaf = list(mesh.get_leaf_indices())
    cache = FemCache(2)
    spaces = [ElementFemSpace2D(cache.get_basis2d(*mesh.get_leaf_orders(i)), np.astype(mesh.get_leaf_corners(i), np.float64, copy=False))
              for i in leaf]
    corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
    orders = np.array([mesh.get_leaf_orders(i) for i in leaf], np.uint32)
    mats = [compute_element_matrix(system.unknown_forms, comp.lhs_codes, s) for s in spaces]
    vectors = [compute_element_rhs(system, s) for s in spaces]
    cons = connect_elements(system.unknown_forms, mesh)
    rows, elems, dofs, vals = [], [], [], []
    for ic, c in enumerate(cons):
        for ec in c.element_constraints:
            for d, cf in zip(ec.dofs, ec.coeffs):
                rows.append(ic), elems.append(ec.i_e), dofs.append(int(d)), vals.append(float(cf))
    # The extension's LinearSystem cannot take these rows (the hanging-side DoF lists are not sorted:
    # "ValueError: Entry indices ... are not sorted", python/mfv2d/solving.py:80), so the reference solution is the one of its
    # production path: the global saddle-point matrix solved by SciPy (python/mfv2d/solve_system_2d.py:258-282)
    import scipy.sparse as sp
    import scipy.sparse.linalg as spl

    sizes = np.array([m.shape[0] for m in mats])
    off = np.concatenate([[0], np.cumsum(sizes)])
    A = sp.block_diag(mats, format="csr")
    cols = off[np.array(elems)] + np.array(dofs)
    N = sp.csr_array((np.array(vals), (np.array(rows), cols)), shape=(len(cons), off[-1]))
    K = sp.block_array([[A, N.T], [N, None]], format="csc")
    rhs = np.concatenate(vectors)
    con_rhs = np.array([c.rhs for c in cons], np.float64)
    sol = spl.spsolve(K, np.concatenate([rhs, con_rhs]))
    resid = np.linalg.norm(K @ sol - np.concatenate([rhs, con_rhs])) / np.linalg.norm(rhs)
    rng = np.random.default_rng(7)
    x = rng.standard_normal(off[-1])
    tin = rng.standard_normal(len(cons))
    Ainv_x = spl.spsolve(A.tocsc(), x)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), corners=corners, orders=orders, k=2,
        con_row=np.array(rows, np.uint32), con_elem=np.array(elems, np.uint32), con_dof=np.array(dofs, np.uint32), con_val=np.array(vals),
        con_rhs=con_rhs, rhs=rhs, x=x, trace_in=tin, apply_diagonal=A @ x, apply_diagonal_inverse=Ainv_x, apply_trace=N @ x,
        apply_trace_transpose=N.T @ tin, sol_dense=sol[:off[-1]], sol_trace=sol[off[-1]:], resid=resid,
        depth=np.array([mesh.get_element_depth(i) for i in leaf]), mat_first=mats[0], mat_last=mats[-1])
    iters = 0
    print(name + ".npz:", len(leaf), "leaves, orders", sorted(set(orders[:, 0].tolist())), "constraints", len(cons), "dense", sum(m.shape[0] for m in mats),
          "iters", iters, "resid", resid)
    per_row = np.bincount(np.array(rows))
    print("entries per constraint row: max", per_row.max(), "rows with > 2 entries", int((per_row > 2).sum()))


if __name__ == "__main__":
    main()  # orders 2..8: hanging-side coefficients up to 4e8, cond(N A^-1 N^T) ~ 1e18
    main("hp_refined_345", 3, 3)  # orders 3..5: neighbouring orders differ by at most 2
