"""The reference's element-block solver driven from arrays.  TEST INFRASTRUCTURE ONLY (checker / CPU baseline).

The arithmetic is the reference's own C code (``LinearSystem``, ``DenseVector``, ``TraceVector`` of the compiled
extension ``oracle/_ref/_mfv2d_ref.so``: src/algebra/sparse_system.c, system_objects.c).  Only the two thin Python
layers above it are restated here, because ``/root/reference/python`` does not travel to the GPU box:

* ``build_system``      = ``LinearSystem.__new__`` (python/mfv2d/solving.py:41-97) fed from a COO constraint matrix
* ``solve_schur``       = ``solve_schur_iterative`` (:439-500) with ``cg_general`` (:350-436), statement for statement,
                          every vector operation a call into the reference's C extension

Only ``tests/`` and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import this module.
"""
from __future__ import annotations

import numpy as np


def build_system(ref, mats, row, col, val, n_trace):
    """``LinearSystem(blocks, offsets, indices)`` of the reference extension (system_objects.c:73-199) from dense blocks
    ``mats[n, N_e, N_e]`` and the constraint matrix in COO form over global dense columns (uniform block size)."""
    n, ne = mats.shape[0], mats.shape[1]
    elem = col // ne
    order = np.lexsort((col, elem, row))
    r_s, e_s, c_s, v_s = row[order], elem[order], col[order] - elem[order] * ne, val[order]
    blocks = tuple((np.ascontiguousarray(mats[e]), ref.MatrixCRS(n_trace, ne)) for e in range(n))
    cuts = np.flatnonzero(np.diff(r_s * n + e_s)) + 1
    starts, stops = np.concatenate([[0], cuts]), np.concatenate([cuts, [r_s.size]])
    next_row = np.zeros(n, np.int64)
    lens, elems = np.zeros(n_trace, np.int64), []
    for lo, hi in zip(starts, stops):
        e, r = int(e_s[lo]), int(r_s[lo])
        for empty in range(next_row[e], r):  # rows must be built in order (crs_matrix.c:106-112)
            blocks[e][1].build_row(empty)
        blocks[e][1].build_row(r, ref.SparseVector.from_entries(ne, c_s[lo:hi].astype(np.uint32), v_s[lo:hi]))
        next_row[e] = r + 1
        lens[r] += 1
        elems.append(e)
    for e in range(n):
        for empty in range(next_row[e], n_trace):
            blocks[e][1].build_row(empty)
    offsets = np.concatenate([[0], np.cumsum(lens)]).astype(np.uint32)
    return ref.LinearSystem(blocks, offsets, np.array(elems, np.uint32))


def solve_schur(ref, system, rhs_blocks, phi, max_iter, atol, rtol):
    """python/mfv2d/solving.py:439-500 + :350-436.  Returns (x blocks merged, lambda merged, residual, iterations)."""
    DV, TV = ref.DenseVector, ref.TraceVector
    rhs = DV(system, *rhs_blocks)
    constraints = TV(system, np.ascontiguousarray(phi, np.float64))
    inv_a_y = DV(system)
    system.apply_diagonal_inverse(rhs, inv_a_y)
    trace_rhs = TV(system)
    system.apply_trace(inv_a_y, trace_rhs)
    TV.subtract(trace_rhs, constraints, trace_rhs, 1.0)
    tmp_d1, tmp_d2 = DV(system), DV(system)

    def apply_system(v):  # apply_full_trace_system (:99-123), in place
        system.apply_trace_transpose(v, tmp_d1)
        system.apply_diagonal_inverse(tmp_d1, tmp_d2)
        system.apply_trace(tmp_d2, v)

    # cg_general
    res_mag2 = TV.dot(trace_rhs, trace_rhs)
    tol = atol if np.sqrt(res_mag2) * rtol > atol else np.sqrt(res_mag2) * rtol
    ap, p, res, x = trace_rhs.copy(), trace_rhs.copy(), trace_rhs.copy(), TV(system)
    iter_cnt = 0
    for iter_cnt in range(max_iter):
        apply_system(ap)
        apa = TV.dot(ap, p)
        alpha = res_mag2 / apa
        TV.add(x, p, x, alpha)
        TV.subtract(res, ap, res, alpha)
        new_res_mag2 = TV.dot(res, res)
        if new_res_mag2 < tol**2:
            res_mag2 = new_res_mag2
            break
        beta = new_res_mag2 / res_mag2
        res_mag2 = new_res_mag2
        ap.set_from(res)
        TV.add(ap, p, ap, beta)
        p.set_from(ap)
    system.apply_trace_transpose(x, tmp_d1)
    system.apply_diagonal_inverse(tmp_d1, tmp_d2)
    DV.subtract(inv_a_y, tmp_d2, inv_a_y, 1.0)
    return inv_a_y.as_merged(), x.as_merged(), float(np.sqrt(res_mag2)), iter_cnt
