"""Array-based structured quad meshes and their continuity constraints.

The reference builds meshes through Python objects (``mfv2d.examples.unit_square_mesh``,
python/mfv2d/examples.py:11-73) and its constraints through a recursive per-edge Python walk
(``connect_elements``, python/mfv2d/continuity.py:610-759).  For the synthetic structured configs of
BASELINE.json (up to 10^6 elements) those are generated here directly as flat arrays, reproducing the
reference's leaf order, row order, DoF order and coefficients exactly (pinned against reference output
in tests/golden/solve_*.npz and constraints_*.npz).
"""
from __future__ import annotations

from collections.abc import Callable

import numpy as np

FORM_ORDER_0, FORM_ORDER_1, FORM_ORDER_2 = 1, 2, 3


def unit_square_corners(nh: int, nv: int, deformation: Callable | None = None) -> np.ndarray:
    """Corner coordinates ``(nh*nv, 4, 2)`` of ``unit_square_mesh(nh, nv, ...)``: element ``i + nh*j``,
    corners counter-clockwise from bottom-left, nodes on the uniform grid of [-1, 1]^2 passed through
    ``deformation(xi, eta) -> (x, y)`` (examples.py:40-47, 55-66)."""
    xi, eta = np.meshgrid(np.linspace(-1, +1, nh + 1), np.linspace(-1, +1, nv + 1))
    if deformation is not None:
        px, py = deformation(xi, eta)
        xi = np.array(px, np.float64)
        eta = np.array(py, np.float64)
    pts = np.stack((xi, eta), axis=-1)  # (nv+1, nh+1, 2)
    c = np.empty((nv, nh, 4, 2))
    c[:, :, 0] = pts[:-1, :-1]
    c[:, :, 1] = pts[:-1, 1:]
    c[:, :, 2] = pts[1:, 1:]
    c[:, :, 3] = pts[1:, :-1]
    return np.ascontiguousarray(c.reshape(nh * nv, 4, 2))


def form_sizes(form_orders, p: int) -> np.ndarray:
    size = {FORM_ORDER_0: (p + 1) ** 2, FORM_ORDER_1: 2 * p * (p + 1), FORM_ORDER_2: p * p}
    return np.array([size[int(o)] for o in form_orders], np.int64)


def continuity_constraints_1form(nh: int, nv: int, p: int, form_orders) -> tuple[np.ndarray, np.ndarray, np.ndarray, int]:
    """Continuity rows of the 1-form unknowns on the conforming ``nh x nv`` mesh of uniform order p, in
    the reference's order (continuity.py:650-683, 498-508): for each 1-form of the system, first the
    interior horizontal lines (line index ``nh*j + i``, j = 1..nv-1), then the interior vertical lines
    (``j*(nh+1) + i``, i = 1..nh-1); ``p`` rows per line.

    Returns COO arrays (row, global dense column, value) sorted by (row, column) and the row count."""
    sizes = form_sizes(form_orders, p)
    offs = np.concatenate([[0], np.cumsum(sizes)])
    ne = int(offs[-1])
    nh_edges = p * (p + 1)
    rows, cols, vals = [], [], []
    row0 = 0
    k = np.arange(p)[::-1]  # DoFs along a side are traversed in reverse
    for f, order in enumerate(form_orders):
        if int(order) != FORM_ORDER_1:
            continue
        o1 = int(offs[f])
        # horizontal interior lines: element below (top side, +1) and element above (bottom side, -1)
        jj, ii = np.meshgrid(np.arange(1, nv), np.arange(nh), indexing="ij")
        e_below = (ii + nh * (jj - 1)).ravel()
        e_above = (ii + nh * jj).ravel()
        n_l = e_below.size
        r = row0 + (np.arange(n_l)[:, None] * p + np.arange(p)[None, :])
        top = o1 + p * p + k  # h(p, c)
        bot = o1 + k  # h(0, c)
        rows += [r.ravel(), r.ravel()]
        cols += [(e_below[:, None] * ne + top[None, :]).ravel(), (e_above[:, None] * ne + bot[None, :]).ravel()]
        vals += [np.full(n_l * p, 1.0), np.full(n_l * p, -1.0)]
        row0 += n_l * p
        # vertical interior lines: element on the right (left side, +1) and on the left (right side, -1)
        jj, ii = np.meshgrid(np.arange(nv), np.arange(1, nh), indexing="ij")
        e_right = (ii + nh * jj).ravel()
        e_left = (ii - 1 + nh * jj).ravel()
        n_l = e_right.size
        r = row0 + (np.arange(n_l)[:, None] * p + np.arange(p)[None, :])
        left = o1 + nh_edges + (p + 1) * k  # v(r, 0)
        right = o1 + nh_edges + (p + 1) * k + p  # v(r, p)
        rows += [r.ravel(), r.ravel()]
        cols += [(e_right[:, None] * ne + left[None, :]).ravel(), (e_left[:, None] * ne + right[None, :]).ravel()]
        vals += [np.full(n_l * p, 1.0), np.full(n_l * p, -1.0)]
        row0 += n_l * p
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0), 0
    row = np.concatenate(rows).astype(np.int64)
    col = np.concatenate(cols).astype(np.int64)
    val = np.concatenate(vals)
    order = np.lexsort((col, row))
    return row[order], col[order], val[order], row0


def strip_partition(nh: int, nv: int, n_ranks: int) -> list[tuple[int, int]]:
    """Contiguous ranges of leaf indices (horizontal strips of whole element rows)."""
    rows = [nv * r // n_ranks for r in range(n_ranks + 1)]
    return [(rows[r] * nh, rows[r + 1] * nh) for r in range(n_ranks)]


class ConstraintPartition:
    """What one rank needs of a global constraint matrix ``N`` (COO over global dense columns) when the
    elements - and with them contiguous ranges of dense columns - are split over ``n_ranks`` processes
    (SURVEY.md 8e).

    ``row, col, val``  the rank's rows of ``N`` restricted to its own columns, with LOCAL row and column
                       indices, sorted by (row, col); ``n_trace`` local multipliers
    ``global_rows``    global multiplier index of every local row (increasing)
    ``shared_index``   local rows that other ranks hold too; ``slot`` their position in the packed
                       interface buffer of ``n_slots`` doubles (same numbering on every rank)
    ``owned``          1 where this rank is the lowest rank holding the multiplier (it alone counts the
                       multiplier in dot products, which replaces TraceVector.dot's division by the sharing
                       count, src/algebra/system_objects.c:1690-1696)
    ``holder_ptr, holder_rank``  CSR over ``shared_index``: the OTHER ranks that hold each shared multiplier, ascending
                       (what the NVLink peer exchange needs: whom to write to, whose contributions to add)
    """

    def __init__(self, row, col, val, n_trace, global_rows, shared_index, slot, n_slots, owned, holder_ptr=None, holder_rank=None):
        self.row, self.col, self.val, self.n_trace = row, col, val, int(n_trace)
        self.global_rows, self.shared_index, self.slot = global_rows, shared_index, slot
        self.n_slots, self.owned = int(n_slots), owned
        self.holder_ptr, self.holder_rank = holder_ptr, holder_rank


def partition_constraints(row: np.ndarray, col: np.ndarray, val: np.ndarray, n_trace: int, dense_bounds,
                          rank: int) -> ConstraintPartition:
    """Split the global constraint matrix for ``rank``.  ``dense_bounds[r]`` is the first global dense
    column of rank ``r`` (``len == n_ranks + 1``).  Every rank calls this with the same global arrays and
    obtains consistent interface slots without any communication: a multiplier is *shared* when its row
    touches the columns of more than one rank (conforming edges on a cut: 2 ranks; a gauge row from
    ``constrained_forms``: every rank), and the shared multipliers are numbered in increasing global order."""
    row = np.asarray(row, np.int64)
    col = np.asarray(col, np.int64)
    val = np.asarray(val, np.float64)
    bounds = np.asarray(dense_bounds, np.int64)
    entry_rank = np.searchsorted(bounds, col, side="right") - 1
    lo = np.full(n_trace, np.iinfo(np.int64).max, np.int64)
    hi = np.full(n_trace, -1, np.int64)
    np.minimum.at(lo, row, entry_rank)
    np.maximum.at(hi, row, entry_rank)
    is_shared = hi > lo
    slot_of = np.cumsum(is_shared) - 1  # valid where is_shared
    mine = entry_rank == rank
    g_rows = np.unique(row[mine])
    local_of = np.full(n_trace, -1, np.int64)
    local_of[g_rows] = np.arange(g_rows.size)
    l_row = local_of[row[mine]]
    l_col = col[mine] - bounds[rank]
    l_val = val[mine]
    order = np.lexsort((l_col, l_row))
    sh = np.flatnonzero(is_shared[g_rows])
    owned = (lo[g_rows] == rank).astype(np.uint8)
    # the other holders of this rank's shared multipliers: unique (row, rank) pairs of the entries of shared rows
    mine_shared = np.zeros(n_trace, bool)
    mine_shared[g_rows[sh]] = True
    sel = mine_shared[row] & (entry_rank != rank)
    pairs = np.unique(np.stack([row[sel], entry_rank[sel]], axis=1), axis=0) if sel.any() else np.zeros((0, 2), np.int64)
    pos_of = np.full(n_trace, -1, np.int64)
    pos_of[g_rows[sh]] = np.arange(sh.size)
    counts = np.bincount(pos_of[pairs[:, 0]], minlength=sh.size) if sh.size else np.zeros(0, np.int64)
    holder_ptr = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    holder_rank = pairs[:, 1].astype(np.int32)  # np.unique sorts by (row, rank): rows increasing = shared order, ranks ascending
    return ConstraintPartition(l_row[order], l_col[order], l_val[order], g_rows.size, g_rows, sh.astype(np.int64),
                               slot_of[g_rows[sh]].astype(np.int64), int(is_shared.sum()), owned, holder_ptr, holder_rank)


def coarse_hierarchy_1form(nh: int, nv: int, p: int, form_orders, mid_sizes=(4,), exact_size: int = 16, strip_rows=None):
    """Aggregate hierarchy of the multilevel preconditioner of the Schur-complement CG (precond.cu) for the multipliers of
    :func:`continuity_constraints_1form` (same global row order).  A multiplier belongs to the aggregate of the element
    below / to the left of its line.  Levels ``mid_sizes`` (aggregates of ``H x H`` elements, confined to one rank's strip
    when ``strip_rows`` = first element row of every rank + ``[nv]`` is given), then the exact level of
    ``exact_size x exact_size`` elements, numbered globally.  ``z`` is the near-null-space vector of ``S = N A^-1 N^T``
    for the mixed formulations, the multipliers of ``u = const``: ``+1`` on horizontal lines, ``-1`` on vertical ones
    (sign convention of continuity.py:498-501 with the outward orientation of the element boundary DoFs).

    Returns ``(agg, z, n_agg)``: ``agg[level, global multiplier]`` (compact ids per level), ``n_agg[level]``."""
    strip_rows = np.asarray(strip_rows if strip_rows is not None else [0, nv], np.int64)
    ei, ej, zs = [], [], []
    for order in form_orders:
        if int(order) != FORM_ORDER_1:
            continue
        jj, ii = np.meshgrid(np.arange(1, nv), np.arange(nh), indexing="ij")  # horizontal lines: element below
        ei.append(np.repeat(ii.ravel(), p))
        ej.append(np.repeat((jj - 1).ravel(), p))
        zs.append(np.ones(jj.size * p))
        jj, ii = np.meshgrid(np.arange(nv), np.arange(1, nh), indexing="ij")  # vertical lines: element to the left
        ei.append(np.repeat((ii - 1).ravel(), p))
        ej.append(np.repeat(jj.ravel(), p))
        zs.append(-np.ones(jj.size * p))
    if not ei:
        return np.zeros((0, 0), np.int32), np.zeros(0), []
    ei, ej, z = np.concatenate(ei), np.concatenate(ej), np.concatenate(zs)
    strip = np.searchsorted(strip_rows, ej, side="right") - 1
    jl = ej - strip_rows[strip]
    levels, n_agg = [], []
    for H in mid_sizes:
        key = (strip * (nv + 1) + jl // H) * (nh + 1) + ei // H
        used, compact = np.unique(key, return_inverse=True)
        levels.append(compact.astype(np.int32))
        n_agg.append(int(used.size))
    if exact_size:
        key = (ej // exact_size) * (nh + 1) + ei // exact_size
        used, compact = np.unique(key, return_inverse=True)  # aggregates without a multiplier would make P^T S P singular
        levels.append(compact.astype(np.int32))
        n_agg.append(int(used.size))
    return np.stack(levels), z, n_agg


def localize_hierarchy(agg: np.ndarray, z: np.ndarray, n_agg, part: "ConstraintPartition", exact_last: bool = True):
    """This rank's view of a global hierarchy: rows = its multipliers; on the rank-local levels the shared multipliers are
    left out (-1) and the ids are compacted; the exact level keeps its global ids."""
    loc = np.ascontiguousarray(agg[:, part.global_rows]).astype(np.int32)
    counts = list(n_agg)
    n_local = loc.shape[0] - (1 if exact_last else 0)
    for lvl in range(n_local):
        a = loc[lvl]
        a[part.shared_index] = -1
        keep = a >= 0
        used, compact = np.unique(a[keep], return_inverse=True)
        a[keep] = compact
        counts[lvl] = int(used.size)
    return loc, np.ascontiguousarray(z[part.global_rows]), counts
