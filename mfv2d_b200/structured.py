"""Array-based structured quad meshes and their continuity constraints.

The reference builds meshes through Python objects (``mfv2d.examples.unit_square_mesh``,
python/mfv2d/examples.py:11-73) and its constraints through a recursive per-edge Python walk
(``connect_elements``, python/mfv2d/continuity.py:610-759).  For the synthetic structured configs of
BASELINE.json (up to 10^6 elements) those are generated here directly as flat arrays, reproducing the
reference's leaf order, row order, DoF order and coefficients exactly (pinned against reference output
in tests/golden/solve_*.npz, constraints.npz and hp_refined*.npz, each written by a committed make_golden*.py).
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


def _form_offsets(form_orders, p: int):
    sizes = form_sizes(form_orders, p)
    return np.concatenate([[0], np.cumsum(sizes)]), int(sizes.sum())


def _interior_lines(nh: int, nv: int):
    """(first element, second element, horizontal?) of every interior line in the reference's line order: horizontal lines
    ``nh*j + i`` (j = 1..nv-1; first = element below, its TOP side; second = element above, its BOTTOM side), then vertical
    lines ``j*(nh+1) + i`` (i = 1..nh-1; first = element to the right, its LEFT side; second = element to the left, its
    RIGHT side) - the ``dual_line.begin / end`` order of continuity.py:650-683 on ``unit_square_mesh``."""
    jj, ii = np.meshgrid(np.arange(1, nv), np.arange(nh), indexing="ij")
    h1, h2 = (ii + nh * (jj - 1)).ravel(), (ii + nh * jj).ravel()
    jj, ii = np.meshgrid(np.arange(nv), np.arange(1, nh), indexing="ij")
    v1, v2 = (ii + nh * jj).ravel(), (ii - 1 + nh * jj).ravel()
    return (h1, h2), (v1, v2)


def continuity_constraints(nh: int, nv: int, p: int, form_orders) -> tuple[np.ndarray, np.ndarray, np.ndarray, int]:
    """All continuity rows of ``connect_elements`` (python/mfv2d/continuity.py:610-759) on the conforming ``nh x nv`` mesh of
    uniform order ``p``, in the reference's order: for every unknown form in turn - a 0-form contributes its edge rows
    (``p - 1`` interior nodes per interior line, coefficient +1 on the first element and -1 on the second, second side
    traversed in reverse, :491-519) followed by its corner chains (one row per consecutive pair of the elements around a
    mesh node in the order top-left, bottom-right, bottom-left, top-right, :321-368, 689-716); a 1-form contributes ``p``
    rows per interior line with sign ``sgn1 * sgn2`` (:498-501); 2-forms contribute nothing.

    Returns COO arrays (row, global dense column, value) sorted by (row, column) and the row count; pinned bit for bit to
    the reference's output in tests/golden/constraints.npz."""
    offs, ne = _form_offsets(form_orders, p)
    (h1, h2), (v1, v2) = _interior_lines(nh, nv)
    np1 = p + 1
    rows, cols, vals = [], [], []
    row0 = 0

    def add(n_rows, e_first, d_first, e_second, d_second):
        nonlocal row0
        r = row0 + np.arange(n_rows)
        rows.extend([r, r])
        cols.extend([e_first * ne + d_first, e_second * ne + d_second])
        vals.extend([np.full(n_rows, 1.0), np.full(n_rows, -1.0)])
        row0 += n_rows

    for f, order in enumerate(form_orders):
        o = int(offs[f])
        order = int(order)
        if order == FORM_ORDER_0:
            c = np.arange(p - 1, 0, -1)  # interior nodes of a side, traversed counter-clockwise on the first element
            if c.size:
                # horizontal lines: top side of the element below, node (p, c); bottom side of the element above, node (0, c)
                add(h1.size * c.size, np.repeat(h1, c.size), np.tile(o + p * np1 + c, h1.size), np.repeat(h2, c.size), np.tile(o + c, h1.size))
                # vertical lines: left side of the right element, node (r, 0); right side of the left element, node (r, p)
                add(v1.size * c.size, np.repeat(v1, c.size), np.tile(o + c * np1, v1.size), np.repeat(v2, c.size), np.tile(o + c * np1 + p, v1.size))
            # corner chains, mesh nodes in index order j*(nh+1) + i: elements around the node in the order
            # top-left, bottom-right, bottom-left, top-right, each with ITS corner DoF at the node
            jj, ii = np.meshgrid(np.arange(nv + 1), np.arange(nh + 1), indexing="ij")
            jj, ii = jj.ravel(), ii.ravel()
            cand = [  # (element column, element row, local DoF)
                (ii - 1, jj, p),  # element up-left of the node: its bottom-right corner
                (ii, jj - 1, p * np1),  # element down-right: its top-left corner
                (ii - 1, jj - 1, np1 * np1 - 1),  # element down-left: its top-right corner
                (ii, jj, 0),  # element up-right: its bottom-left corner
            ]
            elem = np.stack([np.where((ci >= 0) & (ci < nh) & (cj >= 0) & (cj < nv), ci + nh * cj, -1) for ci, cj, _ in cand], axis=1)
            dof = np.array([d for _, _, d in cand])
            present = elem >= 0
            # compact every node's list to the front (stable), then chain consecutive pairs
            order_idx = np.argsort(~present, axis=1, kind="stable")
            elem_s = np.take_along_axis(elem, order_idx, axis=1)
            dof_s = dof[order_idx]
            count = present.sum(axis=1)
            node_ids = np.flatnonzero(count > 1)
            reps = count[node_ids] - 1
            n_rows = int(reps.sum())
            if n_rows:
                node_rep = np.repeat(node_ids, reps)
                k_in = np.arange(n_rows) - np.repeat(np.cumsum(reps) - reps, reps)
                add(n_rows, elem_s[node_rep, k_in], o + dof_s[node_rep, k_in], elem_s[node_rep, k_in + 1], o + dof_s[node_rep, k_in + 1])
        elif order == FORM_ORDER_1:
            k = np.arange(p)[::-1]  # DoFs along a side are traversed in reverse
            nh_edges = p * np1
            # horizontal lines: top side of the element below h(p, c) (+1), bottom side of the element above h(0, c) (-1)
            add(h1.size * p, np.repeat(h1, p), np.tile(o + p * p + k, h1.size), np.repeat(h2, p), np.tile(o + k, h1.size))
            # vertical lines: left side of the right element v(r, 0) (+1), right side of the left element v(r, p) (-1)
            add(v1.size * p, np.repeat(v1, p), np.tile(o + nh_edges + np1 * k, v1.size), np.repeat(v2, p), np.tile(o + nh_edges + np1 * k + p, v1.size))
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0), 0
    row = np.concatenate(rows).astype(np.int64)
    col = np.concatenate(cols).astype(np.int64)
    val = np.concatenate(vals)
    order_rc = np.lexsort((col, row))
    return row[order_rc], col[order_rc], val[order_rc], row0


def continuity_constraints_1form(nh: int, nv: int, p: int, form_orders) -> tuple[np.ndarray, np.ndarray, np.ndarray, int]:
    """The rows of :func:`continuity_constraints` for systems without 0-forms (kept as the historic entry point)."""
    return continuity_constraints(nh, nv, p, form_orders)


def gauge_constraint(n_elem: int, p: int, form_orders, form_index: int) -> tuple[np.ndarray, np.ndarray]:
    """Columns and coefficients of ONE ``constrained_forms`` row (continuity.py:775-794): every DoF of the form in every
    element, coefficient 1."""
    offs, ne = _form_offsets(form_orders, p)
    d = np.arange(int(offs[form_index]), int(offs[form_index + 1]))
    col = (np.arange(n_elem)[:, None] * ne + d[None, :]).ravel()
    return col.astype(np.int64), np.ones(col.size)


def boundary_lines(nh: int, nv: int):
    """``mesh.boundary_indices`` of ``unit_square_mesh`` as (element, side) pairs in the reference's order: bottom lines,
    top lines, then per element row its left and right line.  Sides: 1 bottom, 2 right, 3 top, 4 left."""
    out = [(i, 1) for i in range(nh)] + [(i + nh * (nv - 1), 3) for i in range(nh)]
    for j in range(nv):
        out += [(nh * j, 4), (nh - 1 + nh * j, 2)]
    return out


def _side_dofs(side: int, order: int, p: int) -> np.ndarray:
    """element_boundary_dofs (python/mfv2d/mimetic2d.py:712-800) for ``order_1 == order_2 == p``."""
    np1 = p + 1
    if order == FORM_ORDER_1:
        return {1: np.arange(p), 2: p * np1 + p + np.arange(p) * np1, 3: (p * p + np.arange(p))[::-1], 4: (p * np1 + np.arange(p) * np1)[::-1]}[side]
    return {1: np.arange(np1), 2: p + np.arange(np1) * np1, 3: (np1 * p + np.arange(np1))[::-1], 4: (np.arange(np1) * np1)[::-1]}[side]


def strong_bc_constraints(nh: int, nv: int, p: int, form_orders, form_index: int, func: Callable, corners: np.ndarray, roots: np.ndarray,
                          rule_nodes: np.ndarray, rule_weights: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """Strong boundary-condition rows on the whole boundary of the structured mesh (``mesh_boundary_conditions`` /
    ``_element_strong_boundary_condition``, python/mfv2d/boundary.py:228-498): one row per constrained DoF.

    0-forms take ``func`` at the GLL nodes of the (straight) element side, a mesh node already constrained by an earlier
    line is skipped (:455-470); 1-forms take the flux of ``func`` through each sub-edge, integrated with the 1-D rule of
    the basis.  ``roots`` = Basis1D.roots of order p, ``rule_*`` = its IntegrationRule1D.  Returns (global dense columns,
    values) in the reference's row order."""
    offs, ne = _form_offsets(form_orders, p)
    order = int(form_orders[form_index])
    o = int(offs[form_index])
    set_nodes = np.zeros((nv + 1, nh + 1), bool)
    cols, vals = [], []
    for e, side in boundary_lines(nh, nv):
        ei, ej = e % nh, e // nh
        p0, p1 = corners[e, side - 1], corners[e, side & 3]
        dx, dy = (p1[0] - p0[0]) / 2, (p1[1] - p0[1]) / 2
        xv = (p1[0] + p0[0]) / 2 + dx * roots
        yv = (p1[1] + p0[1]) / 2 + dy * roots
        dofs = _side_dofs(side, order, p) + o
        if order == FORM_ORDER_0:
            v = np.asarray(func(xv, yv), np.float64) * np.ones(p + 1)
            # mesh nodes of the side's first / last corner (corner k of element (ei, ej): 0 BL, 1 BR, 2 TR, 3 TL)
            node_of = [(ej, ei), (ej, ei + 1), (ej + 1, ei + 1), (ej + 1, ei)]
            n0, n1 = node_of[side - 1], node_of[side & 3]
            lo, hi = (1 if set_nodes[n0] else 0), (p if set_nodes[n1] else p + 1)
            set_nodes[n0] = set_nodes[n1] = True
            dofs, v = dofs[lo:hi], v[lo:hi]
        else:
            ndir = 2 * ((side & 2) >> 1) - 1
            v = np.zeros(p)
            for i in range(p):
                xc = (xv[i + 1] + xv[i]) / 2 + (xv[i + 1] - xv[i]) / 2 * rule_nodes
                yc = (yv[i + 1] + yv[i]) / 2 + (yv[i + 1] - yv[i]) / 2 * rule_nodes
                ddx, ddy = (xv[i + 1] - xv[i]) / 2, (yv[i + 1] - yv[i]) / 2
                normal = ndir * np.array((ddy, -ddx))
                fv = np.asarray(func(xc, yc), np.float64)
                v[i] = np.sum((fv[..., 0] * normal[0] + fv[..., 1] * normal[1]) * rule_weights)
        cols.append(e * ne + dofs)
        vals.append(v)
    return np.concatenate(cols).astype(np.int64), np.concatenate(vals)


def weak_bc_contributions(nh: int, nv: int, p: int, form_orders, form_index: int, terms, corners: np.ndarray, node: np.ndarray, edge: np.ndarray,
                          rule_nodes: np.ndarray, rule_weights: np.ndarray, skip_lines=(), out: np.ndarray | None = None) -> np.ndarray:
    """Weak (natural) boundary terms of the equation weighted by form ``form_index`` on the boundary of the structured mesh
    (``_element_weak_boundary_condition``, python/mfv2d/boundary.py:109-214, called per boundary line by
    ``mesh_boundary_conditions`` :439-487): for every boundary side the boundary functions of ``terms`` = ``[(k, func), ...]``
    (the ``KBoundaryProjection`` pairs of the equation's right-hand side) are sampled at the 1-D rule's points on the straight
    side and integrated against the trace basis - 0-forms: ``-(f . t)`` with the nodal basis, 1-forms: ``-n_dir f`` with the
    edge basis.  ``node`` / ``edge`` are ``Basis1D(p, rule).node / .edge``; ``skip_lines`` lists the positions in
    :func:`boundary_lines` that carry a strong condition instead.  Adds into ``out`` ``(n, N_e)`` (created when ``None``):
    the array that is added to ``compute_element_rhs`` to form the explicit right-hand side of ``compute_linear_system``."""
    offs, ne = _form_offsets(form_orders, p)
    order = int(form_orders[form_index])
    if order not in (FORM_ORDER_0, FORM_ORDER_1):
        raise ValueError(f"Unknown/Invalid weak form order form_order={order}.")
    o = int(offs[form_index])
    if out is None:
        out = np.zeros((nh * nv, ne))
    skip = set(int(i) for i in skip_lines)
    for i_line, (e, side) in enumerate(boundary_lines(nh, nv)):
        if i_line in skip:
            continue
        p0, p1 = corners[e, side - 1], corners[e, side & 3]
        dx, dy = (p1[0] - p0[0]) / 2, (p1[1] - p0[1]) / 2
        xv = (p1[0] + p0[0]) / 2 + dx * rule_nodes
        yv = (p1[1] + p0[1]) / 2 + dy * rule_nodes
        ndir = 2 * ((side & 2) >> 1) - 1
        dofs = _side_dofs(side, order, p) + o
        vals = np.zeros(dofs.size)
        for k, func in terms:
            f = np.array(func(xv, yv), np.float64)
            if order == FORM_ORDER_0:
                basis, f = node, -(f[..., 0] * dx + f[..., 1] * dy) * rule_weights
            else:
                basis, f = edge, f * (-rule_weights * ndir)
            vals += np.sum(f[None, ...] * basis, axis=1) * k
        out[e, dofs] += vals
    return out


def system_constraints(nh: int, nv: int, p: int, form_orders, constrained_forms=(), strong_bcs=(), corners=None, roots=None, rule_nodes=None,
                       rule_weights=None):
    """The whole constraint matrix of ``add_system_constraints`` (python/mfv2d/continuity.py:762-872) for a conforming
    structured mesh, as COO over global dense columns plus the right-hand side: continuity rows, then one row per
    ``constrained_forms`` entry ``(value, form index)``, then the strong-BC rows of ``strong_bcs`` entries
    ``(form index, func)`` (one boundary condition per form, on the whole boundary)."""
    row, col, val, n = continuity_constraints(nh, nv, p, form_orders)
    rows, cols, vals, rhs = [row], [col], [val], [np.zeros(n)]
    for value, fi in constrained_forms:
        c, v = gauge_constraint(nh * nv, p, form_orders, fi)
        rows.append(np.full(c.size, n, np.int64))
        cols.append(c)
        vals.append(v)
        rhs.append(np.array([float(value)]))
        n += 1
    if strong_bcs:
        # rows are ordered by boundary line first, by equation second (boundary.py:439-470): merge the per-form row lists
        per_form = []
        for fi, func in sorted(strong_bcs, key=lambda t: t[0]):
            order = int(form_orders[fi])
            per_line = (p + 1) if order == FORM_ORDER_0 else p
            c, v = strong_bc_constraints(nh, nv, p, form_orders, fi, func, corners, roots, rule_nodes, rule_weights)
            # line index of every row: 1-forms have p rows per line; 0-forms a variable number (skipped nodes)
            if order == FORM_ORDER_1:
                line = np.repeat(np.arange(c.size // p), p)
            else:
                line = np.searchsorted(np.cumsum(_bc0_counts(nh, nv, p)), np.arange(c.size), side="right")
            per_form.append((fi, c, v, line))
            del per_line
        allc = np.concatenate([t[1] for t in per_form])
        allv = np.concatenate([t[2] for t in per_form])
        alll = np.concatenate([t[3] for t in per_form])
        allf = np.concatenate([np.full(t[1].size, t[0]) for t in per_form])
        seq = np.concatenate([np.arange(t[1].size) for t in per_form])
        order_bc = np.lexsort((seq, allf, alll))
        c, v = allc[order_bc], allv[order_bc]
        rows.append(n + np.arange(c.size))
        cols.append(c)
        vals.append(np.ones(c.size))
        rhs.append(v)
        n += c.size
    row, col, val = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    order_rc = np.lexsort((col, row))
    return row[order_rc], col[order_rc], val[order_rc], np.concatenate(rhs), n


def _bc0_counts(nh: int, nv: int, p: int) -> np.ndarray:
    """Number of 0-form strong-BC rows contributed by each boundary line (first-come ownership of the mesh nodes)."""
    set_nodes = np.zeros((nv + 1, nh + 1), bool)
    out = []
    for e, side in boundary_lines(nh, nv):
        ei, ej = e % nh, e // nh
        node_of = [(ej, ei), (ej, ei + 1), (ej + 1, ei + 1), (ej + 1, ei)]
        n0, n1 = node_of[side - 1], node_of[side & 3]
        out.append(p + 1 - int(set_nodes[n0]) - int(set_nodes[n1]))
        set_nodes[n0] = set_nodes[n1] = True
    return np.array(out, np.int64)


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


def strip_system_1form(nh: int, nv: int, p: int, form_orders, n_ranks: int, rank: int, mid_sizes=(4,), exact_size: int = 16):
    """One rank's constraint partition AND its view of the aggregate hierarchy in time and memory proportional to the STRIP.

    Equal, field by field, to ``partition_constraints(*continuity_constraints_1form(nh, nv, p, form_orders), bounds, rank)`` and
    ``localize_hierarchy(*coarse_hierarchy_1form(nh, nv, p, form_orders, mid_sizes, exact_size, strip_rows), part)`` for the
    strips of :func:`strip_partition` (checked in tests/test_distributed.py), but built from the closed form of the reference's
    row order (python/mfv2d/continuity.py:610-759 on ``unit_square_mesh``) instead of from the global matrix: on the 10^6-element
    mesh of BASELINE.json configs[4] every rank otherwise spends seconds generating and sorting 24 M global entries.
    Systems without 0-forms only (their corner chains have no closed form here); ``exact_size >= 2`` (every aggregate of the exact
    level then holds a multiplier).  Returns ``(part, (agg, z, n_agg), n_trace_global)``."""
    orders = [int(o) for o in form_orders]
    if any(o == FORM_ORDER_0 for o in orders):
        raise ValueError("strip_system_1form: systems with 0-forms need the global builder")
    if exact_size and exact_size < 2:
        raise ValueError("strip_system_1form: exact_size must be at least 2")
    offs, ne = _form_offsets(form_orders, p)
    strip_rows = np.array([nv * r // n_ranks for r in range(n_ranks + 1)], np.int64)
    j0, j1 = int(strip_rows[rank]), int(strip_rows[rank + 1])
    np1, nh_edges = p + 1, p * (p + 1)
    k = np.arange(p)[::-1]  # DoFs along a side are traversed in reverse
    kk = np.arange(p)
    n_h, n_v = (nv - 1) * nh, nv * (nh - 1)
    per_form = (n_h + n_v) * p
    n1 = sum(1 for o in orders if o == FORM_ORDER_1)
    owner = lambda j: int(np.searchsorted(strip_rows, j, side="right") - 1)  # noqa: E731  rank that holds element row j
    cuts = [j for j in range(1, nv) if owner(j - 1) != owner(j)]  # horizontal lines on a strip boundary, increasing
    cut_index = {j: b for b, j in enumerate(cuts)}
    e_rows, e_cols, e_vals = [], [], []  # entries in this rank's columns
    m_rows, m_ei, m_ej, m_z = [], [], [], []  # one record per multiplier held here
    s_rows, s_slot, s_other = [], [], []  # shared multipliers
    f1 = 0
    for f, order in enumerate(orders):
        if order != FORM_ORDER_1:
            continue
        o, base = int(offs[f]), f1 * per_form
        ja, jb = max(j0, 1), min(j1, nv - 1)
        if jb >= ja and j1 > j0:
            jj, ii = np.meshgrid(np.arange(ja, jb + 1), np.arange(nh), indexing="ij")
            jj, ii = jj.ravel(), ii.ravel()
            grow = base + (nh * (jj - 1) + ii)[:, None] * p + kk[None, :]
            m_rows.append(grow.ravel()), m_ei.append(np.repeat(ii, p)), m_ej.append(np.repeat(jj - 1, p)), m_z.append(np.ones(grow.size))
            below = jj - 1 >= j0  # top side of the element below: h(p, c), +1
            e_rows.append(grow[below].ravel())
            e_cols.append(((ii[below] + nh * (jj[below] - 1 - j0))[:, None] * ne + o + p * p + k[None, :]).ravel())
            e_vals.append(np.ones(int(below.sum()) * p))
            above = jj < j1  # bottom side of the element above: h(0, c), -1
            e_rows.append(grow[above].ravel())
            e_cols.append(((ii[above] + nh * (jj[above] - j0))[:, None] * ne + o + k[None, :]).ravel())
            e_vals.append(-np.ones(int(above.sum()) * p))
            for j in (j0, j1):
                if j in cut_index and ja <= j <= jb:
                    sel = jj == j
                    s_rows.append(grow[sel].ravel())
                    s_slot.append(((f1 * len(cuts) + cut_index[j]) * nh * p + ii[sel][:, None] * p + kk[None, :]).ravel())
                    s_other.append(np.full(int(sel.sum()) * p, owner(j - 1) if j == j0 else owner(j), np.int64))
        if j1 > j0 and nh > 1:
            jj, ii = np.meshgrid(np.arange(j0, j1), np.arange(1, nh), indexing="ij")
            jj, ii = jj.ravel(), ii.ravel()
            grow = base + n_h * p + (jj * (nh - 1) + ii - 1)[:, None] * p + kk[None, :]
            m_rows.append(grow.ravel()), m_ei.append(np.repeat(ii - 1, p)), m_ej.append(np.repeat(jj, p)), m_z.append(-np.ones(grow.size))
            e_rows.append(grow.ravel())  # left side of the right element: v(r, 0), +1
            e_cols.append(((ii + nh * (jj - j0))[:, None] * ne + o + nh_edges + np1 * k[None, :]).ravel())
            e_vals.append(np.ones(grow.size))
            e_rows.append(grow.ravel())  # right side of the left element: v(r, p), -1
            e_cols.append(((ii - 1 + nh * (jj - j0))[:, None] * ne + o + nh_edges + np1 * k[None, :] + p).ravel())
            e_vals.append(-np.ones(grow.size))
        f1 += 1
    n_trace_global = n1 * per_form
    cat = lambda parts, dt: np.concatenate(parts).astype(dt) if parts else np.zeros(0, dt)  # noqa: E731
    g_all = cat(m_rows, np.int64)
    order_m = np.argsort(g_all, kind="stable")
    g_rows = g_all[order_m]
    ei, ej, z = cat(m_ei, np.int64)[order_m], cat(m_ej, np.int64)[order_m], cat(m_z, np.float64)[order_m]
    rows_g, l_col, l_val = cat(e_rows, np.int64), cat(e_cols, np.int64), cat(e_vals, np.float64)
    l_row = np.searchsorted(g_rows, rows_g)
    order_e = np.lexsort((l_col, l_row))
    sg, ss, so = cat(s_rows, np.int64), cat(s_slot, np.int64), cat(s_other, np.int64)
    order_s = np.argsort(sg, kind="stable")
    sg, ss, so = sg[order_s], ss[order_s], so[order_s]
    sh = np.searchsorted(g_rows, sg).astype(np.int64)
    owned = np.ones(g_rows.size, np.uint8)
    owned[sh[so < rank]] = 0
    part = ConstraintPartition(l_row[order_e], l_col[order_e], l_val[order_e], g_rows.size, g_rows, sh, ss, n1 * len(cuts) * nh * p, owned,
                               np.arange(sh.size + 1, dtype=np.int64), so.astype(np.int32))
    # hierarchy: rank-local levels (shared multipliers left out, ids compacted in key order), then the exact level with global ids
    strip = np.searchsorted(strip_rows, ej, side="right") - 1
    jl = ej - strip_rows[strip]
    levels, counts = [], []
    for H in mid_sizes:
        key = (strip * (nv + 1) + jl // H) * (nh + 1) + ei // H
        a = np.full(g_rows.size, -1, np.int32)
        keep = np.ones(g_rows.size, bool)
        keep[sh] = False
        used, compact = np.unique(key[keep], return_inverse=True)
        a[keep] = compact
        levels.append(a)
        counts.append(int(used.size))
    if exact_size:
        nbi = -(-nh // exact_size)
        levels.append(((ej // exact_size) * nbi + ei // exact_size).astype(np.int32))
        counts.append(nbi * (-(-nv // exact_size)))
    agg = np.stack(levels) if levels else np.zeros((0, g_rows.size), np.int32)
    return part, (agg, z, counts), n_trace_global
