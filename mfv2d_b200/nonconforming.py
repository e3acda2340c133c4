"""Continuity constraints of NON-conforming (h/p-refined) quadtree meshes from arrays (BASELINE configs[3]).

The reference builds these rows by walking its ``Mesh`` / ``Manifold2D`` objects (``connect_elements``,
python/mfv2d/continuity.py:610-759, with the hanging-side interpolation of ``_get_side_dofs``, :192-318).  Here the mesh is a
:class:`QuadForest` - plain arrays: an ``nh x nv`` grid of root quads (the layout of ``mfv2d.examples.unit_square_mesh``), the four
children of every split element (bottom-left, bottom-right, top-right, top-left; -1 for leaves), the element ids of the leaves in
leaf order and one polynomial order per leaf - and :func:`continuity_constraints_forest` returns the constraint matrix in COO
form over global dense columns, row for row in the reference's order:

    for every unknown form in turn (0-forms and 1-forms; 2-forms have no boundary DoFs)
        rows inside split root elements (children against each other, depth first)          continuity.py:524-607
        rows across the interior lines of the root grid (hanging sides included)             :394-521
        0-forms only: one chain of corner rows per interior mesh node                        :321-368, 689-716

What one side contributes is described by :class:`_Side` records (which leaf, which DoFs, where along the side); a side whose two
neighbours disagree in order or refinement is expressed in the basis of the HIGHER order through the inverse of the 1-D basis
matrix evaluated at the children's nodes - the same linear algebra as the reference (``numpy.linalg.inv`` of the same matrices, nodes
from the library's ``compute_gll``), so indices agree exactly and coefficients to rounding (tests/test_nonconforming.py against
tests/golden/forest.npz, written by the reference).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from . import _mfv2d as M
from .structured import FORM_ORDER_0, FORM_ORDER_1, FORM_ORDER_2, _side_dofs, form_sizes

BOTTOM, RIGHT, TOP, LEFT = 1, 2, 3, 4


def _next(side: int) -> int:
    return side % 4 + 1


@dataclass
class _Side:
    """One leaf's share of an element side: its DoFs (in side order) and the positions of its nodes on the side, in [-1, 1]."""

    leaf: int
    dofs: np.ndarray
    pos: np.ndarray


class QuadForest:
    """``nh x nv`` root quads (index ``i + nh * j``), every element either a leaf or split into four children."""

    def __init__(self, nh: int, nv: int, children, leaf_elements, leaf_orders):
        self.nh, self.nv = int(nh), int(nv)
        self.children = np.asarray(children, np.int64).reshape(-1, 4)
        self.leaf_elements = np.asarray(leaf_elements, np.int64)
        self.leaf_orders = np.asarray(leaf_orders, np.int64)
        if self.leaf_elements.size != self.leaf_orders.size:
            raise ValueError("one order per leaf is needed")
        self.leaf_index = np.full(self.children.shape[0], -1, np.int64)
        self.leaf_index[self.leaf_elements] = np.arange(self.leaf_elements.size)
        is_leaf = self.children[:, 0] < 0
        if np.any(is_leaf != (self.leaf_index >= 0)):
            raise ValueError("leaf_elements must list exactly the elements without children")
        self._gll: dict[int, np.ndarray] = {}
        self._side_matrix: dict = {}

    # ---- small queries ----------------------------------------------------------------------------------------------
    def is_split(self, e: int) -> bool:
        return self.children[e, 0] >= 0

    def order(self, e: int) -> int:
        return int(self.leaf_orders[self.leaf_index[e]])

    def pair_on_side(self, e: int, side: int) -> tuple[int, int]:
        """The two children that touch ``side``, in the counter-clockwise direction of the side (mimetic2d.py:703-710)."""
        c = self.children[e]
        return int(c[side - 1]), int(c[side & 3])

    def side_order(self, e: int, side: int) -> int:
        """Order seen from outside: a split side carries the sum of its children's orders (mimetic2d.py:824-847)."""
        if not self.is_split(e):
            return self.order(e)
        a, b = self.pair_on_side(e, side)
        return self.side_order(a, side) + self.side_order(b, side)

    def _side_shape(self, e: int, side: int):
        """nested tuple of the leaf orders along ``side``: the key of the side's interpolation matrix"""
        if not self.is_split(e):
            return self.order(e)
        a, b = self.pair_on_side(e, side)
        return (self._side_shape(a, side), self._side_shape(b, side))

    def nodes(self, order: int) -> np.ndarray:
        if order not in self._gll:
            self._gll[order] = np.array(M.compute_gll(order)[0], np.float64)
        return self._gll[order]

    def corner(self, e: int, side: int) -> tuple[int, int]:
        """(leaf, 0-form DoF) of the corner at which ``side`` begins (continuity.py:72-110)."""
        while self.is_split(e):
            e = int(self.children[e, side - 1])
        p = self.order(e)
        return int(self.leaf_index[e]), {BOTTOM: 0, RIGHT: p, TOP: (p + 1) * p + p, LEFT: p * (p + 1)}[side]

    def leaf_corners(self, root_corners) -> np.ndarray:
        """Corners ``(n_leaf, 4, 2)`` of the leaves from those of the roots: a split element is cut at the midpoints of its
        sides and at the mean of its corners, children bottom-left, bottom-right, top-right, top-left (the subdivision of the
        reference's mesh, src/geometry/mesh.c:86-104; same arithmetic, so the coordinates agree bit for bit)."""
        root_corners = np.asarray(root_corners, np.float64).reshape(self.nh * self.nv, 4, 2)
        out = np.empty((self.leaf_elements.size, 4, 2))

        def walk(e: int, c: np.ndarray) -> None:
            if not self.is_split(e):
                out[self.leaf_index[e]] = c
                return
            bottom, top, left, right = (c[0] + c[1]) / 2, (c[2] + c[3]) / 2, (c[0] + c[3]) / 2, (c[1] + c[2]) / 2
            mid = (c[0] + c[1] + c[2] + c[3]) / 4
            kids = self.children[e]
            walk(int(kids[0]), np.array([c[0], bottom, mid, left]))
            walk(int(kids[1]), np.array([bottom, c[1], right, mid]))
            walk(int(kids[2]), np.array([mid, right, c[2], top]))
            walk(int(kids[3]), np.array([left, mid, top, c[3]]))

        for root in range(self.nh * self.nv):
            walk(root, root_corners[root])
        return out

    # ---- what a side is made of ---------------------------------------------------------------------------------------
    def side_pieces(self, e: int, side: int, form: int) -> list[_Side]:
        """The leaves along ``side`` with their DoFs and node positions (continuity.py:113-188).  Where two children meet, the
        second one gives up its first node (0-forms: the DoF too - it is the same point)."""
        if not self.is_split(e):
            p = self.order(e)
            return [_Side(int(self.leaf_index[e]), np.asarray(_side_dofs(side, form, p), np.int64), self.nodes(p))]
        a, b = self.pair_on_side(e, side)
        first, second = self.side_pieces(a, side, form), self.side_pieces(b, side, form)
        head = second[0]
        second[0] = _Side(head.leaf, head.dofs[1:] if form == FORM_ORDER_0 else head.dofs, head.pos[1:])
        return [_Side(s.leaf, s.dofs, (s.pos - 1) / 2) for s in first] + [_Side(s.leaf, s.dofs, (s.pos + 1) / 2) for s in second]

    @staticmethod
    def _edge_basis(nodal: np.ndarray) -> np.ndarray:
        """values of the edge functions between consecutive points: running sums of the differences of the nodal values"""
        diffs = nodal[:-1, :] - nodal[1:, :]
        return np.cumsum(diffs[:, :-1], axis=1)

    def side_rows(self, e: int, side: int, form: int, out_order: int | None = None) -> list[list[tuple[int, np.ndarray, np.ndarray]]]:
        """The side's DoFs in the basis of order ``out_order`` as linear combinations of leaf DoFs (continuity.py:192-318): one
        list of ``(leaf, dofs, coefficients)`` per output DoF."""
        own = self.side_order(e, side)
        out_order = own if out_order is None else int(out_order)
        if not self.is_split(e) and out_order == own:
            leaf = int(self.leaf_index[e])
            return [[(leaf, np.array([d], np.int64), np.ones(1))] for d in _side_dofs(side, form, self.order(e))]
        pieces = self.side_pieces(e, side, form)
        if form not in (FORM_ORDER_0, FORM_ORDER_1):
            raise ValueError("2-forms have no boundary DoFs.")
        # the interpolation matrix depends on the SHAPE of the side only (how it is split and the orders of its leaves), not
        # on which elements they are: computed once per shape (same arithmetic, so the coefficients stay bit-identical)
        key = (form, out_order, self._side_shape(e, side))
        m = self._side_matrix.get(key)
        if m is None:
            own_nodes = self.nodes(own)
            nodal = M.lagrange1d(own_nodes, np.concatenate([s.pos for s in pieces]))
            m = np.linalg.inv(nodal) if form == FORM_ORDER_0 else np.linalg.inv(self._edge_basis(nodal))
            if out_order != own:
                to_out = M.lagrange1d(own_nodes, self.nodes(out_order))
                m = (to_out if form == FORM_ORDER_0 else self._edge_basis(to_out)) @ m
            self._side_matrix[key] = m
        rows = []
        for coeffs in m:
            at, row = 0, []
            for s in pieces:
                row.append((s.leaf, s.dofs, coeffs[at:at + s.dofs.size]))
                at += s.dofs.size
            rows.append(row)
        return rows

    # ---- rows ---------------------------------------------------------------------------------------------------------------
    def corner_chain(self, *corners: tuple[int, int]) -> list:
        """consecutive corners of a node tied together, +1 / -1 (continuity.py:321-368)"""
        where = [self.corner(e, s) for e, s in corners]
        return [[(l1, np.array([d1], np.int64), np.ones(1)), (l2, np.array([d2], np.int64), -np.ones(1))]
                for (l1, d1), (l2, d2) in zip(where[:-1], where[1:])]

    def side_midpoints(self, a: int, b: int, side: int) -> list:
        """0-forms: the node between two children on ``side`` and, recursively, the nodes inside them (continuity.py:371-391)"""
        rows = self.corner_chain((a, _next(side)), (b, side))
        for e in (a, b):
            if self.is_split(e):
                rows += self.side_midpoints(*self.pair_on_side(e, side), side)
        return rows

    def across(self, e1: int, s1: int, e2: int, s2: int, form: int) -> list:
        """Rows that join side ``s1`` of ``e1`` to side ``s2`` of ``e2`` (continuity.py:394-521)."""
        if self.is_split(e1) and self.is_split(e2):
            a1, b1 = self.pair_on_side(e1, s1)
            a2, b2 = self.pair_on_side(e2, s2)
            rows = self.across(a1, s1, b2, s2, form) + self.across(b1, s1, a2, s2, form)
            if form == FORM_ORDER_0:
                rows += self.corner_chain((a1, _next(s1)), (b1, s1), (b2, s2), (a2, _next(s2)))
            return rows
        rows = []
        if form == FORM_ORDER_0:
            if self.is_split(e1):
                rows += self.side_midpoints(*self.pair_on_side(e1, s1), s1)
            elif self.is_split(e2):
                rows += self.side_midpoints(*self.pair_on_side(e2, s2), s2)
        high = max(self.side_order(e1, s1), self.side_order(e2, s2))
        r1, r2 = self.side_rows(e1, s1, form, high), self.side_rows(e2, s2, form, high)
        if form == FORM_ORDER_0:
            r1, r2, sign = r1[1:-1], r2[1:-1], -1.0  # the end nodes belong to the corner chains
        else:
            sign = float((1 - (s1 & 2)) * (1 - (s2 & 2)))  # +1: both sides run the same way around their elements
        for d1, d2 in zip(r1, reversed(r2)):  # the second side is traversed the other way round
            rows.append(list(d1) + [(leaf, dofs, sign * c) for leaf, dofs, c in d2])
        return rows

    def inside(self, e: int, form: int) -> list:
        """Rows between the children of a split element, depth first (continuity.py:524-607)."""
        if not self.is_split(e):
            return []
        bl, br, tr, tl = (int(c) for c in self.children[e])
        rows = []
        for c in (bl, br, tr, tl):
            rows += self.inside(c, form)
        rows += (self.across(bl, RIGHT, br, LEFT, form) + self.across(br, TOP, tr, BOTTOM, form) + self.across(tr, LEFT, tl, RIGHT, form) +
                 self.across(tl, BOTTOM, bl, TOP, form))
        if form == FORM_ORDER_0:
            rows += self.corner_chain((bl, TOP), (br, LEFT), (tr, BOTTOM), (tl, RIGHT))
        return rows

    def form_rows(self, form: int) -> list:
        """All rows of one 0- or 1-form in the reference's order (continuity.py:629-718 on the root grid of unit_square_mesh:
        horizontal interior lines ``nh * j + i`` - element below, its TOP, against element above, its BOTTOM -, then vertical
        lines - right element, its LEFT, against left element, its RIGHT; mesh nodes in index order with the elements around
        them taken top-left, bottom-right, bottom-left, top-right)."""
        nh, nv = self.nh, self.nv
        rows = []
        for root in range(nh * nv):
            rows += self.inside(root, form)
        for j in range(1, nv):
            for i in range(nh):
                rows += self.across(i + nh * (j - 1), TOP, i + nh * j, BOTTOM, form)
        for j in range(nv):
            for i in range(1, nh):
                rows += self.across(i + nh * j, LEFT, i - 1 + nh * j, RIGHT, form)
        if form == FORM_ORDER_0:
            for j in range(nv + 1):
                for i in range(nh + 1):
                    around = [(i - 1, j, RIGHT), (i, j - 1, LEFT), (i - 1, j - 1, TOP), (i, j, BOTTOM)]
                    here = [(ci + nh * cj, s) for ci, cj, s in around if 0 <= ci < nh and 0 <= cj < nv]
                    if len(here) > 1:
                        rows += self.corner_chain(*here)
        return rows


def continuity_constraints_forest(forest: QuadForest, form_orders):
    """``connect_elements`` (python/mfv2d/continuity.py:610-759) for a :class:`QuadForest`: COO ``(row, col, val)`` over global
    dense columns (leaf ``l`` owns ``[offsets[l], offsets[l + 1])``, forms concatenated in ``form_orders`` order), the row
    count and the dense offsets of the leaves.  Entries keep the reference's order inside a row (they are NOT sorted by
    column: pass them through ``numpy.lexsort`` or ``scipy.sparse`` before handing them to ``LinearSystem``)."""
    orders = [int(o) for o in form_orders]
    sizes = np.array([int(form_sizes(orders, int(p)).sum()) for p in forest.leaf_orders], np.int64)
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    cache: dict[int, list] = {}
    rows, cols, vals = [], [], []
    n_rows = 0
    for f, order in enumerate(orders):
        if order == FORM_ORDER_2:
            continue
        if order not in cache:
            cache[order] = forest.form_rows(order)
        for row in cache[order]:
            for leaf, dofs, coeffs in row:
                p = int(forest.leaf_orders[leaf])
                form_off = int(form_sizes(orders[:f], p).sum()) if f else 0
                cols.append(offsets[leaf] + form_off + dofs)
                vals.append(np.asarray(coeffs, np.float64))
                rows.append(np.full(dofs.size, n_rows, np.int64))
            n_rows += 1
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0), 0, offsets
    return np.concatenate(rows), np.concatenate(cols).astype(np.int64), np.concatenate(vals), n_rows, offsets


def boundary_terms_forest(forest: QuadForest, form_orders, root_corners, over_integration: int, strong=(), weak=()):
    """``mesh_boundary_conditions`` (python/mfv2d/boundary.py:382-498) for a :class:`QuadForest`.

    The boundary lines are those of the root grid, in the order of ``structured.boundary_lines`` (= ``mesh.boundary_indices``);
    along a refined side the leaves are visited first child first (:152-172, :264-289).  Per line the equations are taken in
    form order: ``strong`` = ``[(form index, func, line positions or None for all lines), ...]`` wins over ``weak`` =
    ``[(form index, [(k, func), ...]), ...]``.  Strong 0-form rows skip a root-grid node that an earlier line has set (:455-470) -
    the node between two children of one side is set from both (the reference does the same); 1-form rows are the fluxes
    through the sub-edges.  ``over_integration`` is the ``FemCache`` order difference (rule of order ``p + over_integration``).

    Returns ``(leaf, dof, value)`` arrays of the strong rows in the reference's order (``dof`` local to the leaf, form offset
    included) and the weak contributions as one flat vector over the global dense columns."""
    from .structured import boundary_lines

    orders = [int(o) for o in form_orders]
    corners = forest.leaf_corners(root_corners)
    sizes = np.array([int(form_sizes(orders, int(p)).sum()) for p in forest.leaf_orders], np.int64)
    offsets = np.concatenate([[0], np.cumsum(sizes)])
    weak_vec = np.zeros(int(offsets[-1]))
    s_leaf, s_dof, s_val = [], [], []
    tables: dict[int, tuple] = {}

    def basis(p: int):
        if p not in tables:
            rule = M.IntegrationRule1D(p + int(over_integration))
            b = M.Basis1D(p, rule)
            tables[p] = (np.asarray(b.roots, np.float64), np.asarray(rule.nodes, np.float64), np.asarray(rule.weights, np.float64),
                         np.asarray(b.node, np.float64), np.asarray(b.edge, np.float64))
        return tables[p]

    def leaves_on(e: int, side: int, first: bool, last: bool):
        """(leaf element, skip_first, skip_last) along the side, first child first"""
        if not forest.is_split(e):
            return [(e, first, last)]
        a, b = forest.pair_on_side(e, side)
        return leaves_on(a, side, first, False) + leaves_on(b, side, False, last)

    nh, nv = forest.nh, forest.nv
    set_nodes = np.zeros((nv + 1, nh + 1), bool)
    strong_of = {int(fi): (func, None if lines is None else set(int(i) for i in lines)) for fi, func, lines in strong}
    weak_of = {int(fi): terms for fi, terms in weak}
    for i_line, (root, side) in enumerate(boundary_lines(nh, nv)):
        ei, ej = root % nh, root // nh
        node_of = [(ej, ei), (ej, ei + 1), (ej + 1, ei + 1), (ej + 1, ei)]
        n0, n1 = node_of[side - 1], node_of[side & 3]
        ndir = 2 * ((side & 2) >> 1) - 1
        for f, order in enumerate(orders):
            is_strong = f in strong_of and (strong_of[f][1] is None or i_line in strong_of[f][1])
            if not is_strong and not weak_of.get(f):
                continue
            if order == FORM_ORDER_2:
                raise ValueError(f"Unknown/Invalid weak form order form_order={order}." if not is_strong else "2-forms have no boundary DoFs.")
            for e, skip_first, skip_last in leaves_on(root, side, bool(is_strong and set_nodes[n0]), bool(is_strong and set_nodes[n1])):
                leaf = int(forest.leaf_index[e])
                p = int(forest.leaf_orders[leaf])
                roots, rn, rw, node, edge = basis(p)
                c = corners[leaf]
                p0, p1 = c[side - 1], c[side & 3]
                dx, dy = (p1[0] - p0[0]) / 2, (p1[1] - p0[1]) / 2
                form_off = int(form_sizes(orders[:f], p).sum()) if f else 0
                dofs = np.asarray(_side_dofs(side, order, p), np.int64) + form_off
                if is_strong:
                    func = strong_of[f][0]
                    xv = (p1[0] + p0[0]) / 2 + dx * roots
                    yv = (p1[1] + p0[1]) / 2 + dy * roots
                    if order == FORM_ORDER_0:
                        v = np.asarray(func(xv, yv), np.float64) * np.ones(p + 1)
                        lo, hi = (1 if skip_first else 0), (p if skip_last else p + 1)
                        dofs, v = dofs[lo:hi], v[lo:hi]
                    else:
                        v = np.zeros(p)
                        for i in range(p):
                            xc = (xv[i + 1] + xv[i]) / 2 + (xv[i + 1] - xv[i]) / 2 * rn
                            yc = (yv[i + 1] + yv[i]) / 2 + (yv[i + 1] - yv[i]) / 2 * rn
                            ddx, ddy = (xv[i + 1] - xv[i]) / 2, (yv[i + 1] - yv[i]) / 2
                            normal = ndir * np.array((ddy, -ddx))
                            fv = np.asarray(func(xc, yc), np.float64)
                            v[i] = np.sum((fv[..., 0] * normal[0] + fv[..., 1] * normal[1]) * rw)
                    s_leaf.append(np.full(dofs.size, leaf, np.int64)), s_dof.append(dofs), s_val.append(v)
                else:
                    xv = (p1[0] + p0[0]) / 2 + dx * rn
                    yv = (p1[1] + p0[1]) / 2 + dy * rn
                    vals = np.zeros(dofs.size)
                    for k, func in weak_of[f]:
                        fv = np.array(func(xv, yv), np.float64)
                        if order == FORM_ORDER_0:
                            tab, fv = node, -(fv[..., 0] * dx + fv[..., 1] * dy) * rw
                        else:
                            tab, fv = edge, fv * (-rw * ndir)
                        vals += np.sum(fv[None, ...] * tab, axis=1) * k
                    weak_vec[offsets[leaf] + dofs] += vals
            if is_strong:
                set_nodes[n0] = set_nodes[n1] = True
    cat = (lambda parts, dt: np.concatenate(parts).astype(dt) if parts else np.zeros(0, dt))
    return (cat(s_leaf, np.int64), cat(s_dof, np.int64), cat(s_val, np.float64)), weak_vec


def system_constraints_forest(forest: QuadForest, form_orders, root_corners, over_integration: int, constrained_forms=(), strong=(), weak=()):
    """The constraint matrix and right-hand side of ``add_system_constraints`` (python/mfv2d/continuity.py:762-872) for a
    :class:`QuadForest`: continuity rows (:func:`continuity_constraints_forest`), one row per ``constrained_forms`` entry
    ``(value, form index)`` (every DoF of the form in every leaf, coefficient 1), one row per strongly set DoF
    (:func:`boundary_terms_forest`).  Returns ``(row, col, val)`` sorted by row then column with duplicates summed (what
    ``scipy.sparse.csr_array`` does to the reference's triplets), the right-hand side, the row count, the dense offsets of the
    leaves and the weak boundary additions to the dense right-hand side."""
    orders = [int(o) for o in form_orders]
    row, col, val, n, offsets = continuity_constraints_forest(forest, orders)
    rows, cols, vals, rhs = [row], [col], [val], [np.zeros(n)]
    for value, fi in constrained_forms:
        c = []
        for leaf, p in enumerate(forest.leaf_orders):
            fs = form_sizes(orders, int(p))
            start = int(offsets[leaf] + fs[: int(fi)].sum())
            c.append(np.arange(start, start + int(fs[int(fi)])))
        c = np.concatenate(c)
        rows.append(np.full(c.size, n, np.int64)), cols.append(c), vals.append(np.ones(c.size)), rhs.append(np.array([float(value)]))
        n += 1
    (s_leaf, s_dof, s_val), weak_vec = boundary_terms_forest(forest, orders, root_corners, over_integration, strong, weak)
    rows.append(n + np.arange(s_val.size)), cols.append(offsets[s_leaf] + s_dof), vals.append(np.ones(s_val.size)), rhs.append(s_val)
    n += int(s_val.size)
    row, col, val = np.concatenate(rows).astype(np.int64), np.concatenate(cols).astype(np.int64), np.concatenate(vals)
    # sum duplicates (a hanging-side row can name a DoF twice), sort by (row, column)
    key = row * int(offsets[-1]) + col
    uniq, inverse = np.unique(key, return_inverse=True)
    summed = np.zeros(uniq.size)
    np.add.at(summed, inverse, val)
    return uniq // int(offsets[-1]), uniq % int(offsets[-1]), summed, np.concatenate(rhs), n, offsets, weak_vec
