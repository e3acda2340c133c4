"""CPU restatement (numpy) of mfv2d's element-level hot path.  TEST INFRASTRUCTURE ONLY.

This file is the *checker*: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  Nothing under
``mfv2d_b200/`` imports it, and the product path has no CPU fallback.

Parity pin: every function here is checked in ``tests/test_oracle.py (test_oracle_vs_live_reference)`` against the
reference's own C extension compiled from /root/reference by ``oracle/build_ref.sh``
(``oracle/_ref/_mfv2d_ref.so``) and against the committed fixtures in ``tests/golden``
that were generated with the reference (``tests/golden/make_golden.py``).

All file:line citations are into the reference repository (j4nr0th/mfv2d @ b3eecf5).
The arithmetic is written as dense ``einsum`` contractions, i.e. deliberately *not* in the
summation order of either the reference's loops or the CUDA kernels.
"""

from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Sequence

import numpy as np

# ----------------------------------------------------------------------------------------
# enums (src/evaluation/bytecode.h:12-33, src/evaluation/forms.h:7-13, src/algebra/matrix.h:33-40)
# ----------------------------------------------------------------------------------------
MATOP_INVALID, MATOP_IDENTITY, MATOP_MASS, MATOP_INCIDENCE, MATOP_PUSH, MATOP_SCALE, MATOP_SUM, MATOP_INTERPROD = range(8)
FORM_ORDER_0, FORM_ORDER_1, FORM_ORDER_2 = 1, 2, 3
INC_10, INC_21, INC_10_T, INC_21_T = range(4)


class OracleError(RuntimeError):
    """Mirror of ``RuntimeError('Runtime error <MFV2D_...> was encountered.')`` (element_system.c:205-209)."""

    def __init__(self, code: str, msg: str = ""):
        super().__init__(f"Runtime error {code} was encountered. {msg}")
        self.code = code


# ----------------------------------------------------------------------------------------
# polynomials
# ----------------------------------------------------------------------------------------
def _legendre_two(n: int, x: float) -> tuple[float, float]:
    """P_{n-1}(x), P_n(x) by Bonnet recursion (src/polynomials/gauss_lobatto.c:18-33)."""
    v1, v2 = 1.0, x
    for i in range(2, n + 1):
        new = ((2 * i - 1) * x * v2 - (i - 1) * v1) / float(i)
        v1, v2 = v2, new
    return v1, v2


def gll(n: int, tol: float = 1e-15, max_iter: int = 10) -> tuple[np.ndarray, np.ndarray]:
    """``n`` Gauss-Lobatto-Legendre nodes and weights (src/polynomials/gauss_lobatto.c:52-85)."""
    x = np.empty(n)
    w = np.empty(n)
    x[0], x[n - 1] = -1.0, 1.0
    w[0] = w[n - 1] = 2.0 / float(n * (n - 1))
    kx_1 = 1.0 - 3.0 * (n - 2) / (8.0 * (n - 1) * (n - 1) * (n - 1))
    kx_2 = math.pi / (4.0 * (n - 1) + 1)
    for i in range(2, n):
        new_x = kx_1 * math.cos(kx_2 * (4 * i - 3))
        error = 1.0
        it = 0
        while it < max_iter and error > tol:
            p0, p1 = _legendre_two(n - 1, new_x)
            den = 1 - new_x * new_x
            dy = (n - 1) * (p0 - new_x * p1) / den
            d2y = (2 * new_x * dy - (n - 1) * n * p1) / den
            d3y = (4 * new_x * d2y - ((n - 1) * n - 2) * dy) / den
            dx = 2 * dy * d2y / (2 * d2y * d2y - dy * d3y)
            new_x -= dx
            error = abs(dx)
            it += 1
        x[n - i] = new_x
        _, p1 = _legendre_two(n - 1, new_x)
        w[n - i] = 2.0 / (n * (n - 1) * p1 * p1)
    return x, w


def compute_gll(order: int) -> tuple[np.ndarray, np.ndarray]:
    """Python-facing ``compute_gll(order)`` (gauss_lobatto.c:163-271); order 0 -> x=0, w=2 (:251-256)."""
    if order == 0:
        return np.array([0.0]), np.array([2.0])
    return gll(order + 1)


def lagrange1d(roots: np.ndarray, x: np.ndarray) -> np.ndarray:
    """Lagrange basis on ``roots`` at ``x`` -> shape x.shape + (n,) (src/polynomials/lagrange.c:63-107)."""
    roots = np.asarray(roots, float)
    xx = np.asarray(x, float)
    n = roots.size
    out = np.ones(xx.shape + (n,))
    for j in range(n):
        for i in range(n):
            if i != j:
                out[..., j] *= (xx - roots[i]) / (roots[j] - roots[i])
    return out


def dlagrange1d(roots: np.ndarray, x: np.ndarray) -> np.ndarray:
    """First derivative of the Lagrange basis (src/polynomials/lagrange.c:110-171)."""
    roots = np.asarray(roots, float)
    xx = np.asarray(x, float)
    n = roots.size
    out = np.zeros(xx.shape + (n,))
    for j in range(n):
        den = np.prod([roots[j] - roots[i] for i in range(n) if i != j])
        for m in range(n):
            if m == j:
                continue
            term = np.ones(xx.shape)
            for i in range(n):
                if i != j and i != m:
                    term = term * (xx - roots[i])
            out[..., j] += term
        out[..., j] /= den
    return out


def legendre(order: int, x: np.ndarray) -> np.ndarray:
    """P_0..P_order at x -> (order+1, *x.shape) (src/polynomials/legendre.c:12-36)."""
    xx = np.asarray(x, float)
    out = np.empty((order + 1,) + xx.shape)
    out[0] = 1.0
    if order > 0:
        out[1] = xx
    for i in range(2, order + 1):
        out[i] = ((2 * i - 1) * xx * out[i - 1] - (i - 1) * out[i - 2]) / float(i)
    return out


# ----------------------------------------------------------------------------------------
# basis tables (src/fem_space/basis.c:5-94, integration_rule.c:19-67)
# ----------------------------------------------------------------------------------------
@dataclass
class Basis:
    p: int  # element order
    q: int  # integration points per direction (rule order + 1)
    roots: np.ndarray  # (p+1,)
    pnts: np.ndarray  # (q,)
    wgts: np.ndarray  # (q,)
    node: np.ndarray  # (p+1, q)   node[i, m] = l_i(xi_m)
    edge: np.ndarray  # (p, q)     edge[j, m] = -sum_{i<=j} l_i'(xi_m)


_BASIS_CACHE: dict[tuple[int, int], Basis] = {}


def basis(p: int, q: int) -> Basis:
    key = (p, q)
    if key not in _BASIS_CACHE:
        roots, _ = gll(p + 1)
        pnts, wgts = gll(q) if q > 1 else (np.array([0.0]), np.array([2.0]))
        node = lagrange1d(roots, pnts).T.copy()
        d = dlagrange1d(roots, pnts)  # (q, p+1)
        edge = (-np.cumsum(d, axis=1)[:, :p]).T.copy()
        _BASIS_CACHE[key] = Basis(p, q, roots, pnts, wgts, node, edge)
    return _BASIS_CACHE[key]


def form_size(form: int, p: int) -> int:
    """src/evaluation/forms.c:21-34 (square orders)."""
    return {FORM_ORDER_0: (p + 1) ** 2, FORM_ORDER_1: 2 * p * (p + 1), FORM_ORDER_2: p * p}[form]


# ----------------------------------------------------------------------------------------
# geometry (src/fem_space/fem_space.c:16-57)
# ----------------------------------------------------------------------------------------
def jacobian(corners: np.ndarray, pnts: np.ndarray):
    """corners (..., 4, 2) -> j00, j01, j10, j11, det each (..., q_eta, q_xi)."""
    c = np.asarray(corners, float)
    x0, x1, x2, x3 = (c[..., k, 0][..., None, None] for k in range(4))
    y0, y1, y2, y3 = (c[..., k, 1][..., None, None] for k in range(4))
    eta = pnts[:, None]
    xi = pnts[None, :]
    j00 = ((x1 - x0) * (1 - eta) + (x2 - x3) * (1 + eta)) / 4 + 0 * xi
    j01 = ((y1 - y0) * (1 - eta) + (y2 - y3) * (1 + eta)) / 4 + 0 * xi
    j10 = ((x3 - x0) * (1 - xi) + (x2 - x1) * (1 + xi)) / 4 + 0 * eta
    j11 = ((y3 - y0) * (1 - xi) + (y2 - y1) * (1 + xi)) / 4 + 0 * eta
    det = j00 * j11 - j10 * j01
    return j00, j01, j10, j11, det


def physical_coordinates(corners: np.ndarray, pnts: np.ndarray):
    """Bilinear map of the quadrature grid (src/evaluation/integrating_fields.c:92-111)."""
    c = np.asarray(corners, float)
    eta = pnts[:, None]
    xi = pnts[None, :]
    n0 = (1 - xi) / 2 * (1 - eta) / 2
    n1 = (1 + xi) / 2 * (1 - eta) / 2
    n2 = (1 + xi) / 2 * (1 + eta) / 2
    n3 = (1 - xi) / 2 * (1 + eta) / 2
    x = sum(c[..., k, 0][..., None, None] * n for k, n in enumerate((n0, n1, n2, n3)))
    y = sum(c[..., k, 1][..., None, None] * n for k, n in enumerate((n0, n1, n2, n3)))
    return x, y


# ----------------------------------------------------------------------------------------
# 2-D basis function tables at the (eta=a, xi=b) grid
# ----------------------------------------------------------------------------------------
def _phi0(b: Basis):  # (n0, q, q): index i*(p+1)+j, value node[j](xi_b) * node[i](eta_a)  (fem_space.c:121-144)
    return np.einsum("ia,jb->ijab", b.node, b.node).reshape(-1, b.q, b.q)


def _phi1h(b: Basis):  # (p+1)*p: edge_xi(j) * node_eta(i), idx i*p+j   (fem_space.c:147-168)
    return np.einsum("ia,jb->ijab", b.node, b.edge).reshape(-1, b.q, b.q)


def _phi1v(b: Basis):  # p*(p+1): node_xi(j) * edge_eta(i), idx i*(p+1)+j  (fem_space.c:170-190)
    return np.einsum("ia,jb->ijab", b.edge, b.node).reshape(-1, b.q, b.q)


def _phi2(b: Basis):  # p*p: edge_xi(j) * edge_eta(i), idx i*p+j  (fem_space.c:201-217)
    return np.einsum("ia,jb->ijab", b.edge, b.edge).reshape(-1, b.q, b.q)


def _W(b: Basis):
    return b.wgts[:, None] * b.wgts[None, :]


# ----------------------------------------------------------------------------------------
# mass matrices (src/fem_space/fem_space.c:235-411); batched over leading axes of corners
# ----------------------------------------------------------------------------------------
def mass_node(b: Basis, corners):
    *_, det = jacobian(corners, b.pnts)
    f = _phi0(b)
    return np.einsum("wab,vab,...ab->...wv", f, f, det * _W(b), optimize=True)


def mass_edge(b: Basis, corners):
    j00, j01, j10, j11, det = jacobian(corners, b.pnts)
    W = _W(b)
    h, v = _phi1h(b), _phi1v(b)
    khh = (j11 * j11 + j10 * j10) / det * W  # fem_space.c:339
    kvv = (j00 * j00 + j01 * j01) / det * W  # fem_space.c:316
    khv = (j01 * j11 + j00 * j10) / det * W  # fem_space.c:362
    hh = np.einsum("wab,vab,...ab->...wv", h, h, khh, optimize=True)
    vv = np.einsum("wab,vab,...ab->...wv", v, v, kvv, optimize=True)
    hv = np.einsum("wab,vab,...ab->...wv", h, v, khv, optimize=True)
    top = np.concatenate([hh, hv], axis=-1)
    bot = np.concatenate([np.swapaxes(hv, -1, -2), vv], axis=-1)
    return np.concatenate([top, bot], axis=-2)


def mass_surf(b: Basis, corners):
    *_, det = jacobian(corners, b.pnts)
    f = _phi2(b)
    return np.einsum("wab,vab,...ab->...wv", f, f, _W(b) / det, optimize=True)


def mass(b: Basis, corners, form: int, inverse: bool = False):
    m = {FORM_ORDER_0: mass_node, FORM_ORDER_1: mass_edge, FORM_ORDER_2: mass_surf}[form](b, corners)
    if inverse:
        return invert(m)
    return m


# interior-product ("mixed") mass matrices  (src/fem_space/fem_space.c:546-635, 753-846)
def interprod_node_edge(b: Basis, corners, fld, transpose: bool):
    """fld (..., q, q, 2) -> (n0, n1) matrix [or its transpose]."""
    j00, j01, j10, j11, det = jacobian(corners, b.pnts)
    W = _W(b)
    fx, fy = fld[..., 0], fld[..., 1]
    kh = (fx * j11 - fy * j10) * W
    kv = -(fy * j00 - fx * j01) * W
    n, h, v = _phi0(b), _phi1h(b), _phi1v(b)
    mh = np.einsum("wab,vab,...ab->...wv", n, h, kh, optimize=True)
    mv = np.einsum("wab,vab,...ab->...wv", n, v, kv, optimize=True)
    m = np.concatenate([mh, mv], axis=-1)
    return np.swapaxes(m, -1, -2).copy() if transpose else m


def interprod_edge_surf(b: Basis, corners, fld, transpose: bool):
    """fld (..., q, q, 2) -> (n1, n2) matrix [or its transpose]."""
    j00, j01, j10, j11, det = jacobian(corners, b.pnts)
    W = _W(b)
    fx, fy = fld[..., 0], fld[..., 1]
    kh = -(fx * j10 + fy * j11) / det * W
    kv = -(fx * j00 + fy * j01) / det * W
    s, h, v = _phi2(b), _phi1h(b), _phi1v(b)
    mh = np.einsum("wab,vab,...ab->...wv", h, s, kh, optimize=True)
    mv = np.einsum("wab,vab,...ab->...wv", v, s, kv, optimize=True)
    m = np.concatenate([mh, mv], axis=-2)
    return np.swapaxes(m, -1, -2).copy() if transpose else m


# two-basis ("projector") masses (src/fem_space/fem_space.c:847-1055): rows = output basis, cols = input basis
def mass_double(b_out: Basis, b_in: Basis, corners, form: int):
    assert b_out.q == b_in.q
    j00, j01, j10, j11, det = jacobian(corners, b_out.pnts)
    W = _W(b_out)
    if form == FORM_ORDER_0:
        return np.einsum("wab,vab,...ab->...wv", _phi0(b_out), _phi0(b_in), det * W, optimize=True)
    if form == FORM_ORDER_2:
        return np.einsum("wab,vab,...ab->...wv", _phi2(b_out), _phi2(b_in), W / det, optimize=True)
    khh = (j11 * j11 + j10 * j10) / det * W
    kvv = (j00 * j00 + j01 * j01) / det * W
    khv = (j01 * j11 + j00 * j10) / det * W
    ho, vo, hi, vi = _phi1h(b_out), _phi1v(b_out), _phi1h(b_in), _phi1v(b_in)
    hh = np.einsum("wab,vab,...ab->...wv", ho, hi, khh, optimize=True)
    vv = np.einsum("wab,vab,...ab->...wv", vo, vi, kvv, optimize=True)
    hv = np.einsum("wab,vab,...ab->...wv", ho, vi, khv, optimize=True)
    vh = np.einsum("wab,vab,...ab->...wv", vo, hi, khv, optimize=True)
    return np.concatenate([np.concatenate([hh, hv], -1), np.concatenate([vh, vv], -1)], -2)


# ----------------------------------------------------------------------------------------
# dense algebra (src/algebra/matrix.c:361-534)
# ----------------------------------------------------------------------------------------
def static_pivots(a: np.ndarray, tol: float = 1e-12) -> np.ndarray:
    """Greedy row order from the *original* column magnitudes (matrix.c:424-466)."""
    n = a.shape[0]
    piv = np.empty(n, np.int64)
    used = np.zeros(n, bool)
    for i in range(n):
        col = np.abs(a[:, i]).copy()
        col[used] = -2.0
        j = int(np.argmax(col))  # first maximum, as the strict '>' scan does
        if used[j] or col[j] < tol:
            raise OracleError("MFV2D_PIVOT_FAILED")
        piv[i] = j
        used[j] = True
    return piv


def lu_static(a: np.ndarray, tol: float = 1e-12):
    """Unit-L Doolittle LU of the row-permuted matrix (matrix.c:468-498)."""
    piv = static_pivots(a, tol)
    m = a[piv].astype(float).copy()
    n = m.shape[0]
    lu = np.zeros_like(m)
    for i in range(n):
        for j in range(i):
            lu[i, j] = (m[i, j] - lu[i, :j] @ lu[:j, j]) / lu[j, j]
        for j in range(i + 1):
            lu[j, i] = m[j, i] - lu[j, :j] @ lu[:j, i]
    return lu, piv


def lu_solve(lu: np.ndarray, rhs: np.ndarray) -> np.ndarray:
    """matrix.c:500-534."""
    n = lu.shape[0]
    x = np.array(rhs, float, copy=True)
    for i in range(n):
        x[i] = x[i] - lu[i, :i] @ x[:i]
    for i in range(n - 1, -1, -1):
        x[i] = (x[i] - lu[i, i + 1 :] @ x[i + 1 :]) / lu[i, i]
    return x


def invert(a: np.ndarray) -> np.ndarray:
    """matrix_full_invert (matrix.c:361-411), batched over leading axes."""
    a = np.asarray(a, float)
    if a.ndim > 2:
        return np.stack([invert(x) for x in a])
    lu, piv = lu_static(a)
    n = a.shape[0]
    rhs = np.zeros((n, n))
    rhs[np.arange(n), piv] = 1.0
    return lu_solve(lu, rhs)


# ----------------------------------------------------------------------------------------
# incidence (src/evaluation/incidence.c:16-443)
# ----------------------------------------------------------------------------------------
def e10(p: int) -> np.ndarray:
    n0, n1 = (p + 1) ** 2, 2 * p * (p + 1)
    e = np.zeros((n1, n0))
    for r in range(p + 1):
        for c in range(p):
            e[r * p + c, (p + 1) * r + c] += 1.0
            e[r * p + c, (p + 1) * r + c + 1] -= 1.0
    for r in range(p):
        for c in range(p + 1):
            row = p * (p + 1) + r * (p + 1) + c
            e[row, (p + 1) * (r + 1) + c] += 1.0
            e[row, (p + 1) * r + c] -= 1.0
    return e


def e21(p: int) -> np.ndarray:
    n1, n2 = 2 * p * (p + 1), p * p
    e = np.zeros((n2, n1))
    for r in range(p):
        for c in range(p):
            row = r * p + c
            e[row, p * r + c] += 1.0
            e[row, p * (r + 1) + c] -= 1.0
            e[row, p * (p + 1) + (p + 1) * r + c] += 1.0
            e[row, p * (p + 1) + (p + 1) * r + c + 1] -= 1.0
    return e


def incidence(kind: int, p: int) -> np.ndarray:
    return {INC_10: e10(p), INC_21: e21(p), INC_10_T: e10(p).T, INC_21_T: e21(p).T}[kind]


# ----------------------------------------------------------------------------------------
# field reconstruction (src/evaluation/integrating_fields.c:238-352)
# ----------------------------------------------------------------------------------------
def reconstruct_field(b: Basis, corners, form: int, dofs: np.ndarray) -> np.ndarray:
    j00, j01, j10, j11, det = jacobian(corners, b.pnts)
    if form == FORM_ORDER_0:
        return np.einsum("...w,wab->...ab", dofs, _phi0(b)) * det
    if form == FORM_ORDER_2:
        return np.einsum("...w,wab->...ab", dofs, _phi2(b)) / det
    nh = b.p * (b.p + 1)
    v_eta = np.einsum("...w,wab->...ab", dofs[..., :nh], _phi1h(b))
    v_xi = np.einsum("...w,wab->...ab", dofs[..., nh:], _phi1v(b))
    vx = (j00 * v_xi + j10 * v_eta) / det
    vy = (j01 * v_xi + j11 * v_eta) / det
    return np.stack([vx, vy], axis=-1)


# ----------------------------------------------------------------------------------------
# dual degrees of freedom of a function (python/mfv2d/mimetic2d.py:1003-1123)
# ----------------------------------------------------------------------------------------
def dual_dofs(b: Basis, corners, form: int, fvals: np.ndarray) -> np.ndarray:
    """fvals: values of the function at the quadrature points, (q, q) or (q, q, 2) for 1-forms."""
    j00, j01, j10, j11, det = jacobian(corners, b.pnts)
    w = b.wgts[:, None] * b.wgts[None, :]
    if form == FORM_ORDER_0:
        return np.einsum("wab,ab->w", _phi0(b), fvals * w * det)
    if form == FORM_ORDER_2:
        return np.einsum("wab,ab->w", _phi2(b), fvals * w)
    f_xi = (j00 * fvals[..., 0] + j01 * fvals[..., 1]) * w
    f_eta = (j10 * fvals[..., 0] + j11 * fvals[..., 1]) * w
    return np.concatenate([np.einsum("wab,ab->w", _phi1h(b), f_eta), np.einsum("wab,ab->w", _phi1v(b), f_xi)])


# ----------------------------------------------------------------------------------------
# bytecode VM (src/evaluation/element_eval.c:18-479, src/algebra/matrix.c:188-276)
# ----------------------------------------------------------------------------------------
@dataclass
class _State:
    type: str = "invalid"  # invalid | identity | incidence | full
    coef: float = 0.0
    inc: int = -1
    full: np.ndarray | None = None


@dataclass
class ElementContext:
    b: Basis
    corners: np.ndarray  # (4, 2)
    fields: list = field(default_factory=list)  # per interprod op, (q, q, 2)
    _cache: dict = field(default_factory=dict)

    def mass(self, form: int, inv: bool) -> np.ndarray:
        key = (form, inv)
        if key not in self._cache:
            m = self.mass(form, False) if inv else mass(self.b, self.corners, form)
            self._cache[key] = invert(m) if inv else m
        return self._cache[key]


def _inc_to_full_quirk():
    # Q1: incidence_to_full returns 1 on success while callers test != MFV2D_SUCCESS (incidence.c:504,590)
    raise OracleError("MFV2D_FAILED_ALLOC", "dense incidence requested (reference quirk Q1)")


def evaluate_block(ctx: ElementContext, ops: Sequence[tuple], form_row: int, initial: np.ndarray | None):
    """Stack machine of element_eval.c:400-479.  ``ops`` are decoded tuples:
    (IDENTITY,), (MASS, form, inv), (INCIDENCE, form, transpose), (PUSH,), (SCALE, k), (SUM, n),
    (INTERPROD, form, field_index, transpose).  ``initial`` = (n,1) vector in vector mode."""
    p = ctx.b.p
    cur = _State()
    if initial is not None:
        cur = _State("full", 1.0, -1, np.array(initial, float).reshape(-1, 1))
    stack: list[_State] = []

    def left_dense(cur: _State, mat: np.ndarray, extra: float):
        if cur.type in ("invalid", "identity"):
            coef = 1.0 if cur.type == "invalid" else cur.coef
            return _State("full", coef * extra, -1, mat.copy())
        if cur.type == "incidence":  # mat @ E   (apply_incidence_to_full_right)
            return _State("full", cur.coef * extra, -1, mat @ incidence(cur.inc, p))
        # dense @ dense: the accumulated coefficient is dropped (Q2, element_eval.c:269,389)
        if mat.shape[1] != cur.full.shape[0]:
            raise OracleError("MFV2D_DIMS_MISMATCH")
        return _State("full", 1.0 * extra, -1, mat @ cur.full)

    for op in ops:
        code = op[0]
        if code == MATOP_IDENTITY:
            if cur.type == "invalid":
                cur = _State("identity", 1.0)
        elif code == MATOP_SCALE:
            if cur.type == "invalid":
                cur = _State("identity", float(op[1]))
            else:
                cur.coef *= float(op[1])
        elif code == MATOP_PUSH:
            stack.append(cur)
            cur = _State()
            if initial is not None:
                cur = _State("full", 1.0, -1, np.array(initial, float).reshape(-1, 1))
        elif code == MATOP_INCIDENCE:
            form, tr = op[1], bool(op[2])
            if form not in (FORM_ORDER_0, FORM_ORDER_1):
                raise OracleError("MFV2D_BAD_ENUM")
            t = (form - 1) + (2 if tr else 0)
            if cur.type == "invalid":
                cur = _State("incidence", 1.0, t)
            elif cur.type == "identity":
                cur = _State("incidence", cur.coef, t)
            elif cur.type == "incidence":
                _inc_to_full_quirk()
            else:
                e = incidence(t, p)
                if e.shape[1] != cur.full.shape[0]:
                    raise OracleError("MFV2D_DIMS_MISMATCH")
                cur = _State("full", cur.coef, -1, e @ cur.full)
        elif code == MATOP_MASS:
            form, inv = op[1], bool(op[2])
            cur = left_dense(cur, ctx.mass(form, inv), 1.0)
        elif code == MATOP_INTERPROD:
            form, fidx, tr = op[1], op[2], bool(op[3])
            if form == FORM_ORDER_1:
                m = interprod_node_edge(ctx.b, ctx.corners, ctx.fields[fidx], tr)
                k = 1.0
            elif form == FORM_ORDER_2:
                m = interprod_edge_surf(ctx.b, ctx.corners, ctx.fields[fidx], tr)
                k = -1.0
            else:
                raise OracleError("MFV2D_BAD_ENUM")
            cur = left_dense(cur, m, k)
        elif code == MATOP_SUM:
            n = int(op[1])
            if n > len(stack):
                raise OracleError("MFV2D_STACK_UNDERFLOW")
            for _ in range(n):
                left = stack.pop()
                cur = _add(cur, left, p)
        else:
            raise OracleError("MFV2D_BAD_ENUM")

    # matrix_as_full (element_eval.c:18-61)
    if cur.type == "identity":
        n = form_size(form_row, p)
        return np.eye(n) * cur.coef
    if cur.type == "full":
        return cur.full * cur.coef
    if cur.type == "incidence":
        _inc_to_full_quirk()
    raise OracleError("MFV2D_BAD_ENUM")


def _add(right: _State, left: _State, p: int) -> _State:
    """matrix_add (matrix.c:188-276), Q3."""
    if left.type == "invalid" or right.type == "invalid":
        raise OracleError("MFV2D_BAD_ENUM")
    if left.type == right.type:
        if left.type == "identity":
            return _State("identity", left.coef + right.coef)
        if left.type == "incidence" and left.inc == right.inc:
            return _State("incidence", left.coef + right.coef, left.inc)
        if left.type == "full":
            if left.full.shape != right.full.shape:
                raise OracleError("MFV2D_DIMS_MISMATCH")
            return _State("full", 1.0, -1, right.full * right.coef + left.full * left.coef)
    if left.type == "full" and right.type == "full":
        raise OracleError("MFV2D_WRONG_MAT_TYPES")
    _inc_to_full_quirk()


def decode_expressions(expressions, form_names: Sequence[str]):
    """Python tuples -> decoded ops + field specs (src/evaluation/bytecode.c:50-275,
    system_template.c:38-164).  Field spec = ('callable', fn) | ('unknown', form_index)."""
    nf = len(form_names)
    if len(expressions) != nf:
        raise ValueError("number of expression rows does not match the number of forms")
    fields: list[tuple] = []
    blocks: list[list | None] = []
    for row in expressions:
        if len(row) != nf:
            raise ValueError("number of expression columns does not match the number of forms")
        for expr in row:
            if expr is None:
                blocks.append(None)
                continue
            ops = []
            for item in expr:
                code = int(item[0])
                if code == MATOP_INTERPROD:
                    _, form, fld, tr = item
                    if callable(fld):
                        fields.append(("callable", fld))
                    else:
                        fields.append(("unknown", list(form_names).index(fld)))
                    ops.append((code, int(form), len(fields) - 1, bool(tr)))
                elif code == MATOP_SCALE:
                    ops.append((code, float(item[1])))
                elif code == MATOP_SUM:
                    ops.append((code, int(item[1])))
                elif code in (MATOP_MASS, MATOP_INCIDENCE):
                    ops.append((code, int(item[1]), bool(item[2])))
                else:
                    ops.append((code,))
            blocks.append(ops)
    return blocks, fields


def _element_fields(b: Basis, corners, field_specs, form_orders, dofs):
    """compute_fields (integrating_fields.c:13-236); bytecode fields are always 2-component
    (bytecode.c:131-199 forces FORM_ORDER_1)."""
    out = []
    offs = np.concatenate([[0], np.cumsum([form_size(f, b.p) for f in form_orders])])
    x, y = physical_coordinates(corners, b.pnts)
    for kind, what in field_specs:
        if kind == "callable":
            out.append(np.ascontiguousarray(np.asarray(what(x, y), float)).reshape(b.q, b.q, 2))
        else:
            seg = dofs[offs[what] : offs[what + 1]]
            out.append(reconstruct_field(b, corners, FORM_ORDER_1, seg))
    return out


def element_matrix(b: Basis, corners, form_orders: Sequence[int], blocks, field_specs=(), dofs=None,
                   field_values=None) -> np.ndarray:
    """compute_element_matrix (src/evaluation/element_system.c:13-212)."""
    sizes = [form_size(f, b.p) for f in form_orders]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    n = int(offs[-1])
    if dofs is None:
        dofs = np.zeros(n)
    fv = field_values if field_values is not None else _element_fields(b, corners, field_specs, form_orders, dofs)
    ctx = ElementContext(b, np.asarray(corners, float), fv)
    out = np.zeros((n, n))
    nf = len(form_orders)
    for r in range(nf):
        for c in range(nf):
            ops = blocks[r * nf + c]
            if ops is None:
                continue
            m = evaluate_block(ctx, ops, form_orders[r], None)
            if m.shape != (sizes[r], sizes[c]):
                raise OracleError("MFV2D_DIMS_MISMATCH")
            out[offs[r] : offs[r + 1], offs[c] : offs[c + 1]] = m
    return out


def element_vector(b: Basis, corners, form_orders: Sequence[int], blocks, dofs, field_specs=(),
                   field_values=None) -> np.ndarray:
    """compute_element_vector (src/evaluation/element_system.c:245-411)."""
    sizes = [form_size(f, b.p) for f in form_orders]
    offs = np.concatenate([[0], np.cumsum(sizes)])
    n = int(offs[-1])
    dofs = np.asarray(dofs, float)
    fv = field_values if field_values is not None else _element_fields(b, corners, field_specs, form_orders, dofs)
    ctx = ElementContext(b, np.asarray(corners, float), fv)
    out = np.zeros(n)
    nf = len(form_orders)
    for r in range(nf):
        for c in range(nf):
            ops = blocks[r * nf + c]
            if ops is None:
                continue
            m = evaluate_block(ctx, ops, form_orders[r], dofs[offs[c] : offs[c + 1]])
            if m.shape != (sizes[r], 1):
                raise OracleError("MFV2D_DIMS_MISMATCH")
            out[offs[r] : offs[r + 1]] += m[:, 0]
    return out


# ----------------------------------------------------------------------------------------
# element-block linear system (src/algebra/sparse_system.c:3-207, system_objects.c) and the
# Krylov drivers (python/mfv2d/solving.py:350-500).  The trace vector is kept in *merged* form
# (one value per multiplier) - every element copy holds the same value (sparse_system.c:109-161)
# and TraceVector.dot divides by the sharing count (system_objects.c:1690-1696).
# ----------------------------------------------------------------------------------------
@dataclass
class System:
    blocks: list  # dense (n_e, n_e)
    offsets: np.ndarray  # dense offsets per element (n_elem+1)
    con_rows: np.ndarray  # COO of the global constraint matrix N (n_trace x n_dense)
    con_cols: np.ndarray
    con_vals: np.ndarray
    n_trace: int
    _lu: list | None = None

    def lus(self):
        if self._lu is None:
            self._lu = [lu_static(a) for a in self.blocks]
        return self._lu

    def apply_diagonal(self, x):
        return np.concatenate([a @ x[o0:o1] for a, o0, o1 in zip(self.blocks, self.offsets[:-1], self.offsets[1:])])

    def apply_diagonal_inverse(self, x):
        out = []
        for (lu, piv), o0, o1 in zip(self.lus(), self.offsets[:-1], self.offsets[1:]):
            out.append(lu_solve(lu, x[o0:o1][piv]))
        return np.concatenate(out)

    def apply_trace(self, x):
        t = np.zeros(self.n_trace)
        np.add.at(t, self.con_rows, self.con_vals * x[self.con_cols])
        return t

    def apply_trace_transpose(self, t):
        y = np.zeros(int(self.offsets[-1]))
        np.add.at(y, self.con_cols, self.con_vals * t[self.con_rows])
        return y

    def schur(self, lam):
        return self.apply_trace(self.apply_diagonal_inverse(self.apply_trace_transpose(lam)))


def cg(apply: Callable[[np.ndarray], np.ndarray], rhs, x0, max_iter, atol, rtol):
    """cg_general (python/mfv2d/solving.py:350-436)."""
    res_mag2 = float(rhs @ rhs)
    tol = atol if math.sqrt(res_mag2) * rtol > atol else math.sqrt(res_mag2) * rtol
    p = rhs.copy()
    res = rhs.copy()
    x = x0.copy()
    it = 0
    for it in range(max_iter):
        ap = apply(p)
        alpha = res_mag2 / float(ap @ p)
        x += alpha * p
        res -= alpha * ap
        new = float(res @ res)
        if new < tol * tol:
            res_mag2 = new
            break
        beta = new / res_mag2
        res_mag2 = new
        p = res + beta * p
    return x, math.sqrt(res_mag2), it


def solve_schur(system: System, rhs, phi, max_iter, atol, rtol):
    """solve_schur_iterative (python/mfv2d/solving.py:439-500)."""
    inv_a_y = system.apply_diagonal_inverse(rhs)
    trace_rhs = system.apply_trace(inv_a_y) - phi
    lam, r, it = cg(system.schur, trace_rhs, np.zeros(system.n_trace), max_iter, atol, rtol)
    x = inv_a_y - system.apply_diagonal_inverse(system.apply_trace_transpose(lam))
    return x, lam, r, it
