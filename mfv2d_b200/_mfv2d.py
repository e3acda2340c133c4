"""Host-side mirror of the reference's ``mfv2d._mfv2d`` extension for the element hot path.

Same names, argument meaning and error behaviour as the reference's C functions and types (cited per
item), but every computation is a call into libmfv2d_cuda.so - batched CUDA kernels for sm_100a.  The
per-element entry points (``compute_element_matrix`` ...) are batches of one; ``ElementBatch`` and the
``*_batch`` functions are the fast path that the reference's per-element Python loops
(python/mfv2d/solve_system.py:988-1008, 406-421) collapse into.  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Callable, Sequence

import numpy as np

from . import _capi
from ._capi import DEVICE, HOST, OP_DTYPE, check, lib
from .runtime import Context, DeviceArray, addr, get_context, space_of

# ---- constants (src/module.c:193-198) ---------------------------------------------------------
MATOP_INVALID, MATOP_IDENTITY, MATOP_MASS, MATOP_INCIDENCE, MATOP_PUSH, MATOP_SCALE, MATOP_SUM, MATOP_INTERPROD = range(8)
ELEMENT_SIDE_BOTTOM, ELEMENT_SIDE_RIGHT, ELEMENT_SIDE_TOP, ELEMENT_SIDE_LEFT = 1, 2, 3, 4
FORM_ORDER_0, FORM_ORDER_1, FORM_ORDER_2 = 1, 2, 3


def _check_array(a, ndim: int | None, name: str, dtype=np.float64) -> np.ndarray:
    """check_input_array (src/common/common.c:70-117): type, contiguity, dtype and rank."""
    if not isinstance(a, np.ndarray):
        raise TypeError(f"Object {name} is not a numpy array, but is {type(a)!r}.")
    if not (a.flags.c_contiguous and a.flags.aligned):
        raise ValueError(f"Array {name} flags don't contain required flags (C-contiguous, aligned).")
    if ndim is not None and a.ndim != ndim:
        raise ValueError(
            f"Number of dimensions for the array {name} does not match expected value (expected {ndim}, got {a.ndim})."
        )
    if a.dtype != dtype:
        raise ValueError(f"Array {name} does not have the expected type (expected {np.dtype(dtype)}, got {a.dtype}).")
    return a


# =============================================================================================
# polynomials (src/polynomials/*.c)
# =============================================================================================
class GLLCache:
    """Cache of Gauss-Lobatto rules (src/polynomials/gauss_lobatto.c:273-348, 386-452): one entry per order with the tolerance
    it was computed to; an entry serves every request whose tolerance is not larger than the recorded one (:273-294)."""

    def __init__(self):
        self._entries: dict[int, tuple[float, np.ndarray, np.ndarray]] = {}

    def clear(self) -> None:
        self._entries.clear()

    def usage(self) -> tuple[tuple[int, float], ...]:
        return tuple((o, t) for o, (t, _, _) in sorted(self._entries.items()))


DEFAULT_GLL_CACHE = GLLCache()


def compute_gll(order: int, max_iter: int = 10, tol: float = 1e-15, cache: GLLCache | None = DEFAULT_GLL_CACHE) -> tuple[np.ndarray, np.ndarray]:
    """Gauss-Lobatto-Legendre nodes and weights of ``order + 1`` points
    (src/polynomials/gauss_lobatto.c:163-271; order 0 gives x = 0, w = 2, :251-256)."""
    order = int(order)
    if cache is not None and not isinstance(cache, GLLCache):
        raise TypeError("cache must be a GLLCache or None")
    if cache is not None and order in cache._entries and cache._entries[order][0] >= tol and max_iter >= 0 and tol >= 0:
        _, cx, cw = cache._entries[order]
        return cx.copy(), cw.copy()
    if order < 0:
        raise ValueError(f"Order must be non-negative, but was given as {order}.")
    if max_iter < 0:
        raise ValueError(f"Number of maximum iterations must be non-negative, but was given as {max_iter}.")
    if tol < 0:
        raise ValueError(f"Tolerance must be non-negative {tol}.")
    x = np.empty(order + 1)
    w = np.empty(order + 1)
    rc = lib().mfv2d_gll(order + 1, int(max_iter), float(tol), x.ctypes.data, w.ctypes.data)
    if rc == 13:  # MFV2D_NOT_CONVERGED -> the reference issues a RuntimeWarning (gauss_lobatto.c:249-255)
        import warnings

        warnings.warn(f"Not all Gauss-Lobatto nodes converged in {max_iter} iterations to tolerance {tol}", RuntimeWarning)
    else:
        check(rc)
    if cache is not None:
        cache._entries[order] = (float(tol), x.copy(), w.copy())
    return x, w


def lagrange1d(roots, x, out: np.ndarray | None = None) -> np.ndarray:
    """Lagrange basis on ``roots`` at ``x``; result shape ``x.shape + (n,)`` (src/polynomials/lagrange.c:173-265)."""
    roots = np.ascontiguousarray(roots, np.float64)
    xx = np.ascontiguousarray(x, np.float64)
    if roots.ndim != 1:
        raise ValueError("roots must be a 1D array")
    res = np.empty(xx.shape + (roots.size,)) if out is None else _check_array(out, None, "out")
    check(lib().mfv2d_lagrange1d(roots.size, roots.ctypes.data, xx.size, xx.ctypes.data, res.ctypes.data))
    return res


def dlagrange1d(roots, x, out: np.ndarray | None = None) -> np.ndarray:
    """First derivative of the Lagrange basis (src/polynomials/lagrange.c:110-171, 379-...)."""
    roots = np.ascontiguousarray(roots, np.float64)
    xx = np.ascontiguousarray(x, np.float64)
    if roots.ndim != 1:
        raise ValueError("roots must be a 1D array")
    res = np.empty(xx.shape + (roots.size,)) if out is None else _check_array(out, None, "out")
    check(lib().mfv2d_dlagrange1d(roots.size, roots.ctypes.data, xx.size, xx.ctypes.data, res.ctypes.data))
    return res


def compute_legendre(order: int, positions, out: np.ndarray | None = None) -> np.ndarray:
    """Legendre polynomials P_0..P_order -> ``(order + 1, *positions.shape)`` (src/polynomials/legendre.c:38-131)."""
    order = int(order)
    if order < 0:
        raise ValueError(f"Order must be non-negative, got {order}.")
    pos = np.ascontiguousarray(positions, np.float64)
    res = np.empty((order + 1,) + pos.shape) if out is None else _check_array(out, None, "out")
    if res.shape != (order + 1,) + pos.shape:
        raise ValueError("Output array does not have the expected shape.")
    check(lib().mfv2d_legendre(order, pos.size, pos.ctypes.data, res.ctypes.data))
    return res


def compute_legendre_device(order: int, positions: np.ndarray) -> np.ndarray:
    """Bulk variant of :func:`compute_legendre` evaluated on the device (many points)."""
    ctx = get_context()
    pos = np.ascontiguousarray(positions, np.float64)
    res = np.empty((order + 1,) + pos.shape)
    check(lib().mfv2d_cuda_legendre(ctx.handle, int(order), pos.size, pos.ctypes.data, res.ctypes.data, HOST), ctx.handle)
    return res


# =============================================================================================
# integration rule and basis (src/fem_space/integration_rule.c, basis.c)
# =============================================================================================
class IntegrationRule1D:
    """GLL rule with ``order + 1`` points (src/fem_space/integration_rule.c:19-67)."""

    def __init__(self, order: int):
        order = int(order)
        if order < 0:
            raise ValueError(f"Order must be non-negative, got {order}.")
        self.order = order
        self.nodes, self.weights = compute_gll(order)
        self.nodes.setflags(write=False)
        self.weights.setflags(write=False)

    def __repr__(self) -> str:
        return f"IntegrationRule1D({self.order})"


class Basis1D:
    """Nodal and edge basis tables of order ``order`` on ``rule`` (src/fem_space/basis.c:5-136):
    ``node[i, m] = l_i(xi_m)``, ``edge[j, m] = -sum_{i<=j} l_i'(xi_m)``."""

    def __init__(self, order: int, rule: IntegrationRule1D):
        order = int(order)
        if order < 1:
            raise ValueError(f"Order must be positive, got {order}.")
        if not isinstance(rule, IntegrationRule1D):
            raise TypeError("rule must be an IntegrationRule1D")
        self.order = order
        self.rule = rule
        q = rule.order + 1
        self.roots = np.empty(order + 1)
        pn, wg = np.empty(q), np.empty(q)
        self.node = np.empty((order + 1, q))
        self.edge = np.empty((order, q))
        check(lib().mfv2d_basis_tables(order, q, self.roots.ctypes.data, pn.ctypes.data, wg.ctypes.data,
                                       self.node.ctypes.data, self.edge.ctypes.data))
        for a in (self.roots, self.node, self.edge):
            a.setflags(write=False)

    def __repr__(self) -> str:
        return f"Basis1D({self.order}, {self.rule!r})"


class Basis2D:
    """Tensor-product basis (src/fem_space/basis.c:331-361)."""

    def __init__(self, basis_xi: Basis1D, basis_eta: Basis1D):
        if not isinstance(basis_xi, Basis1D) or not isinstance(basis_eta, Basis1D):
            raise TypeError("Basis2D needs two Basis1D objects")
        self.basis_xi = basis_xi
        self.basis_eta = basis_eta

    @property
    def orders(self) -> tuple[int, int]:
        return (self.basis_xi.order, self.basis_eta.order)

    @property
    def integration_orders(self) -> tuple[int, int]:
        return (self.basis_xi.rule.order, self.basis_eta.rule.order)

    def _square(self) -> tuple[int, int]:
        """(p, q) of the bucket; the element hot path requires order_1 == order_2
        (src/evaluation/element_system.c:43-51)."""
        (p1, p2), (i1, i2) = self.orders, self.integration_orders
        if p1 != p2 or i1 != i2:
            raise NotImplementedError(
                f"Element orders must be the same in both directions (got {p1}, {p2} / integration {i1}, {i2})."
            )
        return p1, i1 + 1


# =============================================================================================
# element form specification (src/evaluation/forms.c:150-882)
# =============================================================================================
def _form_size(order: int, o1: int, o2: int) -> int:
    """form_degrees_of_freedom_count (src/evaluation/forms.c:21-34)."""
    if order == FORM_ORDER_0:
        return (o1 + 1) * (o2 + 1)
    if order == FORM_ORDER_1:
        return o1 * (o2 + 1) + (o1 + 1) * o2
    if order == FORM_ORDER_2:
        return o1 * o2
    return 0


class _ElementFormSpecification:
    """Names and orders of the unknown forms of an element system; DoF layout = concatenation of the
    forms in this order (src/evaluation/forms.c:47-72)."""

    def __new__(cls, *specs: tuple[str, int]):
        # like the extension type, everything happens in __new__ (src/evaluation/forms.c:150-260): the reference's own subclass
        # ElementFormSpecification passes converted arguments to super().__new__ (python/mfv2d/system.py:21-24)
        self = super().__new__(cls)
        if not specs:
            raise TypeError("At least one form specification is required.")
        forms = []
        for s in specs:
            if not isinstance(s, tuple) or len(s) != 2:
                raise TypeError("Each form specification must be a tuple (name, order).")
            name, order = s
            if not isinstance(name, str):
                raise TypeError("Form name must be a string.")
            if len(name.encode()) > 31:  # MAXIMUM_FORM_NAME_LENGTH (src/evaluation/forms.c:199)
                raise ValueError(f"Form name {name!r} is longer than 31 bytes.")
            order = int(order)
            if order < FORM_ORDER_0 or order > FORM_ORDER_2:
                raise ValueError(f"Invalid form order: {order}")
            forms.append((name, order))
        names = [f[0] for f in forms]
        if len(set(names)) != len(names):
            raise ValueError("Form names must be unique.")
        self._forms = tuple(forms)
        return self

    @property
    def orders(self) -> tuple[int, ...]:
        return tuple(o for _, o in self._forms)

    @property
    def names(self) -> tuple[str, ...]:
        return tuple(n for n, _ in self._forms)

    def __iter__(self):
        return iter(self._forms)

    def __getitem__(self, idx):
        return self._forms[idx]

    def __len__(self) -> int:
        return len(self._forms)

    def __contains__(self, item) -> bool:
        return item in self._forms

    def form_size(self, idx, /, order_1: int, order_2: int) -> int:
        return _form_size(self._forms[idx][1], order_1, order_2)

    def form_sizes(self, order_1: int, order_2: int) -> tuple[int, ...]:
        return tuple(_form_size(o, order_1, order_2) for _, o in self._forms)

    def form_offset(self, idx, /, order_1: int, order_2: int) -> int:
        idx = range(len(self._forms))[idx]
        return sum(self.form_sizes(order_1, order_2)[:idx])

    def form_offsets(self, order_1: int, order_2: int) -> tuple[int, ...]:
        out = [0]
        for s in self.form_sizes(order_1, order_2):
            out.append(out[-1] + s)
        return tuple(out)

    def total_size(self, order_1: int, order_2: int) -> int:
        return sum(self.form_sizes(order_1, order_2))

    def index(self, value, /) -> int:
        return self._forms.index(value)

    def __repr__(self) -> str:
        return "_ElementFormSpecification(" + ", ".join(repr(f) for f in self._forms) + ")"


# =============================================================================================
# bytecode (src/evaluation/bytecode.c:50-275)
# =============================================================================================
def _encode_expressions(form_specs: _ElementFormSpecification, expressions):
    """Python tuples -> flat mfv2d_op_t array; returns (present, offsets, ops, field_form, callables).

    Mirrors convert_bytecode: operand counts are validated (MFV2D_OUT_OF_INSTRUCTIONS becomes
    ValueError as in the reference), INTERPROD fields are either callables or the name of an unknown."""
    nf = len(form_specs)
    if len(expressions) != nf:
        raise ValueError(f"Expression matrix has {len(expressions)} rows, but {nf} forms were specified.")
    present = np.zeros(nf * nf, np.uint8)
    offsets = np.zeros(nf * nf + 1, np.int32)
    ops: list[tuple] = []
    field_form: list[int] = []
    callables: list[Callable | None] = []
    names = form_specs.names
    for r, row in enumerate(expressions):
        if len(row) != nf:
            raise ValueError(f"Expression row {r} has {len(row)} entries, but {nf} forms were specified.")
        for c, expr in enumerate(row):
            b = r * nf + c
            if expr is not None:
                present[b] = 1
                for item in expr:
                    code = int(item[0])
                    if code in (MATOP_IDENTITY, MATOP_PUSH):
                        ops.append((code, 0, 0, 0, 0.0))
                    elif code == MATOP_SCALE:
                        ops.append((code, 0, 0, 0, float(item[1])))
                    elif code == MATOP_SUM:
                        ops.append((code, int(item[1]), 0, 0, 0.0))
                    elif code in (MATOP_MASS, MATOP_INCIDENCE):
                        ops.append((code, int(item[1]), int(bool(item[2])), 0, 0.0))
                    elif code == MATOP_INTERPROD:
                        _, order, fld, tr = item
                        if callable(fld):
                            field_form.append(-1)
                            callables.append(fld)
                        elif isinstance(fld, str):
                            if fld not in names:
                                raise ValueError(f"Unknown form {fld!r} used as an interior product field.")
                            field_form.append(names.index(fld))
                            callables.append(None)
                        else:
                            raise TypeError("Interior product field must be a callable or the name of an unknown.")
                        ops.append((code, int(order), len(field_form) - 1, int(bool(tr)), 0.0))
                    else:
                        raise ValueError(f"Invalid matrix operation code {code}.")
            offsets[b + 1] = len(ops)
    arr = np.zeros(max(len(ops), 1), OP_DTYPE)
    for i, o in enumerate(ops):
        arr[i] = o
    return present, offsets, arr, np.array(field_form, np.int32), callables


def check_bytecode(form_specs: _ElementFormSpecification, expression):
    """Round trip of one block's bytecode through the flat encoding (src/evaluation/bytecode.c:305-403)."""
    nf = len(form_specs)
    grid = tuple(tuple(expression if (r == 0 and c == 0) else None for c in range(nf)) for r in range(nf))
    _, offsets, ops, field_form, callables = _encode_expressions(form_specs, grid)
    out = []
    for o in ops[: offsets[1]]:
        code = int(o["op"])
        if code in (MATOP_IDENTITY, MATOP_PUSH):
            out.append((code,))
        elif code == MATOP_SCALE:
            out.append((code, float(o["k"])))
        elif code == MATOP_SUM:
            out.append((code, int(o["a"])))
        elif code in (MATOP_MASS, MATOP_INCIDENCE):
            out.append((code, int(o["a"]), bool(o["b"])))
        else:
            f = int(o["b"])
            fld = callables[f] if field_form[f] < 0 else form_specs.names[field_form[f]]
            out.append((code, int(o["a"]), fld, bool(o["c"])))
    return tuple(out)


class Plan:
    """A compiled system of bytecode programs for one (p, q) bucket (mfv2d_cuda_plan_create)."""

    def __init__(self, ctx: Context, p: int, q: int, form_specs: _ElementFormSpecification, expressions):
        self.ctx = ctx
        self.p, self.q = p, q
        self.basis = ctx.basis(p, q)
        present, offsets, ops, field_form, callables = _encode_expressions(form_specs, expressions)
        self.callables = callables
        self.field_form = field_form
        self.n_fields = len(field_form)
        orders = np.array(form_specs.orders, np.int32)
        h = C.c_void_p()
        check(
            lib().mfv2d_cuda_plan_create(ctx.handle, self.basis.handle, len(orders), orders.ctypes.data, present.ctypes.data,
                                         offsets.ctypes.data, ops.ctypes.data, self.n_fields,
                                         field_form.ctypes.data if self.n_fields else None, C.byref(h)),
            ctx.handle,
        )
        self.handle = h
        self.size = int(lib().mfv2d_cuda_plan_size(h))
        self.needs_dofs = bool(np.any(field_form >= 0))

    def __del__(self):
        try:
            if self.handle and self.ctx.handle:
                lib().mfv2d_cuda_plan_destroy(self.handle)
        except Exception:
            pass
        self.handle = None


def _plan_key(p, q, form_specs, expressions):
    def enc(item):
        return tuple(("fn", id(x)) if callable(x) else (int(x) if isinstance(x, (bool, np.bool_)) else x) for x in item)

    return (p, q, tuple(form_specs), tuple(tuple(None if e is None else tuple(enc(i) for i in e) for e in row) for row in expressions))


def get_plan(ctx: Context, p: int, q: int, form_specs, expressions) -> Plan:
    key = _plan_key(p, q, form_specs, expressions)
    plan = ctx._plans.get(key)
    if plan is None:
        plan = Plan(ctx, p, q, form_specs, expressions)
        if len(ctx._plans) > 256:
            ctx._plans.clear()
        ctx._plans[key] = plan
    return plan


# =============================================================================================
# element batches
# =============================================================================================
class ElementBatch:
    """``n`` elements of one (p, q) bucket resident on the device: the batched counterpart of ``n``
    ``ElementFemSpace2D`` objects.  ``corners`` is ``(n, 4, 2)`` float64, counter-clockwise from
    bottom-left (src/common/common.h:110-116); a DeviceArray of that shape is used in place."""

    def __init__(self, basis: Basis2D, corners, ctx: Context | None = None):
        self.ctx = ctx or get_context()
        self.basis_2d = basis
        self.p, self.q = basis._square()
        self.basis = self.ctx.basis(self.p, self.q)
        if isinstance(corners, DeviceArray):
            shape = corners.shape
        else:
            corners = _check_array(corners, 3, "corners")
            shape = corners.shape
        if len(shape) != 3 or shape[1:] != (4, 2):
            raise ValueError(f"corners must have the shape (n, 4, 2), got {shape}")
        self.n = int(shape[0])
        self._corners = corners  # kept for corners_host() (reconstruct_mesh_from_solution needs the point coordinates)
        h = C.c_void_p()
        check(
            lib().mfv2d_cuda_elements_create(self.ctx.handle, self.basis.handle, self.n, addr(corners), space_of(corners), C.byref(h)),
            self.ctx.handle,
        )
        self.handle = h

    def corners_host(self) -> np.ndarray:
        c = self._corners
        return c.get() if isinstance(c, DeviceArray) else np.asarray(c)

    def set_corners(self, corners) -> None:
        self._corners = corners
        check(lib().mfv2d_cuda_elements_set_corners(self.handle, addr(corners), space_of(corners)), self.ctx.handle)

    def __del__(self):
        try:
            if self.handle and self.ctx.handle:
                lib().mfv2d_cuda_elements_destroy(self.handle)
        except Exception:
            pass
        self.handle = None

    # ---- outputs: numpy in -> numpy out; ``out=DeviceArray`` keeps results on the device ----------
    def _out(self, out, shape):
        if out is None:
            return np.empty(shape)
        if tuple(out.shape) != tuple(shape):
            raise ValueError(f"output has the shape {tuple(out.shape)}, expected {tuple(shape)}")
        return out

    def form_size(self, order: int) -> int:
        return _form_size(order, self.p, self.p)

    @staticmethod
    def _in(a, shape, name: str):
        """an input buffer of exactly ``shape`` (numpy: made contiguous float64; DeviceArray: size checked): the kernels index
        that layout, so nothing shorter may reach them"""
        if isinstance(a, DeviceArray):
            if a.size != int(np.prod(shape)) or a.dtype != np.float64:
                raise ValueError(f"{name} must hold {tuple(shape)} float64 values, got {a.shape} {a.dtype}")
            return a
        a = np.ascontiguousarray(a, np.float64)
        if a.size != int(np.prod(shape)):
            raise ValueError(f"{name} must have the shape {tuple(shape)}, got {a.shape}")
        return a

    def mass(self, order: int, inverse: bool = False, out=None):
        """Mass matrices of 0-/1-/2-forms ``(n, n_k, n_k)`` (src/fem_space/fem_space.c:235-411)."""
        order = int(order)
        if order < FORM_ORDER_0 or order > FORM_ORDER_2:
            raise ValueError(f"Invalid form order: {order}")
        nk = self.form_size(order)
        out = self._out(out, (self.n, nk, nk))
        space = space_of(out)
        status = None
        if inverse:  # a singular block must surface as MFV2D_PIVOT_FAILED like in the reference, for device outputs too
            status = np.zeros(self.n, np.int32) if space == HOST else self.ctx.empty(self.n, np.int32)
        check(lib().mfv2d_cuda_mass_batch(self.handle, order, int(bool(inverse)), addr(out), addr(status), space), self.ctx.handle)
        if isinstance(status, DeviceArray):
            status = status.get()
        if status is not None and np.any(status):
            bad = int(np.flatnonzero(status)[0])
            raise _capi.Mfv2dCudaError(int(status[bad]), _capi.result_name(int(status[bad])), f"element {bad}: mass matrix could not be inverted")
        return out

    def jacobian(self, out=None):
        """``(n, 5, q, q)``: j00, j01, j10, j11, det (src/fem_space/fem_space.c:39-53)."""
        out = self._out(out, (self.n, 5, self.q, self.q))
        check(lib().mfv2d_cuda_jacobian_batch(self.handle, addr(out), space_of(out)), self.ctx.handle)
        return out

    def coordinates(self, out=None):
        """Physical coordinates of the quadrature points ``(n, 2, q, q)`` (integrating_fields.c:92-111)."""
        out = self._out(out, (self.n, 2, self.q, self.q))
        check(lib().mfv2d_cuda_coordinates_batch(self.handle, addr(out), space_of(out)), self.ctx.handle)
        return out

    def interior_product(self, start_order: int, field, transpose: bool = False, out=None):
        """Interior-product matrices with ``field`` ``(n, q, q, 2)`` (fem_space.c:546-635, 753-846)."""
        n0, n1, n2 = (self.form_size(o) for o in (1, 2, 3))
        shape = {FORM_ORDER_1: (n0, n1), FORM_ORDER_2: (n1, n2)}.get(int(start_order))
        if shape is None:
            raise ValueError(f"Invalid starting form order {start_order} for an interior product.")
        if transpose:
            shape = shape[::-1]
        out = self._out(out, (self.n,) + shape)
        field = self._in(field, (self.n, self.q, self.q, 2), "field")
        check(lib().mfv2d_cuda_interprod_batch(self.handle, int(start_order), int(bool(transpose)), addr(field), addr(out),
                                               space_of(field, out)), self.ctx.handle)
        return out

    def reconstruct(self, order: int, dofs, out=None):
        """Field values at the quadrature points from DoFs ``(n, n_k)`` (integrating_fields.c:238-352)."""
        shape = (self.n, self.q, self.q, 2) if int(order) == FORM_ORDER_1 else (self.n, self.q, self.q)
        out = self._out(out, shape)
        dofs = self._in(dofs, (self.n, self.form_size(int(order))), "dofs")
        check(lib().mfv2d_cuda_reconstruct_batch(self.handle, int(order), addr(dofs), addr(out), space_of(dofs, out)), self.ctx.handle)
        return out

    def reconstruct_grid(self, order: int, dofs, xi, eta, out=None):
        """Values of a k-form on the tensor grid ``(eta[a], xi[b])`` of reference points from its primal DoFs ``(n, n_k)``:
        ``(n, len(eta), len(xi))`` or ``(..., 2)`` for 1-forms (``reconstruct``, python/mfv2d/mimetic2d.py:1172-1279, for
        all elements of the batch in one launch)."""
        xi = np.ascontiguousarray(xi, np.float64).ravel()
        eta = np.ascontiguousarray(eta, np.float64).ravel()
        shape = (self.n, eta.size, xi.size, 2) if int(order) == FORM_ORDER_1 else (self.n, eta.size, xi.size)
        out = self._out(out, shape)
        dofs = self._in(dofs, (self.n, self.form_size(int(order))), "dofs")
        check(lib().mfv2d_cuda_reconstruct_grid_batch(self.handle, int(order), addr(dofs), xi.size, xi.ctypes.data, eta.size, eta.ctypes.data,
                                                      addr(out), space_of(dofs, out)), self.ctx.handle)
        return out

    def dual_dofs(self, order: int, function, out=None):
        """Dual degrees of freedom ``(n, n_k)`` of ``function(x, y)`` on every element of the batch
        (``element_dual_dofs``, python/mfv2d/mimetic2d.py:1003-1123).  The callable is evaluated ONCE on the stacked
        coordinates ``(n, q, q)``; ``function`` may also be the array of its values at the quadrature points."""
        order = int(order)
        shape = (self.n, self.q, self.q, 2) if order == FORM_ORDER_1 else (self.n, self.q, self.q)
        if callable(function):
            xy = self.coordinates()
            xy = xy.get() if isinstance(xy, DeviceArray) else xy
            vals = np.asarray(function(xy[:, 0], xy[:, 1]), np.float64)
        else:
            vals = function
        if not isinstance(vals, DeviceArray):
            vals = np.ascontiguousarray(np.broadcast_to(np.asarray(vals, np.float64), shape)) if np.shape(vals) != shape else np.ascontiguousarray(vals, np.float64)
        if tuple(vals.shape) != shape:
            raise ValueError(f"function values have the shape {tuple(vals.shape)}, expected {shape}")
        out = self._out(out, (self.n, self.form_size(order)))
        check(lib().mfv2d_cuda_dual_dofs_batch(self.handle, order, addr(vals), addr(out), space_of(vals, out)), self.ctx.handle)
        return out

    def evaluate_callable_fields(self, plan: Plan) -> np.ndarray | None:
        """Evaluate the plan's callable fields on the stacked quadrature coordinates of the whole batch
        (ONE Python call per field instead of one per element, integrating_fields.c:78-126)."""
        if plan.n_fields == 0 or all(f is None for f in plan.callables):
            return None
        xy = self.coordinates()
        vals = np.zeros((plan.n_fields, self.n, self.q, self.q, 2))
        for i, fn in enumerate(plan.callables):
            if fn is None:
                continue
            v = np.asarray(fn(xy[:, 0], xy[:, 1]), np.float64)
            if v.shape != (self.n, self.q, self.q, 2):
                raise ValueError(f"Field callable returned the shape {v.shape}, expected {(self.n, self.q, self.q, 2)}.")
            vals[i] = v
        return vals


def _check_field_values(field_values, plan: "Plan", batch: "ElementBatch"):
    """caller-evaluated fields: ``(n_fields, n, q, q, 2)`` - the kernels index that layout, a smaller buffer must not reach them"""
    if field_values is not None and plan.n_fields:
        want = plan.n_fields * batch.n * batch.q * batch.q * 2
        if int(field_values.size) != want:
            raise ValueError(f"field_values must have the shape {(plan.n_fields, batch.n, batch.q, batch.q, 2)} ({want} entries), "
                             f"got {int(field_values.size)} entries")
    return field_values


def _as_dofs(dofs, n: int, size: int, needed: bool):
    if dofs is None:
        if needed:
            raise ValueError("degrees_of_freedom are required: a field of the system is reconstructed from an unknown.")
        return None
    if isinstance(dofs, DeviceArray):
        if dofs.shape != (n, size):
            raise ValueError(f"degrees_of_freedom must have the shape {(n, size)}, got {dofs.shape}")
        return dofs
    dofs = _check_array(dofs, 2, "degrees_of_freedom")
    if dofs.shape != (n, size):
        raise ValueError(f"degrees_of_freedom must have the shape {(n, size)}, got {dofs.shape}")
    return dofs


def compute_element_matrices(form_specs, expressions, batch: ElementBatch, degrees_of_freedom=None, field_values=None, out=None):
    """Batched ``compute_element_matrix``: ``(n, N_e, N_e)``.  ``out`` may be a DeviceArray."""
    plan = get_plan(batch.ctx, batch.p, batch.q, form_specs, expressions)
    dofs = _as_dofs(degrees_of_freedom, batch.n, plan.size, plan.needs_dofs)
    if field_values is None:
        field_values = batch.evaluate_callable_fields(plan)
    _check_field_values(field_values, plan, batch)
    out = batch._out(out, (batch.n, plan.size, plan.size))
    space = space_of(out)
    if space == DEVICE:
        if isinstance(field_values, np.ndarray):
            field_values = batch.ctx.to_device(field_values)
        if isinstance(dofs, np.ndarray):
            dofs = batch.ctx.to_device(dofs)
    check(lib().mfv2d_cuda_element_matrix_batch(plan.handle, batch.handle, addr(dofs) if plan.needs_dofs else None,
                                                addr(field_values), addr(out), space), batch.ctx.handle)
    return out


def compute_element_vectors(form_specs, expressions, batch: ElementBatch, degrees_of_freedom, field_values=None, out=None):
    """Batched ``compute_element_vector``: ``(n, N_e)``, evaluated matrix-free."""
    plan = get_plan(batch.ctx, batch.p, batch.q, form_specs, expressions)
    dofs = _as_dofs(degrees_of_freedom, batch.n, plan.size, True)
    if field_values is None:
        field_values = batch.evaluate_callable_fields(plan)
    _check_field_values(field_values, plan, batch)
    out = batch._out(out, (batch.n, plan.size))
    space = space_of(out)
    if space == DEVICE:
        if isinstance(field_values, np.ndarray):
            field_values = batch.ctx.to_device(field_values)
        if isinstance(dofs, np.ndarray):
            dofs = batch.ctx.to_device(dofs)
    elif isinstance(dofs, DeviceArray):
        raise TypeError("cannot mix host (numpy) and device (DeviceArray) buffers in one call")
    check(lib().mfv2d_cuda_element_vector_batch(plan.handle, batch.handle, addr(dofs), addr(field_values), addr(out), space),
          batch.ctx.handle)
    return out


# =============================================================================================
# per-element objects and functions (batches of one)
# =============================================================================================
class ElementFemSpace2D:
    """Element space = basis + geometry (src/fem_space/element_fem_space.c:19-58, getters :221-443)."""

    def __init__(self, basis: Basis2D, corners: np.ndarray):
        if not isinstance(basis, Basis2D):
            raise TypeError("basis must be a Basis2D")
        corners = _check_array(corners, 2, "corners")
        if corners.shape != (4, 2):
            raise ValueError(f"Array corners dimension did not match expected value (expected (4, 2), got {corners.shape}).")
        self.basis_2d = basis
        self._corners = corners.copy()
        self._batch: ElementBatch | None = None
        self._cache: dict = {}

    @property
    def batch(self) -> ElementBatch:
        if self._batch is None:
            self._batch = ElementBatch(self.basis_2d, self._corners.reshape(1, 4, 2))
        return self._batch

    basis_xi = property(lambda self: self.basis_2d.basis_xi)
    basis_eta = property(lambda self: self.basis_2d.basis_eta)
    corners = property(lambda self: self._corners.copy())
    orders = property(lambda self: self.basis_2d.orders)
    integration_orders = property(lambda self: self.basis_2d.integration_orders)
    order_1 = property(lambda self: self.basis_2d.orders[0])
    order_2 = property(lambda self: self.basis_2d.orders[1])

    def mass_from_order(self, order, inverse: bool = False) -> np.ndarray:
        key = (int(order), bool(inverse))
        if key not in self._cache:  # the lazy per-element cache of element_fem_space.c:445-524
            self._cache[key] = self.batch.mass(int(order), bool(inverse))[0]
        return self._cache[key].copy()

    mass_node = property(lambda self: self.mass_from_order(FORM_ORDER_0))
    mass_edge = property(lambda self: self.mass_from_order(FORM_ORDER_1))
    mass_surf = property(lambda self: self.mass_from_order(FORM_ORDER_2))
    mass_node_inv = property(lambda self: self.mass_from_order(FORM_ORDER_0, True))
    mass_edge_inv = property(lambda self: self.mass_from_order(FORM_ORDER_1, True))
    mass_surf_inv = property(lambda self: self.mass_from_order(FORM_ORDER_2, True))

    @property
    def jacobian(self):
        """Tuple of the four Jacobian component arrays (q, q) (element_fem_space.c:170-205)."""
        j = self.batch.jacobian()[0]
        return (j[0].copy(), j[1].copy(), j[2].copy(), j[3].copy())

    @property
    def jacobian_determinant(self) -> np.ndarray:
        return self.batch.jacobian()[0, 4].copy()


def compute_element_matrix(form_specs, expressions, element_fem_space: ElementFemSpace2D, degrees_of_freedom=None,
                           stack_memory: int = 1 << 24) -> np.ndarray:
    """src/evaluation/element_system.c:13-212.  ``stack_memory`` is accepted and ignored (no arena)."""
    if not isinstance(form_specs, _ElementFormSpecification):
        raise TypeError("form_specs must be an _ElementFormSpecification")
    if not isinstance(element_fem_space, ElementFemSpace2D):
        raise TypeError("element_fem_space must be an ElementFemSpace2D")
    batch = element_fem_space.batch
    n = form_specs.total_size(batch.p, batch.p)
    dofs = None
    if degrees_of_freedom is not None:
        dofs = _check_array(degrees_of_freedom, 1, "degrees_of_freedom")
        if dofs.shape != (n,):
            raise ValueError(f"Array degrees_of_freedom dimension 0 did not match expected value (expected {n}, got {dofs.shape[0]}).")
        dofs = dofs.reshape(1, n)
    plan = get_plan(batch.ctx, batch.p, batch.q, form_specs, expressions)
    if dofs is None and plan.needs_dofs:
        dofs = np.zeros((1, n))
    return compute_element_matrices(form_specs, expressions, batch, dofs)[0]


def compute_element_vector(form_specs, expressions, element_fem_space: ElementFemSpace2D, degrees_of_freedom,
                           stack_memory: int = 1 << 24) -> np.ndarray:
    """src/evaluation/element_system.c:245-411."""
    if not isinstance(form_specs, _ElementFormSpecification):
        raise TypeError("form_specs must be an _ElementFormSpecification")
    if not isinstance(element_fem_space, ElementFemSpace2D):
        raise TypeError("element_fem_space must be an ElementFemSpace2D")
    batch = element_fem_space.batch
    n = form_specs.total_size(batch.p, batch.p)
    dofs = _check_array(degrees_of_freedom, 1, "degrees_of_freedom")
    if dofs.shape != (n,):
        raise ValueError(f"Array degrees_of_freedom dimension 0 did not match expected value (expected {n}, got {dofs.shape[0]}).")
    return compute_element_vectors(form_specs, expressions, batch, dofs.reshape(1, n))[0]


def compute_element_mass_matrix(order, corners: np.ndarray, basis: Basis2D, inverse: bool = False) -> np.ndarray:
    """src/evaluation/element_system.c:670-758."""
    return ElementFemSpace2D(basis, np.ascontiguousarray(corners, np.float64)).mass_from_order(order, inverse)


def compute_integrating_fields(fem_space: ElementFemSpace2D, form_specs: _ElementFormSpecification, field_orders: tuple,
                               field_information: tuple, degrees_of_freedom: np.ndarray) -> tuple[np.ndarray, ...]:
    """src/evaluation/integrating_fields.c:355-496: callables are evaluated on the mapped quadrature
    points, integers name the unknown whose DoFs are reconstructed."""
    batch = fem_space.batch
    n = form_specs.total_size(batch.p, batch.p)
    dofs = _check_array(degrees_of_freedom, 1, "degrees_of_freedom")
    if dofs.shape != (n,):
        raise ValueError("degrees_of_freedom do not match the form specification")
    if len(field_orders) != len(field_information):
        raise ValueError("field_orders and field_information must have the same length")
    offs = form_specs.form_offsets(batch.p, batch.p)
    out = []
    xy = None
    for order, info in zip(field_orders, field_information):
        order = int(order)
        if callable(info):
            if xy is None:
                xy = batch.coordinates()[0]
            v = np.asarray(info(xy[0], xy[1]), np.float64)
            want = (batch.q, batch.q, 2) if order == FORM_ORDER_1 else (batch.q, batch.q)
            if v.shape != want:
                raise ValueError(f"Field callable returned the shape {v.shape}, expected {want}.")
            out.append(np.ascontiguousarray(v))
        else:
            # the name of an unknown (bytecode.c:131-199 resolves INTERPROD fields the same way) or its index
            idx = form_specs.names.index(info) if isinstance(info, str) else int(info)
            if form_specs.orders[idx] != order:
                raise ValueError("Field order does not match the order of the unknown it is taken from.")
            seg = np.ascontiguousarray(dofs[offs[idx] : offs[idx + 1]]).reshape(1, -1)
            out.append(batch.reconstruct(order, seg)[0])
    return tuple(out)


def check_incidence(x, /, order: int, in_form: int, transpose: bool, right: bool) -> np.ndarray:
    """Apply an incidence matrix without forming it (src/evaluation/incidence.c:593-637).
    ``in_form`` = 0 (E10) or 1 (E21) as in the reference's test helper."""
    ctx = get_context()
    x = np.ascontiguousarray(x, np.float64)
    if x.ndim != 2:
        raise ValueError("x must be a matrix")
    kind = int(in_form) + (2 if transpose else 0)
    p = int(order)
    rows = {0: 2 * p * (p + 1), 1: p * p, 2: (p + 1) ** 2, 3: 2 * p * (p + 1)}[kind]
    cols = {0: (p + 1) ** 2, 1: 2 * p * (p + 1), 2: 2 * p * (p + 1), 3: p * p}[kind]
    if right:
        if x.shape[1] != rows:
            raise ValueError(f"Matrix has {x.shape[1]} columns, the incidence matrix {rows} rows.")
        out = np.empty((x.shape[0], cols))
        n = x.shape[0]
    else:
        if x.shape[0] != cols:
            raise ValueError(f"Matrix has {x.shape[0]} rows, the incidence matrix {cols} columns.")
        out = np.empty((rows, x.shape[1]))
        n = x.shape[1]
    check(lib().mfv2d_cuda_incidence_apply(ctx.handle, p, kind, int(bool(right)), 1, n, x.ctypes.data, out.ctypes.data, HOST), ctx.handle)
    return out


def _compute_matrix_inverse(x, /) -> np.ndarray:
    """src/algebra/matrix.c:546-576."""
    ctx = get_context()
    a = np.ascontiguousarray(x, np.float64)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError("Input must be a square matrix.")
    out = np.empty_like(a)
    status = np.zeros(1, np.int32)
    check(lib().mfv2d_cuda_inverse_batch(ctx.handle, a.shape[0], 1, a.ctypes.data, out.ctypes.data, status.ctypes.data, HOST), ctx.handle)
    if status[0]:
        raise _capi.Mfv2dCudaError(int(status[0]), _capi.result_name(int(status[0])), "matrix could not be inverted")
    return out


def inverse_definite(x, /) -> np.ndarray:
    """Inverse of ONE large definite matrix by the blocked, unpivoted Gauss-Jordan elimination that serves the exact level of
    the multilevel preconditioner (``mfv2d_cuda_inverse_definite``, dense.cu); no counterpart in the reference, whose dense
    inverse is the pivoted LU of :func:`_compute_matrix_inverse`.  Raises when a pivot vanishes."""
    import ctypes as C

    ctx = get_context()
    a = np.ascontiguousarray(x, np.float64)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError("Input must be a square matrix.")
    out = np.empty_like(a)
    status = C.c_int(0)
    check(lib().mfv2d_cuda_inverse_definite(ctx.handle, a.shape[0], a.ctypes.data, out.ctypes.data, C.byref(status), HOST), ctx.handle)
    return out


def _solve_linear_system(mat, rhs, /) -> np.ndarray:
    """src/algebra/matrix.c:578-660: solve ``mat @ x = rhs`` (rhs vector or matrix)."""
    ctx = get_context()
    a = np.ascontiguousarray(mat, np.float64)
    b = np.ascontiguousarray(rhs, np.float64)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError("System matrix must be square.")
    vec = b.ndim == 1
    b2 = b.reshape(a.shape[0], -1) if vec else b
    if b2.ndim != 2 or b2.shape[0] != a.shape[0]:
        raise ValueError("Right-hand side does not match the matrix.")
    b2 = np.ascontiguousarray(b2)
    out = np.empty_like(b2)
    status = np.zeros(1, np.int32)
    check(lib().mfv2d_cuda_solve_batch(ctx.handle, a.shape[0], b2.shape[1], 1, a.ctypes.data, b2.ctypes.data, out.ctypes.data,
                                       status.ctypes.data, HOST), ctx.handle)
    if status[0]:
        raise _capi.Mfv2dCudaError(int(status[0]), _capi.result_name(int(status[0])), "matrix is singular")
    return out.reshape(b.shape)


def compute_element_projector(form_orders, basis_in: Basis2D, space_out: ElementFemSpace2D, dual: bool = False):
    """src/evaluation/element_system.c:480-643: per form, the mass between the output basis (rows) and
    ``basis_in`` (columns) on ``space_out``'s geometry; ``dual=False`` multiplies by the inverse output mass.
    ``form_orders`` is an ``_ElementFormSpecification`` (as in the reference) or a sequence of form orders; the
    result is a tuple aligned with it (forms of equal order share one array, :621-630)."""
    if not isinstance(basis_in, Basis2D):
        raise TypeError("basis_in must be a Basis2D")
    if not isinstance(space_out, ElementFemSpace2D):
        raise TypeError("space_out must be an ElementFemSpace2D")
    orders = form_orders.orders if isinstance(form_orders, _ElementFormSpecification) else tuple(int(o) for o in form_orders)
    batch = space_out.batch
    p_in, q_in = basis_in._square()
    if q_in != batch.q:
        raise ValueError("Input and output bases must share the integration rule.")
    b_in = batch.ctx.basis(p_in, q_in)
    done: dict[int, np.ndarray] = {}
    for order in orders:
        if order in done:
            continue
        if order < FORM_ORDER_0 or order > FORM_ORDER_2:
            raise ValueError(f"Invalid form order: {order}")
        n_out, n_in = _form_size(order, batch.p, batch.p), _form_size(order, p_in, p_in)
        out = np.empty((1, n_out, n_in))
        status = np.zeros(1, np.int32)
        check(lib().mfv2d_cuda_projector_batch(batch.handle, b_in.handle, order, int(bool(dual)), out.ctypes.data,
                                               status.ctypes.data, HOST), batch.ctx.handle)
        if status[0]:
            raise _capi.Mfv2dCudaError(int(status[0]), _capi.result_name(int(status[0])), "output mass matrix could not be inverted")
        done[order] = out[0]
    return tuple(done[o] for o in orders)


def compute_element_projectors(order: int, basis_in: Basis2D, batch: ElementBatch, dual: bool = False, out=None):
    """Batched projector of one form order for every element of ``batch``: ``(n, n_k(out), n_k(in))``."""
    p_in, q_in = basis_in._square()
    if q_in != batch.q:
        raise ValueError("Input and output bases must share the integration rule.")
    b_in = batch.ctx.basis(p_in, q_in)
    n_out, n_in = _form_size(int(order), batch.p, batch.p), _form_size(int(order), p_in, p_in)
    out = batch._out(out, (batch.n, n_out, n_in))
    space = space_of(out)
    status = np.zeros(batch.n, np.int32) if space == HOST else None
    check(lib().mfv2d_cuda_projector_batch(batch.handle, b_in.handle, int(order), int(bool(dual)), addr(out), addr(status), space),
          batch.ctx.handle)
    if status is not None and np.any(status):
        bad = int(np.flatnonzero(status)[0])
        raise _capi.Mfv2dCudaError(int(status[bad]), _capi.result_name(int(status[bad])), f"element {bad}: output mass matrix could not be inverted")
    return out


def element_dual_dofs(order, element_space: ElementFemSpace2D, function) -> np.ndarray:
    """``mfv2d.mimetic2d.element_dual_dofs`` (python/mfv2d/mimetic2d.py:1003-1123) for one element: batch of one."""
    return element_space.batch.dual_dofs(int(order), function)[0]


def compute_element_rhs(system, space_or_batch) -> np.ndarray:
    """``compute_element_rhs`` (python/mfv2d/solve_system.py:129-160): the projections of the right-hand-side functions
    of every equation, concatenated in equation order.  ``system`` is the reference's ``KFormSystem`` (or anything with
    ``equations[i].weight.order`` and ``equations[i].right.explicit_terms`` = ``(k, projection)`` pairs whose
    projections carry ``.func``).  With an :class:`ElementBatch` the result is ``(n, N_e)`` and every function is
    called once on the stacked coordinates; with an :class:`ElementFemSpace2D` it is the reference's ``(N_e,)``.
    The projections of one equation are linear in the function, so the weighted sum is formed on the function values
    and projected once."""
    single = isinstance(space_or_batch, ElementFemSpace2D)
    batch = space_or_batch.batch if single else space_or_batch
    xy = None
    parts = []
    for eq in system.equations:
        order = int(eq.weight.order)
        shape = (batch.n, batch.q, batch.q, 2) if order == FORM_ORDER_1 else (batch.n, batch.q, batch.q)
        vals = None
        for k, f in eq.right.explicit_terms:
            fn = getattr(f, "func", None)  # only element projections carry a function (solve_system.py:117-119)
            if fn is None:
                continue
            if xy is None:
                xy = batch.coordinates()
                xy = xy.get() if isinstance(xy, DeviceArray) else xy
            v = float(k) * np.broadcast_to(np.asarray(fn(xy[:, 0], xy[:, 1]), np.float64), shape)
            vals = v if vals is None else vals + v
        parts.append(np.zeros((batch.n, batch.form_size(order))) if vals is None else batch.dual_dofs(order, np.ascontiguousarray(vals)))
    out = np.concatenate(parts, axis=1)
    return out[0] if single else out


def _mass_expressions(form_specs: _ElementFormSpecification, inverse: bool):
    """block-diagonal bytecode: (MASS, order, inverse) on the diagonal, nothing elsewhere"""
    n = len(form_specs)
    return [[[(MATOP_MASS, int(form_specs[i][1]), bool(inverse))] if i == j else None for j in range(n)] for i in range(n)]


def compute_element_dual_from_primal(form_specs: _ElementFormSpecification, primal, space_or_batch):
    """``compute_element_dual_from_primal`` (python/mfv2d/solve_system.py:302-325): ``dual_k = M_k primal_k`` for every
    form of the element - the matrix-free apply of the block-diagonal mass system (no mass matrix is formed).
    ``primal`` is ``(N_e,)`` with an :class:`ElementFemSpace2D` or ``(n, N_e)`` with an :class:`ElementBatch`."""
    if isinstance(space_or_batch, ElementFemSpace2D):
        return compute_element_vector(form_specs, _mass_expressions(form_specs, False), space_or_batch, primal)
    return compute_element_vectors(form_specs, _mass_expressions(form_specs, False), space_or_batch, primal)


def compute_element_primal_from_dual(form_specs: _ElementFormSpecification, dual, space_or_batch):
    """``compute_element_primal_from_dual`` (python/mfv2d/solve_system.py:328-351): ``primal_k = M_k^-1 dual_k``."""
    if isinstance(space_or_batch, ElementFemSpace2D):
        return compute_element_vector(form_specs, _mass_expressions(form_specs, True), space_or_batch, dual)
    return compute_element_vectors(form_specs, _mass_expressions(form_specs, True), space_or_batch, dual)


def reconstruct(fem_space: "ElementFemSpace2D", form_order: int, dofs, xi, eta) -> np.ndarray:
    """``reconstruct`` (python/mfv2d/mimetic2d.py:1172-1279) for points that form a tensor grid: ``xi`` of shape ``(1, m)``
    (or ``(m,)``) and ``eta`` of shape ``(k, 1)`` as ``reconstruct_mesh_from_solution`` passes them; returns ``(k, m)`` or
    ``(k, m, 2)``."""
    xi, eta = np.asarray(xi, np.float64), np.asarray(eta, np.float64)
    if not ((xi.ndim == 1 or (xi.ndim == 2 and xi.shape[0] == 1)) and (eta.ndim == 1 or (eta.ndim == 2 and eta.shape[1] == 1))):
        raise ValueError("reconstruct: xi must be (m,) or (1, m) and eta (k,) or (k, 1) - a tensor grid of reference points")
    c = np.ascontiguousarray(dofs, np.float64)
    if c.ndim != 1:
        raise ValueError("Coefficient array must be one dimensional.")
    if c.size != _form_size(int(form_order), fem_space.batch.p, fem_space.batch.p):
        raise ValueError("The number of degrees of freedom is not correct.")
    return fem_space.batch.reconstruct_grid(int(form_order), c[None, :], xi.ravel(), eta.ravel())[0]


def reconstruct_mesh_from_solution(form_specs: _ElementFormSpecification, recon_order: int | None, batch: "ElementBatch", solution):
    """``reconstruct_mesh_from_solution`` (python/mfv2d/solve_system.py:163-271) without the graphics stack (pyvista / VTK:
    unavailable on this box): the point coordinates and the reconstructed unknowns on the uniform
    ``(recon_order + 1)^2`` grid of every element as plain arrays - ``{"x", "y": (n, m, m), name: (n, m, m[, 2]), ...}`` in the
    element / row-major point order in which the reference concatenates them into its ``UnstructuredGrid``."""
    m = (batch.p if recon_order is None else int(recon_order)) + 1
    nodes = np.linspace(-1.0, +1.0, m)
    sol = solution if isinstance(solution, DeviceArray) else np.ascontiguousarray(solution, np.float64).reshape(batch.n, -1)
    host = sol.get() if isinstance(sol, DeviceArray) else sol
    corners = batch.corners_host()
    xi, eta = nodes[None, None, :], nodes[None, :, None]
    w = [(1 - xi) * (1 - eta) / 4, (1 + xi) * (1 - eta) / 4, (1 + xi) * (1 + eta) / 4, (1 - xi) * (1 + eta) / 4]
    out = {"x": sum(corners[:, k, 0][:, None, None] * w[k] for k in range(4)), "y": sum(corners[:, k, 1][:, None, None] * w[k] for k in range(4))}
    for idx, (name, order) in enumerate(zip(form_specs.names, form_specs.orders)):
        o = form_specs.form_offset(idx, batch.p, batch.p)
        dofs = np.ascontiguousarray(host[:, o:o + form_specs.form_size(idx, batch.p, batch.p)])
        out[name] = batch.reconstruct_grid(int(order), dofs, nodes, nodes)
    return out


# host containers of the module (src/algebra/svector.c, src/algebra/crs_matrix.c): same names as ``mfv2d._mfv2d``
from .sparse import MatrixCRS, SparseVector  # noqa: E402,F401
