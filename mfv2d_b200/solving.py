"""Element-block linear system and Krylov solvers on the device.

Mirror of the reference's ``LinearSystem`` / ``DenseVector`` / ``TraceVector`` (src/algebra/
system_objects.c, sparse_system.c) and of python/mfv2d/solving.py (``solve_schur_iterative``,
``solve_pcg_iterative``, ``ConvergenceSettings``).  The reference crosses Python -> C once per vector
operation; here a whole solve is one call into libmfv2d_cuda.so and the vectors never leave the GPU.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Sequence
from dataclasses import dataclass

import numpy as np

from ._capi import DEVICE, HOST, check, lib
from .runtime import Context, DeviceArray, addr, get_context, space_of
from .sparse import MatrixCRS, SparseVector, constraint_blocks_to_coo


@dataclass(frozen=True)
class ConvergenceSettings:
    """python/mfv2d/solving.py:157-171."""

    maximum_iterations: int = 100
    absolute_tolerance: float = 1e-6
    relative_tolerance: float = 1e-5


@dataclass(frozen=True)
class ElementConstraint:
    """python/mfv2d/mimetic2d.py:804-821."""

    i_e: int
    dofs: np.ndarray
    coeffs: np.ndarray


@dataclass(frozen=True, init=False)
class Constraint:
    """python/mfv2d/mimetic2d.py:853-873."""

    rhs: float
    element_constraints: tuple

    def __init__(self, rhs: float, *element_constraints: ElementConstraint) -> None:
        object.__setattr__(self, "rhs", float(rhs))
        object.__setattr__(self, "element_constraints", element_constraints)


def constraints_to_coo(constraints: Sequence[Constraint], offsets: np.ndarray):
    """Flatten Constraint objects into the global COO (row, column, value) of N, sorted by (row, col)."""
    rows, cols, vals = [], [], []
    for ic, con in enumerate(constraints):
        for ec in con.element_constraints:
            d = np.asarray(ec.dofs, np.int64)
            rows.append(np.full(d.size, ic, np.int64))
            cols.append(d + int(offsets[ec.i_e]))
            vals.append(np.asarray(ec.coeffs, np.float64))
    if not rows:
        return np.zeros(0, np.int64), np.zeros(0, np.int64), np.zeros(0)
    row, col, val = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    order = np.lexsort((col, row))
    return row[order], col[order], val[order]


def _is_block_pairs(obj) -> bool:
    return (isinstance(obj, (tuple, list)) and len(obj) > 0
            and all(isinstance(b, (tuple, list)) and len(b) == 2 and isinstance(b[1], MatrixCRS) for b in obj))


def _check_trace_lists(pairs, n_trace: int, offsets: np.ndarray, indices: np.ndarray) -> None:
    """``offsets`` / ``indices`` of the extension's constructor (sparse_system.h:27-32) against the constraint blocks:
    multiplier ``i`` must list exactly the elements whose block has a non-empty row ``i``."""
    if offsets.ndim != 1 or offsets.size != n_trace + 1 or indices.ndim != 1 or (offsets.size and offsets[-1] != indices.size):
        raise ValueError("trace offsets / indices do not describe one element list per constraint row")
    rows = [np.stack([m.nonempty_rows.astype(np.int64), np.full(m.nonempty_rows.size, e, np.int64)], 1) for e, (_, m) in enumerate(pairs)]
    have = np.concatenate(rows) if rows else np.zeros((0, 2), np.int64)
    want = np.stack([np.repeat(np.arange(n_trace), np.diff(offsets.astype(np.int64))), indices.astype(np.int64)], 1)
    key = lambda a: a[np.lexsort((a[:, 1], a[:, 0]))]
    if have.shape != want.shape or not np.array_equal(key(have), key(want)):
        raise ValueError("trace offsets / indices disagree with the non-empty rows of the constraint blocks")


class LinearSystem:
    """``[[A, N^T], [N, 0]]`` with block-diagonal ``A`` (one dense block per element) and the
    constraint matrix ``N`` (system_new / system_fill_out, src/algebra/system_objects.c:73-199).

    ``element_matrices``: sequence of ``(N_e, N_e)`` arrays, a ``(n, N_e, N_e)`` array, or a DeviceArray of
    that shape (e.g. straight out of ``compute_element_matrices(..., out=DeviceArray)``).
    ``constraints``: sequence of :class:`Constraint`, or a COO triple ``(row, col, val)`` over global
    dense columns together with ``n_trace``."""

    def __new__(cls, *args, **kwargs):
        # the reference's Python subclass builds the blocks in ITS __new__ and hands them to super().__new__
        # (python/mfv2d/solving.py:41-97); Python then calls __init__ with the subclass's own arguments, which the
        # five-argument form below accepts
        return object.__new__(cls)

    def __init__(self, element_matrices, constraints=None, n_trace: int | None = None, *args, ctx: Context | None = None):
        if isinstance(element_matrices, (int, np.integer)):
            # the reference's own call, LinearSystem(n_elem, form_specs, orders, element_matrices, constraints)
            # (python/mfv2d/solving.py:41-60): block sizes follow from the matrices themselves
            n_elem, (element_matrices, constraints), n_trace = int(element_matrices), args[:2], None
            if len(element_matrices) != n_elem:
                raise ValueError(f"expected {n_elem} element matrices, got {len(element_matrices)}")
        elif _is_block_pairs(element_matrices):
            # the extension type's own call, LinearSystem(blocks, offsets, indices) with blocks = ((A_e, MatrixCRS), ...)
            # (system_new, src/algebra/system_objects.c:73-199): ``offsets`` / ``indices`` list the elements of
            # every multiplier and are implied by the constraint blocks; they are checked, not needed
            pairs, trace_offsets, trace_indices = element_matrices, constraints, n_trace
            element_matrices = [a for a, _ in pairs]
            n_trace = pairs[0][1].shape[0] if pairs else 0
            sizes0 = np.array([np.shape(a)[0] for a in element_matrices], np.int64)
            for a, m in pairs:
                if m.shape != (n_trace, np.shape(a)[1]):
                    raise ValueError(f"constraint block of shape {m.shape} does not match ({n_trace}, {np.shape(a)[1]})")
            constraints = constraint_blocks_to_coo([m for _, m in pairs], np.concatenate([[0], np.cumsum(sizes0)]))
            if trace_offsets is not None:
                _check_trace_lists(pairs, n_trace, np.asarray(trace_offsets), np.asarray(trace_indices))
        if ctx is None and args and isinstance(args[0], Context):
            ctx = args[0]  # LinearSystem(matrices, constraints, n_trace, ctx)
        self.ctx = ctx or get_context()
        if isinstance(element_matrices, DeviceArray):
            n, ne, ne2 = element_matrices.shape
            assert ne == ne2
            sizes = np.full(n, ne, np.int64)
            blocks, space = element_matrices, DEVICE
        else:
            mats = [np.ascontiguousarray(m, np.float64) for m in element_matrices]
            for m in mats:
                if m.ndim != 2 or m.shape[0] != m.shape[1]:
                    raise ValueError("element matrices must be square")
            sizes = np.array([m.shape[0] for m in mats], np.int64)
            blocks = np.concatenate([m.ravel() for m in mats]) if mats else np.zeros(0)
            space = HOST
        self._blocks = blocks  # keep a borrowed device buffer alive
        self.offsets = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        self.n_elem = int(sizes.size)
        if isinstance(constraints, tuple) and len(constraints) == 3 and isinstance(constraints[0], np.ndarray):
            row, col, val = (np.ascontiguousarray(constraints[0], np.int64), np.ascontiguousarray(constraints[1], np.int64),
                             np.ascontiguousarray(constraints[2], np.float64))
            if n_trace is None:
                n_trace = int(row.max()) + 1 if row.size else 0
            self.constraint_rhs = None
        else:
            row, col, val = constraints_to_coo(constraints, self.offsets)
            n_trace = len(constraints)
            self.constraint_rhs = np.array([c.rhs for c in constraints], np.float64)
        if not (row.size == col.size == val.size):
            raise ValueError(f"constraint triplets of different lengths: {row.size} rows, {col.size} columns, {val.size} values")
        self.con = (row, col, val)
        self.n_trace = int(n_trace)
        self.n_dense = int(self.offsets[-1])
        h = C.c_void_p()
        check(
            lib().mfv2d_cuda_system_create(self.ctx.handle, self.n_elem, self.offsets.ctypes.data, addr(blocks), space, self.n_trace,
                                           row.size, row.ctypes.data, col.ctypes.data, val.ctypes.data, C.byref(h)),
            self.ctx.handle,
        )
        self.handle = h

    def __del__(self):
        try:
            if self.handle and self.ctx.handle:
                lib().mfv2d_cuda_system_destroy(self.handle)
        except Exception:
            pass
        self.handle = None

    def get_dense_blocks(self) -> tuple:
        """The element matrices ``A_e`` as host arrays (system_objects.c:201-244)."""
        if isinstance(self._blocks, DeviceArray):
            flat = self._blocks.get().ravel()
        else:
            flat = self._blocks
        o, sq = self.offsets, np.concatenate([[0], np.cumsum(np.diff(self.offsets) ** 2)])
        return tuple(flat[sq[e]:sq[e + 1]].reshape(o[e + 1] - o[e], o[e + 1] - o[e]).copy() for e in range(self.n_elem))

    def get_constraint_blocks(self) -> tuple:
        """The per-element constraint blocks ``N_e`` as ``n_trace x N_e`` MatrixCRS (system_objects.c:246-290)."""
        row, col, val = self.con
        elem = np.searchsorted(self.offsets, col, side="right") - 1
        order = np.lexsort((col, row, elem))
        row, col, val, elem = row[order], col[order], val[order], elem[order]
        first = np.searchsorted(elem, np.arange(self.n_elem + 1))
        out = []
        for e in range(self.n_elem):
            lo, hi = first[e], first[e + 1]
            m = MatrixCRS(self.n_trace, int(self.offsets[e + 1] - self.offsets[e]))
            m.set_from_data(val[lo:hi], col[lo:hi] - self.offsets[e], np.bincount(row[lo:hi], minlength=self.n_trace))
            out.append(m)
        return tuple(out)

    # ---- operators (sparse_system.c:3-207); numpy in -> numpy out, DeviceArray in -> DeviceArray out ----
    def _apply(self, fn, x, n_in: int, n_out: int, out=None, dense_out: bool = True):
        if isinstance(x, _SystemVector):  # DenseVector / TraceVector in -> vector object out (``out`` is written in place)
            if x.parent is not self or (out is not None and out.parent is not self):
                raise ValueError("vector does not belong to this system (MFV2D_BAD_VECTOR_PARENT)")
            if out is None:
                out = DenseVector(self) if dense_out else TraceVector(self)
            self._apply(fn, x.data, n_in, n_out, out.data)
            return out
        if isinstance(x, DeviceArray):
            out = out if out is not None else self.ctx.empty(n_out)
        else:
            x = np.ascontiguousarray(x, np.float64)
            out = out if out is not None else np.empty(n_out)
        if x.size != n_in or out.size != n_out:
            raise ValueError("vector does not belong to this system (MFV2D_BAD_VECTOR_PARENT)")
        check(fn(self.handle, addr(x), addr(out), space_of(x, out)), self.ctx.handle)
        return out

    def apply_diagonal(self, x, out=None):
        return self._apply(lib().mfv2d_cuda_system_apply_diagonal, x, self.n_dense, self.n_dense, out)

    def apply_diagonal_inverse(self, x, out=None):
        return self._apply(lib().mfv2d_cuda_system_apply_diagonal_inverse, x, self.n_dense, self.n_dense, out)

    def apply_trace(self, x, out=None):
        return self._apply(lib().mfv2d_cuda_system_apply_trace, x, self.n_dense, self.n_trace, out, dense_out=False)

    def apply_trace_transpose(self, t, out=None):
        return self._apply(lib().mfv2d_cuda_system_apply_trace_transpose, t, self.n_trace, self.n_dense, out)

    def apply_full_trace_system(self, t, out=None):
        """``N A^-1 N^T t`` (solving.py:99-123) through the condensed per-element blocks."""
        return self._apply(lib().mfv2d_cuda_system_apply_schur, t, self.n_trace, self.n_trace, out, dense_out=False)

    def set_interface(self, shared_index: np.ndarray, slot: np.ndarray, n_slots: int, owned: np.ndarray) -> None:
        """Multi-GPU: declare the multipliers shared with other ranks (see include/mfv2d_cuda.h)."""
        si = np.ascontiguousarray(shared_index, np.int64)
        sl = np.ascontiguousarray(slot, np.int64)
        ow = np.ascontiguousarray(owned, np.uint8)
        if ow.size != self.n_trace or si.size != sl.size:
            raise ValueError(f"set_interface: owned must have {self.n_trace} entries (got {ow.size}), shared_index and slot the same length "
                             f"({si.size}, {sl.size})")
        check(lib().mfv2d_cuda_system_set_interface(self.handle, si.size, si.ctypes.data, sl.ctypes.data, int(n_slots), ow.ctypes.data),
              self.ctx.handle)


    def connect_peers(self, holder_ptr: np.ndarray, holder_rank: np.ndarray, all_gather, n_coarse: int = 0) -> None:
        """Multi-GPU: map the exchange blocks of the other ranks so that the solver's kernels write interface sums and
        Krylov scalars straight into peer memory over NVLink (no NCCL call inside an iteration).  ``all_gather(bytes) ->
        list[bytes]`` (rank-major) is any host-side gather of the IPC handles, e.g. ``torch.distributed.all_gather_object``
        over gloo - plumbing, like the NCCL unique id.  Collective: every rank calls it after ``set_interface``."""
        n = lib().mfv2d_cuda_peer_handle_size()
        buf = (C.c_char * n)()
        check(lib().mfv2d_cuda_system_peer_export(self.handle, int(n_coarse), buf), self.ctx.handle)
        handles = all_gather(bytes(buf))
        blob = b"".join(handles)
        hp = np.ascontiguousarray(holder_ptr, np.int64)
        hr = np.ascontiguousarray(holder_rank, np.int32)
        if hp.size < 1 or int(hp[-1]) != hr.size or len(blob) != n * len(handles):
            raise ValueError("connect_peers: holder_ptr must be a CSR pointer over holder_rank and every rank must contribute one handle")
        hbuf = (C.c_char * len(blob)).from_buffer_copy(blob)
        check(lib().mfv2d_cuda_system_peer_connect(self.handle, hbuf, hp.size - 1, hp.ctypes.data, hr.ctypes.data if hr.size else None),
              self.ctx.handle)

    def peer_error(self) -> bool:
        return bool(lib().mfv2d_cuda_system_peer_error(self.handle))

    def peer_stats(self) -> np.ndarray:
        """Timeline of this rank's distributed CG (``MFV2D_PEER_DEBUG`` bit 8): nanoseconds accumulated before each of the
        16 stamps since the last call (``mfv2d_cuda_system_peer_stats``)."""
        out = np.zeros(16, np.uint64)
        check(lib().mfv2d_cuda_system_peer_stats(self.handle, out.ctypes.data), self.ctx.handle)
        return out

    def set_coarse_space(self, aggregates: np.ndarray, z: np.ndarray, n_agg, exact_last: bool = True) -> None:
        """Aggregate hierarchy of the multilevel preconditioner (``precondition="multilevel"``): ``aggregates[level, multiplier]``
        (aggregate id or -1) and the near-null-space value ``z`` of every multiplier of this rank
        (``structured.coarse_hierarchy_1form`` + ``localize_hierarchy``).  Collective when partitioned."""
        ag = np.ascontiguousarray(np.atleast_2d(aggregates), np.int32)
        zz = np.ascontiguousarray(z, np.float64)
        na = np.ascontiguousarray(n_agg, np.int64)
        if ag.shape[1] != self.n_trace or zz.size != self.n_trace or na.size != ag.shape[0]:
            raise ValueError("aggregates must be (levels, n_trace), z (n_trace,), n_agg (levels,)")
        check(lib().mfv2d_cuda_system_set_coarse_space(self.handle, ag.shape[0], na.ctypes.data, ag.ctypes.data, zz.ctypes.data,
                                                       int(bool(exact_last))), self.ctx.handle)


PRECONDITIONERS = {False: 0, None: 0, "none": 0, True: 1, "jacobi": 1, "multilevel": 2}


def _check_solver_vectors(system: "LinearSystem", rhs, constraints) -> None:
    """The C entry points read ``n_dense`` / ``n_trace`` doubles: a shorter buffer must never reach them."""
    if int(rhs.size) != system.n_dense or int(constraints.size) != system.n_trace:
        raise ValueError(f"right-hand sides of {int(rhs.size)} and {int(constraints.size)} entries do not match the system "
                         f"({system.n_dense} dense unknowns, {system.n_trace} multipliers) (MFV2D_BAD_VECTOR_PARENT)")


def solve_schur_iterative(system: LinearSystem, rhs, constraints, convergence: ConvergenceSettings, precondition=False, out=None):
    """python/mfv2d/solving.py:439-500: CG on ``N A^-1 N^T`` for the multipliers, then back substitution.
    Returns ``(x, lambda, residual, iterations)`` like the reference.  ``precondition``: ``False`` (the reference's
    method), ``True`` / ``"jacobi"``, or ``"multilevel"`` (Jacobi + aggregate hierarchy, see ``LinearSystem.set_coarse_space``).
    ``out=(x, lambda)``: buffers for the solution (e.g. page-locked arrays of ``Context.pinned`` on the host-buffer path)."""
    if isinstance(rhs, _SystemVector):  # DenseVector / TraceVector in and out, like the reference
        xd, ld, r, i = solve_schur_iterative(system, rhs.data, constraints.data, convergence, precondition)
        return DenseVector.__new__(DenseVector)._init(system, xd), TraceVector.__new__(TraceVector)._init(system, ld), r, i
    dev = isinstance(rhs, DeviceArray)
    if dev:
        _check_solver_vectors(system, rhs, constraints)
        x, lam = out if out is not None else (system.ctx.empty(system.n_dense), system.ctx.empty(system.n_trace))
        if int(x.size) != system.n_dense or int(lam.size) != system.n_trace:
            raise ValueError("out must be device arrays of the dense and trace sizes")
    else:
        rhs = np.ascontiguousarray(rhs, np.float64)
        constraints = np.ascontiguousarray(constraints, np.float64)
        _check_solver_vectors(system, rhs, constraints)
        x, lam = out if out is not None else (np.empty(system.n_dense), np.empty(system.n_trace))
        if x.shape != (system.n_dense,) or lam.shape != (system.n_trace,) or x.dtype != np.float64 or lam.dtype != np.float64:
            raise ValueError("out must be float64 arrays of the dense and trace sizes")
    res, it = C.c_double(), C.c_int()
    check(
        lib().mfv2d_cuda_solve_schur_cg(system.handle, addr(rhs), addr(constraints), int(convergence.maximum_iterations),
                                        float(convergence.absolute_tolerance), float(convergence.relative_tolerance),
                                        PRECONDITIONERS[precondition], addr(x), addr(lam), C.byref(res), C.byref(it),
                                        space_of(rhs, constraints, x, lam)),
        system.ctx.handle,
    )
    return x, lam, res.value, it.value


def solve_schur_bicgstab(system: LinearSystem, rhs, constraints, convergence: ConvergenceSettings, precondition=True):
    """The Schur-complement solve for NON-symmetric element blocks (advection terms): BiCGStab on ``N A^-1 N^T`` on the
    device, then the back substitution.  Same arguments and return value as :func:`solve_schur_iterative`; the reference
    has no iterative solver for such systems (production path: SuperLU, python/mfv2d/solve_system_2d.py:258-282)."""
    if isinstance(rhs, _SystemVector):
        xd, ld, r, i = solve_schur_bicgstab(system, rhs.data, constraints.data, convergence, precondition)
        return DenseVector.__new__(DenseVector)._init(system, xd), TraceVector.__new__(TraceVector)._init(system, ld), r, i
    if isinstance(rhs, DeviceArray):
        _check_solver_vectors(system, rhs, constraints)
        x, lam = system.ctx.empty(system.n_dense), system.ctx.empty(system.n_trace)
    else:
        rhs = np.ascontiguousarray(rhs, np.float64)
        constraints = np.ascontiguousarray(constraints, np.float64)
        _check_solver_vectors(system, rhs, constraints)
        x, lam = np.empty(system.n_dense), np.empty(system.n_trace)
    res, it = C.c_double(), C.c_int()
    check(
        lib().mfv2d_cuda_solve_schur_bicgstab(system.handle, addr(rhs), addr(constraints), int(convergence.maximum_iterations),
                                              float(convergence.absolute_tolerance), float(convergence.relative_tolerance),
                                              PRECONDITIONERS[precondition], addr(x), addr(lam), C.byref(res), C.byref(it),
                                              space_of(rhs, constraints, x, lam)),
        system.ctx.handle,
    )
    return x, lam, res.value, it.value


def solve_pcg_iterative(system: LinearSystem, rhs, constraints, convergence: ConvergenceSettings):
    """python/mfv2d/solving.py:634-684: PCG on the saddle system with the block-Jacobi ``A^-1 (+) I``."""
    if isinstance(rhs, _SystemVector):
        xd, ld, r, i = solve_pcg_iterative(system, rhs.data, constraints.data, convergence)
        return DenseVector.__new__(DenseVector)._init(system, xd), TraceVector.__new__(TraceVector)._init(system, ld), r, i
    dev = isinstance(rhs, DeviceArray)
    if dev:
        _check_solver_vectors(system, rhs, constraints)
        x, lam = system.ctx.empty(system.n_dense), system.ctx.empty(system.n_trace)
    else:
        rhs = np.ascontiguousarray(rhs, np.float64)
        constraints = np.ascontiguousarray(constraints, np.float64)
        _check_solver_vectors(system, rhs, constraints)
        x, lam = np.empty(system.n_dense), np.empty(system.n_trace)
    res, it = C.c_double(), C.c_int()
    check(
        lib().mfv2d_cuda_solve_pcg(system.handle, addr(rhs), addr(constraints), int(convergence.maximum_iterations),
                                   float(convergence.absolute_tolerance), float(convergence.relative_tolerance), addr(x), addr(lam),
                                   C.byref(res), C.byref(it), space_of(rhs, constraints, x, lam)),
        system.ctx.handle,
    )
    return x, lam, res.value, it.value


def vec_axpby(a: float, x, b: float, y, out=None, ctx: Context | None = None):
    """``out = a x + b y`` (DenseVector/TraceVector add, subtract, scale; system_objects.c:945-1300)."""
    ctx = ctx or get_context()
    if out is None:
        out = ctx.empty(x.size) if isinstance(x, DeviceArray) else np.empty(x.size)
    if int(y.size) != int(x.size) or int(out.size) != int(x.size):
        raise ValueError(f"vec_axpby: operands of {int(x.size)}, {int(y.size)} and {int(out.size)} entries")
    check(lib().mfv2d_cuda_vec_axpby(ctx.handle, x.size, float(a), addr(x), float(b), addr(y), addr(out), space_of(x, y, out)), ctx.handle)
    return out


def vec_dot(x, y, ctx: Context | None = None) -> float:
    ctx = ctx or get_context()
    if int(y.size) != int(x.size):
        raise ValueError(f"vec_dot: operands of {int(x.size)} and {int(y.size)} entries")
    r = C.c_double()
    check(lib().mfv2d_cuda_vec_dot(ctx.handle, x.size, addr(x), addr(y), C.byref(r), space_of(x, y)), ctx.handle)
    return r.value


# The reference also keeps ``solve_cg_iterative`` / ``solve_gmres_iterative`` (python/mfv2d/solving.py:553-631) as Python
# loops; nothing in it calls or tests them, its CG divides by zero on its own example system (b^T K b = 0) and its
# GMRES applies the operator in place and returns a wrong vector (recorded in tests/golden/pcg.json).  They are not
# mirrored: the device solvers are ``solve_schur_iterative`` (CG on the Schur complement) and ``solve_pcg_iterative``.
# ---- DenseVector / TraceVector (src/algebra/system_objects.c:725-2152): device-resident counterparts --------------
class _SystemVector:
    """Common part of the two vector types: a DeviceArray that belongs to one LinearSystem.  The static methods write
    into ``out`` and return it, like the reference's; operands of another system raise ``ValueError``
    (MFV2D_BAD_VECTOR_PARENT, sparse_system.c:6-9)."""

    _n_attr = ""

    def __init__(self, parent: LinearSystem, data: DeviceArray):
        self._parent, self.data = parent, data

    @property
    def parent(self) -> LinearSystem:
        return self._parent

    @classmethod
    def _length(cls, parent: LinearSystem) -> int:
        return int(getattr(parent, cls._n_attr))

    @classmethod
    def _from_host(cls, parent: LinearSystem, values):
        n = cls._length(parent)
        if values is None:
            values = np.zeros(n)
        values = np.ascontiguousarray(values, np.float64).ravel()
        if values.size != n:
            raise ValueError(f"expected {n} values for this system, got {values.size}")
        return cls.__new__(cls)._init(parent, parent.ctx.to_device(values))

    def _init(self, parent, data):
        _SystemVector.__init__(self, parent, data)
        return self

    @staticmethod
    def _check(*vs) -> LinearSystem:
        parent = vs[0]._parent
        for v in vs:
            if type(v) is not type(vs[0]) or v._parent is not parent:
                raise ValueError("vectors do not belong to the same system (MFV2D_BAD_VECTOR_PARENT)")
        return parent

    def as_merged(self) -> np.ndarray:
        return self.data.get()

    def copy(self):
        out = type(self).__new__(type(self))._init(self._parent, self._parent.ctx.empty(self.data.size))
        out.set_from(self)
        return out

    def set_from(self, other) -> None:
        self._check(self, other)
        if self.data.size:
            vec_axpby(1.0, other.data, 0.0, other.data, out=self.data, ctx=self._parent.ctx)

    @classmethod
    def dot(cls, v1, v2, /) -> float:
        parent = cls._check(v1, v2)
        return vec_dot(v1.data, v2.data, ctx=parent.ctx) if v1.data.size else 0.0

    @classmethod
    def add(cls, v1, v2, out, k: float, /):
        parent = cls._check(v1, v2, out)
        if out.data.size:
            vec_axpby(1.0, v1.data, float(k), v2.data, out=out.data, ctx=parent.ctx)
        return out

    @classmethod
    def subtract(cls, v1, v2, out, k: float, /):
        return cls.add(v1, v2, out, -float(k))

    @classmethod
    def scale(cls, v, x: float, out, /):
        parent = cls._check(v, out)
        if out.data.size:
            vec_axpby(float(x), v.data, 0.0, v.data, out=out.data, ctx=parent.ctx)
        return out

    def __str__(self) -> str:
        return f"{type(self).__name__}({self.as_merged()})"


class DenseVector(_SystemVector):
    """``DenseVector(parent, *blocks)`` (system_objects.c:725-830): one block of values per element, stored
    concatenated on the device; no blocks = zero vector."""

    _n_attr = "n_dense"

    def __init__(self, parent: LinearSystem, /, *blocks):
        if blocks and len(blocks) != parent.n_elem:
            raise ValueError(f"expected {parent.n_elem} blocks, got {len(blocks)}")
        for i, b in enumerate(blocks):
            if np.size(b) != parent.offsets[i + 1] - parent.offsets[i]:
                raise ValueError(f"block {i} does not have the size of the system's block")
        values = np.concatenate([np.asarray(b, np.float64).ravel() for b in blocks]) if blocks else None
        self._init(parent, self._from_host(parent, values).data)

    def as_split(self) -> tuple:
        merged, o = self.as_merged(), self._parent.offsets
        return tuple(merged[o[i]:o[i + 1]].copy() for i in range(self._parent.n_elem))


class TraceVector(_SystemVector):
    """``TraceVector(parent[, values])`` (system_objects.c:1408-1523): one value per multiplier (the reference keeps a
    copy per sharing element and divides its dot products by the sharing count; the merged form needs neither)."""

    _n_attr = "n_trace"

    def __init__(self, parent: LinearSystem, value=None, /):
        self._init(parent, self._from_host(parent, value).data)

    def as_split(self) -> tuple:
        """per element: a SparseVector of dimension ``n_trace`` with the values of its multipliers
        (trace_vector_as_split, system_objects.c:1560-1600)"""
        merged = self.as_merged()
        row, col, _ = self._parent.con
        elem = np.searchsorted(self._parent.offsets, col, side="right") - 1
        order = np.lexsort((row, elem))
        row, elem = row[order], elem[order]
        first = np.searchsorted(elem, np.arange(self._parent.n_elem + 1))
        out = []
        for e in range(self._parent.n_elem):
            idx = np.unique(row[first[e]:first[e + 1]])
            out.append(SparseVector._make(self._parent.n_trace, idx.astype(np.uint64), merged[idx]))
        return tuple(out)


def compute_linear_system(form_specs, expressions, batch, constraints, n_trace: int | None = None, degrees_of_freedom=None,
                          field_values=None) -> LinearSystem:
    """The per-element loop of ``compute_linear_system`` (python/mfv2d/solve_system.py:988-1008) collapsed: all element
    matrices of one ``(p, q)`` bucket in ONE launch, left on the device and handed to :class:`LinearSystem` without a
    host copy.  ``batch`` is an ``ElementBatch``; ``constraints`` as for ``LinearSystem``.
    Mixed-order meshes: one call per bucket for the matrices (``compute_element_matrices``), then
    ``LinearSystem(list_of_host_matrices, constraints)``."""
    from . import _mfv2d as M

    ne = form_specs.total_size(batch.p, batch.p)
    out = batch.ctx.empty((batch.n, ne, ne))
    M.compute_element_matrices(form_specs, expressions, batch, degrees_of_freedom, field_values, out=out)
    return LinearSystem(out, constraints, n_trace, ctx=batch.ctx)


class ConstraintConditioner:
    """Well-conditioned replacement of ill-scaled constraint rows (the hanging-side rows of non-conforming meshes).

    The reference expresses a hanging side through inverse interpolation matrices (``np.linalg.inv``,
    python/mfv2d/continuity.py:256-265); between elements whose orders differ by several units those rows carry coefficients
    up to 1e8 and are nearly linearly dependent, which its direct solver (SuperLU with partial pivoting) absorbs but which
    puts cond(N A^-1 N^T) beyond 1e16 for any Krylov method on the multipliers.  Rows that couple the same set of elements
    span the same constraint space as any basis of their row space, so each such group ``B`` (g x c) is replaced by the
    orthonormal rows ``Q^T`` of its QR factorisation ``B^T = Q R``:  ``B = R^T Q^T``.  The solution ``x`` is unchanged;
    right-hand sides and multipliers map as ``phi' = R^-T phi`` and ``lambda = R^-1 lambda'``.  Groups whose rows are plain
    conforming rows (two entries of magnitude one) are left untouched, bit for bit."""

    def __init__(self, row, col, val, n_trace: int, offsets):
        row, col, val = np.asarray(row, np.int64), np.asarray(col, np.int64), np.asarray(val, np.float64)
        offsets = np.asarray(offsets, np.int64)
        elem = np.searchsorted(offsets, col, side="right") - 1
        starts = np.searchsorted(row, np.arange(n_trace + 1))
        keys: dict = {}
        plain = np.ones(n_trace, bool)
        for i in range(n_trace):
            lo, hi = starts[i], starts[i + 1]
            keys.setdefault(tuple(np.unique(elem[lo:hi]).tolist()), []).append(i)
            plain[i] = hi - lo <= 2 and bool(np.all(np.abs(val[lo:hi]) == 1.0))
        new_rows, new_cols, new_vals = [], [], []
        self.blocks: list[tuple[np.ndarray, np.ndarray]] = []  # (rows of the group, R^T)
        keep = np.ones(row.size, bool)
        for rows_g in keys.values():
            rows_g = np.array(rows_g)
            if np.all(plain[rows_g]):
                continue
            sel = np.concatenate([np.arange(starts[i], starts[i + 1]) for i in rows_g])
            keep[sel] = False
            cols_g, inv = np.unique(col[sel], return_inverse=True)
            local_row = np.concatenate([np.full(starts[i + 1] - starts[i], k) for k, i in enumerate(rows_g)])
            B = np.zeros((rows_g.size, cols_g.size))
            B[local_row, inv] = val[sel]
            Q, R = np.linalg.qr(B.T)  # B^T = Q R
            if np.min(np.abs(np.diag(R))) <= 1e-13 * np.max(np.abs(np.diag(R))):
                raise ValueError("a group of constraint rows is rank deficient")
            self.blocks.append((rows_g, R.T.copy()))
            rr, cc = np.nonzero(np.abs(Q.T) > 1e-15)
            new_rows.append(rows_g[rr])
            new_cols.append(cols_g[cc])
            new_vals.append(Q.T[rr, cc])
        r = np.concatenate([row[keep]] + new_rows)
        c = np.concatenate([col[keep]] + new_cols)
        v = np.concatenate([val[keep]] + new_vals)
        o = np.lexsort((c, r))
        self.row, self.col, self.val = r[o], c[o], v[o]

    def rhs(self, phi: np.ndarray) -> np.ndarray:
        """phi' = R^-T phi"""
        out = np.array(phi, np.float64)
        for rows_g, rt in self.blocks:
            out[rows_g] = np.linalg.solve(rt, out[rows_g])
        return out

    def multipliers(self, lam: np.ndarray) -> np.ndarray:
        """lambda = R^-1 lambda'"""
        out = np.array(lam, np.float64)
        for rows_g, rt in self.blocks:
            out[rows_g] = np.linalg.solve(rt.T, out[rows_g])
        return out
