"""The callers' loops of the hot path, batched and device-backed.

The reference drives its C extension from per-element Python loops and solves the global system with SciPy's SuperLU:

* ``non_linear_solve_run``  python/mfv2d/solve_system.py:353-485   residual with one ``compute_element_vector`` per element,
                                                                   correction from ``system_decomp.solve``
* the trapezoidal time march python/mfv2d/solve_system_2d.py:325-397  carry terms, one ``compute_element_dual_from_primal`` per
                                                                   element and step

Here the per-element loops are ONE launch per bucket (``compute_element_vectors``), the element matrices never leave the
device (``solving.compute_linear_system``) and the global factorisation is replaced by the Schur-complement Krylov solvers of
``solving`` (CG for symmetric element blocks, BiCGStab for systems with advection).  Same update formulas, same stopping
rule; the functions work on arrays (``ElementBatch``) instead of ``Mesh`` objects.
"""
from __future__ import annotations

from collections.abc import Callable, Sequence

import numpy as np

from . import _mfv2d as M
from . import solving as SV
from ._capi import check, lib
from .runtime import DeviceArray


def _apply_codes(form_specs, codes, batch: M.ElementBatch, dofs: np.ndarray, field_values, cache: dict | None):
    """``compute_element_vectors`` with the callable fields of ``codes`` evaluated ONCE per (codes, batch) and kept on the device:
    the reference evaluates them per element and call (integrating_fields.c:78-126); on a 512 x 512 order-5 mesh one evaluation
    of a numpy callable on the 26 M quadrature points costs more than the Krylov solve of a whole time step."""
    ctx = batch.ctx
    if field_values is None and cache is not None:
        key = id(codes)
        if key not in cache:
            plan = M.get_plan(ctx, batch.p, batch.q, form_specs, codes)
            fv = batch.evaluate_callable_fields(plan)
            cache[key] = (codes, ctx.to_device(fv) if fv is not None else None)  # the codes object is kept alive with its id
        field_values = cache[key][1]
    if field_values is None or not isinstance(field_values, np.ndarray):
        out = ctx.empty((batch.n, dofs.shape[1]))
        return M.compute_element_vectors(form_specs, codes, batch, dofs, field_values, out=out).get()
    return np.asarray(M.compute_element_vectors(form_specs, codes, batch, dofs, field_values))


def non_linear_solve_run(form_specs, lhs_codes, rhs_codes, batch: M.ElementBatch, system: SV.LinearSystem, explicit_vec: np.ndarray,
                         constraint_rhs: np.ndarray, solution: np.ndarray, lagrange: np.ndarray, *, max_iterations: int, atol: float,
                         rtol: float, relax: float = 1.0, max_mag: float | None = None, time_carry_index: np.ndarray | None = None,
                         time_carry_term: np.ndarray | None = None, solve: Callable | None = None, field_values=None,
                         reduce_max: Callable[[float], float] | None = None, field_cache: dict | None = None, sg_operator=None,
                         fine_scales: np.ndarray | None = None, residual_history: list | None = None):
    """``non_linear_solve_run`` (python/mfv2d/solve_system.py:353-485) on one bucket of elements.

    ``explicit_vec`` ``(n * N_e,)`` is the dense right-hand side (weak boundary terms included), ``constraint_rhs`` the one of
    the constraint rows; ``solution`` / ``lagrange`` are updated in place and returned with the iteration count and the last
    maximum residual.  ``system`` holds the element matrices of ``lhs_codes`` (the linearised operator of the correction
    solve, ``system_decomp`` of the reference); ``solve(system, r_dense, r_trace) -> (dx, dlambda)`` defaults to the
    Schur-complement CG, pass ``solving.solve_schur_bicgstab`` based solvers for non-symmetric blocks.  On a partitioned
    system (one strip of the mesh per rank, ``system`` connected to its peers) every rank passes its local vectors, the
    GLOBAL ``max_mag`` and ``reduce_max`` = the host's max-all-reduce: the stopping decision is then the same on all ranks.

    ``sg_operator`` (:class:`mfv2d_b200.vms.SuyashGreenOperator`) adds the variational-multiscale closure of
    solve_system.py:432-440: before the residual is tested the unresolved contributions are recomputed from the current
    solution (starting from ``fine_scales``) and their coarse image is taken off the residual; the function then returns a
    fifth value, the unresolved scales it ended with, like the reference.  ``residual_history``: a list that receives the
    maximum residual of every iteration (``return_all_residuals`` of the reference)."""
    n, ne = batch.n, form_specs.total_size(batch.p, batch.p)
    base = np.array(explicit_vec, np.float64)
    if time_carry_term is not None:
        base[time_carry_index] += time_carry_term
    if max_mag is None:
        max_mag = float(np.max(np.abs(np.concatenate([explicit_vec, constraint_rhs])))) if explicit_vec.size else 0.0
    if solve is None:
        conv = SV.ConvergenceSettings(20000, 1e300, 1e-13)

        def solve(sysm, rd, rt):
            return SV.solve_schur_iterative(sysm, rd, rt, conv, precondition="jacobi")[:2]

    if field_cache is None:
        field_cache = {}
    iter_cnt, max_residual = 0, 0.0
    unresolved = fine_scales
    while iter_cnt < max_iterations:
        lhs = _apply_codes(form_specs, lhs_codes, batch, solution.reshape(n, ne), field_values, field_cache)
        if rhs_codes is not None:
            lhs = lhs - _apply_codes(form_specs, rhs_codes, batch, solution.reshape(n, ne), field_values, field_cache)
        main_value = np.asarray(lhs).ravel() + system.apply_trace_transpose(lagrange)
        r_dense = base - main_value
        r_trace = constraint_rhs - system.apply_trace(solution)
        if sg_operator is not None:
            sg_operator.update_nonlinear_advection(solution)
            unresolved = sg_operator.compute_unresolved_contributions(solution, unresolved)
            image = sg_operator.fine_results_to_coarse_dofs(unresolved, dual=True)
            r_dense -= image[: r_dense.size]
            r_trace -= image[r_dense.size:]
        max_residual = float(max(np.max(np.abs(r_dense), initial=0.0), np.max(np.abs(r_trace), initial=0.0)))
        if reduce_max is not None:
            max_residual = float(reduce_max(max_residual))
        if residual_history is not None:
            residual_history.append(max_residual)
        if not (max_residual > atol and max_residual > max_mag * rtol):
            break
        dx, dl = solve(system, r_dense, r_trace)
        solution += relax * dx
        lagrange += relax * dl
        iter_cnt += 1
    if sg_operator is not None:
        return solution, lagrange, iter_cnt, max_residual, unresolved
    return solution, lagrange, iter_cnt, max_residual


def vec_absmax(x, ctx) -> float:
    """``np.max(np.abs(x))`` of a host or device vector (0 for an empty one, NaN if any entry is)."""
    import ctypes as C

    r = C.c_double()
    dev = isinstance(x, DeviceArray)
    check(lib().mfv2d_cuda_vec_absmax(ctx.handle, x.size, x.ptr if dev else x.ctypes.data, C.byref(r), 1 if dev else 0), ctx.handle)
    return r.value


def _form_slices(form_specs, p: int, march_forms: Sequence[int]):
    return [(form_specs.form_offset(int(u), p, p), form_specs.form_size(int(u), p, p)) for u in sorted(march_forms)]


def _masked_copy(ctx, n: int, ne: int, slices, a: float, x: DeviceArray, b: float, y: DeviceArray, out: DeviceArray):
    """``out[:, form] = a x[:, form] + b y[:, form]`` for the marched forms only (all three ``(n, N_e)`` on the device)."""
    for off, size in slices:
        o = 8 * off
        check(lib().mfv2d_cuda_strided_axpby(ctx.handle, n, size, float(a), x.ptr + o, ne, float(b), y.ptr + o, ne, out.ptr + o, ne), ctx.handle)


def non_linear_solve_run_device(form_specs, lhs_codes, rhs_codes, batch: M.ElementBatch, system: SV.LinearSystem, base: DeviceArray,
                                constraint_rhs: DeviceArray, solution: DeviceArray, lagrange: DeviceArray, *, max_iterations: int,
                                atol: float, rtol: float, max_mag: float, relax: float = 1.0, solve: Callable, field_cache: dict,
                                work: dict, reduce_max: Callable[[float], float] | None = None):
    """:func:`non_linear_solve_run` with every vector resident in HBM: ``base`` is the right-hand side with the time-carry term
    already added, ``solution`` ``(n, N_e)`` / ``lagrange`` are updated in place.  The only values that cross PCIe per iteration
    are the two residual maxima and the Krylov solver's scalars.  ``work`` keeps the scratch vectors between calls."""
    ctx, n, ne = batch.ctx, batch.n, solution.shape[1]
    if not work:
        work.update(lhs=ctx.empty((n, ne)), rhs=ctx.empty((n, ne)) if rhs_codes is not None else None, nt=ctx.empty((n, ne)),
                    r_dense=ctx.empty((n, ne)), r_trace=ctx.empty(system.n_trace))

    def codes_apply(codes, out):
        key = id(codes)
        if key not in field_cache:
            fv = batch.evaluate_callable_fields(M.get_plan(ctx, batch.p, batch.q, form_specs, codes))
            field_cache[key] = (codes, ctx.to_device(fv) if fv is not None else None)
        return M.compute_element_vectors(form_specs, codes, batch, solution, field_cache[key][1], out=out)

    iter_cnt, max_residual = 0, 0.0
    lhs, nt, r_dense, r_trace = work["lhs"], work["nt"], work["r_dense"], work["r_trace"]
    while iter_cnt < max_iterations:
        codes_apply(lhs_codes, lhs)
        if rhs_codes is not None:
            SV.vec_axpby(1.0, lhs, -1.0, codes_apply(rhs_codes, work["rhs"]), lhs, ctx)
        system.apply_trace_transpose(lagrange, nt)
        SV.vec_axpby(1.0, lhs, 1.0, nt, nt, ctx)                  # main_value
        SV.vec_axpby(1.0, base, -1.0, nt, r_dense, ctx)
        system.apply_trace(solution, r_trace)
        SV.vec_axpby(1.0, constraint_rhs, -1.0, r_trace, r_trace, ctx)
        max_residual = max(vec_absmax(r_dense, ctx), vec_absmax(r_trace, ctx))
        if max_residual != max_residual:
            max_residual = float("nan")
        if reduce_max is not None:
            max_residual = float(reduce_max(max_residual))
        if not (max_residual > atol and max_residual > max_mag * rtol):
            break
        dx, dl = solve(system, r_dense, r_trace)
        SV.vec_axpby(1.0, solution, relax, dx, solution, ctx)
        SV.vec_axpby(1.0, lagrange, relax, dl, lagrange, ctx)
        iter_cnt += 1
    return solution, lagrange, iter_cnt, max_residual


def time_carry_indices(form_specs, p: int, n: int, march_forms: Sequence[int]) -> np.ndarray:
    """``find_time_carry_indices`` (python/mfv2d/solve_system.py:604-617) for ``n`` elements of order ``p``: global dense
    indices of the DoFs of the marched unknowns."""
    ne = form_specs.total_size(p, p)
    local = np.concatenate([form_specs.form_offset(int(u), p, p) + np.arange(form_specs.form_size(int(u), p, p)) for u in sorted(march_forms)])
    return (np.arange(n)[:, None] * ne + local[None, :]).ravel().astype(np.int64)


def trapezoidal_march(form_specs, lhs_codes, rhs_codes, batch: M.ElementBatch, constraints, n_trace: int, constraint_rhs: np.ndarray,
                      explicit_vec: np.ndarray, march_forms: Sequence[int], dt: float, nt: int, *, max_iterations: int = 20, atol: float = 1e-13,
                      rtol: float = 0.0, symmetric: bool = False, initial_solution: np.ndarray | None = None,
                      initial_dual: np.ndarray | None = None, field_values=None, krylov: SV.ConvergenceSettings | None = None,
                      system: SV.LinearSystem | None = None, reduce_max: Callable[[float], float] | None = None, precondition="jacobi",
                      device: bool = False, keep: Callable[[int], bool] | None = None):
    """The trapezoidal time march of ``solve_system_2d`` (python/mfv2d/solve_system_2d.py:289-397).

    ``lhs_codes`` are those of the system AFTER ``update_system_for_time_march`` (the ``2 / dt`` mass terms are part of the
    operator).  Returns the list of DoF vectors after every step (the arrays the reference hands to
    ``reconstruct_mesh_from_solution``).  ``symmetric=False`` solves the corrections with BiCGStab on the Schur complement
    (advection makes the element blocks non-symmetric), ``True`` with CG.  Multi-GPU: pass this rank's strip as ``batch``, its
    partitioned ``constraints`` / vectors, a ``system`` that is already connected to its peers (``set_interface`` +
    ``connect_peers``) and ``reduce_max``.  ``precondition``: ``"jacobi"`` or ``"multilevel"`` (the system must then carry an aggregate
    hierarchy, ``LinearSystem.set_coarse_space``).  ``device=True`` keeps the solution, the multipliers, the residuals and the
    carry terms in HBM for the whole march (:func:`non_linear_solve_run_device`); a step then downloads its solution only when
    ``keep(step)`` says so (default: every step), the other entries of the returned list are ``None``."""
    n, ne = batch.n, form_specs.total_size(batch.p, batch.p)
    if system is None:
        system = SV.compute_linear_system(form_specs, lhs_codes, batch, constraints, n_trace, field_values=field_values)
    krylov = krylov or SV.ConvergenceSettings(20000, 1e300, 1e-13)
    if symmetric:
        def solve(sysm, rd, rt):
            return SV.solve_schur_iterative(sysm, rd, rt, krylov, precondition=precondition)[:2]
    else:
        def solve(sysm, rd, rt):
            return SV.solve_schur_bicgstab(sysm, rd, rt, krylov, precondition=precondition)[:2]

    if device:
        return _trapezoidal_march_device(form_specs, lhs_codes, rhs_codes, batch, system, n_trace, constraint_rhs, explicit_vec, march_forms, dt, nt,
                                         max_iterations, atol, rtol, initial_solution, initial_dual, field_values, solve, reduce_max, keep)
    carry = time_carry_indices(form_specs, batch.p, n, march_forms)
    solution = np.zeros(n * ne) if initial_solution is None else np.array(initial_solution, np.float64).ravel()
    lagrange = np.zeros(n_trace)
    old_carry = np.zeros(carry.size) if initial_dual is None else np.asarray(initial_dual, np.float64).ravel()[carry]
    time_carry_term = np.asarray(explicit_vec, np.float64)[carry].copy()
    max_mag = float(np.max(np.abs(np.concatenate([explicit_vec, constraint_rhs]))))
    if reduce_max is not None:
        max_mag = float(reduce_max(max_mag))
    out, iters = [], []
    field_cache: dict = {}
    for _ in range(nt):
        current = 2.0 / dt * old_carry + time_carry_term
        solution, lagrange, it, _res = non_linear_solve_run(form_specs, lhs_codes, rhs_codes, batch, system, explicit_vec, constraint_rhs, solution,
                                                           lagrange, max_iterations=max_iterations, atol=atol, rtol=rtol, max_mag=max_mag,
                                                           time_carry_index=carry, time_carry_term=current, solve=solve, field_values=field_values,
                                                           reduce_max=reduce_max, field_cache=field_cache)
        projected = np.asarray(M.compute_element_dual_from_primal(form_specs, solution.reshape(n, ne), batch)).ravel()
        new_carry = projected[carry]
        time_carry_term = 2.0 / dt * (new_carry - old_carry) - time_carry_term
        old_carry = new_carry
        out.append(solution.copy())
        iters.append(it)
    return out, iters


def _trapezoidal_march_device(form_specs, lhs_codes, rhs_codes, batch, system, n_trace, constraint_rhs, explicit_vec, march_forms, dt, nt,
                              max_iterations, atol, rtol, initial_solution, initial_dual, field_values, solve, reduce_max, keep):
    """The loop of :func:`trapezoidal_march` on device vectors.  The carry terms are kept as full ``(n, N_e)`` vectors that are
    zero outside the marched forms, so that adding them to the right-hand side is one ``axpby``."""
    ctx, n, ne = batch.ctx, batch.n, form_specs.total_size(batch.p, batch.p)
    slices = _form_slices(form_specs, batch.p, march_forms)
    explicit_h = np.ascontiguousarray(explicit_vec, np.float64).reshape(n, ne)
    max_mag = float(np.max(np.abs(np.concatenate([explicit_h.ravel(), constraint_rhs]))))
    if reduce_max is not None:
        max_mag = float(reduce_max(max_mag))
    explicit = ctx.to_device(explicit_h)
    con_rhs = ctx.to_device(np.ascontiguousarray(constraint_rhs, np.float64))
    solution = ctx.to_device(np.zeros((n, ne)) if initial_solution is None else np.array(initial_solution, np.float64).reshape(n, ne))
    lagrange = ctx.empty(n_trace)
    lagrange.zero()
    old_carry, new_carry, term, base, projected = (ctx.empty((n, ne)) for _ in range(5))
    for v in (old_carry, new_carry, term):
        v.zero()
    if initial_dual is not None:
        dual = ctx.to_device(np.array(initial_dual, np.float64).reshape(n, ne))
        _masked_copy(ctx, n, ne, slices, 1.0, dual, 0.0, dual, old_carry)
    _masked_copy(ctx, n, ne, slices, 1.0, explicit, 0.0, explicit, term)
    field_cache: dict = {}
    if field_values is not None:  # the caller's values serve every code list
        fv = field_values if isinstance(field_values, DeviceArray) else ctx.to_device(field_values)
        field_cache = {id(c): (c, fv) for c in (lhs_codes, rhs_codes) if c is not None}
    mass = M._mass_expressions(form_specs, False)
    work: dict = {}
    out, iters = [], []
    for step in range(nt):
        SV.vec_axpby(2.0 / dt, old_carry, 1.0, term, base, ctx)        # current carry term ...
        SV.vec_axpby(1.0, explicit, 1.0, base, base, ctx)              # ... added to the explicit right-hand side
        _, _, it, _res = non_linear_solve_run_device(form_specs, lhs_codes, rhs_codes, batch, system, base, con_rhs, solution, lagrange,
                                                     max_iterations=max_iterations, atol=atol, rtol=rtol, max_mag=max_mag, solve=solve,
                                                     field_cache=field_cache, work=work, reduce_max=reduce_max)
        M.compute_element_vectors(form_specs, mass, batch, solution, None, out=projected)
        _masked_copy(ctx, n, ne, slices, 1.0, projected, 0.0, projected, new_carry)
        SV.vec_axpby(1.0, new_carry, -1.0, old_carry, base, ctx)       # base is free again: new - old
        SV.vec_axpby(2.0 / dt, base, -1.0, term, term, ctx)
        old_carry, new_carry = new_carry, old_carry
        out.append(solution.get().ravel() if keep is None or keep(step) else None)
        iters.append(it)
    return out, iters
