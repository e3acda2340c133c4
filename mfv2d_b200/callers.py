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
                         reduce_max: Callable[[float], float] | None = None, field_cache: dict | None = None):
    """``non_linear_solve_run`` (python/mfv2d/solve_system.py:353-485) on one bucket of elements.

    ``explicit_vec`` ``(n * N_e,)`` is the dense right-hand side (weak boundary terms included), ``constraint_rhs`` the one of
    the constraint rows; ``solution`` / ``lagrange`` are updated in place and returned with the iteration count and the last
    maximum residual.  ``system`` holds the element matrices of ``lhs_codes`` (the linearised operator of the correction
    solve, ``system_decomp`` of the reference); ``solve(system, r_dense, r_trace) -> (dx, dlambda)`` defaults to the
    Schur-complement CG, pass ``solving.solve_schur_bicgstab`` based solvers for non-symmetric blocks.  On a partitioned
    system (one strip of the mesh per rank, ``system`` connected to its peers) every rank passes its local vectors, the
    GLOBAL ``max_mag`` and ``reduce_max`` = the host's max-all-reduce: the stopping decision is then the same on all ranks."""
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
    while iter_cnt < max_iterations:
        lhs = _apply_codes(form_specs, lhs_codes, batch, solution.reshape(n, ne), field_values, field_cache)
        if rhs_codes is not None:
            lhs = lhs - _apply_codes(form_specs, rhs_codes, batch, solution.reshape(n, ne), field_values, field_cache)
        main_value = np.asarray(lhs).ravel() + system.apply_trace_transpose(lagrange)
        r_dense = base - main_value
        r_trace = constraint_rhs - system.apply_trace(solution)
        max_residual = float(max(np.max(np.abs(r_dense), initial=0.0), np.max(np.abs(r_trace), initial=0.0)))
        if reduce_max is not None:
            max_residual = float(reduce_max(max_residual))
        if not (max_residual > atol and max_residual > max_mag * rtol):
            break
        dx, dl = solve(system, r_dense, r_trace)
        solution += relax * dx
        lagrange += relax * dl
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
                      system: SV.LinearSystem | None = None, reduce_max: Callable[[float], float] | None = None, precondition="jacobi"):
    """The trapezoidal time march of ``solve_system_2d`` (python/mfv2d/solve_system_2d.py:289-397).

    ``lhs_codes`` are those of the system AFTER ``update_system_for_time_march`` (the ``2 / dt`` mass terms are part of the
    operator).  Returns the list of DoF vectors after every step (the arrays the reference hands to
    ``reconstruct_mesh_from_solution``).  ``symmetric=False`` solves the corrections with BiCGStab on the Schur complement
    (advection makes the element blocks non-symmetric), ``True`` with CG.  Multi-GPU: pass this rank's strip as ``batch``, its
    partitioned ``constraints`` / vectors, a ``system`` that is already connected to its peers (``set_interface`` +
    ``connect_peers``) and ``reduce_max``.  ``precondition``: ``"jacobi"`` or ``"multilevel"`` (the system must then carry an aggregate
    hierarchy, ``LinearSystem.set_coarse_space``)."""
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
