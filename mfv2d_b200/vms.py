"""The fine-scale Green's operator of the variational multiscale closure, batched and device resident.

Reference: ``SuyashGreenOperator`` (python/mfv2d/solve_system.py:653-961).  There, per element and in Python: the fine
space of order ``p + order_increase``, the two projectors, four element matrices, a triple product ``P^T A_f P``; then two
SuperLU factorisations (fine and coarse symmetric saddle systems) and a fixed-point loop over SciPy sparse products.

Here every per-element object is ONE batched launch on an :class:`ElementBatch` and stays in HBM (projectors, advection
blocks, the condensed fine / coarse systems); the two factorisations are replaced by the Schur-complement Krylov solver of
``solving``; the fixed-point loop works on device vectors and brings two maxima per iteration to the host.  Same formulas,
same stopping rule, same method names.  Like the other callers (``callers.py``) the operator takes arrays instead of
``Mesh`` / ``KFormSystem`` objects: the bytecode of the systems (``CompiledSystem(...).lhs_codes`` etc.), the fine forcing
``(n, N_f)`` (``compute_element_rhs(system, operator_fine_batch)``) and the constraint rows of the fine and the coarse
mesh as COO triples (``structured`` / ``constraints`` / ``nonconforming`` build them).
"""
from __future__ import annotations

import numpy as np

from . import _mfv2d as M
from . import solving as SV
from ._capi import check, lib
from .callers import vec_absmax
from .refinement import batched_matvec
from .runtime import DeviceArray, addr, space_of


def fine_batch_of(coarse: M.ElementBatch, order_increase: int) -> M.ElementBatch:
    """The elements of ``coarse`` with the basis of order ``p + order_increase`` on the SAME integration rule
    (solve_system.py:701-708: ``get_basis2d(order + increase, *fem_space.integration_orders)``)."""
    rule = coarse.basis_2d.basis_xi.rule
    b1 = M.Basis1D(coarse.p + int(order_increase), rule)
    return M.ElementBatch(M.Basis2D(b1, b1), coarse._corners, coarse.ctx)


def batched_matmul(ctx, a, b, trans_a: bool = False, out=None):
    """``c[e] = op(a[e]) @ b[e]`` for stacks of row-major matrices (numpy or device arrays)."""
    n, k, l = b.shape
    m = a.shape[2] if trans_a else a.shape[1]
    assert a.shape == ((n, k, m) if trans_a else (n, m, k)), (a.shape, b.shape, trans_a)
    if out is None:
        out = ctx.empty((n, m, l)) if isinstance(a, DeviceArray) else np.empty((n, m, l))
    check(lib().mfv2d_cuda_batched_gemm(ctx.handle, n, m, k, l, int(bool(trans_a)), addr(a), addr(b), addr(out), space_of(a, b, out)), ctx.handle)
    return out


def _block_diagonal_projector(form_specs, basis_in: M.Basis2D, batch_out: M.ElementBatch, p_in: int) -> np.ndarray:
    """``sp.block_diag(compute_element_projector(forms, basis_in, space_out))`` for every element, dense ``(n, N_out, N_in)``."""
    p_out = batch_out.p
    n_out, n_in = form_specs.total_size(p_out, p_out), form_specs.total_size(p_in, p_in)
    full = np.zeros((batch_out.n, n_out, n_in))
    done: dict[int, np.ndarray] = {}
    for i, order in enumerate(form_specs.orders):
        order = int(order)
        if order not in done:
            done[order] = M.compute_element_projectors(order, basis_in, batch_out)
        ro, co = form_specs.form_offset(i, p_out, p_out), form_specs.form_offset(i, p_in, p_in)
        blk = done[order]
        full[:, ro:ro + blk.shape[1], co:co + blk.shape[2]] = blk
    return full


class SuyashGreenOperator:
    """``SuyashGreenOperator`` (python/mfv2d/solve_system.py:653-961) on one bucket of elements.

    ``symmetric_codes`` / ``advection_linear_codes`` / ``advection_nonlin_codes``: bytecode of
    ``CompiledSystem(settings.symmetric_system).lhs_codes`` and ``CompiledSystem(settings.nonsymmetric_system).linear_codes`` /
    ``.nonlin_codes`` (``None`` when the advection is linear).  ``fine_constraints`` / ``coarse_constraints``:
    ``(row, col, val, n_rows)`` of ``add_system_constraints`` on the p-enriched and the original mesh.  ``convergence`` is
    ``settings.fine_scale_convergence``; ``krylov`` bounds the two inner saddle-point solves (the reference factorises).
    ``solver``: ``"cg"`` (symmetric system, the default) or ``"bicgstab"``."""

    def __init__(self, form_specs, symmetric_codes, advection_linear_codes, advection_nonlin_codes, coarse_batch: M.ElementBatch,
                 order_increase: int, fine_forcing, fine_constraints, coarse_constraints, convergence: SV.ConvergenceSettings,
                 relaxation: float = 1.0, krylov: SV.ConvergenceSettings | None = None, solver: str = "cg", field_values=None):
        ctx = self.ctx = coarse_batch.ctx
        self.unknown_forms = form_specs
        self.convergence, self.relaxation = convergence, float(relaxation)
        self.krylov = krylov or SV.ConvergenceSettings(20000, 1e300, 1e-13)
        self.solver = {"cg": SV.solve_schur_iterative, "bicgstab": SV.solve_schur_bicgstab}[solver]
        self.coarse_batch = coarse_batch
        self.fine_batch = fine = fine_batch_of(coarse_batch, order_increase)
        n, pc, pf = coarse_batch.n, coarse_batch.p, fine.p
        self.n, self.n_coarse, self.n_fine = n, form_specs.total_size(pc, pc), form_specs.total_size(pf, pf)
        self._nonlin_codes, self._field_values = advection_nonlin_codes, field_values
        # projectors, dense per element (block diagonal over the forms): coarse -> fine, its transpose, fine -> coarse
        c2f = _block_diagonal_projector(form_specs, coarse_batch.basis_2d, fine, pc)
        f2c = _block_diagonal_projector(form_specs, fine.basis_2d, coarse_batch, pf)
        self.projector_c2f = ctx.to_device(c2f)
        self.projector_c2f_t = ctx.to_device(np.ascontiguousarray(c2f.transpose(0, 2, 1)))
        self.projector_f2c = ctx.to_device(f2c)
        # advection blocks on the fine space (the coarse ones of the reference, :741-744, are never used by its methods)
        self.fine_linear_advection = ctx.empty((n, self.n_fine, self.n_fine))
        M.compute_element_matrices(form_specs, advection_linear_codes, fine, None, field_values, out=self.fine_linear_advection)
        self.fine_advection = self.fine_linear_advection
        # the symmetric systems: fine blocks from the bytecode, coarse blocks = P^T A_f P (:746-751)
        fine_blocks = ctx.empty((n, self.n_fine, self.n_fine))
        M.compute_element_matrices(form_specs, symmetric_codes, fine, None, field_values, out=fine_blocks)
        tmp = batched_matmul(ctx, fine_blocks, self.projector_c2f)
        coarse_blocks = batched_matmul(ctx, self.projector_c2f, tmp, trans_a=True)
        tmp.free()
        fr, fc, fv, self.fine_padding = fine_constraints
        cr, cc, cv, self.coarse_padding = coarse_constraints
        self.fine_system = SV.LinearSystem(fine_blocks, (np.asarray(fr), np.asarray(fc), np.asarray(fv, np.float64)), int(self.fine_padding), ctx=ctx)
        self.coarse_system = SV.LinearSystem(coarse_blocks, (np.asarray(cr), np.asarray(cc), np.asarray(cv, np.float64)), int(self.coarse_padding),
                                             ctx=ctx)
        self.fine_forcing = ctx.to_device(np.zeros((n, self.n_fine)) if fine_forcing is None
                                          else np.ascontiguousarray(fine_forcing, np.float64).reshape(n, self.n_fine))
        self._zero_fine_trace = ctx.empty(int(self.fine_padding))
        self._zero_fine_trace.zero()
        self._zero_coarse_trace = ctx.empty(int(self.coarse_padding))
        self._zero_coarse_trace.zero()
        self.krylov_iterations = 0

    # ---- device helpers -------------------------------------------------------------------------------------------
    def _dev(self, x, width):
        if isinstance(x, DeviceArray):
            return x
        return self.ctx.to_device(np.ascontiguousarray(x, np.float64).reshape(self.n, width))

    def _advect(self, v, out=None):
        return batched_matvec(self.ctx, self.fine_advection, v, out)

    def _green(self, x):
        """fine_scale_greens_function on device vectors (:947-961)."""
        fine, _, _, it1 = self.solver(self.fine_system, x, self._zero_fine_trace, self.krylov)
        xc = batched_matvec(self.ctx, self.projector_c2f_t, x)              # x @ projector_c2f
        coarse, _, _, it2 = self.solver(self.coarse_system, xc, self._zero_coarse_trace, self.krylov)
        self.krylov_iterations += int(it1) + int(it2)
        back = batched_matvec(self.ctx, self.projector_c2f, coarse)
        return SV.vec_axpby(1.0, fine, -1.0, back, fine, self.ctx)

    # ---- the reference's methods ----------------------------------------------------------------------------------
    def fine_scale_greens_function(self, x) -> np.ndarray:
        """Apply the fine-scale Green's function to a fine dual vector (:947-961)."""
        return self._green(self._dev(x, self.n_fine)).get().ravel()

    def update_nonlinear_advection(self, coarse_dofs) -> None:
        """:922-945: advection blocks = linear part + non-linear part evaluated at the projected coarse solution
        (``coarse_dofs`` carries the multipliers at its end, like the reference's)."""
        if self._nonlin_codes is None:
            return
        c = np.asarray(coarse_dofs, np.float64).ravel()[: self.n * self.n_coarse]
        fine_dofs = batched_matvec(self.ctx, self.projector_c2f, self._dev(c, self.n_coarse))
        if self.fine_advection is self.fine_linear_advection:
            self.fine_advection = self.ctx.empty((self.n, self.n_fine, self.n_fine))
        M.compute_element_matrices(self.unknown_forms, self._nonlin_codes, self.fine_batch, fine_dofs, self._field_values, out=self.fine_advection)
        SV.vec_axpby(1.0, self.fine_advection, 1.0, self.fine_linear_advection, self.fine_advection, self.ctx)

    def compute_unresolved_contributions(self, coarse_solution, initial_guess=None) -> np.ndarray:
        """:858-890: fixed point ``u <- A G(residual) - A G(u)`` for the unresolved forcing."""
        ctx = self.ctx
        proj = batched_matvec(ctx, self.projector_c2f, self._dev(coarse_solution, self.n_coarse))
        residual = SV.vec_axpby(1.0, self.fine_forcing, -1.0, self._advect(proj), None, ctx)
        agr = self._advect(self._green(residual))
        u = ctx.empty((self.n, self.n_fine))
        SV.vec_axpby(1.0, agr if initial_guess is None else self._dev(initial_guess, self.n_fine), 0.0, agr, u, ctx)
        diff = ctx.empty((self.n, self.n_fine))
        for _ in range(self.convergence.maximum_iterations):
            u_new = SV.vec_axpby(1.0, agr, -1.0, self._advect(self._green(u)), None, ctx)
            max_du = vec_absmax(SV.vec_axpby(1.0, u, -1.0, u_new, diff, ctx), ctx)
            max_u = vec_absmax(u_new, ctx)
            if self.relaxation == 1.0:
                u = u_new
            else:
                SV.vec_axpby(1.0 - self.relaxation, u, self.relaxation, u_new, u, ctx)
            if max_u == 0 or max_du < max_u * self.convergence.relative_tolerance or max_du < self.convergence.absolute_tolerance:
                break
        return u.get().ravel()

    def recover_unresolved(self, coarse_solution, unresolved_contribution) -> np.ndarray:
        """:892-907 (``coarse_solution`` with the multipliers at its end)."""
        ctx = self.ctx
        c = np.asarray(coarse_solution, np.float64).ravel()
        c = c[: c.size - int(self.coarse_padding)]
        proj = batched_matvec(ctx, self.projector_c2f, self._dev(c, self.n_coarse))
        residual = SV.vec_axpby(1.0, self.fine_forcing, -1.0, self._advect(proj), None, ctx)
        SV.vec_axpby(1.0, residual, -1.0, self._dev(unresolved_contribution, self.n_fine), residual, ctx)
        return self._green(residual).get().ravel()

    def fine_results_to_coarse_dofs(self, x, *, dual: bool) -> np.ndarray:
        """:909-920: ``x @ P_c2f`` for dual vectors, ``P_f2c @ x`` for primal ones, padded for the coarse multipliers."""
        xd = self._dev(x, self.n_fine)
        y = batched_matvec(self.ctx, self.projector_c2f_t if dual else self.projector_f2c, xd)
        return np.pad(y.get().ravel(), (0, int(self.coarse_padding)))
