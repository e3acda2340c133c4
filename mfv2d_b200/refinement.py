"""Error estimators of the reference's ``mfv2d.refinement`` as batched device passes.

The reference loops over the leaves in Python and calls, per element, primitives that this library has as batched
kernels; the estimators are therefore written here over an :class:`ElementBatch` (all elements of one order):

=============================================  =====================================================================
reference (python/mfv2d/refinement.py)          here
=============================================  =====================================================================
``compute_legendre_coefficients``      :40-85   same name, per element (mirror of the interface)
``compute_legendre_error_estimates``   :88-152  same name, per element; ``legendre_error_estimates`` for a whole batch
                                                (one launch of ``legendre_estimate_kernel``, refine.cu)
``error_estimate_with_order_reduction`` :1095   same name over a batch: projectors (projector.cu) -> batched mat-vecs
                                                (refine.cu) -> reconstruction -> Legendre estimates, all on the device
``error_estimate_with_local_inversion`` :832    same name over a batch: higher-order batch, projected solution, residual
                                                through the matrix-free apply, local (bordered) systems solved by the
                                                batched dense solver, reconstruction, Legendre estimates
=============================================  =====================================================================

The weak-boundary-condition contribution to the residual (:902-958) walks the reference's ``Mesh`` / dual manifold; here
it is an array argument (``boundary_residual``, built by ``structured.weak_bc_contributions``).  The fine-scale Green's
operator the VMS estimator (:1387-) needs is ``mfv2d_b200.vms.SuyashGreenOperator``.
"""
from __future__ import annotations

from collections.abc import Sequence

import numpy as np

from . import _capi
from . import _mfv2d as M
from ._capi import check, lib
from .runtime import DeviceArray, addr, space_of
from .structured import _side_dofs

FORM_ORDER_0, FORM_ORDER_1, FORM_ORDER_2 = 1, 2, 3


def compute_legendre_coefficients(order_1: int, order_2: int, nodes_xi, nodes_eta, weighted_function, det) -> np.ndarray:
    """python/mfv2d/refinement.py:40-85: ``(order_2 + 1, order_1 + 1)`` Legendre coefficients of a function given as
    ``function * weight * det`` at the tensor grid ``(eta, xi)`` of integration nodes."""
    leg1 = M.compute_legendre(int(order_1), np.ascontiguousarray(nodes_xi, np.float64).ravel())
    leg2 = M.compute_legendre(int(order_2), np.ascontiguousarray(nodes_eta, np.float64).ravel())
    wf = np.asarray(weighted_function, np.float64) / np.sqrt(np.asarray(det, np.float64))
    rleg = np.einsum("ab,ib,ja->ji", wf, leg1, leg2)
    n1, n2 = np.arange(int(order_1) + 1), np.arange(int(order_2) + 1)
    return rleg / ((2 / (2 * n1 + 1))[None, :] * (2 / (2 * n2 + 1))[:, None])


def compute_legendre_error_estimates(order_1: int, order_2: int, xi, eta, w, det, u, err) -> tuple[float, float]:
    """python/mfv2d/refinement.py:88-152 for one element: ``(sum err^2 w det, |h-refinement cost|)``."""
    u, err = np.asarray(u, np.float64), np.asarray(err, np.float64)
    if err.shape != u.shape:
        raise ValueError("u and err must have the same shape")
    if err.ndim == 3:
        err, u = np.linalg.norm(err, axis=-1), np.linalg.norm(u, axis=-1)
    w, det = np.asarray(w, np.float64), np.asarray(det, np.float64)
    c_err = compute_legendre_coefficients(order_1, order_2, xi, eta, err * w * det, det)
    c_u = compute_legendre_coefficients(order_1, order_2, xi, eta, u * w * det, det)
    norm = 4 / ((2 * np.arange(order_1 + 1) + 1)[None, :] * (2 * np.arange(order_2 + 1) + 1)[:, None])
    measure = c_u * (c_u + 2 * c_err) * norm
    estimate = np.sum(measure) - np.sum(measure[: order_2 // 2, : order_1 // 2])
    return float(np.sum(err**2 * w * det)), float(abs(estimate))


def legendre_error_estimates(batch: M.ElementBatch, form_order: int, u, err, orders: tuple[int, int] | None = None, out=None):
    """``compute_legendre_error_estimates`` for every element of ``batch`` in one launch: ``u`` and ``err`` are the
    fields at the batch's quadrature points (as returned by :meth:`ElementBatch.reconstruct`), ``orders`` the orders of
    the Legendre spectrum (default: the batch's).  Returns ``(n, 2)``: error norm squared and h-refinement cost."""
    o1, o2 = (batch.p, batch.p) if orders is None else (int(orders[0]), int(orders[1]))
    out = batch._out(out, (batch.n, 2))
    check(lib().mfv2d_cuda_legendre_estimates_batch(batch.handle, o1, o2, int(form_order), addr(u), addr(err), addr(out),
                                                    space_of(u, err, out)), batch.ctx.handle)
    return out


def batched_matvec(ctx, mats, x, out=None):
    """``y[e] = mats[e] @ x[e]`` for ``(n, r, c)`` matrices and ``(n, c)`` vectors (host arrays or device arrays)."""
    n, r, c = mats.shape
    if out is None:
        out = ctx.empty((n, r)) if isinstance(mats, DeviceArray) else np.empty((n, r))
    if int(x.size) != n * c or int(out.size) != n * r:
        raise ValueError(f"batched_matvec: {n} matrices of {r} x {c} against vectors of {int(x.size)} and {int(out.size)} entries")
    check(lib().mfv2d_cuda_batched_gemv(ctx.handle, n, r, c, addr(mats), addr(x), addr(out), space_of(mats, x, out)), ctx.handle)
    return out


def _form_slice(form_specs, form_index: int, p: int) -> slice:
    off = form_specs.form_offset(form_index, p, p)
    return slice(off, off + form_specs.form_size(form_index, p, p))


def error_estimate_with_order_reduction(solution, batch: M.ElementBatch, reduction_order: int, form_specs, form_index: int,
                                        alternative: str = "ignore") -> tuple[np.ndarray, np.ndarray]:
    """python/mfv2d/refinement.py:1095-1203 for all elements of ``batch``: the target form is projected to the basis of
    order ``p - reduction_order`` and back; the difference is the error field.  ``solution`` is ``(n, N_e)`` (or flat).
    Returns ``(element_error, href_cost)``."""
    if alternative not in ("ignore", "prioritize"):
        raise ValueError(f"Invalid alternative strategy specified {alternative}")
    n, p, q = batch.n, batch.p, batch.q
    if p <= reduction_order:
        fill = 0.0 if alternative == "ignore" else np.inf
        return np.full(n, fill), np.full(n, fill)
    ctx = batch.ctx
    order = int(form_specs[form_index][1])
    sol = np.ascontiguousarray(np.asarray(solution, np.float64).reshape(n, -1)[:, _form_slice(form_specs, form_index, p)])
    d_dofs = ctx.to_device(sol)
    d_corners = batch._corners
    rule = M.IntegrationRule1D(q - 1)
    p_lo = p - int(reduction_order)
    lower = M.ElementBatch(M.Basis2D(M.Basis1D(p_lo, rule), M.Basis1D(p_lo, rule)), d_corners, ctx)
    forward = M.compute_element_projectors(order, batch.basis_2d, lower, out=ctx.empty((n, lower.form_size(order), batch.form_size(order))))
    backward = M.compute_element_projectors(order, lower.basis_2d, batch, out=ctx.empty((n, batch.form_size(order), lower.form_size(order))))
    low = batched_matvec(ctx, forward, d_dofs)
    back = batched_matvec(ctx, backward, low)
    d_err = ctx.empty(sol.shape)
    check(lib().mfv2d_cuda_vec_axpby(ctx.handle, sol.size, 1.0, addr(d_dofs), -1.0, addr(back), addr(d_err), _capi.DEVICE), ctx.handle)
    shape = (n, q, q, 2) if order == FORM_ORDER_1 else (n, q, q)
    rec_err = batch.reconstruct(order, d_err, out=ctx.empty(shape))
    rec_u = batch.reconstruct(order, d_dofs, out=ctx.empty(shape))
    est = legendre_error_estimates(batch, order, rec_u, rec_err, out=ctx.empty((n, 2))).get()
    return est[:, 0].copy(), np.abs(est[:, 1])


def error_estimate_with_local_inversion(solution, batch: M.ElementBatch, recon_order: int, form_specs, lhs_codes, rhs_codes, system,
                                        form_index: int, zeroed_forms: Sequence[int] = (), constrained_forms: Sequence[int] = (),
                                        boundary_residual=None):
    """python/mfv2d/refinement.py:832-1093 for all elements of ``batch`` (without the weak-boundary-condition part of the
    residual, :902-958): the solution is projected to order ``p + recon_order``, the residual of the system is formed
    there with the matrix-free apply, the local system - bordered with Lagrange rows that zero the boundary DoFs of
    ``zeroed_forms`` and pin the sums of ``constrained_forms`` (form indices) - is solved for every element, and the
    target form's part of the local solution is the error field.

    ``system`` provides the right-hand-side functions as in :func:`mfv2d_b200._mfv2d.compute_element_rhs`;
    ``lhs_codes`` / ``rhs_codes`` are the bytecode of ``CompiledSystem`` (``rhs_codes`` may be ``None``).
    ``boundary_residual`` ``(n, N_fine)``: the natural-boundary-condition terms of the right-hand side on the space of order
    ``p + recon_order`` (``structured.weak_bc_contributions`` with the fine order), added to the residual where the
    reference intends to (:902-958; its own loop runs over a boolean mask instead of line indices and takes the side DoFs
    with the coarse orders, so it cannot serve as a fixture - the argument is the array-side replacement)."""
    n, p, q = batch.n, batch.p, batch.q
    ctx = batch.ctx
    ph = p + int(recon_order)
    rule = M.IntegrationRule1D(q - 1)
    fine = M.ElementBatch(M.Basis2D(M.Basis1D(ph, rule), M.Basis1D(ph, rule)), batch._corners, ctx)
    ne_c, ne_f = form_specs.total_size(p, p), form_specs.total_size(ph, ph)
    sol = np.ascontiguousarray(np.asarray(solution, np.float64).reshape(n, ne_c))
    # fine_solution = blockdiag(projectors) @ element_solution (:878-886)
    d_fine = ctx.empty((n, ne_f))
    fine_host = np.empty((n, ne_f))
    done: dict[int, DeviceArray] = {}
    for fi in range(len(form_specs)):
        order = int(form_specs[fi][1])
        if order not in done:
            done[order] = M.compute_element_projectors(order, batch.basis_2d, fine, out=ctx.empty((n, fine.form_size(order), batch.form_size(order))))
        d_part = ctx.to_device(np.ascontiguousarray(sol[:, _form_slice(form_specs, fi, p)]))
        part = batched_matvec(ctx, done[order], d_part)
        fine_host[:, _form_slice(form_specs, fi, ph)] = part.get()
    d_fine.set(fine_host)
    # residual = fine_rhs - (lhs - rhs)(fine_solution) (:876-899)
    residual = np.asarray(M.compute_element_rhs(system, fine), np.float64).copy()
    forcing = M.compute_element_vectors(form_specs, lhs_codes, fine, d_fine, out=ctx.empty((n, ne_f))).get()
    residual -= forcing
    if rhs_codes is not None and any(any(c for c in row) for row in rhs_codes):
        residual += M.compute_element_vectors(form_specs, rhs_codes, fine, d_fine, out=ctx.empty((n, ne_f))).get()
    if boundary_residual is not None:
        residual += np.asarray(boundary_residual, np.float64).reshape(n, ne_f)
    # local systems (:989-1053)
    d_lhs = M.compute_element_matrices(form_specs, lhs_codes, fine, d_fine, out=ctx.empty((n, ne_f, ne_f)))
    target_order = int(form_specs[form_index][1])
    lrow, lcol = [], []
    pad = 0
    if zeroed_forms:
        if target_order == FORM_ORDER_2:
            raise ValueError("2-forms have no boundary DoFs.")
        cols = []
        for fi in zeroed_forms:
            for side in (1, 2, 3, 4):
                # the reference takes the side DoFs with the TARGET form's order (:1003-1007)
                cols.append(form_specs.form_offset(fi, ph, ph) + _side_dofs(side, target_order, ph))
        cols = np.unique(np.concatenate(cols))
        lrow.append(np.arange(cols.size)), lcol.append(cols)
        pad = int(cols.size)
    for fi in constrained_forms:
        cols = form_specs.form_offset(fi, ph, ph) + np.arange(form_specs.form_size(fi, ph, ph))
        lrow.append(np.full(cols.size, pad)), lcol.append(cols)
        pad += 1
    if pad:
        lr, lc = np.concatenate(lrow).astype(np.int32), np.concatenate(lcol).astype(np.int32)
        lv = np.ones(lr.size)
        d_mat = ctx.empty((n, ne_f + pad, ne_f + pad))
        check(lib().mfv2d_cuda_border_batch(ctx.handle, n, ne_f, pad, addr(d_lhs), lr.size, lr.ctypes.data, lc.ctypes.data, lv.ctypes.data,
                                            addr(d_mat), _capi.DEVICE), ctx.handle)
        residual = np.pad(residual, ((0, 0), (0, pad)))
    else:
        d_mat = d_lhs
    size = ne_f + pad
    d_x = ctx.empty((n, size))
    d_status = ctx.empty((n,), np.int32)
    d_res = ctx.to_device(np.ascontiguousarray(residual))
    check(lib().mfv2d_cuda_solve_batch(ctx.handle, size, 1, n, addr(d_mat), addr(d_res), addr(d_x), addr(d_status), _capi.DEVICE), ctx.handle)
    status = d_status.get()
    if np.any(status):
        bad = int(np.flatnonzero(status)[0])
        raise _capi.Mfv2dCudaError(int(status[bad]), _capi.result_name(int(status[bad])), f"element {bad}: the local system is singular")
    local = d_x.get()
    sl = _form_slice(form_specs, form_index, ph)
    d_err = ctx.to_device(np.ascontiguousarray(local[:, sl]))
    d_u = ctx.to_device(np.ascontiguousarray(fine_host[:, sl]))
    shape = (n, q, q, 2) if target_order == FORM_ORDER_1 else (n, q, q)
    rec_err = fine.reconstruct(target_order, d_err, out=ctx.empty(shape))
    rec_u = fine.reconstruct(target_order, d_u, out=ctx.empty(shape))
    est = legendre_error_estimates(fine, target_order, rec_u, rec_err, orders=(p, p), out=ctx.empty((n, 2))).get()
    return est[:, 0].copy(), np.abs(est[:, 1])
