"""Fixture for the batched error estimators (mfv2d_b200/refinement.py): the reference's own estimators.

Run in the build container (needs /root/reference):  python tests/golden/make_golden_refinement.py
Writes tests/golden/refinement.npz: on the deformed 3 x 2 mesh of order 4 (FemCache(2): rule order 6) with a fixed
pseudo-random solution vector of the mixed Poisson system (q 1-form, u 2-form, forcing f), the outputs of
  * compute_legendre_coefficients / compute_legendre_error_estimates   (python/mfv2d/refinement.py:40-152) on element 0
  * error_estimate_with_order_reduction                                (:1095-1203) for both unknowns, reduction 1 and 2
  * error_estimate_with_local_inversion                                (:832-1093) with recon_order 1, plain / with the
    2-form constrained / with the 1-form strongly zeroed on the element boundary and the 2-form constrained (the weak-BC loop of :902-958 needs a
    Mesh and is not exercised: the system has no boundary projection with a function)
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.ref_loader import load_ref_python  # noqa: E402

mfv2d = load_ref_python()
import fields as F  # noqa: E402
from mfv2d._mfv2d import ElementFemSpace2D  # noqa: E402
from mfv2d.eval import CompiledSystem  # noqa: E402
from mfv2d.examples import unit_square_mesh  # noqa: E402
from mfv2d.kform import KFormUnknown, UnknownFormOrder  # noqa: E402
from mfv2d.mimetic2d import FemCache, jacobian, reconstruct  # noqa: E402
from mfv2d.refinement import (  # noqa: E402
    compute_legendre_coefficients,
    compute_legendre_error_estimates,
    error_estimate_with_local_inversion,
    error_estimate_with_order_reduction,
)
from mfv2d.system import KFormSystem  # noqa: E402

O1, O2 = UnknownFormOrder.FORM_ORDER_1, UnknownFormOrder.FORM_ORDER_2


def forcing(x, y):
    return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)


def main():
    nh, nv, p = 3, 2, 4
    u, q = KFormUnknown("u", O2), KFormUnknown("q", O1)
    v, w = u.weight, q.weight
    system = KFormSystem(w @ q + w.derivative @ u == 0, v @ q.derivative == v @ forcing, sorting=lambda f: f.order)
    comp = CompiledSystem(system)
    specs = system.unknown_forms
    mesh = unit_square_mesh(nh, nv, p, F.deform)
    cache = FemCache(2)
    leaf = list(mesh.get_leaf_indices())
    corners = np.array([np.array(mesh.get_leaf_corners(i), np.float64) for i in leaf])
    spaces = [ElementFemSpace2D(cache.get_basis2d(p, p), corners[i]) for i in range(len(leaf))]
    ne = specs.total_size(p, p)
    offsets = (np.arange(len(leaf) + 1) * ne).astype(np.uint32)
    rng = np.random.default_rng(99)
    # a smooth part plus noise, so that both the low and the high Legendre modes carry weight
    solution = np.concatenate([np.linspace(-1, 1, ne) ** 2 + 0.1 * rng.standard_normal(ne) for _ in leaf])
    data = dict(corners=corners, order=p, k=2, solution=solution, form_orders=np.array([int(o) for o in specs.orders]))

    # element 0: the two Legendre helpers on the reconstructed 2-form and a synthetic error field
    sp0 = spaces[0]
    r1, r2 = sp0.basis_xi.rule, sp0.basis_eta.rule
    xi, eta = np.asarray(r1.nodes, np.float64), np.asarray(r2.nodes, np.float64)
    jac = jacobian(corners[0], xi[None, :], eta[:, None])
    det = jac[0][0] * jac[1][1] - jac[0][1] * jac[1][0]
    wts = np.asarray(r1.weights)[None, :] * np.asarray(r2.weights)[:, None]
    off_u = specs.form_offset(specs.index(u), p, p)
    dofs_u = solution[off_u : off_u + specs.form_size(specs.index(u), p, p)]
    vals = reconstruct(sp0, O2, dofs_u, xi[None, :], eta[:, None])
    errv = 0.05 * np.cos(3 * xi[None, :] + eta[:, None])
    data["leg_coeffs"] = compute_legendre_coefficients(p, p, xi, eta, vals * wts * det, det)
    data["leg_coeffs_32"] = compute_legendre_coefficients(3, 2, xi, eta, vals * wts * det, det)
    data["leg_estimates"] = np.array(compute_legendre_error_estimates(p, p, xi, eta, wts, det, vals, errv))
    data["leg_u"], data["leg_err"], data["leg_det"], data["leg_w"] = vals, errv, det, wts

    for target, name in ((u, "u"), (q, "q")):
        for red in (1, 2):
            e, h = error_estimate_with_order_reduction(solution, offsets, spaces, red, cache, specs, target, "ignore")
            data[f"reduction_{name}_{red}"] = np.stack([e, h], axis=1)
    e, h = error_estimate_with_order_reduction(solution, offsets, spaces, 4, cache, specs, u, "prioritize")
    data["reduction_u_4_prioritize"] = np.stack([e, h], axis=1)

    stub = types.SimpleNamespace(leaf_count=len(leaf), boundary_indices=np.zeros(0, np.uint32))
    # (zeroing q alone leaves u free up to a constant - a singular local system; the well-posed case pins u as well)
    for tag, zeroed, constrained in (("plain", (), ()), ("constrained_u", (), (u,)), ("zeroed_q_constrained_u", (q,), (u,))):
        for target, name in ((u, "u"), (q, "q")):
            if zeroed and target is u:
                continue  # the reference takes the boundary DoFs with the TARGET's order (:1003-1007): a 2-form has none and it raises
            e, h = error_estimate_with_local_inversion(stub, solution, offsets, (), spaces, 1, cache, system, comp, target, zeroed, constrained)
            data[f"local_{tag}_{name}"] = np.stack([e, h], axis=1)
    np.savez_compressed(os.path.join(HERE, "refinement.npz"), **data)
    for k_, v_ in data.items():
        print(k_, np.asarray(v_).shape)


if __name__ == "__main__":
    main()
