"""Dual degrees of freedom of test functions from the REFERENCE's element_dual_dofs
(python/mfv2d/mimetic2d.py:1003-1123).  Build container only:

    python tests/golden/make_golden_dual.py        ->  tests/golden/dual_dofs.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

import fields as F  # noqa: E402
from oracle.ref_loader import load_ref_python  # noqa: E402


def f_scalar(x, y):
    return np.sin(1.3 * x + 0.2) * np.cos(0.7 * y) + 0.5 * x * y


def f_vector(x, y):
    return np.stack((np.cos(x) * y + 1.0, np.sin(0.5 * y) - 0.25 * x * x), axis=-1)


def main():
    assert load_ref_python() is not None
    from mfv2d._mfv2d import Basis1D, Basis2D, ElementFemSpace2D, IntegrationRule1D
    from mfv2d.kform import UnknownFormOrder
    from mfv2d.mimetic2d import element_dual_dofs

    out = {}
    for p, k in ((1, 2), (3, 3), (4, 2), (6, 3)):
        b1 = Basis1D(p, IntegrationRule1D(p + k))
        basis = Basis2D(b1, b1)
        for iq in (2, 4, 5):
            sp = ElementFemSpace2D(basis, np.ascontiguousarray(F.QUADS[iq], np.float64))
            for fo, fn in ((1, f_scalar), (2, f_vector), (3, f_scalar)):
                out[f"p{p}k{k}q{iq}_f{fo}"] = element_dual_dofs(UnknownFormOrder(fo), sp, fn)
    np.savez_compressed(os.path.join(HERE, "dual_dofs.npz"), **out)
    print(len(out), "vectors")


if __name__ == "__main__":
    main()
