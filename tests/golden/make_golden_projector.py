"""Golden fixture of compute_element_projector, generated from the REFERENCE's own C extension
(oracle/_ref/_mfv2d_ref.so, built by oracle/build_ref.sh).  Build container only:

    python tests/golden/make_golden_projector.py        ->  tests/golden/projector.npz
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(HERE))

import fields as F  # noqa: E402
from oracle.ref_loader import load_ref  # noqa: E402

CASES = [(3, 2, 6), (2, 4, 6), (4, 4, 7), (5, 3, 8)]  # (p_out, p_in, rule order)


def main():
    ref = load_ref()
    assert ref is not None, "build oracle/_ref first"
    specs = ref._ElementFormSpecification(("a", 1), ("b", 2), ("c", 3))
    data = {"cases": np.array(CASES)}
    for po, pi, ro in CASES:
        rule = ref.IntegrationRule1D(ro)
        bo1, bi1 = ref.Basis1D(po, rule), ref.Basis1D(pi, rule)
        bo, bi = ref.Basis2D(bo1, bo1), ref.Basis2D(bi1, bi1)
        for iq in (4, 5):
            sp = ref.ElementFemSpace2D(bo, np.ascontiguousarray(F.QUADS[iq]))
            for dual in (False, True):
                res = ref.compute_element_projector(specs, bi, sp, dual)
                for fo, m in zip((1, 2, 3), res):
                    data[f"po{po}_pi{pi}_r{ro}_q{iq}_d{int(dual)}_f{fo}"] = np.array(m)
    np.savez_compressed(os.path.join(HERE, "projector.npz"), **data)
    print("wrote projector.npz with", len(data), "arrays")


if __name__ == "__main__":
    main()
