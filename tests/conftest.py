import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN = os.path.join(ROOT, "tests", "golden")
os.environ.setdefault("MFV2D_TEST_HOOKS", "1")  # lets _capi._use_library_for_testing load the host-emulation build


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_systems():
    import fields as F

    with open(os.path.join(GOLDEN, "systems.json")) as f:
        raw = json.load(f)
    with open(os.path.join(GOLDEN, "systems_examples.json")) as f:  # the remaining examples/unsteady systems (make_golden_examples.py)
        raw.update(json.load(f))

    def dec(codes):
        if codes is None:
            return None
        out = []
        for row in codes:
            r = []
            for blk in row:
                if blk is None:
                    r.append(None)
                    continue
                ops = []
                for item in blk:
                    item = list(item)
                    if item[0] == 7 and isinstance(item[2], str) and item[2].startswith("fn:"):
                        item[2] = F.FIELDS[item[2][3:]]
                    ops.append(tuple(item))
                r.append(tuple(ops))
            out.append(tuple(r))
        return tuple(out)

    return {
        k: dict(form_names=v["form_names"], form_orders=v["form_orders"], **{c: dec(v[c]) for c in ("lhs", "rhs", "linear", "nonlin")})
        for k, v in raw.items()
    }


@pytest.fixture(scope="session")
def systems():
    return load_systems()


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def relerr(a, b):
    a = np.asarray(a)
    b = np.asarray(b)
    s = np.max(np.abs(b))
    return float(np.max(np.abs(a - b)) / (s if s > 0 else 1.0))
