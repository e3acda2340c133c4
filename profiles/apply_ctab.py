"""1-D tables of the fused apply: launch parameters read as uniform constant operands (MFV2D_APPLY_CTAB=1, default) against shared memory (=0)."""
import json
import os
import sys

import numpy as np

sys.path.insert(0, "/root/repo")
sys.path.insert(0, "/root/repo/tests")
import bench as B
from mfv2d_b200 import _mfv2d as M, structured as ST
import mfv2d_b200.runtime as rt
from mfv2d_b200.runtime import Context

ctx = Context(0)
rt._context = ctx
res = {}
for name, p, n_side in (("mixed_poisson", 6, 354), ("mixed_poisson", 8, 256), ("stokes", 6, 256), ("direct_poisson", 5, 256), ("mixed_poisson", 5, 256)):
    names, orders, lhs = B.load_system(name)
    specs = M._ElementFormSpecification(*zip(names, orders))
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + 3))
    basis = M.Basis2D(b1, b1)
    corners = ST.unit_square_corners(n_side, n_side, B.deform)
    n = corners.shape[0]
    batch = M.ElementBatch(basis, ctx.to_device(corners), ctx)
    ne = specs.total_size(p, p)
    x = ctx.to_device(np.random.default_rng(0).standard_normal((n, ne)))
    y = ctx.empty((n, ne))
    out, ref = {}, None
    for mode in ("0", "1"):
        os.environ["MFV2D_APPLY_CTAB"] = mode
        for _ in range(3):
            M.compute_element_vectors(specs, lhs, batch, x, out=y)
        ctx.synchronize()
        ctx.timer_start()
        for _ in range(10):
            M.compute_element_vectors(specs, lhs, batch, x, out=y)
        t = ctx.timer_stop() / 10
        yh = y.get()
        ref = yh if ref is None else ref
        out[mode] = round(n * ne / (t * 1e-3) / 1e9, 2)
        assert np.array_equal(yh, ref), (name, p, mode)
    res[f"{name}_p{p}_N{ne}"] = {"shared_tables": out["0"], "param_tables": out["1"]}
    print(name, p, ne, res[f"{name}_p{p}_N{ne}"], flush=True)
json.dump(res, open("/root/repo/gpurun_out/apply_ctab_r02.json", "w"), indent=1)
