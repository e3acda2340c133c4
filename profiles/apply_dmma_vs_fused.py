import os, sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import bench as B
from mfv2d_b200 import _mfv2d as M, structured as ST
import mfv2d_b200.runtime as rt
from mfv2d_b200.runtime import Context
ctx = Context(0); rt._context = ctx
for name, p in (("stokes", 4), ("mixed_poisson", 4), ("stokes", 3)):
    names, orders, lhs = B.load_system(name)
    specs = M._ElementFormSpecification(*zip(names, orders))
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + 3)); basis = M.Basis2D(b1, b1)
    corners = ST.unit_square_corners(256, 256, B.deform); n = corners.shape[0]
    batch = M.ElementBatch(basis, ctx.to_device(corners), ctx)
    ne = specs.total_size(p, p)
    x = ctx.to_device(np.random.default_rng(0).standard_normal((n, ne))); y = ctx.empty((n, ne))
    out = {}
    for mode in ("1", "0"):
        os.environ["MFV2D_APPLY_DMMA"] = mode
        for _ in range(3): M.compute_element_vectors(specs, lhs, batch, x, out=y)
        ctx.synchronize(); ctx.timer_start()
        for _ in range(10): M.compute_element_vectors(specs, lhs, batch, x, out=y)
        t = ctx.timer_stop() / 10
        out["dmma" if mode == "1" else "fused"] = round(n * ne / (t * 1e-3) / 1e9, 2)
    print(name, p, out, flush=True)
