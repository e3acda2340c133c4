import sys, os, json, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import bench as B
from mfv2d_b200 import _mfv2d as M, structured as ST
from mfv2d_b200.runtime import Context
import mfv2d_b200.runtime as rt
ctx = Context(0); rt._context = ctx
res = {}
for name, p, n_side in (("mixed_poisson", 6, 354), ("stokes", 4, 256), ("mixed_poisson", 4, 256), ("mixed_poisson", 3, 256), ("mixed_poisson", 8, 256), ("direct_poisson", 5, 256)):
    names, orders, lhs = B.load_system(name)
    specs = M._ElementFormSpecification(*zip(names, orders))
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + 3)); basis = M.Basis2D(b1, b1)
    corners = ST.unit_square_corners(n_side, n_side, B.deform); n = corners.shape[0]
    batch = M.ElementBatch(basis, ctx.to_device(corners), ctx)
    ne = specs.total_size(p, p)
    rng = np.random.default_rng(0)
    xh = rng.standard_normal((n, ne)); x = ctx.to_device(xh); y = ctx.empty((n, ne))
    out = {}
    for mode in ("1", "0"):
        os.environ["MFV2D_APPLY_WARP"] = mode
        for _ in range(3): M.compute_element_vectors(specs, lhs, batch, x, out=y)
        ctx.synchronize(); ctx.timer_start()
        for _ in range(10): M.compute_element_vectors(specs, lhs, batch, x, out=y)
        t = ctx.timer_stop() / 10
        out[mode] = (n * ne / (t * 1e-3) / 1e9, y.get())
    err = np.abs(out["1"][1] - out["0"][1]).max() / np.abs(out["0"][1]).max()
    macs = B.apply_macs(name, p, p + 4) if name in ("mixed_poisson", "stokes") else 0
    res[f"{name}_p{p}"] = {"warp_gdofs": out["1"][0], "old_gdofs": out["0"][0], "relerr_vs_old": err, "frac_fp64_warp": 2 * macs * n / ne * 0 if not macs else out["1"][0] * 1e9 / ne * 2 * macs / 37.2e12}
    print(name, p, res[f"{name}_p{p}"], flush=True)
json.dump(res, open("/root/repo/gpurun_out/r2_apply_warp.json", "w"), indent=1)
