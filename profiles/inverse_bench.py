"""Batched inverse: register-resident kernel (dense.cu, inverse_reg_kernel) against the shared-memory kernel and numpy."""
import os, sys, json, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mfv2d_b200._capi import lib, check, DEVICE
from mfv2d_b200.runtime import Context
import mfv2d_b200.runtime as rt
ctx = Context(0); rt._context = ctx
res = {}
rng = np.random.default_rng(0)
for n, batch in ((7, 4096), (33, 4096), (56, 65536), (81, 20000), (120, 40000), (128, 8192)):
    a = rng.standard_normal((min(batch, 512), n, n)); a = a + a.transpose(0, 2, 1)
    a[:, np.arange(n // 2, n), np.arange(n // 2, n)] = 0.0  # zero diagonal tail: pivoting is exercised
    reps = -(-batch // a.shape[0])
    big = np.tile(a, (reps, 1, 1))[:batch]
    d_in = ctx.to_device(big); d_out = ctx.empty((batch, n, n)); d_st = ctx.empty(batch, np.int32)
    out = {}
    for mode in ("1", "0"):
        os.environ["MFV2D_INVERSE_REG"] = mode
        check(lib().mfv2d_cuda_inverse_batch(ctx.handle, n, batch, d_in.ptr, d_out.ptr, d_st.ptr, DEVICE), ctx.handle)
        ctx.synchronize(); ctx.timer_start()
        check(lib().mfv2d_cuda_inverse_batch(ctx.handle, n, batch, d_in.ptr, d_out.ptr, d_st.ptr, DEVICE), ctx.handle)
        t = ctx.timer_stop()
        inv = d_out.get()[: a.shape[0]]
        err = np.abs(inv @ a - np.eye(n)).max()
        out[mode] = {"ms": t, "tflops": 2.0 * n**3 * batch / (t * 1e-3) / 1e12, "max|inv A - I|": float(err), "status_bad": int(np.count_nonzero(d_st.get()))}
    res[f"n{n}_b{batch}"] = {"reg": out["1"], "smem": out["0"], "numpy_cond_max": float(np.linalg.cond(a).max())}
    print(n, batch, res[f"n{n}_b{batch}"], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(res, open("gpurun_out/r2_inverse_bench.json", "w"), indent=1)
