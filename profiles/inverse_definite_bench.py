"""One large definite matrix: blocked Gauss-Jordan (default) against one pivot per pass (MFV2D_INVERSE_BLOCKED=0), device buffers."""
import ctypes as C
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, "/root/repo")
from mfv2d_b200._capi import DEVICE, check, lib
from mfv2d_b200.runtime import Context

ctx = Context(0)
res = {}
for n in (350, 1024, 4096):
    rng = np.random.default_rng(n)
    g = rng.standard_normal((n, n))
    a = g @ g.T + n * np.eye(n)
    d_in, d_out = ctx.to_device(a), ctx.empty((n, n))
    out = {}
    for mode in ("0", "1"):
        os.environ["MFV2D_INVERSE_BLOCKED"] = mode
        st = C.c_int(0)
        for rep in range(2):
            ctx.synchronize()
            t = time.perf_counter()
            check(lib().mfv2d_cuda_inverse_definite(ctx.handle, n, d_in.ptr, d_out.ptr, C.byref(st), DEVICE), ctx.handle)
            ctx.synchronize()
            dt = time.perf_counter() - t
        inv = d_out.get()
        out["blocked" if mode == "1" else "per_pivot"] = {"ms": round(dt * 1e3, 3), "max_abs_inv_a_minus_i": float(np.max(np.abs(inv @ a - np.eye(n)))),
                                                           "tflops": round(2.0 * n**3 / dt / 1e12, 3)}
    res[f"n{n}"] = out
    print(n, out, flush=True)
json.dump(res, open("/root/repo/gpurun_out/inverse_definite_r02.json", "w"), indent=1)
