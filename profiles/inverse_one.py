import os, sys, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mfv2d_b200._capi import lib, check, DEVICE
from mfv2d_b200.runtime import Context
ctx = Context(0)
n, batch = 120, 4000
rng = np.random.default_rng(0)
a = rng.standard_normal((256, n, n)); a = a + a.transpose(0, 2, 1)
big = np.tile(a, (-(-batch // 256), 1, 1))[:batch]
d_in = ctx.to_device(big); d_out = ctx.empty((batch, n, n)); d_st = ctx.empty(batch, np.int32)
for _ in range(2):
    check(lib().mfv2d_cuda_inverse_batch(ctx.handle, n, batch, d_in.ptr, d_out.ptr, d_st.ptr, DEVICE), ctx.handle)
ctx.synchronize()
