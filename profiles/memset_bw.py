import sys; sys.path.insert(0,'.')
from mfv2d_b200.runtime import Context
from mfv2d_b200._capi import lib, check
ctx=Context(0)
n=65536*81*81
buf=ctx.empty((n,))
from mfv2d_b200.runtime import addr
for rep in range(2):
    ctx.synchronize(); ctx.timer_start()
    for _ in range(10):
        check(lib().mfv2d_cuda_memset(ctx.handle, addr(buf), 0, n*8), ctx.handle)
    ms=ctx.timer_stop()/10
    print("memset %.3f ms  %.1f GB/s"%(ms, n*8/ms/1e6))
