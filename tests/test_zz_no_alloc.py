"""The allocation-free entry point ``mfv2d_cuda_inverse_definite_noalloc`` (mfv2d_b200/csrc/noalloc.cu; it has no
counterpart in the reference) against the existing entry point ``mfv2d_cuda_inverse_batch``, through the C-ABI.

Collected LAST on purpose (``test_zz_``): a failure here must not hide the
parity tests.  Two calls in a row share the library's static scratch array: the second call - a smaller matrix, then the
first one again - must not see what the previous call left behind.  The file passes alone in a fresh process and after
the other tests (it only needs a context, which ``get_context`` creates or reuses).
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

TOL = 1e-10  # the tolerance of the inverse tests (parity_cases.check_linalg_inverse)
MAX_N = 4096  # MFV2D_NOALLOC_INVERSE_MAX_N of include/mfv2d_cuda.h
DEVICE = 1  # MFV2D_DEVICE


def relerr(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def definite(n, seed):
    rng = np.random.default_rng(seed)
    q = rng.standard_normal((n, n))
    return q @ q.T / n + np.eye(n) * (1.0 + rng.random())


class DeviceBuffer:
    """The CALLER's buffers: the entry point under test allocates nothing, its caller does."""

    def __init__(self, capi, ctx, nbytes):
        self.capi, self.ctx, self.ptr = capi, ctx, C.c_void_p()
        capi.check(capi.lib().mfv2d_cuda_malloc(ctx, nbytes, C.byref(self.ptr)), ctx)

    def upload(self, a):
        self.capi.check(self.capi.lib().mfv2d_cuda_upload(self.ctx, self.ptr, a.ctypes.data, a.nbytes), self.ctx)

    def download(self, a):
        self.capi.check(self.capi.lib().mfv2d_cuda_download(self.ctx, a.ctypes.data, self.ptr, a.nbytes), self.ctx)
        return a

    def free(self):
        self.capi.check(self.capi.lib().mfv2d_cuda_free(self.ctx, self.ptr), self.ctx)


def noalloc_inverse(capi, ctx, a, in_place=False):
    n = a.shape[0]
    lib = capi.lib()
    d_in, d_status = DeviceBuffer(capi, ctx, a.nbytes), DeviceBuffer(capi, ctx, 4)
    d_out = d_in if in_place else DeviceBuffer(capi, ctx, a.nbytes)
    try:
        d_in.upload(np.ascontiguousarray(a))
        d_status.upload(np.full(1, 77, np.int32))  # the entry point has to clear it
        capi.check(lib.mfv2d_cuda_inverse_definite_noalloc(ctx, n, d_in.ptr, d_out.ptr, d_status.ptr), ctx)
        capi.check(lib.mfv2d_cuda_synchronize(ctx), ctx)
        out = d_out.download(np.empty_like(a))
        status = d_status.download(np.empty(1, np.int32))
    finally:
        d_in.free()
        d_status.free()
        if not in_place:
            d_out.free()
    return out, int(status[0])


def existing_inverse(capi, ctx, a):
    out, status = np.empty_like(a), np.zeros(1, np.int32)
    HOST = 0
    capi.check(capi.lib().mfv2d_cuda_inverse_batch(ctx, a.shape[0], 1, a.ctypes.data, out.ctypes.data, status.ctypes.data, HOST), ctx)
    assert status[0] == 0
    return out


def run_checks(capi, ctx, sizes):
    mats = [definite(n, 100 + n) for n in sizes]
    want = [existing_inverse(capi, ctx, a) for a in mats]
    # two (and more) calls in a row on the shared scratch: large, small, large again, the last one in place
    order = list(range(len(sizes))) + [0]
    for step, i in enumerate(order):
        got, status = noalloc_inverse(capi, ctx, mats[i], in_place=(step == len(order) - 1))
        assert status == 0
        assert relerr(got, want[i]) < TOL, (sizes[i], step)
        assert relerr(got @ mats[i], np.eye(sizes[i])) < TOL
    # a vanishing pivot is reported through the caller's status word, not by a host error
    _, status = noalloc_inverse(capi, ctx, np.zeros((5, 5)))
    assert status == 20  # MFV2D_PIVOT_FAILED
    # ... and the next call is not affected by it
    got, status = noalloc_inverse(capi, ctx, mats[-1])
    assert status == 0 and relerr(got, want[-1]) < TOL
    # sizes beyond the static scratch are refused with a clear error, before anything is launched
    launches = capi.lib().mfv2d_cuda_launch_count(ctx)
    code = capi.lib().mfv2d_cuda_inverse_definite_noalloc(ctx, MAX_N + 1, 8, 8, 8)
    assert code == 102  # MFV2D_CUDA_UNSUPPORTED
    message = capi.lib().mfv2d_cuda_last_error(ctx).decode()
    assert "MFV2D_NOALLOC_INVERSE_MAX_N" in message and str(MAX_N + 1) in message
    assert capi.lib().mfv2d_cuda_launch_count(ctx) == launches
    with pytest.raises(ValueError):
        capi.check(capi.lib().mfv2d_cuda_inverse_definite_noalloc(ctx, 0, 8, 8, 8), ctx)


@pytest.mark.gpu
def test_no_alloc_inverse_matches_inverse_batch():
    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(capi.LIB_PATH)
    from mfv2d_b200.runtime import get_context

    ctx = get_context()  # raises loudly without a B200 / without the built library
    assert ctx.handle
    # 300 and 200 do not fit the shared memory of inverse_batch's kernels (its global-memory path), 96 and 33 do
    run_checks(capi, ctx.handle, sizes=(300, 96, 200, 33))


def test_no_alloc_inverse_host_emulation():
    """The same calls through the host emulation of the kernels (index arithmetic and argument checks only)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
    from build_emu import build

    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(build())
    try:
        from mfv2d_b200.runtime import get_context

        run_checks(capi, get_context().handle, sizes=(40, 17))
    finally:
        capi._use_library_for_testing(capi.LIB_PATH)
