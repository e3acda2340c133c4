"""Process-wide device context, device buffers and handle caches (basis tables, bytecode plans)."""
from __future__ import annotations

import ctypes as C
import os
import weakref

import numpy as np

from . import _capi
from ._capi import DEVICE, HOST, check, lib, ptr

_context = None


class Context:
    """One CUDA device + stream (mfv2d_cuda_init).  One per process (one process per GPU)."""

    def __init__(self, device: int | None = None):
        if device is None:
            device = int(os.environ.get("LOCAL_RANK", "0"))
        h = C.c_void_p()
        check(lib().mfv2d_cuda_init(C.byref(h), device), None)
        self.handle = h
        self.device = device
        self._basis: dict[tuple[int, int], "BasisHandle"] = {}
        self._plans: dict = {}
        self.n_ranks = 1
        self.rank = 0

    # -- memory ------------------------------------------------------------------------------
    def empty(self, shape, dtype=np.float64) -> "DeviceArray":
        return DeviceArray(self, shape, dtype)

    def to_device(self, a: np.ndarray) -> "DeviceArray":
        a = np.ascontiguousarray(a)
        d = DeviceArray(self, a.shape, a.dtype)
        if a.nbytes:
            check(lib().mfv2d_cuda_upload(self.handle, d.ptr, a.ctypes.data, a.nbytes), self.handle)
        return d

    def pinned(self, shape, dtype=np.float64) -> np.ndarray:
        """numpy array backed by page-locked host memory (for the host-buffer / e2e path)."""
        dtype = np.dtype(dtype)
        n = int(np.prod(shape)) * dtype.itemsize
        h = C.c_void_p()
        check(lib().mfv2d_cuda_host_alloc(self.handle, n, C.byref(h)), self.handle)
        buf = (C.c_char * max(n, 1)).from_address(h.value)
        arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
        ctx_handle, addr = self.handle, h.value
        weakref.finalize(buf, lambda: lib().mfv2d_cuda_host_free(ctx_handle, addr))
        return arr

    def synchronize(self) -> None:
        check(lib().mfv2d_cuda_synchronize(self.handle), self.handle)

    def timer_start(self) -> None:
        check(lib().mfv2d_cuda_timer_start(self.handle), self.handle)

    def timer_stop(self) -> float:
        ms = C.c_double()
        check(lib().mfv2d_cuda_timer_stop(self.handle, C.byref(ms)), self.handle)
        return ms.value

    def measure_fp64_peak(self) -> tuple[float, float]:
        """(DMMA, DFMA) fp64 TFLOP/s of this device, measured now (mfv2d_cuda_measure_fp64_peak)."""
        a, b = C.c_double(), C.c_double()
        check(lib().mfv2d_cuda_measure_fp64_peak(self.handle, C.byref(a), C.byref(b)), self.handle)
        return a.value, b.value

    def flush_l2(self) -> None:
        check(lib().mfv2d_cuda_flush_l2(self.handle), self.handle)

    @property
    def launch_count(self) -> int:
        return int(lib().mfv2d_cuda_launch_count(self.handle))

    # -- caches ------------------------------------------------------------------------------
    def basis(self, p: int, q: int) -> "BasisHandle":
        key = (int(p), int(q))
        if key not in self._basis:
            self._basis[key] = BasisHandle(self, *key)
        return self._basis[key]

    # -- multi-GPU ---------------------------------------------------------------------------
    def init_comm(self, n_ranks: int, rank: int, id_bytes: bytes | None) -> None:
        """Join the NCCL communicator; ``id_bytes`` comes from ``comm_unique_id()`` on rank 0 and is
        distributed by the caller (e.g. a torch.distributed / gloo broadcast - plumbing only)."""
        # id_bytes None: ranks only - every exchange then goes through LinearSystem.connect_peers (NVLink peer memory)
        buf = (C.c_char * len(id_bytes)).from_buffer_copy(id_bytes) if id_bytes else None
        check(lib().mfv2d_cuda_comm_init(self.handle, n_ranks, rank, buf), self.handle)
        self.n_ranks, self.rank = n_ranks, rank

    def allreduce_sum(self, d: "DeviceArray") -> None:
        check(lib().mfv2d_cuda_comm_allreduce_sum(self.handle, d.ptr, d.size), self.handle)

    def close(self) -> None:
        if self.handle:
            for b in self._basis.values():
                b.close()
            self._basis.clear()
            lib().mfv2d_cuda_destroy(self.handle)
            self.handle = None


def comm_unique_id() -> bytes:
    n = lib().mfv2d_cuda_comm_id_size()
    buf = (C.c_char * n)()
    check(lib().mfv2d_cuda_comm_get_id(buf), None)
    return bytes(buf)


def get_context() -> Context:
    """The process's default context; created on first use.  Raises when no B200 is visible."""
    global _context
    if _context is None:
        _context = Context()
    return _context


def _reset_for_testing() -> None:
    global _context
    _context = None


class DeviceArray:
    """A typed, shaped block of device memory (mfv2d_cuda_malloc)."""

    def __init__(self, ctx: Context, shape, dtype=np.float64):
        self.ctx = ctx
        self.shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list)) else (shape,)))
        self.dtype = np.dtype(dtype)
        self.size = int(np.prod(self.shape)) if self.shape else 1
        self.nbytes = self.size * self.dtype.itemsize
        h = C.c_void_p()
        check(lib().mfv2d_cuda_malloc(ctx.handle, self.nbytes, C.byref(h)), ctx.handle)
        self.ptr = h.value
        self._fin = weakref.finalize(self, DeviceArray._release, ctx, h.value)

    @staticmethod
    def _release(ctx, p):
        if ctx.handle:
            lib().mfv2d_cuda_free(ctx.handle, p)

    def free(self) -> None:
        self._fin()

    def get(self) -> np.ndarray:
        out = np.empty(self.shape, self.dtype)
        if self.nbytes:
            check(lib().mfv2d_cuda_download(self.ctx.handle, out.ctypes.data, self.ptr, self.nbytes), self.ctx.handle)
        return out

    def get_into(self, out: np.ndarray) -> np.ndarray:
        assert out.nbytes == self.nbytes and out.flags.c_contiguous
        check(lib().mfv2d_cuda_download(self.ctx.handle, out.ctypes.data, self.ptr, self.nbytes), self.ctx.handle)
        return out

    def set(self, a: np.ndarray) -> None:
        a = np.ascontiguousarray(a, self.dtype)
        assert a.nbytes == self.nbytes
        check(lib().mfv2d_cuda_upload(self.ctx.handle, self.ptr, a.ctypes.data, a.nbytes), self.ctx.handle)

    def zero(self) -> None:
        check(lib().mfv2d_cuda_memset(self.ctx.handle, self.ptr, 0, self.nbytes), self.ctx.handle)


class BasisHandle:
    """Device-resident Basis2D tables of one (p, q) bucket (mfv2d_cuda_basis_create)."""

    def __init__(self, ctx: Context, p: int, q: int):
        self.ctx, self.p, self.q = ctx, p, q
        h = C.c_void_p()
        check(lib().mfv2d_cuda_basis_create(ctx.handle, p, q, C.byref(h)), ctx.handle)
        self.handle = h

    def close(self):
        if self.handle:
            lib().mfv2d_cuda_basis_destroy(self.handle)
            self.handle = None


def space_of(*arrays) -> int:
    """HOST if every non-None argument is a numpy array, DEVICE if every one is a DeviceArray."""
    kinds = {isinstance(a, DeviceArray) for a in arrays if a is not None}
    if len(kinds) > 1:
        raise TypeError("cannot mix host (numpy) and device (DeviceArray) buffers in one call")
    return DEVICE if kinds == {True} else HOST


def addr(a) -> int | None:
    if a is None:
        return None
    if isinstance(a, DeviceArray):
        return a.ptr
    return ptr(a)


__all__ = ["Context", "DeviceArray", "BasisHandle", "get_context", "comm_unique_id", "space_of", "addr", "HOST", "DEVICE"]
