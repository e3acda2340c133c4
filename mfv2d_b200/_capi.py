"""ctypes binding of libmfv2d_cuda.so (include/mfv2d_cuda.h).

There is no CPU fallback: if the shared library has not been built (``python -m mfv2d_b200.build``) or no
B200 is visible, every compute entry point raises.  Importing this module does not need a GPU.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libmfv2d_cuda.so")

HOST, DEVICE = 0, 1

# result codes (src/common/error.h:10-36 plus the library's own)
RESULT_NAMES = {}
BAD_ARGUMENT = 18
CUDA_NO_DEVICE = 101
CUDA_UNSUPPORTED = 102


class Mfv2dCudaError(RuntimeError):
    """``RuntimeError('Runtime error <MFV2D_...> was encountered.')`` as raised by the reference
    (src/evaluation/element_system.c:205-209), with the library's message attached."""

    def __init__(self, code: int, name: str, message: str):
        super().__init__(f"Runtime error {name} was encountered. {message}".strip())
        self.code = code
        self.name = name
        self.message = message


class Op(C.Structure):
    """mfv2d_op_t"""

    _fields_ = [("op", C.c_int32), ("a", C.c_int32), ("b", C.c_int32), ("c", C.c_int32), ("k", C.c_double)]


OP_DTYPE = np.dtype([("op", np.int32), ("a", np.int32), ("b", np.int32), ("c", np.int32), ("k", np.float64)], align=True)

_lib = None
_lib_path = LIB_PATH

vp = C.c_void_p
dp = C.c_void_p  # double* (host numpy pointer or device pointer, passed as integer address)
i32 = C.c_int
sz = C.c_size_t
dbl = C.c_double

# name -> (restype, argtypes); must list EVERY function declared in include/mfv2d_cuda.h
SIGNATURES = {
    "mfv2d_cuda_init": (i32, [C.POINTER(vp), i32]),
    "mfv2d_cuda_destroy": (None, [vp]),
    "mfv2d_cuda_last_error": (C.c_char_p, [vp]),
    "mfv2d_cuda_result_str": (C.c_char_p, [i32]),
    "mfv2d_cuda_synchronize": (i32, [vp]),
    "mfv2d_cuda_device_count": (i32, []),
    "mfv2d_cuda_malloc": (i32, [vp, sz, C.POINTER(vp)]),
    "mfv2d_cuda_free": (i32, [vp, vp]),
    "mfv2d_cuda_memset": (i32, [vp, vp, i32, sz]),
    "mfv2d_cuda_upload": (i32, [vp, vp, vp, sz]),
    "mfv2d_cuda_download": (i32, [vp, vp, vp, sz]),
    "mfv2d_cuda_host_alloc": (i32, [vp, sz, C.POINTER(vp)]),
    "mfv2d_cuda_host_free": (i32, [vp, vp]),
    "mfv2d_cuda_timer_start": (i32, [vp]),
    "mfv2d_cuda_timer_stop": (i32, [vp, C.POINTER(dbl)]),
    "mfv2d_cuda_launch_count": (C.c_uint64, [vp]),
    "mfv2d_cuda_flush_l2": (i32, [vp]),
    "mfv2d_cuda_measure_fp64_peak": (i32, [vp, C.POINTER(dbl), C.POINTER(dbl)]),
    "mfv2d_gll": (i32, [i32, i32, dbl, dp, dp]),
    "mfv2d_lagrange1d": (i32, [i32, dp, i32, dp, dp]),
    "mfv2d_dlagrange1d": (i32, [i32, dp, i32, dp, dp]),
    "mfv2d_legendre": (i32, [i32, i32, dp, dp]),
    "mfv2d_basis_tables": (i32, [i32, i32, dp, dp, dp, dp, dp]),
    "mfv2d_cuda_legendre": (i32, [vp, i32, sz, dp, dp, i32]),
    "mfv2d_cuda_basis_create": (i32, [vp, i32, i32, C.POINTER(vp)]),
    "mfv2d_cuda_basis_destroy": (None, [vp]),
    "mfv2d_cuda_elements_create": (i32, [vp, vp, sz, dp, i32, C.POINTER(vp)]),
    "mfv2d_cuda_elements_destroy": (None, [vp]),
    "mfv2d_cuda_elements_set_corners": (i32, [vp, dp, i32]),
    "mfv2d_cuda_jacobian_batch": (i32, [vp, dp, i32]),
    "mfv2d_cuda_coordinates_batch": (i32, [vp, dp, i32]),
    "mfv2d_cuda_mass_batch": (i32, [vp, i32, i32, dp, vp, i32]),
    "mfv2d_cuda_interprod_batch": (i32, [vp, i32, i32, dp, dp, i32]),
    "mfv2d_cuda_reconstruct_batch": (i32, [vp, i32, dp, dp, i32]),
    "mfv2d_cuda_reconstruct_grid_batch": (i32, [vp, i32, dp, i32, dp, i32, dp, dp, i32]),
    "mfv2d_cuda_dual_dofs_batch": (i32, [vp, i32, dp, dp, i32]),
    "mfv2d_cuda_projector_batch": (i32, [vp, vp, i32, i32, dp, vp, i32]),
    "mfv2d_cuda_incidence_apply": (i32, [vp, i32, i32, i32, sz, i32, dp, dp, i32]),
    "mfv2d_cuda_plan_create": (i32, [vp, vp, i32, vp, vp, vp, vp, i32, vp, C.POINTER(vp)]),
    "mfv2d_cuda_plan_destroy": (None, [vp]),
    "mfv2d_cuda_plan_size": (i32, [vp]),
    "mfv2d_cuda_element_matrix_batch": (i32, [vp, vp, dp, dp, dp, i32]),
    "mfv2d_cuda_element_vector_batch": (i32, [vp, vp, dp, dp, dp, i32]),
    "mfv2d_cuda_inverse_batch": (i32, [vp, i32, sz, dp, dp, vp, i32]),
    "mfv2d_cuda_inverse_definite": (i32, [vp, i32, dp, dp, C.POINTER(i32), i32]),
    "mfv2d_cuda_inverse_definite_noalloc": (i32, [vp, i32, vp, vp, vp]),
    "mfv2d_cuda_batched_gemv": (i32, [vp, sz, i32, i32, dp, dp, dp, i32]),
    "mfv2d_cuda_batched_gemm": (i32, [vp, sz, i32, i32, i32, i32, dp, dp, dp, i32]),
    "mfv2d_cuda_border_batch": (i32, [vp, sz, i32, i32, dp, i32, vp, vp, vp, dp, i32]),
    "mfv2d_cuda_legendre_estimates_batch": (i32, [vp, i32, i32, i32, dp, dp, dp, i32]),
    "mfv2d_cuda_solve_batch": (i32, [vp, i32, i32, sz, dp, dp, dp, vp, i32]),
    "mfv2d_cuda_system_create": (i32, [vp, sz, vp, dp, i32, sz, sz, vp, vp, dp, C.POINTER(vp)]),
    "mfv2d_cuda_system_destroy": (None, [vp]),
    "mfv2d_cuda_system_dense_size": (C.c_int64, [vp]),
    "mfv2d_cuda_system_trace_size": (C.c_int64, [vp]),
    "mfv2d_cuda_system_apply_diagonal": (i32, [vp, dp, dp, i32]),
    "mfv2d_cuda_system_apply_diagonal_inverse": (i32, [vp, dp, dp, i32]),
    "mfv2d_cuda_system_apply_trace": (i32, [vp, dp, dp, i32]),
    "mfv2d_cuda_system_apply_trace_transpose": (i32, [vp, dp, dp, i32]),
    "mfv2d_cuda_system_apply_schur": (i32, [vp, dp, dp, i32]),
    "mfv2d_cuda_vec_axpby": (i32, [vp, sz, dbl, dp, dbl, dp, dp, i32]),
    "mfv2d_cuda_vec_dot": (i32, [vp, sz, dp, dp, C.POINTER(dbl), i32]),
    "mfv2d_cuda_vec_absmax": (i32, [vp, sz, dp, C.POINTER(dbl), i32]),
    "mfv2d_cuda_strided_axpby": (i32, [vp, sz, sz, dbl, dp, sz, dbl, dp, sz, dp, sz]),
    "mfv2d_cuda_solve_schur_cg": (i32, [vp, dp, dp, i32, dbl, dbl, i32, dp, dp, C.POINTER(dbl), C.POINTER(i32), i32]),
    "mfv2d_cuda_solve_schur_bicgstab": (i32, [vp, dp, dp, i32, dbl, dbl, i32, dp, dp, C.POINTER(dbl), C.POINTER(i32), i32]),
    "mfv2d_cuda_solve_pcg": (i32, [vp, dp, dp, i32, dbl, dbl, dp, dp, C.POINTER(dbl), C.POINTER(i32), i32]),
    "mfv2d_cuda_comm_id_size": (i32, []),
    "mfv2d_cuda_comm_get_id": (i32, [vp]),
    "mfv2d_cuda_comm_init": (i32, [vp, i32, i32, vp]),
    "mfv2d_cuda_comm_destroy": (i32, [vp]),
    "mfv2d_cuda_comm_allreduce_sum": (i32, [vp, dp, sz]),
    "mfv2d_cuda_system_set_interface": (i32, [vp, sz, vp, vp, sz, vp]),
    "mfv2d_cuda_peer_handle_size": (i32, []),
    "mfv2d_cuda_system_peer_export": (i32, [vp, sz, vp]),
    "mfv2d_cuda_system_peer_connect": (i32, [vp, vp, sz, vp, vp]),
    "mfv2d_cuda_system_peer_error": (i32, [vp]),
    "mfv2d_cuda_system_peer_stats": (i32, [vp, vp]),
    "mfv2d_cuda_system_set_coarse_space": (i32, [vp, i32, vp, vp, vp, i32]),
}


def _use_library_for_testing(path: str) -> None:
    """Point the binding at another build of the same C-ABI.  Only tests/emu uses this (the host
    emulation of the kernels, see tests/emu/emu_cuda.h); nothing in the package calls it."""
    global _lib, _lib_path
    if os.path.abspath(path) != os.path.abspath(LIB_PATH) and os.environ.get("MFV2D_TEST_HOOKS") != "1":
        raise RuntimeError("only libmfv2d_cuda.so may back mfv2d_b200; another build of the C-ABI is accepted under "
                           "MFV2D_TEST_HOOKS=1 alone (tests/conftest.py sets it; the product never does)")
    _lib = None
    _lib_path = path
    from . import runtime

    runtime._reset_for_testing()


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_lib_path):
            raise RuntimeError(
                f"{_lib_path} is missing: build it with `python -m mfv2d_b200.build` (nvcc, sm_100a). "
                "mfv2d_b200 has no CPU fallback."
            )
        handle = C.CDLL(_lib_path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # raises AttributeError if the library lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def result_name(code: int) -> str:
    return lib().mfv2d_cuda_result_str(code).decode()


def check(code: int, ctx=None) -> None:
    if code == 0:
        return
    msg = lib().mfv2d_cuda_last_error(ctx).decode(errors="replace")
    name = result_name(code)
    if code == BAD_ARGUMENT:
        raise ValueError(msg)
    raise Mfv2dCudaError(code, name, msg)


def ptr(a) -> int | None:
    """Address of a numpy array's buffer (host pointer) or pass-through of an int device pointer."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    return int(a)
