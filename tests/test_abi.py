"""`-m "not gpu"`: the C-ABI library builds for sm_100a, loads, exports every symbol that
include/mfv2d_cuda.h declares, the ctypes table covers exactly those, and - with no GPU in this
container - the compute path fails loudly instead of falling back to anything."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    text = open(os.path.join(ROOT, "include", "mfv2d_cuda.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"MFV2D_API\s+[\w\s\*]+?\b(mfv2d_\w+)\s*\(", text)))


@pytest.fixture(scope="module")
def libpath():
    from mfv2d_b200.build import build

    return build()


def test_header_symbols_exported(libpath):
    syms = header_symbols()
    assert len(syms) > 50
    handle = ctypes.CDLL(libpath)
    for s in syms:
        assert hasattr(handle, s), f"{s} is declared in include/mfv2d_cuda.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", libpath], capture_output=True, text=True).stdout
    exported = sorted(set(re.findall(r" T (mfv2d_\w+)", out)))
    assert exported == syms, "exported symbols and the header differ"


def test_ctypes_table_matches_header(libpath):
    from mfv2d_b200 import _capi

    assert sorted(_capi.SIGNATURES) == header_symbols()


def test_built_for_sm_100a(libpath):
    out = subprocess.run(["cuobjdump", "-lelf", libpath], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


def test_no_torch_in_library(libpath):
    out = subprocess.run(["ldd", libpath], capture_output=True, text=True).stdout
    names = [ln.split()[0] for ln in out.splitlines() if ln.strip()]  # library names only (load addresses are random hex)
    assert not any("torch" in n or "c10" in n for n in names), names


def test_fails_loudly_without_gpu(libpath):
    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(capi.LIB_PATH)
    if capi.lib().mfv2d_cuda_device_count() > 0:
        pytest.skip("a GPU is visible")
    from mfv2d_b200.runtime import Context

    with pytest.raises(RuntimeError, match="MFV2D_CUDA_NO_DEVICE"):
        Context(0)


def test_host_tables_need_no_gpu(libpath):
    import numpy as np

    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(capi.LIB_PATH)
    from mfv2d_b200 import _mfv2d as M

    x, w = M.compute_gll(5)
    assert abs(w.sum() - 2) < 1e-14 and x[0] == -1 and x[-1] == 1
    b = M.Basis1D(3, M.IntegrationRule1D(6))
    assert b.node.shape == (4, 7) and b.edge.shape == (3, 7)
    assert np.allclose(b.node.sum(axis=0), 1.0)
