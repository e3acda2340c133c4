"""Loader for the compiled reference extension (oracle/_ref/_mfv2d_ref.so).  TEST INFRASTRUCTURE ONLY.

``load_ref()`` returns the reference's own ``_mfv2d`` module (built by oracle/build_ref.sh from the
sources under /root/reference) or ``None`` when it has not been built.  It is used by tests as the
strongest checker and by ``bench.py``'s CPU-baseline legs (``cpu_baseline.kind == "reference"``).

``load_ref_python()`` additionally imports the reference's *Python* package from /root/reference
(only possible in the build container; used by tests/golden/make_golden.py to generate fixtures).
"""
from __future__ import annotations

import importlib.machinery
import importlib.util
import os
import sys
import types

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "_mfv2d_ref.so")
_REF = None


def load_ref():
    global _REF
    if _REF is not None:
        return _REF
    if not os.path.exists(_SO):
        return None
    loader = importlib.machinery.ExtensionFileLoader("_mfv2d", _SO)
    spec = importlib.util.spec_from_file_location("_mfv2d", _SO, loader=loader)
    mod = importlib.util.module_from_spec(spec)
    loader.exec_module(mod)
    _REF = mod
    return mod


def load_ref_python(reference_root: str = "/root/reference"):
    """Import ``mfv2d`` from the reference checkout with the compiled extension injected and a
    minimal stand-in for the (absent) ``pyvista`` package.  Build-container only."""
    ref = load_ref()
    pydir = os.path.join(reference_root, "python")
    if ref is None or not os.path.isdir(pydir):
        return None
    if "mfv2d" in sys.modules and getattr(sys.modules["mfv2d"], "__file__", "").startswith(pydir):
        return sys.modules["mfv2d"]
    if "pyvista" not in sys.modules:
        pv = types.ModuleType("pyvista")

        class CellType:
            LAGRANGE_QUADRILATERAL = 70

        class UnstructuredGrid:
            def __init__(self, cells=None, celltypes=None, points=None):
                self.cells, self.celltypes, self.points = cells, celltypes, points
                self.point_data, self.cell_data, self.field_data = {}, {}, {}

        pv.CellType = CellType
        pv.UnstructuredGrid = UnstructuredGrid
        sys.modules["pyvista"] = pv
    sys.path.insert(0, pydir)
    sys.modules["mfv2d._mfv2d"] = ref
    import mfv2d  # noqa: E402

    return mfv2d
