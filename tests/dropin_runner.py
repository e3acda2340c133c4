"""Runs the REFERENCE's own test files against its Python package with the hot-path names of `mfv2d._mfv2d` replaced by
this repository's mirror, in a fresh interpreter.

    python tests/dropin_runner.py emu|gpu <result.json> <test file> ...

`emu`: the kernels run through the host emulation build (tests/emu) - the `-m "not gpu"` acceptance run;
`gpu`: through libmfv2d_cuda.so.  Needs /root/reference and oracle/_ref (the reference extension supplies the types that
are out of scope: Mesh, Manifold2D, GeoID, Line, Surface, ...).  TEST INFRASTRUCTURE ONLY.
"""
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tests", "emu"))

HOT_PATH = ("compute_element_matrix", "compute_element_vector", "compute_element_mass_matrix", "compute_integrating_fields",
            "compute_element_projector", "ElementFemSpace2D", "Basis1D", "Basis2D", "IntegrationRule1D", "_ElementFormSpecification",
            "compute_gll", "GLLCache", "DEFAULT_GLL_CACHE", "lagrange1d", "dlagrange1d", "compute_legendre", "check_bytecode", "check_incidence", "_compute_matrix_inverse",
            "_solve_linear_system")
SOLVER = ("LinearSystem", "DenseVector", "TraceVector", "SparseVector", "MatrixCRS")


def main():
    mode, result_path, files = sys.argv[1], sys.argv[2], sys.argv[3:]
    if mode == "emu":
        from build_emu import build

        import mfv2d_b200._capi as capi

        os.environ.setdefault("MFV2D_TEST_HOOKS", "1")
        capi._use_library_for_testing(build())
    from oracle.ref_loader import load_ref

    ref = load_ref()
    import mfv2d_b200._mfv2d as gpu
    import mfv2d_b200.solving as gpu_solving

    patched = types.ModuleType("mfv2d._mfv2d")
    patched.__dict__.update({k: v for k, v in ref.__dict__.items() if not k.startswith("__")})
    replaced = []
    for name in HOT_PATH:
        if hasattr(gpu, name):
            setattr(patched, name, getattr(gpu, name))
            replaced.append(name)
    for name in SOLVER:
        setattr(patched, name, getattr(gpu_solving, name) if hasattr(gpu_solving, name) else getattr(gpu, name))
        replaced.append(name)
    # pyvista stand-in (graphics stack: unavailable on this box)
    pv = types.ModuleType("pyvista")
    pv.CellType = type("CellType", (), {"LAGRANGE_QUADRILATERAL": 70})

    class UnstructuredGrid:
        def __init__(self, cells=None, celltypes=None, points=None):
            self.cells, self.celltypes, self.points = cells, celltypes, points
            self.point_data, self.cell_data, self.field_data = {}, {}, {}

    pv.UnstructuredGrid = UnstructuredGrid
    sys.modules["pyvista"] = pv
    sys.path.insert(0, "/root/reference/python")
    sys.modules["mfv2d._mfv2d"] = patched
    import mfv2d  # noqa: F401  (every `from mfv2d._mfv2d import ...` of the package now sees the mirror)
    import pytest

    class Collect:
        def __init__(self):
            self.outcomes = {}

        def pytest_runtest_logreport(self, report):
            if report.when == "call" or (report.when == "setup" and report.outcome != "passed"):
                self.outcomes[report.nodeid] = report.outcome

    c = Collect()
    rc = pytest.main(["-c", "/dev/null", "--rootdir", "/tmp", "-p", "no:cacheprovider", "-q", "-x" if os.environ.get("DROPIN_X") else "-q",
                      "--tb=short", *files], plugins=[c])
    summary = {"mode": mode, "replaced": replaced, "rc": int(rc), "passed": sum(v == "passed" for v in c.outcomes.values()),
               "failed": sorted(k for k, v in c.outcomes.items() if v == "failed"), "skipped": sum(v == "skipped" for v in c.outcomes.values())}
    with open(result_path, "w") as f:
        json.dump(summary, f, indent=1)
    print(json.dumps({k: (v if k != "failed" else len(v)) for k, v in summary.items()}))


if __name__ == "__main__":
    main()
