"""Drop-in acceptance run (SURVEY.md section 4 "reuse plan"): the REFERENCE's own test files, unmodified,
executed against its Python package with the hot-path names of `mfv2d._mfv2d` replaced by this repository's mirror
(tests/dropin_runner.py builds the patched module in a fresh interpreter before `import mfv2d`).

`-m "not gpu"`: kernels through the host emulation build.  Skipped where /root/reference is absent (the GPU box); the
`-m gpu` twin would need the reference there as well and is therefore skipped on the box too - parity of the CUDA path
itself is what tests/test_gpu_parity.py covers with the same check functions as the emulated run.

Known, documented gap: 21 parametrisations use DIFFERENT orders in the two directions (`Basis2D(Basis1D(2, .), Basis1D(4, .))`)
for mass matrices and field evaluation.  The reference supports that outside its hot path only (compute_element_matrix
itself requires order_1 == order_2, src/evaluation/element_system.c:43-51); the kernels here are written for the square
case and raise NotImplementedError.  Everything else must pass."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_TESTS = "/root/reference/tests"
FILES = ["test_basis.py", "test_gll.py", "test_legendre.py", "test_interior.py", "test_explicit.py", "test_compute_fields.py",
         "test_projections.py", "test_solving.py", "test_apply_incidence-c.py", "test_linalg_inverse.py", "test_element_form_specs.py",
         "test_svector.py", "test_crs_matrix.py", "test_bytecode.py", "test_dofs.py", "test_enums.py", "test_incidence_python.py",
         "test_apply_incidence.py"]
ANISOTROPIC = {"test_explicit": {"1-2", "1-7", "2-1", "3-5", "5-3", "7-1"}, "test_mixed": {"1-2", "1-7", "2-1", "3-5", "5-3", "7-1"},
               "test_solution_based": {"1-2", "1-7", "2-1", "3-5", "5-3", "7-1"}, "test_mass_matrix_inverses": {"2-4", "4-2", "8-3"}}


def _run(mode, tmp_path):
    if not os.path.isdir(REF_TESTS) or not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "_mfv2d_ref.so")):
        pytest.skip("needs /root/reference and oracle/_ref (build container only)")
    out = str(tmp_path / "dropin.json")
    res = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dropin_runner.py"), mode, out, *[os.path.join(REF_TESTS, f) for f in FILES]],
                         capture_output=True, text=True, timeout=1500, cwd=ROOT)
    assert os.path.exists(out), res.stdout[-3000:] + res.stderr[-3000:]
    with open(out) as f:
        return json.load(f)


def _check(summary):
    unexpected = []
    for nodeid in summary["failed"]:
        name, _, param = nodeid.split("::")[-1].partition("[")
        if param.rstrip("]") not in ANISOTROPIC.get(name, ()):
            unexpected.append(nodeid)
    assert not unexpected, unexpected
    assert summary["passed"] >= 255, summary["passed"]
    assert {"compute_element_matrix", "compute_element_vector", "ElementFemSpace2D", "LinearSystem", "MatrixCRS"} <= set(summary["replaced"])


def test_reference_tests_against_the_mirror_emulated(tmp_path):
    _check(_run("emu", tmp_path))


@pytest.mark.gpu
def test_reference_tests_against_the_mirror_gpu(tmp_path):
    _check(_run("gpu", tmp_path))
