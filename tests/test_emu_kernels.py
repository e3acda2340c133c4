"""`-m "not gpu"`: the kernels' sources executed through the host emulation (tests/emu/emu_cuda.h) and
checked against the oracle and the reference's golden outputs.  This validates the HOST logic of the
library (bytecode planner, shared-memory layouts, incidence maps, constraint structure) and the
kernels' index arithmetic in the GPU-less container.  It is not a parity claim for the CUDA path - the
`-m gpu` tests in test_gpu_parity.py run the same checks through libmfv2d_cuda.so on the B200."""
import os
import sys

import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

import parity_cases as P


@pytest.fixture(scope="module")
def emu():
    from build_emu import build

    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(build())
    from mfv2d_b200 import _mfv2d as M
    from mfv2d_b200 import solving as SV
    from mfv2d_b200 import structured as ST

    yield M, SV, ST
    capi._use_library_for_testing(capi.LIB_PATH)


def test_tables(emu):
    P.check_tables(emu[0])


def test_masses_golden(emu):
    P.check_masses_golden(emu[0])


def test_masses_oracle(emu):
    P.check_masses_oracle(emu[0], cases=((1, 0), (2, 3), (3, 2), (6, 3)), n_random=4)


@pytest.mark.parametrize("name", sorted(P.SYSTEMS))
def test_system_golden(emu, name):
    P.check_system_golden(emu[0], name, cases=((1, 2), (3, 3), (4, 2), (5, 0)))


@pytest.mark.parametrize("mesh,gate,sol_tol", [("b", 1e-10, 1e-7), ("a", 2e-7, 1e-5)])
def test_hp_forest(emu, mesh, gate, sol_tol):
    """configs[3] from arrays alone: own quadtree, own hanging-side rows, mixed block sizes, Schur CG.  Mesh a (93 leaves,
    orders 2..8) has hanging-side coefficients up to 4e8: gates as for the reference-generated twin of it (test_hp_refined)"""
    print(P.check_hp_forest(*emu, mesh, gate, sol_tol))


def test_hp_forest_boundary_conditions(emu):
    """the same solve with boundary conditions built from arrays: strong flux rows on half of the boundary (single-entry rows
    next to the hanging-side rows), natural condition on the rest"""
    print(P.check_hp_forest(*emu, "b", 1e-10, 1e-7, bc=True))


def test_degenerate_element(emu):
    P.check_degenerate_element(emu[0])


def test_dual_dofs(emu):
    P.check_dual_dofs(emu[0])


def test_mixed_order_buckets(emu):
    P.check_mixed_order_buckets(emu[0], orders=(2, 3, 5), per_order=2)


def test_matrix_vector_consistency(emu):
    P.check_matrix_vector_consistency(emu[0], "stokes", 3, 3, n=3)
    P.check_matrix_vector_consistency(emu[0], "adv_dif", 2, 2, n=3)


def test_incidence(emu):
    P.check_incidence(emu[0])


def test_dense(emu):
    P.check_dense(emu[0])


def test_fields(emu):
    P.check_fields(emu[0])


def test_projectors(emu):
    P.check_projectors(emu[0])


def test_interior_products(emu):
    P.check_interior_products(emu[0], p=3, k=2)


def test_bytecode_errors(emu):
    P.check_bytecode_errors(emu[0])


def test_generic_chain(emu):
    P.check_generic_chain(emu[0], p=2, k=2)


@pytest.mark.parametrize("fix,nh,nv", [("solve_3x4_p4", 3, 4), ("solve_5x2_p5_def", 5, 2), ("solve_6x6_p3_def", 6, 6)])
def test_linear_system(emu, fix, nh, nv):
    P.check_linear_system(*emu, fix, nh, nv)


def test_saddle_residual_small_elements(emu):
    P.check_saddle_residual_small_elements(*emu, nh=6, nv=5, p=3)


@pytest.mark.parametrize("fixture,gate", [("hp_refined_345", 1e-10), ("hp_refined", 2e-7)])
def test_hp_refined(emu, fixture, gate):
    print(P.check_hp_refined(*emu, fixture, gate))


def test_time_march(emu):
    print(P.check_time_march(*emu))


def test_reconstruct_grid(emu):
    P.check_reconstruct_grid(emu[0])


def test_inverse_definite(emu):
    P.check_inverse_definite(emu[0])


def test_time_march_heat(emu):
    print(P.check_time_march_heat(*emu))


@pytest.mark.parametrize("name", ["reaction", "reaction_mixed", "vector_reaction", "heat_mixed"])
def test_time_march_reaction(emu, name):
    print(P.check_time_march_reaction(*emu, name))


def test_empty_and_unconstrained(emu):
    P.check_empty_and_unconstrained(emu[0], emu[1])
