"""`-m gpu`: parity of the CUDA path (libmfv2d_cuda.so on cuda:0, called through the C-ABI by the host
mirror) against the oracle and the reference's golden outputs, plus size-independent properties at the
BASELINE.json sizes.  Nothing here reads /root/reference."""
import numpy as np
import pytest

import parity_cases as P

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu():
    import mfv2d_b200._capi as capi

    capi._use_library_for_testing(capi.LIB_PATH)
    from mfv2d_b200 import _mfv2d as M
    from mfv2d_b200 import solving as SV
    from mfv2d_b200 import structured as ST
    from mfv2d_b200.runtime import get_context

    ctx = get_context()  # raises loudly without a B200 / without the built library
    assert ctx.handle
    return M, SV, ST


def test_tables(gpu):
    P.check_tables(gpu[0])


def test_masses_golden(gpu):
    P.check_masses_golden(gpu[0])


@pytest.mark.parametrize("k", [0, 2, 3])
def test_masses_oracle(gpu, k):
    P.check_masses_oracle(gpu[0], cases=[(p, k) for p in range(1, 9)], n_random=256)


@pytest.mark.parametrize("name", sorted(P.SYSTEMS))
def test_system_golden(gpu, name):
    P.check_system_golden(gpu[0], name)


@pytest.mark.parametrize("name,p,k", [("stokes", 4, 3), ("mixed_poisson", 6, 3), ("adv_dif", 5, 3), ("direct_poisson", 8, 3),
                                      ("navier_stokes", 3, 3), ("heat_mixed", 7, 2), ("stokes", 8, 3)])
def test_matrix_vector_consistency(gpu, name, p, k):
    P.check_matrix_vector_consistency(gpu[0], name, p, k, n=300)


@pytest.mark.parametrize("name", ["mixed_poisson", "stokes"])
def test_mixed_order_buckets(gpu, name):
    P.check_mixed_order_buckets(gpu[0], orders=(2, 3, 4, 5, 6, 7, 8), per_order=9, name=name)


@pytest.mark.parametrize("mesh,gate,sol_tol", [("b", 1e-10, 1e-7), ("a", 2e-7, 1e-5)])
def test_hp_forest(gpu, mesh, gate, sol_tol):
    """configs[3] from arrays alone: own quadtree, own hanging-side rows, mixed block sizes, Schur CG.  Mesh a (93 leaves,
    orders 2..8) has hanging-side coefficients up to 4e8: gates as for the reference-generated twin of it (test_hp_refined)"""
    print(P.check_hp_forest(*gpu, mesh, gate, sol_tol))


def test_hp_forest_boundary_conditions(gpu):
    """the same solve with boundary conditions built from arrays: strong flux rows on half of the boundary (single-entry rows
    next to the hanging-side rows), natural condition on the rest"""
    print(P.check_hp_forest(*gpu, "b", 1e-10, 1e-7, bc=True))


def test_degenerate_element(gpu):
    P.check_degenerate_element(gpu[0])


def test_dual_dofs(gpu):
    P.check_dual_dofs(gpu[0])


def test_incidence(gpu):
    P.check_incidence(gpu[0], orders=(1, 2, 3, 4, 5, 8))


def test_dense(gpu):
    P.check_dense(gpu[0])


def test_fields(gpu):
    P.check_fields(gpu[0])


def test_projectors(gpu):
    P.check_projectors(gpu[0])


def test_interior_products(gpu):
    P.check_interior_products(gpu[0], p=4, k=2)
    P.check_interior_products(gpu[0], p=6, k=3)


def test_bytecode_errors(gpu):
    P.check_bytecode_errors(gpu[0])


def test_generic_chain(gpu):
    P.check_generic_chain(gpu[0], p=3, k=2)


@pytest.mark.parametrize("fix,nh,nv", [("solve_10x10_p3", 10, 10), ("solve_3x4_p4", 3, 4), ("solve_5x2_p5_def", 5, 2),
                                       ("solve_6x6_p3_def", 6, 6)])
def test_linear_system(gpu, fix, nh, nv):
    P.check_linear_system(*gpu, fix, nh, nv)


def test_saddle_residual_small_elements(gpu):
    P.check_saddle_residual_small_elements(*gpu, nh=6, nv=5, p=6)


@pytest.mark.parametrize("fixture,gate", [("hp_refined_345", 1e-10), ("hp_refined", 2e-7)])
def test_hp_refined(gpu, fixture, gate):
    """BASELINE.json configs[3]: non-conforming h/p-refined mesh, blocks of several sizes in one LinearSystem"""
    P.check_hp_refined(*gpu, fixture, gate)


def test_reconstruct_grid(gpu):
    P.check_reconstruct_grid(gpu[0])


def test_apply_warp_kernel(gpu, monkeypatch):
    """the warp-autonomous apply kernel (apply_warp.cuh; opt-in, slower than the CTA-phase kernel on B200) stays parity-green"""
    monkeypatch.setenv("MFV2D_APPLY_WARP", "1")
    P.check_matrix_vector_consistency(gpu[0], "mixed_poisson", 6, 3)
    P.check_matrix_vector_consistency(gpu[0], "stokes", 4, 3)
    P.check_matrix_vector_consistency(gpu[0], "direct_poisson", 5, 3)


def test_apply_dmma_kernel(gpu, monkeypatch):
    """the fp64 tensor-core apply kernel (q <= 8; opt-in since the DFMA kernel overtook it) stays parity-green"""
    monkeypatch.setenv("MFV2D_APPLY_DMMA", "1")
    P.check_matrix_vector_consistency(gpu[0], "stokes", 4, 3)
    P.check_matrix_vector_consistency(gpu[0], "mixed_poisson", 4, 3)
    P.check_matrix_vector_consistency(gpu[0], "stokes", 3, 3)


def test_time_march(gpu):
    """BASELINE.json configs[2] (slice): trapezoidal march of unsteady advection-diffusion against the reference's own march"""
    P.check_time_march(*gpu)


def test_empty_batches(gpu):
    M = gpu[0]
    b = P.make_basis(M, 3, 3)
    batch = M.ElementBatch(b, np.zeros((0, 4, 2)))
    assert batch.mass(2).shape == (0, 24, 24)
    s = P.SYSTEMS["mixed_poisson"]
    assert M.compute_element_matrices(P.specs_of(M, s), s["lhs"], batch).shape == (0, 33, 33)


def test_full_size_properties(gpu):
    """Config 2 (Stokes 256 x 256, p = 4) at full size, device resident: symmetry of the assembled blocks,
    matrix @ x == matrix-free vector, and agreement of a strided sample with the oracle."""
    import fields as F
    from oracle import mfv2d_oracle as O

    M, SV, ST = gpu
    from mfv2d_b200.runtime import get_context

    ctx = get_context()
    p, k, n_side = 4, 3, 256
    s = P.SYSTEMS["stokes"]
    specs = P.specs_of(M, s)
    corners = ST.unit_square_corners(n_side, n_side, F.deform)
    batch = M.ElementBatch(P.make_basis(M, p, k), corners)
    ne = specs.total_size(p, p)
    out = ctx.empty((batch.n, ne, ne))
    M.compute_element_matrices(specs, s["lhs"], batch, out=out)
    mats = out.get()
    assert np.array_equal(mats, np.swapaxes(mats, 1, 2)) or P.relerr(mats, np.swapaxes(mats, 1, 2)) < 1e-13
    rng = np.random.default_rng(0)
    x = rng.standard_normal((batch.n, ne))
    y = M.compute_element_vectors(specs, s["lhs"], batch, x)
    assert P.relerr(np.einsum("eij,ej->ei", mats, x), y) < 1e-12
    # the fused kernels hand data between phases through shared memory and asynchronous copies: a missing barrier
    # would show up as run-to-run differences (compute-sanitizer is not available on this pool)
    for _ in range(3):
        assert np.array_equal(M.compute_element_vectors(specs, s["lhs"], batch, x), y)
    out2 = ctx.empty((batch.n, ne, ne))
    M.compute_element_matrices(specs, s["lhs"], batch, out=out2)
    assert np.array_equal(out2.get(), mats)
    ob = O.basis(p, p + k + 1)
    blocks, fs = O.decode_expressions(s["lhs"], s["form_names"])
    for e in range(0, batch.n, 4099):
        assert P.relerr(mats[e], O.element_matrix(ob, corners[e], s["form_orders"], blocks, fs)) < 1e-12


def test_inverse_definite(gpu):
    P.check_inverse_definite(gpu[0])


def test_time_march_heat(gpu):
    print(P.check_time_march_heat(*gpu))


@pytest.mark.parametrize("name", ["reaction", "reaction_mixed", "vector_reaction", "heat_mixed"])
def test_time_march_reaction(gpu, name):
    print(P.check_time_march_reaction(*gpu, name))


def test_empty_and_unconstrained(gpu):
    P.check_empty_and_unconstrained(gpu[0], gpu[1])
