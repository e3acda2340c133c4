"""world_size-2 tests of the multi-GPU path.

`-m "not gpu"`: two gloo processes run the partitioned Schur CG with numpy as the local engine
(tests/dist_worker.py, mode numpy) - this covers the host logic the N > 1 path adds (strip partition,
constraint partition, interface slots, ownership weights) against the reference's golden solution.
`-m gpu`: the same through libmfv2d_cuda.so with NCCL, one process per GPU; skipped on a one-GPU box.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WORKER = os.path.join(ROOT, "tests", "dist_worker.py")


def _run(mode, fix, nh, nv, world, port, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), WORKER, mode, fix, str(nh), str(nv)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env={**os.environ, **(env or {})})
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert "dist_worker ok" in res.stdout, res.stdout[-2000:]


def test_partition_is_consistent():
    from mfv2d_b200 import structured as ST

    nh, nv, p = 4, 6, 3
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, [2, 3])
    ne = 2 * p * (p + 1) + p * p
    rng = np.random.default_rng(0)
    x = rng.standard_normal(nh * nv * ne)
    full = np.zeros(nt)
    np.add.at(full, row, val * x[col])
    for world in (1, 2, 3):
        parts = ST.strip_partition(nh, nv, world)
        bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
        tot, cnt, slots = np.zeros(nt), np.zeros(nt), {}
        for r in range(world):
            P = ST.partition_constraints(row, col, val, nt, bounds, r)
            loc = np.zeros(P.n_trace)
            np.add.at(loc, P.row, P.val * x[bounds[r]:bounds[r + 1]][P.col])
            tot[P.global_rows] += loc
            cnt[P.global_rows] += P.owned
            for g, s in zip(P.global_rows[P.shared_index], P.slot):
                assert slots.setdefault(int(g), int(s)) == int(s)  # same slot on every rank
            assert np.all(np.diff(P.global_rows) > 0)
        assert np.array_equal(tot, full) or np.allclose(tot, full, rtol=0, atol=1e-13)
        assert np.all(cnt == 1)  # every multiplier has exactly one owner
        assert len(slots) == (world - 1) * nh * p  # p multipliers per edge on every cut (one 1-form)


@pytest.mark.parametrize("nh,nv,p,orders,world,mids,exact", [(6, 6, 3, [2, 3], 2, (2,), 3), (7, 9, 2, [2, 3], 3, (2,), 4), (5, 8, 4, [3, 2], 4, (2,), 2),
                                                             (8, 8, 3, [2, 3], 1, (2,), 4), (4, 12, 2, [2, 2, 3], 3, (2, 3), 4), (10, 5, 3, [2, 3], 8, (2,), 2)])
def test_strip_local_builder_equals_global(nh, nv, p, orders, world, mids, exact):
    """structured.strip_system_1form (closed form of the reference's row order, O(strip)) against the global path:
    partition_constraints of the reference-ordered global matrix + localize_hierarchy, field by field, every rank
    (including ranks with an empty strip)."""
    from mfv2d_b200 import structured as ST

    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, orders)
    ne = int(ST.form_sizes(orders, p).sum())
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    strip_rows = [parts[r][0] // nh for r in range(world)] + [nv]
    agg, zc, n_agg = ST.coarse_hierarchy_1form(nh, nv, p, orders, mids, exact, strip_rows)
    for rank in range(world):
        ref = ST.partition_constraints(row, col, val, nt, bounds, rank)
        rl, rz, rc = ST.localize_hierarchy(agg, zc, n_agg, ref)
        got, (gl, gz, gc), ntg = ST.strip_system_1form(nh, nv, p, orders, world, rank, mids, exact)
        assert ntg == nt and ref.n_trace == got.n_trace and ref.n_slots == got.n_slots
        for name in ("row", "col", "val", "global_rows", "shared_index", "slot", "owned", "holder_ptr", "holder_rank"):
            assert np.array_equal(np.asarray(getattr(ref, name)), np.asarray(getattr(got, name))), (rank, name)
        assert np.array_equal(rl, gl) and np.array_equal(rz, gz) and list(rc) == list(gc)


@pytest.mark.parametrize("fix,nh,nv", [("solve_3x4_p4", 3, 4), ("solve_6x6_p3_def", 6, 6)])
def test_gloo_world2_schur_cg(fix, nh, nv):
    _run("numpy", fix, nh, nv, 2, 29611)


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["peer", "peer-only", "nccl"])  # NVLink peer stores inside the CG kernels / the same without NCCL loaded / ncclAllReduce
@pytest.mark.parametrize("fix,nh,nv", [("solve_10x10_p3", 10, 10), ("solve_6x6_p3_def", 6, 6)])
def test_nccl_world2_schur_cg(fix, nh, nv, exchange):
    import mfv2d_b200._capi as capi

    if capi.lib().mfv2d_cuda_device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    env = {"MFV2D_CG_EXCHANGE": "nccl" if exchange == "nccl" else "peer", "MFV2D_TEST_NCCL": "0" if exchange == "peer-only" else "1"}
    _run("gpu", fix, nh, nv, 2, 29612 + ["peer", "peer-only", "nccl"].index(exchange), env=env)
