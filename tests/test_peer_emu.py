"""`-m "not gpu"`: the NVLink peer exchange (mfv2d_b200/csrc/peer.cu, the fused CG kernels of schur_cg.cu and the coarse
space of precond.cu) executed through the host emulation with one THREAD per rank.  In the emulation an "IPC handle" is
the pointer itself and a peer store is a plain store, so this covers the protocol (epochs, parities, flags, holder
lists, rank-ordered sums, the two exchange types) and the host logic - not the memory model of the real device, which
the `-m gpu` world-2 tests (tests/test_distributed.py) exercise on two B200s."""
import os
import sys
import threading

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


class ThreadGather:
    """all_gather between the threads of one process"""

    def __init__(self, world):
        self.world, self.barrier, self.slots = world, threading.Barrier(world), [None] * world

    def bind(self, rank):
        def gather(obj):
            self.slots[rank] = obj
            self.barrier.wait(timeout=120)
            out = list(self.slots)
            self.barrier.wait(timeout=120)
            return out

        return gather


def run_ranks(emu, world, check, *args):
    M, SV, ST = emu
    from mfv2d_b200.runtime import Context

    tg = ThreadGather(world)
    results, errors = [None] * world, []

    def work(rank):
        try:
            ctx = Context(0)
            if world > 1:
                ctx.init_comm(world, rank, None)  # ranks only: no NCCL anywhere
            results[rank] = check(M, SV, ST, ctx, rank, world, tg.bind(rank), *args)
        except BaseException as exc:  # noqa: BLE001
            errors.append((rank, exc))
            tg.barrier.abort()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=600)
    assert not errors, errors
    return results


@pytest.mark.parametrize("world", [1, 2, 3])
def test_peer_exchange_schur_cg(emu, world):
    run_ranks(emu, world, P.check_partitioned_solve, "solve_6x6_p3_def", 6, 6)


@pytest.mark.parametrize("world", [1, 2, 3])
def test_multilevel(emu, world):
    res = run_ranks(emu, world, P.check_multilevel)
    assert all(r == res[0] for r in res)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_peer_exchange_bicgstab(emu, world):
    """BiCGStab on a non-symmetric (advection-diffusion) system, partitioned into strips, against a direct sparse solve"""
    res = run_ranks(emu, world, P.check_partitioned_bicgstab)
    assert all(r[0] == res[0][0] for r in res)


@pytest.mark.parametrize("world", [1, 2, 4])
def test_partitioned_time_march(emu, world):
    """the trapezoidal march on strips (host vectors and device-resident vectors), every strip against the reference's march"""
    res = run_ranks(emu, world, P.check_partitioned_march)
    assert all(r == res[0] for r in res)
