"""Worker of the world_size > 1 tests (launched with ``python -m torch.distributed.run``).

    dist_worker.py numpy <fixture> <nh> <nv>     CPU / gloo: the N > 1 ALGORITHM with numpy as the local engine
    dist_worker.py gpu   <fixture> <nh> <nv>     one process per GPU through libmfv2d_cuda.so + NCCL

Both modes split the elements of a golden mixed-Poisson system (tests/golden/solve_*.npz, produced by the
reference) into horizontal strips with ``structured.strip_partition`` / ``partition_constraints``, run the
Schur-complement CG with one interface sum per operator application and one all-reduced scalar per dot
product, and compare the gathered solution with the reference's.  The numpy mode mirrors, operation by
operation, what system.cu does on the device (pack shared multipliers -> all-reduce -> unpack; ownership-
weighted dots), with gloo standing in for NCCL, so that the host-side logic of the multi-GPU path (index maps,
slots, ownership) is covered in the GPU-less container.
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    mode, fix, nh, nv = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
    import torch
    import torch.distributed as dist

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group(backend="gloo")
    rank, world = dist.get_rank(), dist.get_world_size()

    from conftest import golden, load_systems, relerr
    from mfv2d_b200 import structured as ST

    g = golden(fix + ".npz")
    p, k = int(g["order"]), int(g["k"])
    s = load_systems()["mixed_poisson_sorted"]
    ne = int(ST.form_sizes(s["form_orders"], p).sum())
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    parts = ST.strip_partition(nh, nv, world)
    bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
    P = ST.partition_constraints(row, col, val, nt, bounds, rank)
    lo, hi = parts[rank]
    rhs_l = np.ascontiguousarray(g["rhs"][bounds[rank]:bounds[rank + 1]])
    phi_l = np.ascontiguousarray(g["con_rhs"][P.global_rows])

    def allreduce(a: np.ndarray) -> np.ndarray:
        t = torch.from_numpy(np.ascontiguousarray(a, np.float64).copy())
        dist.all_reduce(t)
        return t.numpy()

    if mode == "numpy":
        from oracle import mfv2d_oracle as O

        ob = O.basis(p, p + k + 1)
        blocks, fs = O.decode_expressions(s["lhs"], s["form_names"])
        mats = [O.element_matrix(ob, g["corners"][e], s["form_orders"], blocks, fs) for e in range(lo, hi)]
        nd = (hi - lo) * ne
        A = np.zeros((nd, nd))
        for i, m in enumerate(mats):
            A[i * ne:(i + 1) * ne, i * ne:(i + 1) * ne] = m
        Ainv = np.linalg.inv(A)
        N = np.zeros((P.n_trace, nd))
        N[P.row, P.col] = P.val

        def iface_sum(t):  # iface_pack_kernel -> ncclAllReduce -> iface_unpack_kernel
            buf = np.zeros(P.n_slots)
            buf[P.slot] = t[P.shared_index]
            buf = allreduce(buf)
            t = t.copy()
            t[P.shared_index] = buf[P.slot]
            return t

        def dot(a, b):  # ownership-weighted partial + all-reduce of one scalar
            return float(allreduce(np.array([np.sum(P.owned * a * b)]))[0])

        def schur_and_dot(t):
            # system.cu, N > 1: the local un-summed product S_r t and t . (S_r t) travel in ONE all-reduce
            # (t . S t = sum_r t_r . S_r t_r because shared entries of t are identical on both ranks)
            loc = N @ (Ainv @ (N.T @ t))
            buf = np.zeros(P.n_slots + 1)
            buf[P.slot] = loc[P.shared_index]
            buf[-1] = t @ loc
            buf = allreduce(buf)
            loc[P.shared_index] = buf[P.slot]
            return loc, float(buf[-1])

        inv_a_y = Ainv @ rhs_l
        r = iface_sum(N @ inv_a_y) - phi_l
        lam = np.zeros(P.n_trace)
        rr = dot(r, r)
        tol2 = (min(1e-30, np.sqrt(rr) * 1e-12)) ** 2
        d = r.copy()
        it = 0
        while it < 2000 and rr >= tol2:
            ad, dad = schur_and_dot(d)
            assert abs(dad - dot(d, ad)) <= 1e-9 * abs(dad)  # same number as the ownership-weighted dot of the summed product
            alpha = rr / dad
            lam += alpha * d
            r -= alpha * ad
            rr_new = dot(r, r)
            d = r + (rr_new / rr) * d
            rr = rr_new
            it += 1
        x = inv_a_y - Ainv @ (N.T @ lam)
    else:
        from mfv2d_b200 import _mfv2d as M
        from mfv2d_b200 import solving as SV
        from mfv2d_b200.runtime import Context, comm_unique_id
        import mfv2d_b200.runtime as rt
        import parity_cases as PC

        ctx = Context(int(os.environ.get("LOCAL_RANK", "0")))
        rt._context = ctx
        # MFV2D_TEST_NCCL=0: ranks only - every exchange goes through the NVLink peer blocks, NCCL is never loaded
        ids = [comm_unique_id() if rank == 0 and os.environ.get("MFV2D_TEST_NCCL", "1") != "0" else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.init_comm(world, rank, ids[0])

        def all_gather(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out

        it, itj, _, worst = PC.check_partitioned_solve(M, SV, ST, ctx, rank, world, all_gather, fix, nh, nv)
        itj2, it2 = PC.check_multilevel(M, SV, ST, ctx, rank, world, all_gather, nh=32, nv=32, p=3, k=3, mid_sizes=(2,), exact_size=8)
        itb, errb = PC.check_partitioned_bicgstab(M, SV, ST, ctx, rank, world, all_gather)
        if rank == 0:
            print(f"dist_worker ok: mode {mode}, world {world}, iterations {it} (jacobi {itj}), err {worst:.2e}; "
                  f"multilevel {it2} vs jacobi {itj2} on 32x32 p3; BiCGStab (advection-diffusion) {itb} iterations, err {errb:.1e}; "
                  f"exchange {os.environ.get('MFV2D_CG_EXCHANGE', 'peer')}", flush=True)
        dist.barrier()
        dist.destroy_process_group()
        return

    err_x = relerr(x, g["sol_dense"][bounds[rank]:bounds[rank + 1]])
    err_l = relerr(lam, g["sol_trace"][P.global_rows])
    worst = float(allreduce(np.array([max(err_x, err_l)]) / world)[0])  # not a max, but fails if any rank is off
    assert err_x < 1e-8 and err_l < 1e-8, (rank, err_x, err_l)
    assert abs(it - int(g["iters"])) <= 12, (it, int(g["iters"]))
    if rank == 0:
        print(f"dist_worker ok: mode {mode}, world {world}, iterations {it}, err {worst:.2e}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
