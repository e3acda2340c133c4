"""BASELINE.json configs[4]: synthetic 1000 x 1000 curvilinear quad mesh, mixed Poisson, order 6 — assembly,
matrix-free apply and the element-by-element Krylov solve (CG on the Schur complement of the multipliers,
SURVEY.md §3.3) across the GPUs of one box.  STRONG scaling: the mesh is fixed, rank r owns the r-th strip.

    torchrun --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
        profiles/config4_solve.py --gpus 8                   # the full 1M-element mesh (needs >= 4 B200: 115 GB of A_e)
    python profiles/config4_solve.py --gpus 1 --rows 125     # one rank's share of it on a single GPU

Prints one JSON line on rank 0.  Not the judged bench (bench.py is); evidence for the configs[4] row of DESIGN.md §7.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench as B  # noqa: E402


def source(x, y):
    return 2.0 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--cols", type=int, default=1000)
    ap.add_argument("--rows", type=int, default=1000)
    ap.add_argument("--order", type=int, default=6)
    ap.add_argument("--rtol", type=float, default=1e-10)
    ap.add_argument("--max-iter", type=int, default=60000)
    ap.add_argument("--precondition", default="none", choices=["none", "jacobi", "multilevel"])
    ap.add_argument("--agg", type=int, default=25, help="aggregate size (elements per side) of the exact coarse level")
    ap.add_argument("--mid", default="5", help="comma-separated aggregate sizes of the diagonal (rank-local) levels")
    ap.add_argument("--source", default="smooth", choices=["smooth", "random"])
    ap.add_argument("--repeat", type=int, default=2)
    args = ap.parse_args()
    rank, world, local, dist = B.dist_setup(args.gpus)
    from mfv2d_b200 import _mfv2d as M
    from mfv2d_b200 import solving as SV
    from mfv2d_b200 import structured as ST
    import mfv2d_b200.runtime as rt

    ctx = rt.Context(local)
    rt._context = ctx
    names, orders, lhs = B.load_system("mixed_poisson")
    specs = M._ElementFormSpecification(*zip(names, orders))
    p, k = args.order, 3
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + k))
    basis = M.Basis2D(b1, b1)
    nh, nv = args.cols, args.rows
    t_host = time.perf_counter()
    full = ST.unit_square_corners(nh, nv, B.deform)
    parts = ST.strip_partition(nh, nv, world)
    lo, hi = parts[rank]
    corners = np.ascontiguousarray(full[lo:hi])
    del full
    n = corners.shape[0]
    ne = specs.total_size(p, p)
    batch = M.ElementBatch(basis, ctx.to_device(corners), ctx)
    peak, _ = B.measured_peaks()

    # ---- assembly: this rank's element matrices, kept on the device for the system ----
    mats = ctx.empty((n, ne, ne))
    M.compute_element_matrices(specs, lhs, batch, out=mats)
    ctx.synchronize()
    B.barrier(dist)
    ctx.timer_start()
    reps = 3
    for _ in range(reps):
        M.compute_element_matrices(specs, lhs, batch, out=mats)
    asm_ms = max(B.max_over_ranks(dist, ctx.timer_stop() / reps), 1e-9)
    total = B.sum_over_ranks(dist, float(n))

    # ---- matrix-free apply (recompute) ----
    rng = np.random.default_rng(rank)
    x = ctx.to_device(rng.standard_normal((n, ne)))
    y = ctx.empty((n, ne))
    M.compute_element_vectors(specs, lhs, batch, x, out=y)
    ctx.synchronize()
    B.barrier(dist)
    ctx.timer_start()
    for _ in range(reps):
        M.compute_element_vectors(specs, lhs, batch, x, out=y)
    app_ms = max(B.max_over_ranks(dist, ctx.timer_stop() / reps), 1e-9)
    del x, y

    # ---- constraints of the GLOBAL mesh, this rank's part ----
    mids = tuple(int(v) for v in args.mid.split(",") if v)
    part, hier, nt_glob = ST.strip_system_1form(nh, nv, p, orders, world, rank, mids, args.agg)  # O(strip): closed form of the reference's row order
    if world > 1:
        ids = [rt.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.init_comm(world, rank, ids[0])
    host_s = time.perf_counter() - t_host
    t0 = time.perf_counter()
    system = SV.LinearSystem(mats, (part.row, part.col, part.val), part.n_trace, ctx)
    n_coarse = 0
    if args.precondition == "multilevel":
        n_coarse = hier[2][-1]
    if world > 1:
        system.set_interface(part.shared_index, part.slot, part.n_slots, part.owned)

        def all_gather(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out

        system.connect_peers(part.holder_ptr, part.holder_rank, all_gather, n_coarse=n_coarse)
    ctx.synchronize()
    setup_s = B.max_over_ranks(dist, time.perf_counter() - t0)
    coarse_s = 0.0
    if n_coarse:
        t0 = time.perf_counter()
        system.set_coarse_space(*hier)
        ctx.synchronize()
        coarse_s = B.max_over_ranks(dist, time.perf_counter() - t0)

    # ---- right-hand side: dual DoFs of the source on the 2-form rows, zero on the 1-form rows ----
    rhs_h = np.zeros((n, ne))
    off_u = specs.form_offset(1, p, p) if hasattr(specs, "form_offset") else ne - p * p
    if args.source == "random":
        rhs_h[:] = np.random.default_rng(100 + rank).standard_normal(rhs_h.shape)
    else:
        f = batch.dual_dofs(orders[1], source)
        rhs_h[:, off_u:] = f.get() if hasattr(f, "get") else f
    rhs = ctx.to_device(rhs_h.ravel())
    phi = ctx.to_device(np.zeros(part.n_trace))
    solve_times = []
    for _ in range(args.repeat):  # the first solve pays NCCL's lazy set-up and the graph instantiation
        B.barrier(dist)
        ctx.synchronize()
        if world > 1:
            system.peer_stats()  # reset the timeline counters (MFV2D_PEER_DEBUG bit 8)
        t0 = time.perf_counter()
        xs, lam, res, it = SV.solve_schur_iterative(system, rhs, phi, SV.ConvergenceSettings(args.max_iter, 1e300, args.rtol),
                                                    precondition=args.precondition)
        assert not system.peer_error()
        ctx.synchronize()
        solve_times.append(B.max_over_ranks(dist, time.perf_counter() - t0))
    solve_s = solve_times[-1]
    timeline = None
    if world > 1 and int(os.environ.get("MFV2D_PEER_DEBUG", "0")) & 8:
        mine = (system.peer_stats()[:10].astype(np.float64) / max(int(it), 1) / 1e3).round(2).tolist()
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        timeline = {"us_per_iteration_by_stamp": gathered,
                    "stamps": "0 product+tail | 1 gather+publish | 2 gap | 3 wait peers | 4 update (CTA 0) | 5 gap+chunk sums | 6 aggregate+publish | "
                              "7 gap | 8 wait coarse | 9 coarse solve (CTA 0); one row per rank"}

    # ---- residual of the saddle-point system [[A, N^T], [N, 0]] (x, lam) = (rhs, 0), summed over the ranks ----
    r1 = system.apply_diagonal(xs).get() + system.apply_trace_transpose(lam).get() - rhs_h.ravel()
    r2 = system.apply_trace(xs).get()
    own = np.asarray(part.owned, bool) if world > 1 else np.ones(part.n_trace, bool)
    num = B.sum_over_ranks(dist, float(r1 @ r1 + r2[own] @ r2[own]))
    den = B.sum_over_ranks(dist, float(rhs_h.ravel() @ rhs_h.ravel()))
    line = {
        "config": f"BASELINE.json configs[4]: mixed Poisson (q1, u2) on a deformed {nh}x{nv} quad mesh, order {p}, integration order p+3, "
                  f"{world} strip(s), strong scaling",
        "n_gpus": world, "elements": int(total), "element_dofs": ne, "dense_dofs": int(total) * ne, "multipliers": int(nt_glob),
        "interface_multipliers": int(part.n_slots),
        "assembly": {"ms": asm_ms, "element_matrices_per_s": total / (asm_ms * 1e-3),
                     "hbm_frac_per_gpu": (64 + 8 * ne * ne) * n / (asm_ms * 1e-3) / 1e9 / peak},
        "apply_recompute": {"ms": app_ms, "gdofs": total * ne / (app_ms * 1e-3) / 1e9},
        "system_setup_s": setup_s, "coarse_setup_s": coarse_s, "n_coarse": int(n_coarse), "host_mesh_and_constraints_s": host_s,
        "solve": {"method": "CG on the Schur complement N A^-1 N^T" + f", preconditioner {args.precondition}", "rtol": args.rtol, "s": solve_s, "s_all_solves": solve_times,
                  "exchange": os.environ.get("MFV2D_CG_EXCHANGE", "peer" if world > 1 else "-"), "graph": os.environ.get("MFV2D_CG_GRAPH", "default"),
                  "iterations": int(it), "us_per_iteration": solve_s / max(int(it), 1) * 1e6, "cg_residual": float(res),
                  "saddle_relative_residual": float(np.sqrt(num / den)), "peer_debug": os.environ.get("MFV2D_PEER_DEBUG", "0"),
                  "timeline": timeline},
    }
    if rank == 0:
        B.emit(line)
    B.barrier(dist)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
