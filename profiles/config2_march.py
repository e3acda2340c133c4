"""BASELINE.json configs[2] at full size on one B200: unsteady linear advection-diffusion, n x n deformed elements of order p
(default 512 x 512, p = 5), trapezoidal time march with the device-backed callers (mfv2d_b200/callers.py).

    python profiles/config2_march.py [--n 512] [--order 5] [--steps 3]
    python -m torch.distributed.run --nproc-per-node N ... profiles/config2_march.py --gpus N      (strong scaling: N strips)

The operator is the time-march-updated system of tests/golden/timemarch_advdif.npz (bytecode written by the reference's own
CompiledSystem after update_system_for_time_march, dt = 0.05: q 1-form, u 2-form, nu = 0.05, the advection field of
tests/fields.py); constraints = the continuity rows of the structured mesh (natural boundary conditions); right-hand side =
dual DoFs of a smooth source on the u rows.  Every step: residual by one matrix-free apply over all elements, correction by
BiCGStab on the (non-symmetric) Schur complement, dual <- primal by the matrix-free mass apply.  Prints one JSON line:
set-up and per-step times, Krylov iterations, and the relative residual of the saddle-point system of the last step,
`||[b - A x - N^T l ; -N x]|| / ||b||`.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bench as B  # noqa: E402
import fields as F  # noqa: E402
from mfv2d_b200 import _mfv2d as M  # noqa: E402
from mfv2d_b200 import callers as CL  # noqa: E402
from mfv2d_b200 import solving as SV  # noqa: E402
from mfv2d_b200 import structured as ST  # noqa: E402
import mfv2d_b200.runtime as rt  # noqa: E402


def source(x, y):
    return 3.0 * np.cos(np.pi * x) * np.sin(0.7 * np.pi * y)


def decode(codes):
    out = []
    for rowc in codes:
        r = []
        for blk in rowc:
            if blk is None:
                r.append(None)
                continue
            ops = []
            for item in blk:
                item = list(item)
                if item[0] == 7 and isinstance(item[2], str) and item[2].startswith("fn:"):
                    item[2] = F.FIELDS[item[2][3:]]
                ops.append(tuple(item))
            r.append(tuple(ops))
        out.append(tuple(r))
    return tuple(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--n", type=int, default=512)
    ap.add_argument("--order", type=int, default=5)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--rtol", type=float, default=1e-12)
    ap.add_argument("--max-iter", type=int, default=40000)
    ap.add_argument("--nl-rtol", type=float, default=1e-10, help="stop the corrections of a step when max|residual| <= nl_rtol * max|rhs|")
    ap.add_argument("--nl-max", type=int, default=3)
    ap.add_argument("--precondition", default="jacobi", choices=["jacobi", "multilevel"])
    ap.add_argument("--agg", type=int, default=16)
    ap.add_argument("--mid", default="4")
    ap.add_argument("--device", action="store_true", help="keep every vector of the march in HBM (callers.trapezoidal_march(device=True))")
    args = ap.parse_args()
    g = np.load(os.path.join(ROOT, "tests", "golden", "timemarch_advdif.npz"), allow_pickle=True)
    raw = json.loads(str(g["codes"]))
    lhs = decode(raw["lhs"])
    dt, k = float(g["dt"]), int(g["k"])
    p, n = args.order, args.n
    specs = M._ElementFormSpecification(*zip(raw["form_names"], raw["form_orders"]))
    rank, world, local, dist = B.dist_setup(args.gpus)
    ctx = rt.Context(local)
    rt._context = ctx
    t0 = time.perf_counter()
    parts = ST.strip_partition(n, n, world)
    lo, hi = parts[rank]
    corners = np.ascontiguousarray(ST.unit_square_corners(n, n, F.deform)[lo:hi])
    rule = M.IntegrationRule1D(p + k)
    batch = M.ElementBatch(M.Basis2D(M.Basis1D(p, rule), M.Basis1D(p, rule)), corners, ctx)
    ne = specs.total_size(p, p)
    row, col, val, n_trace_global = ST.continuity_constraints(n, n, p, raw["form_orders"])
    if world > 1:
        bounds = [parts[r][0] * ne for r in range(world)] + [parts[-1][1] * ne]
        part = ST.partition_constraints(row, col, val, n_trace_global, bounds, rank)
        row, col, val, n_trace = part.row, part.col, part.val, part.n_trace
        owned = np.asarray(part.owned, bool)
    else:
        part, n_trace, owned = None, n_trace_global, np.ones(n_trace_global, bool)
    explicit = np.zeros((batch.n, ne))
    iu = raw["form_names"].index("u")
    off_u = specs.form_offset(iu, p, p)
    explicit[:, off_u:off_u + specs.form_size(iu, p, p)] = batch.dual_dofs(3, source)
    explicit = explicit.ravel()
    con_rhs = np.zeros(n_trace)
    host_s = B.max_over_ranks(dist, time.perf_counter() - t0)
    t0 = time.perf_counter()
    n_coarse = 0
    if args.precondition == "multilevel":
        mids = tuple(int(v) for v in args.mid.split(",") if v)
        _, hier, _ = ST.strip_system_1form(n, n, p, raw["form_orders"], world, rank, mids, args.agg)
        n_coarse = hier[2][-1]
    if world > 1:
        ctx.init_comm(world, rank, None)
    system = SV.compute_linear_system(specs, lhs, batch, (row, col, val), n_trace)
    if world > 1:
        system.set_interface(part.shared_index, part.slot, part.n_slots, part.owned)

        def all_gather(obj):
            outl = [None] * world
            dist.all_gather_object(outl, obj)
            return outl

        system.connect_peers(part.holder_ptr, part.holder_rank, all_gather, n_coarse=n_coarse)
    if args.precondition == "multilevel":
        system.set_coarse_space(*hier)
    ctx.synchronize()
    setup_s = B.max_over_ranks(dist, time.perf_counter() - t0)
    reduce_max = (lambda v: B.max_over_ranks(dist, v)) if world > 1 else None

    # the march (callers.trapezoidal_march builds the LinearSystem: element matrices stay on the device)
    stats = {"krylov_iterations": [], "krylov_residual": []}
    copy_s = [0.0]
    orig = SV.solve_schur_bicgstab

    def counting(system, rd, rt_, conv, precondition="jacobi"):
        x, lam, res, it = orig(system, rd, rt_, conv, precondition)
        stats["krylov_iterations"].append(int(it))
        stats["krylov_residual"].append(float(res))
        if isinstance(rd, rt.DeviceArray):  # the march reuses its residual buffers: copy now, and keep the copy out of the step time
            ctx.synchronize()
            tc = time.perf_counter()
            stats["last"] = (system,) + tuple(v.get().ravel() for v in (rd, rt_, x, lam))
            copy_s[0] += time.perf_counter() - tc
        else:
            stats["last"] = (system, np.array(rd), np.array(rt_), np.array(x), np.array(lam))
        return x, lam, res, it

    SV.solve_schur_bicgstab = counting
    step_times = []
    t_all = time.perf_counter()

    def timed_version(fn):
        def timed(*a, **kw):
            ctx.synchronize()
            t, c0 = time.perf_counter(), copy_s[0]
            out = fn(*a, **kw)
            ctx.synchronize()
            step_times.append(time.perf_counter() - t - (copy_s[0] - c0))
            return out
        return timed

    CL.non_linear_solve_run = timed_version(CL.non_linear_solve_run)
    CL.non_linear_solve_run_device = timed_version(CL.non_linear_solve_run_device)
    sols, iters = CL.trapezoidal_march(specs, lhs, None, batch, (row, col, val), n_trace, con_rhs, explicit, [iu], dt, args.steps,
                                       max_iterations=args.nl_max, rtol=args.nl_rtol, krylov=SV.ConvergenceSettings(args.max_iter, 1e300, args.rtol),
                                       system=system, reduce_max=reduce_max, precondition=args.precondition, device=args.device,
                                       keep=lambda step: step == args.steps - 1)
    sols = [sv for sv in sols if sv is not None]
    total_s = B.max_over_ranks(dist, time.perf_counter() - t_all - copy_s[0])
    assert not (world > 1 and system.peer_error())
    _, rd, rtr, x, lam = stats.pop("last")
    r1 = rd - system.apply_diagonal(x) - system.apply_trace_transpose(lam)
    r2 = (rtr - system.apply_trace(x))[owned]
    num = B.sum_over_ranks(dist, float(r1 @ r1 + r2 @ r2))
    den = B.sum_over_ranks(dist, float(rd @ rd + rtr[owned] @ rtr[owned]))
    rel = float(np.sqrt(num / den))
    sol_norms = [float(np.sqrt(B.sum_over_ranks(dist, float(sv @ sv)))) for sv in sols]
    step_times = [B.max_over_ranks(dist, t) for t in step_times]
    if rank == 0:
        B.emit({
        "config": f"BASELINE.json configs[2]: unsteady advection-diffusion (q1, u2), {n}x{n} deformed elements, order {p}, integration order p+{k}, "
                  f"trapezoidal march dt = {dt}, {args.steps} steps, {world} GPU(s), strong scaling over strips",
            "n_gpus": world, "elements": n * n, "element_dofs": ne, "dense_dofs": n * n * ne, "multipliers": int(n_trace_global),
            "host_mesh_constraints_rhs_s": host_s, "system_setup_s": setup_s, "march_total_s": total_s, "march_s_per_step": total_s / args.steps, "device_resident": bool(args.device),
            "step_s": step_times, "newton_iterations_per_step": [int(i) for i in iters], **stats,
            "last_correction_saddle_relative_residual": rel, "solution_norms": sol_norms,
            "method": f"BiCGStab on the Schur complement N A^-1 N^T (full condensed blocks, {args.precondition} right preconditioner), device resident; "
                      "interface sums and dot products over NVLink peer memory",
        })
    B.barrier(dist)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
