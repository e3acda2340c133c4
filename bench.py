#!/usr/bin/env python
"""Benchmark of the element hot path (BASELINE.json metric: element matrices/sec, fp64).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: the assembly of ALL element matrices of the
workload (config 2 of BASELINE.json: steady Stokes, 256 x 256 elements, order 4, integration order p + 3,
sinusoidally deformed unit square, N_e = 81).  With N > 1 (torchrun, one process per GPU) every rank
assembles its own copy-sized partition of an N-times larger mesh ("weak" scaling); assembly needs no
collective.

`value`     element matrices/s, whole job, inputs (corners) and outputs resident in HBM, CUDA events.
`e2e`       the same through the C-ABI with HOST buffers: corners copied H2D and the element matrices
            copied D2H to pinned memory inside the timed region.
`roofline`  HBM-write roofline of the assembly kernel: (64 + 8 N_e^2) bytes per element / kernel time.
`cpu_baseline`  the reference's own C path (oracle/_ref, built from /root/reference by oracle/build_ref.sh)
            or the numpy oracle port, timed on a bounded sample on this box's host cores.
`extra`     matrix-free apply GDoF/s (recompute and stored-block variants) and Schur-CG solve seconds.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

P_ORDER, K_OVER, N_SIDE = 4, 3, 256
SYSTEM = "stokes"


def load_system(name):
    with open(os.path.join(ROOT, "tests", "golden", "systems.json")) as f:
        raw = json.load(f)[name]

    def dec(codes):
        return tuple(tuple(None if b is None else tuple(tuple(i) for i in b) for b in row) for row in codes)

    return raw["form_names"], raw["form_orders"], dec(raw["lhs"])


def deform(xi, eta):
    return (xi + 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta), eta - 0.1 * np.sin(np.pi * xi) * np.sin(np.pi * eta))


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,utilization.gpu")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [x.strip() for x in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                clk, mxc, util = float(parts[0]), float(parts[1]), float(parts[6])
            except ValueError:
                continue
            mx = mxc
            if util >= 50:
                sm.append(clk)
            for nm, val in zip(names, parts[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        allc = sm or [float(x.split(",")[0]) for x in self.lines if x.split(",")[0].strip().replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(allc)) if allc else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples_under_load": len(sm)}


def dist_setup(n_gpus):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist_mod  # plumbing only: barrier and max over ranks on host scalars

        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist_mod.init_process_group(backend="gloo", rank=rank, world_size=world)
        dist = dist_mod
    return rank, world, local, dist


def barrier(dist):
    if dist is not None:
        dist.barrier()


def max_over_ranks(dist, x: float) -> float:
    if dist is None:
        return x
    import torch

    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def sum_over_ranks(dist, x: float) -> float:
    if dist is None:
        return x
    import torch

    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t[0])


# ------------------------------------------------------------------------------------------------
# CPU baseline: the reference's own C path (or the numpy oracle port) on a bounded sample
# ------------------------------------------------------------------------------------------------
def _cpu_worker(args):
    kind, corners, seconds = args
    names, orders, lhs = load_system(SYSTEM)
    t_end = time.perf_counter() + seconds
    done = 0
    t0 = time.perf_counter()
    if kind == "reference":
        from oracle.ref_loader import load_ref

        ref = load_ref()
        rule = ref.IntegrationRule1D(P_ORDER + K_OVER)
        b1 = ref.Basis1D(P_ORDER, rule)
        b2 = ref.Basis2D(b1, b1)
        specs = ref._ElementFormSpecification(*zip(names, orders))
        for c in corners:
            sp = ref.ElementFemSpace2D(b2, np.ascontiguousarray(c))
            ref.compute_element_matrix(specs, lhs, sp)
            done += 1
            if time.perf_counter() > t_end:
                break
    else:
        from oracle import mfv2d_oracle as O

        ob = O.basis(P_ORDER, P_ORDER + K_OVER + 1)
        blocks, fs = O.decode_expressions(lhs, names)
        for c in corners:
            O.element_matrix(ob, c, orders, blocks, fs)
            done += 1
            if time.perf_counter() > t_end:
                break
    return done, time.perf_counter() - t0


def cpu_baseline(cores: int, seconds: float):
    from oracle.ref_loader import load_ref

    kind = "reference" if load_ref() is not None else "port"
    from mfv2d_b200.structured import unit_square_corners

    corners = unit_square_corners(N_SIDE, N_SIDE, deform)
    rng = np.random.default_rng(0)
    per = 20000
    chunks = [corners[rng.choice(len(corners), per)] for _ in range(cores)]
    if cores == 1:
        res = [_cpu_worker((kind, chunks[0], seconds))]
    else:
        import multiprocessing as mp

        with mp.get_context("spawn").Pool(cores) as pool:
            res = pool.map(_cpu_worker, [(kind, c, seconds) for c in chunks])
    n = sum(r[0] for r in res)
    t = max(r[1] for r in res)
    return {"value": n / t, "unit": "element matrices/s", "cores": cores, "kind": kind,
            "sample": f"{n} randomly drawn elements of the {N_SIDE}x{N_SIDE} p{P_ORDER} Stokes mesh, {t:.1f} s wall on {cores} process(es), "
                      "per-element cost is independent of mesh size"}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation with all host cores, bounded sample per step."""
    rank, world, local, dist = dist_setup(args.gpus)
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    per_step = max(2.0, min(20.0, 90.0 / max(args.steps + args.warmup, 1)))
    vals = []
    for i in range(args.warmup + args.steps):
        r = cpu_baseline(cores, per_step)
        if i >= args.warmup:
            vals.append(r)
    v = float(np.mean([r["value"] for r in vals]))
    n_elem = N_SIDE * N_SIDE
    line = {
        "impl": "reference", "metric": "element_matrices_per_s", "value": v, "unit": "element matrices/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * n_elem * max(args.gpus, 1) / v, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.gpus),
        "cpu_baseline": {"value": v, "unit": "element matrices/s", "cores": cores, "kind": vals[-1]["kind"], "sample": vals[-1]["sample"]},
        "e2e": {"value": v, "unit": "element matrices/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def workload_config(n_gpus):
    return {"workload": f"steady {'Stokes (vor0, vel1, prs2)' if SYSTEM == 'stokes' else 'mixed Poisson (q1, u2)'} element matrices, {N_SIDE}x{N_SIDE} elements per GPU, order {P_ORDER}, "
                        f"integration order p+{K_OVER}, deformed unit square " + ("(BASELINE.json configs[1])" if SYSTEM == "stokes" else "(per-GPU slice of BASELINE.json configs[4])"),
            "elements_per_gpu": N_SIDE * N_SIDE, "element_dofs": {"stokes": 81, "mixed_poisson": 120}.get(SYSTEM, 0) if P_ORDER in (4, 6) else 0, "parallelism": f"elements x{n_gpus}, no collective",
            "l2": "outputs (3.4 GB per step or more) are larger than L2; no flush needed"}


def run_ours(args):
    rank, world, local, dist = dist_setup(args.gpus)
    from mfv2d_b200 import _mfv2d as M
    from mfv2d_b200 import solving as SV
    from mfv2d_b200 import structured as ST
    from mfv2d_b200.runtime import Context

    ctx = Context(local)
    import mfv2d_b200.runtime as rt

    rt._context = ctx
    names, orders, lhs = load_system(SYSTEM)
    specs = M._ElementFormSpecification(*zip(names, orders))
    p, k = P_ORDER, K_OVER
    b1 = M.Basis1D(p, M.IntegrationRule1D(p + k))
    basis = M.Basis2D(b1, b1)
    # weak scaling: rank r owns the r-th horizontal strip of a (N_SIDE) x (N_SIDE * world) mesh
    full = ST.unit_square_corners(N_SIDE, N_SIDE * world, deform)
    lo, hi = ST.strip_partition(N_SIDE, N_SIDE * world, world)[rank]
    corners = np.ascontiguousarray(full[lo:hi])
    n = corners.shape[0]
    ne = specs.total_size(p, p)
    d_corners = ctx.to_device(corners)
    batch = M.ElementBatch(basis, d_corners, ctx)
    out = ctx.empty((n, ne, ne))

    def step_device():
        M.compute_element_matrices(specs, lhs, batch, out=out)

    for _ in range(args.warmup):
        step_device()
    ctx.synchronize()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier(dist)
    ctx.synchronize()
    l0 = ctx.launch_count
    ctx.timer_start()
    for _ in range(args.steps):
        step_device()
    ms = ctx.timer_stop()
    ctx.synchronize()
    barrier(dist)
    launches = ctx.launch_count - l0
    clocks = None
    ms = max_over_ranks(dist, ms)
    total_elems = sum_over_ranks(dist, float(n))
    ms_per_step = ms / args.steps
    value = total_elems / (ms_per_step * 1e-3)

    # per-launch kernel duration (the step is exactly one kernel launch): roofline of the assembly kernel
    kern_ms = ms_per_step
    alg_bytes = (64 + 8 * ne * ne) * n
    peak, peak_src = measured_peaks()
    achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "element_matrix_kernel", "peak_source": peak_src,
                "algorithmic_bytes_per_element": 64 + 8 * ne * ne}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        with open(tr_path) as f:
            roofline["traffic"] = json.load(f).get("element_matrix_kernel_bytes_per_launch")

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region ----
    h_corners = ctx.pinned((n, 4, 2))
    h_corners[:] = corners
    h_out = ctx.pinned((n, ne, ne))
    hbatch = M.ElementBatch(basis, h_corners, ctx)

    def step_host():
        hbatch.set_corners(h_corners)
        M.compute_element_matrices(specs, lhs, hbatch, out=h_out)

    if args.skip_e2e:
        e2e_steps = 0
    else:
        step_host()
        e2e_steps = max(1, min(args.steps, 3))
    ctx.synchronize()
    barrier(dist)
    t0 = time.perf_counter()
    ctx.timer_start()
    for _ in range(e2e_steps):
        step_host()
    e2e_ms = ctx.timer_stop()
    ctx.synchronize()
    e2e_wall = (time.perf_counter() - t0) * 1e3
    e2e_ms = max_over_ranks(dist, max(e2e_ms, e2e_wall))
    e2e_value = total_elems / (e2e_ms / e2e_steps * 1e-3) if e2e_steps else None
    check_sum = float(h_out[0].sum()) if e2e_steps else None
    if rank == 0:
        clocks = sampler.stop()
    del h_out, hbatch

    # ---- extra metrics on rank-local data: matrix-free apply and the Schur-CG solve ----
    extra = {}
    if not args.skip_extra:
        rng = np.random.default_rng(rank)
        x = ctx.to_device(rng.standard_normal((n, ne)))
        y = ctx.empty((n, ne))
        M.compute_element_vectors(specs, lhs, batch, x, out=y)
        ctx.synchronize()
        ctx.timer_start()
        reps = 5
        for _ in range(reps):
            M.compute_element_vectors(specs, lhs, batch, x, out=y)
        t = ctx.timer_stop() / reps
        extra["apply_recompute_gdofs"] = sum_over_ranks(dist, n * ne / (t * 1e-3) / 1e9)
        # stored blocks: y_e = A_e x_e with the assembled matrices (LinearSystem.apply_diagonal)
        nh = N_SIDE
        s_names, s_orders, s_lhs = load_system("mixed_poisson")
        s_specs = M._ElementFormSpecification(*zip(s_names, s_orders))
        s_ne = s_specs.total_size(p, p)
        s_out = ctx.empty((n, s_ne, s_ne))
        M.compute_element_matrices(s_specs, s_lhs, batch, out=s_out)
        # the GLOBAL Darcy system of the nh x (nh * world) mesh, split into the ranks' strips: interface multipliers
        # are summed and the CG scalars reduced over NCCL (one process per GPU); with one rank nothing is shared
        row, col, val, nt_glob = ST.continuity_constraints_1form(nh, (n // nh) * world, p, s_orders)
        parts = ST.strip_partition(nh, (n // nh) * world, world)
        bounds = [parts[r][0] * s_ne for r in range(world)] + [parts[-1][1] * s_ne]
        part = ST.partition_constraints(row, col, val, nt_glob, bounds, rank)
        del row, col, val
        nt = part.n_trace
        if world > 1:
            from mfv2d_b200.runtime import comm_unique_id

            ids = [comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ids, src=0)
            ctx.init_comm(world, rank, ids[0])
        t0 = time.perf_counter()
        system = SV.LinearSystem(s_out, (part.row, part.col, part.val), nt, ctx)
        if world > 1:
            system.set_interface(part.shared_index, part.slot, part.n_slots, part.owned)
        ctx.synchronize()
        extra["system_setup_s"] = time.perf_counter() - t0
        xv = ctx.to_device(rng.standard_normal(n * s_ne))
        yv = ctx.empty(n * s_ne)
        system.apply_diagonal(xv, yv)
        ctx.synchronize()
        ctx.timer_start()
        for _ in range(reps):
            system.apply_diagonal(xv, yv)
        t = ctx.timer_stop() / reps
        extra["apply_stored_gdofs"] = sum_over_ranks(dist, n * s_ne / (t * 1e-3) / 1e9)
        extra["apply_stored_gbs_per_gpu"] = (8 * s_ne * s_ne + 16 * s_ne) * n / (t * 1e-3) / 1e9
        # Darcy (mixed Poisson) solve, f = random, rtol 1e-10
        rhs = ctx.to_device(rng.standard_normal(n * s_ne))
        phi = ctx.to_device(np.zeros(nt))
        barrier(dist)
        ctx.synchronize()
        t0 = time.perf_counter()
        xs, lam, res, it = SV.solve_schur_iterative(system, rhs, phi, SV.ConvergenceSettings(20000, 1e300, 1e-10))
        ctx.synchronize()
        solve_s = max_over_ranks(dist, time.perf_counter() - t0)
        extra["solve"] = {"system": f"mixed Poisson (Darcy) {nh}x{(n // nh) * world} p{p}, {world} strip(s), random rhs, rtol 1e-10, "
                                    "Schur-complement CG on the device" + (", interface sums and dots over NCCL" if world > 1 else ""),
                          "s": solve_s, "iterations": it, "residual": res, "multipliers": int(nt_glob),
                          "interface_multipliers": int(part.n_slots), "dense_dofs": int(n * s_ne * world)}

        # slice of BASELINE.json configs[4]: mixed Poisson, order 6, the same number of elements per GPU
        del system, xv, yv, rhs, phi, xs, lam, s_out
        p6 = 6
        b6 = M.Basis1D(p6, M.IntegrationRule1D(p6 + k))
        batch6 = M.ElementBatch(M.Basis2D(b6, b6), d_corners, ctx)
        ne6 = s_specs.total_size(p6, p6)
        out6 = ctx.empty((n, ne6, ne6))
        M.compute_element_matrices(s_specs, s_lhs, batch6, out=out6)
        ctx.synchronize()
        ctx.timer_start()
        for _ in range(reps):
            M.compute_element_matrices(s_specs, s_lhs, batch6, out=out6)
        t = ctx.timer_stop() / reps
        x6 = ctx.to_device(rng.standard_normal((n, ne6)))
        y6 = ctx.empty((n, ne6))
        M.compute_element_vectors(s_specs, s_lhs, batch6, x6, out=y6)
        ctx.synchronize()
        ctx.timer_start()
        for _ in range(reps):
            M.compute_element_vectors(s_specs, s_lhs, batch6, x6, out=y6)
        t2 = ctx.timer_stop() / reps
        extra["order6_mixed_poisson"] = {
            "elements_per_gpu": n, "element_dofs": ne6,
            "element_matrices_per_s": sum_over_ranks(dist, n / (t * 1e-3)),
            "assembly_hbm_frac_per_gpu": (64 + 8 * ne6 * ne6) * n / (t * 1e-3) / 1e9 / peak,
            "apply_recompute_gdofs": sum_over_ranks(dist, n * ne6 / (t2 * 1e-3) / 1e9)}
        del out6, x6, y6

    line = {
        "metric": "element_matrices_per_s", "value": value, "unit": "element matrices/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(world), "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "element matrices/s", "h2d_bytes_per_step": int(corners.nbytes),
                "d2h_bytes_per_step": int(n * ne * ne * 8), "checksum_first_matrix": check_sum},
        "gpu_launches": int(launches), "roofline": roofline, "extra": extra,
    }
    if rank == 0:
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(1, args.cpu_seconds)
        emit(line)
    barrier(dist)
    if dist is not None:
        dist.destroy_process_group()


_JSON_FD = None


def quiet_stdout():
    """Native libraries (NCCL's version banner, ...) write to fd 1: keep the process's stdout for the ONE JSON line
    and send everything else to stderr."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--skip-extra", action="store_true", help="skip the apply / solve extras (profiling runs)")
    ap.add_argument("--skip-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--workload", default="stokes4", choices=["stokes4", "poisson6"],
                    help="stokes4 = BASELINE.json configs[1] (the default and the reported metric); poisson6 = per-GPU slice of "
                         "configs[4] (mixed Poisson, order 6), for profiling")
    args = ap.parse_args()
    if args.workload == "poisson6":
        global P_ORDER, SYSTEM
        P_ORDER, SYSTEM = 6, "mixed_poisson"
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
