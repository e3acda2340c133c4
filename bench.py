#!/usr/bin/env python
"""Benchmark of the element hot path (BASELINE.json metric: element matrices/sec and matrix-free apply GDoF/s (fp64); solve s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch: the assembly of ALL element matrices of the workload
(BASELINE.json configs[1]: steady Stokes, 256 x 256 elements, order 4, integration order p + 3, sinusoidally
deformed unit square, N_e = 81).  With N > 1 (torchrun, one process per GPU) every rank assembles its own
copy-sized partition of an N-times larger mesh ("weak" scaling); assembly needs no collective.

Top-level keys of the ONE JSON line:
`value`         element matrices/s, whole job, inputs (corners) and outputs resident in HBM, CUDA events.
`e2e`           the same through the C-ABI with HOST buffers: corners copied H2D and the element matrices copied D2H to
                pinned memory inside the timed region.
`roofline`      HBM-write roofline of the assembly kernel: (64 + 8 N_e^2) bytes per element / kernel time.
`apply`         matrix-free apply (compute_element_vector semantics, recompute variant), mixed Poisson order 6 (the
                north_star order) and Stokes order 4: GDoF/s, fp64 roofline (flops per element from SURVEY.md 8d's MAC
                formula against the DMMA / DFMA peak measured in this process), CPU baseline.
`apply_stored`  y_e = A_e x_e with the stored blocks (LinearSystem.apply_diagonal): GDoF/s, HBM-read roofline, CPU baseline.
`solve`         BASELINE.json configs[4], weak-scaled: mixed Poisson order 6 on a deformed 1000 x (125 N) mesh (N = 8 is the
                full 10^6-element mesh), smooth right-hand side, Schur-complement CG with the multilevel preconditioner,
                interface sums and CG scalars through NVLink peer stores inside the CG kernels: seconds, iterations,
                us / iteration, and the relative residual of the SADDLE-POINT system summed over the ranks.  The run
                EXITS NON-ZERO when that residual exceeds north_star's 1e-10.
`cpu_baseline`  the reference's own C path (oracle/_ref, built from /root/reference by oracle/build_ref.sh) or the numpy
                oracle port, timed on a bounded sample on this box's host cores.
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
            "elements_per_gpu": N_SIDE * N_SIDE, "element_dofs": {"stokes": 81, "mixed_poisson": 120}.get(SYSTEM, 0) if P_ORDER in (4, 6) else 0,
            "parallelism": f"elements x{n_gpus}: assembly and apply need no collective; the solve (key `solve`) exchanges interface sums and CG scalars "
                           "through NVLink peer stores",
            "l2": "outputs (3.4 GB per step or more) are larger than L2; no flush needed"}


def apply_macs(system: str, p: int, q: int) -> int:
    """MACs of one matrix-free element apply, SURVEY.md 8(d): I(A, B) = A B q + A q^2 per tensor interpolation (its transpose
    costs the same); M0 x = 2 I(p+1, p+1) + q^2, M2 x = 2 I(p, p) + q^2, M1 x = 2 [I(p+1, p) + I(p, p+1)] + 4 q^2, geometry
    15 q^2 once per element; incidences and scalings are not counted."""
    def interp(a, b):
        return a * b * q + a * q * q

    m0 = 2 * interp(p + 1, p + 1) + q * q
    m2 = 2 * interp(p, p) + q * q
    m1 = 2 * (interp(p + 1, p) + interp(p, p + 1)) + 4 * q * q
    geo = 15 * q * q
    if system == "mixed_poisson":  # M1 x_q, M2 x_u, M2 (E21 x_q)
        return m1 + 2 * m2 + geo
    if system == "stokes":  # M0 x_vor, M1 (E10 x_vor), E10^T (M1 x_vel), M2 (E21 x_vel), E21^T (M2 x_prs)
        return m0 + 2 * m1 + 2 * m2 + geo
    raise ValueError(system)


def source_exact(x, y):
    """right-hand side of examples/steady/plot_mixed_poisson.py:125-143 (smooth)"""
    return 2.0 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)


# ------------------------------------------------------------------------------------------------
# CPU baselines of the apply / stored apply / solve legs: the reference's own C code on one host core
# ------------------------------------------------------------------------------------------------
def cpu_baseline_apply(seconds: float):
    """compute_element_vector (src/evaluation/element_system.c:245-411), mixed Poisson order 6, one core."""
    from oracle.ref_loader import load_ref

    ref = load_ref()
    p, k = 6, K_OVER
    names, orders, lhs = load_system("mixed_poisson")
    from mfv2d_b200.structured import unit_square_corners

    corners = unit_square_corners(64, 64, deform)
    rng = np.random.default_rng(3)
    ne = 2 * p * (p + 1) + p * p
    done, t0 = 0, time.perf_counter()
    if ref is not None:
        rule = ref.IntegrationRule1D(p + k)
        b1 = ref.Basis1D(p, rule)
        b2 = ref.Basis2D(b1, b1)
        specs = ref._ElementFormSpecification(*zip(names, orders))
        x = rng.standard_normal(ne)
        for c in corners:
            sp = ref.ElementFemSpace2D(b2, np.ascontiguousarray(c))
            ref.compute_element_vector(specs, lhs, sp, x)
            done += 1
            if time.perf_counter() - t0 > seconds:
                break
        kind = "reference"
    else:
        from oracle import mfv2d_oracle as O

        ob = O.basis(p, p + k + 1)
        blocks, fs = O.decode_expressions(lhs, names)
        x = rng.standard_normal(ne)
        for c in corners:
            O.element_vector(ob, c, orders, blocks, x, fs)
            done += 1
            if time.perf_counter() - t0 > seconds:
                break
        kind = "port"
    t = time.perf_counter() - t0
    return {"value": done * ne / t / 1e9, "unit": "GDoF/s", "cores": 1, "kind": kind,
            "sample": f"{done} elements of a deformed 64x64 p6 mixed-Poisson mesh (ElementFemSpace2D + compute_element_vector per element, "
                      f"as the reference's callers do), {t:.1f} s on one core; per-element cost is independent of mesh size"}


def cpu_baseline_system(seconds: float):
    """LinearSystem.apply_diagonal (src/algebra/sparse_system.c:3-40) and solve_schur_iterative (python/mfv2d/solving.py:439-500)
    of the reference on a small mixed-Poisson mesh: its C operators driven by oracle/ref_solve.py."""
    from oracle.ref_loader import load_ref

    ref = load_ref()
    if ref is None:
        return None, None
    from mfv2d_b200 import structured as ST
    from oracle import ref_solve as RS

    p, k, nh = 6, K_OVER, 12
    names, orders, lhs = load_system("mixed_poisson")
    rule = ref.IntegrationRule1D(p + k)
    b1 = ref.Basis1D(p, rule)
    b2 = ref.Basis2D(b1, b1)
    specs = ref._ElementFormSpecification(*zip(names, orders))
    corners = ST.unit_square_corners(nh, nh, deform)
    mats = np.stack([ref.compute_element_matrix(specs, lhs, ref.ElementFemSpace2D(b2, np.ascontiguousarray(c))) for c in corners])
    ne = mats.shape[1]
    row, col, val, nt = ST.continuity_constraints_1form(nh, nh, p, orders)
    t0 = time.perf_counter()
    system = RS.build_system(ref, mats, row, col, val, nt)
    build_s = time.perf_counter() - t0
    rng = np.random.default_rng(4)
    x = ref.DenseVector(system, *rng.standard_normal((nh * nh, ne)))
    y = ref.DenseVector(system)
    reps, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < min(seconds, 3.0):
        system.apply_diagonal(x, y)
        reps += 1
    t = (time.perf_counter() - t0) / reps
    stored = {"value": nh * nh * ne / t / 1e9, "unit": "GDoF/s", "cores": 1, "kind": "reference",
              "sample": f"LinearSystem.apply_diagonal on {nh}x{nh} p{p} mixed-Poisson blocks (N_e = {ne}), {reps} applications, one core"}
    # right-hand side as in the GPU leg: dual DoFs of the smooth source on the 2-form rows
    from oracle import mfv2d_oracle as O

    ob = O.basis(p, p + k + 1)
    rhs = np.zeros((nh * nh, ne))
    for e, c in enumerate(corners):
        xx, yy = O.physical_coordinates(c, ob.pnts)
        rhs[e, ne - p * p:] = O.dual_dofs(ob, c, 3, source_exact(xx, yy))
    t0 = time.perf_counter()
    xs, lam, res, it = RS.solve_schur(ref, system, list(rhs), np.zeros(nt), 20000, 1e300, 1e-10)
    t = time.perf_counter() - t0
    solve = {"value": t, "unit": "s", "cores": 1, "kind": "reference",
             "iterations": int(it) + 1, "us_per_iteration_per_element": t / (it + 1) / (nh * nh) * 1e6,
             "sample": f"solve_schur_iterative on a deformed {nh}x{nh} p{p} mixed-Poisson mesh ({nh * nh} elements, {nt} multipliers), rtol 1e-10, "
                       f"the reference's C operators (LU solves of the {ne}x{ne} blocks per application) driven by its own CG loop; "
                       f"LinearSystem construction took {build_s:.1f} s on top; cost per iteration is proportional to the element count"}
    return stored, solve


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
    del full
    n = corners.shape[0]
    ne = specs.total_size(p, p)
    d_corners = ctx.to_device(corners)
    batch = M.ElementBatch(basis, d_corners, ctx)
    out = ctx.empty((n, ne, ne))
    peak, peak_src = measured_peaks()

    def step_device():
        M.compute_element_matrices(specs, lhs, batch, out=out)

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()  # sampled over ALL timed regions of this run (assembly, apply, solve)
    for _ in range(args.warmup):
        step_device()
    ctx.synchronize()
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
    ms = max_over_ranks(dist, ms)
    total_elems = sum_over_ranks(dist, float(n))
    ms_per_step = ms / args.steps
    value = total_elems / (ms_per_step * 1e-3)

    # per-launch kernel duration (the step is exactly one kernel launch): roofline of the assembly kernel
    alg_bytes = (64 + 8 * ne * ne) * n
    achieved = alg_bytes / (ms_per_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "kernel": "element_matrix_kernel", "peak_source": peak_src, "algorithmic_bytes_per_element": 64 + 8 * ne * ne}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        with open(tr_path) as f:
            roofline["traffic"] = json.load(f).get("element_matrix_kernel_bytes_per_launch")

    # ---- e2e: host buffers through the C-ABI, copies inside the timed region ----
    e2e_value, check_sum = None, None
    if not args.skip_e2e:
        h_corners = ctx.pinned((n, 4, 2))
        h_corners[:] = corners
        h_out = ctx.pinned((n, ne, ne))
        hbatch = M.ElementBatch(basis, h_corners, ctx)

        def step_host():
            hbatch.set_corners(h_corners)
            M.compute_element_matrices(specs, lhs, hbatch, out=h_out)

        step_host()
        e2e_steps = max(1, min(args.steps, 5))
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
        e2e_value = total_elems / (e2e_ms / e2e_steps * 1e-3)
        check_sum = float(h_out[0].sum())
        del h_out, hbatch
    del out

    line = {
        "metric": "element_matrices_per_s", "value": value, "unit": "element matrices/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(world), "clocks": None,
        "e2e": {"value": e2e_value, "unit": "element matrices/s", "h2d_bytes_per_step": int(corners.nbytes),
                "d2h_bytes_per_step": int(n * ne * ne * 8), "checksum_first_matrix": check_sum},
        "gpu_launches": int(launches), "roofline": roofline,
    }
    failed = None
    if not args.skip_extra:
        dmma_tf, dfma_tf = ctx.measure_fp64_peak()
        fp64_peak = max(dmma_tf, dfma_tf)
        reps = 10
        rng = np.random.default_rng(rank)

        def time_apply(sys_name, specs_, lhs_, batch_, ne_, p_):
            x = ctx.to_device(rng.standard_normal((n, ne_)))
            y = ctx.empty((n, ne_))
            for _ in range(3):
                M.compute_element_vectors(specs_, lhs_, batch_, x, out=y)
            ctx.synchronize()
            barrier(dist)
            ctx.timer_start()
            for _ in range(reps):
                M.compute_element_vectors(specs_, lhs_, batch_, x, out=y)
            t = max_over_ranks(dist, ctx.timer_stop() / reps)
            macs = apply_macs(sys_name, p_, p_ + k + 1)
            tf = 2.0 * macs * n / (t * 1e-3) / 1e12
            return {"value": total_elems * ne_ / (t * 1e-3) / 1e9, "unit": "GDoF/s", "ms": t, "elements_per_gpu": n, "element_dofs": ne_,
                    "roofline": {"bound": "fp64", "achieved": tf, "peak": fp64_peak, "unit": "TFLOP/s", "frac": tf / fp64_peak, "traffic": None,
                                 "flops_per_element": 2 * macs, "algorithmic_bytes_per_element": 64 + 16 * ne_,
                                 "peak_source": f"measured in this process: mma.sync.m8n8k4.f64 chains {dmma_tf:.1f}, DFMA chains {dfma_tf:.1f} TFLOP/s "
                                                "(mfv2d_cuda_measure_fp64_peak)"}}

        # matrix-free apply, Stokes order 4 (the assembly workload) and mixed Poisson order 6 (north_star's order)
        apply4 = time_apply("stokes", specs, lhs, batch, ne, p) if SYSTEM == "stokes" else None
        s_names, s_orders, s_lhs = load_system("mixed_poisson")
        s_specs = M._ElementFormSpecification(*zip(s_names, s_orders))
        p6 = 6
        b6 = M.Basis1D(p6, M.IntegrationRule1D(p6 + k))
        basis6 = M.Basis2D(b6, b6)
        ne6 = s_specs.total_size(p6, p6)
        batch6 = M.ElementBatch(basis6, d_corners, ctx)
        apply6 = time_apply("mixed_poisson", s_specs, s_lhs, batch6, ne6, p6)
        line["apply"] = dict(apply6, workload=f"compute_element_vector semantics (recompute), mixed Poisson order 6, {n} deformed elements per GPU",
                             stokes_order4=apply4)
        del batch6

        # ---- BASELINE.json configs[4], weak-scaled: 1000 x (125 world) elements, mixed Poisson order 6 ----
        nh, rows_per = args.solve_cols, args.solve_rows
        nv = rows_per * world
        t_host = time.perf_counter()
        parts = ST.strip_partition(nh, nv, world)
        flo, fhi = parts[rank]
        xi, eta = np.meshgrid(np.linspace(-1, 1, nh + 1), np.linspace(-1, 1, nv + 1)[flo // nh:fhi // nh + 1])
        px, py = deform(xi, eta)
        pts = np.stack((px, py), axis=-1)
        c6 = np.empty((fhi // nh - flo // nh, nh, 4, 2))
        c6[:, :, 0], c6[:, :, 1], c6[:, :, 2], c6[:, :, 3] = pts[:-1, :-1], pts[:-1, 1:], pts[1:, 1:], pts[1:, :-1]
        c6 = np.ascontiguousarray(c6.reshape(-1, 4, 2))
        n6 = c6.shape[0]
        # this rank's constraint rows, interface lists and aggregate hierarchy straight from the closed form of the reference's
        # row order: O(strip) per rank (the global builder + partition_constraints give the same arrays, tests/test_distributed.py)
        mids = tuple(int(v) for v in args.solve_mid.split(",") if v)
        part, hier, nt_glob = ST.strip_system_1form(nh, nv, p6, s_orders, world, rank, mids, args.solve_agg)
        n_agg = hier[2]
        host_s = time.perf_counter() - t_host
        if world > 1:
            ctx.init_comm(world, rank, None)  # ranks only: every exchange goes through the NVLink peer blocks, NCCL is not loaded
        t0 = time.perf_counter()
        solve_batch = M.ElementBatch(basis6, ctx.to_device(c6), ctx)
        mats6 = ctx.empty((n6, ne6, ne6))
        M.compute_element_matrices(s_specs, s_lhs, solve_batch, out=mats6)
        ctx.synchronize()
        t_asm = time.perf_counter() - t0
        system = SV.LinearSystem(mats6, (part.row, part.col, part.val), part.n_trace, ctx)
        t_sys = time.perf_counter() - t0 - t_asm
        if world > 1:
            system.set_interface(part.shared_index, part.slot, part.n_slots, part.owned)

            def all_gather(obj):
                outl = [None] * world
                dist.all_gather_object(outl, obj)
                return outl

            system.connect_peers(part.holder_ptr, part.holder_rank, all_gather, n_coarse=n_agg[-1])
        ctx.synchronize()
        setup_rank0 = time.perf_counter() - t0
        setup_s = max_over_ranks(dist, setup_rank0)
        t0 = time.perf_counter()
        system.set_coarse_space(*hier)
        ctx.synchronize()
        coarse_s = max_over_ranks(dist, time.perf_counter() - t0)

        # stored-block apply on the same blocks (LinearSystem.apply_diagonal): HBM read bound
        xv = ctx.to_device(rng.standard_normal(n6 * ne6))
        yv = ctx.empty(n6 * ne6)
        system.apply_diagonal(xv, yv)
        ctx.synchronize()
        barrier(dist)
        ctx.timer_start()
        for _ in range(5):
            system.apply_diagonal(xv, yv)
        t = max_over_ranks(dist, ctx.timer_stop() / 5)
        sb = (8 * ne6 * ne6 + 16 * ne6) * n6 / (t * 1e-3) / 1e9
        line["apply_stored"] = {"value": sum_over_ranks(dist, float(n6)) * ne6 / (t * 1e-3) / 1e9, "unit": "GDoF/s", "ms": t,
                                "workload": f"y_e = A_e x_e, stored mixed-Poisson order-6 blocks (N_e = {ne6}), {n6} elements per GPU",
                                "roofline": {"bound": "hbm", "achieved": sb, "peak": peak, "unit": "GB/s", "frac": sb / peak, "traffic": None,
                                             "kernel": "block_gemv_kernel", "algorithmic_bytes_per_element": 8 * ne6 * ne6 + 16 * ne6,
                                             "peak_source": peak_src}}
        del xv, yv

        # smooth right-hand side: dual DoFs of source_exact on the 2-form rows; HOST buffers in, HOST solution out
        # (page-locked host memory, as the e2e contract asks: the copies run at PCIe speed instead of through a staging buffer)
        rhs_h = ctx.pinned((n6, ne6))
        rhs_h[:] = 0.0
        fdual = solve_batch.dual_dofs(s_orders[1], source_exact)
        rhs_h[:, ne6 - p6 * p6:] = fdual.get() if hasattr(fdual, "get") else fdual
        rhs_h = rhs_h.reshape(-1)
        phi_h = ctx.pinned(part.n_trace)
        phi_h[:] = 0.0
        sol_out = (ctx.pinned(n6 * ne6), ctx.pinned(part.n_trace))
        conv = SV.ConvergenceSettings(20000, 1e300, args.solve_rtol)
        times = []
        for _ in range(3):  # the first solve instantiates the CUDA graph of 16 iterations; the best of the others is reported
            barrier(dist)
            ctx.synchronize()
            t0 = time.perf_counter()
            xs, lam, res, it = SV.solve_schur_iterative(system, rhs_h, phi_h, conv, precondition="multilevel", out=sol_out)
            ctx.synchronize()
            times.append(max_over_ranks(dist, time.perf_counter() - t0))
        peer_bad = system.peer_error() if world > 1 else False
        # residual of the saddle-point system [[A, N^T], [N, 0]] (x, lambda) = (rhs, 0), with the library's own operators,
        # ownership-weighted over the ranks (north_star: <= 1e-10)
        r1 = system.apply_diagonal(xs) + system.apply_trace_transpose(lam) - rhs_h
        r2 = system.apply_trace(xs)
        own = np.asarray(part.owned, bool) if world > 1 else np.ones(part.n_trace, bool)
        num = sum_over_ranks(dist, float(r1 @ r1 + r2[own] @ r2[own]))
        den = sum_over_ranks(dist, float(rhs_h @ rhs_h))
        saddle = float(np.sqrt(num / den))
        solve_s = min(times[1:])
        # per-iteration byte model: packed condensed blocks + slot vectors + gather / update vectors of the CG
        m_e = 4 * p6
        it_bytes = n6 * (m_e * (m_e + 1) // 2 * 8 + m_e * (8 + 4 + 8)) + part.n_trace * (48 + 88)
        line["solve"] = {
            "system": f"mixed Poisson (q1, u2) order {p6} on a deformed {nh}x{nv} quad mesh, {world} strip(s) of {rows_per} rows "
                      + ("= BASELINE.json configs[4] (10^6 elements)" if nh * nv == 1000000 else "(per-GPU share of BASELINE.json configs[4], weak-scaled)"),
            "method": "CG on the Schur complement N A^-1 N^T (packed symmetric condensed blocks), additive multilevel preconditioner "
                      f"(Jacobi + aggregates of {args.solve_mid}^2 elements + exact level of {args.solve_agg}^2 elements, {n_agg[-1]} global aggregates)"
                      + (", interface sums and CG scalars by NVLink peer stores inside the CG kernels (no NCCL)" if world > 1 else ""),
            "rhs": "dual DoFs of 2 pi^2 cos(pi x) cos(pi y) (plot_mixed_poisson.py's source_exact), page-locked host buffers in / out", "rtol": args.solve_rtol,
            "s": solve_s, "s_first_solve": times[0], "s_all_solves": times, "iterations": int(it), "us_per_iteration": solve_s / max(int(it), 1) * 1e6,
            "cg_residual": float(res), "saddle_relative_residual": saddle, "gate": 1e-10, "peer_error": bool(peer_bad),
            "elements": int(nh * nv), "dense_dofs": int(nh * nv * ne6), "multipliers": int(nt_glob), "interface_multipliers": int(part.n_slots),
            "system_setup_s": setup_s, "system_setup_breakdown_rank0_s": {"assembly": t_asm, "system_create": t_sys, "interface_and_peers": max(setup_rank0 - t_asm - t_sys, 0.0)},
            "preconditioner_setup_s": coarse_s, "host_mesh_constraints_partition_s": host_s,
            "pipeline_s": host_s + setup_s + coarse_s + times[0],
            "roofline": {"bound": "hbm", "achieved": it_bytes / (solve_s / max(int(it), 1)) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": it_bytes / (solve_s / max(int(it), 1)) / 1e9 / peak, "traffic": None, "kernel": "one CG iteration (schur_local_packed + "
                         "cg1_gather + cg1_update; the preconditioner's kernels are not counted as useful bytes)", "algorithmic_bytes_per_iteration": int(it_bytes),
                         "peak_source": peak_src}}
        if not (saddle <= 1e-10) or peer_bad:
            failed = f"solve: saddle-point relative residual {saddle:.3e} exceeds 1e-10" if not peer_bad else "solve: peer exchange timed out"
        del system, mats6, xs, lam
    if rank == 0:
        line["clocks"] = sampler.stop()
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(1, args.cpu_seconds)
            line["vs_cpu_baseline"] = {"value": value / line["cpu_baseline"]["value"], "e2e": (e2e_value or 0.0) / line["cpu_baseline"]["value"]}
            if not args.skip_extra:
                line["apply"]["cpu_baseline"] = cpu_baseline_apply(min(args.cpu_seconds, 6.0))
                stored_b, solve_b = cpu_baseline_system(min(args.cpu_seconds, 6.0))
                line["apply_stored"]["cpu_baseline"] = stored_b
                line["solve"]["cpu_baseline"] = solve_b
        emit(line)
    barrier(dist)
    if dist is not None:
        dist.destroy_process_group()
    if failed:
        sys.stderr.write("bench.py: FAILED parity gate: " + failed + "\n")
        sys.exit(3)


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
    ap.add_argument("--skip-extra", action="store_true", help="skip the apply / solve legs (profiling runs)")
    ap.add_argument("--skip-e2e", action="store_true", help="skip the host-buffer leg (profiling runs)")
    ap.add_argument("--workload", default="stokes4", choices=["stokes4", "poisson6"],
                    help="stokes4 = BASELINE.json configs[1] (the default and the reported metric); poisson6 = per-GPU slice of "
                         "configs[4] (mixed Poisson, order 6), for profiling")
    ap.add_argument("--solve-cols", type=int, default=1000, help="elements per row of the solve leg's mesh (configs[4]: 1000)")
    ap.add_argument("--solve-rows", type=int, default=125, help="element rows PER GPU of the solve leg's mesh (8 x 125 = configs[4])")
    ap.add_argument("--solve-agg", type=int, default=20, help="aggregate size of the exact coarse level (elements per side)")
    ap.add_argument("--solve-mid", default="5", help="aggregate sizes of the diagonal levels, comma separated")
    ap.add_argument("--solve-rtol", type=float, default=1e-11, help="relative tolerance of the Schur CG (the saddle-point gate is 1e-10)")
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
