"""BASELINE.json configs[3] beyond the fixture size: an h/p-refined NON-conforming quadtree mesh built from arrays alone
(mfv2d_b200/nonconforming.py), mixed orders, mixed Poisson, on one B200.

    python profiles/config3_forest.py [--roots 32] [--depth 2] [--split 0.3] [--orders 3,4,5,6] [--seed 1]

A `--roots x --roots` deformed root grid; every element is split with probability `--split` down to `--depth` levels; every
leaf gets one of `--orders` at random.  Pipeline: leaf corners from the tree -> element matrices per order bucket (one launch
per order) -> hanging-side continuity rows + natural boundary terms (`system_constraints_forest`) -> ConstraintConditioner
(orthonormal bases of the hanging-side row groups) -> LinearSystem with non-uniform blocks -> CG on the Schur complement ->
`||[b - A x - N^T l ; phi - N x]|| / ||b||` with the ORIGINAL (unconditioned) rows.  Prints one JSON line.
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
from mfv2d_b200 import nonconforming as NC  # noqa: E402
from mfv2d_b200 import solving as SV  # noqa: E402
from mfv2d_b200 import structured as ST  # noqa: E402


def random_forest(nh, nv, depth, split, orders, seed):
    """children (-1 for leaves), leaves in depth-first order, one order per leaf"""
    rng = np.random.default_rng(seed)
    children = [[-1, -1, -1, -1] for _ in range(nh * nv)]
    leaves = []

    def visit(e, level):
        if level < depth and rng.random() < split:
            kids = list(range(len(children), len(children) + 4))
            children.extend([[-1, -1, -1, -1] for _ in range(4)])
            children[e] = kids
            for k in kids:
                visit(k, level + 1)
        else:
            leaves.append(e)

    for root in range(nh * nv):
        visit(root, 0)
    leaf_orders = rng.choice(np.asarray(orders), size=len(leaves))
    return NC.QuadForest(nh, nv, np.array(children), np.array(leaves), leaf_orders)


def source(x, y):
    return -2 * np.pi**2 * np.cos(np.pi * x) * np.cos(np.pi * y)


def u_exact(x, y):
    return np.cos(np.pi * x) * np.cos(np.pi * y)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--roots", type=int, default=32)
    ap.add_argument("--depth", type=int, default=2)
    ap.add_argument("--split", type=float, default=0.3)
    ap.add_argument("--orders", default="3,4,5,6")
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--k", type=int, default=2)
    ap.add_argument("--rtol", type=float, default=1e-14)
    args = ap.parse_args()
    orders_pool = [int(v) for v in args.orders.split(",")]
    names, form_orders, lhs = B.load_system("mixed_poisson")
    specs = M._ElementFormSpecification(*zip(names, form_orders))
    nh = nv = args.roots
    t0 = time.perf_counter()
    forest = random_forest(nh, nv, args.depth, args.split, orders_pool, args.seed)
    root_corners = ST.unit_square_corners(nh, nv, F.deform)
    corners = forest.leaf_corners(root_corners)
    orders = forest.leaf_orders.astype(int)
    tree_s = time.perf_counter() - t0
    # element matrices and the source term, one bucket (one launch) per order
    t0 = time.perf_counter()
    iu = names.index("u")
    mats, rhs_parts = [None] * orders.size, [None] * orders.size
    for p in np.unique(orders):
        idx = np.flatnonzero(orders == p)
        rule = M.IntegrationRule1D(int(p) + args.k)
        b1 = M.Basis1D(int(p), rule)
        batch = M.ElementBatch(M.Basis2D(b1, b1), np.ascontiguousarray(corners[idx]))
        out = M.compute_element_matrices(specs, lhs, batch)
        f = np.asarray(batch.dual_dofs(3, source))
        ou, nu = specs.form_offset(iu, int(p), int(p)), specs.form_size(iu, int(p), int(p))
        for j, e in enumerate(idx):
            mats[e] = out[j]
            r = np.zeros(out[j].shape[0])
            r[ou:ou + nu] = f[j]
            rhs_parts[e] = r
    assembly_s = time.perf_counter() - t0
    # constraint rows: hanging sides + the natural condition u = u_exact on the whole boundary (weak term of the q equation)
    t0 = time.perf_counter()
    iq = names.index("q")
    row, col, val, phi, nt, off, weak_vec = NC.system_constraints_forest(forest, form_orders, root_corners, args.k, (), (), [(iq, [(1.0, u_exact)])])
    rhs = np.concatenate(rhs_parts) + weak_vec
    constraints_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    cond = SV.ConstraintConditioner(row, col, val, nt, off)
    conditioner_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    sysc = SV.LinearSystem(mats, (cond.row, cond.col, cond.val), nt)
    ctx = sysc.ctx
    ctx.synchronize()
    system_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    x, lam_c, res, it = SV.solve_schur_iterative(sysc, rhs, cond.rhs(phi), SV.ConvergenceSettings(50000, 1e300, args.rtol), precondition="jacobi")
    ctx.synchronize()
    solve_s = time.perf_counter() - t0
    lam = cond.multipliers(lam_c)
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    r1 = rhs - sysm.apply_diagonal(x) - sysm.apply_trace_transpose(lam)
    r2 = phi - sysm.apply_trace(x)
    rel = float(np.sqrt(r1 @ r1 + r2 @ r2) / np.sqrt(rhs @ rhs))
    # the same residual with the conditioned rows (an equivalent formulation of the same constraints, same x): it is free of the
    # up-to-1e7 coefficients of the raw hanging-side rows that amplify the rounding of N x in the figure above
    phic = cond.rhs(phi)
    c1 = rhs - sysc.apply_diagonal(x) - sysc.apply_trace_transpose(lam_c)
    c2 = phic - sysc.apply_trace(x)
    rel_c = float(np.sqrt(c1 @ c1 + c2 @ c2) / np.sqrt(rhs @ rhs))
    per_row = np.bincount(row, minlength=nt)
    B.emit({
        "config": f"BASELINE.json configs[3] at scale: {nh}x{nv} deformed roots, depth <= {args.depth}, split probability {args.split}, "
                  f"orders {orders_pool}, mixed Poisson (q1, u2), integration order p+{args.k}, natural boundary condition, one GPU",
        "elements_total": int(forest.children.shape[0]), "leaves": int(orders.size), "leaves_per_order": {int(p): int(np.sum(orders == p)) for p in np.unique(orders)},
        "dense_dofs": int(off[-1]), "multipliers": int(nt), "hanging_rows": int(np.sum(per_row > 2)), "max_entries_per_row": int(per_row.max()),
        "max_abs_coefficient": float(np.abs(val).max()), "conditioned_groups": len(cond.blocks),
        "tree_s": tree_s, "assembly_incl_context_and_plans_s": assembly_s, "constraints_host_s": constraints_s, "conditioner_host_s": conditioner_s, "system_setup_s": system_s,
        "solve_s": solve_s, "iterations": int(it), "cg_residual": float(res), "saddle_relative_residual": rel, "saddle_relative_residual_conditioned_rows": rel_c,
    })


if __name__ == "__main__":
    main()
