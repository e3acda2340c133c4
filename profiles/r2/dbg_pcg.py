import sys, os, json
sys.path.insert(0, "tests")
import numpy as np
import parity_cases as P
from mfv2d_b200 import _mfv2d as M, solving as SV, structured as ST
for fix, nh, nv in [("solve_6x6_p3_def",6,6),("solve_10x10_p3",10,10),("solve_3x4_p4",3,4),("solve_5x2_p5_def",5,2)]:
    g = P.golden(fix + ".npz"); s = P.SYSTEMS["mixed_poisson_sorted"]
    p, k = int(g["order"]), int(g["k"])
    specs = P.specs_of(M, s)
    batch = M.ElementBatch(P.make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
    mats = M.compute_element_matrices(specs, s["lhs"], batch)
    row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
    sysm = SV.LinearSystem(mats, (row, col, val), nt)
    for mi in (5000, 20000):
        xp, lp, resp, itp = SV.solve_pcg_iterative(sysm, g["rhs"], g["con_rhs"], SV.ConvergenceSettings(mi, 1e-30, 1e-10))
        print(fix, mi, resp, itp, P.relerr(xp, g["sol_dense"]), P.relerr(lp, g["sol_trace"]), flush=True)
    ainv = np.stack(sysm.get_dense_blocks()) if False else None
    xr = np.random.default_rng(0).standard_normal(sysm.n_dense)
    y = sysm.apply_diagonal(sysm.apply_diagonal_inverse(xr))
    print("  A A^-1 x - x:", P.relerr(y, xr), " vs golden inverse:", P.relerr(sysm.apply_diagonal_inverse(g["x"]), g["apply_diagonal_inverse"]), flush=True)
