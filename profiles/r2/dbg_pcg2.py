import sys, os, json
sys.path.insert(0, "tests")
import numpy as np
import parity_cases as P
from mfv2d_b200 import _mfv2d as M, solving as SV, structured as ST
fix, nh, nv = "solve_6x6_p3_def", 6, 6
g = P.golden(fix + ".npz"); s = P.SYSTEMS["mixed_poisson_sorted"]
p, k = int(g["order"]), int(g["k"])
specs = P.specs_of(M, s)
batch = M.ElementBatch(P.make_basis(M, p, k), np.ascontiguousarray(g["corners"]))
mats = M.compute_element_matrices(specs, s["lhs"], batch)
row, col, val, nt = ST.continuity_constraints_1form(nh, nv, p, s["form_orders"])
sysm = SV.LinearSystem(mats, (row, col, val), nt)
for mi in (1, 2, 3, 5, 10, 20, 30, 40, 50, 60, 70, 80, 90, 100, 150, 200, 250, 270, 280, 290, 295, 300, 305, 310, 315, 320, 330, 350, 400, 500, 700, 1000, 2000, 4000):
    xp, lp, resp, itp = SV.solve_pcg_iterative(sysm, g["rhs"], g["con_rhs"], SV.ConvergenceSettings(mi, 1e-30, 1e-10))
    print(mi, resp, itp, P.relerr(xp, g["sol_dense"]), flush=True)
