import sys, os, time
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/tests/emu"); sys.path.insert(0, "/root/repo")
os.environ["MFV2D_TEST_HOOKS"] = "1"
from build_emu import build
import mfv2d_b200._capi as capi
capi._use_library_for_testing(build())
import numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spl
import parity_cases as P
import fields as F
from mfv2d_b200 import _mfv2d as M, solving as SV, structured as ST
from oracle import mfv2d_oracle as O
def bc_vec(x, y):
    return np.stack((np.cos(0.9 * x) * (1 + y), np.sin(1.1 * y + 0.3) * (2 - x)), axis=-1)
n, p = int(sys.argv[1]), int(sys.argv[2])
s = P.SYSTEMS["stokes"]; specs = P.specs_of(M, s)
corners = ST.unit_square_corners(n, n, None)
batch = M.ElementBatch(P.make_basis(M, p, 2), np.ascontiguousarray(corners))
mats = M.compute_element_matrices(specs, s["lhs"], batch)
fo = [1, 2, 3]
roots, _ = O.gll(p + 1); rn, rw = O.gll(p + 3)
row, col, val, phi, nt = ST.system_constraints(n, n, p, fo, [(0.0, 2)], [(1, bc_vec)], corners, roots, rn, rw)
ne = mats.shape[1]; N = sp.coo_matrix((val, (row, col)), shape=(nt, n*n*ne)).tocsr()
A = sp.block_diag([sp.csr_matrix(m) for m in mats]).tocsr()
K = sp.bmat([[A, N.T], [N, None]]).tocsc()
rhs = np.zeros(n*n*ne)
b = np.concatenate([rhs, phi])
t = time.time(); sol = spl.spsolve(K, b); print("spsolve", time.time() - t, "resid", np.linalg.norm(K @ sol - b) / np.linalg.norm(b))
asym = max(np.max(np.abs(m - m.T)) for m in mats[:4]); print("block asymmetry", asym, "ne", ne, "nt", nt)
sysm = SV.LinearSystem(mats, (row, col, val), nt)

Ainv = sp.block_diag([sp.csr_matrix(np.linalg.inv(m)) for m in mats]).tocsr()
S = (N @ Ainv @ N.T).tocsr()
g = N @ (Ainv @ rhs) - phi
d = np.abs(S.diagonal()); d[d == 0] = 1.0
print("diag S: min", S.diagonal().min(), "max", S.diagonal().max(), "neg", (S.diagonal() < 0).sum(), "zero", (S.diagonal()==0).sum())
for prec in (None, sp.diags(1.0 / d)):
    cnt = [0]
    def cb(xk): cnt[0] += 1
    t = time.time(); lam, info = spl.minres(S, g, M=prec, rtol=1e-12, maxiter=200000, callback=cb); dt = time.time() - t
    x = Ainv @ (rhs - N.T @ lam)
    r = np.linalg.norm(K @ np.concatenate([x, lam]) - b) / np.linalg.norm(b)
    print("minres", "jacobi" if prec is not None else "none", "it", cnt[0], "info", info, "saddle", r, "err", np.linalg.norm(x - sol[:x.size]) / np.linalg.norm(sol[:x.size]), "t", dt)
