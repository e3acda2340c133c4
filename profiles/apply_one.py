import sys, os, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import bench as B
from mfv2d_b200 import _mfv2d as M, structured as ST
from mfv2d_b200.runtime import Context
import mfv2d_b200.runtime as rt
ctx = Context(0); rt._context = ctx
name, p, n_side = "mixed_poisson", 6, 354
names, orders, lhs = B.load_system(name)
specs = M._ElementFormSpecification(*zip(names, orders))
b1 = M.Basis1D(p, M.IntegrationRule1D(p + 3)); basis = M.Basis2D(b1, b1)
corners = ST.unit_square_corners(n_side, n_side, B.deform); n = corners.shape[0]
batch = M.ElementBatch(basis, ctx.to_device(corners), ctx)
ne = specs.total_size(p, p)
x = ctx.to_device(np.random.default_rng(0).standard_normal((n, ne))); y = ctx.empty((n, ne))
for _ in range(4): M.compute_element_vectors(specs, lhs, batch, x, out=y)
ctx.synchronize()
