import os, sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, __graft_entry__ as ge
pkg = ge.load_package(); pkg.load_library(); ctx = pkg.Context(0)
rng = np.random.Generator(np.random.PCG64(5))
for n in (33, 64, 75, 144, 150, 153, 303, 384, 400):
    M = rng.normal(size=(n, n)); A = M @ M.T + np.diag(rng.uniform(1, 50, n)); b = rng.normal(size=n)
    x = pkg.ldlt_solve(ctx, A, b)
    print(n, np.abs(A @ x - b).max() / np.abs(b).max())
