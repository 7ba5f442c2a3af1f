"""vx_ldlt_solve at n = 2400 (top level of C4): residual check and per-kernel split, DMMA trailing tiles vs the round-1 DFMA tiles
(VX_LDLT_DFMA=1 in the environment selects the latter)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
n = int(os.environ.get("VX_N", "2400"))
rng = np.random.Generator(np.random.PCG64(5))
M = rng.normal(size=(n, n // 2)); A = M @ M.T + np.diag(rng.uniform(1, 50, n)); b = rng.normal(size=n)
x = pkg.ldlt_solve(ctx, A, b)
print("n", n, "residual rel", np.abs(A @ x - b).max() / np.abs(b).max(), "vs numpy", np.abs(x - np.linalg.solve(A, b)).max() / np.abs(x).max())
ctx.profile(True)
for _ in range(3):
    pkg.ldlt_solve(ctx, A, b)
prof = ctx.profile_read(); ctx.profile(False)
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print("%-24s %8.3f ms per solve  (%d launches)" % (k, v[0] / 3, v[1] // 3))
tr = [v for k, v in prof.items() if k.startswith("k_ldlt_trailing")][0]
print("trailing: %.2f TFLOP/s (n^3/3 flop)" % ((n ** 3 / 3) / (tr[0] / 3 * 1e-3) / 1e12))
