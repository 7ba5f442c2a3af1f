#!/usr/bin/env python
"""Times the B1 pass (k_hess_small + k_fin_small + k_cost) on the bench workload; used for A/B runs and as the ncu target.
usage: python scripts/hess_prof.py [steps]"""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
import bench

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
pkg = ge.load_package(); pkg.load_library()
args = type("A", (), {"windows": bench.N_WINDOWS, "voxels_per_window": bench.VOX_PER_WIN})()
fac = bench.build_workload(args, 0)
ctx = pkg.Context(0)
f = pkg.Factor(ctx, bench.WIN, args.windows)
f.upload_csr(fac); f.set_poses(fac["poses"])
for _ in range(3):
    f.hessian_async()
ctx.sync()
ctx.profile(True)
for _ in range(steps):
    f.hessian_async()
prof = ctx.profile_read(); ctx.profile(False)
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print("%-16s %.4f ms/step (%d launches)" % (k, v[0] / steps, v[1]))
H, J, c = f.hessian(fac["poses"])
print("checksum H %.17g J %.17g c %.17g" % (np.abs(H).sum(), np.abs(J).sum(), c.sum() if hasattr(c, "sum") else c))
