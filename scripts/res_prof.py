#!/usr/bin/env python
"""Times the B2 pass (k_residual + k_cost) on the bench workload; ncu target."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
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
    f.residual_async()
ctx.sync()
ctx.profile(True)
for _ in range(steps):
    f.residual_async()
prof = ctx.profile_read(); ctx.profile(False)
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0]):
    print("%-16s %.4f ms/step (%d launches)" % (k, v[0] / steps, v[1]))
print("cost %.17g" % f.residual(fac["poses"]).sum())
