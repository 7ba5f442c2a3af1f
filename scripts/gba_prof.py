#!/usr/bin/env python
"""gba_build kernel times on the bench's C4-shaped input (128 windows); VX_SEG_LANES picks the lanes per leaf / voxel-frame of k_leaf_stats / k_vf_fill."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
import scene, synth
pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
clouds, (R, p) = scene.config1(seed=20241008 + 4, W=25, pts_per_scan=20000)
Rs, ps_ = synth.perturb_poses((R, p), 4, 0.004, 0.03)
W = 10
base = np.concatenate([np.arange(s0, s0 + W) for s0 in (0, 5, 10, 15)]).astype(np.int32)
fc = np.tile(base, 32); NW = fc.size // W
f = pkg.Factor(ctx, W, NW)
go = lambda: f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)
go()
ctx.profile(True)
t0 = time.perf_counter()
for _ in range(3):
    go()
t = (time.perf_counter() - t0) / 3
pr = ctx.profile_read(); ctx.profile(False)
print("lanes %s wall %.3f ms voxels %d vf %d | %s" % (os.environ.get("VX_SEG_LANES", "8"), 1e3 * t, f.num_voxels, f.num_voxel_frames,
      ", ".join("%s %.3f/%d" % (k, v[0] / 3, v[1] // 3) for k, v in sorted(pr.items(), key=lambda kv: -kv[1][0])[:5])))
H, J, c = f.hessian((Rs[fc], ps_[fc]))
print("   checksum |H| %.17g cost %.17g" % (np.abs(H).sum(), c.sum()))
