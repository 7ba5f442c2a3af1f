#!/usr/bin/env python
"""Wall clock against summed kernel time for the non-headline rows (gba_build, hba bottom level, LIO update, local map):
the gap is host overhead (launch latency, synchronisations, copies)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
import scene, synth
pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
rng = np.random.Generator(np.random.PCG64(20241008 + 4))


def prof(name, fn, n=3):
    fn()
    ctx.profile(True)
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t = (time.perf_counter() - t0) / n
    p = ctx.profile_read(); ctx.profile(False)
    ks = sorted(p.items(), key=lambda kv: -kv[1][0])
    tot = sum(v[0] for v in p.values()) / n
    nl = sum(v[1] for v in p.values()) / n
    print("%-20s wall %.3f ms   kernels %.3f ms in %d launches   top: %s" % (name, 1e3 * t, tot, nl, ", ".join("%s %.3f/%d" % (k, v[0] / n, v[1] // n) for k, v in ks[:6])))


clouds, (R, p) = scene.config1(seed=20241008 + 4, W=25, pts_per_scan=20000)
Rs, ps_ = synth.perturb_poses((R, p), 4, 0.004, 0.03)
W = 10
base = np.concatenate([np.arange(s0, s0 + W) for s0 in (0, 5, 10, 15)]).astype(np.int32)
fc = np.tile(base, 32); NW = fc.size // W
f = pkg.Factor(ctx, W, NW)
prof("gba_build", lambda: f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc))
poses = pkg.as_poses((Rs[fc], ps_[fc]))
prof("ba_lidar(128 win)", lambda: f.ba_lidar((Rs[fc], ps_[fc])) if hasattr(f, "ba_lidar") else None)
LOCAL = (0.5, 0.0025, [0.25] * 4); GBA = (1.0, 0.01, [0.5] * 4)
prof("hba_add_edge", lambda: pkg.hba_add_edge(ctx, clouds, (Rs[fc], ps_[fc]), W, NW, LOCAL, GBA, 2, 1, frame_cloud=fc), n=2)

# device LM step against the host sweep: the same bits
f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)
a = f.ba_lidar((Rs[fc], ps_[fc]))
os.environ["VX_LM_HOST"] = "1"
f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)
b = f.ba_lidar((Rs[fc], ps_[fc]))
prof("ba_lidar host LDLT", lambda: f.ba_lidar((Rs[fc], ps_[fc])))
del os.environ["VX_LM_HOST"]
for k in a:
    same = np.ascontiguousarray(a[k]).tobytes() == np.ascontiguousarray(b[k]).tobytes()
    print("device vs host LM step: %-10s bitwise equal: %s" % (k, same))
