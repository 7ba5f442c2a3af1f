#!/usr/bin/env python
"""Wall-clock of vx_factor_hessian (host poses in, host Hess/JacT/cost out) on the resident bench factor."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import __graft_entry__ as ge
import bench
pkg = ge.load_package(); pkg.load_library()
args = type("A", (), {"windows": bench.N_WINDOWS, "voxels_per_window": bench.VOX_PER_WIN})()
fac = bench.build_workload(args, 0)
ctx = pkg.Context(0)
f = pkg.Factor(ctx, bench.WIN, args.windows); f.upload_csr(fac)
ps, o0 = bench.pinned_like(pkg.as_poses(fac["poses"]))
jl = 6 * bench.WIN
Hh, o1 = bench.pinned_like(np.zeros((args.windows, jl, jl))); Jh, o2 = bench.pinned_like(np.zeros((args.windows, jl))); ch, o3 = bench.pinned_like(np.zeros(args.windows))
for _ in range(3):
    f.hessian_raw(ps, Hh, Jh, ch)
t0 = time.perf_counter()
for _ in range(30):
    f.hessian_raw(ps, Hh, Jh, ch)
t = (time.perf_counter() - t0) / 30
H, J, c = f.hessian(fac["poses"])
print("e2e %.4f ms  checksum H %.17g J %.17g c %.17g  | pinned-out max diff %.3g" % (1e3 * t, np.abs(Hh).sum(), np.abs(Jh).sum(), ch.sum(), np.abs(np.asarray(H).reshape(Hh.shape) - Hh).max()))
