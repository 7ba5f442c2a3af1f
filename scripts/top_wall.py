#!/usr/bin/env python
"""Top-level GBA window (W = 399): wall-clock breakdown of vx_ba_lidar — with / without the Hessian snapshot, per max_iter."""
import os, sys, time, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
import synth
pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
Wt, Vt = 399, 20000
top = synth.batched_factor_csr(20241008 + 8, 1, Vt, Wt, occupancy=4.0 / Wt, min_frames=2, pts_max=12)
start = synth.perturb_poses(top["poses"], 7, 0.002, 0.02)
ft = pkg.Factor(ctx, Wt); ft.upload_csr(top)
_p = lambda a: a.ctypes.data_as(C.c_void_p)
jl = 6 * Wt
hess = np.ones((1, jl, jl)); resis = np.zeros((1, 2)); conv = np.zeros(1, np.int32); trace = np.zeros((1, 3), np.int32); nit = np.zeros(1, np.int32)
def call(mi, with_hess):
    ps = pkg.as_poses(start).copy()
    ft.residual(start)
    t0 = time.perf_counter()
    rc = ft.lib.vx_ba_lidar(ctx.h, ft.h, _p(ps), _p(hess) if with_hess else None, _p(resis), C.c_int(mi), C.c_int(2), _p(conv), _p(trace), _p(nit))
    t = time.perf_counter() - t0
    assert rc == 0
    return 1e3 * t
for mi in (1, 2, 3):
    for wh in (True, False):
        call(mi, wh)
        ts = sorted(call(mi, wh) for _ in range(5))
        print("max_iter %d hess_out %-5s  wall %.2f ms (min %.2f)" % (mi, wh, ts[2], ts[0]))
t0 = time.perf_counter(); r = ft.ba_lidar(start, max_iter=3); print("python wrapper ba_lidar: %.2f ms" % (1e3 * (time.perf_counter() - t0)))
for name, fn in (("residual", lambda: ft.residual(start)), ("hessian(host out)", lambda: ft.hessian(start))):
    fn(); t0 = time.perf_counter(); fn(); print("%s: %.2f ms" % (name, 1e3 * (time.perf_counter() - t0)))
