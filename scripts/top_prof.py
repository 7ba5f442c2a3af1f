#!/usr/bin/env python
"""Top-level GBA window (W = 399, 2394 unknowns): vx_ba_lidar wall clock and kernel times, device LM step vs VX_LM_HOST=1."""
import os, sys, time
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
res = {}
for mode in ("device", "host"):
    if mode == "host":
        os.environ["VX_LM_HOST"] = "1"
    ft = pkg.Factor(ctx, Wt); ft.upload_csr(top)
    ft.residual(start); r = ft.ba_lidar(start, max_iter=3)
    ft.residual(start)
    ctx.profile(True)
    t0 = time.perf_counter()
    r = ft.ba_lidar(start, max_iter=3)
    t = time.perf_counter() - t0
    pr = ctx.profile_read(); ctx.profile(False)
    print("%-6s wall %.2f ms accept %s | %s" % (mode, 1e3 * t, r["accept"][0], ", ".join("%s %.3f/%d" % (k, v[0], v[1]) for k, v in sorted(pr.items(), key=lambda kv: -kv[1][0])[:8])))
    res[mode] = r
    ft.free()
a, b = res["device"], res["host"]
print("accept equal", a["accept"] == b["accept"], " pose diff R %.3e p %.3e  hess rel %.3e" % (
    np.abs(a["poses"]["R"] - b["poses"]["R"]).max(), np.abs(a["poses"]["p"] - b["poses"]["p"]).max(),
    np.abs(a["hess"] - b["hess"]).max() / np.abs(b["hess"]).max()))
