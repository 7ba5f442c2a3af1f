"""Full C4 job (2,000 keyframes x 20,000 points) through vx_hba_global on one GPU: phase timings + kernel profile."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as ge
import synthgen as sg
n_kf = int(os.environ.get("VX_KF", "2000")); n_az = int(os.environ.get("VX_AZ", "625")); tmi = int(os.environ.get("VX_TMI", "3"))
pkg = ge.load_package(); pkg.load_library()
t = time.time(); d = sg.c4_session(n_kf=n_kf, n_az=n_az); print("gen %.1fs, %d points" % (time.time() - t, d["off"][-1]), file=sys.stderr)
LOCAL = (0.5, 0.0025, [0.25] * 4); GBA = (1.0, 0.01, [0.5] * 4)
ctx = pkg.Context(0)
if os.environ.get("VX_PIN"):     # pinned host clouds, as bench.py's e2e leg passes them
    import bench
    xyz_pin, _pin_owner = bench.pinned_like(d["xyz"]); d["xyz"] = xyz_pin
job = pkg.HbaGlobalJob(d["xyz"], d["off"], np.zeros(n_kf, dtype=np.int32), d["start"], pkg.hba_global_params(LOCAL, GBA, 2, tmi))
for k in range(2):
    t = time.perf_counter(); out = job.run(ctx); dt = time.perf_counter() - t
    import hashlib
    sha = lambda a: hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()[:12]
    e1, e2 = out["edges1"], out["edges2"]
    print(json.dumps({"run": k, "s": dt, "stats": out["stats"], "edges1": len(e1), "edges2": len(e2),
                      "sha_edges1": sha(np.stack([e1[f] for f in ("id1", "id2")] , 1)) + sha(e1["tra"]) + sha(e1["cov"]),
                      "sha_edges2": sha(e2["tra"]) + sha(e2["cov"]), "sha_top": sha(out["top_poses"]["p"])}))
ctx.profile(True); out = job.run(ctx); prof = ctx.profile_read(); ctx.profile(False)
tot = sum(v[0] for v in prof.values())
print(json.dumps({"profiled_stats": out["stats"], "kernel_ms_total": tot,
                  "kernels": {k: [round(v[0], 3), v[1]] for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:25]}}))
truth = pkg.as_poses(d["truth"]); start = pkg.as_poses(d["start"])
tk = out["top_kf"]
# relative pose error between consecutive submaps before / after (gauge-free)
def rel_err(ps):
    e = []
    for a, b in zip(range(len(tk) - 1), range(1, len(tk))):
        Ra = ps[a]["R"].reshape(3, 3).T; Rb = ps[b]["R"].reshape(3, 3).T
        Ta = truth[tk[a]]["R"].reshape(3, 3).T; Tb = truth[tk[b]]["R"].reshape(3, 3).T
        e.append(np.linalg.norm(Ra.T @ (ps[b]["p"] - ps[a]["p"]) - Ta.T @ (truth[tk[b]]["p"] - truth[tk[a]]["p"])))
    return float(np.mean(e))
print(json.dumps({"rel_trans_err_start": rel_err(start[tk]), "rel_trans_err_opt": rel_err(out["top_poses"])}))
ctx.close()
