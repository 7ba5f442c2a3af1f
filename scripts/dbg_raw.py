import sys; sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import __graft_entry__ as ge
import oracle_bind as ob
import test_map_gpu as T
pkg = ge.load_package(); ctx = pkg.Context(0)
W = 6
dev = pkg.VoxelMap(ctx, win_size=W, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2)
ref = ob.LocalMap(win_size=W, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2, dept_err=0.02, beam_err=0.05)
f = pkg.Factor(ctx, W)
cov = ob.default_cov()
ident = (np.eye(3)[None], np.zeros((1, 3)))
win = np.zeros(0, dtype=pkg.pose_dtype)
def stats(tag):
    ds, dn = dev.export(); rs, rn = ref.export_states()
    di, ri = T._index(ds), T._index(rs)
    a = np.array([di[k] for k in ri]); b = np.array([ri[k] for k in ri])
    d, r = ds[a], rs[b]
    for cl in ("pcr_add", "pcr_fix"):
        sc = np.maximum(np.abs(r[cl]["P"]).max(axis=1, keepdims=True), 1e-300)
        e = (np.abs(d[cl]["P"] - r[cl]["P"]) / sc).max(axis=1)
        i = int(np.argmax(e))
        print(tag, cl, "max rel", e.max(), "node", r[i]["root_key"], r[i]["layer"], "N", r[cl]["N"][i], d[cl]["N"][i], "fixN", r["pcr_fix"]["N"][i])
for k, (xyz, pose) in enumerate(T._scans(W + 3, dirs=500, seed=21)):
    pv, pw = ref.insert(xyz, pose, cov)
    win = np.concatenate([win, pkg.as_poses(pose)])
    dev.cut_voxel_xyz(xyz, ident, 0.02, 0.05, pose, cov, win.shape[0] - 1)
    stats("ins%d" % k)
    ref.recut(); dev.recut(win, f)
    stats("rec%d" % k)
    if ref.margi():
        dev.margi(win, f); dev.shift_window(1); win = win[1:].copy()
        stats("mar%d" % k)
