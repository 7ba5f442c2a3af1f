"""GPU parity for VOXEL_SLAM::HBA_add_edge on BASELINE.json config 1:
synthetic multi-plane room, 10-frame LiDAR-only BA, Avia-shaped scans, 0.5 m voxels.

Bars: converged poses <= 1e-8 abs after identical outer/LM iteration counts; edges: same set, cov <= 1e-7 rel
(1/|hess| of a 1e-9-rel Hessian), relative poses <= 1e-8."""
import numpy as np
import pytest

import scene
import synth

pytestmark = pytest.mark.gpu

LOCAL = (0.5, 0.0025, [0.25] * 4)       # config-1 voxel_size 0.5, avia.yaml min_eigen_value / plane_eigen_value_thre 4 -> 1/4
GBA = (1.0, 0.01, [0.5] * 4)            # hesai.yaml GBA block


def pose_err(a, b):
    return max(np.max(np.abs(a["R"] - b["R"])), np.max(np.abs(a["p"] - b["p"])))


def check_edges(eg, eo):
    assert len(eg) == len(eo)
    assert np.array_equal(eg["id1"], eo["id1"]) and np.array_equal(eg["id2"], eo["id2"])
    if len(eg):
        assert np.max(np.abs(eg["rot"] - eo["rot"])) < 1e-8 and np.max(np.abs(eg["tra"] - eo["tra"])) < 1e-8
        assert np.max(np.abs(eg["cov"] - eo["cov"]) / np.abs(eo["cov"])) < 1e-6


@pytest.mark.parametrize("max_iter,pts", [(1, 24000), (3, 8000)])
def test_hba_config1(pkg, ob, ctx, max_iter, pts):
    clouds, (R, p) = scene.config1(pts_per_scan=pts)
    W = len(clouds)
    start = synth.perturb_poses((R, p), 3, np.deg2rad(0.5), 0.05)
    ref = ob.hba_add_edge(clouds, start, LOCAL, GBA, 2, max_iter, ids=np.arange(100, 100 + W), mps=np.zeros(W))
    out = pkg.hba_add_edge(ctx, clouds, start, W, 1, LOCAL, GBA, 2, max_iter, ids=np.arange(100, 100 + W), mps=np.zeros(W))
    assert ref["n_edges"] >= 0 and out["status"][0] == 0
    assert out["n_voxels"][0] == ref["n_voxels"][-1]
    assert abs(out["resis"][0, 0] - ref["resis_first"][-1]) <= 1e-8 * ref["resis_first"][-1]
    assert abs(out["resis"][0, 1] - ref["resis_last"][-1]) <= 1e-8 * ref["resis_last"][-1]
    assert pose_err(out["poses"], ref["poses"]) < 1e-8
    check_edges(out["edges"][0], ref["edges"])
    truth = ob.as_poses((R, p))
    assert pose_err(out["poses"], truth) < 0.5 * pose_err(ob.as_poses(start), truth)     # the BA converged toward the planted truth


def test_hba_batched_bottom_level(pkg, ob, ctx):
    # bottom-level schedule: windows of 10 keyframes, stride 5 (voxelslam.cpp:3033-3034, :3137-3138), max_iter = 1
    clouds, (R, p) = scene.config1(W=25, pts_per_scan=5000)
    W = 10
    starts = [0, 5, 10, 15]
    NW = len(starts)
    Rs, ps_ = synth.perturb_poses((R, p), 4, 0.004, 0.03)
    fc = np.concatenate([np.arange(s, s + W) for s in starts]).astype(np.int32)
    out = pkg.hba_add_edge(ctx, clouds, (Rs[fc], ps_[fc]), W, NW, LOCAL, GBA, 2, 1, frame_cloud=fc, ids=fc, mps=np.zeros(fc.size))
    for b, s in enumerate(starts):
        ref = ob.hba_add_edge([clouds[i] for i in range(s, s + W)], (Rs[s:s + W], ps_[s:s + W]), LOCAL, GBA, 2, 1, ids=np.arange(s, s + W), mps=np.zeros(W))
        assert out["n_voxels"][b] == ref["n_voxels"][-1]
        assert pose_err(out["poses"][b * W:(b + 1) * W], ref["poses"]) < 1e-8
        check_edges(out["edges"][b], ref["edges"])


def test_hba_merged_submap_cloud(pkg, ob, ctx):
    # voxelslam.cpp:2954-2985: clouds moved into keyframe 0's frame with the optimized poses, intensity = session id,
    # down_sampling_voxel(voxel_size / 8).  Same cells, same counts, bit-identical float centroids (order: sorted by cell).
    clouds, (R, p) = scene.config1(W=15, pts_per_scan=4000)
    W = 10
    start = synth.perturb_poses((R, p), 3, np.deg2rad(0.5), 0.05)
    mps = np.arange(15) % 3
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    for s0 in (0, 5):
        sl = slice(s0, s0 + W)
        ref = ob.hba_add_edge(clouds[sl], (start[0][sl], start[1][sl]), LOCAL, GBA, 2, 1, ids=np.arange(W), mps=mps[sl], want_cloud=True)
        # the cloud is built from the poses HBA_add_edge leaves behind: feed the oracle's optimized poses to the device path
        got = pkg.hba_merge_clouds(ctx, clouds[sl], ref["poses"], W, 1, LOCAL[0], mps=mps[sl])[0]
        r = ref["merged"]
        assert got.shape == r.shape and r.shape[0] > 1000
        g = got[key(got)]; r = r[key(r)]
        for fld in ("x", "y", "z", "intensity", "curvature"):
            assert np.array_equal(g[fld].view(np.uint32), r[fld].view(np.uint32)), fld
        assert int(got["curvature"].sum()) == sum(c.shape[0] for c in clouds[sl])
    # two windows in one call == each alone
    fc = np.concatenate([np.arange(0, W), np.arange(5, 5 + W)]).astype(np.int32)
    both = pkg.hba_merge_clouds(ctx, clouds, (start[0][fc], start[1][fc]), W, 2, LOCAL[0], frame_cloud=fc, mps=mps[fc])
    for b, s0 in enumerate((0, 5)):
        sl = slice(s0, s0 + W)
        one = pkg.hba_merge_clouds(ctx, clouds[sl], (start[0][sl], start[1][sl]), W, 1, LOCAL[0], mps=mps[sl])[0]
        assert np.array_equal(both[b].view(np.uint8), one.view(np.uint8))
    # ragged input: an empty scan and a one-point scan inside the window (the key fields are sized from the clouds' bounding boxes)
    rag = [c for c in clouds[:W]]
    rag[3] = rag[3][:0]; rag[6] = rag[6][:1]
    sl = slice(0, W)
    ref = ob.hba_add_edge(clouds[sl], (start[0][sl], start[1][sl]), LOCAL, GBA, 2, 1, ids=np.arange(W), mps=mps[sl], want_cloud=True)
    got = pkg.hba_merge_clouds(ctx, rag, ref["poses"], W, 1, LOCAL[0], mps=mps[sl])[0]
    assert int(got["curvature"].sum()) == sum(c.shape[0] for c in rag)
    full = pkg.hba_merge_clouds(ctx, clouds[sl], ref["poses"], W, 1, LOCAL[0], mps=mps[sl])[0]
    assert 0 < got.shape[0] <= full.shape[0]
