"""Keyframe assembly of a stored session on the device (voxelslam.cpp:345-378) against the oracle, fed
through the session files both sides would read: same cells, same counts, bit-identical float centroids."""
import numpy as np
import pytest

import scene
import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("win,pts,voxel", [(10, 3000, 1.0), (5, 8000, 0.5), (1, 2000, 1.0)])
def test_keyframe_assemble_from_session_files(pkg, ob, ctx, tmp_path, win, pts, voxel):
    clouds, (R, p) = scene.config1(seed=77 + win, W=win, pts_per_scan=pts)
    # write the session the way the reference stores it, read it back the way it loads it
    st = np.zeros(win, dtype=pkg.scan_state_dtype)
    st["t"] = np.arange(win) * 0.1; st["p"] = p; st["R"] = np.transpose(R, (0, 2, 1)).reshape(win, 9); st["g"] = [0, 0, -9.8]
    pkg.write_lidarstate(str(tmp_path / "alidarState.txt"), st)
    for i, c in enumerate(clouds):
        pkg.write_pcd_xyzi(str(tmp_path / ("%d.pcd" % i)), np.concatenate([c.astype(np.float32), np.zeros((c.shape[0], 1), np.float32)], axis=1))
    rd = pkg.read_lidarstate(str(tmp_path / "alidarState.txt"))
    Rr = np.transpose(rd["R"].reshape(win, 3, 3), (0, 2, 1)); pr = rd["p"]
    cl = [pkg.read_pcd_xyzi(str(tmp_path / ("%d.pcd" % i)))[:, :3] for i in range(win)]
    assert all(np.array_equal(a, b.astype(np.float32)) for a, b in zip(cl, clouds))
    ref = ob.keyframe_assemble(cl, (Rr, pr), voxel)
    got = pkg.keyframe_assemble(ctx, cl, (Rr, pr), voxel)
    assert got.shape == ref.shape and ref.shape[0] > 200
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    g = got[key(got)]; r = ref[key(ref)]
    for fld in ("x", "y", "z", "intensity", "curvature"):
        assert np.array_equal(g[fld].view(np.uint32), r[fld].view(np.uint32)), fld
    assert int(got["curvature"].sum()) == sum(c.shape[0] for c in cl)      # down_sampling_voxel leaves the cell count there


def test_keyframe_assemble_empty_and_capacity(pkg, ctx):
    R, p = synth.random_poses(4, 3)
    empty = [np.zeros((0, 3), np.float32)] * 3
    assert pkg.keyframe_assemble(ctx, empty, (R, p), 1.0).shape[0] == 0
