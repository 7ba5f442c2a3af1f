"""BASELINE.json configs[2] (C3) as a chain, reduced size for the oracle: W = 20 scans through the sliding-window map
(cut_voxel_multi -> multi_recut per scan, voxelslam.cpp:1917-1926) -> the LidarFactor tras_opt leaves on the device ->
LI_BA_OptimizerGravity::damping_iter(max_iter = 5) (voxelslam.cpp:1956-1965, voxel_map.hpp:878-975), against the oracle session.

Bars: factor identical in structure (voxels, voxel-frames, N exact), LM accept trace identical, states <= 1e-8, Hess <= 1e-9 rel."""
import numpy as np
import pytest

import c3_scene
import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_az,W", [(190, 20), (120, 10)])
def test_c3_chain_matches_oracle(pkg, ob, ctx, n_az, W):
    c = c3_scene.build(pkg, W=W, n_az=n_az)
    kw = dict(voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2)            # hesai.yaml Odometry / LocalBA
    om = ob.LocalMap(win_size=W, dept_err=0.01, beam_err=0.01, **kw)
    vm = pkg.VoxelMap(ctx, win_size=W, **kw)
    f = pkg.Factor(ctx, W)
    cov = ob.default_cov()
    Rs, ps = synth.states_poses(c["start"])
    win = np.zeros(0, dtype=pkg.pose_dtype)
    for i in range(W):
        pose = (Rs[i:i + 1], ps[i:i + 1])
        pv, pw = om.insert(c["scans"][i], pose, cov); om.recut()
        win = np.concatenate([win, pkg.as_poses(pose)])
        vm.cut_voxel(pv, pw, i)
        vm.recut(win, f)
    of = om.factor(W)
    assert f.num_voxels == of.num_voxels > 200 and f.num_voxel_frames == of.num_voxel_frames
    a = f.ba_li_gravity(c["start"], c["imus"], imu_coef=1e-4, max_iter=5)
    b = of.ba_li(c["start"], c["imus"], imu_coef=1e-4, gravity=True, max_iter=5)
    assert a["accept"] == b["accept"] and len(a["accept"]) >= 1
    for name in ("R", "p", "v", "bg", "ba", "g"):
        assert np.abs(a["states"][name] - b["states"][name]).max() <= 1e-8, name
    assert np.abs(a["hess"] - b["hess"]).max() <= 1e-9 * np.abs(b["hess"]).max()
    assert np.abs(a["resis"] - b["resis"]).max() <= 1e-9 * np.abs(b["resis"]).max()
    # the LiDAR-inertial BA pulled the perturbed window towards the planted trajectory
    e0 = np.abs(c["start"]["p"] - c["truth"]["p"]).max(); e1 = np.abs(a["states"]["p"] - c["truth"]["p"]).max()
    assert 1 in a["accept"] and e1 < e0
    f.free(); vm.free()
