"""B4 LI_BA_Optimizer / B5 LI_BA_OptimizerGravity ::damping_iter (voxel_map.hpp:504-976) over the device LidarFactor,
against the oracle.  Tolerances: accept/reject sequence identical; states <= 1e-8 abs; Hess <= 1e-9 rel."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def _problem(pkg, seed, W, V, rot=0.004, tra=0.03, vel=0.02, g_tilt=0.0):
    states, segs = synth.imu_trajectory(seed, W, gyr_sigma=1e-3, acc_sigma=1e-2)
    fac = synth.random_factor(seed + 1, V, W, poses=synth.states_poses(states), occupancy=0.7, min_frames=2)
    nz = pkg.imu_noise(0.01, 0.1, 1e-4, 1e-3)
    imus = np.concatenate([pkg.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"]) for i, s in enumerate(segs)])
    rng = np.random.Generator(np.random.PCG64(seed + 2))
    start = states.copy()
    for i in range(1, W):
        R = start[i]["R"].reshape(3, 3).T @ synth.rodrigues(rng.normal(0, rot, 3))
        start[i]["R"] = R.T.reshape(9)
        start[i]["p"] += rng.normal(0, tra, 3); start[i]["v"] += rng.normal(0, vel, 3)
    if g_tilt:
        start["g"] = (synth.rodrigues([g_tilt, -g_tilt, 0]) @ states[0]["g"])[None, :]
    return fac, states, start, imus


def _rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def _check(a, b, truth, start):
    assert a["accept"] == b["accept"]
    for name in ("R", "p", "v", "bg", "ba", "g"):
        assert np.abs(a["states"][name] - b["states"][name]).max() <= 1e-8, name
    assert _rel(a["hess"], b["hess"]) < 1e-9
    for name in ("dbg", "dba", "dbg_buf", "dba_buf"):
        assert np.abs(a["imus"][name] - b["imus"][name]).max() <= 1e-8, name
    # sanity: the optimizer moved the window towards the truth
    e0 = np.abs(start["p"] - truth["p"]).max(); e1 = np.abs(a["states"]["p"] - truth["p"]).max()
    assert 1 in a["accept"] and e1 < e0


@pytest.mark.parametrize("W,V", [(10, 400), (5, 150)])
def test_ba_li_parity(pkg, ob, ctx, W, V):
    fac, truth, start, imus = _problem(pkg, 700 + W, W, V)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    of = ob.Factor(W); of.set_dense(fac)
    a = f.ba_li(start, imus, imu_coef=1e-4)
    b = of.ba_li(start, imus, imu_coef=1e-4, gravity=False)
    _check(a, b, truth, start)
    assert a["hess"].shape == (15 * W, 15 * W)
    # the device cache follows the oracle's stateful cache (last residual pass)
    ev, evec, mrg = f.download_cache()
    o = of.get_dense()
    scale = (np.abs(mrg["P"]).max(axis=1) / mrg["N"]).max()
    assert np.abs(ev - o["eigval"]).max() <= 1e-12 * scale
    f.free()


@pytest.mark.parametrize("W,V,iters", [(20, 600, 5), (10, 300, 2)])
def test_ba_li_gravity_parity(pkg, ob, ctx, W, V, iters):
    fac, truth, start, imus = _problem(pkg, 900 + W, W, V, g_tilt=np.deg2rad(2.0))
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    of = ob.Factor(W); of.set_dense(fac)
    a = f.ba_li_gravity(start, imus, imu_coef=1e-4, max_iter=iters)
    b = of.ba_li(start, imus, imu_coef=1e-4, gravity=True, max_iter=iters)
    _check(a, b, truth, start)
    assert a["hess"].shape == (15 * W + 3, 15 * W + 3)
    assert np.abs(a["resis"] - b["resis"]).max() <= 1e-9 * np.abs(b["resis"]).max()
    assert (a["states"]["g"] == a["states"]["g"][0]).all()          # one gravity shared by all frames (:921-930)
    f.free()


def test_ba_li_rejects_batched_factor(pkg, ctx):
    f = pkg.Factor(ctx, 5, 2)
    x = np.zeros(5, dtype=pkg.imu_state_dtype); pr = np.zeros(4, dtype=pkg.imu_pre_dtype)
    with pytest.raises(pkg.VxError):
        f.ba_li(x, pr)
    f.free()
