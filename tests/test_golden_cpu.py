"""The CPU oracle against the committed fixtures of tests/golden/ (made by tests/golden/make_golden.py; they pin the oracle
restatement — the reference ships no vectors).  Tolerances allow for a different libm / compiler."""
import os

import numpy as np
import pytest

import scene
import synth

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name))


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_keys_two_codings_and_fixture(ob):
    z = load("keys.npz")
    for v in z["voxel_sizes"]:
        xyz = z["xyz_%g" % v]; want = z["keys_%g" % v]
        assert np.array_equal(ob.keys_f64(xyz, v), want)
        assert np.array_equal(scene.np_keys(xyz, v), want)          # the independent numpy float32 coding
        assert np.array_equal(ob.keys_f32(xyz.astype(np.float32), v), scene.np_keys(xyz.astype(np.float32).astype(np.float64), v))


def factor_of(z):
    return dict(plvec=z["plvec"], sig=z["sig"], coe=z["coe"], eigval=z["eigval"], eigvec=z["eigvec"], merged=z["merged"])


def test_factor_fixture(ob):
    z = load("factor.npz")
    W = int(z["W"])
    f = ob.Factor(W); f.set_dense(factor_of(z))
    poses = (z["poses_R"], z["poses_p"])
    r = f.residual(poses)
    assert abs(r - float(z["residual"])) <= 1e-12 * abs(float(z["residual"]))
    H, J, c = f.hessian(poses)
    assert rel(H, z["H"]) < 1e-12 and rel(J, z["J"]) < 1e-12 and abs(c - float(z["cost"])) <= 1e-12 * abs(float(z["cost"]))
    f2 = ob.Factor(W); f2.set_dense(factor_of(z))
    start = (z["start_R"], z["start_p"])
    f2.residual(start)
    ba = f2.ba_lidar(start, max_iter=4, thd_num=2)
    assert ba["accept"] == z["ba_accept"].tolist() and ba["rc"] == int(z["ba_rc"])
    assert np.abs(ba["poses"]["R"] - z["ba_poses"]["R"]).max() < 1e-10 and np.abs(ba["poses"]["p"] - z["ba_poses"]["p"]).max() < 1e-10


def test_imu_fixture(ob):
    z = load("imu.npz")
    st, st2, nz = z["states"], z["states2"], z["noise"]
    for i in range(z["samples"].shape[0]):
        pre = ob.imu_pre_new(z["samples"][i], nz, st[i]["bg"], st[i]["ba"])
        for name in pre.dtype.names:
            assert rel(pre[name], z["pres"][i:i + 1][name]) < 1e-12, name
        c, h, g = ob.imu_pre_evaluate(z["pres"][i:i + 1], st[i:i + 1], st2[i + 1:i + 2], True)
        assert abs(c - z["cost"][i]) <= 1e-10 * abs(z["cost"][i]) and rel(h, z["jtj"][i]) < 1e-10 and rel(g, z["gg"][i]) < 1e-10


def test_downsample_fixture(ob):
    z = load("downsample.npz")
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    for name, fn, src, vs in (("voxel", ob.downsample_voxel, z["pts"], 0.25), ("close", ob.downsample_close, z["pts"], 0.25), ("pvec", ob.downsample_pvec, z["pv"], 0.1)):
        out = fn(src, vs)
        assert np.array_equal(out[key(out)].view(np.uint8), z[name].view(np.uint8)), name
