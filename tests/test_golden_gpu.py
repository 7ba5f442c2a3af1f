"""The CUDA path against the committed fixtures of tests/golden/ (numbers that exist independently of the kernels)."""
import os

import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return np.load(os.path.join(G, name))


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_keys(pkg, ctx):
    z = load("keys.npz")
    for v in z["voxel_sizes"]:
        assert np.array_equal(pkg.keys(ctx, z["xyz_%g" % v], v), z["keys_%g" % v])


def test_factor(pkg, ctx):
    z = load("factor.npz")
    W = int(z["W"])
    fac = dict(plvec=z["plvec"], sig=z["sig"], coe=z["coe"], eigval=z["eigval"], eigvec=z["eigvec"], merged=z["merged"])
    poses = (z["poses_R"], z["poses_p"])
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    r = f.residual(poses)
    assert abs(r - float(z["residual"])) <= 1e-9 * abs(float(z["residual"]))
    ev, _, mrg = f.download_cache()
    scale = (np.abs(z["cache_merged"]["P"]).max(axis=1) / z["cache_merged"]["N"]).max()
    assert np.abs(ev - z["cache_eigval"]).max() <= 1e-12 * scale and np.array_equal(mrg["N"], z["cache_merged"]["N"])
    H, J, c = f.hessian(poses)
    assert rel(H, z["H"]) < 1e-9 and rel(J, z["J"]) < 1e-9 and abs(c - float(z["cost"])) <= 1e-9 * abs(float(z["cost"]))
    f.upload_dense(fac)
    start = (z["start_R"], z["start_p"])
    f.residual(start)
    ba = f.ba_lidar(start, max_iter=4, thd_num=2)
    assert ba["accept"][0] == z["ba_accept"].tolist()
    assert np.abs(ba["poses"]["R"] - z["ba_poses"]["R"]).max() < 1e-8 and np.abs(ba["poses"]["p"] - z["ba_poses"]["p"]).max() < 1e-8
    assert rel(ba["hess"][0], z["ba_hess"]) < 1e-9
    f.free()


def test_downsample(pkg, ctx):
    z = load("downsample.npz")
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    for name, fn, src, vs in (("voxel", pkg.downsample_voxel, z["pts"], 0.25), ("close", pkg.downsample_close, z["pts"], 0.25), ("pvec", pkg.downsample_pvec, z["pv"], 0.1)):
        out = fn(ctx, src, vs)
        assert np.array_equal(out[key(out)].view(np.uint8), z[name].view(np.uint8)), name
