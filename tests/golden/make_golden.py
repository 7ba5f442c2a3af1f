"""Regenerates the committed fixtures under tests/golden/ from the CPU oracle (oracle/, built by __graft_entry__.build_oracle).

The reference ships no golden vectors, known-answer tests or fixtures for this path and cannot be compiled
here (Eigen / PCL / ROS / GTSAM absent), so these files pin the ORACLE RESTATEMENT, not the reference: parity stays
"unpinned" in the sense of DESIGN.md §2.  What they buy: (1) the oracle cannot drift silently between rounds (CPU test),
(2) the CUDA path is checked against numbers that were committed before the kernel under test was written (GPU test),
(3) the key rule is pinned by an independent numpy-float32 coding (tests/scene.py: np_keys), not by the oracle alone.

    python tests/golden/make_golden.py        # rewrites the .npz files; inputs are seeded (PCG64, base 20241008)
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_bind as ob   # noqa: E402
import scene               # noqa: E402
import synth               # noqa: E402

SEED = 20241008


def keys_fixture():
    rng = np.random.Generator(np.random.PCG64(SEED))
    vs = np.array([0.5, 1.0, 0.1, 0.3])
    cells = rng.integers(-2000, 2000, 1000).astype(np.float64)
    parts = [cells, np.nextafter(cells, np.inf), np.nextafter(cells, -np.inf), cells + 0.5, rng.uniform(-250, 250, 1000),
             np.array([0.0, -0.0, -1.0, 1.0, -1e-30, 1e-30, -0.9999999, -1.0000001, 16777216.0, -16777217.0, 5e-324, -5e-324])]
    x = np.concatenate(parts)
    out = {}
    for v in vs:
        xyz = np.stack([x * v, np.roll(x, 1) * v, np.roll(x, 2) * v], axis=1)
        k_np = scene.np_keys(xyz, v)                       # independent coding (numpy float32 arithmetic)
        k_or = ob.keys_f64(xyz, v)
        assert np.array_equal(k_np, k_or), "the two codings of the key rule disagree"
        out["xyz_%g" % v] = xyz; out["keys_%g" % v] = k_or
    np.savez_compressed(os.path.join(HERE, "keys.npz"), voxel_sizes=vs, **out)


def factor_fixture():
    W, V = 5, 48
    fac = synth.random_factor(SEED + 1, V, W, occupancy=0.7, min_frames=1)
    poses = synth.perturb_poses(fac["poses"], SEED + 2, 0.01, 0.05)
    f = ob.Factor(W); f.set_dense(fac)
    r = f.residual(poses)
    cache = f.get_dense()
    H, J, c = f.hessian(poses)
    start = synth.perturb_poses(fac["poses"], SEED + 3, 0.004, 0.03)
    f2 = ob.Factor(W); f2.set_dense(fac)
    f2.residual(start)
    ba = f2.ba_lidar(start, max_iter=4, thd_num=2)
    np.savez_compressed(os.path.join(HERE, "factor.npz"), W=W, plvec=fac["plvec"], sig=fac["sig"], coe=fac["coe"], eigval=fac["eigval"], eigvec=fac["eigvec"],
                        merged=fac["merged"], poses_R=poses[0], poses_p=poses[1], residual=r, cache_eigval=cache["eigval"], cache_merged=cache["merged"],
                        H=H, J=J, cost=c, start_R=start[0], start_p=start[1], ba_poses=ba["poses"], ba_hess=ba["hess"], ba_resis=ba["resis"],
                        ba_accept=np.array(ba["accept"], dtype=np.int32), ba_rc=ba["rc"])


def imu_fixture():
    states, segs = synth.imu_trajectory(SEED + 4, 4, gyr_sigma=1e-3, acc_sigma=1e-2)
    nz = np.zeros(1, dtype=ob.imu_noise_dtype)
    nz["scale_gravity"] = 1.0
    nz["noise_meas"][0] = np.diag([0.01] * 3 + [0.1] * 3).reshape(-1); nz["noise_walk"][0] = np.diag([1e-4] * 3 + [1e-3] * 3).reshape(-1)
    pres = np.concatenate([ob.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"]) for i, s in enumerate(segs)])
    s2 = states.copy(); s2["p"] += 0.01; s2["v"] -= 0.02
    ev = [ob.imu_pre_evaluate(pres[i:i + 1], states[i:i + 1], s2[i + 1:i + 2], True) for i in range(len(segs))]
    np.savez_compressed(os.path.join(HERE, "imu.npz"), states=states, states2=s2, noise=nz, samples=np.stack(segs), pres=pres,
                        cost=np.array([e[0] for e in ev]), jtj=np.stack([e[1] for e in ev]), gg=np.stack([e[2] for e in ev]))


def downsample_fixture():
    rng = np.random.Generator(np.random.PCG64(SEED + 5))
    n = 1200
    pts = np.zeros(n, dtype=ob.point48_dtype)
    pts["x"] = rng.uniform(-4, 4, n); pts["y"] = rng.uniform(-3, 3, n); pts["z"] = rng.normal(0, 0.05, n)
    pts["intensity"] = rng.uniform(0, 255, n); pts["curvature"] = np.arange(n) * 1e-5
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    dv = ob.downsample_voxel(pts, 0.25); dc = ob.downsample_close(pts, 0.25)
    pv = ob.var_init(np.stack([pts["x"], pts["y"], pts["z"] + 2], axis=1), 0.02, 0.05)
    dp = ob.downsample_pvec(pv, 0.1)
    np.savez_compressed(os.path.join(HERE, "downsample.npz"), pts=pts, voxel=dv[key(dv)], close=dc[key(dc)], pv=pv, pvec=dp[key(dp)])


if __name__ == "__main__":
    keys_fixture(); factor_fixture(); imu_fixture(); downsample_fixture()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
