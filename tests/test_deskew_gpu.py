"""IMU de-skew: vx_deskew against the oracle's statement-for-statement loop of IMUEKF::motion_blur
(ekf_imu.hpp:138-163).  Bars: the pose each point is compensated with — exact (incl. the reference's walk quirks: unsorted stamps,
untouched points, the repeatedly compensated first point); coordinates: float, <= 2 ulp (device sin / cos vs libm), >= 99.9 % bitwise."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def _sweep(seed, n, n_imu, sorted_stamps=True, sweep_s=0.1):
    rng = np.random.Generator(np.random.PCG64(seed))
    # IMU-propagated poses across the sweep (what ekf_imu.hpp:57-120 leaves in imu_poses): a turning, accelerating sensor
    t = np.sort(rng.uniform(0, sweep_s, n_imu)); t[0] = 0.0
    w = np.array([0.05, -0.1, 0.9]); acc = np.array([0.8, -0.3, 0.1]); v0 = np.array([1.5, 0.2, 0.0]); p0 = np.array([3.0, -2.0, 1.0])
    poses = np.zeros((n_imu, 22))
    for i in range(n_imu):
        R = synth.rodrigues(w * t[i])
        poses[i, 0] = t[i]; poses[i, 1:10] = R.T.reshape(9)
        poses[i, 10:13] = p0 + v0 * t[i] + 0.5 * acc * t[i] ** 2; poses[i, 13:16] = v0 + acc * t[i]
        poses[i, 16:19] = w + rng.normal(0, 1e-3, 3); poses[i, 19:22] = acc + rng.normal(0, 1e-2, 3)
    te = sweep_s
    xc = (synth.rodrigues(w * te)[None], (p0 + v0 * te + 0.5 * acc * te ** 2)[None])
    ext = (synth.rodrigues([0.01, -0.02, 0.03])[None], np.array([[0.05, -0.02, 0.1]]))
    pts = np.zeros((n, 4), dtype=np.float32)
    pts[:, :3] = rng.uniform(-30, 30, (n, 3))
    st = rng.uniform(-0.002, sweep_s, n)                # a few stamps before the first pose: those points stay untouched
    pts[:, 3] = np.sort(st) if sorted_stamps else st
    return pts, poses, xc, ext


@pytest.mark.parametrize("n,n_imu,sorted_stamps", [(60000, 21, True), (5000, 7, False), (1, 3, True), (300, 1, True)])
def test_deskew_matches_oracle(pkg, ob, ctx, n, n_imu, sorted_stamps):
    pts, poses, xc, ext = _sweep(100 + n, n, n_imu, sorted_stamps)
    ref, ra = ob.deskew(pts, poses, xc, ext)
    p48 = np.zeros(n, dtype=pkg.point48_dtype)
    p48["x"], p48["y"], p48["z"], p48["curvature"] = pts[:, 0], pts[:, 1], pts[:, 2], pts[:, 3]
    ip = np.zeros(n_imu, dtype=pkg.imu_pose_dtype)
    ip["t"] = poses[:, 0]; ip["R"] = poses[:, 1:10]; ip["p"] = poses[:, 10:13]; ip["v"] = poses[:, 13:16]; ip["angvel"] = poses[:, 16:19]; ip["acc"] = poses[:, 19:22]
    out, a = pkg.deskew(ctx, p48, ip, xc, ext)
    assert np.array_equal(a, ra)
    got = np.stack([out["x"], out["y"], out["z"]], axis=1)
    ulp = np.abs(got.view(np.int32).astype(np.int64) - ref[:, :3].view(np.int32).astype(np.int64))
    assert ulp.max() <= 2
    assert (ulp == 0).mean() >= 0.999 or n < 1000
    assert np.array_equal(out["curvature"], pts[:, 3])
    untouched = ra < 0
    assert np.array_equal(got[untouched], pts[untouched, :3])
    if n >= 5000 and sorted_stamps:
        assert untouched.sum() > 0 and (ra >= 0).sum() > 0.9 * n
    if not sorted_stamps:      # the walk stops at the first stamp from the back that is not after the pose: most of an unsorted sweep stays put
        assert 0 < (ra >= 0).sum() < n


def test_deskew_static_plane_becomes_planar(pkg, ctx):
    # planted truth: a wall seen from a moving sensor; each raw point is expressed in the sensor frame AT ITS OWN stamp, so the raw
    # sweep is smeared; after de-skew into the end-of-sweep frame the wall is a plane again
    rng = np.random.Generator(np.random.PCG64(9))
    n, n_imu, T = 20000, 41, 0.1
    w = np.array([0.0, 0.0, 1.2]); v0 = np.array([2.0, 0.5, 0.0]); p0 = np.zeros(3)
    t_imu = np.linspace(0, T, n_imu)
    ip = np.zeros(n_imu, dtype=pkg.imu_pose_dtype)
    for i, t in enumerate(t_imu):
        ip[i]["t"] = t; ip[i]["R"] = synth.rodrigues(w * t).T.reshape(9); ip[i]["p"] = p0 + v0 * t; ip[i]["v"] = v0; ip[i]["angvel"] = w
    stamps = np.sort(rng.uniform(1e-4, T, n)).astype(np.float32)
    wall = np.stack([np.full(n, 12.0), rng.uniform(-5, 5, n), rng.uniform(-1, 2, n)], axis=1)        # world plane x = 12
    raw = np.zeros(n, dtype=pkg.point48_dtype)
    for k in range(n):
        t = float(stamps[k]); R = synth.rodrigues(w * t)
        q = R.T @ (wall[k] - (p0 + v0 * t))
        raw[k]["x"], raw[k]["y"], raw[k]["z"] = q
    raw["curvature"] = stamps
    xc = (synth.rodrigues(w * T)[None], (p0 + v0 * T)[None])
    ident = (np.eye(3)[None], np.zeros((1, 3)))
    out, _ = pkg.deskew(ctx, raw, ip, xc, ident)
    def thickness(P):
        c = P - P.mean(0); return np.sqrt(np.linalg.eigvalsh(c.T @ c / len(P))[0])
    before = thickness(np.stack([raw["x"], raw["y"], raw["z"]], 1).astype(float))
    after = thickness(np.stack([out["x"], out["y"], out["z"]], 1).astype(float))
    assert before > 0.02 and after < 1e-4
