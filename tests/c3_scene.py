"""C3 of BASELINE.json: "dense indoor" room (x2 scale, 12 extra partitions), W = 20 frames on an analytic IMU
trajectory, XT32-shaped scans, synthetic 200 Hz IMU, gravity tilted 2 deg.  Shared by tests/test_c3_gpu.py and bench / scripts."""
import numpy as np

import scene
import synth


def build(pkg, W=20, n_az=1563, seed=20241008 + 3, rot=0.004, tra=0.03, vel=0.02, g_tilt=np.deg2rad(2.0)):
    """-> dict(scans [W x (n,3) float32], truth states, start states, imus, segs)"""
    states, segs = synth.imu_trajectory(seed, W, gyr_sigma=1e-3, acc_sigma=1e-2)
    states["p"] += np.array([-8.0, -3.0, 1.3])           # into the room, sensor 1.3 m above the floor
    rects = scene.room(scale=2.0, extra_partitions=12, seed=seed)
    rng = np.random.Generator(np.random.PCG64(seed + 11))
    R, p = synth.states_poses(states)
    dirs = scene.xt32_dirs(n_az)
    scans = [scene.scan(rects, R[i], p[i], dirs, rng, range_noise=0.01, blind=0.7) for i in range(W)]
    nz = pkg.imu_noise(0.01, 0.1, 1e-4, 1e-3)
    imus = np.concatenate([pkg.imu_pre_new(s, nz, states[i]["bg"], states[i]["ba"]) for i, s in enumerate(segs)])
    start = states.copy()
    for i in range(1, W):
        Ri = start[i]["R"].reshape(3, 3).T @ synth.rodrigues(rng.normal(0, rot, 3))
        start[i]["R"] = Ri.T.reshape(9)
        start[i]["p"] += rng.normal(0, tra, 3); start[i]["v"] += rng.normal(0, vel, 3)
    if g_tilt:
        start["g"] = (synth.rodrigues([g_tilt, -g_tilt, 0]) @ states[0]["g"])[None, :]
    return dict(scans=scans, truth=states, start=start, imus=imus, segs=segs)
