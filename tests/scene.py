"""Seeded synthetic ray-cast scenes for the BASELINE.json configs.

Scene "multi-plane room": axis-aligned box 20 x 15 x 4 m (6 planes) + 4 interior partitions
(2 axis-aligned, 2 rotated 30 / -20 deg about z, 6 x 3 m each) + a 3 x 3 m table top at z = 0.8.
Every surface is a finite rectangle; rays are cast to the nearest rectangle and range noise is added
along the ray.  Scan shapes: Livox-Avia-like (70.4 x 77.2 deg rectangular FoV about body x) and
Hesai-XT32-like (32 rings, -16..+15 deg, 2000 azimuth steps)."""
import numpy as np

from synth import _rodrigues_batch


def _rect(center, u, v, hu, hv):
    u = np.asarray(u, float); v = np.asarray(v, float)
    u /= np.linalg.norm(u); v /= np.linalg.norm(v)
    return dict(c=np.asarray(center, float), u=u, v=v, hu=hu, hv=hv, n=np.cross(u, v))


def room(scale=1.0, extra_partitions=0, seed=0):
    L, Wd, H = 20.0 * scale, 15.0 * scale, 4.0
    R = []
    R.append(_rect([0, 0, 0], [1, 0, 0], [0, 1, 0], L / 2, Wd / 2))          # floor
    R.append(_rect([0, 0, H], [1, 0, 0], [0, 1, 0], L / 2, Wd / 2))          # ceiling
    R.append(_rect([L / 2, 0, H / 2], [0, 1, 0], [0, 0, 1], Wd / 2, H / 2))
    R.append(_rect([-L / 2, 0, H / 2], [0, 1, 0], [0, 0, 1], Wd / 2, H / 2))
    R.append(_rect([0, Wd / 2, H / 2], [1, 0, 0], [0, 0, 1], L / 2, H / 2))
    R.append(_rect([0, -Wd / 2, H / 2], [1, 0, 0], [0, 0, 1], L / 2, H / 2))
    R.append(_rect([4 * scale, 2 * scale, 1.5], [1, 0, 0], [0, 0, 1], 3, 1.5))
    R.append(_rect([-5 * scale, -3 * scale, 1.5], [0, 1, 0], [0, 0, 1], 3, 1.5))
    for ang, c in ((30.0, [-3 * scale, 4 * scale, 1.5]), (-20.0, [6 * scale, -4 * scale, 1.5])):
        a = np.deg2rad(ang)
        R.append(_rect(c, [np.cos(a), np.sin(a), 0], [0, 0, 1], 3, 1.5))
    R.append(_rect([0, -1 * scale, 0.8], [1, 0, 0], [0, 1, 0], 1.5, 1.5))    # table top
    rng = np.random.Generator(np.random.PCG64(seed + 991))
    for _ in range(extra_partitions):
        a = rng.uniform(0, np.pi)
        c = [rng.uniform(-L / 2 + 2, L / 2 - 2), rng.uniform(-Wd / 2 + 2, Wd / 2 - 2), 1.5]
        R.append(_rect(c, [np.cos(a), np.sin(a), 0], [0, 0, 1], rng.uniform(1.5, 3), 1.5))
    return R


def avia_dirs(n, rng):
    az = np.deg2rad(rng.uniform(-35.2, 35.2, n)); el = np.deg2rad(rng.uniform(-38.6, 38.6, n))
    return np.stack([np.cos(el) * np.cos(az), np.cos(el) * np.sin(az), np.sin(el)], axis=1)


def xt32_dirs(n_az=2000):
    el = np.deg2rad(np.arange(-16, 16, 1.0)); az = np.linspace(-np.pi, np.pi, n_az, endpoint=False)
    E, A = np.meshgrid(el, az, indexing="ij")
    return np.stack([np.cos(E) * np.cos(A), np.cos(E) * np.sin(A), np.sin(E)], axis=1).reshape(3, -1).T.copy() \
        if False else np.stack([(np.cos(E) * np.cos(A)).ravel(), (np.cos(E) * np.sin(A)).ravel(), np.sin(E).ravel()], axis=1)


def raycast(rects, origin, dirs_world):
    """nearest hit range per ray (inf when none)"""
    best = np.full(dirs_world.shape[0], np.inf)
    for r in rects:
        denom = dirs_world @ r["n"]
        with np.errstate(divide="ignore", invalid="ignore"):
            t = ((r["c"] - origin) @ r["n"]) / denom
        hit = origin[None, :] + t[:, None] * dirs_world - r["c"][None, :]
        ok = (np.abs(denom) > 1e-9) & (t > 1e-6) & (np.abs(hit @ r["u"]) <= r["hu"]) & (np.abs(hit @ r["v"]) <= r["hv"])
        best = np.where(ok & (t < best), t, best)
    return best


def scan(rects, R, p, dirs_body, rng, range_noise=0.02, blind=0.5):
    """-> (n,3) float32 points in the BODY frame (what the reference stores per keyframe)."""
    dw = dirs_body @ R.T
    t = raycast(rects, p, dw)
    ok = np.isfinite(t) & (t > blind)
    t = t[ok] + rng.normal(0, range_noise, ok.sum())
    return (dirs_body[ok] * t[:, None]).astype(np.float32)


def arc_trajectory(W, seed, step=1.0, yaw_step_deg=4.0, start=(-7.0, 0.5, 1.2), yaw0=-0.35):
    rng = np.random.Generator(np.random.PCG64(seed))
    yaw = yaw0 + np.deg2rad(yaw_step_deg) * np.arange(W) + rng.normal(0, 0.005, W)
    rv = np.zeros((W, 3)); rv[:, 2] = yaw; rv[:, :2] = rng.normal(0, 0.01, (W, 2))
    R = _rodrigues_batch(rv)
    p = np.zeros((W, 3)); p[0] = start
    for i in range(1, W):
        p[i] = p[i - 1] + step * np.array([np.cos(yaw[i]), np.sin(yaw[i]), 0.0]) + rng.normal(0, 0.01, 3)
    return R, p


def config1(seed=20241008 + 1, W=10, pts_per_scan=24000, shape="avia"):
    """C1: 10 poses on an arc, Avia-shaped scans, returns clouds (body frame, float32) + true poses."""
    rng = np.random.Generator(np.random.PCG64(seed))
    rects = room()
    R, p = arc_trajectory(W, seed + 17)
    clouds = []
    for i in range(W):
        d = avia_dirs(pts_per_scan, rng) if shape == "avia" else xt32_dirs()
        clouds.append(scan(rects, R[i], p[i], d, rng))
    return clouds, (R, p)


def np_keys(xyz, vs):
    """numpy restatement of the key rule (second, independent coding of A1; float32 arithmetic)."""
    loc = (np.asarray(xyz, dtype=np.float64) / vs).astype(np.float32)
    loc = np.where(loc < 0, loc - np.float32(1.0), loc).astype(np.float32)
    return loc.astype(np.int64)


def synthetic_plane_map(points_world, voxel_size=1.0, min_eigen_value=0.0025, node_dtype=None):
    """A layer-0 voxel-plane map (root voxels only, every root a leaf) from world points, in the flattened form
    vx_planemap_create takes: per occupied cell the centroid / smallest-eigenvector plane, plane_var = 1e-6 I.
    Bench input only (no parity claim): the real map comes from the host-side OctoTree."""
    k = np_keys(points_world, voxel_size)
    uk, inv = np.unique(k, axis=0, return_inverse=True)
    inv = inv.reshape(-1)
    n = uk.shape[0]
    cnt = np.bincount(inv, minlength=n).astype(np.float64)
    cen = np.stack([np.bincount(inv, weights=points_world[:, a], minlength=n) for a in range(3)], axis=1) / cnt[:, None]
    d = points_world - cen[inv]
    cov = np.zeros((n, 3, 3))
    for a in range(3):
        for b in range(3):
            cov[:, a, b] = np.bincount(inv, weights=d[:, a] * d[:, b], minlength=n) / cnt
    lam, U = np.linalg.eigh(cov)
    nodes = np.zeros(n, dtype=node_dtype)
    nodes["center"] = (uk + 0.5) * voxel_size
    nodes["quater_length"] = voxel_size / 4
    nodes["octo_state"] = 0
    nodes["child"] = -1
    nodes["is_plane"] = ((cnt > 5) & (lam[:, 0] < min_eigen_value)).astype(np.int32)
    nodes["radius"] = lam[:, 2]
    nodes["p_center"] = cen
    nodes["p_normal"] = U[:, :, 0]
    nodes["plane_var"] = (np.eye(6) * 1e-6).reshape(36)[None, :]
    return uk, np.arange(n, dtype=np.int32), nodes
