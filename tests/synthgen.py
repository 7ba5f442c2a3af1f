"""Large synthetic inputs: a tile world of rooms joined by doors, a loop trajectory through it,
XT32-shaped keyframe clouds ray-cast by tests/synthgen.c (gcc + OpenMP, seeded, thread-count independent).

Bench / test INPUT generation only.  The world: room tiles of 20 x 15 x 4 m on an nx x ny grid, each with the four interior
partitions and the table top of scene.room(); neighbouring tiles share a wall with a 3 m door in its middle; one floor, one
ceiling, four outer walls."""
import ctypes as C
import os
import subprocess

import numpy as np

from synth import _rodrigues_batch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "synthgen.c")
_LIB = os.path.join(_HERE, "libsynthgen.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(_SRC):
        tmp = _LIB + ".%d.tmp" % os.getpid()
        subprocess.run(["gcc", "-O3", "-fopenmp", "-shared", "-fPIC", "-o", tmp, _SRC, "-lm"], check=True)
        os.replace(tmp, _LIB)
    return _LIB


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


TL, TW, TH = 20.0, 15.0, 4.0


def _rect(c, u, v, hu, hv):
    u = np.asarray(u, float); v = np.asarray(v, float)
    return np.concatenate([np.asarray(c, float), u / np.linalg.norm(u), v / np.linalg.norm(v), [hu, hv]])


def tile_world(nx=10, ny=4, door=3.0):
    """-> (n_rects, 11) float64: centre, u, v, half-extents."""
    R = []
    x0, x1 = -TL / 2, TL * nx - TL / 2
    y0, y1 = -TW / 2, TW * ny - TW / 2
    cx, cy = (x0 + x1) / 2, (y0 + y1) / 2
    R.append(_rect([cx, cy, 0], [1, 0, 0], [0, 1, 0], (x1 - x0) / 2, (y1 - y0) / 2))
    R.append(_rect([cx, cy, TH], [1, 0, 0], [0, 1, 0], (x1 - x0) / 2, (y1 - y0) / 2))
    R.append(_rect([x0, cy, TH / 2], [0, 1, 0], [0, 0, 1], (y1 - y0) / 2, TH / 2))
    R.append(_rect([x1, cy, TH / 2], [0, 1, 0], [0, 0, 1], (y1 - y0) / 2, TH / 2))
    R.append(_rect([cx, y0, TH / 2], [1, 0, 0], [0, 0, 1], (x1 - x0) / 2, TH / 2))
    R.append(_rect([cx, y1, TH / 2], [1, 0, 0], [0, 0, 1], (x1 - x0) / 2, TH / 2))
    seg_y = (TW - door) / 4      # half-length of each of the two wall pieces left and right of a door
    seg_x = (TL - door) / 4
    for i in range(nx):
        for j in range(ny):
            ox, oy = TL * i, TW * j
            if i + 1 < nx:       # wall towards tile (i+1, j), door in the middle
                for sgn in (-1, 1):
                    R.append(_rect([ox + TL / 2, oy + sgn * (door / 2 + seg_y), TH / 2], [0, 1, 0], [0, 0, 1], seg_y, TH / 2))
            if j + 1 < ny:
                for sgn in (-1, 1):
                    R.append(_rect([ox + sgn * (door / 2 + seg_x), oy + TW / 2, TH / 2], [1, 0, 0], [0, 0, 1], seg_x, TH / 2))
            R.append(_rect([ox + 4, oy + 2, 1.5], [1, 0, 0], [0, 0, 1], 3, 1.5))
            R.append(_rect([ox - 5, oy - 3, 1.5], [0, 1, 0], [0, 0, 1], 3, 1.5))
            for ang, c in ((30.0, [ox - 3, oy + 4, 1.5]), (-20.0, [ox + 6, oy - 4, 1.5])):
                a = np.deg2rad(ang)
                R.append(_rect(c, [np.cos(a), np.sin(a), 0], [0, 0, 1], 3, 1.5))
            R.append(_rect([ox, oy - 1, 0.8], [1, 0, 0], [0, 1, 0], 1.5, 1.5))
    return np.ascontiguousarray(np.stack(R))


def xt32_dirs(n_az=625):
    el = np.deg2rad(np.arange(-16, 16, 1.0)); az = np.linspace(-np.pi, np.pi, n_az, endpoint=False)
    E, A = np.meshgrid(el, az, indexing="ij")
    return np.ascontiguousarray(np.stack([(np.cos(E) * np.cos(A)).ravel(), (np.cos(E) * np.sin(A)).ravel(), np.sin(E).ravel()], axis=1))


def loop_path(n_kf, nx=10, seed=0, rows=(0, 1), phase=0.0):
    """Keyframe poses along the loop: east through the doors of row rows[0] (y = 0.5), north through the door of the last tile,
    west along row rows[1], south through the door of tile 0.  -> R (n,3,3), p (n,3)."""
    ya, yb = TW * rows[0] + 0.5, TW * rows[1] + 0.5
    xa, xb = -7.0, TL * (nx - 1)
    corners = np.array([[xa, ya], [xb, ya], [xb, yb], [0.0, yb], [0.0, ya], [xa + 1e-3, ya]])
    seg = np.linalg.norm(np.diff(corners, axis=0), axis=1)
    cum = np.concatenate([[0], np.cumsum(seg)])
    total = cum[-1]
    s = (phase + np.arange(n_kf) * (total / n_kf)) % total
    k = np.clip(np.searchsorted(cum, s, side="right") - 1, 0, len(seg) - 1)
    t = (s - cum[k]) / seg[k]
    xy = corners[k] + (corners[k + 1] - corners[k]) * t[:, None]
    d = (corners[k + 1] - corners[k]) / seg[k][:, None]
    yaw = np.unwrap(np.arctan2(d[:, 1], d[:, 0]))
    # round the corners: heading low-pass filtered over ~2 m of travel
    w = max(1, int(round(2.0 / (total / n_kf))))
    ker = np.ones(2 * w + 1) / (2 * w + 1)
    yaw = np.convolve(np.pad(yaw, w, mode="edge"), ker, mode="valid")
    rng = np.random.Generator(np.random.PCG64(seed + 4177))
    rv = np.zeros((n_kf, 3)); rv[:, 2] = yaw + rng.normal(0, 0.003, n_kf); rv[:, :2] = rng.normal(0, 0.005, (n_kf, 2))
    p = np.zeros((n_kf, 3)); p[:, :2] = xy + rng.normal(0, 0.01, (n_kf, 2)); p[:, 2] = 1.2 + rng.normal(0, 0.01, n_kf)
    return _rodrigues_batch(rv), p


def odometry_drift(R, p, seed, rot_deg=0.02, trans=0.005):
    """Odometry-like start poses: every step's relative motion is corrupted by a small random rotation / translation and the
    chain is re-integrated (random walk)."""
    n = R.shape[0]
    rng = np.random.Generator(np.random.PCG64(seed + 733))
    dR = _rodrigues_batch(rng.normal(0, np.deg2rad(rot_deg), (n, 3)))
    dt = rng.normal(0, trans, (n, 3))
    Ro = np.zeros_like(R); po = np.zeros_like(p)
    Ro[0] = R[0]; po[0] = p[0]
    for i in range(1, n):
        rel_R = R[i - 1].T @ R[i]
        rel_p = R[i - 1].T @ (p[i] - p[i - 1])
        Ro[i] = Ro[i - 1] @ rel_R @ dR[i]
        po[i] = po[i - 1] + Ro[i - 1] @ (rel_p + dt[i])
    return Ro, po


def cast(rects, R, p, dirs, seed, range_noise=0.02, blind=0.5, max_range=30.0, az_jitter=None):
    """-> xyz (N,3) float32 body-frame points of all scans back to back, off (n+1,) int64."""
    n = R.shape[0]
    poses = np.ascontiguousarray(np.concatenate([R.reshape(n, 9), p], axis=1))
    dirs = np.ascontiguousarray(dirs, dtype=np.float64)
    nd = dirs.shape[0]
    out = np.empty((n, nd, 3), dtype=np.float32)
    cnt = np.zeros(n, dtype=np.int32)
    if az_jitter is None:
        az_jitter = 2 * np.pi / 625
    vp = lambda a: a.ctypes.data_as(C.c_void_p)
    lib().synthgen_scans(vp(rects), C.c_int(rects.shape[0]), vp(poses), C.c_int(n), vp(dirs), C.c_int(nd), C.c_double(az_jitter),
                         C.c_double(range_noise), C.c_double(blind), C.c_double(max_range), C.c_uint64(seed), vp(out), vp(cnt))
    off = np.zeros(n + 1, dtype=np.int64); off[1:] = np.cumsum(cnt)
    xyz = np.empty((int(off[-1]), 3), dtype=np.float32)
    for i in range(n):
        xyz[off[i]:off[i + 1]] = out[i, :cnt[i]]
    return xyz, off


def c4_session(n_kf=2000, n_az=625, seed=20241008 + 4, nx=10, ny=4, rows=(0, 1), phase=0.0, max_range=30.0):
    """One session of C4: n_kf keyframes of 32 * n_az rays on the loop through an nx x ny tile world.
    -> dict(xyz, off, truth=(R, p), start=(R, p))."""
    rects = tile_world(nx, ny)
    R, p = loop_path(n_kf, nx, seed, rows, phase)
    xyz, off = cast(rects, R, p, xt32_dirs(n_az), seed, max_range=max_range)
    Rs, ps = odometry_drift(R, p, seed)
    return dict(xyz=xyz, off=off, truth=(R, p), start=(Rs, ps), rects=rects)


def clouds_of(d):
    return [d["xyz"][d["off"][i]:d["off"][i + 1]] for i in range(d["off"].size - 1)]
