"""Seeded synthetic inputs for the BA factor tests (PCG64 via numpy).

random_factor builds a LidarFactor directly (planar patches seen from random subsets of frames);
tests/scene.py builds the ray-cast room scenes of the BASELINE.json configs.
"""
import numpy as np

cluster_dtype = np.dtype([("P", "<f8", (9,)), ("v", "<f8", (3,)), ("N", "<i4"), ("_pad", "<i4")], align=True)


def rodrigues(w):
    w = np.asarray(w, dtype=np.float64)
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.eye(3)
    k = w / th
    K = np.array([[0, -k[2], k[1]], [k[2], 0, -k[0]], [-k[1], k[0], 0]])
    return np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)


def random_poses(seed, W, rot_sigma=0.3, tra_sigma=2.0):
    rng = np.random.Generator(np.random.PCG64(seed))
    R = np.stack([rodrigues(rng.normal(0, rot_sigma, 3)) for _ in range(W)])
    p = rng.normal(0, tra_sigma, (W, 3))
    return R, p


def perturb_poses(poses, seed, rot_sigma, tra_sigma, fix_first=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    R, p = poses
    R2, p2 = R.copy(), p.copy()
    for i in range(R.shape[0]):
        if fix_first and i == 0:
            continue
        R2[i] = R[i] @ rodrigues(rng.normal(0, rot_sigma, 3))
        p2[i] = p[i] + rng.normal(0, tra_sigma, 3)
    return R2, p2


def clusters_from_points(pts):
    """(n,3) -> (P 3x3, v 3, N)"""
    return pts.T @ pts, pts.sum(axis=0), pts.shape[0]


def set_cluster(c, P, v, N):
    c["P"] = np.asarray(P).T.reshape(9)  # column-major (symmetric anyway)
    c["v"] = v
    c["N"] = N


def random_factor(seed, V, W, poses=None, occupancy=0.6, noise=0.01, with_fix=0.3, pts_per_frame=(6, 40),
                  extent=0.4, min_frames=1):
    """V planar patches; patch a is seen by a random subset of frames.  Returns the dense factor dict
    (plvec, sig, coe, eigval, eigvec, merged) with the cache consistent with `poses` (truth)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    if poses is None:
        poses = random_poses(seed + 1000003, W)
    R, p = poses
    plvec = np.zeros((V, W), dtype=cluster_dtype)
    sig = np.zeros(V, dtype=cluster_dtype)
    merged = np.zeros(V, dtype=cluster_dtype)
    coe = np.ones(V)
    eigval = np.zeros((V, 3)); eigvec = np.zeros((V, 9))
    for a in range(V):
        n = rng.normal(size=3); n /= np.linalg.norm(n)
        t1 = np.cross(n, [1.0, 0, 0]); t1 = t1 / np.linalg.norm(t1) if np.linalg.norm(t1) > 0.1 else np.cross(n, [0, 1.0, 0])
        t1 /= np.linalg.norm(t1)
        t2 = np.cross(n, t1)
        c = rng.uniform(-10, 10, 3)
        ext = extent * rng.uniform(0.5, 1.5, 2)
        occ = rng.random(W) < occupancy
        while occ.sum() < min_frames:
            occ[rng.integers(W)] = True
        Pm = np.zeros((3, 3)); vm = np.zeros(3); Nm = 0

        def sample(k):
            uv = rng.uniform(-1, 1, (k, 2)) * ext
            return c + uv[:, :1] * t1 + uv[:, 1:] * t2 + rng.normal(0, noise, (k, 1)) * n

        if rng.random() < with_fix:
            w = sample(int(rng.integers(5, 60)))
            P_, v_, N_ = clusters_from_points(w)
            set_cluster(sig[a], P_, v_, N_)
            Pm += P_; vm += v_; Nm += N_
        for i in range(W):
            if not occ[i]:
                continue
            w = sample(int(rng.integers(pts_per_frame[0], pts_per_frame[1])))
            loc = (w - p[i]) @ R[i]          # R^T (w - p)
            P_, v_, N_ = clusters_from_points(loc)
            set_cluster(plvec[a, i], P_, v_, N_)
            Pw, vw, _ = clusters_from_points(w)
            Pm += Pw; vm += vw; Nm += N_
        set_cluster(merged[a], Pm, vm, Nm)
        cov = Pm / Nm - np.outer(vm / Nm, vm / Nm)
        wv, U = np.linalg.eigh(cov)
        eigval[a] = wv
        eigvec[a] = U.T.reshape(9)   # column-major
    return dict(plvec=plvec, sig=sig, coe=coe, eigval=eigval, eigvec=eigvec, merged=merged, poses=poses)


def dense_to_csr(fac):
    V, W = fac["plvec"].shape
    occ = fac["plvec"]["N"] != 0
    offsets = np.zeros(V + 1, dtype=np.int64)
    offsets[1:] = np.cumsum(occ.sum(axis=1))
    vi, fi = np.nonzero(occ)
    out = dict(fac)
    out["offsets"] = offsets
    out["frames"] = fi.astype(np.int32)
    out["clusters"] = np.ascontiguousarray(fac["plvec"][vi, fi])
    return out


def _rodrigues_batch(w):
    """(n,3) rotation vectors -> (n,3,3)."""
    w = np.asarray(w, dtype=np.float64)
    th = np.linalg.norm(w, axis=1)
    k = w / np.maximum(th, 1e-300)[:, None]
    K = np.zeros((w.shape[0], 3, 3))
    K[:, 0, 1] = -k[:, 2]; K[:, 0, 2] = k[:, 1]
    K[:, 1, 0] = k[:, 2]; K[:, 1, 2] = -k[:, 0]
    K[:, 2, 0] = -k[:, 1]; K[:, 2, 1] = k[:, 0]
    s = np.sin(th)[:, None, None]; c = (1 - np.cos(th))[:, None, None]
    return np.eye(3)[None] + s * K + c * (K @ K)


def batched_factor_csr(seed, n_windows, voxels_per_window, W, occupancy=0.6, noise=0.01, pts_max=24,
                       min_frames=2, chunk_windows=16):
    """Vectorised generator of a BATCH of independent LidarFactors in the CSR upload form
    (include/voxelba.h: vx_factor_upload_csr): `n_windows` bottom-level GBA windows
    (voxelslam.cpp:3096) of `W` keyframes, `voxels_per_window` plane voxels each; voxel a of window b
    is seen by a random subset (>= min_frames, the loop_refine.hpp:372-376 gate) of the window's frames.
    Clusters are exact moments of sampled planar points; the cache (eigval/eigvec/merged) is
    consistent with the returned poses, as OctreeGBA::recut leaves it (loop_refine.hpp:363-387)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    F = n_windows * W
    # per-window trajectory: a gentle arc (C1-like: ~1 m / ~4 deg between poses) about a random start
    yaw = np.cumsum(rng.normal(0.07, 0.02, (n_windows, W)), axis=1)
    rv = np.zeros((F, 3)); rv[:, 2] = yaw.reshape(-1); rv[:, :2] = rng.normal(0, 0.02, (F, 2))
    R = _rodrigues_batch(rv)
    step = np.stack([np.cos(yaw), np.sin(yaw), np.zeros_like(yaw)], axis=2) * rng.uniform(0.8, 1.2, (n_windows, W, 1))
    p = (np.cumsum(step, axis=1) + rng.uniform(-100, 100, (n_windows, 1, 3)) * np.array([1, 1, 0.02])).reshape(F, 3)

    offs = [np.zeros(1, dtype=np.int64)]
    out = {k: [] for k in ("frames", "clusters", "sig", "coe", "eigval", "eigvec", "merged")}
    base = 0
    for w0 in range(0, n_windows, chunk_windows):
        nw = min(chunk_windows, n_windows - w0)
        V = nw * voxels_per_window
        win = np.repeat(np.arange(w0, w0 + nw), voxels_per_window)
        n = rng.normal(size=(V, 3)); n /= np.linalg.norm(n, axis=1, keepdims=True)
        ref = np.where(np.abs(n[:, :1]) < 0.9, np.array([[1.0, 0, 0]]), np.array([[0, 1.0, 0]]))
        t1 = np.cross(n, ref); t1 /= np.linalg.norm(t1, axis=1, keepdims=True)
        t2 = np.cross(n, t1)
        ctr = p[win * W + W // 2] + rng.uniform(-12, 12, (V, 3)) * np.array([1, 1, 0.2])
        ext = 0.4 * rng.uniform(0.5, 1.5, (V, 2))
        occ = rng.random((V, W)) < occupancy
        short = occ.sum(axis=1) < min_frames
        while short.any():
            idx = np.nonzero(short)[0]
            occ[idx, rng.integers(0, W, idx.size)] = True
            short = occ.sum(axis=1) < min_frames
        vi, fi = np.nonzero(occ)
        nnz = vi.size
        g = win[vi] * W + fi
        cnt = rng.integers(6, pts_max + 1, nnz)
        mask = (np.arange(pts_max)[None, :] < cnt[:, None]).astype(np.float64)
        uv = rng.uniform(-1, 1, (nnz, pts_max, 2)) * ext[vi][:, None, :]
        wpts = ctr[vi][:, None, :] + uv[:, :, :1] * t1[vi][:, None, :] + uv[:, :, 1:] * t2[vi][:, None, :] \
            + rng.normal(0, noise, (nnz, pts_max, 1)) * n[vi][:, None, :]
        loc = np.einsum("npi,nij->npj", wpts - p[g][:, None, :], R[g])      # R^T (w - p)
        loc_m = loc * mask[:, :, None]
        w_m = wpts * mask[:, :, None]
        cl = np.zeros(nnz, dtype=cluster_dtype)
        cl["P"] = np.einsum("npi,npj->nij", loc_m, loc).reshape(nnz, 9)
        cl["v"] = loc_m.sum(axis=1)
        cl["N"] = cnt
        off = np.zeros(V + 1, dtype=np.int64); off[1:] = np.cumsum(occ.sum(axis=1))
        Pw = np.add.reduceat(np.einsum("npi,npj->nij", w_m, wpts).reshape(nnz, 9), off[:-1], axis=0)
        vw = np.add.reduceat(w_m.sum(axis=1), off[:-1], axis=0)
        Nw = np.add.reduceat(cnt, off[:-1])
        mg = np.zeros(V, dtype=cluster_dtype)
        mg["P"] = Pw; mg["v"] = vw; mg["N"] = Nw
        cen = vw / Nw[:, None]
        cov = Pw.reshape(V, 3, 3) / Nw[:, None, None] - cen[:, :, None] * cen[:, None, :]
        lam, U = np.linalg.eigh(cov)
        out["frames"].append(g.astype(np.int32)); out["clusters"].append(cl)
        out["sig"].append(np.zeros(V, dtype=cluster_dtype)); out["coe"].append(np.ones(V))
        out["eigval"].append(lam); out["eigvec"].append(np.transpose(U, (0, 2, 1)).reshape(V, 9))
        out["merged"].append(mg)
        offs.append(off[1:] + base); base += nnz
    fac = {k: np.concatenate(v) for k, v in out.items()}
    fac["offsets"] = np.concatenate(offs)
    fac["poses"] = (R, p)
    fac["W"] = W; fac["n_windows"] = n_windows
    return fac


def csr_window(fac, b):
    """Extract window b of a batched CSR factor as a dense single-window factor dict (for the oracle)."""
    W = fac["W"]
    off = fac["offsets"]
    fr = fac["frames"]
    first = fr[off[:-1]] // W
    vs = np.nonzero(first == b)[0]
    V = vs.size
    plvec = np.zeros((V, W), dtype=cluster_dtype)
    for k, v in enumerate(vs):
        e0, e1 = off[v], off[v + 1]
        plvec[k, fr[e0:e1] - b * W] = fac["clusters"][e0:e1]
    R, p = fac["poses"]
    return dict(plvec=plvec, sig=fac["sig"][vs], coe=fac["coe"][vs], eigval=fac["eigval"][vs], eigvec=fac["eigvec"][vs],
                merged=fac["merged"][vs], poses=(R[b * W:(b + 1) * W], p[b * W:(b + 1) * W]))


# ---- synthetic IMU for the LiDAR-inertial optimizers (200 Hz IMU from an analytic trajectory) ----
imu_sample_dtype = np.dtype([("t", "<f8"), ("gyr", "<f8", (3,)), ("acc", "<f8", (3,))])
imu_state_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,)), ("v", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)), ("g", "<f8", (3,))])


def _euler_R(psi, th, phi):
    cz, sz, cy, sy, cx, sx = np.cos(psi), np.sin(psi), np.cos(th), np.sin(th), np.cos(phi), np.sin(phi)
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1.0]]); Ry = np.array([[cy, 0, sy], [0, 1.0, 0], [-sy, 0, cy]])
    Rx = np.array([[1.0, 0, 0], [0, cx, -sx], [0, sx, cx]])
    return Rz @ Ry @ Rx


def imu_trajectory(seed, W, frame_dt=0.1, rate=200, gyr_sigma=0.0, acc_sigma=0.0, bg=(0.002, -0.001, 0.0015), ba=(0.02, -0.01, 0.015),
                   g=(0, 0, -9.8)):
    """Analytic trajectory sampled at W frame times + IMU samples between consecutive frames (each segment includes both
    end points, as push_imu expects).  Measurement model of the reference: gyr = w_body + bg, acc = R^T (a_world - g) + ba."""
    rng = np.random.Generator(np.random.PCG64(seed))
    A = rng.uniform(0.5, 1.5, 3) * np.array([2.0, 1.0, 0.2]); wv = rng.uniform(0.5, 1.2, 3); v0 = np.array([1.0, 0.2, 0.0])
    a_psi, a_th, a_phi = 0.4, 0.05 * rng.uniform(0.5, 1.5), 0.04 * rng.uniform(0.5, 1.5)
    g = np.asarray(g, dtype=np.float64); bg = np.asarray(bg, dtype=np.float64); ba = np.asarray(ba, dtype=np.float64)

    def pos(t): return v0 * t + A * np.sin(wv * t)
    def vel(t): return v0 + A * wv * np.cos(wv * t)
    def acc(t): return -A * wv * wv * np.sin(wv * t)
    def eul(t): return a_psi * t + 0.1 * np.sin(t), a_th * np.sin(0.9 * t), a_phi * np.sin(1.3 * t)
    def eul_d(t): return a_psi + 0.1 * np.cos(t), 0.9 * a_th * np.cos(0.9 * t), 1.3 * a_phi * np.cos(1.3 * t)

    def body_rate(t):
        psi, th, phi = eul(t); dpsi, dth, dphi = eul_d(t)
        return np.array([dphi - dpsi * np.sin(th), dth * np.cos(phi) + dpsi * np.sin(phi) * np.cos(th),
                         -dth * np.sin(phi) + dpsi * np.cos(phi) * np.cos(th)])

    states = np.zeros(W, dtype=imu_state_dtype)
    segs = []
    n_sub = int(round(frame_dt * rate))
    for i in range(W):
        t = i * frame_dt
        R = _euler_R(*eul(t))
        states[i]["R"] = R.T.reshape(9); states[i]["p"] = pos(t); states[i]["v"] = vel(t)
        states[i]["bg"] = bg; states[i]["ba"] = ba; states[i]["g"] = g
        if i + 1 < W:
            s = np.zeros(n_sub + 1, dtype=imu_sample_dtype)
            for k in range(n_sub + 1):
                tk = t + k * frame_dt / n_sub
                Rk = _euler_R(*eul(tk))
                s[k]["t"] = tk
                s[k]["gyr"] = body_rate(tk) + bg + rng.normal(0, gyr_sigma, 3)
                s[k]["acc"] = Rk.T @ (acc(tk) - g) + ba + rng.normal(0, acc_sigma, 3)
            segs.append(s)
    return states, segs


def states_poses(states):
    R = np.transpose(states["R"].reshape(-1, 3, 3), (0, 2, 1)).copy()
    return R, states["p"].copy()
