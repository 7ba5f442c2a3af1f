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
