"""Self-consistency known-answer tests that pin the oracle: the reference ships
no tests or golden vectors, so the restatement is validated against LAPACK and finite differences."""
import numpy as np
import pytest

import synth


def test_eig3_vs_lapack(ob):
    rng = np.random.Generator(np.random.PCG64(20241008))
    n = 20000
    B = rng.normal(size=(n, 3, 3))
    A = B @ np.transpose(B, (0, 2, 1))
    # near-degenerate and plane-like cases
    A[: n // 4] *= np.array([1.0, 1.0, 1e-4])[None, :, None] * np.array([1.0, 1.0, 1e-4])[None, None, :]
    A[n // 4: n // 2] += np.eye(3)[None] * 1e3
    w, U = ob.eig3(A)
    wl, _ = np.linalg.eigh(A)
    scale = np.abs(wl).max(axis=1, keepdims=True)
    assert np.all(np.diff(w, axis=1) >= 0)
    assert np.max(np.abs(w - wl) / scale) < 5e-15
    # A U = U diag(w), U orthonormal
    res = A @ U - U * w[:, None, :]
    assert np.max(np.abs(res) / scale[:, :, None]) < 1e-14
    assert np.max(np.abs(np.transpose(U, (0, 2, 1)) @ U - np.eye(3))) < 1e-14


def test_ldlt_vs_numpy(ob):
    rng = np.random.Generator(np.random.PCG64(1))
    for n in (6, 60, 153):
        B = rng.normal(size=(n, n + 5))
        A = B @ B.T + 0.1 * np.eye(n)
        b = rng.normal(size=n)
        x = ob.ldlt_solve(A, b)
        assert np.max(np.abs(x - np.linalg.solve(A, b))) < 1e-9 * np.max(np.abs(x))
        # only the LOWER triangle is read (Eigen LDLT default): garbage above the diagonal is ignored
        A2 = A.copy(); A2[np.triu_indices(n, 1)] = 123.0
        assert np.array_equal(ob.ldlt_solve(A2, b), x)


def test_inverse15_and_so3(ob):
    rng = np.random.Generator(np.random.PCG64(2))
    B = rng.normal(size=(15, 15)); A = B @ B.T + np.eye(15)
    assert np.max(np.abs(ob.inverse15(A) @ A - np.eye(15))) < 1e-11
    for _ in range(20):
        w = rng.normal(0, 0.7, 3)
        R = ob.so3_exp(w)
        assert np.max(np.abs(R - synth.rodrigues(w))) < 1e-14
        assert np.max(np.abs(ob.so3_log(R) - w)) < 1e-9
    assert np.array_equal(ob.so3_exp(np.zeros(3)), np.eye(3))


def test_key_rule_quirks(ob):
    # not floor().  (-1,0) -> -1, exactly -1.0 -> -2, -0.0 -> 0
    xyz = np.array([[-0.5, -1.0, -0.0], [0.0, 0.999, 1.0], [-2.0, -2.5, 2.5]])
    k = ob.keys_f64(xyz, 1.0)
    assert k.tolist() == [[-1, -2, 0], [0, 0, 1], [-3, -3, 2]]
    # independent numpy float32 coding of the same rule on adversarial coordinates
    rng = np.random.Generator(np.random.PCG64(3))
    n = 200000
    vs = 0.5
    cells = rng.integers(-3000, 3000, (n, 3)).astype(np.float64)
    eps = rng.choice([0.0, 1e-9, -1e-9, 1e-7, -1e-7, 0.25, -0.25], (n, 3))
    pts = (cells + eps) * vs
    loc = (pts / vs).astype(np.float32)
    loc = np.where(loc < 0, loc - np.float32(1.0), loc).astype(np.float32)
    ref = np.trunc(loc).astype(np.int64)
    assert np.array_equal(ob.keys_f64(pts, vs), ref)
    p32 = pts.astype(np.float32)
    loc = (p32.astype(np.float64) / vs).astype(np.float32)
    loc = np.where(loc < 0, loc - np.float32(1.0), loc).astype(np.float32)
    assert np.array_equal(ob.keys_f32(p32, vs), np.trunc(loc).astype(np.int64))


def test_transform_equals_reaccumulate(ob):
    # KAT 3: PointCluster::transform == re-accumulating the transformed raw points
    rng = np.random.Generator(np.random.PCG64(4))
    pts = rng.normal(size=(50, 3))
    R, p = synth.random_poses(5, 1)
    c = np.zeros(1, dtype=ob.cluster_dtype)[0]
    P, v, N = synth.clusters_from_points(pts)
    synth.set_cluster(c, P, v, N)
    out = ob.cluster_transform(c, (R, p))
    w = pts @ R[0].T + p[0]
    Pw, vw, _ = synth.clusters_from_points(w)
    assert np.max(np.abs(out["P"].reshape(3, 3).T - Pw)) < 1e-12 * np.max(np.abs(Pw))
    assert np.max(np.abs(out["v"] - vw)) < 1e-12 * np.max(np.abs(vw))
    assert out["N"] == 50


def _fd_setup(ob, seed=11, V=40, W=5):
    fac = synth.random_factor(seed, V, W, noise=0.02)
    poses = synth.perturb_poses(fac["poses"], seed + 1, 0.01, 0.02, fix_first=False)
    f = ob.Factor(W)
    f.set_dense(fac)
    return f, poses, W


def _retract(poses, dx):
    R, p = poses
    R2 = np.stack([R[i] @ synth.rodrigues(dx[6 * i: 6 * i + 3]) for i in range(R.shape[0])])
    p2 = p + dx.reshape(-1, 6)[:, 3:]
    return R2, p2


def test_gradient_matches_finite_differences(ob):
    # KAT 1: JacT == d/dx sum coe*lambda_min under R*Exp(dtheta), p+dp (voxel_map.hpp:462-463)
    f, poses, W = _fd_setup(ob)
    f.residual(poses)            # refresh cache at the linearisation point
    H, J, c = f.hessian(poses)
    h = 1e-6
    g = np.zeros(6 * W)
    for k in range(6 * W):
        dx = np.zeros(6 * W); dx[k] = h
        g[k] = (f.residual(_retract(poses, dx)) - f.residual(_retract(poses, -dx))) / (2 * h)
    assert np.max(np.abs(g - J)) < 1e-6 * np.max(np.abs(J))


def test_hessian_matches_fd_of_gradient(ob):
    # KAT 2: Hess == d JacT / dx up to the documented antisymmetric -0.5*hat(jjt_rot) on rot-rot diagonals
    f, poses, W = _fd_setup(ob, seed=21, V=30, W=4)
    f.residual(poses)
    H, J, _ = f.hessian(poses)
    h = 1e-6
    Hfd = np.zeros((6 * W, 6 * W))
    for k in range(6 * W):
        dx = np.zeros(6 * W); dx[k] = h
        pp = _retract(poses, dx); f.residual(pp); _, Jp, _ = f.hessian(pp)
        pm = _retract(poses, -dx); f.residual(pm); _, Jm, _ = f.hessian(pm)
        Hfd[:, k] = (Jp - Jm) / (2 * h)
    f.residual(poses)
    Hs = 0.5 * (H + H.T)
    Hfds = 0.5 * (Hfd + Hfd.T)
    assert np.max(np.abs(Hs - Hfds)) < 2e-5 * np.max(np.abs(Hs))
    # lower blocks mirror upper blocks (voxel_map.hpp:279-281)
    for i in range(W):
        for j in range(i):
            assert np.array_equal(H[6 * i:6 * i + 6, 6 * j:6 * j + 6], H[6 * j:6 * j + 6, 6 * i:6 * i + 6].T)


def test_ba_recovers_planted_poses(ob):
    # KAT (v): noiseless planes, perturbed start -> LM returns to the truth
    W = 5
    fac = synth.random_factor(31, 200, W, noise=0.0, occupancy=0.8, with_fix=0.0, min_frames=2)
    truth = fac["poses"]
    start = synth.perturb_poses(truth, 32, 0.004, 0.01)
    f = ob.Factor(W)
    f.set_dense(fac)
    f.residual(start)
    out = f.ba_lidar(start, max_iter=12, thd_num=2)
    assert out["rc"] in (0, 1)
    R, p = ob.poses_to_Rp(out["poses"])
    assert out["resis"][1] < 1e-6 * out["resis"][0]
    # gauge: pose 0 is fixed and was exact
    assert np.max(np.abs(p - truth[1])) < 1e-5
    assert np.max(np.abs(R - truth[0])) < 1e-5


def test_thread_split_matches_single_range(ob):
    f, poses, W = _fd_setup(ob, seed=41, V=101, W=6)
    f.residual(poses)
    H1, J1, r1 = f.hessian(poses)
    H5, J5, r5 = f.divide_thread(poses, thd_num=5, use_threads=True)
    assert np.max(np.abs(H1 - H5)) < 1e-12 * np.max(np.abs(H1))
    assert abs(r1 - r5) < 1e-12 * abs(r1)


def test_btc_plane_cloud_vs_numpy(ob):
    # independent numpy restatement of BTC.cpp:279-334 / :90-143 on a small cloud: key rule (double quotient, -1 when negative,
    # truncation), per-voxel mean / covariance / eigh, plane when N > init_num and lambda_min < thre, std::map key order
    rng = np.random.Generator(np.random.PCG64(77))
    n0 = np.array([0.3, -0.5, 0.81]); n0 /= np.linalg.norm(n0)
    a = rng.uniform(-3.0, 3.0, size=(6000, 2))
    e1 = np.cross(n0, [1.0, 0, 0]); e1 /= np.linalg.norm(e1); e2 = np.cross(n0, e1)
    plane = a[:, :1] * e1 + a[:, 1:] * e2 + rng.normal(size=(6000, 1)) * 0.01 * n0
    clutter = rng.uniform(4.0, 7.0, size=(3000, 3))            # dense non-planar voxels: must be rejected by the eigenvalue test
    xyz = np.concatenate([plane, clutter]).astype(np.float32)
    thre = float(np.float32(0.01))
    out = ob.btc_plane_cloud(xyz, 1.0, thre, 10)
    q = xyz.astype(np.float64) / 1.0
    q[q < 0] -= 1.0
    keys = np.trunc(q).astype(np.int64)
    want = []
    for key in sorted({tuple(k) for k in keys}):
        m = np.all(keys == np.array(key), axis=1)
        pts = xyz[m].astype(np.float64)
        if pts.shape[0] <= 10:
            continue
        c = pts.mean(axis=0)
        cov = pts.T @ pts / pts.shape[0] - np.outer(c, c)
        w, U = np.linalg.eigh(cov)
        if w[0] < thre:
            want.append((c, U[:, 0], w[0], pts.shape[0]))
    assert len(want) > 20 and out.shape[0] == len(want)
    for row, (c, nrm, w0, cnt) in zip(out, want):
        assert np.abs(row[:3] - c).max() < 1e-12 and int(row[9]) == cnt
        assert abs(row[6] - w0) < 1e-12
        if w0 < 1e-3:        # a clean plane voxel: the normal is well conditioned
            assert min(np.abs(row[3:6] - nrm).max(), np.abs(row[3:6] + nrm).max()) < 1e-8
    clean = out[out[:, 6] < 3e-4]
    assert clean.shape[0] > 10 and np.all(np.abs(np.abs(clean[:, 3:6] @ n0) - 1.0) < 2e-2)   # planted normal recovered (border voxels hold few points)


def test_icp_normal_recovers_planted_pose(ob):
    # three orthogonal walls sampled on a jittered grid with exact normals; the source is the target moved by a known small
    # transform: icp_normal must bring it back, converge within its 20 iterations and accept (eigenvalue gate on sum n n^T)
    rng = np.random.Generator(np.random.PCG64(5))
    g = np.stack(np.meshgrid(np.arange(0.0, 10.0, 0.5), np.arange(0.0, 10.0, 0.5), indexing="ij"), -1).reshape(-1, 2)
    g = g + rng.uniform(-0.05, 0.05, size=g.shape)
    z = np.zeros((g.shape[0], 1))
    walls = [np.hstack([g, z]), np.hstack([g[:, :1], z, g[:, 1:]]), np.hstack([z, g])]
    nrms = [np.tile([0, 0, 1.0], (g.shape[0], 1)), np.tile([0, 1.0, 0], (g.shape[0], 1)), np.tile([1.0, 0, 0], (g.shape[0], 1))]
    tar = np.concatenate(walls).astype(np.float32); tn = np.concatenate(nrms).astype(np.float32)
    Rt, pt = synth.random_poses(3, 1, rot_sigma=0.03, tra_sigma=0.15)
    Rt, pt = Rt[0], pt[0]
    src = ((tar.astype(np.float64) - pt) @ Rt).astype(np.float32)       # R^T (x - p): local frame of the source
    sn = (tn.astype(np.float64) @ Rt).astype(np.float32)
    out = ob.icp_normal(src, sn, tar, tn, (np.eye(3), np.zeros(3)), 0.1)
    assert out["ok"] and out["is_converge"] == 1 and out["iters"] <= 20
    assert out["match_nums"][-1] > 0.8 * src.shape[0]
    assert np.abs(out["R"] - Rt).max() < 2e-3 and np.abs(out["p"] - pt).max() < 2e-2
    # an empty target can never be accepted
    bad = ob.icp_normal(src, sn, tar[:0], tn[:0], (np.eye(3), np.zeros(3)), 0.1)
    assert not bad["ok"]
