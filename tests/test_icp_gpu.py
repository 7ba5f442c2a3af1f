"""GPU parity for icp_normal (loop_refine.hpp:47-145) through vx_icp_normal.

Bars: the per-iteration match counts, the iteration count, the converged flag and the accept decision are exact (the nearest
neighbour and every gate are evaluated in the oracle's arithmetic; only the 35 sums are re-ordered); residuals <= 1e-10 rel;
final pose <= 1e-9 abs; eigenvalues of sum n n^T <= 1e-9 rel."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def plane_clouds(seed, n_tar, n_src, noise=0.01, planes=5):
    """Points with unit normals on a few finite planes of a room corner; the source is a different sample of the same surfaces,
    expressed in a frame displaced by a small rigid motion."""
    rng = np.random.Generator(np.random.PCG64(seed))
    defs = [(np.array([0, 0, 1.0]), 0.0), (np.array([1.0, 0, 0]), -4.0), (np.array([0, 1.0, 0]), -3.0),
            (np.array([0.6, 0.8, 0]), -2.0), (np.array([0, 0.6, 0.8]), -2.5)][:planes]

    def sample(n):
        pts = []; nrm = []
        for k, (nv, d) in enumerate(defs):
            m = n // len(defs)
            a = np.cross(nv, [0.3, -0.5, 0.8]); a /= np.linalg.norm(a); b = np.cross(nv, a)
            uv = rng.uniform(-3, 3, size=(m, 2))
            p = -d * nv + uv[:, :1] * a + uv[:, 1:] * b + noise * rng.normal(size=(m, 1)) * nv
            pts.append(p); nrm.append(np.tile(nv, (m, 1)) + 0.01 * rng.normal(size=(m, 3)))
        nrm = np.concatenate(nrm); nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
        return np.concatenate(pts), nrm

    tar, tn = sample(n_tar)
    src_w, sn_w = sample(n_src)
    return tar.astype(np.float32), tn.astype(np.float32), src_w, sn_w


@pytest.mark.parametrize("seed,n_tar,n_src,rot,tra,noise", [(1, 4000, 3000, 0.02, 0.15, 0.01), (2, 6000, 2500, 0.05, 0.3, 0.01),
                                                                (3, 3000, 3000, 0.0, 0.0, 0.01), (4, 5000, 4000, 0.08, 0.5, 0.03),
                                                                (5, 5000, 4000, 0.12, 0.8, 0.03)])
def test_icp_normal_matches_oracle(pkg, ob, ctx, seed, n_tar, n_src, rot, tra, noise):
    tar, tn, src_w, sn_w = plane_clouds(seed, n_tar, n_src, noise=noise)
    Rt, pt = synth.random_poses(seed + 50, 1, rot_sigma=0.3, tra_sigma=2.0)       # true source pose
    Rt, pt = Rt[0], pt[0]
    src = ((src_w - pt) @ Rt).astype(np.float32)                                   # R^T (x - p)
    sn = (sn_w @ Rt).astype(np.float32)
    R0, p0 = synth.perturb_poses((Rt[None], pt[None]), seed + 7, rot, tra, fix_first=False)
    ref = ob.icp_normal(src, sn, tar, tn, (R0[0], p0[0]), 0.1)
    out = pkg.icp_normal(ctx, src, sn, tar, tn, (R0[0], p0[0]), 0.1)
    assert ref["iters"] >= 2 and ref["match_nums"][0] > 500
    assert out["iters"] == ref["iters"] and out["match_nums"] == ref["match_nums"]
    assert out["is_converge"] == ref["is_converge"] and out["ok"] == ref["ok"]
    assert np.max(np.abs(out["resis"] - ref["resis"]) / np.maximum(np.abs(ref["resis"]), 1e-300)) < 1e-10
    assert np.abs(out["R"] - ref["R"]).max() < 1e-9 and np.abs(out["p"] - ref["p"]).max() < 1e-9
    assert np.abs(out["eig"] - ref["eig"]).max() <= 1e-9 * np.abs(ref["eig"]).max()
    if rot > 0:     # the ICP pulled the perturbed pose back onto the planted one (noise-limited)
        assert np.abs(out["p"] - pt).max() < 0.3 * max(np.abs(p0[0] - pt).max(), 1e-9) + 0.01
        assert ref["ok"]


def test_icp_normal_edge_cases(pkg, ob, ctx):
    tar, tn, src_w, sn_w = plane_clouds(9, 2000, 1000)
    src = src_w.astype(np.float32); sn = sn_w.astype(np.float32)
    I = (np.eye(3), np.zeros(3))
    # empty source / empty target: no match, singular system -> dx = 0 twice -> "converged" with a zero normal matrix
    for a, b, c, d in ((src[:0], sn[:0], tar, tn), (src, sn, tar[:0], tn[:0])):
        ref = ob.icp_normal(a, b, c, d, I, 0.1)
        out = pkg.icp_normal(ctx, a, b, c, d, I, 0.1)
        assert out["match_nums"] == ref["match_nums"] == [0, 0] and out["ok"] == ref["ok"] is False
        assert np.array_equal(out["R"], ref["R"]) and np.array_equal(out["p"], ref["p"])
    # a source far away from the target (beyond the 3 m gate): nothing is associated
    far = (src + np.float32(50.0)).astype(np.float32)
    ref = ob.icp_normal(far, sn, tar, tn, I, 0.1)
    out = pkg.icp_normal(ctx, far, sn, tar, tn, I, 0.1)
    assert out["match_nums"] == ref["match_nums"] and sum(ref["match_nums"]) == 0
    # exact duplicates in the target: the lowest index wins on both sides (normals differ between the copies)
    tar2 = np.concatenate([tar, tar]); tn2 = np.concatenate([tn, -tn])
    ref = ob.icp_normal(src, sn, tar2, tn2, I, 0.1)
    out = pkg.icp_normal(ctx, src, sn, tar2, tn2, I, 0.1)
    assert out["match_nums"] == ref["match_nums"] and out["iters"] == ref["iters"]
    assert np.abs(out["p"] - ref["p"]).max() < 1e-9


# ---- BTC plane voxels (BTC.cpp:279-334): the plane cloud icp_normal is fed with ------------------------------------
def _world_cloud(seed, frames, pts):
    import scene
    clouds, (R, p) = scene.config1(seed=seed, W=frames, pts_per_scan=pts)
    return np.concatenate([c @ R[i].T + p[i] for i, c in enumerate(clouds)]).astype(np.float32), clouds, (R, p)


@pytest.mark.parametrize("seed,voxel", [(31, 1.0), (32, 0.5), (33, 2.0)])
def test_btc_plane_cloud_matches_oracle(pkg, ob, ctx, seed, voxel):
    xyz, _, _ = _world_cloud(seed, 6, 8000)
    xyz[:50] *= -1.0                                     # negative coordinates: the "-1.0 on the double quotient" branch
    thre = float(np.float32(0.01))                       # config_setting_.plane_detection_thre_ is a float (BTC.h:29)
    ref = ob.btc_plane_cloud(xyz, voxel, thre, 10)
    pts, aux = pkg.btc_plane_cloud(ctx, xyz, voxel, thre, 10)
    assert ref.shape[0] > 30 and aux.shape == ref.shape  # same voxels pass N > 10 and lambda_min < thre, in the same (key) order
    assert np.array_equal(aux[:, :3], ref[:, :3]) and np.array_equal(aux[:, 9], ref[:, 9])     # centers: same sums, bit for bit
    sgn = np.sign(np.sum(aux[:, 3:6] * ref[:, 3:6], axis=1, keepdims=True))                   # eigenvector sign is unspecified
    assert np.abs(aux[:, 3:6] * sgn - ref[:, 3:6]).max() < 1e-7      # lambda_0 vs lambda_1 gaps of ~1e-3 amplify eps
    assert np.abs(aux[:, 6] - ref[:, 6]).max() <= 1e-12 * np.abs(ref[:, 7] ** 2).max()
    assert np.abs(aux[:, 7] - ref[:, 7]).max() < 1e-12 and np.abs(aux[:, 8] * sgn[:, 0] - ref[:, 8]).max() < 1e-6
    assert np.array_equal(pts["x"], aux[:, 0].astype(np.float32)) and np.array_equal(pts["normal_z"], aux[:, 5].astype(np.float32))
    assert pkg.btc_plane_cloud(ctx, xyz[:0], voxel, thre, 10)[0].shape[0] == 0
    assert pkg.btc_plane_cloud(ctx, xyz[:10], voxel, thre, 10)[0].shape[0] == 0        # no voxel with more than 10 points


def test_loop_closure_geometry_chain(pkg, ob, ctx):
    # keyframe clouds -> BTC plane clouds -> icp_normal, device against oracle end to end (voxelslam.cpp:2434)
    xyz_a, _, _ = _world_cloud(41, 6, 8000)
    xyz_b, _, _ = _world_cloud(42, 6, 8000)             # another sample of the same room
    Rt, pt = synth.random_poses(9, 1, rot_sigma=0.2, tra_sigma=1.0); Rt, pt = Rt[0], pt[0]
    src_local = ((xyz_b.astype(np.float64) - pt) @ Rt).astype(np.float32)
    thre = float(np.float32(0.01))
    pa, _ = pkg.btc_plane_cloud(ctx, xyz_a, 1.0, thre, 10)
    pb, _ = pkg.btc_plane_cloud(ctx, src_local, 1.0, thre, 10)
    assert pa.shape[0] > 50 and pb.shape[0] > 50
    f3 = lambda a, names: np.stack([a[n] for n in names], axis=1)
    tar, tn = f3(pa, "xyz"), f3(pa, ("normal_x", "normal_y", "normal_z"))
    src, sn = f3(pb, "xyz"), f3(pb, ("normal_x", "normal_y", "normal_z"))
    R0, p0 = synth.perturb_poses((Rt[None], pt[None]), 3, 0.02, 0.1, fix_first=False)
    ref = ob.icp_normal(src, sn, tar, tn, (R0[0], p0[0]), 0.1)
    out = pkg.icp_normal(ctx, src, sn, tar, tn, (R0[0], p0[0]), 0.1)
    assert ref["match_nums"][0] > 20
    assert out["match_nums"] == ref["match_nums"] and out["iters"] == ref["iters"] and out["ok"] == ref["ok"]
    assert np.abs(out["R"] - ref["R"]).max() < 1e-9 and np.abs(out["p"] - ref["p"]).max() < 1e-9
