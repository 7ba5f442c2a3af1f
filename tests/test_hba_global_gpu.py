"""GPU parity for the two-level hierarchical global BA (BASELINE.json configs[3] reduced):
vx_hba_global (schedule -> batched bottom windows -> submaps -> top-level HBA_add_edge) against the oracle chain of
tests/oracle_global.py (thd_globalmapping, voxelslam.cpp:3018-3141, over orc_hba_add_edge).

Bars: same windows, same edge sets (ids exact), relative poses <= 1e-8, cov <= 1e-6 rel; top-level poses <= 1e-8 abs."""
import numpy as np
import pytest

import oracle_global as og
import synthgen as sg

pytestmark = pytest.mark.gpu

LOCAL = (0.5, 0.0025, [0.25] * 4)
GBA = (1.0, 0.01, [0.5] * 4)


def pose_err(a, b):
    return max(np.max(np.abs(a["R"] - b["R"])), np.max(np.abs(a["p"] - b["p"])))


def check_edges(eg, eo, cov_tol=1e-6):
    assert len(eg) == len(eo)
    for f in ("m1", "m2", "id1", "id2"):
        assert np.array_equal(eg[f], eo[f]), f
    if len(eg):
        assert np.max(np.abs(eg["rot"] - eo["rot"])) < 1e-8 and np.max(np.abs(eg["tra"] - eo["tra"])) < 1e-8
        assert np.max(np.abs(eg["cov"] - eo["cov"]) / np.abs(eo["cov"])) < cov_tol


def session(n_kf, n_az, seed, nx=2, phase=0.0):
    d = sg.c4_session(n_kf=n_kf, n_az=n_az, seed=seed, nx=nx, ny=2, phase=phase)
    return sg.clouds_of(d), d


@pytest.mark.parametrize("n_kf,n_az,tmi", [(35, 150, 2), (22, 100, 1), (120, 60, 3)])
def test_two_level_matches_oracle(pkg, ob, ctx, n_kf, n_az, tmi):
    # 35 keyframes -> 6 full windows + the 5-keyframe tail -> top level over 7 submaps (k_lm_solve path);
    # 120 keyframes -> 23 + 1 windows -> 144 unknowns at the top (host sweep between LM_MAXN and VX_LDLT_DEVICE_MIN)
    clouds, d = session(n_kf, n_az, 20241008 + 40 + n_kf)
    sess = np.zeros(n_kf, dtype=np.int32)
    ids = 7 + 10 * np.arange(n_kf)                      # Keyframe::id = index of the scan the keyframe was cut at
    ref = og.hba_global(clouds, sess, d["start"], LOCAL, GBA, 2, tmi, kf_id=ids)
    out = pkg.hba_global(ctx, clouds, sess, d["start"], LOCAL, GBA, 2, tmi, kf_id=ids)
    assert out["stats"]["windows"] == len(ref["windows"])
    assert np.array_equal(out["top_kf"], ref["top_kf"])
    check_edges(out["edges1"], ref["edges1"])
    check_edges(out["edges2"], ref["edges2"], cov_tol=1e-5)
    assert pose_err(out["top_poses"], ref["top_poses"]) < 1e-8
    assert len(out["edges1"]) > 20 * (len(ref["windows"]) - 1) and len(out["edges2"]) >= len(ref["windows"]) - 1
    assert out["stats"]["voxel_frame_evals"] > 0


def test_two_sessions_and_cnct_filter(pkg, ob, ctx):
    # two stored sessions (20 + 25 keyframes); the top level takes only the submaps of the connected sessions (voxelslam.cpp:2836-2857)
    ca, da = session(20, 100, 20241008 + 61)
    cb, db = session(25, 100, 20241008 + 62, phase=3.0)
    clouds = ca + cb
    sess = np.array([0] * 20 + [1] * 25, dtype=np.int32)
    x0 = (np.concatenate([da["start"][0], db["start"][0]]), np.concatenate([da["start"][1], db["start"][1]]))
    for cn in (None, [1], [0, 1]):
        ref = og.hba_global(clouds, sess, x0, LOCAL, GBA, 2, 2, cnct_map=cn)
        out = pkg.hba_global(ctx, clouds, sess, x0, LOCAL, GBA, 2, 2, cnct_map=cn)
        assert np.array_equal(out["top_kf"], ref["top_kf"])
        check_edges(out["edges1"], ref["edges1"])
        check_edges(out["edges2"], ref["edges2"], cov_tol=1e-5)
        assert pose_err(out["top_poses"], ref["top_poses"]) < 1e-8
    assert set(np.unique(out["edges2"]["m1"])) <= {0, 1}


def test_sharded_build_on_one_rank_is_the_plain_build(pkg, ctx):
    # the exchange path (owner keys, stable bucket partition, self-copy, explicit-frame build) with a single rank must reproduce
    # vx_gba_build bit for bit: same voxels in the same order, same moments
    clouds, d = session(12, 200, 20241008 + 70)
    W = 12
    f0 = pkg.Factor(ctx, W); f0.gba_build(clouds, d["start"], 1.0, 0.01, [0.5] * 4, 2)
    comm = pkg.Comm(ctx, None, 0, 1)
    f1 = pkg.Factor(ctx, W)
    # clouds handed over in a scrambled order with their frame ids
    perm = np.array([5, 0, 11, 3, 4, 1, 6, 7, 8, 9, 10, 2])
    n = f1.gba_build_sharded(comm, [clouds[i] for i in perm], perm, d["start"], 1.0, 0.01, [0.5] * 4, 2)
    assert n == sum(c.shape[0] for c in clouds)
    a, b = f0.download_csr(), f1.download_csr()
    assert f0.num_voxels == f1.num_voxels > 100
    for k in ("offsets", "frames"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(a["clusters"].view(np.uint8), b["clusters"].view(np.uint8))
    ps = pkg.as_poses(d["start"])
    H0, J0, c0 = f0.hessian(ps); H1, J1, c1 = f1.hessian(ps)
    assert np.array_equal(H0, H1) and np.array_equal(J0, J1)
    r = f1.ba_lidar_sharded(comm, ps, max_iter=2)
    r0 = f0.ba_lidar(ps, max_iter=2)
    assert pose_err(r["poses"], r0["poses"]) == 0.0
    comm.free(); f0.free(); f1.free()
