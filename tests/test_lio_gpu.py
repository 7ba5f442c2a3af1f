"""GPU parity for the LIO ESIKF point-to-plane update on a BASELINE.json config-2
shaped case: Hesai-XT32-like sweeps against a voxel map prebuilt from earlier sweeps (A5 -> A6 -> A9 -> A12
run by the oracle's map session; the flattened map is what the drop-in boundary receives).

Bars: per-point association flags, cached leaves, match_num: exact.  HTH / HTz / sum n n^T:
<= 1e-12 relative (re-ordered FP64 sums).  State after lio_state_estimation: <= 1e-10 abs, same iteration count."""
import numpy as np
import pytest

import scene
import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world(ob):
    rng = np.random.Generator(np.random.PCG64(20241008 + 2))
    rects = scene.room()
    R, p = scene.arc_trajectory(16, 5, step=0.5, yaw_step_deg=3.0)
    m = ob.LocalMap(win_size=10, voxel_size=1.0, thre=(1, 1, 1, 1), dept_err=0.01, beam_err=0.01)     # hesai.yaml
    for i in range(14):
        m.push(scene.scan(rects, R[i], p[i], scene.xt32_dirs(700), rng, range_noise=0.01, blind=0.7), (R[i:i + 1], p[i:i + 1]), ob.default_cov())
    q = scene.scan(rects, R[14], p[14], scene.xt32_dirs(2000), rng, range_noise=0.01, blind=0.7)        # ~60k points
    return dict(map=m, export=m.export(), pv=ob.var_init(q, 0.01, 0.01), truth=(R[14], p[14]))


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def test_lio_pass_parity(pkg, ob, ctx, world):
    keys, rn, nodes = world["export"]
    gm = pkg.PlaneMap(ctx, keys, rn, nodes, 1.0)
    pv = world["pv"]
    assert pv.shape[0] > 50000
    R0, p0 = synth.perturb_poses((world["truth"][0][None], world["truth"][1][None]), 3, np.deg2rad(0.3), 0.03, fix_first=False)
    cov = ob.default_cov()
    ref = world["map"].lio_pass(pv, (R0, p0), cov)
    out = gm.lio_pass(pv, (R0, p0), cov)
    assert ref["match_num"] > 0.8 * pv.shape[0]
    assert out["match_num"] == ref["match_num"]
    assert np.array_equal(out["flags"], ref["flags"])                 # point -> plane association, bit-exact gates
    assert np.array_equal(out["leaf_cache"], ref["leaf_cache"])
    assert rel(out["HTH"], ref["HTH"]) < 1e-12 and rel(out["HTz"], ref["HTz"]) < 1e-11 and rel(out["nnt"], ref["nnt"]) < 1e-12
    # second pass from another pose re-uses the leaf cache (inside() fast path, voxelslam.cpp:1020)
    R1, p1 = synth.perturb_poses((world["truth"][0][None], world["truth"][1][None]), 4, np.deg2rad(0.1), 0.01, fix_first=False)
    ref2 = world["map"].lio_pass(pv, (R1, p1), cov, leaf_cache=ref["leaf_cache"])
    out2 = gm.lio_pass(pv, (R1, p1), cov, leaf_cache=out["leaf_cache"])
    assert out2["match_num"] == ref2["match_num"] and np.array_equal(out2["flags"], ref2["flags"])
    assert np.array_equal(out2["leaf_cache"], ref2["leaf_cache"])
    assert rel(out2["HTH"], ref2["HTH"]) < 1e-12
    # run-to-run bitwise reproducible (no atomics)
    out3 = gm.lio_pass(pv, (R1, p1), cov, leaf_cache=out["leaf_cache"])
    assert np.array_equal(out3["HTH"], out2["HTH"]) and np.array_equal(out3["HTz"], out2["HTz"])
    gm.free()


def test_lio_empty_and_unmapped(pkg, ob, ctx, world):
    keys, rn, nodes = world["export"]
    gm = pkg.PlaneMap(ctx, keys, rn, nodes, 1.0)
    cov = ob.default_cov()
    o = gm.lio_pass(np.zeros((0, 12)), (np.eye(3)[None], np.zeros((1, 3))), cov)
    assert o["match_num"] == 0 and not o["HTH"].any()
    far = world["pv"][:100].copy()
    o = gm.lio_pass(far, (np.eye(3)[None], np.array([[500.0, 500.0, 50.0]])), cov)     # no root voxel there
    assert o["match_num"] == 0 and not o["flags"].any() and np.all(o["leaf_cache"] == -1)
    gm.free()


@pytest.mark.parametrize("rot_deg,tra", [(0.3, 0.03), (1.0, 0.1)])
def test_lio_state_estimation_parity(pkg, ob, ctx, world, rot_deg, tra):
    keys, rn, nodes = world["export"]
    gm = pkg.PlaneMap(ctx, keys, rn, nodes, 1.0)
    pv = world["pv"]
    Rt, pt = world["truth"]
    R0, p0 = synth.perturb_poses((Rt[None], pt[None]), 9, np.deg2rad(rot_deg), tra, fix_first=False)
    cov = ob.default_cov()
    ref = world["map"].lio_state_estimation(pv, R0[0], p0[0], cov)
    out = gm.lio_state_estimation(pv, R0[0], p0[0], cov)
    assert out["iters"] == ref["iters"] and out["match_nums"] == ref["match_nums"] and out["ok"] == ref["ok"]
    assert np.max(np.abs(out["R"] - ref["R"])) < 1e-10 and np.max(np.abs(out["p"] - ref["p"])) < 1e-10
    assert rel(out["cov"], ref["cov"]) < 1e-9
    assert np.max(np.abs(out["p"] - pt)) < 0.2 * np.max(np.abs(p0[0] - pt))     # the update converged toward the planted truth
    gm.free()


# ---- A5: calcBodyVar / var_init / pvec_update on the device (voxelslam.hpp:180-265) ----------------------------
def _raw_scan(seed, n=20000):
    rng = np.random.Generator(np.random.PCG64(seed))
    q = scene.scan(scene.room(), np.eye(3), np.array([-7.0, 0.5, 1.2]), scene.xt32_dirs(700), rng, range_noise=0.01, blind=0.7)[:n].astype(np.float32)
    q[5, 2] = 0.0; q[17, 2] = -0.0          # calcBodyVar's pb[2] == 0 patch (:182-183) changes the stored point too
    return q


@pytest.mark.parametrize("dept,beam", [(0.01, 0.01), (0.02, 0.05)])
def test_var_init_parity(pkg, ob, ctx, dept, beam):
    q = _raw_scan(41)
    ext = synth.random_poses(8, 1, rot_sigma=0.05, tra_sigma=0.1)
    ref = ob.var_init_ext(q, ext, dept, beam)
    out = pkg.var_init(ctx, q, ext, dept, beam)
    assert np.abs(out[:, :3] - ref[:, :3]).max() <= 1e-14 * np.abs(ref[:, :3]).max()
    scale = np.abs(ref[:, 3:]).max(axis=1, keepdims=True)
    assert (np.abs(out[:, 3:] - ref[:, 3:]) <= 1e-12 * scale).all()
    ident = ob.var_init(q, dept, beam)      # the identity-extrinsic entry the map tests use
    out_i = pkg.var_init(ctx, q, (np.eye(3)[None], np.zeros((1, 3))), dept, beam)
    assert (np.abs(out_i - ident) <= 1e-12 * np.maximum(np.abs(ident).max(axis=1, keepdims=True), 1e-300)).all()
    assert out_i[5, 2] == 0.0001 and out_i[17, 2] == 0.0001
    assert pkg.var_init(ctx, q[:0], ext, dept, beam).shape == (0, 12)


def test_pvec_update_parity(pkg, ob, ctx):
    q = _raw_scan(42, n=7001)               # ragged last block
    pv = ob.var_init(q, 0.01, 0.01)
    pose = synth.random_poses(9, 1, rot_sigma=0.4, tra_sigma=3.0)
    rng = np.random.Generator(np.random.PCG64(3))
    A = rng.normal(size=(15, 15)); cov = A @ A.T * 1e-4
    ref, ref_w = ob.pvec_update(pv, pose, cov)
    out, out_w = pkg.pvec_update(ctx, pv, pose, cov)
    assert (out[:, :3] == ref[:, :3]).all()                                   # pnt is carried through untouched
    scale = np.abs(ref[:, 3:]).max(axis=1, keepdims=True)
    assert (np.abs(out[:, 3:] - ref[:, 3:]) <= 1e-12 * scale).all()
    assert np.abs(out_w - ref_w).max() <= 1e-14 * np.abs(ref_w).max()


def test_lio_state_estimation_from_raw_scan(pkg, ob, ctx, world):
    # 12 B/point upload + var_init on the device == host var_init + 96 B/point upload (same association, same state)
    keys, rn, nodes = world["export"]
    gm = pkg.PlaneMap(ctx, keys, rn, nodes, 1.0)
    rng = np.random.Generator(np.random.PCG64(77))
    Rt, pt = world["truth"]
    q = scene.scan(scene.room(), Rt, pt, scene.xt32_dirs(1000), rng, range_noise=0.01, blind=0.7).astype(np.float32)
    R0, p0 = synth.perturb_poses((Rt[None], pt[None]), 9, np.deg2rad(0.3), 0.03, fix_first=False)
    cov = ob.default_cov()
    ref = world["map"].lio_state_estimation(ob.var_init(q, 0.01, 0.01), R0[0], p0[0], cov)
    out = gm.lio_state_estimation(q, R0[0], p0[0], cov, xyz_ext=((np.eye(3)[None], np.zeros((1, 3))), 0.01, 0.01))
    assert out["iters"] == ref["iters"] and out["match_nums"] == ref["match_nums"] and out["ok"] == ref["ok"]
    assert np.max(np.abs(out["R"] - ref["R"])) < 1e-10 and np.max(np.abs(out["p"] - ref["p"])) < 1e-10
    gm.free()
