"""GPU parity for Lidar_BA_Optimizer::damping_iter through vx_ba_lidar.

Tolerances: LM accept/reject sequence identical; converged poses <= 1e-8 abs (rad, m)
after identical iteration counts; first/last cost and the pre-gauge Hessian snapshot <= 1e-9 rel."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def pose_err(ps_a, ps_b):
    Ra = ps_a["R"].reshape(-1, 3, 3); Rb = ps_b["R"].reshape(-1, 3, 3)
    return max(np.max(np.abs(Ra - Rb)), np.max(np.abs(ps_a["p"] - ps_b["p"])))


@pytest.mark.parametrize("V,W,occ,rot,tra,iters,stale", [(1500, 10, 0.7, 0.005, 0.03, 4, False), (800, 6, 0.8, 0.01, 0.05, 3, False),
                                                          (600, 20, 0.4, 0.003, 0.02, 4, False), (900, 10, 0.9, 0.03, 0.3, 6, False),
                                                          (1500, 10, 0.7, 0.005, 0.03, 4, True)])
def test_ba_lidar_matches_oracle(pkg, ob, ctx, V, W, occ, rot, tra, iters, stale):
    fac = synth.random_factor(900 + V, V, W, occupancy=occ, min_frames=2)
    start = synth.perturb_poses(fac["poses"], 11, rot, tra)
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    if not stale:
        # GBA path: the cache is consistent with the starting poses (loop_refine.hpp:363-387)
        of.residual(start); f.residual(start)
    # stale: the first Hessian trusts a cache computed at other poses
    ref = of.ba_lidar(start, max_iter=iters, thd_num=2)
    out = f.ba_lidar(start, max_iter=iters, thd_num=2)
    assert ref["rc"] in (0, 1)
    assert out["accept"][0] == ref["accept"]
    assert int(out["converged"][0]) == ref["rc"]
    assert rel(out["resis"][0], ref["resis"]) < 1e-9
    assert rel(out["hess"][0], ref["hess"]) < 1e-9
    assert pose_err(out["poses"], ref["poses"]) < 1e-8
    # the BA actually moved toward the truth when started close to it
    if rot <= 0.005 and not stale:
        assert 1 in ref["accept"]
        truth = pkg.as_poses(fac["poses"])
        assert pose_err(out["poses"], truth) < pose_err(pkg.as_poses(start), truth)
    # stateful cache after the last step equals the oracle's
    ev, evec, mrg = f.download_cache()
    cpu = of.get_dense()
    scale = np.max(np.abs(cpu["merged"]["P"]), axis=1, keepdims=True) / cpu["merged"]["N"][:, None]
    assert np.max(np.abs(ev - cpu["eigval"]) / scale) < 1e-9
    f.free()


def test_ba_lidar_too_few_voxels(pkg, ctx):
    fac = synth.random_factor(5, 1, 4, min_frames=2)
    f = pkg.Factor(ctx, 4); f.upload_dense(fac)
    with pytest.raises(pkg.VxError) as e:
        f.ba_lidar(fac["poses"], max_iter=2, thd_num=2)
    assert e.value.code == -3      # VX_ERR_TOO_FEW: reference prints "Too Less Voxel" and exits (voxel_map.hpp:399-403)
    f.free()


def test_ba_lidar_batched_windows(pkg, ob, ctx):
    # 6 independent windows in one factor: each follows its own LM trajectory, equal to the oracle run alone
    W, NW = 10, 6
    big = synth.batched_factor_csr(77, NW, 300, W, occupancy=0.7)
    R, p = big["poses"]
    Rs, ps_ = synth.perturb_poses((R, p), 5, 0.004, 0.03, fix_first=False)
    for b in range(NW):   # pose 0 of every window is its gauge
        Rs[b * W] = R[b * W]; ps_[b * W] = p[b * W]
    f = pkg.Factor(ctx, W, NW); f.upload_csr(big)
    out = f.ba_lidar((Rs, ps_), max_iter=4, thd_num=2)
    for b in range(NW):
        d = synth.csr_window(big, b)
        of = ob.Factor(W); of.set_dense(d)
        ref = of.ba_lidar((Rs[b * W:(b + 1) * W], ps_[b * W:(b + 1) * W]), max_iter=4, thd_num=2)
        assert out["accept"][b] == ref["accept"]
        assert rel(out["resis"][b], ref["resis"]) < 1e-8
        assert pose_err(out["poses"][b * W:(b + 1) * W], ref["poses"]) < 1e-8
    f.free()


@pytest.mark.parametrize("n", [1, 31, 33, 64, 200, 777])
def test_device_ldlt_matches_oracle(pkg, ob, ctx, n):
    # the blocked device LDL^T (csrc/vx_ldlt.cu) against the oracle's sequential Eigen-style sweep, on a matrix that is NOT
    # exactly symmetric (only the lower triangle may be read) with a diagonal that forces non-trivial pivot orders and ties
    rng = np.random.Generator(np.random.PCG64(n))
    M = rng.normal(size=(n, n + 3))
    A = M @ M.T + np.diag(rng.uniform(0.1, 50.0, n))
    if n > 4:
        A[3, 3] = A[1, 1]                       # a tie on the diagonal: first index wins
    A_low = A.copy()
    A_bad = A.copy(); A_bad[np.triu_indices(n, 1)] += rng.normal(size=n * (n - 1) // 2)   # garbage above the diagonal
    b = rng.normal(size=n)
    want = ob.ldlt_solve(A_low, b)
    got = pkg.ldlt_solve(ctx, A_bad, b)
    assert np.abs(got - want).max() <= 1e-10 * np.abs(want).max()
    sym = np.tril(A) + np.tril(A, -1).T
    assert np.abs(sym @ got - b).max() <= 1e-9 * np.abs(b).max() * np.linalg.cond(sym) ** 0.5


@pytest.mark.parametrize("n", [70, 300])
def test_device_ldlt_tied_pivots(pkg, ob, ctx, n):
    # runs of equal diagonal entries (the six gauge rows 1 + u of the LM system are such a run): Eigen takes the first POSITION
    # among equals at every step, and positions move with the swaps; the O(n log n) pivot order of vx_ldlt.cu must reproduce it.
    # A is diagonally dominant with well separated groups, so a different order among equals changes the answer beyond rounding
    # only through the permutation bookkeeping (x comes back un-permuted) - which is exactly what is being checked.
    rng = np.random.Generator(np.random.PCG64(1000 + n))
    M = rng.normal(size=(n, n)) * 0.05
    A = M @ M.T
    levels = rng.choice([1.01, 3.0, 3.0, 7.5, 20.0], size=n)
    levels[:6] = 1.01
    A[np.arange(n), np.arange(n)] = levels            # exact ties, many per level, scattered positions
    b = rng.normal(size=n)
    want = ob.ldlt_solve(A, b)
    got = pkg.ldlt_solve(ctx, A, b)
    assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max()
    sym = np.tril(A) + np.tril(A, -1).T
    assert np.abs(sym @ got - b).max() <= 1e-10 * np.abs(b).max()


def test_ba_large_window_uses_device_solve(pkg, ob, ctx):
    # W = 64 -> 384 unknowns: vx_ba_lidar factorises on the device; same LM trace and poses as the oracle
    W, V = 64, 1500
    fac = synth.random_factor(3100, V, W, occupancy=0.06, min_frames=2, pts_per_frame=(6, 20))
    start = synth.perturb_poses(fac["poses"], 5, 0.003, 0.02)
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    of.residual(start); f.residual(start)
    a = f.ba_lidar(start, max_iter=3, thd_num=2)
    b = of.ba_lidar(start, max_iter=3, thd_num=2)
    assert a["accept"][0] == b["accept"] and 1 in b["accept"]
    assert np.abs(a["poses"]["R"] - b["poses"]["R"]).max() < 1e-8 and np.abs(a["poses"]["p"] - b["poses"]["p"]).max() < 1e-8
    f.free()


@pytest.mark.parametrize("V,W,occ,iters", [(1200, 10, 0.7, 4), (500, 20, 0.5, 4), (300, 21, 0.6, 3), (400, 4, 0.9, 5)])
def test_device_lm_step_is_the_host_sweep(pkg, ctx, V, W, occ, iters, monkeypatch):
    # k_lm_solve (gauge fix, damped LDL^T on the lower triangle, q1: csrc/vx_ba.cu) restates hm::ldlt_lower_solve operation for
    # operation: poses, Hess snapshot, costs, flags and the accept trace must be the same BYTES as with the host sweep
    # (VX_LM_HOST=1), including W = 20 / 21 (120 / 126 unknowns: the 128-thread instantiation, > 48 KB of shared memory).
    fac = synth.random_factor(4100 + V, V, W, occupancy=occ, min_frames=2)
    start = synth.perturb_poses(fac["poses"], 5, 0.01, 0.05)
    outs = []
    for host in (False, True):
        if host:
            monkeypatch.setenv("VX_LM_HOST", "1")
        else:
            monkeypatch.delenv("VX_LM_HOST", raising=False)
        f = pkg.Factor(ctx, W); f.upload_dense(fac); f.residual(start)
        outs.append(f.ba_lidar(start, max_iter=iters, thd_num=2))
        f.free()
    dev, hst = outs
    assert 1 in dev["accept"][0]
    for k in dev:
        assert np.ascontiguousarray(dev[k]).tobytes() == np.ascontiguousarray(hst[k]).tobytes(), k
