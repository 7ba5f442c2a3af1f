"""GPU parity for the BA evaluation passes, through the C ABI.

Tolerances: cost, Hess, JacT <= 1e-9 relative to max|.| (north_star's FP64 figure); merged cluster
<= 1e-13 relative (reordered FP64 sums).  Eigenvalues: the reference forms cov = P/N - c c^T
(tools.hpp:333-337), which cancels |c|^2 against the plane thickness, so lambda is only defined to
eps * max|P/N| ABSOLUTE; the test bounds |dlambda| by 1e-12 * max|P/N| per voxel, not relative to
lambda_0 itself."""
import numpy as np
import pytest

import synth

pytestmark = pytest.mark.gpu


def rel(a, b):
    return np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)


def check_eigvec(U_gpu, U_cpu, lam):
    # columns equal up to sign where eigenvalues are separated
    for k in range(3):
        d = np.abs(np.sum(U_gpu[:, :, k] * U_cpu[:, :, k], axis=1))
        gap = np.min(np.abs(lam[:, [k]] - np.delete(lam, k, axis=1)), axis=1) / np.max(np.abs(lam), axis=1)
        ok = gap > 1e-6
        assert np.all(d[ok] > 1 - 1e-9), (k, d[ok].min())


@pytest.mark.parametrize("V,W,occ", [(1, 2, 1.0), (257, 4, 0.5), (3000, 10, 0.6), (500, 20, 0.3)])
def test_residual_parity(pkg, ob, ctx, V, W, occ):
    fac = synth.random_factor(100 + V, V, W, occupancy=occ)
    poses = synth.perturb_poses(fac["poses"], 7, 0.01, 0.03)
    of = ob.Factor(W); of.set_dense(fac)
    r_cpu = of.residual(poses)
    cpu = of.get_dense()
    f = pkg.Factor(ctx, W)
    f.upload_dense(fac)
    assert f.num_voxels == V and f.num_voxel_frames == of.num_voxel_frames
    r_gpu = f.residual(poses)
    assert abs(r_gpu - r_cpu) <= 1e-9 * abs(r_cpu)
    ev, evec, mrg = f.download_cache()
    scale = np.max(np.abs(cpu["merged"]["P"]), axis=1, keepdims=True) / cpu["merged"]["N"][:, None]
    assert np.max(np.abs(ev - cpu["eigval"]) / scale) < 1e-12
    assert np.array_equal(mrg["N"], cpu["merged"]["N"])
    assert rel(mrg["P"], cpu["merged"]["P"]) < 1e-13 and rel(mrg["v"], cpu["merged"]["v"]) < 1e-13
    check_eigvec(evec.reshape(V, 3, 3).transpose(0, 2, 1), cpu["eigvec"].reshape(V, 3, 3).transpose(0, 2, 1), cpu["eigval"])
    f.free()


@pytest.mark.parametrize("V,W,occ", [(1, 2, 1.0), (2, 3, 1.0), (257, 4, 0.5), (3000, 10, 0.6), (500, 20, 0.3), (4000, 10, 1.0),
                                     (1000, 5, 0.7), (1000, 8, 0.6), (800, 12, 0.5), (600, 15, 0.5), (600, 16, 0.4)])
def test_hessian_parity(pkg, ob, ctx, V, W, occ):
    fac = synth.random_factor(200 + V, V, W, occupancy=occ)
    poses = synth.perturb_poses(fac["poses"], 9, 0.01, 0.03)
    of = ob.Factor(W); of.set_dense(fac)
    Ho, Jo, co = of.hessian(poses)          # uses the uploaded (push-time) cache, voxel_map.hpp:187-190
    f = pkg.Factor(ctx, W)
    f.upload_dense(fac)
    H, J, c = f.hessian(poses)
    assert abs(c - co) <= 1e-9 * abs(co)
    assert rel(J, Jo) < 1e-9
    assert rel(H, Ho) < 1e-9
    # structure: lower blocks are exact transposes of upper blocks, diagonal blocks are NOT symmetric
    for i in range(W):
        for j in range(i):
            assert np.array_equal(H[6 * i:6 * i + 6, 6 * j:6 * j + 6], H[6 * j:6 * j + 6, 6 * i:6 * i + 6].T)
    # CSR upload gives bit-identical results (same kernels, same plan)
    g = pkg.Factor(ctx, W)
    g.upload_csr(synth.dense_to_csr(fac))
    H2, J2, c2 = g.hessian(poses)
    assert np.array_equal(H, H2) and np.array_equal(J, J2) and c == c2
    # the generic pair-list path and the fused tensor-core path (W = 10, 20) agree to rounding,
    # and each is bitwise reproducible run to run (no atomics)
    g.force_generic(True)
    H3, J3, c3 = g.hessian(poses)
    assert rel(H3, Ho) < 1e-9 and rel(J3, Jo) < 1e-9 and rel(H3, H) < 1e-11
    H4, J4, c4 = f.hessian(poses)
    assert np.array_equal(H, H4) and np.array_equal(J, J4)
    f.free(); g.free()


@pytest.mark.parametrize("W", [8, 10])
def test_hessian_after_residual_uses_refreshed_cache(pkg, ob, ctx, W):
    # stateful cache semantics: B1 reads what the LAST B2 wrote
    V = 800
    fac = synth.random_factor(301, V, W)
    p1 = synth.perturb_poses(fac["poses"], 1, 0.02, 0.05)
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    of.residual(p1); f.residual(p1)
    Ho, Jo, co = of.hessian(p1)
    H, J, c = f.hessian(p1)
    assert abs(c - co) <= 1e-9 * abs(co) and rel(J, Jo) < 1e-9 and rel(H, Ho) < 1e-9
    f.free()


def test_partial_residual_then_full_hessian(pkg, ob, ctx):
    # the fused W = 10 path keeps per-voxel division planes next to the cache: a residual pass over PART of the voxels
    # refreshes them for that part only, and a Hessian over a wider range must rebuild them from the (mixed) cache
    V, W = 900, 10
    fac = synth.random_factor(311, V, W)
    p1 = synth.perturb_poses(fac["poses"], 2, 0.02, 0.05)
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    for head, end in [(0, 450), (200, 700)]:
        of.residual(p1, head, end); f.residual(p1, head, end)
        for h2, e2 in [(0, V), (head, end), (head + 10, end - 10)]:
            Ho, Jo, co = of.hessian(p1, h2, e2)
            H, J, c = f.hessian(p1, h2, e2)
            assert abs(c - co) <= 1e-9 * abs(co) and rel(J, Jo) < 1e-9 and rel(H, Ho) < 1e-9, (head, end, h2, e2)
    f.free()


@pytest.mark.parametrize("W", [5, 10])
def test_head_end_subrange(pkg, ob, ctx, W):
    V = 600
    fac = synth.random_factor(401, V, W)
    poses = fac["poses"]
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    for head, end in [(0, 300), (300, 600), (17, 18), (0, 0)]:
        Ho, Jo, co = of.hessian(poses, head, end)
        H, J, c = f.hessian(poses, head, end)
        if end > head:
            assert abs(c - co) <= 1e-9 * abs(co) and rel(J, Jo) < 1e-9 and rel(H, Ho) < 1e-9
        else:
            assert c == 0 and not H.any() and not J.any()
        assert abs(f.residual(poses, head, end) - of.residual(poses, head, end)) <= 1e-9 * max(abs(co), 1e-300)
    f.free()


def test_empty_and_reupload(pkg, ctx):
    f = pkg.Factor(ctx, 4)
    R, p = synth.random_poses(1, 4)
    H, J, c = f.hessian((R, p))
    assert c == 0 and not H.any() and not J.any()
    assert f.residual((R, p)) == 0
    fac = synth.random_factor(5, 50, 4)
    f.upload_dense(fac)
    r1 = f.residual(fac["poses"])
    f.upload_dense(synth.random_factor(6, 20, 4))
    assert f.num_voxels == 20
    f.upload_dense(fac)
    assert f.residual(fac["poses"]) == r1     # deterministic
    f.free()


@pytest.mark.parametrize("W,NW", [(5, 7), (10, 5), (20, 3)])
def test_batched_windows_match_single(pkg, ctx, W, NW):
    # n_windows independent factors in one launch == each evaluated alone
    facs = [synth.random_factor(500 + b, 120 + 10 * b, W) for b in range(NW)]
    csr = [synth.dense_to_csr(fc) for fc in facs]
    off = [0]; frames = []; parts = {k: [] for k in ("clusters", "sig", "coe", "eigval", "eigvec", "merged")}
    for b, c in enumerate(csr):
        off += (c["offsets"][1:] + off[-1]).tolist()
        frames.append(c["frames"] + b * W)
        for k in parts:
            parts[k].append(c[k])
    big = {k: np.concatenate(v) for k, v in parts.items()}
    big["offsets"] = np.array(off, dtype=np.int64)
    big["frames"] = np.concatenate(frames)
    R = np.concatenate([fc["poses"][0] for fc in facs]); p = np.concatenate([fc["poses"][1] for fc in facs])
    fb = pkg.Factor(ctx, W, NW)
    fb.upload_csr(big)
    rb = fb.residual((R, p))
    Hb, Jb, cb = fb.hessian((R, p))
    for b in range(NW):
        f1 = pkg.Factor(ctx, W); f1.upload_dense(facs[b])
        r1 = f1.residual(facs[b]["poses"])
        H1, J1, c1 = f1.hessian(facs[b]["poses"])
        assert abs(rb[b] - r1) <= 1e-13 * abs(r1)
        assert rel(Hb[b], H1) < 1e-12 and rel(Jb[b], J1) < 1e-12 and abs(cb[b] - c1) <= 1e-13 * abs(c1)
        f1.free()
    fb.free()


def test_upload_csr_validation_on_device(pkg, ctx):
    # malformed CSR input is rejected by the device-side validation pass (k_validate_vox / k_unpack_vf), error text kept
    W, NW = 5, 3
    good = synth.batched_factor_csr(11, NW, 40, W, occupancy=0.7)
    f = pkg.Factor(ctx, W, NW)
    f.upload_csr(good)
    assert f.num_voxels == 120

    def broken(mut):
        bad = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in good.items()}
        mut(bad)
        with pytest.raises(pkg.VxError) as e:
            f.upload_csr(bad)
        assert e.value.code == -2
        return str(e.value)

    o = good["offsets"]
    def swap(b): b["frames"][[o[3], o[3] + 1]] = b["frames"][[o[3] + 1, o[3]]]
    assert "ascending" in broken(swap)
    def span(b): b["frames"][o[4] - 1] = W            # last frame of voxel 3 (window 0) points into window 1
    assert "spans" in broken(span)
    def oob(b): b["frames"][o[120] - 1] = W * NW
    assert "bad frame" in broken(oob)
    def order(b): b["frames"][o[0]:o[1]] = W * 2 + np.arange(o[1] - o[0])   # voxel 0 in window 2 before window 0 voxels
    assert "grouped" in broken(order)
    def zero(b): b["clusters"]["N"][7] = 0
    assert "N == 0" in broken(zero)
    f.upload_csr(good)                                  # still usable afterwards
    assert np.isfinite(f.residual(good["poses"])).all()
    f.free()


def test_top_level_window_regime(pkg, ob, ctx):
    # C4-top: one problem over W ~ 399 submaps (voxelslam.cpp:3102-3127); every voxel is seen by a handful of frames
    # (k << W), Hess is 2394 x 2394.  The generic pair-list path carries it (the fused small-window kernel is W <= 20).
    W, V = 399, 900
    fac = synth.random_factor(9001, V, W, occupancy=0.008, min_frames=2, pts_per_frame=(6, 20))
    poses = synth.perturb_poses(fac["poses"], 11, 0.01, 0.03)
    of = ob.Factor(W); of.set_dense(fac)
    f = pkg.Factor(ctx, W); f.upload_dense(fac)
    k = f.num_voxel_frames / V
    assert 2 <= k < 8
    rc, rg = of.residual(poses), f.residual(poses)
    assert abs(rg - rc) <= 1e-9 * abs(rc)
    Ho, Jo, co = of.hessian(poses)
    H, J, c = f.hessian(poses)
    assert H.shape == (6 * W, 6 * W)
    assert abs(c - co) <= 1e-9 * abs(co) and rel(J, Jo) < 1e-9 and rel(H, Ho) < 1e-9
    # block sparsity: a frame pair without a common voxel has an exactly zero block
    occ = fac["plvec"]["N"] != 0
    common = (occ.T.astype(np.int32) @ occ.astype(np.int32)) > 0
    Hb = np.abs(H.reshape(W, 6, W, 6)).max(axis=(1, 3))
    assert not Hb[~common].any() and (Hb[common] > 0).all()
    f.free()
