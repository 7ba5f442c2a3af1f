"""GPU parity for group A: A1 keys, A2 down_sampling_voxel, A8+A10 GBA voxel map -> LidarFactor.

Bars: keys, cell sets, per-cell counts, point->leaf assignment, per-(leaf,frame) N and the
plane/accept decisions: exact.  A2 centroids: bit-identical (the reference's sequential float mean is
replayed in input order).  Moments P, v: <= 1e-12 relative (reordered FP64 sums)."""
import numpy as np
import pytest

import scene
import synth

pytestmark = pytest.mark.gpu


def adversarial_coords(rng, n, vs):
    k = rng.integers(-2000, 2000, n).astype(np.float64)
    base = k * vs
    jit = rng.choice([0.0, 1e-12, -1e-12, 1e-7, -1e-7, 0.5 * vs, -0.25 * vs], n)
    x = base + jit
    x[: n // 50] = rng.choice([0.0, -0.0, -vs, vs, -1e-300, 1e-300, -2.0 ** 24 * vs, 2.0 ** 24 * vs + vs], n // 50)
    return x


@pytest.mark.parametrize("vs", [0.5, 1.0, 0.1, 0.3])
def test_keys_bit_exact(pkg, ob, ctx, vs):
    rng = np.random.Generator(np.random.PCG64(5))
    n = 300000
    xyz = np.stack([adversarial_coords(rng, n, vs) for _ in range(3)], axis=1)
    xyz[n // 2:] = rng.uniform(-250, 250, (n - n // 2, 3))
    assert np.array_equal(pkg.keys(ctx, xyz, vs), ob.keys_f64(xyz, vs))
    xf = xyz.astype(np.float32)
    assert np.array_equal(pkg.keys(ctx, xf, vs), ob.keys_f32(xf, vs))


def test_keys_empty(pkg, ctx):
    assert pkg.keys(ctx, np.zeros((0, 3)), 0.5).shape == (0, 3)


def cell_sorted(a, vs, ob):
    k = ob.keys_f32(np.stack([a["x"], a["y"], a["z"]], axis=1), vs)
    return a[np.lexsort((k[:, 2], k[:, 1], k[:, 0]))]


@pytest.mark.parametrize("n,vs", [(1, 0.1), (5000, 0.1), (60000, 0.1), (24000, 0.5)])
def test_downsample_voxel(pkg, ob, ctx, n, vs):
    rng = np.random.Generator(np.random.PCG64(n))
    pts = np.zeros(n, dtype=ob.point48_dtype)
    pts["x"] = rng.uniform(-8, 8, n); pts["y"] = rng.uniform(-6, 6, n); pts["z"] = rng.normal(0, 0.05, n)
    pts["intensity"] = rng.uniform(0, 255, n); pts["curvature"] = rng.uniform(0, 0.1, n)
    ref = ob.downsample_voxel(pts, vs)
    out = pkg.downsample_voxel(ctx, pts, vs)
    assert out.shape == ref.shape
    # the running mean drifts out of its birth cell now and then, so order both by the cell of the FIRST point,
    # which both sides keep in `intensity`-independent fields: sort by (count, intensity, x) instead
    key = lambda a: np.lexsort((a["x"], a["intensity"], a["curvature"]))
    r = ref[key(ref)]; o = out[key(out)]
    assert np.array_equal(o["curvature"], r["curvature"])          # per-cell counts
    for fld in ("x", "y", "z", "intensity", "normal_x"):
        assert np.array_equal(o[fld].view(np.uint32), r[fld].view(np.uint32)), fld     # bit-identical
    assert int(out["curvature"].sum()) == n


def test_downsample_small_voxel_is_identity(pkg, ob, ctx):
    pts = np.zeros(10, dtype=ob.point48_dtype); pts["x"] = np.arange(10)
    out = pkg.downsample_voxel(ctx, pts, 0.0005)       # tools.hpp:203
    assert np.array_equal(out, pts)


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("n,vs", [(1, 0.1), (5000, 0.1), (60000, 0.2), (24000, 0.5)])
def test_downsample_close(pkg, ob, ctx, n, vs):
    # A4 (tools.hpp:240-302): the kept record is an INPUT point, so compare whole 48-byte records as sets
    rng = np.random.Generator(np.random.PCG64(100 + n))
    pts = np.zeros(n, dtype=ob.point48_dtype)
    pts["x"] = rng.uniform(-8, 8, n); pts["y"] = rng.uniform(-6, 6, n); pts["z"] = rng.normal(0, 0.05, n)
    pts["intensity"] = rng.uniform(0, 255, n); pts["curvature"] = np.arange(n) * 1e-5          # per-point timestamp
    ref = ob.downsample_close(pts, vs)
    out = pkg.downsample_close(ctx, pts, vs)
    assert out.shape == ref.shape
    key = lambda a: np.argsort(a["curvature"], kind="stable")
    assert np.array_equal(out[key(out)].view(np.uint8), ref[key(ref)].view(np.uint8))
    assert np.array_equal(pkg.downsample_close(ctx, pts, 0.0005), pts)                          # tools.hpp:242
    assert pkg.downsample_close(ctx, pts[:0], vs).shape == (0,)


@pytest.mark.parametrize("n,vs", [(1, 0.05), (7000, 0.05), (64000, 0.1)])
def test_downsample_pvec(pkg, ob, ctx, n, vs):
    # A3 (voxel_map.hpp:39-81): double running means, bit-identical after the float cast; order differs (sorted by cell)
    rng = np.random.Generator(np.random.PCG64(200 + n))
    pv = np.zeros((n, 12))
    pv[:, 0] = rng.uniform(-8, 8, n); pv[:, 1] = rng.uniform(-6, 6, n); pv[:, 2] = rng.normal(0, 0.05, n)
    A = rng.normal(0, 0.01, (n, 3, 3)); pv[:, 3:] = (A @ np.transpose(A, (0, 2, 1))).reshape(n, 9)
    ref = ob.downsample_pvec(pv, vs)
    out = pkg.downsample_pvec(ctx, pv, vs)
    assert out.shape == ref.shape
    key = lambda a: np.lexsort((a["z"], a["y"], a["x"]))
    r = ref[key(ref)]; o = out[key(out)]
    for fld in ("x", "y", "z", "normal_x", "normal_y", "normal_z", "_p0", "intensity", "curvature"):
        assert np.array_equal(_bits(o[fld]), _bits(r[fld])), fld
    assert pkg.downsample_pvec(ctx, pv[:0], vs).shape == (0,)


def canon(fac, W):
    """order voxels by merged centroid (unique per voxel) and expand to dense (V, W) N / moments"""
    m = fac["merged"]
    cen = m["v"] / m["N"][:, None]
    order = np.lexsort((np.round(cen[:, 2], 6), np.round(cen[:, 1], 6), np.round(cen[:, 0], 6)))
    return order


def dense_from_csr(c, W):
    V = c["merged"].shape[0]
    pl = np.zeros((V, W), dtype=synth.cluster_dtype)
    vi = np.repeat(np.arange(V), np.diff(c["offsets"]))
    pl[vi, c["frames"] % W] = c["clusters"]
    return pl


@pytest.mark.parametrize("vs,min_eig,thr,L", [(0.5, 0.0025, 0.25, 2), (1.0, 0.01, 0.5, 2), (2.0, 0.1, 0.25, 3), (0.5, 0.0025, 0.25, 0), (1.0, 0.01, 0.5, 1)])
def test_gba_build_matches_oracle(pkg, ob, ctx, vs, min_eig, thr, L):
    clouds, (R, p) = scene.config1(pts_per_scan=6000)
    W = len(clouds)
    start = synth.perturb_poses((R, p), 3, 0.004, 0.03)
    of = ob.gba_build(clouds, start, vs, min_eig, [thr] * 4, L)
    cpu = of.get_dense()
    f = pkg.Factor(ctx, W)
    f.gba_build(clouds, start, vs, min_eig, [thr] * 4, L)
    assert f.num_voxels == of.num_voxels and f.num_voxel_frames == of.num_voxel_frames
    g = f.download_csr()
    pl = dense_from_csr(g, W)
    og, oc = canon(g, W), canon(cpu, W)
    assert np.array_equal(g["merged"]["N"][og], cpu["merged"]["N"][oc])
    assert np.array_equal(pl["N"][og], cpu["plvec"]["N"][oc])                       # per-(leaf, frame) membership counts
    rel = lambda a, b: np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)
    assert rel(pl["P"][og], cpu["plvec"]["P"][oc]) < 1e-12 and rel(pl["v"][og], cpu["plvec"]["v"][oc]) < 1e-12
    assert rel(g["merged"]["P"][og], cpu["merged"]["P"][oc]) < 1e-12 and rel(g["merged"]["v"][og], cpu["merged"]["v"][oc]) < 1e-12
    scale = np.max(np.abs(cpu["merged"]["P"][oc]), axis=1, keepdims=True) / cpu["merged"]["N"][oc][:, None]
    assert np.max(np.abs(g["eigval"][og] - cpu["eigval"][oc]) / scale) < 1e-12
    assert np.all(g["coe"] == 1.0)
    # the factor built on the device evaluates like the oracle's
    Ho, Jo, co = of.hessian(start)
    H, J, c = f.hessian(start)
    assert abs(c - co) <= 1e-9 * abs(co) and rel(J, Jo) < 1e-8 and rel(H, Ho) < 1e-8
    f.free()


def test_gba_build_ragged_and_empty_clouds(pkg, ob, ctx):
    # empty scans, a one-point scan and a seven-point scan inside a window: the per-cloud bounding boxes, the frame walk of the key
    # kernel and the compact index arrays must cope with frames that own no instance at all
    clouds, (R, p) = scene.config1(pts_per_scan=4000)
    W = len(clouds)
    clouds = list(clouds)
    clouds[0] = clouds[0][:0]; clouds[2] = clouds[2][:0]; clouds[5] = clouds[5][:1]; clouds[7] = clouds[7][:7]; clouds[W - 1] = clouds[W - 1][:0]
    start = synth.perturb_poses((R, p), 5, 0.004, 0.03)
    of = ob.gba_build(clouds, start, 1.0, 0.01, [0.5] * 4, 2)
    cpu = of.get_dense()
    f = pkg.Factor(ctx, W)
    f.gba_build(clouds, start, 1.0, 0.01, [0.5] * 4, 2)
    assert f.num_voxels == of.num_voxels and f.num_voxel_frames == of.num_voxel_frames and f.num_voxels > 0
    g = f.download_csr()
    pl = dense_from_csr(g, W)
    og, oc = canon(g, W), canon(cpu, W)
    assert np.array_equal(g["merged"]["N"][og], cpu["merged"]["N"][oc])
    assert np.array_equal(pl["N"][og], cpu["plvec"]["N"][oc])
    assert not np.any(pl["N"][:, [0, 2, W - 1]])                     # the empty scans own nothing
    rel = lambda a, b: np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300)
    assert rel(pl["P"][og], cpu["plvec"]["P"][oc]) < 1e-12 and rel(g["merged"]["P"][og], cpu["merged"]["P"][oc]) < 1e-12
    f.free()
    # every scan empty: an empty factor, not an error
    f0 = pkg.Factor(ctx, W)
    f0.gba_build([c[:0] for c in clouds], start, 1.0, 0.01, [0.5] * 4, 2)
    assert f0.num_voxels == 0 and f0.num_voxel_frames == 0
    f0.free()


def test_gba_build_batched_windows(pkg, ob, ctx):
    # 3 overlapping windows (stride 5, size 10 over 20 keyframes) in one call == each window built alone
    clouds, (R, p) = scene.config1(W=20, pts_per_scan=3000)
    W, NW = 10, 3
    fc = np.concatenate([np.arange(5 * b, 5 * b + W) for b in range(NW)]).astype(np.int32)
    Rf, pf = R[fc], p[fc]
    fb = pkg.Factor(ctx, W, NW)
    fb.gba_build(clouds, (Rf, pf), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)
    gb = fb.download_csr()
    first = gb["frames"][gb["offsets"][:-1]] // W
    tot = 0
    for b in range(NW):
        f1 = pkg.Factor(ctx, W)
        f1.gba_build([clouds[i] for i in fc[b * W:(b + 1) * W]], (Rf[b * W:(b + 1) * W], pf[b * W:(b + 1) * W]), 1.0, 0.01, [0.5] * 4, 2)
        g1 = f1.download_csr()
        vs_ = np.nonzero(first == b)[0]
        assert vs_.size == g1["merged"].shape[0]
        assert np.array_equal(gb["merged"][vs_], g1["merged"])            # same order, bitwise
        e0, e1 = gb["offsets"][vs_[0]], gb["offsets"][vs_[-1] + 1]
        assert np.array_equal(gb["frames"][e0:e1] - b * W, g1["frames"]) and np.array_equal(gb["clusters"][e0:e1], g1["clusters"])
        tot += vs_.size
        f1.free()
    assert tot == gb["merged"].shape[0]
    fb.free()
