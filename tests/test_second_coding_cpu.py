"""Independent SECOND codings of the rows whose quirks matter most, compared with the oracle the way
test_golden_cpu.py::test_keys_two_codings_and_fixture does for A1.  Nothing here is derived from oracle/*.hpp:

* B2 / B1 (voxel_map.hpp:150-325): the cost is written from the MATHS — sum over voxels of coe * lambda_min of the covariance of the
  voxel's points, the points of frame i moved by (R_i Exp(dtheta_i), p_i + dp_i) — in torch float64 on the second moments, and
  differentiated by autograd: the exact gradient must equal JacT, the exact (symmetric) Hessian must equal the symmetric part of the
  reference's Hess (its only asymmetry is the -0.5 hat(jjt) term of voxel_map.hpp:242).
* D1 (voxel_map.hpp:1649-1721, :2167-2205): root lookup, descent and the float-narrowed gates in numpy float32 / float64, per point,
  against the oracle's association flags on a map the oracle built."""
import numpy as np
import pytest
import torch

import scene
import synth


def _hat(w):
    z = torch.zeros((), dtype=w.dtype)
    return torch.stack([torch.stack([z, -w[2], w[1]]), torch.stack([w[2], z, -w[0]]), torch.stack([-w[1], w[0], z])])


def _exp(w):
    # Rodrigues as a smooth function of w (series near 0 so that autograd is exact at w = 0)
    th2 = (w * w).sum()
    K = _hat(w)
    a = 1 - th2 / 6 + th2 * th2 / 120          # sin(th)/th
    b = 0.5 - th2 / 24 + th2 * th2 / 720       # (1 - cos(th))/th^2
    return torch.eye(3, dtype=w.dtype) + a * K + b * (K @ K)


def _cost(delta, fac, R, p):
    """delta (W,6) = [dtheta, dp] per frame; fac: the dense factor (second moments per voxel and frame)."""
    V, W = fac["plvec"].shape
    total = torch.zeros((), dtype=torch.float64)
    for a in range(V):
        P = torch.tensor(fac["sig"][a]["P"].reshape(3, 3).T.copy()); v = torch.tensor(fac["sig"][a]["v"].copy()); N = float(fac["sig"][a]["N"])
        for i in range(W):
            c = fac["plvec"][a, i]
            if c["N"] == 0:
                continue
            Ri = torch.tensor(R[i]) @ _exp(delta[i, :3]); pi = torch.tensor(p[i]) + delta[i, 3:]
            Pi = torch.tensor(c["P"].reshape(3, 3).T.copy()); vi = torch.tensor(c["v"].copy()); n = float(c["N"])
            Rv = Ri @ vi
            # second moments of the moved points: sum (R x + p)(R x + p)^T
            P = P + Ri @ Pi @ Ri.T + torch.outer(Rv, pi) + torch.outer(pi, Rv) + n * torch.outer(pi, pi)
            v = v + Rv + n * pi
            N += n
        C = P / N - torch.outer(v, v) / (N * N)
        total = total + float(fac["coe"][a]) * torch.linalg.eigvalsh(C)[0]
    return total


@pytest.mark.parametrize("seed,V,W", [(11, 40, 5), (12, 25, 8)])
def test_b1_b2_from_the_maths_autograd(ob, seed, V, W):
    fac = synth.random_factor(seed, V, W, occupancy=0.7, min_frames=2)
    R, p = synth.perturb_poses(fac["poses"], seed + 1, 0.01, 0.03)
    of = ob.Factor(W); of.set_dense(fac)
    r_o = of.residual((R, p))                       # refreshes the cache acc_evaluate2 reads
    H_o, J_o, c_o = of.hessian((R, p))
    z = torch.zeros((W, 6), dtype=torch.float64, requires_grad=True)
    c_t = _cost(z, fac, R, p)
    assert abs(float(c_t) - r_o) <= 1e-9 * abs(r_o) and abs(c_o - r_o) <= 1e-12 * abs(r_o)
    (g,) = torch.autograd.grad(c_t, z, create_graph=True)
    g_flat = g.reshape(-1)
    assert np.abs(g_flat.detach().numpy() - J_o).max() <= 1e-8 * np.abs(J_o).max()
    H_t = np.stack([torch.autograd.grad(g_flat[k], z, retain_graph=True)[0].reshape(-1).numpy() for k in range(6 * W)])
    H_sym = 0.5 * (H_o + H_o.T)
    assert np.abs(H_t - H_sym).max() <= 1e-7 * np.abs(H_sym).max()
    # the reference's Hess is NOT symmetric: its antisymmetric part lives in the rotation-rotation diagonal blocks only (voxel_map.hpp:242)
    A = 0.5 * (H_o - H_o.T)
    mask = np.zeros_like(A, dtype=bool)
    for i in range(W):
        mask[6 * i:6 * i + 3, 6 * i:6 * i + 3] = True
    assert np.abs(A[~mask]).max() <= 1e-12 * np.abs(H_o).max() and np.abs(A[mask]).max() > 0


def _np_keys(w, vs):
    loc = (np.asarray(w, dtype=np.float64) / vs).astype(np.float32)
    loc = np.where(loc < 0, loc - np.float32(1.0), loc)
    return loc.astype(np.int64)


def _match_numpy(keys, root_node, nodes, vs, wld, var_wld):
    """match() + OctoTree::match for ONE world point, float-narrowed gates as the reference declares its locals (float dis_to_plane,
    float dis_to_center, float range_dis against 3 * plane.radius... see voxel_map.hpp:1652-1691) -> flag."""
    k = tuple(_np_keys(wld, vs).tolist())
    node = root_node.get(k, -1)
    while node >= 0:
        nd = nodes[node]
        if nd["octo_state"] == 0:
            if not nd["is_plane"]:
                return 0
            c = nd["p_center"]; n = nd["p_normal"]
            dis_to_plane = np.float32(abs(float(n @ (wld - c))))
            dis_to_center = np.float32(float((c - wld) @ (c - wld)))
            range_dis = np.float32(dis_to_center - np.float32(dis_to_plane * dis_to_plane))
            if not (range_dis <= np.float32(3 * 3) * np.float32(nd["radius"])):
                return 0
            J = np.concatenate([wld - c, -n])
            sigma_l = float(J @ nd["plane_var"].reshape(6, 6).T @ J) + float(n @ var_wld @ n)
            return int(float(dis_to_plane) < 3 * np.sqrt(sigma_l))
        ctr = nd["center"]
        node = int(nd["child"][4 * int(wld[0] > ctr[0]) + 2 * int(wld[1] > ctr[1]) + int(wld[2] > ctr[2])])
    return 0


def test_d1_gates_numpy_second_coding(ob):
    rng = np.random.Generator(np.random.PCG64(20241008 + 77))
    rects = scene.room()
    W = 6
    R, p = scene.arc_trajectory(W + 1, 5, step=0.5, yaw_step_deg=3.0)
    om = ob.LocalMap(win_size=W, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2, dept_err=0.01, beam_err=0.01)
    cov = ob.default_cov()
    for i in range(W):
        om.push(scene.scan(rects, R[i], p[i], scene.xt32_dirs(300), rng, range_noise=0.01, blind=0.7), (R[i:i + 1], p[i:i + 1]), cov)
    keys, rn, nodes = om.export()
    root_node = {tuple(k.tolist()): int(n) for k, n in zip(keys, rn)}
    q = scene.scan(rects, R[W], p[W], scene.xt32_dirs(120), rng, range_noise=0.01, blind=0.7)
    pv = ob.var_init(q, 0.01, 0.01)
    R0, p0 = synth.perturb_poses((R[W:W + 1], p[W:W + 1]), 3, 0.003, 0.02, fix_first=False)
    ref = om.lio_pass(pv, (R0, p0), cov)
    Rq, pq = R0[0], p0[0]
    rot_var, tsl_var = cov[0:3, 0:3], cov[3:6, 3:6]
    flags = np.zeros(pv.shape[0], dtype=np.int32)
    for i in range(pv.shape[0]):
        x = pv[i, :3]; var = pv[i, 3:].reshape(3, 3).T
        hx = np.array([[0, -x[2], x[1]], [x[2], 0, -x[0]], [-x[1], x[0], 0]])
        vw = Rq @ var @ Rq.T + hx @ rot_var @ hx.T + tsl_var                    # voxelslam.cpp:1011
        flags[i] = _match_numpy(keys, root_node, nodes, 1.0, Rq @ x + pq, vw)
    assert ref["match_num"] > 500
    diff = int((flags != ref["flags"]).sum())
    assert diff <= max(1, pv.shape[0] // 2000), diff      # numpy sums in another order: a gate may flip on a handful of borderline points
