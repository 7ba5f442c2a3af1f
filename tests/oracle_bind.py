"""ctypes binding over oracle/liboracle.so — the CPU restatement of the reference (test infrastructure).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(ROOT, "oracle", "liboracle.so")

cluster_dtype = np.dtype([("P", "<f8", (9,)), ("v", "<f8", (3,)), ("N", "<i4"), ("_pad", "<i4")], align=True)
pose_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,))], align=True)

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            import sys
            sys.path.insert(0, ROOT)
            import __graft_entry__ as ge
            ge.build_oracle()
        _lib = C.CDLL(LIB_PATH)
        _lib.orc_factor_create.restype = C.c_void_p
        _lib.orc_factor_num_voxels.restype = C.c_int64
        _lib.orc_factor_num_voxel_frames.restype = C.c_int64
        _lib.orc_ba_divide_thread.restype = C.c_double
        _lib.orc_now.restype = C.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def as_poses(poses):
    if isinstance(poses, np.ndarray) and poses.dtype == pose_dtype:
        return np.ascontiguousarray(poses)
    R, p = poses
    R = np.asarray(R, dtype=np.float64)
    out = np.zeros(R.shape[0], dtype=pose_dtype)
    out["R"] = np.transpose(R, (0, 2, 1)).reshape(-1, 9)
    out["p"] = np.asarray(p, dtype=np.float64)
    return out


def poses_to_Rp(ps):
    R = np.transpose(ps["R"].reshape(-1, 3, 3), (0, 2, 1)).copy()
    return R, ps["p"].copy()


def keys_f64(xyz, voxel_size):
    xyz = np.ascontiguousarray(xyz, dtype=np.float64)
    out = np.zeros(xyz.shape, dtype=np.int64)
    lib().orc_keys_f64(_p(xyz), C.c_int64(xyz.size // 3), C.c_double(voxel_size), _p(out))
    return out


def keys_f32(xyz, voxel_size):
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    out = np.zeros(xyz.shape, dtype=np.int64)
    lib().orc_keys_f32(_p(xyz), C.c_int64(xyz.size // 3), C.c_double(voxel_size), _p(out))
    return out


def eig3(A):
    """A: (n,3,3) symmetric -> (w (n,3) ascending, U (n,3,3) columns = eigenvectors)."""
    A = np.ascontiguousarray(np.transpose(np.asarray(A, dtype=np.float64), (0, 2, 1)))
    n = A.shape[0]
    w = np.zeros((n, 3)); U = np.zeros((n, 9))
    lib().orc_eig3(_p(A), C.c_int64(n), _p(w), _p(U))
    return w, np.transpose(U.reshape(n, 3, 3), (0, 2, 1)).copy()


def ldlt_solve(A, b):
    A = np.asarray(A, dtype=np.float64)
    n = A.shape[0]
    Ac = np.ascontiguousarray(A.T)  # column-major
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros(n)
    lib().orc_ldlt_solve(_p(Ac), _p(b), C.c_int(n), _p(x))
    return x


def inverse15(A):
    Ac = np.ascontiguousarray(np.asarray(A, dtype=np.float64).T)
    out = np.zeros((15, 15))
    lib().orc_inverse15(_p(Ac), _p(out))
    return out.T.copy()


def so3_exp(w):
    w = np.ascontiguousarray(w, dtype=np.float64)
    R = np.zeros(9)
    lib().orc_exp(_p(w), _p(R))
    return R.reshape(3, 3).T.copy()


def so3_log(R):
    Rc = np.ascontiguousarray(np.asarray(R, dtype=np.float64).T)
    w = np.zeros(3)
    lib().orc_log(_p(Rc), _p(w))
    return w


def cluster_transform(cl, pose):
    cin = np.zeros(1, dtype=cluster_dtype); cin[0] = cl
    ps = as_poses(pose)
    out = np.zeros(1, dtype=cluster_dtype)
    lib().orc_cluster_transform(_p(cin), _p(ps), _p(out))
    return out[0]


point48_dtype = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("_p0", "<f4"), ("normal_x", "<f4"), ("normal_y", "<f4"),
                          ("normal_z", "<f4"), ("_p1", "<f4"), ("intensity", "<f4"), ("curvature", "<f4"), ("_p2", "<f4"), ("_p3", "<f4")])
assert point48_dtype.itemsize == 48


def downsample_voxel(pts, voxel_size):
    pts = np.ascontiguousarray(pts, dtype=point48_dtype)
    out = np.zeros(pts.shape[0], dtype=point48_dtype)
    lib().orc_downsample_voxel.restype = C.c_int64
    n = lib().orc_downsample_voxel(_p(pts), C.c_int64(pts.shape[0]), C.c_double(voxel_size), _p(out))
    return out[:n]


def downsample_close(pts, voxel_size):
    pts = np.ascontiguousarray(pts, dtype=point48_dtype)
    out = np.zeros(pts.shape[0], dtype=point48_dtype)
    lib().orc_downsample_close.restype = C.c_int64
    n = lib().orc_downsample_close(_p(pts), C.c_int64(pts.shape[0]), C.c_double(voxel_size), _p(out))
    return out[:n]


def downsample_pvec(pv, voxel_size):
    pv = np.ascontiguousarray(pv, dtype=np.float64)
    out = np.zeros(pv.shape[0], dtype=point48_dtype)
    lib().orc_downsample_pvec.restype = C.c_int64
    n = lib().orc_downsample_pvec(_p(pv), C.c_int64(pv.shape[0]), C.c_double(voxel_size), _p(out))
    return out[:n]


def gba_build(clouds, poses, gba_voxel_size, gba_min_eigen_value, eig_array, max_layer):
    """clouds: list of (n_i,3) float32 local-frame clouds, one per keyframe -> orc LidarFactor (A8 + A10)."""
    W = len(clouds)
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(W + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses)
    ea = np.ascontiguousarray(eig_array, dtype=np.float64)
    lib().orc_gba_build.restype = C.c_void_p
    h = lib().orc_gba_build(_p(xyz), _p(off), _p(ps), C.c_int(W), C.c_double(gba_voxel_size), C.c_double(gba_min_eigen_value),
                            _p(ea), C.c_int(max_layer))
    return Factor(W, handle=h)


edge_dtype = np.dtype([("m1", "<i4"), ("m2", "<i4"), ("id1", "<i4"), ("id2", "<i4"), ("rot", "<f8", (9,)), ("tra", "<f8", (3,)), ("cov", "<f8", (6,))])
assert edge_dtype.itemsize == 160


def hba_add_edge(clouds, poses, local, gba, max_layer, max_iter, thread_num=2, ids=None, mps=None, want_cloud=False):
    """HBA_add_edge (voxelslam.cpp:2822-3015) for one window.  local = (voxel_size, min_eig, thre4), gba likewise."""
    W = len(clouds)
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(W + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses).copy()
    jl = 6 * W
    edges = np.zeros(W * (W - 1) // 2, dtype=edge_dtype); hess = np.zeros((jl, jl))
    rf = np.zeros(max(max_iter, 1)); rl = np.zeros(max(max_iter, 1)); nv = np.zeros(max(max_iter, 1), dtype=np.int32); no = C.c_int(0)
    merged = np.zeros(xyz.shape[0] if want_cloud else 1, dtype=point48_dtype); mn = C.c_int64(0)
    ids_ = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32)
    mps_ = None if mps is None else np.ascontiguousarray(mps, dtype=np.int32)
    lt = np.ascontiguousarray(local[2], dtype=np.float64); gt = np.ascontiguousarray(gba[2], dtype=np.float64)
    lib().orc_hba_add_edge.restype = C.c_int64
    ne = lib().orc_hba_add_edge(_p(xyz), _p(off), _p(ids_), _p(mps_), _p(ps), C.c_int(W), C.c_double(local[0]), C.c_double(local[1]), _p(lt),
                                C.c_double(gba[0]), C.c_double(gba[1]), _p(gt), C.c_int(max_layer), C.c_int(max_iter), C.c_int(thread_num),
                                _p(edges), _p(hess), _p(rf), _p(rl), _p(nv), C.byref(no), C.c_int(int(want_cloud)), _p(merged), C.byref(mn))
    k = no.value
    return dict(n_edges=int(ne), edges=edges[:max(ne, 0)], poses=ps, hess=hess.T.copy(), resis_first=rf[:k], resis_last=rl[:k], n_voxels=nv[:k],
                merged=merged[:mn.value])


class Factor:
    """orc::LidarFactor (voxel_map.hpp:124-339 restated)."""

    def __init__(self, win_size, handle=None, borrowed=False):
        self.W = win_size
        self.borrowed = borrowed
        self.h = C.c_void_p(handle if handle is not None else lib().orc_factor_create(C.c_int(win_size)))

    def __del__(self):
        if getattr(self, "h", None):
            if not self.borrowed:
                lib().orc_factor_free(self.h)
            self.h = None

    @property
    def num_voxels(self):
        return int(lib().orc_factor_num_voxels(self.h))

    @property
    def num_voxel_frames(self):
        return int(lib().orc_factor_num_voxel_frames(self.h))

    def set_dense(self, fac):
        a = [np.ascontiguousarray(fac[k]) for k in ("plvec", "sig", "coe", "eigval", "eigvec", "merged")]
        lib().orc_factor_set_dense(self.h, C.c_int64(a[1].shape[0]), *[_p(x) for x in a])

    def get_dense(self):
        V, W = self.num_voxels, self.W
        fac = dict(plvec=np.zeros((V, W), dtype=cluster_dtype), sig=np.zeros(V, dtype=cluster_dtype), coe=np.zeros(V),
                   eigval=np.zeros((V, 3)), eigvec=np.zeros((V, 9)), merged=np.zeros(V, dtype=cluster_dtype))
        lib().orc_factor_get_dense(self.h, *[_p(fac[k]) for k in ("plvec", "sig", "coe", "eigval", "eigvec", "merged")])
        return fac

    def hessian(self, poses, head=0, end=-1):
        ps = as_poses(poses)
        if end < 0:
            end = self.num_voxels
        jl = 6 * self.W
        H = np.zeros((jl, jl)); J = np.zeros(jl); c = C.c_double(0)
        lib().orc_factor_hessian(self.h, _p(ps), C.c_int64(head), C.c_int64(end), _p(H), _p(J), C.byref(c))
        return H.T.copy(), J, c.value

    def residual(self, poses, head=0, end=-1):
        ps = as_poses(poses)
        if end < 0:
            end = self.num_voxels
        c = C.c_double(0)
        lib().orc_factor_residual(self.h, _p(ps), C.c_int64(head), C.c_int64(end), C.byref(c))
        return c.value

    def divide_thread(self, poses, thd_num=2, use_threads=False, want=True):
        ps = as_poses(poses)
        jl = 6 * self.W
        H = np.zeros((jl, jl)) if want else None
        J = np.zeros(jl) if want else None
        r = lib().orc_ba_divide_thread(self.h, _p(ps), C.c_int(thd_num), C.c_int(int(use_threads)), _p(H), _p(J))
        return (H.T.copy() if want else None), J, r

    def ba_lidar(self, poses, max_iter=3, thd_num=2, use_threads=False):
        ps = as_poses(poses).copy()
        jl = 6 * self.W
        hess = np.zeros((jl, jl)); resis = np.zeros(2)
        trace = np.zeros(max_iter + 1, dtype=np.int32); nt = C.c_int(0)
        rc = lib().orc_ba_lidar(self.h, _p(ps), _p(hess), _p(resis), C.c_int(max_iter), C.c_int(thd_num),
                                C.c_int(int(use_threads)), _p(trace), C.byref(nt))
        return dict(rc=rc, poses=ps, hess=hess.T.copy(), resis=resis, accept=trace[:nt.value].tolist())


    def ba_li(self, states, imus, imu_coef=1e-4, gravity=False, max_iter=2, use_threads=False):
        """LI_BA_Optimizer (gravity False) / LI_BA_OptimizerGravity (True) ::damping_iter."""
        x = np.ascontiguousarray(states, dtype=imu_state_dtype).copy(); pr = np.ascontiguousarray(imus, dtype=imu_pre_dtype).copy()
        n = 15 * self.W + (3 if gravity else 0)
        hess = np.zeros((n, n)); resis = np.zeros(2); trace = np.zeros(max(max_iter, 3) + 1, dtype=np.int32); nt = C.c_int(0)
        lib().orc_ba_li(self.h, _p(x), _p(pr), C.c_double(imu_coef), C.c_int(int(gravity)), C.c_int(max_iter), C.c_int(int(use_threads)),
                        _p(resis), _p(hess), _p(trace), C.byref(nt))
        return dict(states=x, imus=pr, hess=hess.T.copy(), resis=resis, accept=trace[:nt.value].tolist())


imu_pre_dtype = np.dtype([("R_delta", "<f8", (9,)), ("p_delta", "<f8", (3,)), ("v_delta", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)),
                          ("R_bg", "<f8", (9,)), ("p_bg", "<f8", (9,)), ("p_ba", "<f8", (9,)), ("v_bg", "<f8", (9,)), ("v_ba", "<f8", (9,)),
                          ("dtime", "<f8"), ("dbg", "<f8", (3,)), ("dba", "<f8", (3,)), ("dbg_buf", "<f8", (3,)), ("dba_buf", "<f8", (3,)),
                          ("cov", "<f8", (225,))])
imu_noise_dtype = np.dtype([("scale_gravity", "<f8"), ("noise_meas", "<f8", (36,)), ("noise_walk", "<f8", (36,))])
imu_sample_dtype = np.dtype([("t", "<f8"), ("gyr", "<f8", (3,)), ("acc", "<f8", (3,))])
imu_state_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,)), ("v", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)), ("g", "<f8", (3,))])


def imu_pre_new(samples, noise, bg=(0, 0, 0), ba=(0, 0, 0)):
    pre = np.zeros(1, dtype=imu_pre_dtype)
    bg = np.ascontiguousarray(bg, dtype=np.float64); ba = np.ascontiguousarray(ba, dtype=np.float64)
    lib().orc_imu_pre_init(_p(pre), _p(bg), _p(ba))
    s = np.ascontiguousarray(samples, dtype=imu_sample_dtype)
    lib().orc_imu_pre_push(_p(pre), _p(np.ascontiguousarray(noise, dtype=imu_noise_dtype)), _p(s), C.c_int(s.shape[0]))
    return pre


def imu_pre_evaluate(pre, st1, st2, with_gravity=False, jac=True):
    nc = 33 if with_gravity else 30
    jtj = np.zeros((nc, nc)); gg = np.zeros(nc); cost = np.zeros(1)
    lib().orc_imu_pre_evaluate(_p(pre), _p(st1), _p(st2), C.c_int(int(with_gravity)), _p(jtj) if jac else None, _p(gg) if jac else None, _p(cost))
    return float(cost[0]), jtj.T.copy(), gg


map_node_dtype = np.dtype([("center", "<f8", (3,)), ("quater_length", "<f4"), ("octo_state", "<i4"), ("child", "<i4", (8,)), ("is_plane", "<i4"),
                           ("radius", "<f4"), ("p_center", "<f8", (3,)), ("p_normal", "<f8", (3,)), ("plane_var", "<f8", (36,))])
assert map_node_dtype.itemsize == 408


def btc_plane_cloud(xyz, voxel_size, plane_thre, init_num):
    """orc_btc_plane_cloud (BTC.cpp:279-334) -> (m, 10): center 3 | normal 3 | min_eigen_value | radius | d | points_size."""
    a = np.ascontiguousarray(xyz, dtype=np.float32); cap = max(a.shape[0], 1)
    out = np.zeros((cap, 10)); n = C.c_int64(0)
    lib().orc_btc_plane_cloud(_p(a), C.c_int64(a.shape[0]), C.c_double(voxel_size), C.c_double(plane_thre), C.c_int(init_num), _p(out), C.c_int64(cap), C.byref(n))
    return out[:n.value].copy()


def keyframe_assemble(clouds, poses, voxel_size):
    """orc_keyframe_assemble: voxelslam.cpp:345-378 (scans into the frame of the last one, down_sampling_voxel(voxel_size / 10))."""
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses)
    out = np.zeros(max(int(off[-1]), 1), dtype=point48_dtype); n = C.c_int64(0)
    lib().orc_keyframe_assemble(_p(xyz), _p(off), C.c_int(len(clouds)), _p(ps), C.c_double(voxel_size), _p(out), C.byref(n))
    return out[:n.value].copy()


def icp_normal(src_xyz, src_nrm, tar_xyz, tar_nrm, pose, icp_eigval):
    """orc_icp_normal: the oracle's icp_normal (loop_refine.hpp:47-145), brute-force nearest neighbour."""
    sx = np.ascontiguousarray(src_xyz, dtype=np.float32); sn = np.ascontiguousarray(src_nrm, dtype=np.float32)
    tx = np.ascontiguousarray(tar_xyz, dtype=np.float32); tn = np.ascontiguousarray(tar_nrm, dtype=np.float32)
    ps = as_poses((np.asarray(pose[0], dtype=np.float64)[None], np.asarray(pose[1], dtype=np.float64)[None])).copy()
    it = C.c_int(0); conv = C.c_int(0)
    mn = np.zeros(20, dtype=np.int32); rs = np.zeros(20); eig = np.zeros(3)
    f = lib().orc_icp_normal; f.restype = C.c_int
    ok = f(_p(sx), _p(sn), C.c_int(sx.shape[0]), _p(tx), _p(tn), C.c_int(tx.shape[0]), _p(ps), C.c_double(icp_eigval),
           C.byref(it), _p(mn), _p(rs), _p(eig), C.byref(conv))
    return dict(ok=bool(ok), R=ps["R"][0].reshape(3, 3).T.copy(), p=ps["p"][0].copy(), iters=it.value, match_nums=mn[:it.value].tolist(),
                resis=rs[:it.value].copy(), eig=eig, is_converge=conv.value)


def deskew(pts4, poses22, xc, lidar_to_imu):
    """orc_deskew: pts4 (n,4) float32 x y z stamp; poses22 (m,22) float64 -> (points (n,4), assign (n,))."""
    out = np.ascontiguousarray(pts4, dtype=np.float32).copy()
    ps = np.ascontiguousarray(poses22, dtype=np.float64)
    a = np.full(out.shape[0], -2, dtype=np.int32)
    lib().orc_deskew(_p(out), C.c_int64(out.shape[0]), _p(ps), C.c_int(ps.shape[0]), _p(as_poses(xc)), _p(as_poses(lidar_to_imu)), _p(a))
    return out, a


def var_init(xyz, dept_err, beam_err):
    """calcBodyVar + var_init (voxelslam.hpp:180-234), identity extrinsic -> (n,12): pnt 3 + var 9 (col-major)."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    out = np.zeros((xyz.shape[0], 12))
    lib().orc_var_init(_p(xyz), C.c_int(xyz.shape[0]), C.c_double(dept_err), C.c_double(beam_err), _p(out))
    return out


def var_init_ext(xyz, pose, dept_err, beam_err):
    xyz = np.ascontiguousarray(xyz, dtype=np.float32)
    out = np.zeros((xyz.shape[0], 12))
    lib().orc_var_init_ext(_p(xyz), C.c_int(xyz.shape[0]), _p(as_poses(pose)), C.c_double(dept_err), C.c_double(beam_err), _p(out))
    return out


def pvec_update(pv, pose, cov):
    pv = np.ascontiguousarray(pv, dtype=np.float64)
    out = np.zeros_like(pv); pw = np.zeros((pv.shape[0], 3))
    lib().orc_pvec_update(_p(pv), C.c_int(pv.shape[0]), _p(as_poses(pose)), _p(np.ascontiguousarray(cov.T)), _p(out), _p(pw))
    return out, pw


def default_cov():
    """IMUST::setZero covariance (tools.hpp:188-197)."""
    cov = np.eye(15) * 0.0001
    cov[9:, 9:] = np.eye(6) * 0.00001
    return cov


node_state_dtype = np.dtype([("root_key", "<i8", (3,)), ("layer", "<i4"), ("path", "<i4"), ("octo_state", "<i4"), ("isexist", "<i4"), ("has_sw", "<i4"),
                             ("is_plane", "<i4"), ("last_num", "<i4"), ("opt_state", "<i4"), ("in_slide", "<i4"), ("n_window_pts", "<i4"),
                             ("n_fix_pts", "<i4"), ("_pad", "<i4"), ("pcr_add", cluster_dtype), ("pcr_fix", cluster_dtype), ("eig_value", "<f8", (3,)),
                             ("cov_add_trace", "<f8")])


class LocalMap:
    """The sliding-window voxel map session of voxelslam.cpp:1888-2043 (poses given, no LIO/BA)."""

    def __init__(self, win_size=10, voxel_size=1.0, min_eigen_value=0.0025, thre=(1, 1, 1, 1), max_layer=2, thread_num=5, dept_err=0.01, beam_err=0.01):
        t = np.ascontiguousarray(thre, dtype=np.float64)
        lib().orc_local_create.restype = C.c_void_p
        self.h = C.c_void_p(lib().orc_local_create(C.c_int(win_size), C.c_double(voxel_size), C.c_double(min_eigen_value), _p(t), C.c_int(max_layer),
                                                   C.c_int(thread_num), C.c_double(dept_err), C.c_double(beam_err)))
        self.voxel_size = voxel_size

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_local_free(self.h); self.h = None

    def push(self, xyz, pose, cov=None):
        xyz = np.ascontiguousarray(xyz, dtype=np.float32)
        ps = as_poses(pose)
        cv = None if cov is None else np.ascontiguousarray(np.asarray(cov, dtype=np.float64).T)
        lib().orc_local_push(self.h, _p(xyz), C.c_int(xyz.shape[0]), _p(ps), _p(cv), None)

    def export(self):
        nr = C.c_int64(0)
        lib().orc_local_export_count.restype = C.c_int64
        nn = lib().orc_local_export_count(self.h, C.byref(nr))
        keys = np.zeros((nr.value, 3), dtype=np.int64); rn = np.zeros(nr.value, dtype=np.int32); nodes = np.zeros(nn, dtype=map_node_dtype)
        lib().orc_local_export(self.h, _p(keys), _p(rn), _p(nodes))
        return keys, rn, nodes

    # ---- step-wise session (voxelslam.cpp:1905-2040 without the optimizer) ----
    def insert(self, xyz, pose, cov=None):
        """window bookkeeping + cut_voxel_multi; returns (pv (n,12): local pnt + world var, pwld (n,3))."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32); n = xyz.shape[0]
        ps = as_poses(pose)
        cv = None if cov is None else np.ascontiguousarray(np.asarray(cov, dtype=np.float64).T)
        pv = np.zeros((n, 12)); pw = np.zeros((n, 3))
        lib().orc_local_insert(self.h, _p(xyz), C.c_int(n), _p(ps), _p(cv), _p(pv), _p(pw))
        return pv, pw

    def recut(self):
        lib().orc_local_recut(self.h)

    def margi(self):
        return int(lib().orc_local_margi(self.h))

    @property
    def win_count(self):
        return int(lib().orc_local_win_count(self.h))

    def get_poses(self):
        ps = np.zeros(self.win_count, dtype=pose_dtype)
        lib().orc_local_get_poses(self.h, _p(ps))
        return ps

    def set_poses(self, poses):
        ps = as_poses(poses)
        assert ps.shape[0] == self.win_count
        lib().orc_local_set_poses(self.h, _p(ps))

    def factor(self, win_size):
        lib().orc_local_factor.restype = C.c_void_p
        return Factor(win_size, handle=lib().orc_local_factor(self.h), borrowed=True)

    def insert_fix(self, pv, jour=0.0):
        pv = np.ascontiguousarray(pv, dtype=np.float64)
        lib().orc_local_insert_fix(self.h, _p(pv), C.c_int(pv.shape[0]), C.c_double(jour))

    def export_states(self):
        nr = C.c_int64(0)
        lib().orc_local_export_count.restype = C.c_int64
        nn = lib().orc_local_export_count(self.h, C.byref(nr))
        keys = np.zeros((nr.value, 3), dtype=np.int64); rn = np.zeros(nr.value, dtype=np.int32); nodes = np.zeros(nn, dtype=map_node_dtype)
        lib().orc_local_export(self.h, _p(keys), _p(rn), _p(nodes))
        st = np.zeros(nn, dtype=node_state_dtype)
        lib().orc_local_export_states(self.h, _p(st))
        return st, nodes

    def lio_pass(self, pv, pose, cov, leaf_cache=None):
        pv = np.ascontiguousarray(pv, dtype=np.float64); n = pv.shape[0]
        ps = as_poses(pose); cv = np.ascontiguousarray(np.asarray(cov, dtype=np.float64).T)
        lc = np.full(n, -1, dtype=np.int32) if leaf_cache is None else np.ascontiguousarray(leaf_cache, dtype=np.int32).copy()
        HTH = np.zeros((6, 6)); HTz = np.zeros(6); nnt = np.zeros((3, 3)); mn = C.c_int(0); fl = np.zeros(n, dtype=np.int32)
        lib().orc_lio_pass(self.h, _p(pv), C.c_int(n), _p(ps), _p(cv), _p(lc), _p(HTH), _p(HTz), _p(nnt), C.byref(mn), _p(fl))
        return dict(HTH=HTH.T.copy(), HTz=HTz, nnt=nnt.T.copy(), match_num=mn.value, flags=fl, leaf_cache=lc)

    def lio_state_estimation(self, pv, R, p, cov, v=None, bg=None, ba=None):
        pv = np.ascontiguousarray(pv, dtype=np.float64); n = pv.shape[0]
        st = np.zeros(21); st[:9] = np.asarray(R).T.reshape(9); st[9:12] = p
        if v is not None: st[12:15] = v
        if bg is not None: st[15:18] = bg
        if ba is not None: st[18:21] = ba
        cv = np.ascontiguousarray(np.asarray(cov, dtype=np.float64).T).copy()
        it = C.c_int(0); mn = np.zeros(4, dtype=np.int32)
        ok = lib().orc_lio_state_estimation(self.h, _p(pv), C.c_int(n), _p(st), _p(cv), C.byref(it), _p(mn))
        return dict(ok=bool(ok), R=st[:9].reshape(3, 3).T.copy(), p=st[9:12].copy(), v=st[12:15].copy(), bg=st[15:18].copy(), ba=st[18:21].copy(),
                    cov=cv.reshape(15, 15).T.copy(), iters=it.value, match_nums=mn[:it.value].tolist())
