"""voxel_slam_b200 — ctypes host binding over libvoxelba.so (the C ABI in include/voxelba.h).

This is plumbing for tests/ and bench.py: every compute call goes through the C ABI into the
hand-written sm_100a kernels under csrc/.  There is no Python or CPU implementation of the hot
path here; if the shared library is missing or no CUDA device is present the calls raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvoxelba.so")
_lib = None

# numpy mirrors of the POD structs in include/voxelba.h
cluster_dtype = np.dtype([("P", "<f8", (9,)), ("v", "<f8", (3,)), ("N", "<i4"), ("_pad", "<i4")], align=True)
pose_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,))], align=True)
assert cluster_dtype.itemsize == 104 and pose_dtype.itemsize == 96

gba_params_dtype = np.dtype([("voxel_size", "<f8"), ("min_eigen_value", "<f8"), ("eigen_value_array", "<f8", (4,)), ("max_layer", "<i4"), ("_pad", "<i4")])
point48_dtype = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("_p0", "<f4"), ("normal_x", "<f4"), ("normal_y", "<f4"),
                          ("normal_z", "<f4"), ("_p1", "<f4"), ("intensity", "<f4"), ("curvature", "<f4"), ("_p2", "<f4"), ("_p3", "<f4")])
assert gba_params_dtype.itemsize == 56 and point48_dtype.itemsize == 48

VX_OK = 0
ERR_NAMES = {-1: "VX_ERR_CUDA", -2: "VX_ERR_ARG", -3: "VX_ERR_TOO_FEW", -4: "VX_ERR_OPT_STATE", -5: "VX_ERR_NOMEM", -6: "VX_ERR_NCCL"}


class VxError(RuntimeError):
    def __init__(self, code, text):
        super().__init__("%s (%d): %s" % (ERR_NAMES.get(code, "VX_ERR"), code, text))
        self.code = code


def load_library():
    """dlopen libvoxelba.so; fails loudly when the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("libvoxelba.so is missing at %s — run `python __graft_entry__.py` (nvcc, sm_100a). "
                          "There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    lib.vx_last_error.restype = C.c_char_p
    lib.vx_ctx_stream.restype = C.c_void_p
    lib.vx_ctx_stream.argtypes = [C.c_void_p]
    lib.vx_ctx_launch_count.restype = C.c_int64
    lib.vx_ctx_launch_count.argtypes = [C.c_void_p]
    for name in ("vx_factor_num_voxels", "vx_factor_num_voxel_frames"):
        getattr(lib, name).restype = C.c_int64
        getattr(lib, name).argtypes = [C.c_void_p]
    _lib = lib
    return lib


def _chk(code):
    if code != VX_OK:
        raise VxError(code, load_library().vx_last_error().decode("utf-8", "replace"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def as_poses(poses):
    """(F,3,3) rotations + (F,3) translations, or a pose_dtype array -> contiguous pose_dtype array."""
    if isinstance(poses, np.ndarray) and poses.dtype == pose_dtype:
        return np.ascontiguousarray(poses)
    R, p = poses
    R = np.asarray(R, dtype=np.float64)
    out = np.zeros(R.shape[0], dtype=pose_dtype)
    out["R"] = np.transpose(R, (0, 2, 1)).reshape(-1, 9)  # column-major
    out["p"] = np.asarray(p, dtype=np.float64)
    return out


class Context:
    def __init__(self, device=0):
        self.lib = load_library()
        h = C.c_void_p()
        _chk(self.lib.vx_ctx_create(C.c_int(device), C.byref(h)))
        self.h = h
        self.device = device
        self.launch_count_last = 0

    @property
    def stream(self):
        return self.lib.vx_ctx_stream(self.h)

    @property
    def launch_count(self):
        return int(self.lib.vx_ctx_launch_count(self.h))

    def sync(self):
        _chk(self.lib.vx_ctx_sync(self.h))

    def fp64_peak(self, mode=0):
        """Measured FP64 TFLOP/s of this device: mode 0 DFMA (CUDA cores), mode 1 DMMA (tensor pipe)."""
        t = C.c_double(0)
        _chk(self.lib.vx_bench_fp64(self.h, C.c_int(mode), C.byref(t)))
        return t.value

    def fp64_latency(self, mma=0, ilp=1):
        """Cycles per instruction of one warp running `ilp` independent dependent chains of DFMA / DMMA."""
        t = C.c_double(0)
        _chk(self.lib.vx_bench_fp64_latency(self.h, C.c_int(mma), C.c_int(ilp), C.byref(t)))
        return t.value

    def profile(self, enable=True):
        """Start (and clear) / stop per-kernel CUDA-event timing of this context's launches."""
        _chk(self.lib.vx_ctx_profile(self.h, C.c_int(int(enable))))

    def profile_read(self):
        """-> {kernel name: (total ms, launches)} since profile(True)."""
        names = C.create_string_buffer(8192)
        ms = np.zeros(128); cnt = np.zeros(128, dtype=np.int64); n = C.c_int(0)
        _chk(self.lib.vx_ctx_profile_read(self.h, names, C.c_size_t(8192), _p(ms), _p(cnt), C.c_int(128), C.byref(n)))
        nm = names.value.decode().split("\n")
        return {nm[i]: (float(ms[i]), int(cnt[i])) for i in range(n.value)}

    def close(self):
        if self.h:
            self.launch_count_last = self.launch_count
            self.lib.vx_ctx_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


def keys(ctx, xyz, voxel_size):
    """A1: (n,3) float64 or float32 coordinates -> (n,3) int64 VOXEL_LOC (bit-exact reference rule)."""
    xyz = np.ascontiguousarray(xyz)
    assert xyz.dtype in (np.float32, np.float64)
    out = np.zeros(xyz.shape, dtype=np.int64)
    _chk(ctx.lib.vx_keys(ctx.h, _p(xyz), C.c_int64(xyz.size // 3), C.c_double(voxel_size), C.c_int(int(xyz.dtype == np.float32)), _p(out)))
    return out


KEYS_STATIC_MAX_POINTS = 262144   # VX_KEYS_STATIC_MAX_POINTS, include/voxelba.h


def keys_static(ctx, xyz, voxel_size):
    """vx_keys_static: vx_keys on the library's fixed device scratch (no allocation); at most KEYS_STATIC_MAX_POINTS points."""
    xyz = np.ascontiguousarray(xyz)
    assert xyz.dtype in (np.float32, np.float64)
    out = np.zeros(xyz.shape, dtype=np.int64)
    _chk(ctx.lib.vx_keys_static(ctx.h, _p(xyz), C.c_int64(xyz.size // 3), C.c_double(voxel_size), C.c_int(int(xyz.dtype == np.float32)), _p(out)))
    return out


def downsample_voxel(ctx, pts, voxel_size):
    """A2: down_sampling_voxel (tools.hpp:201-238); returns the cells in ascending packed-key order."""
    pts = np.ascontiguousarray(pts, dtype=point48_dtype)
    out = np.zeros(pts.shape[0], dtype=point48_dtype)
    n = C.c_int64(0)
    _chk(ctx.lib.vx_downsample_voxel(ctx.h, _p(pts), C.c_int64(pts.shape[0]), C.c_double(voxel_size), _p(out), C.byref(n)))
    return out[:n.value]


def downsample_close(ctx, pts, voxel_size):
    """A4: down_sampling_close (tools.hpp:240-302); cells in ascending packed-key order."""
    pts = np.ascontiguousarray(pts, dtype=point48_dtype)
    out = np.zeros(pts.shape[0], dtype=point48_dtype)
    n = C.c_int64(0)
    _chk(ctx.lib.vx_downsample_close(ctx.h, _p(pts), C.c_int64(pts.shape[0]), C.c_double(voxel_size), _p(out), C.byref(n)))
    return out[:n.value]


def downsample_pvec(ctx, pv, voxel_size):
    """A3: down_sampling_pvec (voxel_map.hpp:39-81); pv (n,12) pointVar rows; cells in ascending packed-key order."""
    pv = np.ascontiguousarray(pv, dtype=np.float64)
    out = np.zeros(pv.shape[0], dtype=point48_dtype)
    n = C.c_int64(0)
    _chk(ctx.lib.vx_downsample_pvec(ctx.h, _p(pv), C.c_int64(pv.shape[0]), C.c_double(voxel_size), _p(out), C.byref(n)))
    return out[:n.value]


hba_params_dtype = np.dtype([("gba", gba_params_dtype), ("voxel_size", "<f8"), ("min_eigen_value", "<f8"), ("plane_eigen_value_thre", "<f8", (4,))])
edge_dtype = np.dtype([("m1", "<i4"), ("m2", "<i4"), ("id1", "<i4"), ("id2", "<i4"), ("rot", "<f8", (9,)), ("tra", "<f8", (3,)), ("cov", "<f8", (6,))])
assert hba_params_dtype.itemsize == 104 and edge_dtype.itemsize == 160


def hba_add_edge(ctx, clouds, poses, W, n_windows, local, gba, max_layer, max_iter, thread_num=2, frame_cloud=None, ids=None, mps=None):
    """VOXEL_SLAM::HBA_add_edge (voxelslam.cpp:2822-3015) for n_windows independent windows of W frames.
    local = (voxel_size, min_eigen_value, thre4 as stored = 1/yaml); gba = (gba_voxel_size, gba_min_eigen_value, eig4)."""
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses).copy()
    prm = np.zeros(1, dtype=hba_params_dtype)
    prm["gba"]["voxel_size"] = gba[0]; prm["gba"]["min_eigen_value"] = gba[1]; prm["gba"]["eigen_value_array"] = gba[2]; prm["gba"]["max_layer"] = max_layer
    prm["voxel_size"] = local[0]; prm["min_eigen_value"] = local[1]; prm["plane_eigen_value_thre"] = local[2]
    cap = W * (W - 1) // 2
    jl = 6 * W
    edges = np.zeros(n_windows * cap, dtype=edge_dtype); ne = np.zeros(n_windows, dtype=np.int64)
    hess = np.zeros((n_windows, jl, jl)); resis = np.zeros((n_windows, 2)); nv = np.zeros(n_windows, dtype=np.int32); st = np.zeros(n_windows, dtype=np.int32)
    fc = None if frame_cloud is None else np.ascontiguousarray(frame_cloud, dtype=np.int32)
    ids_ = None if ids is None else np.ascontiguousarray(ids, dtype=np.int32)
    mps_ = None if mps is None else np.ascontiguousarray(mps, dtype=np.int32)
    _chk(ctx.lib.vx_hba_add_edge(ctx.h, xyz.ctypes.data_as(C.c_void_p), _p(off), C.c_int(len(clouds)), _p(fc), _p(ids_), _p(mps_), _p(ps), C.c_int(W),
                                 C.c_int(n_windows), _p(prm), C.c_int(max_iter), C.c_int(thread_num), _p(edges), _p(ne), _p(hess), _p(resis), _p(nv), _p(st)))
    return dict(poses=ps, edges=[edges[b * cap:b * cap + ne[b]] for b in range(n_windows)], hess=np.transpose(hess, (0, 2, 1)), resis=resis,
                n_voxels=nv, status=st)


def hba_merge_clouds(ctx, clouds, poses, W, n_windows, voxel_size, frame_cloud=None, mps=None):
    """The merged submap cloud of HBA_add_edge (voxelslam.cpp:2954-2985), one per window -> list of point48 arrays."""
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses)
    fc = None if frame_cloud is None else np.ascontiguousarray(frame_cloud, dtype=np.int32)
    mps_ = None if mps is None else np.ascontiguousarray(mps, dtype=np.int32)
    sizes = np.diff(off)
    idx = np.arange(W * n_windows) if fc is None else fc
    cap = int(sizes[idx].sum())
    out = np.zeros(max(cap, 1), dtype=point48_dtype); wo = np.zeros(n_windows + 1, dtype=np.int64)
    _chk(ctx.lib.vx_hba_merge_clouds(ctx.h, _p(xyz), _p(off), C.c_int(len(clouds)), _p(fc), _p(mps_), _p(ps), C.c_int(W), C.c_int(n_windows),
                                     C.c_double(voxel_size), _p(out), C.c_int64(cap), _p(wo)))
    return [out[wo[b]:wo[b + 1]].copy() for b in range(n_windows)]



# ---- multi-GPU (vx_comm_*) and the two-level hierarchical GBA ----------------------------------------------------------------
hba_global_params_dtype = np.dtype([("hba", hba_params_dtype), ("wdsize", "<i4"), ("mgsize", "<i4"), ("bottom_max_iter", "<i4"),
                                    ("bottom_thread_num", "<i4"), ("total_max_iter", "<i4"), ("top_thread_num", "<i4"), ("xyz_on_device", "<i4"),
                                    ("keep_factors", "<i4")])
assert hba_global_params_dtype.itemsize == 104 + 32


def comm_unique_id():
    """128-byte NCCL id (rank 0 creates it, the other ranks receive it through the host program's own channel)."""
    buf = np.zeros(128, dtype=np.uint8)
    _chk(load_library().vx_comm_unique_id(_p(buf)))
    return buf


class Comm:
    """vx_comm: NCCL communicator of the library, one per (process, GPU)."""

    def __init__(self, ctx, uid, rank, world):
        self.ctx = ctx
        h = C.c_void_p()
        uid = None if uid is None else np.ascontiguousarray(uid, dtype=np.uint8)
        _chk(ctx.lib.vx_comm_create(ctx.h, _p(uid), C.c_int(rank), C.c_int(world), C.byref(h)))
        self.h = h
        self.rank, self.world = rank, world

    @classmethod
    def from_torch(cls, ctx, group=None):
        """Bootstrap over an initialised torch.distributed process group (any backend): rank 0's id is broadcast."""
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        obj = [comm_unique_id().tobytes() if rank == 0 else None]
        if world > 1:
            dist.broadcast_object_list(obj, src=0, group=group)
        return cls(ctx, np.frombuffer(obj[0], dtype=np.uint8), rank, world)

    def stats(self):
        a, b, c = C.c_int64(0), C.c_int64(0), C.c_int64(0)
        _chk(self.ctx.lib.vx_comm_stats(self.h, C.byref(a), C.byref(b), C.byref(c)))
        return {"bytes_allreduced": a.value, "bytes_exchanged": b.value, "collectives": c.value}

    def free(self):
        if self.h:
            self.ctx.lib.vx_comm_free(self.h)
            self.h = None


def hba_schedule(session_sizes, wdsize=10, mgsize=5):
    """vx_hba_schedule: the bottom windows of thd_globalmapping -> (session, start, size) int32 arrays."""
    ss = np.ascontiguousarray(session_sizes, dtype=np.int32)
    n = C.c_int32(0)
    lib = load_library()
    _chk(lib.vx_hba_schedule(_p(ss), C.c_int(ss.size), C.c_int(wdsize), C.c_int(mgsize), None, None, None, C.c_int(0), C.byref(n)))
    a = np.zeros(n.value, dtype=np.int32); b = np.zeros(n.value, dtype=np.int32); c = np.zeros(n.value, dtype=np.int32)
    _chk(lib.vx_hba_schedule(_p(ss), C.c_int(ss.size), C.c_int(wdsize), C.c_int(mgsize), _p(a), _p(b), _p(c), C.c_int(n.value), C.byref(n)))
    return a, b, c


def hba_window_range(win_size, wdsize, rank, world):
    """vx_hba_window_range: the windows [lo, hi) of the schedule that rank `rank` of `world` optimises in vx_hba_global."""
    ws = np.ascontiguousarray(win_size, dtype=np.int32)
    lo, hi = C.c_int32(0), C.c_int32(0)
    _chk(load_library().vx_hba_window_range(_p(ws), C.c_int(ws.size), C.c_int(wdsize), C.c_int(rank), C.c_int(world), C.byref(lo), C.byref(hi)))
    return lo.value, hi.value


def hba_global_params(local, gba, max_layer, total_max_iter=1, wdsize=10, mgsize=5, bottom_max_iter=1, bottom_thread_num=2, top_thread_num=5):
    prm = np.zeros(1, dtype=hba_global_params_dtype)
    h = prm["hba"]
    h["gba"]["voxel_size"] = gba[0]; h["gba"]["min_eigen_value"] = gba[1]; h["gba"]["eigen_value_array"] = gba[2]; h["gba"]["max_layer"] = max_layer
    h["voxel_size"] = local[0]; h["min_eigen_value"] = local[1]; h["plane_eigen_value_thre"] = local[2]
    prm["wdsize"] = wdsize; prm["mgsize"] = mgsize; prm["bottom_max_iter"] = bottom_max_iter; prm["bottom_thread_num"] = bottom_thread_num
    prm["total_max_iter"] = total_max_iter; prm["top_thread_num"] = top_thread_num
    return prm


class HbaGlobalJob:
    """Host buffers of one vx_hba_global call, prepared once (bench.py times run() only)."""

    def __init__(self, xyz, off, kf_session, kf_x0, prm, kf_x=None, kf_id=None, cnct_map=None, xyz_device_ptr=None):
        """xyz: host float32 (n,3) array (pinned or pageable), or None with xyz_device_ptr = device address of all clouds."""
        prm = prm.copy()
        if xyz_device_ptr is not None:
            self.xyz = None; self.xyz_arg = C.c_void_p(int(xyz_device_ptr)); prm["xyz_on_device"] = 1
        else:
            self.xyz = xyz if (isinstance(xyz, np.ndarray) and xyz.dtype == np.float32 and xyz.flags.c_contiguous) else np.ascontiguousarray(xyz, dtype=np.float32)
            self.xyz_arg = _p(self.xyz)
        self.off = np.ascontiguousarray(off, dtype=np.int64)
        self.n_kf = self.off.size - 1
        self.sess = np.ascontiguousarray(kf_session, dtype=np.int32)
        self.x0 = as_poses(kf_x0)
        self.x = None if kf_x is None else as_poses(kf_x)
        self.ids = None if kf_id is None else np.ascontiguousarray(kf_id, dtype=np.int32)
        self.cnct = None if cnct_map is None else np.ascontiguousarray(cnct_map, dtype=np.int32)
        self.prm = prm
        sizes = np.bincount(self.sess)
        ws, wst, wsz = hba_schedule(sizes, int(prm["wdsize"][0]), int(prm["mgsize"][0]))
        self.n_windows = ws.size
        wd = int(prm["wdsize"][0])
        self.cap1 = self.n_windows * wd * (wd - 1) // 2
        self.cap2 = self.n_windows * (self.n_windows - 1) // 2
        self.e1 = np.zeros(max(self.cap1, 1), dtype=edge_dtype); self.e2 = np.zeros(max(self.cap2, 1), dtype=edge_dtype)
        self.tp = np.zeros(self.n_windows, dtype=pose_dtype); self.tk = np.zeros(self.n_windows, dtype=np.int32)
        self.stats = np.zeros(16)

    def run(self, ctx, comm=None):
        n1, n2, nt = C.c_int64(0), C.c_int64(0), C.c_int32(0)
        _chk(ctx.lib.vx_hba_global(ctx.h, comm.h if comm is not None else None, self.xyz_arg, _p(self.off), C.c_int(self.n_kf), _p(self.sess), _p(self.ids),
                                   _p(self.x0), _p(self.x), _p(self.cnct), C.c_int(0 if self.cnct is None else self.cnct.size), _p(self.prm),
                                   _p(self.e1), C.c_int64(self.cap1), C.byref(n1), _p(self.e2), C.c_int64(self.cap2), C.byref(n2), _p(self.tp), _p(self.tk),
                                   C.byref(nt), _p(self.stats)))
        st = self.stats
        return dict(edges1=self.e1[:n1.value], edges2=self.e2[:n2.value], top_poses=self.tp[:nt.value], top_kf=self.tk[:nt.value],
                    stats=dict(windows=int(st[0]), top_frames=int(st[1]), voxel_frame_evals=int(st[2]), ms_upload=st[3], ms_bottom=st[4], ms_merge=st[5],
                               ms_gather=st[6], ms_top=st[7], ms_total=st[8], top_points=int(st[9]), vf_evals_bottom=int(st[10]), vf_evals_top=int(st[11]), vf_res_bottom=int(st[12]),
                               vf_res_top=int(st[13]), top_hess_passes=int(st[14]), top_res_passes=int(st[15])))


def hba_global(ctx, clouds, kf_session, kf_x0, local, gba, max_layer, total_max_iter=1, kf_x=None, kf_id=None, cnct_map=None, comm=None, **kw):
    """vx_hba_global: bottom windows -> submaps -> top-level HBA_add_edge (voxelslam.cpp:3018-3141)."""
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    job = HbaGlobalJob(xyz, off, kf_session, kf_x0, hba_global_params(local, gba, max_layer, total_max_iter, **kw), kf_x, kf_id, cnct_map)
    return job.run(ctx, comm)

map_node_dtype = np.dtype([("center", "<f8", (3,)), ("quater_length", "<f4"), ("octo_state", "<i4"), ("child", "<i4", (8,)), ("is_plane", "<i4"),
                           ("radius", "<f4"), ("p_center", "<f8", (3,)), ("p_normal", "<f8", (3,)), ("plane_var", "<f8", (36,))])
lio_state_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,)), ("v", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)), ("cov", "<f8", (225,))])
assert map_node_dtype.itemsize == 408 and lio_state_dtype.itemsize == 8 * (21 + 225)


def ldlt_solve(ctx, A, b):
    """vx_ldlt_solve: Eigen-style LDL^T solve of the system defined by the LOWER triangle of A, factorised on the device."""
    A = np.asarray(A, dtype=np.float64); n = A.shape[0]
    Ac = np.ascontiguousarray(A.T); bb = np.ascontiguousarray(b, dtype=np.float64); x = np.zeros(n)
    _chk(ctx.lib.vx_ldlt_solve(ctx.h, _p(Ac), C.c_int(n), _p(bb), _p(x)))
    return x


# ---- stored sessions: host-side formats, no device needed ------------------------------------
scan_state_dtype = np.dtype([("t", "<f8"), ("p", "<f8", (3,)), ("R", "<f8", (9,)), ("v", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)),
                             ("g", "<f8", (3,)), ("v6", "<f8", (6,))])
pgo_edge_dtype = np.dtype([("m1", "<i4"), ("m2", "<i4"), ("id1", "<i4"), ("id2", "<i4"), ("rot", "<f8", (9,)), ("tra", "<f8", (3,)), ("cov", "<f8", (6,))])
assert scan_state_dtype.itemsize == 8 * 31 and pgo_edge_dtype.itemsize == 16 + 8 * 18


def _chk_host(lib, rc):
    if rc != 0:
        raise VxError(rc, lib.vx_last_error().decode())


def write_lidarstate(path, states):
    lib = load_library(); st = np.ascontiguousarray(states, dtype=scan_state_dtype)
    _chk_host(lib, lib.vx_session_write_lidarstate(path.encode(), _p(st), C.c_int64(st.shape[0])))


def read_lidarstate(path, cap=1 << 20):
    lib = load_library(); out = np.zeros(cap, dtype=scan_state_dtype); n = C.c_int64(0)
    _chk_host(lib, lib.vx_session_read_lidarstate(path.encode(), _p(out), C.c_int64(cap), C.byref(n)))
    return out[:n.value].copy()


def write_pcd_xyzi(path, xyzi):
    lib = load_library(); a = np.ascontiguousarray(xyzi, dtype=np.float32).reshape(-1, 4)
    _chk_host(lib, lib.vx_session_write_pcd_xyzi(path.encode(), _p(a), C.c_int64(a.shape[0])))


def read_pcd_xyzi(path, cap=1 << 22):
    lib = load_library(); out = np.zeros((cap, 4), dtype=np.float32); n = C.c_int64(0)
    _chk_host(lib, lib.vx_session_read_pcd_xyzi(path.encode(), _p(out), C.c_int64(cap), C.byref(n)))
    return out[:n.value].copy()


def _names(fnames):
    arr = (C.c_char_p * max(len(fnames), 1))(*[f.encode() for f in fnames])
    return arr


def write_edges(path, fnames, edges, absent=None):
    lib = load_library(); e = np.ascontiguousarray(edges, dtype=pgo_edge_dtype)
    _chk_host(lib, lib.vx_session_write_edges(path.encode(), _names(fnames), C.c_int(len(fnames)), _p(e), C.c_int64(e.shape[0]),
                                              None if absent is None else absent.encode()))


def read_edges(path, fnames, bagname, cap=1 << 16):
    lib = load_library(); out = np.zeros(cap, dtype=pgo_edge_dtype); n = C.c_int64(0)
    buf = C.create_string_buffer(1 << 20); blen = C.c_int64(0)
    _chk_host(lib, lib.vx_session_read_edges(path.encode(), _names(fnames), C.c_int(len(fnames)), None if bagname is None else bagname.encode(),
                                             _p(out), C.c_int64(cap), C.byref(n), buf, C.c_int64(len(buf)), C.byref(blen)))
    return out[:n.value].copy(), buf.value.decode()


def keyframe_assemble(ctx, clouds, poses, voxel_size):
    """voxelslam.cpp:345-378: the scans of one keyframe into the frame of the last scan, down_sampling_voxel(voxel_size / 10)."""
    xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
    off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
    ps = as_poses(poses); cap = int(off[-1])
    out = np.zeros(max(cap, 1), dtype=point48_dtype); n = C.c_int64(0)
    _chk(ctx.lib.vx_keyframe_assemble(ctx.h, _p(xyz), _p(off), C.c_int(len(clouds)), _p(ps), C.c_double(voxel_size), _p(out), C.c_int64(cap), C.byref(n)))
    return out[:n.value].copy()


def btc_plane_cloud(ctx, xyz, voxel_size=1.0, plane_thre=float(np.float32(0.01)), init_num=10):
    """STDescManager::init_voxel_map + get_plane (BTC.cpp:279-334) -> (planes as point48, aux (m, 10))."""
    a = np.ascontiguousarray(xyz, dtype=np.float32); cap = max(a.shape[0], 1)
    out = np.zeros(cap, dtype=point48_dtype); aux = np.zeros((cap, 10)); n = C.c_int64(0)
    _chk(ctx.lib.vx_btc_plane_cloud(ctx.h, _p(a), C.c_int64(a.shape[0]), C.c_double(voxel_size), C.c_double(plane_thre), C.c_int(init_num), _p(out), _p(aux),
                                    C.c_int64(cap), C.byref(n)))
    return out[:n.value].copy(), aux[:n.value].copy()


def icp_normal(ctx, src_xyz, src_nrm, tar_xyz, tar_nrm, pose, icp_eigval):
    """icp_normal (loop_refine.hpp:47-145).  pose = (R (3,3), p (3,)); returns dict(ok, R, p, iters, match_nums, resis, eig, is_converge)."""
    sx = np.ascontiguousarray(src_xyz, dtype=np.float32); sn = np.ascontiguousarray(src_nrm, dtype=np.float32)
    tx = np.ascontiguousarray(tar_xyz, dtype=np.float32); tn = np.ascontiguousarray(tar_nrm, dtype=np.float32)
    ps = as_poses((np.asarray(pose[0], dtype=np.float64)[None], np.asarray(pose[1], dtype=np.float64)[None])).copy()
    ok = C.c_int32(0); it = C.c_int32(0); conv = C.c_int32(0)
    mn = np.zeros(20, dtype=np.int32); rs = np.zeros(20); eig = np.zeros(3)
    _chk(ctx.lib.vx_icp_normal(ctx.h, _p(sx), _p(sn), C.c_int64(sx.shape[0]), _p(tx), _p(tn), C.c_int64(tx.shape[0]), _p(ps), C.c_double(icp_eigval),
                               C.byref(ok), C.byref(it), _p(mn), _p(rs), _p(eig), C.byref(conv)))
    return dict(ok=bool(ok.value), R=ps["R"][0].reshape(3, 3).T.copy(), p=ps["p"][0].copy(), iters=it.value, match_nums=mn[:it.value].tolist(),
                resis=rs[:it.value].copy(), eig=eig, is_converge=conv.value)


imu_pose_dtype = np.dtype([("t", "<f8"), ("R", "<f8", (9,)), ("p", "<f8", (3,)), ("v", "<f8", (3,)), ("angvel", "<f8", (3,)), ("acc", "<f8", (3,))])
assert imu_pose_dtype.itemsize == 22 * 8


def deskew(ctx, pts, poses, xc, lidar_to_imu, want_assign=True):
    """vx_deskew: IMUEKF::motion_blur's per-point loop (ekf_imu.hpp:138-163).  pts: point48 array (x, y, z, curvature = stamp);
    poses: imu_pose_dtype array (ascending t); xc / lidar_to_imu: (R (1,3,3), p (1,3)) or pose_dtype.  -> (points, assign)."""
    out = np.ascontiguousarray(pts, dtype=point48_dtype).copy()
    ps = np.ascontiguousarray(poses, dtype=imu_pose_dtype)
    a = np.full(out.shape[0], -2, dtype=np.int32) if want_assign else None
    _chk(ctx.lib.vx_deskew(ctx.h, _p(out), C.c_int64(out.shape[0]), _p(ps), C.c_int(ps.shape[0]), _p(as_poses(xc)), _p(as_poses(lidar_to_imu)), _p(a)))
    return out, a


def var_init(ctx, xyz, ext, dept_err, beam_err):
    """calcBodyVar + var_init (voxelslam.hpp:180-234) on the device -> (n,12): pnt 3 + var 9 (col-major)."""
    xyz = np.ascontiguousarray(xyz, dtype=np.float32); n = xyz.shape[0]
    out = np.zeros((n, 12))
    _chk(ctx.lib.vx_var_init(ctx.h, _p(xyz), C.c_int64(n), _p(as_poses(ext)), C.c_double(dept_err), C.c_double(beam_err), _p(out)))
    return out


def pvec_update(ctx, pv, pose, cov):
    """pvec_update (voxelslam.hpp:242-265) -> (pv_out (n,12), pwld (n,3)).  cov: 15x15 state covariance."""
    pv = np.ascontiguousarray(pv, dtype=np.float64); n = pv.shape[0]
    cov = np.asarray(cov, dtype=np.float64)
    rv = np.ascontiguousarray(cov[0:3, 0:3].T); tv = np.ascontiguousarray(cov[3:6, 3:6].T)
    out = np.zeros((n, 12)); pw = np.zeros((n, 3))
    _chk(ctx.lib.vx_pvec_update(ctx.h, _p(pv), C.c_int64(n), _p(as_poses(pose)), _p(rv), _p(tv), _p(out), _p(pw)))
    return out, pw


class PlaneMap:
    """Flattened OctoTree voxel map on the device (what OctoTree::match reads, voxel_map.hpp:1649-1721)."""

    def __init__(self, ctx, root_keys, root_node, nodes, voxel_size):
        self.ctx, self.lib = ctx, ctx.lib
        rk = np.ascontiguousarray(root_keys, dtype=np.int64); rn = np.ascontiguousarray(root_node, dtype=np.int32)
        nd = np.ascontiguousarray(nodes, dtype=map_node_dtype)
        h = C.c_void_p()
        _chk(self.lib.vx_planemap_create(ctx.h, _p(rk), _p(rn), C.c_int64(rk.shape[0]), _p(nd), C.c_int64(nd.shape[0]), C.c_double(voxel_size), C.byref(h)))
        self.h = h

    def free(self):
        if self.h and not getattr(self, "view", False):
            self.lib.vx_planemap_free(self.h)
        self.h = None

    def lio_pass(self, pv, pose, cov, leaf_cache=None):
        """One pass of voxelslam.cpp:1005-1050.  pv: (n,12) pnt + var (col-major); cov: 15x15 state covariance."""
        pv = np.ascontiguousarray(pv, dtype=np.float64); n = pv.shape[0]
        ps = as_poses(pose); cov = np.asarray(cov, dtype=np.float64)
        rv = np.ascontiguousarray(cov[0:3, 0:3].T); tv = np.ascontiguousarray(cov[3:6, 3:6].T)
        lc = np.full(n, -1, dtype=np.int32) if leaf_cache is None else np.ascontiguousarray(leaf_cache, dtype=np.int32).copy()
        HTH = np.zeros((6, 6)); HTz = np.zeros(6); nnt = np.zeros((3, 3)); mn = C.c_int32(0); fl = np.zeros(n, dtype=np.int32)
        _chk(self.lib.vx_lio_match_accumulate(self.ctx.h, self.h, _p(pv), C.c_int64(n), _p(ps), _p(rv), _p(tv), _p(lc), _p(HTH), _p(HTz), _p(nnt),
                                              C.byref(mn), _p(fl)))
        return dict(HTH=HTH.T.copy(), HTz=HTz, nnt=nnt.T.copy(), match_num=mn.value, flags=fl, leaf_cache=lc)

    def lio_state_estimation(self, pv, R, p, cov, v=None, bg=None, ba=None, xyz_ext=None):
        """VOXEL_SLAM::lio_state_estimation (voxelslam.cpp:962-1098).  xyz_ext=(ext_pose, dept_err, beam_err): `pv` is the raw
        (n,3) float scan and var_init runs on the device (vx_lio_state_estimation_xyz)."""
        if xyz_ext is None:
            pv = np.ascontiguousarray(pv, dtype=np.float64)
        else:
            pv = np.ascontiguousarray(pv, dtype=np.float32)
        n = pv.shape[0]
        st = np.zeros(1, dtype=lio_state_dtype)
        st["R"] = np.asarray(R, dtype=np.float64).T.reshape(9); st["p"] = p
        if v is not None: st["v"] = v
        if bg is not None: st["bg"] = bg
        if ba is not None: st["ba"] = ba
        st["cov"] = np.asarray(cov, dtype=np.float64).T.reshape(225)
        ok = C.c_int32(0); it = C.c_int32(0); mn = np.zeros(4, dtype=np.int32)
        if xyz_ext is None:
            _chk(self.lib.vx_lio_state_estimation(self.ctx.h, self.h, _p(pv), C.c_int64(n), _p(st), C.byref(ok), C.byref(it), _p(mn)))
        else:
            ext = as_poses(xyz_ext[0])
            _chk(self.lib.vx_lio_state_estimation_xyz(self.ctx.h, self.h, _p(pv), C.c_int64(n), _p(ext), C.c_double(xyz_ext[1]),
                                                      C.c_double(xyz_ext[2]), _p(st), C.byref(ok), C.byref(it), _p(mn)))
        s = st[0]
        return dict(ok=bool(ok.value), R=s["R"].reshape(3, 3).T.copy(), p=s["p"].copy(), v=s["v"].copy(), bg=s["bg"].copy(), ba=s["ba"].copy(),
                    cov=s["cov"].reshape(15, 15).T.copy(), iters=it.value, match_nums=mn[:it.value].tolist())


# ---- A6 / A7 / A9 / A11 / A12: the sliding-window voxel map on the device ------------------------------------------
map_params_dtype = np.dtype([("voxel_size", "<f8"), ("min_eigen_value", "<f8"), ("plane_eigen_value_thre", "<f8", (4,)), ("min_point", "<f8", (4,)),
                             ("max_layer", "<i4"), ("max_points", "<i4"), ("win_size", "<i4"), ("thread_num", "<i4")])
node_state_dtype = np.dtype([("root_key", "<i8", (3,)), ("layer", "<i4"), ("path", "<i4"), ("octo_state", "<i4"), ("isexist", "<i4"), ("has_sw", "<i4"),
                             ("is_plane", "<i4"), ("last_num", "<i4"), ("opt_state", "<i4"), ("in_slide", "<i4"), ("n_window_pts", "<i4"),
                             ("n_fix_pts", "<i4"), ("_pad", "<i4"), ("pcr_add", cluster_dtype), ("pcr_fix", cluster_dtype), ("eig_value", "<f8", (3,)),
                             ("cov_add_trace", "<f8")])


class VoxelMap:
    """vx_map: surf_map + surf_map_slide of voxelslam.cpp (OctoTree / SlideWindow, voxel_map.hpp:1009-1881) in device memory."""

    def __init__(self, ctx, win_size=10, voxel_size=1.0, min_eigen_value=0.0025, thre=(1, 1, 1, 1), max_layer=2, max_points=100, thread_num=5,
                 min_point=(5, 5, 5, 5)):
        self.ctx = ctx; self.lib = ctx.lib; self.W = win_size; self.voxel_size = voxel_size
        pr = np.zeros(1, dtype=map_params_dtype)
        pr["voxel_size"] = voxel_size; pr["min_eigen_value"] = min_eigen_value; pr["plane_eigen_value_thre"] = thre; pr["min_point"] = min_point
        pr["max_layer"] = max_layer; pr["max_points"] = max_points; pr["win_size"] = win_size; pr["thread_num"] = thread_num
        h = C.c_void_p()
        _chk(self.lib.vx_map_create(ctx.h, _p(pr), C.byref(h)))
        self.h = h
        self.lib.vx_map_num_nodes.restype = C.c_int64; self.lib.vx_map_num_nodes.argtypes = [C.c_void_p]
        self.lib.vx_map_num_roots.restype = C.c_int64; self.lib.vx_map_num_roots.argtypes = [C.c_void_p]

    def free(self):
        if self.h:
            _chk(self.lib.vx_map_free(self.h)); self.h = None

    @property
    def num_nodes(self):
        return int(self.lib.vx_map_num_nodes(self.h))

    @property
    def num_roots(self):
        return int(self.lib.vx_map_num_roots(self.h))

    def cut_voxel(self, pv, pwld, ord_):
        """cut_voxel_multi (voxel_map.hpp:1964-2096): pv (n,12) local pnt + world var, pwld (n,3), ord_ = win_count - 1."""
        pv = np.ascontiguousarray(pv, dtype=np.float64); pw = np.ascontiguousarray(pwld, dtype=np.float64)
        _chk(self.lib.vx_map_cut_voxel(self.ctx.h, self.h, _p(pv), _p(pw), C.c_int64(pv.shape[0]), C.c_int(ord_)))

    def cut_voxel_xyz(self, xyz, ext, dept_err, beam_err, pose, cov, ord_):
        """var_init + pvec_update + cut_voxel_multi on the device from the raw (n,3) float scan."""
        xyz = np.ascontiguousarray(xyz, dtype=np.float32); cov = np.asarray(cov, dtype=np.float64)
        rv = np.ascontiguousarray(cov[0:3, 0:3].T); tv = np.ascontiguousarray(cov[3:6, 3:6].T)
        _chk(self.lib.vx_map_cut_voxel_xyz(self.ctx.h, self.h, _p(xyz), C.c_int64(xyz.shape[0]), _p(as_poses(ext)), C.c_double(dept_err), C.c_double(beam_err),
                                           _p(as_poses(pose)), _p(rv), _p(tv), C.c_int(ord_)))

    def cut_voxel_fix(self, pv, jour=0.0):
        pv = np.ascontiguousarray(pv, dtype=np.float64)
        _chk(self.lib.vx_map_cut_voxel_fix(self.ctx.h, self.h, _p(pv), C.c_int64(pv.shape[0]), C.c_double(jour)))

    def recut(self, poses, factor):
        ps = as_poses(poses)
        _chk(self.lib.vx_map_recut(self.ctx.h, self.h, _p(ps), C.c_int(ps.shape[0]), factor.h))

    def margi(self, poses, factor, mgsize=1, jour=0.0):
        ps = as_poses(poses)
        _chk(self.lib.vx_map_margi(self.ctx.h, self.h, _p(ps), C.c_int(ps.shape[0]), C.c_int(mgsize), factor.h, C.c_double(jour)))

    def shift_window(self, mgsize=1):
        _chk(self.lib.vx_map_shift_window(self.h, C.c_int(mgsize)))

    def planemap(self):
        """non-owning PlaneMap view for lio_pass / lio_state_estimation (valid until the map changes)."""
        v = C.c_void_p()
        _chk(self.lib.vx_map_planemap(self.h, C.byref(v)))
        pm = PlaneMap.__new__(PlaneMap)
        pm.ctx = self.ctx; pm.lib = self.lib; pm.h = v; pm.view = True
        return pm

    def export(self):
        n = self.num_nodes
        st = np.zeros(n, dtype=node_state_dtype); nd = np.zeros(n, dtype=map_node_dtype)
        _chk(self.lib.vx_map_export(self.h, _p(st), _p(nd)))
        return st, nd


# ---- B6 / B4 / B5: IMU preintegration factor (host) and the LiDAR-inertial optimizers -------------------------
imu_pre_dtype = np.dtype([("R_delta", "<f8", (9,)), ("p_delta", "<f8", (3,)), ("v_delta", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)),
                          ("R_bg", "<f8", (9,)), ("p_bg", "<f8", (9,)), ("p_ba", "<f8", (9,)), ("v_bg", "<f8", (9,)), ("v_ba", "<f8", (9,)),
                          ("dtime", "<f8"), ("dbg", "<f8", (3,)), ("dba", "<f8", (3,)), ("dbg_buf", "<f8", (3,)), ("dba_buf", "<f8", (3,)),
                          ("cov", "<f8", (225,))])
imu_noise_dtype = np.dtype([("scale_gravity", "<f8"), ("noise_meas", "<f8", (36,)), ("noise_walk", "<f8", (36,))])
imu_sample_dtype = np.dtype([("t", "<f8"), ("gyr", "<f8", (3,)), ("acc", "<f8", (3,))])
imu_state_dtype = np.dtype([("R", "<f8", (9,)), ("p", "<f8", (3,)), ("v", "<f8", (3,)), ("bg", "<f8", (3,)), ("ba", "<f8", (3,)), ("g", "<f8", (3,))])


def imu_noise(cov_gyr, cov_acc, rand_walk_gyr, rand_walk_acc, scale_gravity=1.0):
    """voxelslam.cpp:934-939: diagonal noiseMeas / noiseWalk."""
    nz = np.zeros(1, dtype=imu_noise_dtype)
    nz["scale_gravity"] = scale_gravity
    nz["noise_meas"][0] = np.diag([cov_gyr] * 3 + [cov_acc] * 3).reshape(-1)
    nz["noise_walk"][0] = np.diag([rand_walk_gyr] * 3 + [rand_walk_acc] * 3).reshape(-1)
    return nz


def imu_pre_new(samples, noise, bg=(0, 0, 0), ba=(0, 0, 0), lib=None):
    """IMU_PRE(bg, ba) + push_imu(samples) (preintegration.hpp:32-73); needs no GPU."""
    lib = lib or load_library()
    pre = np.zeros(1, dtype=imu_pre_dtype)
    bg = np.ascontiguousarray(bg, dtype=np.float64); ba = np.ascontiguousarray(ba, dtype=np.float64)
    _chk(lib.vx_imu_pre_init(_p(pre), _p(bg), _p(ba)))
    s = np.ascontiguousarray(samples, dtype=imu_sample_dtype)
    _chk(lib.vx_imu_pre_push(_p(pre), _p(noise), _p(s), C.c_int(s.shape[0])))
    return pre


def imu_pre_evaluate(pre, st1, st2, with_gravity=False, jac=True, lib=None):
    """give_evaluate / give_evaluate_g (preintegration.hpp:137-294) -> (cost, jtj (nc,nc), gg (nc,))."""
    lib = lib or load_library()
    nc = 33 if with_gravity else 30
    jtj = np.zeros((nc, nc)); gg = np.zeros(nc); cost = np.zeros(1)
    _chk(lib.vx_imu_pre_evaluate(_p(pre), _p(st1), _p(st2), C.c_int(int(with_gravity)), _p(jtj) if jac else None, _p(gg) if jac else None, _p(cost)))
    return float(cost[0]), jtj.T.copy(), gg


def imu_pre_evaluate_device(ctx, pres, states, with_gravity=False, jac=True):
    """vx_imu_pre_evaluate_device: the chain of len(pres) IMU factors on the device -> (costs (n,), jtj (n,nc,nc), gg (n,nc))."""
    pres = np.ascontiguousarray(pres, dtype=imu_pre_dtype); states = np.ascontiguousarray(states, dtype=imu_state_dtype)
    n = pres.shape[0]
    assert states.shape[0] == n + 1
    nc = 33 if with_gravity else 30
    jtj = np.zeros((n, nc, nc)); gg = np.zeros((n, nc)); cost = np.zeros(n)
    _chk(ctx.lib.vx_imu_pre_evaluate_device(ctx.h, _p(pres), _p(states), C.c_int(n), C.c_int(int(with_gravity)), _p(jtj) if jac else None,
                                            _p(gg) if jac else None, _p(cost)))
    return cost, np.transpose(jtj, (0, 2, 1)).copy(), gg


class DevArray:
    """__cuda_array_interface__ view of a device buffer owned by the library (for torch.as_tensor)."""

    def __init__(self, ptr, shape, typestr="<f8"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class Factor:
    """Device-resident LidarFactor / VOX_HESS (voxel_map.hpp:124-339)."""

    def __init__(self, ctx, win_size, n_windows=1, handle=None):
        self.ctx, self.lib = ctx, ctx.lib
        self.W, self.n_win = win_size, n_windows
        if handle is not None:
            self.h = handle
            return
        h = C.c_void_p()
        _chk(self.lib.vx_factor_create(ctx.h, C.c_int(win_size), C.c_int(n_windows), C.byref(h)))
        self.h = h

    @classmethod
    def taken_from_hba_global(cls, ctx, which):
        """vx_hba_global_take_factor: the factor a keep_factors run left behind (0 bottom batch, 1 top level) or None."""
        h = C.c_void_p()
        _chk(ctx.lib.vx_hba_global_take_factor(ctx.h, C.c_int(which), C.byref(h)))
        if not h.value:
            return None
        ctx.lib.vx_factor_win_size.argtypes = [C.c_void_p]; ctx.lib.vx_factor_num_windows.argtypes = [C.c_void_p]
        return cls(ctx, int(ctx.lib.vx_factor_win_size(h)), int(ctx.lib.vx_factor_num_windows(h)), handle=h)

    def hessian_sharded_async(self, comm):
        _chk(self.lib.vx_factor_hessian_sharded_async(self.ctx.h, comm.h if comm is not None else None, self.h))

    def free(self):
        if self.h:
            self.lib.vx_factor_free(self.h)
            self.h = None

    @property
    def num_voxels(self):
        return int(self.lib.vx_factor_num_voxels(self.h))

    @property
    def num_voxel_frames(self):
        return int(self.lib.vx_factor_num_voxel_frames(self.h))

    def force_generic(self, on=True):
        _chk(self.lib.vx_factor_force_generic(self.h, C.c_int(int(on))))

    def upload_dense(self, fac):
        """fac: dict with plvec (V,W) cluster, sig (V,), coe (V,), eigval (V,3), eigvec (V,9), merged (V,)."""
        V = fac["sig"].shape[0]
        self._keep = [np.ascontiguousarray(fac[k]) for k in ("plvec", "sig", "coe", "eigval", "eigvec", "merged")]
        a = self._keep
        _chk(self.lib.vx_factor_upload_dense(self.h, C.c_int64(V), _p(a[0]), _p(a[1]), _p(a[2]), _p(a[3]), _p(a[4]), _p(a[5])))

    def upload_csr(self, fac):
        """fac: dict with offsets (V+1,) i8, frames (nnz,) i4, clusters (nnz,), sig, coe, eigval, eigvec, merged."""
        V = fac["sig"].shape[0]
        a = [np.ascontiguousarray(fac["offsets"], dtype=np.int64), np.ascontiguousarray(fac["frames"], dtype=np.int32)] + \
            [np.ascontiguousarray(fac[k]) for k in ("clusters", "sig", "coe", "eigval", "eigvec", "merged")]
        self._keep = a
        _chk(self.lib.vx_factor_upload_csr(self.h, C.c_int64(V), *[_p(x) for x in a]))

    def upload_csr_raw(self, a):
        """Same as upload_csr but takes already-contiguous (e.g. pinned) arrays and does no copies."""
        V = a["sig"].shape[0]
        _chk(self.lib.vx_factor_upload_csr(self.h, C.c_int64(V), *[_p(a[k]) for k in
                                           ("offsets", "frames", "clusters", "sig", "coe", "eigval", "eigvec", "merged")]))

    def hessian_raw(self, ps, H, J, c, head=0, end=-1):
        """vx_factor_hessian into caller-owned buffers (H stays column-major per window)."""
        _chk(self.lib.vx_factor_hessian(self.ctx.h, self.h, _p(ps), C.c_int64(head), C.c_int64(end), _p(H), _p(J), _p(c)))

    def download_csr(self):
        V, nnz = self.num_voxels, self.num_voxel_frames
        off = np.zeros(V + 1, dtype=np.int64); fr = np.zeros(nnz, dtype=np.int32); cl = np.zeros(nnz, dtype=cluster_dtype); coe = np.zeros(V)
        _chk(self.lib.vx_factor_download_csr(self.h, _p(off), _p(fr), _p(cl), _p(coe)))
        ev, evec, mrg = self.download_cache()
        return dict(offsets=off, frames=fr, clusters=cl, coe=coe, eigval=ev, eigvec=evec, merged=mrg)

    def gba_build(self, clouds, poses, voxel_size, min_eigen_value, eig_array, max_layer, frame_cloud=None):
        """OctreeGBA::cut_voxel per frame + OctreeGBA_multi_recut (loop_refine.hpp:446-537) into this factor.
        clouds: list of (n,3) float32 local-frame clouds; frame g uses clouds[frame_cloud[g]] (default g)."""
        xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32)
        off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
        ps = as_poses(poses)
        prm = np.zeros(1, dtype=gba_params_dtype)
        prm["voxel_size"] = voxel_size; prm["min_eigen_value"] = min_eigen_value; prm["eigen_value_array"] = eig_array; prm["max_layer"] = max_layer
        fc = None if frame_cloud is None else np.ascontiguousarray(frame_cloud, dtype=np.int32)
        _chk(self.lib.vx_gba_build(self.ctx.h, xyz.ctypes.data_as(C.c_void_p), _p(off), C.c_int(len(clouds)), _p(fc), _p(ps), _p(prm), self.h))

    def download_cache(self):
        V = self.num_voxels
        ev = np.zeros((V, 3)); evec = np.zeros((V, 9)); mrg = np.zeros(V, dtype=cluster_dtype)
        _chk(self.lib.vx_factor_download_cache(self.h, _p(ev), _p(evec), _p(mrg)))
        return ev, evec, mrg

    def hessian(self, poses, head=0, end=-1):
        ps = as_poses(poses)
        jl = 6 * self.W
        H = np.zeros((self.n_win, jl, jl)); J = np.zeros((self.n_win, jl)); c = np.zeros(self.n_win)
        _chk(self.lib.vx_factor_hessian(self.ctx.h, self.h, _p(ps), C.c_int64(head), C.c_int64(end), _p(H), _p(J), _p(c)))
        H = np.transpose(H, (0, 2, 1))  # column-major -> numpy [row, col]
        if self.n_win == 1:
            return H[0], J[0], float(c[0])
        return H, J, c

    def residual(self, poses, head=0, end=-1):
        ps = as_poses(poses)
        c = np.zeros(self.n_win)
        _chk(self.lib.vx_factor_residual(self.ctx.h, self.h, _p(ps), C.c_int64(head), C.c_int64(end), _p(c)))
        return float(c[0]) if self.n_win == 1 else c

    def ba_lidar(self, poses, max_iter=3, thd_num=2, hess_buf=None):
        """Lidar_BA_Optimizer::damping_iter (voxel_map.hpp:422-497) on every window of the factor.  hess_buf: optional
        C-contiguous float64 array (n_win, 6W, 6W) to receive the Hessian snapshots (the reference's MatrixXd *hess is the
        caller's too); a fresh one costs a page fault per 4 KB, 30 ms at W = 399."""
        ps = as_poses(poses).copy()
        jl = 6 * self.W
        if hess_buf is None:
            hess = np.zeros((self.n_win, jl, jl))
        else:
            hess = hess_buf
            if hess.shape != (self.n_win, jl, jl) or hess.dtype != np.float64 or not hess.flags.c_contiguous:
                raise ValueError("hess_buf must be a C-contiguous float64 array of shape (%d, %d, %d)" % (self.n_win, jl, jl))
        resis = np.zeros((self.n_win, 2))
        conv = np.zeros(self.n_win, dtype=np.int32); trace = np.full((self.n_win, max(max_iter, 1)), -1, dtype=np.int32)
        nit = np.zeros(self.n_win, dtype=np.int32)
        _chk(self.lib.vx_ba_lidar(self.ctx.h, self.h, _p(ps), _p(hess), _p(resis), C.c_int(max_iter), C.c_int(thd_num),
                                  _p(conv), _p(trace), _p(nit)))
        return dict(poses=ps, hess=np.transpose(hess, (0, 2, 1)), resis=resis, converged=conv,
                    accept=[trace[w, :nit[w]].tolist() for w in range(self.n_win)])

    def ba_li(self, states, imus, imu_coef=1e-4):
        """LI_BA_Optimizer::damping_iter (voxel_map.hpp:624-713).  states: (W,) imu_state_dtype; imus: (W-1,) imu_pre_dtype."""
        x = np.ascontiguousarray(states, dtype=imu_state_dtype).copy(); pr = np.ascontiguousarray(imus, dtype=imu_pre_dtype).copy()
        n = 15 * self.W
        hess = np.zeros((n, n)); trace = np.full(3, -1, dtype=np.int32); nit = np.zeros(1, dtype=np.int32)
        _chk(self.lib.vx_ba_li(self.ctx.h, self.h, _p(x), _p(pr), C.c_double(imu_coef), _p(hess), _p(trace), _p(nit)))
        return dict(states=x, imus=pr, hess=hess.T.copy(), accept=trace[:nit[0]].tolist())

    def ba_li_gravity(self, states, imus, imu_coef=1e-4, max_iter=2):
        """LI_BA_OptimizerGravity::damping_iter (voxel_map.hpp:878-975)."""
        x = np.ascontiguousarray(states, dtype=imu_state_dtype).copy(); pr = np.ascontiguousarray(imus, dtype=imu_pre_dtype).copy()
        n = 15 * self.W + 3
        hess = np.zeros((n, n)); trace = np.full(max(max_iter, 1), -1, dtype=np.int32); nit = np.zeros(1, dtype=np.int32); resis = np.zeros(2)
        _chk(self.lib.vx_ba_li_gravity(self.ctx.h, self.h, _p(x), _p(pr), C.c_double(imu_coef), _p(resis), _p(hess), C.c_int(max_iter), _p(trace), _p(nit)))
        return dict(states=x, imus=pr, hess=hess.T.copy(), resis=resis, accept=trace[:nit[0]].tolist())

    # device-resident path
    def gba_build_sharded(self, comm, clouds, cloud_frame, poses, voxel_size, min_eigen_value, eig_array, max_layer):
        """vx_gba_build_sharded: this rank's clouds (cloud c = frame cloud_frame[c]); poses = ALL frames.  -> points owned."""
        xyz = np.ascontiguousarray(np.concatenate(clouds), dtype=np.float32) if len(clouds) else np.zeros((0, 3), dtype=np.float32)
        off = np.zeros(len(clouds) + 1, dtype=np.int64); off[1:] = np.cumsum([c.shape[0] for c in clouds])
        cf = np.ascontiguousarray(cloud_frame, dtype=np.int32)
        prm = np.zeros(1, dtype=gba_params_dtype)
        prm["voxel_size"] = voxel_size; prm["min_eigen_value"] = min_eigen_value; prm["eigen_value_array"] = eig_array; prm["max_layer"] = max_layer
        ps = as_poses(poses)
        n = C.c_int64(0)
        _chk(self.lib.vx_gba_build_sharded(self.ctx.h, comm.h, _p(xyz), _p(off), C.c_int(len(clouds)), _p(cf), _p(ps), _p(prm), self.h, C.byref(n)))
        return n.value

    def ba_lidar_sharded(self, comm, poses, max_iter=3, thd_num=2):
        ps = as_poses(poses).copy()
        NW, W = self.n_win, self.W
        jl = 6 * W
        hess = np.zeros((NW, jl, jl)); resis = np.zeros((NW, 2)); conv = np.zeros(NW, dtype=np.int32)
        tr = np.full((NW, max(max_iter, 1)), -1, dtype=np.int32); it = np.zeros(NW, dtype=np.int32)
        _chk(self.lib.vx_ba_lidar_sharded(self.ctx.h, comm.h, self.h, _p(ps), _p(hess), _p(resis), C.c_int(max_iter), C.c_int(thd_num), _p(conv), _p(tr), _p(it)))
        return dict(poses=ps, hess=np.transpose(hess, (0, 2, 1)), resis=resis, converged=conv, accept=[tr[w, :it[w]].tolist() for w in range(NW)], iters=it)

    def set_poses(self, poses):
        self._poses_keep = as_poses(poses)
        _chk(self.lib.vx_factor_set_poses(self.ctx.h, self.h, _p(self._poses_keep)))

    def hessian_async(self, head=0, end=-1):
        _chk(self.lib.vx_factor_hessian_async(self.ctx.h, self.h, C.c_int64(head), C.c_int64(end)))

    def residual_async(self, head=0, end=-1):
        _chk(self.lib.vx_factor_residual_async(self.ctx.h, self.h, C.c_int64(head), C.c_int64(end)))

    def device_buffers(self):
        ptrs = [C.c_void_p() for _ in range(4)]
        _chk(self.lib.vx_factor_device_buffers(self.h, *[C.byref(p) for p in ptrs]))
        jl = 6 * self.W
        return (DevArray(ptrs[0].value, (self.n_win, jl, jl)), DevArray(ptrs[1].value, (self.n_win, jl)),
                DevArray(ptrs[2].value, (self.n_win,)), DevArray(ptrs[3].value, (self.n_win,)))
