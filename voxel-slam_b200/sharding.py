"""Multi-GPU partitioning of the BA hot path.  One process per GPU, torch.distributed for plumbing.

* Bottom-level GBA windows (voxelslam.cpp:3137-3138) are independent: `block_range` gives each rank a contiguous
  block of windows; no data-path collective.
* One BA whose voxels are sharded (top-level GBA, W ~ 399; C3): Hess, JacT and cost are plain sums over voxels —
  the reference already sums per-thread partials (voxel_map.hpp:382-384) — so every rank evaluates its voxel range
  [head, end) (the same head/end arguments LidarFactor::acc_evaluate2 takes) and the results are summed with ONE
  all-reduce per array over NCCL (gloo on CPU in the tests); the LM solve is replicated (deterministic)."""
import numpy as np


def block_range(n, rank, world):
    """Contiguous block distribution of n units: rank r gets [lo, hi); sizes differ by at most one."""
    base, extra = divmod(int(n), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def window_shard(n_windows, rank, world):
    return block_range(n_windows, rank, world)


def voxel_shard(n_voxels, rank, world):
    """(head, end) for LidarFactor::acc_evaluate2(xs, head, end, ...) on this rank."""
    return block_range(n_voxels, rank, world)


def allreduce_sum_(tensors, group=None):
    """In-place SUM all-reduce of a list of torch tensors (CUDA -> NCCL over NVLink, CPU -> gloo)."""
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return tensors
    for t in tensors:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return tensors


class ShardedEvaluator:
    """acc_evaluate2 / evaluate_only_residual of ONE factor replicated on every rank, each rank evaluating its own
    voxel range on its GPU, followed by the all-reduce of [Hess | JacT | cost] on the library's stream."""

    def __init__(self, factor, rank, world):
        self.f, self.rank, self.world = factor, rank, world
        self.head, self.end = voxel_shard(factor.num_voxels, rank, world)

    def _on_lib_stream(self):
        import torch
        return torch.cuda.stream(torch.cuda.ExternalStream(self.f.ctx.stream, device=torch.device("cuda", self.f.ctx.device)))

    def hessian(self, poses):
        import torch
        self.f.set_poses(poses)
        self.f.hessian_async(self.head, self.end)
        H, J, c, _ = self.f.device_buffers()
        dev = torch.device("cuda", self.f.ctx.device)
        with self._on_lib_stream():
            ts = [torch.as_tensor(x, device=dev) for x in (H, J, c)]
            allreduce_sum_(ts)
            out = [t.cpu().numpy() for t in ts]
        self.f.ctx.sync()
        return np.transpose(out[0], (0, 2, 1)), out[1], out[2]

    def residual(self, poses):
        import torch
        self.f.set_poses(poses)
        self.f.residual_async(self.head, self.end)
        _, _, _, cr = self.f.device_buffers()
        dev = torch.device("cuda", self.f.ctx.device)
        with self._on_lib_stream():
            t = torch.as_tensor(cr, device=dev)
            allreduce_sum_([t])
            out = t.cpu().numpy()
        self.f.ctx.sync()
        return out


# ---- voxelization across GPUs: one exchange step ------------------------------------------------
def route_by_owner(payloads, owner, world, group=None):
    """All-to-all of per-point payload tensors by destination rank.
    payloads: list of tensors whose first dimension is the local point count; owner: int tensor (n,) in [0, world).
    Points are sent grouped by owner in their original order (stable sort), and arrive ordered by (source rank, original
    order) — both deterministic.  Works on CUDA tensors (NCCL) and CPU tensors (gloo, the tests).  Returns the received
    payloads and the per-source receive counts."""
    import torch
    import torch.distributed as dist
    order = torch.sort(owner.to(torch.int64), stable=True).indices
    send_counts = torch.bincount(owner.to(torch.int64), minlength=world).to(torch.int64)
    if world == 1 or not dist.is_initialized():
        return [p[order] for p in payloads], send_counts
    recv_counts = torch.empty_like(send_counts)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    sc, rc = send_counts.tolist(), recv_counts.tolist()
    out = []
    for p in payloads:
        src = p[order].contiguous()
        dst = torch.empty((int(sum(rc)),) + tuple(p.shape[1:]), dtype=p.dtype, device=p.device)
        dist.all_to_all_single(dst, src, output_split_sizes=rc, input_split_sizes=sc, group=group)
        out.append(dst)
    return out, recv_counts


def clouds_by_frame(xyz, frame, n_frames):
    """Regroups routed points into one cloud per frame (stable, so the order inside a frame stays deterministic):
    returns xyz sorted by frame and the host offsets (n_frames + 1) vx_gba_build_device expects."""
    import torch
    order = torch.sort(frame.to(torch.int64), stable=True).indices
    counts = torch.bincount(frame.to(torch.int64), minlength=n_frames)
    off = np.zeros(n_frames + 1, dtype=np.int64)
    off[1:] = np.cumsum(counts.cpu().numpy())
    return xyz[order].contiguous(), off


def sharded_gba_build(pkg, ctx, factor, xyz, frame, poses, gba_params, rank, world, group=None):
    """OctreeGBA::cut_voxel + multi_recut with the ROOT voxels partitioned over `world` GPUs.
    xyz (n,3) float32 / frame (n,) int32: this rank's slice of the keyframe points, CUDA tensors on ctx's device.
    After the call `factor` holds the LidarFactor of the root voxels this rank owns; Hess / JacT / cost of the whole problem are
    the all-reduced sums (ShardedEvaluator with head=0, end=-1 on every rank, or `allreduce_sum_`)."""
    import ctypes as C
    import torch
    n = int(xyz.shape[0])
    ps = pkg.as_poses(poses)
    F = ps.shape[0]
    xyz = xyz.contiguous(); frame = frame.to(torch.int32).contiguous()
    owner = torch.empty(n, dtype=torch.int32, device=xyz.device)
    torch.cuda.current_stream(xyz.device).synchronize()      # the cast / copy above ran on torch's stream; the library reads on its own
    lib = ctx.lib
    rc = lib.vx_gba_point_owner(ctx.h, C.c_void_p(xyz.data_ptr()), C.c_void_p(frame.data_ptr()), C.c_int64(n), ps.ctypes.data_as(C.c_void_p), C.c_int(F),
                                C.c_double(float(gba_params["voxel_size"][0])), C.c_int(world), C.c_void_p(owner.data_ptr()))
    if rc != 0:
        raise pkg.VxError(rc, lib.vx_last_error().decode())
    (rx, rf), _ = route_by_owner([xyz, frame], owner, world, group)
    sx, off = clouds_by_frame(rx, rf, F)
    torch.cuda.synchronize(xyz.device)
    rc = lib.vx_gba_build_device(ctx.h, C.c_void_p(sx.data_ptr() if sx.numel() else 0), off.ctypes.data_as(C.c_void_p), C.c_int(F), None,
                                 ps.ctypes.data_as(C.c_void_p), gba_params.ctypes.data_as(C.c_void_p), factor.h)
    if rc != 0:
        raise pkg.VxError(rc, lib.vx_last_error().decode())
    return int(sx.shape[0])
