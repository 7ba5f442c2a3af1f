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
