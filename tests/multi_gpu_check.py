"""Multi-GPU checks that need real GPUs and NCCL; run under torchrun, one process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_check.py

1. Voxelization with ONE exchange step: every rank keys its slice of the keyframe points (vx_gba_point_owner), an NCCL
   all-to-all routes each point to the owner of its root voxel, OctreeGBA cut_voxel + multi_recut run locally
   (vx_gba_build_device); the all-reduced [Hess | JacT | cost] must equal the single-GPU build on all points.
2. One BA with its voxels sharded by range (ShardedEvaluator): all-reduced Hess == unsharded Hess.
Rank 0 prints one JSON line (and exits non-zero on a mismatch)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    import scene
    import synth
    import importlib.util
    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pkg = ge.load_package()
    spec = importlib.util.spec_from_file_location("vx_sharding", os.path.join(ge.PKG_DIR, "sharding.py"))
    sh = importlib.util.module_from_spec(spec); spec.loader.exec_module(sh)
    ctx = pkg.Context(local)
    out = {"world": world}

    # ---- 1. exchange-based voxelization ---------------------------------------------------------------------------
    clouds, (R, p) = scene.config1(pts_per_scan=20000)
    W = len(clouds)
    poses = synth.perturb_poses((R, p), 3, 0.004, 0.03)
    all_xyz = np.concatenate(clouds).astype(np.float32)
    all_frame = np.concatenate([np.full(c.shape[0], g, dtype=np.int32) for g, c in enumerate(clouds)])
    lo, hi = sh.block_range(all_xyz.shape[0], rank, world)
    xyz = torch.from_numpy(all_xyz[lo:hi]).to(dev); frame = torch.from_numpy(all_frame[lo:hi]).to(dev)
    gp = np.zeros(1, dtype=pkg.gba_params_dtype)
    gp["voxel_size"] = 1.0; gp["min_eigen_value"] = 0.01; gp["eigen_value_array"] = [0.5] * 4; gp["max_layer"] = 2
    f = pkg.Factor(ctx, W)
    sh.sharded_gba_build(pkg, ctx, f, xyz, frame, poses, gp, rank, world)            # warm-up (allocations, NCCL channels)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    n_mine = sh.sharded_gba_build(pkg, ctx, f, xyz, frame, poses, gp, rank, world)
    torch.cuda.synchronize(); t_build = time.perf_counter() - t0
    H, J, c = f.hessian(poses)
    ts = [torch.from_numpy(np.ascontiguousarray(H)).to(dev), torch.from_numpy(J).to(dev), torch.tensor([float(c)], dtype=torch.float64, device=dev),
          torch.tensor([f.num_voxels, f.num_voxel_frames, n_mine], dtype=torch.int64, device=dev)]
    sh.allreduce_sum_(ts)
    full = pkg.Factor(ctx, W)
    full.gba_build(clouds, poses, 1.0, 0.01, [0.5] * 4, 2)
    Hf, Jf, cf = full.hessian(poses)
    rel = lambda a, b: float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
    out["exchange_voxelization"] = {
        "points_total": int(all_xyz.shape[0]), "points_after_exchange_sum": int(ts[3][2].item()), "voxels_sum": int(ts[3][0].item()), "voxels_single_gpu": full.num_voxels,
        "voxel_frames_sum": int(ts[3][1].item()), "voxel_frames_single_gpu": full.num_voxel_frames,
        "rel_err_H": rel(ts[0].cpu().numpy(), Hf), "rel_err_J": rel(ts[1].cpu().numpy(), Jf), "rel_err_cost": abs(ts[2].item() - cf) / abs(cf),
        "ms_build_incl_all_to_all": 1e3 * t_build, "bytes_per_point_exchanged": 16}
    ok = (out["exchange_voxelization"]["voxels_sum"] == full.num_voxels and out["exchange_voxelization"]["voxel_frames_sum"] == full.num_voxel_frames
          and out["exchange_voxelization"]["points_after_exchange_sum"] == all_xyz.shape[0]
          and out["exchange_voxelization"]["rel_err_H"] < 1e-11 and out["exchange_voxelization"]["rel_err_J"] < 1e-11 and out["exchange_voxelization"]["rel_err_cost"] < 1e-11)

    # ---- 2. voxel-range sharded evaluation of one replicated factor ----------------------------------------------------
    ev = sh.ShardedEvaluator(full, rank, world)
    Hs, Js, cs = ev.hessian(poses)
    out["voxel_range_allreduce"] = {"head_end": [ev.head, ev.end], "rel_err_H": rel(Hs[0], Hf), "rel_err_J": rel(Js[0] if Js.ndim > 1 else Js, Jf),
                                    "rel_err_cost": abs(float(np.ravel(cs)[0]) - cf) / abs(cf)}
    ok = ok and out["voxel_range_allreduce"]["rel_err_H"] < 1e-11 and out["voxel_range_allreduce"]["rel_err_cost"] < 1e-11
    flag = torch.tensor([1 if ok else 0], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["ok"] = bool(flag.item())
    if rank == 0:
        print(json.dumps(out))
    f.free(); full.free(); ctx.close()
    if world > 1:
        dist.destroy_process_group()
    sys.exit(0 if out["ok"] else 1)


if __name__ == "__main__":
    main()
