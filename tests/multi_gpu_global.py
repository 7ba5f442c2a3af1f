"""Multi-GPU check of the in-library collectives (run under torchrun on N >= 2 GPUs of one box, not a pytest):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 tests/multi_gpu_global.py
Every rank runs vx_hba_global with a vx_comm (bottom windows block-distributed, top level voxel-sharded: owner exchange +
all-reduce); rank 0 also runs the single-GPU call and compares: same edge sets, relative poses / top poses <= 1e-9, and
the sharded map build + sharded LM against the plain ones.  Prints one JSON line."""
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
    import synthgen as sg
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); lr = int(os.environ.get("LOCAL_RANK", "0"))
    n_kf = int(os.environ.get("VX_MG_KF", "300")); n_az = int(os.environ.get("VX_MG_AZ", "150"))
    torch.cuda.set_device(lr)
    dist.init_process_group("gloo")          # plumbing only (id broadcast, barriers): the data path is the library's own NCCL communicator
    pkg = ge.load_package(); pkg.load_library()
    ctx = pkg.Context(lr)
    comm = pkg.Comm.from_torch(ctx)
    LOCAL = (0.5, 0.0025, [0.25] * 4); GBA = (1.0, 0.01, [0.5] * 4)
    d = sg.c4_session(n_kf=n_kf, n_az=n_az, seed=20241008 + 90, nx=4, ny=2)
    sess = np.zeros(n_kf, dtype=np.int32)
    prm = pkg.hba_global_params(LOCAL, GBA, 2, 3)
    job = pkg.HbaGlobalJob(d["xyz"], d["off"], sess, d["start"], prm)
    out = job.run(ctx, comm)                  # warm-up (allocations, NCCL channels)
    dist.barrier(); t0 = time.perf_counter()
    out = job.run(ctx, comm)
    e1 = out["edges1"].copy(); e2 = out["edges2"].copy(); tp = out["top_poses"].copy(); st = dict(out["stats"])
    dist.barrier(); t_multi = time.perf_counter() - t0
    res = {"world": world, "n_kf": n_kf, "points": int(d["off"][-1]), "ms_sharded": 1e3 * t_multi, "stats_rank0": st, "comm": comm.stats(),
           "nccl_version": int(pkg.load_library().vx_comm_nccl_version())}
    # every rank must hold the same answer
    h = np.array([float(np.sum(e1["cov"])), float(np.sum(e2["cov"])), float(np.sum(tp["p"])), float(len(e1)), float(len(e2))])
    hs = [None] * world
    dist.all_gather_object(hs, h.tolist())
    res["ranks_agree"] = bool(all(x == hs[0] for x in hs))
    if rank == 0:
        job1 = pkg.HbaGlobalJob(d["xyz"], d["off"], sess, d["start"], prm)
        job1.run(ctx, None)
        t0 = time.perf_counter()
        ref = job1.run(ctx, None)
        res["ms_single"] = 1e3 * (time.perf_counter() - t0)
        res["stats_single"] = ref["stats"]
        ok = len(ref["edges1"]) == len(e1) and len(ref["edges2"]) == len(e2)
        res["edges1"] = [len(e1), len(ref["edges1"])]; res["edges2"] = [len(e2), len(ref["edges2"])]
        if ok:
            for f in ("id1", "id2"):
                ok = ok and bool(np.array_equal(ref["edges1"][f], e1[f]) and np.array_equal(ref["edges2"][f], e2[f]))
            res["edges1_tra_err"] = float(np.max(np.abs(ref["edges1"]["tra"] - e1["tra"])))      # bottom windows are not sharded: exact
            res["edges2_tra_err"] = float(np.max(np.abs(ref["edges2"]["tra"] - e2["tra"]))) if len(e2) else 0.0
            res["edges2_cov_rel"] = float(np.max(np.abs(ref["edges2"]["cov"] - e2["cov"]) / np.abs(ref["edges2"]["cov"]))) if len(e2) else 0.0
            res["top_pose_err"] = float(max(np.max(np.abs(ref["top_poses"]["R"] - tp["R"])), np.max(np.abs(ref["top_poses"]["p"] - tp["p"]))))
            ok = ok and res["edges1_tra_err"] == 0.0 and res["edges2_tra_err"] < 1e-9 and res["top_pose_err"] < 1e-9 and res["edges2_cov_rel"] < 1e-6
        res["match_single_gpu"] = bool(ok)
    # sharded map build + sharded LM on one W = 40 window against the plain calls
    W = 40
    cl = sg.clouds_of(d)[:W]
    poses = (d["start"][0][:W], d["start"][1][:W])
    mine = [i for i in range(W) if i % world == rank]
    fs = pkg.Factor(ctx, W)
    owned = fs.gba_build_sharded(comm, [cl[i] for i in mine], mine, poses, 1.0, 0.01, [0.5] * 4, 2)
    rs = fs.ba_lidar_sharded(comm, poses, max_iter=3, thd_num=5)
    cnt = [None] * world
    dist.all_gather_object(cnt, [int(owned), int(fs.num_voxels), int(fs.num_voxel_frames)])
    if rank == 0:
        f1 = pkg.Factor(ctx, W); f1.gba_build(cl, poses, 1.0, 0.01, [0.5] * 4, 2)
        r1 = f1.ba_lidar(poses, max_iter=3, thd_num=5)
        res["build"] = {"points": [c[0] for c in cnt], "voxels": [c[1] for c in cnt], "voxel_frames": [c[2] for c in cnt],
                        "single": [int(sum(c.shape[0] for c in cl)), int(f1.num_voxels), int(f1.num_voxel_frames)]}
        res["build_counts_match"] = bool(sum(c[0] for c in cnt) == res["build"]["single"][0] and sum(c[1] for c in cnt) == f1.num_voxels
                                         and sum(c[2] for c in cnt) == f1.num_voxel_frames)
        res["lm_hess_rel"] = float(np.max(np.abs(rs["hess"] - r1["hess"])) / np.max(np.abs(r1["hess"])))
        res["lm_pose_err"] = float(max(np.max(np.abs(rs["poses"]["R"] - r1["poses"]["R"])), np.max(np.abs(rs["poses"]["p"] - r1["poses"]["p"]))))
        res["lm_accept"] = [rs["accept"], r1["accept"]]
        res["ok"] = bool(res["ranks_agree"] and res["match_single_gpu"] and res["build_counts_match"] and res["lm_hess_rel"] < 1e-10 and res["lm_pose_err"] < 1e-9
                         and rs["accept"] == r1["accept"])
        print(json.dumps(res), flush=True)
        f1.free()
    fs.free(); comm.free(); ctx.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
