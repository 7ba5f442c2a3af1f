"""BASELINE.json configs[4] (C5) at size: 5 synthetic sessions x 500 keyframes x 20,000 points (~5 x 10^7 points) — voxelization +
cut_voxel + hierarchical global BA across the connected sessions, sharded over the GPUs of one box (run under torchrun, N = 8):
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29555 scripts/c5_run.py
Every rank calls vx_hba_global with the library's NCCL communicator (bottom windows block-sharded; top level: points exchanged by
root-voxel owner, Hess/JacT/cost all-reduced).  Rank 0 then runs the same call on ONE GPU and compares.  One JSON line."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); lr = int(os.environ.get("LOCAL_RANK", "0"))
    n_sess = int(os.environ.get("VX_C5_SESSIONS", "5")); n_kf = int(os.environ.get("VX_C5_KF", "500")); n_az = int(os.environ.get("VX_C5_AZ", "625"))
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // world))
    import synthgen as sg
    torch.cuda.set_device(lr)
    dist.init_process_group("gloo")
    pkg = ge.load_package(); pkg.load_library()
    ctx = pkg.Context(lr)
    comm = pkg.Comm.from_torch(ctx) if world > 1 else None
    # sessions: loops through different row pairs of the tile world, each recorded in its OWN frame (rigid offset), aligned by kf_x
    xyz, off, sess, R0, p0, Rx, px = [], [0], [], [], [], [], []
    rows = [(0, 1), (1, 2), (2, 3), (0, 1), (1, 2)]
    for s in range(n_sess):
        d = sg.c4_session(n_kf=n_kf, n_az=n_az, seed=20241008 + 50 + s, rows=rows[s % 5], phase=7.0 * s)
        xyz.append(d["xyz"]); off += (off[-1] + d["off"][1:]).tolist(); sess += [s] * n_kf
        Rs, ps = d["start"]
        Rx.append(Rs); px.append(ps)                                # the pose graph has already aligned the sessions (multimap_scanPoses)
        yaw = 0.3 * s; c, sn = np.cos(yaw), np.sin(yaw)
        Ro = np.array([[c, -sn, 0], [sn, c, 0], [0, 0, 1.0]]); to = np.array([3.0 * s, -2.0 * s, 0.1 * s])
        R0.append(np.einsum("ij,njk->nik", Ro, Rs)); p0.append(ps @ Ro.T + to)   # Keyframe::x0: the session's own odometry frame
    xyz = np.concatenate(xyz); off = np.array(off, dtype=np.int64); sess = np.array(sess, dtype=np.int32)
    x0 = (np.concatenate(R0), np.concatenate(p0)); x = (np.concatenate(Rx), np.concatenate(px))
    LOCAL = (0.5, 0.0025, [0.25] * 4); GBA = (1.0, 0.01, [0.5] * 4)
    prm = pkg.hba_global_params(LOCAL, GBA, 2, 3)
    job = pkg.HbaGlobalJob(xyz, off, sess, x0, prm, kf_x=x, cnct_map=np.arange(n_sess))
    job.run(ctx, comm)
    dist.barrier(); t0 = time.perf_counter()
    out = job.run(ctx, comm)
    e1 = out["edges1"].copy(); e2 = out["edges2"].copy(); tp = out["top_poses"].copy(); st = dict(out["stats"])
    dist.barrier(); t_multi = time.perf_counter() - t0
    res = {"config": "C5: %d sessions x %d keyframes x %d rays" % (n_sess, n_kf, 32 * n_az), "world": world, "points": int(off[-1]), "ms_sharded": 1e3 * t_multi,
           "stats_rank0": st, "comm": comm.stats() if comm else None, "edges1": int(len(e1)), "edges2": int(len(e2)),
           "inter_session_edges": int(np.sum(e2["m1"] != e2["m2"]))}
    allst = [None] * world
    dist.all_gather_object(allst, [st["top_points"], st["vf_evals_top"], st["vf_evals_bottom"]])
    res["top_points_per_rank"] = [a[0] for a in allst]
    if rank == 0 and world > 1 and os.environ.get("VX_C5_NO_SINGLE") is None:
        job1 = pkg.HbaGlobalJob(xyz, off, sess, x0, prm, kf_x=x, cnct_map=np.arange(n_sess))
        job1.run(ctx, None)
        t0 = time.perf_counter(); ref = job1.run(ctx, None); res["ms_single_gpu"] = 1e3 * (time.perf_counter() - t0)
        res["stats_single"] = ref["stats"]
        res["edges_single"] = [int(len(ref["edges1"])), int(len(ref["edges2"]))]
        res["top_points_single"] = ref["stats"]["top_points"]
        res["top_points_sum_matches"] = bool(sum(res["top_points_per_rank"]) == ref["stats"]["top_points"])
        res["edges1_identical"] = bool(len(ref["edges1"]) == len(e1) and np.array_equal(ref["edges1"].view(np.uint8), e1.view(np.uint8)))
        if len(ref["edges1"]) == len(e1):
            res["edges1_ids_equal"] = bool(np.array_equal(ref["edges1"]["id1"], e1["id1"]) and np.array_equal(ref["edges1"]["id2"], e1["id2"]))
            res["edges1_tra_maxdiff"] = float(np.abs(ref["edges1"]["tra"] - e1["tra"]).max()); res["edges1_rot_maxdiff"] = float(np.abs(ref["edges1"]["rot"] - e1["rot"]).max())
            res["edges1_cov_rel"] = float((np.abs(ref["edges1"]["cov"] - e1["cov"]) / np.abs(ref["edges1"]["cov"])).max())
        res["vf_evals_bottom_sum"] = int(sum(a[2] for a in allst)); res["vf_evals_top_sum"] = int(sum(a[1] for a in allst))
        d0 = np.abs(ref["top_poses"]["p"] - x[1][ref["top_kf"]]).max()
        res["top_correction_max_m"] = float(d0)
        res["top_pose_err"] = float(max(np.max(np.abs(ref["top_poses"]["R"] - tp["R"])), np.max(np.abs(ref["top_poses"]["p"] - tp["p"]))))
        # edges2: the sets may differ in a handful of borderline entries (|hess| against 1e-6 on sums in a different order)
        k1 = set(zip(ref["edges2"]["id1"].tolist(), ref["edges2"]["id2"].tolist(), ref["edges2"]["m1"].tolist(), ref["edges2"]["m2"].tolist()))
        k2 = set(zip(e2["id1"].tolist(), e2["id2"].tolist(), e2["m1"].tolist(), e2["m2"].tolist()))
        res["edges2_symmetric_difference"] = len(k1 ^ k2)
    if rank == 0:
        print(json.dumps(res), flush=True)
    if comm:
        comm.free()
    ctx.close()
    dist.barrier(); dist.destroy_process_group()


if __name__ == "__main__":
    main()
