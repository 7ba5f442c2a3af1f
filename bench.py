#!/usr/bin/env python
"""bench.py — LiDAR-BA Hessian build throughput (voxel-frames/s) on B200, one JSON line.

Workload (config.workload): the bottom level of the hierarchical global BA of BASELINE.json
configs[3] — 399 independent 10-keyframe windows (voxelslam.cpp:3033-3034, :3096) evaluated as ONE
batched LidarFactor (include/voxelba.h: vx_factor_create(n_windows)), synthetic plane voxels.
A step = one LidarFactor::acc_evaluate2 pass (voxel_map.hpp:150-282) over every window: Hess, JacT and
cost of all 399 windows.  Per rank the work is fixed (weak scaling, no data-path collective: the
windows are independent); rank r evaluates its own 2,000-keyframe job.

  value      voxel-frames/s, inputs resident in HBM, CUDA-event timed on the library's stream
  e2e        same metric through the C ABI with HOST buffers: vx_factor_hessian on a LidarFactor built beforehand
             (H2D poses from pinned memory, D2H Hess/JacT/cost) — the call the reference arm times;
             e2e.with_upload adds vx_factor_upload_csr (H2D of the reference-layout PointCluster records) to every step
  roofline   dominant kernel of the step, algorithmic bytes / its event-timed duration
  cpu_baseline  the oracle (CPU restatement of the reference) on a bounded sample of the same windows

--impl reference times the oracle's Lidar_BA_Optimizer::divide_thread on host threads (the reference
itself needs Eigen/PCL/ROS/GTSAM and cannot be built here: DESIGN.md "Oracle").
"""
import argparse
import contextlib
import io
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "lidar_ba_hessian_build_voxel_frames_per_s"
UNIT = "voxel-frames/s"
N_WINDOWS = 399          # (2000 - 10) / 5 + 1, voxelslam.cpp:3137-3138
WIN = 10                 # wdsize, voxelslam.cpp:3033
VOX_PER_WIN = 1000
OCC = 0.6


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--windows", type=int, default=N_WINDOWS)
    ap.add_argument("--voxels-per-window", type=int, default=VOX_PER_WIN)
    ap.add_argument("--cpu-sample-windows", type=int, default=24)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the other-rows throughput legs")
    return ap.parse_args()


def workload_config(args, fac):
    nnz = int(fac["offsets"][-1]); V = int(fac["sig"].shape[0])
    return {
        "workload": "C4 bottom level: %d independent %d-keyframe GBA windows batched in one LidarFactor, %d plane voxels/window, "
                    "frame occupancy %.1f (synthetic planar patches)" % (args.windows, WIN, args.voxels_per_window, OCC),
        "windows": args.windows, "win_size": WIN, "voxels": V, "voxel_frames": nnz,
        "pairs": int(np.sum((np.diff(fac["offsets"]) * (np.diff(fac["offsets"]) - 1)) // 2)),
        "parallelism": "windows sharded, no collective",
    }


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 200 ms while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index; self.proc = None; self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True); self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for k, nm in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def build_workload(args, rank):
    import synth
    return synth.batched_factor_csr(20241008 + 3 + 1000 * rank, args.windows, args.voxels_per_window, WIN, occupancy=OCC, pts_max=12)


def algorithmic_bytes(nnz, V, n_win, W):
    """76 B per voxel-frame + 132 B per voxel + poses + the upper block triangle,
    gradient and cost written once per window."""
    jl = 6 * W
    return 76 * nnz + 132 * V + 96 * W * n_win + (8 * jl * (jl + 1) // 2 + 8 * jl + 8) * n_win


def algorithmic_flops(fac):
    k = np.diff(fac["offsets"])
    return int(900 * k.sum() + 216 * np.sum(k * (k - 1) // 2))


def cpu_hessian_sample(fac, windows, threads, repeats=1):
    """Oracle acc_evaluate2 over `windows` (list of window ids) on `threads` host threads
    (one window per task).  Returns (voxel-frames processed per pass, best seconds per pass)."""
    import concurrent.futures as cf
    import oracle_bind as ob
    import synth
    items = []
    nvf = 0
    for b in windows:
        d = synth.csr_window(fac, b)
        of = ob.Factor(WIN); of.set_dense(d)
        items.append((of, ob.as_poses(d["poses"])))
        nvf += of.num_voxel_frames
    best = float("inf")
    with cf.ThreadPoolExecutor(max_workers=threads) as ex:
        for _ in range(repeats):
            t0 = time.perf_counter()
            list(ex.map(lambda it: it[0].hessian(it[1]), items))
            best = min(best, time.perf_counter() - t0)
    return nvf, best


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle restatement) on all host threads."""
    if rank != 0:
        return
    fac = build_workload(args, 0)
    cores = os.cpu_count() or 1
    nwin = min(args.windows, max(args.cpu_sample_windows, 8 * cores))
    windows = list(range(nwin))
    # one step = one pass over the sample; warm-up passes first
    import oracle_bind as ob  # noqa: F401  (builds/loads the oracle)
    # time K passes back to back on a persistent pool
    import concurrent.futures as cf
    import synth
    items = []
    for b in windows:
        d = synth.csr_window(fac, b)
        of = ob.Factor(WIN); of.set_dense(d)
        items.append((of, ob.as_poses(d["poses"])))
    nvf = sum(it[0].num_voxel_frames for it in items)
    with cf.ThreadPoolExecutor(max_workers=cores) as ex:
        for _ in range(max(args.warmup, 1)):
            list(ex.map(lambda it: it[0].hessian(it[1]), items))
        t0 = time.perf_counter()
        for _ in range(args.steps):
            list(ex.map(lambda it: it[0].hessian(it[1]), items))
        tot = time.perf_counter() - t0
    val = nvf * args.steps / tot
    sample = "%d of %d windows (%d voxel-frames) per step, oracle acc_evaluate2, one window per task" % (nwin, args.windows, nvf)
    cfg = workload_config(args, fac)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference needs Eigen/PCL/ROS/GTSAM (absent): timed the oracle restatement of its CPU path",
    }))


def extra_benches(pkg, ctx):
    """Throughput of the other rows of the path on config-shaped synthetic inputs, each through the C ABI with HOST
    buffers (wall clock, copies included).  Reported beside the headline, not part of `value`."""
    import scene
    import synth
    out = {}
    rng = np.random.Generator(np.random.PCG64(20241008 + 4))
    # --- C4-shaped: 25 keyframes, 20k-ray Avia scans, windows of 10 / stride 5, replicated to 128 windows
    clouds, (R, p) = scene.config1(seed=20241008 + 4, W=25, pts_per_scan=20000)
    Rs, ps_ = synth.perturb_poses((R, p), 4, 0.004, 0.03)
    W = 10
    base = np.concatenate([np.arange(s0, s0 + W) for s0 in (0, 5, 10, 15)]).astype(np.int32)
    fc = np.tile(base, 32); NW = fc.size // W
    inst = int(sum(clouds[c].shape[0] for c in fc))
    f = pkg.Factor(ctx, W, NW)
    f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)      # warm-up
    t0 = time.perf_counter()
    for _ in range(3):
        f.gba_build(clouds, (Rs[fc], ps_[fc]), 1.0, 0.01, [0.5] * 4, 2, frame_cloud=fc)
    t = (time.perf_counter() - t0) / 3
    out["gba_build"] = {"points_per_s": inst / t, "ms": 1e3 * t, "windows": NW, "point_instances": inst, "voxels": f.num_voxels,
                        "voxel_frames": f.num_voxel_frames, "what": "vx_gba_build (OctreeGBA::cut_voxel + multi_recut), host clouds in, device LidarFactor out"}
    f.free()
    LOCAL = (0.5, 0.0025, [0.25] * 4); GBA = (1.0, 0.01, [0.5] * 4)
    pkg.hba_add_edge(ctx, clouds, (Rs[fc], ps_[fc]), W, NW, LOCAL, GBA, 2, 1, frame_cloud=fc)
    t0 = time.perf_counter()
    r = pkg.hba_add_edge(ctx, clouds, (Rs[fc], ps_[fc]), W, NW, LOCAL, GBA, 2, 1, frame_cloud=fc)
    t = time.perf_counter() - t0
    out["hba_bottom_level"] = {"windows_per_s": NW / t, "ms": 1e3 * t, "windows": NW, "edges": int(sum(len(e) for e in r["edges"])),
                               "what": "vx_hba_add_edge, max_iter 1: map build + damping_iter(up=4) + edges for every window"}
    # --- C2-shaped: one ~60k-point XT32 sweep against a plane map of earlier sweeps
    rects = scene.room()
    Rl, pl = scene.arc_trajectory(12, 5, step=0.5, yaw_step_deg=3.0)
    world = []
    for i in range(10):
        c = scene.scan(rects, Rl[i], pl[i], scene.xt32_dirs(700), rng, range_noise=0.01, blind=0.7).astype(np.float64)
        world.append(c @ Rl[i].T + pl[i])
    keys, rn, nodes = scene.synthetic_plane_map(np.concatenate(world), 1.0, node_dtype=pkg.map_node_dtype)
    q = scene.scan(rects, Rl[10], pl[10], scene.xt32_dirs(2000), rng, range_noise=0.01, blind=0.7).astype(np.float64)
    pv = np.zeros((q.shape[0], 12)); pv[:, :3] = q; pv[:, 3] = pv[:, 7] = pv[:, 11] = 1e-4
    gm = pkg.PlaneMap(ctx, keys, rn, nodes, 1.0)
    cov = np.eye(15) * 1e-4
    R0, p0 = synth.perturb_poses((Rl[10:11], pl[10:11]), 3, 0.003, 0.02, fix_first=False)
    o = gm.lio_pass(pv, (R0, p0), cov)
    ctx.profile(True)
    t0 = time.perf_counter()
    for _ in range(10):
        o = gm.lio_pass(pv, (R0, p0), cov)
    t = (time.perf_counter() - t0) / 10
    prof = ctx.profile_read(); ctx.profile(False)
    kms = prof.get("k_lio_pass", (0.0, 1))[0] / max(prof.get("k_lio_pass", (0.0, 1))[1], 1)
    out["lio_pass"] = {"points_per_s_e2e": q.shape[0] / t, "ms_e2e": 1e3 * t, "points": int(q.shape[0]), "matched": int(o["match_num"]),
                       "kernel_ms": kms, "points_per_s_kernel": q.shape[0] / (kms * 1e-3) if kms > 0 else None,
                       "kernel_hbm_gbs": (96 * q.shape[0] + 408 * o["match_num"]) / (kms * 1e-3) / 1e9 if kms > 0 else None,
                       "what": "vx_lio_match_accumulate: one pass of voxelslam.cpp:1005-1050 over a ~60k-point sweep (synthetic layer-0 plane map)"}
    # the call the LIO loop makes per sweep (voxelslam.cpp:962-1098): raw float xyz in (12 B/point), var_init on the device,
    # up to 4 match/update iterations, state + covariance out
    q32 = q.astype(np.float32)
    ext = (np.eye(3)[None], np.zeros((1, 3)))
    se = gm.lio_state_estimation(q32, R0[0], p0[0], cov, xyz_ext=(ext, 0.01, 0.01))
    t0 = time.perf_counter()
    for _ in range(10):
        se = gm.lio_state_estimation(q32, R0[0], p0[0], cov, xyz_ext=(ext, 0.01, 0.01))
    t = (time.perf_counter() - t0) / 10
    out["lio_state_estimation"] = {"ms_e2e": 1e3 * t, "points": int(q.shape[0]), "iterations": int(se["iters"]),
                                   "points_per_s_e2e": q.shape[0] / t, "point_iterations_per_s_e2e": q.shape[0] * se["iters"] / t,
                                   "h2d_bytes": int(q32.nbytes),
                                   "what": "vx_lio_state_estimation_xyz: whole ESIKF update of one ~60k-point sweep, host scan in, host state out (wall clock)"}
    gm.free()
    # the reference's CPU path for the same update (oracle restatement, one thread as in the reference's LIO loop), on a map
    # the oracle builds from the same sweeps
    try:
        import oracle_bind as ob
        rng2 = np.random.Generator(np.random.PCG64(20241008 + 5))
        om = ob.LocalMap(win_size=10, voxel_size=1.0, thre=(1, 1, 1, 1), dept_err=0.01, beam_err=0.01)
        for i in range(10):
            om.push(scene.scan(rects, Rl[i], pl[i], scene.xt32_dirs(700), rng2, range_noise=0.01, blind=0.7), (Rl[i:i + 1], pl[i:i + 1]), ob.default_cov())
        opv = ob.var_init(q, 0.01, 0.01)
        best = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            oref = om.lio_state_estimation(opv, R0[0], p0[0], ob.default_cov())
            best = min(best, time.perf_counter() - t0)
        out["lio_state_estimation"]["cpu_baseline"] = {"ms": 1e3 * best, "iterations": int(oref["iters"]), "points_per_s": q.shape[0] / best,
                                                       "point_iterations_per_s": q.shape[0] * oref["iters"] / best, "cores": 1, "kind": "port",
                                                       "sample": "the same sweep, oracle lio_state_estimation on an oracle-built map of the same 10 sweeps, best of 3"}
    except Exception as e:  # the CPU leg is a reported baseline only
        out["lio_state_estimation"]["cpu_baseline"] = {"error": str(e)[:200]}
    # --- C2-shaped local mapping: the device sliding-window map fed with ~60k-point sweeps (A5 -> A6 -> A9/A11 -> A12)
    Wl = 10
    Rm, pm = scene.arc_trajectory(Wl + 4, 5, step=0.5, yaw_step_deg=3.0)
    mposes = pkg.as_poses((Rm, pm))
    vm = pkg.VoxelMap(ctx, win_size=Wl, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2)
    fl = pkg.Factor(ctx, Wl)
    ident = (np.eye(3)[None], np.zeros((1, 3)))
    win = np.zeros(0, dtype=pkg.pose_dtype)
    tt = {"front_end": [], "cut_voxel": [], "recut": [], "margi": []}
    npts = 0
    ctx.profile(True)
    for i in range(Wl + 4):
        xyz = scene.scan(rects, Rm[i], pm[i], scene.xt32_dirs(2000), rng, range_noise=0.01, blind=0.7).astype(np.float32)
        t0 = time.perf_counter()
        t1 = t0
        win = np.concatenate([win, mposes[i:i + 1]])
        vm.cut_voxel_xyz(xyz, ident, 0.01, 0.01, mposes[i:i + 1], cov, win.shape[0] - 1)
        t2 = time.perf_counter()
        vm.recut(win, fl)
        t3 = time.perf_counter()
        if i >= 4:      # skip the cold scans
            tt["front_end"].append(t1 - t0); tt["cut_voxel"].append(t2 - t1); tt["recut"].append(t3 - t2); npts += xyz.shape[0]
        if win.shape[0] >= Wl:
            t4 = time.perf_counter()
            vm.margi(win, fl); vm.shift_window(1); win = win[1:].copy()
            tt["margi"].append(time.perf_counter() - t4)
    prof = ctx.profile_read(); ctx.profile(False)
    n_meas = len(tt["cut_voxel"])
    med = lambda v: float(np.median(v)) if len(v) else None
    kern = {k: {"ms_total": v[0], "launches": v[1]} for k, v in prof.items() if k.startswith("k_") and v[0] > 0}
    out["local_map"] = {"points_per_scan": npts // max(n_meas, 1), "nodes": vm.num_nodes, "roots": vm.num_roots, "factor_voxels": fl.num_voxels,
                        "ms_cut_voxel": 1e3 * med(tt["cut_voxel"]), "ms_recut_tras_opt": 1e3 * med(tt["recut"]),
                        "ms_margi": 1e3 * med(tt["margi"]) if tt["margi"] else None,
                        "cut_voxel_points_per_s": (npts / max(n_meas, 1)) / med(tt["cut_voxel"]),
                        "kernel_ms_over_session": {k: round(v["ms_total"], 4) for k, v in sorted(kern.items(), key=lambda kv: -kv[1]["ms_total"])[:8]},
                        "what": "vx_map_cut_voxel_xyz (var_init + pvec_update + cut_voxel_multi from the raw scan), vx_map_recut (+tras_opt into the device factor), vx_map_margi per 64k-ray "
                                "XT32 sweep, W=10, voxel 1 m, host buffers in (wall clock, median over the steady-state scans)"}
    # --- B4: LiDAR-inertial LM on the factor the map just produced (W = 10), synthetic IMU
    states, segs = synth.imu_trajectory(20241008 + 6, Wl, gyr_sigma=1e-3, acc_sigma=1e-2)
    fli = synth.random_factor(20241008 + 7, 2000, Wl, poses=synth.states_poses(states), occupancy=0.7, min_frames=2)
    f2 = pkg.Factor(ctx, Wl); f2.upload_dense(fli)
    nz = pkg.imu_noise(0.01, 0.1, 1e-4, 1e-3)
    imus = np.concatenate([pkg.imu_pre_new(sg, nz, states[k]["bg"], states[k]["ba"]) for k, sg in enumerate(segs)])
    start = states.copy(); start["p"][1:] += rng.normal(0, 0.03, (Wl - 1, 3))
    f2.ba_li(start, imus)
    t0 = time.perf_counter()
    for _ in range(5):
        r = f2.ba_li(start, imus)
    t = (time.perf_counter() - t0) / 5
    out["li_ba"] = {"ms_per_damping_iter_call": 1e3 * t, "lm_iterations": len(r["accept"]), "voxels": 2000, "win_size": Wl,
                    "what": "vx_ba_li (LI_BA_Optimizer::damping_iter, 3 LM iterations): device B1/B2 + host IMU factors and 150x150 LDL^T"}
    f2.free(); fl.free(); vm.free()
    # --- C4-top shaped: one window over W = 399 submaps (2394 unknowns): B1 through the generic pair-list path + the blocked
    #     device LDL^T of the LM solve (voxel_map.hpp:457-458)
    Wt, Vt = 399, 20000
    top = synth.batched_factor_csr(20241008 + 8, 1, Vt, Wt, occupancy=4.0 / Wt, min_frames=2, pts_max=12)
    ft = pkg.Factor(ctx, Wt); ft.upload_csr(top)
    ft.residual(top["poses"])
    Ht, Jt, ct = ft.hessian(top["poses"])
    t0 = time.perf_counter()
    for _ in range(3):
        Ht, Jt, ct = ft.hessian(top["poses"])
    t_h = (time.perf_counter() - t0) / 3
    n_t = 6 * Wt
    At = Ht.copy(); At[:6, :] = 0; At[:, :6] = 0; At[np.arange(6), np.arange(6)] = 1.0
    At[np.arange(n_t), np.arange(n_t)] *= 1.01
    bt = -Jt.copy(); bt[:6] = 0
    xs = pkg.ldlt_solve(ctx, At, bt)
    t0 = time.perf_counter()
    xs = pkg.ldlt_solve(ctx, At, bt)
    t_s = time.perf_counter() - t0
    sym = np.tril(At) + np.tril(At, -1).T
    # the whole top-level damping_iter (voxel_map.hpp:422-497): B1 + LM solve + B2 per iteration, Hess / JacT resident
    start_t = synth.perturb_poses(top["poses"], 7, 0.002, 0.02)
    ft.residual(start_t)
    hbuf = np.ones((1, n_t, n_t))                       # the caller's MatrixXd, resident (a fresh 45.8 MB array page-faults for ~30 ms)
    rt = ft.ba_lidar(start_t, max_iter=3, hess_buf=hbuf)
    ft.residual(start_t)
    t0 = time.perf_counter()
    rt = ft.ba_lidar(start_t, max_iter=3, hess_buf=hbuf)
    t_ba = time.perf_counter() - t0
    ft.residual(start_t)
    t0 = time.perf_counter()
    ft.ba_lidar(start_t, max_iter=1, hess_buf=hbuf)
    t_ba1 = time.perf_counter() - t0
    t_it = (t_ba - t_ba1) / max(len(rt["accept"][0]) - 1, 1)      # iterations 2, 3 (rejected here): solve + B2, no new Hessian
    out["top_level_gba"] = {"ms_damping_iter_call": 1e3 * t_ba, "ms_per_later_lm_iteration": 1e3 * t_it,
                            "device_solve_gflops": (n_t ** 3 / 3) / t_it / 1e9, "lm_iterations": len(rt["accept"][0]),
                            "accept": [int(a) for a in rt["accept"][0]],"win_size": Wt, "unknowns": n_t, "voxels": Vt, "voxel_frames": ft.num_voxel_frames,
                            "ms_hessian_e2e": 1e3 * t_h, "ms_ldlt_solve_from_host_matrix": 1e3 * t_s,
                            "solve_residual_rel": float(np.abs(sym @ xs - bt).max() / max(np.abs(bt).max(), 1e-300)),
                            "what": "vx_ba_lidar at W=399 (Hess/JacT resident, blocked device LDL^T + device substitution per LM iteration; the call also "
                                    "returns the 45.8 MB Hessian snapshot); vx_factor_hessian alone with the 45.8 MB Hess to the host; vx_ldlt_solve of "
                                    "the gauge-fixed damped system from a HOST matrix (H2D 45.8 MB + the same device solve), wall clock; ms_per_later_lm_iteration = "
                                    "(call with max_iter 3 - call with max_iter 1) / 2: pivot order + permutation + graph-replayed factorisation + substitutions + B2"}
    # the reference's CPU path for the LM solve of that window: Eigen's unblocked diagonal-pivot LDL^T (oracle restatement, one
    # thread as in the reference: voxel_map.hpp:457-458 runs on the optimizer's own thread), once; also the parity check
    try:
        import oracle_bind as ob
        t0 = time.perf_counter()
        xo = ob.ldlt_solve(At, bt)
        t_cpu = time.perf_counter() - t0
        out["top_level_gba"]["cpu_baseline"] = {"ms_ldlt_solve": 1e3 * t_cpu, "gflops": (n_t ** 3 / 3) / t_cpu / 1e9, "cores": 1, "kind": "port",
                                                "device_vs_cpu_rel": float(np.abs(xs - xo).max() / max(np.abs(xo).max(), 1e-300)),
                                                "sample": "one (Hess + u D).ldlt().solve(-JacT) of the same 2394-unknown system (the device "
                                                          "iteration above also contains the B2 pass and the retraction)"}
    except Exception as e:  # the CPU leg is a reported baseline only
        out["top_level_gba"]["cpu_baseline"] = {"error": str(e)[:200]}
    ft.free()
    return out


def pinned_like(a):
    """Copy a numpy array into page-locked host memory (torch-owned), same dtype/shape view."""
    import torch
    raw = torch.empty(max(a.nbytes, 1), dtype=torch.uint8).pin_memory()
    view = raw.numpy()[:a.nbytes].view(a.dtype).reshape(a.shape)
    view[...] = a
    return view, raw


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    pkg = ge.load_package()
    pkg.load_library()
    fac = build_workload(args, rank)
    nnz = int(fac["offsets"][-1]); V = int(fac["sig"].shape[0])
    cfg = workload_config(args, fac)
    ctx = pkg.Context(local_rank)
    f = pkg.Factor(ctx, WIN, args.windows)
    f.upload_csr(fac)
    f.set_poses(fac["poses"])
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local_rank))
    alg_bytes = algorithmic_bytes(nnz, V, args.windows, WIN)
    resident = 76 * nnz + 8 * nnz + 132 * V + 104 * V       # what one pass touches besides its own scratch
    cfg["l2"] = "inputs larger than L2: %.0f MB read per step vs 126 MB L2, no flush" % (alg_bytes / 1e6)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ---------------------------------------------------------------
    for _ in range(args.warmup):
        f.hessian_async()
    barrier()
    l0 = ctx.launch_count
    clocks = ClockSampler(local_rank); clocks.start()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        f.hessian_async()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count - l0
    # keep the GPU busy a little longer so nvidia-smi gets several samples under load
    t_end = time.time() + 1.5
    while time.time() < t_end:
        for _ in range(10):
            f.hessian_async()
        ctx.sync()
    clk = clocks.stop()
    if world > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_step = ms / args.steps
    value = nnz * world / (ms_step * 1e-3)

    # ---- per-kernel timing (roofline) ---------------------------------------------------------
    ctx.profile(True)
    for _ in range(args.steps):
        f.hessian_async()
    prof = ctx.profile_read()
    ctx.profile(False)
    kern = {k: {"ms_per_step": v[0] / args.steps, "launches_per_step": v[1] / args.steps} for k, v in prof.items()}
    top = max(kern, key=lambda k: kern[k]["ms_per_step"])
    pass_ms = sum(v["ms_per_step"] for v in kern.values())
    import json as _j
    peaks = {}
    try:
        peaks = _j.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = alg_bytes / (pass_ms * 1e-3) / 1e9
    flops = algorithmic_flops(fac)
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of k_hess_small from the committed ncu --set full capture of this workload
                # (profiles/r01m_k_hess_small_ncu.txt: 249.45 + 24.55 MB); only valid for the default workload size
                "traffic": 274003968 if (args.windows, args.voxels_per_window) == (N_WINDOWS, VOX_PER_WIN) else None,
                "traffic_source": "profiles/r01m_k_hess_small_ncu.txt (ncu --set full, one launch of k_hess_small)",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650 GB/s",
                "kernel": "whole acc_evaluate2 pass (sum of its kernels' event times); top kernel = " + top,
                "algorithmic_bytes_per_step": alg_bytes, "kernels": kern,
                "fp64": {"algorithmic_flops_per_step": flops, "achieved_tflops": flops / (pass_ms * 1e-3) / 1e12,
                         "note": "B1 is FP64-FMA bound for k>=3 occupied frames per voxel (DESIGN.md); no FP64 peak in MEASURED_PEAKS.json"}}

    # ---- B2 residual pass, same factor (reported beside the headline) ------------------------
    for _ in range(args.warmup):
        f.residual_async()
    barrier()
    e0.record(stream)
    for _ in range(args.steps):
        f.residual_async()
    e1.record(stream)
    barrier()
    ms_res = e0.elapsed_time(e1) / args.steps
    res_bytes = 76 * nnz + (104 + 200) * V
    extra = {"residual_pass": {"voxel_frames_per_s": nnz / (ms_res * 1e-3), "ms_per_step": ms_res,
                               "hbm_gbs": res_bytes / (ms_res * 1e-3) / 1e9, "hbm_frac": res_bytes / (ms_res * 1e-3) / 1e9 / peak}}

    # ---- end to end through the C ABI with host buffers ---------------------------------------
    keys = ("offsets", "frames", "clusters", "sig", "coe", "eigval", "eigvec", "merged")
    host = {}; owners = []
    for k in keys:
        a = np.ascontiguousarray(fac[k], dtype=np.int64 if k == "offsets" else (np.int32 if k == "frames" else None))
        host[k], o = pinned_like(a); owners.append(o)
    ps, o = pinned_like(pkg.as_poses(fac["poses"])); owners.append(o)
    jl = 6 * WIN
    Hh, o1 = pinned_like(np.zeros((args.windows, jl, jl))); Jh, o2 = pinned_like(np.zeros((args.windows, jl))); ch, o3 = pinned_like(np.zeros(args.windows))
    owners += [o1, o2, o3]
    h2d = sum(host[k].nbytes for k in keys) + ps.nbytes
    d2h = Hh.nbytes + Jh.nbytes + ch.nbytes
    g = pkg.Factor(ctx, WIN, args.windows)

    def timed(fn, n):
        for _ in range(2):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(n):
            fn()
        barrier()
        t = (time.perf_counter() - t0) / n
        if world > 1:
            tt = torch.tensor([t], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        return t

    n_e2e = max(3, min(args.steps, 10))
    # (1) the evaluation call on a LidarFactor built beforehand — what Lidar_BA_Optimizer::damping_iter does every LM iteration
    #     (voxel_map.hpp:437) and what --impl reference times: poses in from pinned host memory, Hess / JacT / cost out to it
    g.upload_csr_raw(host)
    t_eval = timed(lambda: g.hessian_raw(ps, Hh, Jh, ch), max(n_e2e, args.steps))
    # (2) the same with the factor pushed from the host every step (LidarFactor::push_voxel for all voxels + the evaluation)
    def e2e_step():
        g.upload_csr_raw(host)
        g.hessian_raw(ps, Hh, Jh, ch)
    t_e2e = timed(e2e_step, n_e2e)
    e2e = {"value": nnz * world / t_eval, "unit": UNIT, "h2d_bytes_per_step": ps.nbytes, "d2h_bytes_per_step": d2h,
           "ms_per_step": 1e3 * t_eval, "steps": max(n_e2e, args.steps),
           "what": "vx_factor_hessian on a resident LidarFactor: poses H2D from pinned host memory, Hess/JacT/cost D2H into it, wall clock "
                   "(the reference arm times the same call, LidarFactor::acc_evaluate2 on factors built beforehand)",
           "with_upload": {"value": nnz * world / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                           "ms_per_step": 1e3 * t_e2e, "steps": n_e2e,
                           "what": "vx_factor_upload_csr (pinned host PointCluster records -> device, validation) + vx_factor_hessian every step"}}
    # sanity: e2e result equals the resident result
    Hd = np.zeros_like(Hh); Jd = np.zeros_like(Jh); cd = np.zeros_like(ch)
    f.hessian_raw(ps, Hd, Jd, cd)
    # (f's cache was refreshed by the B2 pass above, g's is the uploaded one: equal to rounding, not bitwise)
    extra["e2e_vs_resident"] = {"rel_err_H": float(np.max(np.abs(Hd - Hh)) / np.max(np.abs(Hh))),
                                "rel_err_cost": float(np.max(np.abs(cd - ch) / np.abs(ch)))}
    assert extra["e2e_vs_resident"]["rel_err_H"] < 1e-6, extra["e2e_vs_resident"]

    # ---- cpu baseline (rank 0, N=1 only) ------------------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        nwin = min(args.windows, args.cpu_sample_windows)
        nvf, t = cpu_hessian_sample(fac, list(range(nwin)), 2, repeats=3)
        cpu = {"value": nvf / t, "unit": UNIT, "cores": 2, "kind": "port",
               "sample": "%d of %d windows (%d voxel-frames), oracle acc_evaluate2 on 2 host threads (the reference's thd_num for "
                         "LiDAR-only BA, voxel_map.hpp:345), best of 3; host has %d cores" % (nwin, args.windows, nvf, os.cpu_count() or 1)}
        # parity spot check on the sampled windows (the oracle here is the checker, not the product)
        import oracle_bind as ob
        import synth
        d = synth.csr_window(fac, 0)
        of = ob.Factor(WIN); of.set_dense(d)
        Ho, Jo, co = of.hessian(d["poses"])
        Hg = Hh[0].T
        err = float(np.max(np.abs(Hg - Ho)) / np.max(np.abs(Ho)))
        extra["parity_window0_rel_err_H"] = err
        assert err < 1e-7, err

    if rank == 0 and world == 1 and not args.no_extras:
        extra["other_rows"] = extra_benches(pkg, ctx)
    extra["fp64_peak_tflops"] = {"dfma": ctx.fp64_peak(0), "dmma": ctx.fp64_peak(1), "mixed_half_half": ctx.fp64_peak(2),
                                 "how": "vx_bench_fp64 (csrc/vx_peaks.cu), measured in this run; mixed < either alone: DFMA and DMMA share the FP64 datapath"}
    roofline["fp64"]["peak_tflops"] = extra["fp64_peak_tflops"]["dfma"]
    roofline["fp64"]["frac"] = roofline["fp64"]["achieved_tflops"] / extra["fp64_peak_tflops"]["dfma"]
    # Which ceiling binds?  B1's algorithmic intensity (flops / bytes of this workload) against the ridge of the measured peaks:
    # above the ridge the FP64 datapath is the roofline (DFMA and the m8n8k4 DMMA share it, see fp64_peak_tflops), below it HBM.
    intensity = flops / alg_bytes
    ridge = extra["fp64_peak_tflops"]["dfma"] * 1e12 / (peak * 1e9)
    hbm_view = {k: roofline[k] for k in ("achieved", "peak", "unit", "frac", "peak_source")}
    roofline["intensity_flop_per_byte"] = intensity
    roofline["ridge_flop_per_byte"] = ridge
    if intensity > ridge:
        roofline.update({"bound": "tensor", "achieved": roofline["fp64"]["achieved_tflops"], "peak": roofline["fp64"]["peak_tflops"], "unit": "TFLOP/s",
                         "frac": roofline["fp64"]["frac"],
                         "peak_source": "FP64 datapath peak measured in this run by vx_bench_fp64 (DFMA; the DMMA figure is within 9 % and the two share "
                                        "the datapath); MEASURED_PEAKS.json carries no FP64 number",
                         "hbm": hbm_view})
    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": cfg, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
               "roofline": roofline, "cpu_baseline": cpu}
        out.update(extra)
        print(json.dumps(out))
    f.free(); g.free(); ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    # stdout carries exactly ONE JSON line: libraries that write to file descriptor 1 themselves (NCCL prints its version
    # banner there) are pointed at stderr for the duration of the run
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    buf = io.StringIO()
    try:
        with contextlib.redirect_stdout(buf):
            if args.impl == "reference":
                run_reference(args, rank, world)
            else:
                run_b200(args, rank, world, local_rank)
    finally:
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        os.close(real_stdout)
    lines = [ln for ln in buf.getvalue().splitlines() if ln.strip()]
    for ln in lines:
        if ln.lstrip().startswith("{"):
            print(ln, flush=True)
        else:
            print(ln, file=sys.stderr, flush=True)


if __name__ == "__main__":
    main()
