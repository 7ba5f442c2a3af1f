#!/usr/bin/env python
"""bench.py — the hierarchical global LiDAR BA of BASELINE.json configs[3] on B200, one JSON line.

Workload (config.workload): ONE global BA over 2,000 synthetic keyframes (XT32-shaped, 20,000 points each, a loop through a
10 x 4 tile world, odometry-like drift): VOXEL_SLAM::thd_globalmapping (voxelslam.cpp:3018-3141) — 400 bottom windows of
10 keyframes / stride 5 through HBA_add_edge (max_iter 1), one submap per window, the top-level HBA_add_edge over the 400
submaps (GBA/total_max_iter = 3).  The total work is FIXED as N grows ("scaling": "strong"): bottom windows are
block-distributed over the ranks (no collective), the top level is voxel-sharded (exchange of 16 B/point by root-voxel owner
per map build, one all-reduce of [tri(Hess) | JacT | cost] = 23 MB per evaluation), the LM solve is replicated.  All of it
runs inside libvoxelba.so (vx_hba_global + vx_comm, NCCL); torch.distributed only launches the ranks and carries the NCCL id.

  step       one whole job (every map build, acc_evaluate2 / evaluate_only_residual pass, LM solve, edge set)
  value      voxel-frames through acc_evaluate2 (LidarFactor Hessian builds, voxel_map.hpp:150-282) per second, keyframe clouds
             RESIDENT in HBM, CUDA-event timed on the library's stream, max over ranks
  e2e        the same job through the C ABI from PINNED HOST keyframe clouds (H2D inside the timed region), edges and poses out
  roofline   the kernel with the largest share of the step, algorithmic bytes / its event-timed duration
  cpu_baseline / --impl reference   the oracle's HBA_add_edge (CPU restatement of the reference) on a bounded sample of the same
             bottom windows, all host threads, same metric (voxel-frames through its acc_evaluate2 per second)
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
LOCAL = (0.5, 0.0025, [0.25] * 4)     # Odometry/voxel_size (config string 0.5), min_eigen_value, 1/plane_eigen_value_thre (avia.yaml)
GBA = (1.0, 0.01, [0.5] * 4)          # GBA block of hesai.yaml: voxel_size 1, min_eigen_value 0.01, 1/eigen_value_array
MAX_LAYER = 2
SEED = 20241008 + 4


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=12)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--keyframes", type=int, default=2000)
    ap.add_argument("--n-az", type=int, default=625, help="azimuth steps of the 32-ring scan: 625 -> 20,000 rays per keyframe")
    ap.add_argument("--total-max-iter", type=int, default=3, help="GBA/total_max_iter (hesai.yaml)")
    ap.add_argument("--cpu-sample-windows", type=int, default=0, help="bottom windows of the CPU arm's mini-job (0: 40)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the per-row roofline / other-rows legs")
    return ap.parse_args()


def c4_workload(args, world):
    """Seeded C4 input, identical on every rank (tests/synthgen.c ray caster)."""
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // max(world, 1)))   # read by libgomp when synthgen loads
    import synthgen as sg
    d = sg.c4_session(n_kf=args.keyframes, n_az=args.n_az, seed=SEED)
    d["session"] = np.zeros(args.keyframes, dtype=np.int32)
    return d


def n_windows_of(n_kf):
    n_full = max(0, (n_kf - 10) // 5 + 1)
    tail = n_kf - 5 * n_full if n_full else n_kf      # keyframes left in the queue when the global pass fires
    return n_full, n_full + (1 if tail >= 2 else 0)


def workload_config(args, d):
    """Identical for both arms: depends on the arguments and the generated input only."""
    n_kf = args.keyframes
    n_full, n_win = n_windows_of(n_kf)
    jl = 6 * n_win
    packed = 8 * (jl * (jl + 1) // 2 + 15 * n_win + jl + 1)
    return {
        "workload": "C4: one hierarchical global BA over %d synthetic keyframes (%d points, XT32-shaped 32 x %d rays, 10 x 4 tile world loop): "
                    "thd_globalmapping schedule, bottom windows of 10 / stride 5 (HBA_add_edge max_iter 1) -> one submap each -> top-level HBA_add_edge "
                    "(total_max_iter %d)" % (n_kf, int(d["off"][-1]), args.n_az, args.total_max_iter),
        "keyframes": n_kf, "points": int(d["off"][-1]), "bottom_windows": n_win, "top_level_frames": n_win, "win_size": 10,
        "voxel_size": {"gba": GBA[0], "local": LOCAL[0]}, "max_layer": MAX_LAYER, "total_max_iter": args.total_max_iter,
        "parallelism": "strong scaling of ONE job: bottom windows block-sharded over the ranks (no collective, edges gathered by one integer all-reduce); "
                       "top level voxel-sharded: grouped ncclSend/ncclRecv exchange of 16 B/point by root-voxel owner per map build + one ncclAllReduce of "
                       "[tri(Hess) | diag-block upper | JacT | cost] = %.1f MB per acc_evaluate2; LM solve replicated" % (packed / 1e6),
        "l2": "inputs larger than L2: %.0f MB of keyframe clouds per job vs 126 MB L2, no flush" % (12 * int(d["off"][-1]) / 1e6),
    }


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled every 100 ms while the timed regions run."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index; self.proc = None; self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
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
                "samples": len(sm), "reasons": sorted(reasons), "when": "sampled only while the timed regions (resident + e2e) ran"}


# ---- CPU arm: a complete, smaller hierarchical job through the oracle (CPU restatement of the reference) -----------------------------
def cpu_windows(args, d, n_sample):
    """The first n_sample full bottom windows (consecutive: their submaps form a connected top level)."""
    n_full, _ = n_windows_of(args.keyframes)
    return list(range(max(2, min(n_sample, n_full))))


def cpu_pass(d, windows, threads):
    """thd_globalmapping over the keyframes of `windows` with the oracle: every bottom window through HBA_add_edge(max_iter 1, thread_num 2,
    plptr) (voxelslam.cpp:3096; one window per host thread), then the top-level HBA_add_edge over their submaps (total_max_iter 3, all
    host threads inside).  -> (voxel-frames through the oracle's acc_evaluate2, seconds, (bottom s, top s))."""
    import concurrent.futures as cf
    import ctypes
    import oracle_bind as ob
    lib = ob.lib()
    lib.orc_vf_counter.restype = ctypes.c_int64
    Rs, ps = d["start"]

    def one(w):
        s = 5 * w
        cl = [d["xyz"][d["off"][i]:d["off"][i + 1]] for i in range(s, s + 10)]
        r = ob.hba_add_edge(cl, (Rs[s:s + 10], ps[s:s + 10]), LOCAL, GBA, MAX_LAYER, 1, thread_num=2, ids=np.arange(s, s + 10), mps=np.zeros(10), want_cloud=True)
        m = r["merged"]
        return np.stack([m["x"], m["y"], m["z"]], axis=1).astype(np.float32)

    lib.orc_vf_counter(1)
    t0 = time.perf_counter()
    with cf.ThreadPoolExecutor(max_workers=threads) as ex:
        subs = list(ex.map(one, windows))
    t1 = time.perf_counter()
    tk = [5 * w for w in windows]
    ob.hba_add_edge(subs, (Rs[tk], ps[tk]), LOCAL, GBA, MAX_LAYER, 3, thread_num=threads, ids=tk, mps=np.zeros(len(tk)))
    t2 = time.perf_counter()
    return int(lib.orc_vf_counter(1)), t2 - t0, (t1 - t0, t2 - t1)


def cpu_sample_text(args, n_items, cores):
    return "a complete hierarchical job over the first %d of the %d bottom windows (%d keyframes): every window through the oracle's HBA_add_edge(max_iter 1) " \
           "(map build + damping_iter(up=4) + edges + submap cloud), one window per task on %d threads, then the top-level HBA_add_edge over the %d submaps " \
           "(total_max_iter 3, %d threads inside); the full job's top level (W = 400, 2,400 unknowns, 35 M points) costs the host relatively more" % \
           (n_items, n_windows_of(args.keyframes)[0], 5 * n_items + 5, cores, n_items, cores)


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path (oracle restatement; the reference itself needs Eigen/PCL/ROS/GTSAM) on all host threads."""
    if rank != 0:
        return
    import __graft_entry__ as ge
    ge.build_oracle()
    d = c4_workload(args, 1)
    cores = os.cpu_count() or 1
    items = cpu_windows(args, d, args.cpu_sample_windows or 40)
    for _ in range(max(1, min(args.warmup, 1))):
        cpu_pass(d, items[:max(2, len(items) // 4)], cores)
    vf = 0; tot = 0.0
    for _ in range(args.steps):
        v, t, _ = cpu_pass(d, items, cores)
        vf += v; tot += t
    val = vf / tot
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(args, d),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": cpu_sample_text(args, len(items), cores)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "voxel_frames_per_step": vf // max(args.steps, 1),
        "note": "reference needs Eigen/PCL/ROS/GTSAM (absent): timed the oracle restatement of its CPU path",
    }))


def pinned_like(a):
    """Copy a numpy array into page-locked host memory (torch-owned), same dtype/shape view."""
    import torch
    raw = torch.empty(max(a.nbytes, 1), dtype=torch.uint8).pin_memory()
    view = raw.numpy()[:a.nbytes].view(a.dtype).reshape(a.shape)
    view[...] = a
    return view, raw


# ---- algorithmic bytes / flops per step of the kernels that can dominate it (SURVEY §8d figures) --------------------------------
def factor_ratios(f):
    """Shape of a kept LidarFactor: voxels and occupied pairs per voxel-frame."""
    c = f.download_csr()
    k = np.diff(c["offsets"])
    nnz = max(int(k.sum()), 1)
    return {"W": f.W, "n_win": f.n_win, "V": int(k.size), "nnz": nnz, "pairs": int(np.sum(k * (k - 1) // 2)),
            "v_per_vf": k.size / nnz, "pairs_per_vf": float(np.sum(k * (k - 1) // 2)) / nnz, "mean_frames_per_voxel": float(k.mean()) if k.size else 0.0}


def kernel_models(st, rb, rt, inst_bottom, top_points, top_builds):
    """-> {kernel: (algorithmic bytes per step, flops per step, what)} on THIS rank.  st: job stats (voxel-frames through acc_evaluate2 /
    evaluate_only_residual per level); rb / rt: factor_ratios of the bottom batch / top-level factors."""
    m = {}
    Eb, Et, Rb, Rt = st["vf_evals_bottom"], st["vf_evals_top"], st["vf_res_bottom"], st["vf_res_top"]
    if rb:
        jl = 6 * rb["W"]
        passes = Eb / rb["nnz"]
        out_b = (96 * rb["W"] + 8 * jl * (jl + 1) // 2 + 8 * jl + 8) * rb["n_win"] * passes
        m["k_hess_small"] = (76 * Eb + 132 * rb["v_per_vf"] * Eb + out_b, 900 * Eb + 216 * rb["pairs_per_vf"] * Eb,
                             "B1 acc_evaluate2 of the bottom windows (fused DMMA SYRK): 76 B/voxel-frame + 132 B/voxel + poses and the upper block triangle per window")
    if rt:
        n = 6 * rt["W"]
        m["k_diag"] = (76 * Et + 132 * rt["v_per_vf"] * Et, 900 * Et, "B1 acc_evaluate2 at the top level, per-voxel-frame part: 76 B/voxel-frame + 132 B/voxel")
        m["k_offdiag"] = (2 * 144 * rt["pairs_per_vf"] * Et, 216 * rt["pairs_per_vf"] * Et,
                          "B1 acc_evaluate2 at the top level, pair part: two 18-double rows per occupied frame pair, 216 flop")
        m["k_ldlt_trailing"] = (0, (n ** 3 / 3.0) * st["top_res_passes"], "trailing update of the blocked LDL^T (voxel_map.hpp:457-458, %d unknowns): n^3/3 flop per LM solve" % n)
    vpv = (rb or rt or {"v_per_vf": 0.25})["v_per_vf"]
    m["k_residual"] = (76 * (Rb + Rt) + 304 * vpv * (Rb + Rt), 0, "B2 evaluate_only_residual, both levels: 76 B/voxel-frame + 104 B read + 200 B written per voxel")
    inst = inst_bottom + top_points * top_builds
    m["k_leaf_stats"] = (16 * inst, 0, "A8 / A10 finest-layer PointCluster accumulation straight from the sorted points: 12 B xyz + 4 B point index per point instance read")
    m["k_vf_fill"] = (16 * inst, 0, "A10 scan-local cluster of every (voxel, frame): 12 B xyz + 4 B point index per point instance of an accepted plane (upper bound: all instances)")
    m["k_gba_key64"] = (12 * inst + 12 * inst, 0, "A1 / A8 keys: 12 B/point instance read, 8 B key + 4 B index written")
    m["k_hb_emit"] = (1 * inst, 0, "A8 fused head-counter scan: 1 B of head bits per point instance read (compact index arrays written at the heads)")
    return m


def traffic_of(kernel):
    """DRAM bytes per launch of `kernel` from the committed ncu capture summary (profiles/ncu_traffic.json), or None."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return t.get(kernel, {}).get("dram_bytes_per_launch")
    except (OSError, ValueError):
        return None


def extra_benches(pkg, ctx, peak_hbm, fp64):
    """The other rows of the path on config-shaped inputs, each with its own roofline (algorithmic bytes of SURVEY §8d / event-timed
    kernel / measured peak).  Reported beside the headline, not part of `value`."""
    import scene
    import synth
    import torch
    out = {}
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", ctx.device))
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def timed_async(fn, n=20, warm=3):
        for _ in range(warm):
            fn()
        ctx.sync()
        a, b = ev(), ev()
        a.record(stream)
        for _ in range(n):
            fn()
        b.record(stream)
        ctx.sync()
        return a.elapsed_time(b) / n

    # --- the round-1 headline workload, kept as a labelled second line: B1 / B2 kernels on a SYNTHETIC batched factor (399 windows x
    #     1000 random plane voxels, occupancy 0.6), inputs resident
    fac = synth.batched_factor_csr(20241008 + 3, 399, 1000, 10, occupancy=0.6, pts_max=12)
    nnz = int(fac["offsets"][-1]); V = int(fac["sig"].shape[0]); k = np.diff(fac["offsets"])
    f = pkg.Factor(ctx, 10, 399); f.upload_csr(fac); f.set_poses(fac["poses"])
    f.residual_async(); ctx.sync()
    ms_h = timed_async(f.hessian_async)
    ms_r = timed_async(f.residual_async)
    jl = 60
    b1 = 76 * nnz + 132 * V + 96 * 10 * 399 + (8 * jl * (jl + 1) // 2 + 8 * jl + 8) * 399
    fl = int(900 * k.sum() + 216 * np.sum(k * (k - 1) // 2))
    b2 = 76 * nnz + 304 * V
    out["synthetic_r01_workload"] = {
        "what": "round-1 headline workload (synthetic random factor, NOT produced by the map build): whole acc_evaluate2 / evaluate_only_residual pass, inputs resident",
        "voxel_frames": nnz, "B1": {"ms": ms_h, "voxel_frames_per_s": nnz / (ms_h * 1e-3), "fp64_tflops": fl / (ms_h * 1e-3) / 1e12,
                                    "fp64_frac": fl / (ms_h * 1e-3) / 1e12 / fp64["dfma"], "hbm_gbs": b1 / (ms_h * 1e-3) / 1e9, "hbm_frac": b1 / (ms_h * 1e-3) / 1e9 / peak_hbm},
        "B2": {"ms": ms_r, "voxel_frames_per_s": nnz / (ms_r * 1e-3), "hbm_gbs": b2 / (ms_r * 1e-3) / 1e9, "hbm_frac": b2 / (ms_r * 1e-3) / 1e9 / peak_hbm}}
    f.free()

    # --- C2: LIO on ONE map for both arms: 14 XT32 sweeps (64,000 rays) through the device sliding-window map and, the same sweeps
    #     with the same covariances, through the oracle's map; the 11th sweep is registered against each
    import oracle_bind as ob
    rng = np.random.Generator(np.random.PCG64(20241008 + 2))
    rects = scene.room()
    Wl = 10
    n_sw = Wl + 4
    Rl, pl = scene.arc_trajectory(n_sw + 1, 5, step=0.5, yaw_step_deg=3.0)
    cov = ob.default_cov()
    vm = pkg.VoxelMap(ctx, win_size=Wl, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2)
    om = ob.LocalMap(win_size=Wl, voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2, dept_err=0.01, beam_err=0.01)
    fl_ = pkg.Factor(ctx, Wl)
    win = np.zeros(0, dtype=pkg.pose_dtype)
    t_ins, t_rec, t_mar, t_cpu_ins, npts = [], [], [], [], []
    ctx.profile(True)
    for i in range(n_sw):     # the local-mapping loop without the optimizer: insert, recut, marginalise the oldest scan once the window is full
        xyz = scene.scan(rects, Rl[i], pl[i], scene.xt32_dirs(2000), rng, range_noise=0.01, blind=0.7).astype(np.float32)
        t0 = time.perf_counter()
        pv, pw = om.insert(xyz, (Rl[i:i + 1], pl[i:i + 1]), cov); om.recut()
        t_cpu_ins.append(time.perf_counter() - t0)
        win = np.concatenate([win, pkg.as_poses((Rl[i:i + 1], pl[i:i + 1]))])
        t0 = time.perf_counter()
        vm.cut_voxel(pv, pw, win.shape[0] - 1)
        t1 = time.perf_counter()
        vm.recut(win, fl_)
        t2 = time.perf_counter()
        t_ins.append(t1 - t0); t_rec.append(t2 - t1); npts.append(xyz.shape[0])
        if om.margi():
            t3 = time.perf_counter()
            vm.margi(win, fl_); vm.shift_window(1); win = win[1:].copy()
            t_mar.append(time.perf_counter() - t3)
    prof = ctx.profile_read(); ctx.profile(False)
    n_pt = float(np.mean(npts[2:]))
    ins_kernel_ms = sum(v[0] for kname, v in prof.items() if kname.startswith("k_")) / n_sw
    out["local_map"] = {"points_per_scan": int(n_pt), "nodes": vm.num_nodes, "roots": vm.num_roots, "factor_voxels": fl_.num_voxels,
                        "ms_cut_voxel_e2e": 1e3 * float(np.median(t_ins[2:])), "ms_recut_tras_opt_e2e": 1e3 * float(np.median(t_rec[2:])),
                        "ms_margi_e2e": 1e3 * float(np.median(t_mar)) if t_mar else None,
                        "cut_voxel_points_per_s_e2e": n_pt / float(np.median(t_ins[2:])),
                        "roofline": {"bound": "hbm", "kernel": "all map kernels of one sweep (insertion + recut + tras_opt + margi)",
                                     "achieved": 124 * n_pt / (ins_kernel_ms * 1e-3) / 1e9, "peak": peak_hbm, "unit": "GB/s",
                                     "frac": 124 * n_pt / (ins_kernel_ms * 1e-3) / 1e9 / peak_hbm, "ms_kernels_per_scan": ins_kernel_ms,
                                     "note": "A1+A6: 120 B/point read + 4 B label written (SURVEY §8d); 64 k points per call: launch-latency bound, tens of kernels of a few us"},
                        "cpu_baseline": {"ms_insert_recut": 1e3 * float(np.median(t_cpu_ins[2:])), "cores": 1, "kind": "port",
                                         "sample": "the same sweeps through the oracle's cut_voxel + recut (single-threaded; the reference uses 5 threads)"},
                        "what": "vx_map_cut_voxel + vx_map_recut (+ tras_opt into the device factor) per 64k-ray XT32 sweep, W=10, voxel 1 m, host buffers in (wall clock)"}
    q = scene.scan(rects, Rl[n_sw], pl[n_sw], scene.xt32_dirs(2000), rng, range_noise=0.01, blind=0.7).astype(np.float32)
    opv = ob.var_init(q, 0.01, 0.01)
    R0, p0 = synth.perturb_poses((Rl[n_sw:n_sw + 1], pl[n_sw:n_sw + 1]), 3, 0.003, 0.02, fix_first=False)
    gm = vm.planemap()
    best = 1e30
    for _ in range(3):
        t0 = time.perf_counter()
        oref = om.lio_state_estimation(opv, R0[0], p0[0], cov)
        best = min(best, time.perf_counter() - t0)
    se = gm.lio_state_estimation(opv, R0[0], p0[0], cov)
    ctx.profile(True)
    t0 = time.perf_counter()
    for _ in range(10):
        se = gm.lio_state_estimation(opv, R0[0], p0[0], cov)
    t = (time.perf_counter() - t0) / 10
    prof = ctx.profile_read(); ctx.profile(False)
    kp = prof.get("k_lio_pass", (0.0, 1))
    kms = kp[0] / max(kp[1], 1)
    nmatch = int(se["match_nums"][0]) if len(se["match_nums"]) else 0
    lio_bytes = 96 * q.shape[0] + 228 * nmatch
    out["lio_state_estimation"] = {
        "points": int(q.shape[0]), "iterations": int(se["iters"]), "cpu_iterations": int(oref["iters"]), "matched_first_pass": nmatch,
        "same_result_as_cpu": bool(se["iters"] == oref["iters"] and list(se["match_nums"]) == list(oref["match_nums"])),
        "ms_e2e": 1e3 * t, "point_iterations_per_s_e2e": q.shape[0] * se["iters"] / t, "h2d_bytes": int(opv.nbytes),
        "roofline": {"bound": "hbm", "kernel": "k_lio_pass", "achieved": lio_bytes / (kms * 1e-3) / 1e9 if kms > 0 else None, "peak": peak_hbm, "unit": "GB/s",
                     "frac": lio_bytes / (kms * 1e-3) / 1e9 / peak_hbm if kms > 0 else None, "ms_per_launch": kms, "points_per_s_kernel": q.shape[0] / (kms * 1e-3) if kms > 0 else None,
                     "note": "D2: 96 B/point-iteration + 228 B plane record per matched point (SURVEY §8d); 60 k points do not fill a B200: launch-latency bound"},
        "cpu_baseline": {"ms": 1e3 * best, "point_iterations_per_s": q.shape[0] * oref["iters"] / best, "cores": 1, "kind": "port",
                         "sample": "the same sweep against the oracle's map of the SAME 14 sweeps and covariances (the device map equals it node for node), best of 3"},
        "what": "vx_lio_state_estimation on the device-built sliding-window map: whole ESIKF update of one ~60k-point sweep, host pointVar in, state out (wall clock)"}
    fl_.free(); vm.free()

    # --- C3 at its stated size: W = 20 scans of ~50,000 points (1 M points in the window) through the device sliding-window map
    #     (raw scan in: var_init + pvec_update + cut_voxel_multi, then multi_recut + tras_opt per scan) and
    #     LI_BA_OptimizerGravity::damping_iter(max_iter 5) on the factor it leaves (voxelslam.cpp:1917-1965)
    import c3_scene
    c3 = c3_scene.build(pkg, W=20, n_az=1563)
    W3 = 20
    kw = dict(voxel_size=1.0, thre=(1, 1, 1, 1), max_layer=2)
    Rs3, ps3 = synth.states_poses(c3["start"])
    ident = pkg.as_poses((np.eye(3)[None], np.zeros((1, 3))))

    def c3_device():
        vm3 = pkg.VoxelMap(ctx, win_size=W3, **kw); f3 = pkg.Factor(ctx, W3)
        win3 = np.zeros(0, dtype=pkg.pose_dtype)
        t0 = time.perf_counter()
        for i in range(W3):
            pose = pkg.as_poses((Rs3[i:i + 1], ps3[i:i + 1]))
            win3 = np.concatenate([win3, pose])
            vm3.cut_voxel_xyz(c3["scans"][i], ident, 0.01, 0.01, pose, cov, i)
            vm3.recut(win3, f3)
        t_map = time.perf_counter() - t0
        t0 = time.perf_counter()
        r = f3.ba_li_gravity(c3["start"], c3["imus"], imu_coef=1e-4, max_iter=5)
        t_ba = time.perf_counter() - t0
        nv, nvf = f3.num_voxels, f3.num_voxel_frames
        f3.free(); vm3.free()
        return t_map, t_ba, r, nv, nvf
    c3_device()
    t_map, t_ba, r3, nv3, nvf3 = c3_device()
    om3 = ob.LocalMap(win_size=W3, dept_err=0.01, beam_err=0.01, **kw)
    t0 = time.perf_counter()
    for i in range(W3):
        om3.insert(c3["scans"][i], (Rs3[i:i + 1], ps3[i:i + 1]), cov); om3.recut()
    t_map_cpu = time.perf_counter() - t0
    of3 = om3.factor(W3)
    t0 = time.perf_counter()
    b3 = of3.ba_li(c3["start"], c3["imus"], imu_coef=1e-4, gravity=True, max_iter=5)
    t_ba_cpu = time.perf_counter() - t0
    npts3 = int(sum(sc.shape[0] for sc in c3["scans"]))
    out["c3_li_ba_gravity"] = {
        "points_in_window": npts3, "win_size": W3, "voxels": nv3, "voxel_frames": nvf3, "lm_iterations": len(r3["accept"]), "accept": r3["accept"],
        "ms_map_20_scans_e2e": 1e3 * t_map, "ms_damping_iter_e2e": 1e3 * t_ba, "points_per_s_map_e2e": npts3 / t_map,
        "pos_err_start_m": float(np.abs(c3["start"]["p"] - c3["truth"]["p"]).max()), "pos_err_opt_m": float(np.abs(r3["states"]["p"] - c3["truth"]["p"]).max()),
        "same_trace_as_cpu": bool(r3["accept"] == b3["accept"]), "state_diff_vs_cpu": float(np.abs(r3["states"]["p"] - b3["states"]["p"]).max()),
        "cpu_baseline": {"ms_map_20_scans": 1e3 * t_map_cpu, "ms_damping_iter": 1e3 * t_ba_cpu, "cores": 1, "kind": "port",
                         "sample": "the same 20 scans and IMU factors through the oracle session (single-threaded; the reference uses 5 threads for the map and the LI-BA)"},
        "what": "C3 at size: vx_map_cut_voxel_xyz + vx_map_recut per ~50k-point scan, then vx_ba_li_gravity(max_iter 5), host buffers in / out (wall clock); the device "
                "front end differs from the oracle's var_init at 1e-12, so the CPU comparison is to rounding, not bitwise (tests/test_c3_gpu.py is the parity test)"}
    return out


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback (use --impl reference for the CPU leg)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pkg = ge.load_package()
    pkg.load_library()
    d = c4_workload(args, world)
    cfg = workload_config(args, d)
    ctx = pkg.Context(local_rank)
    comm = pkg.Comm.from_torch(ctx) if world > 1 else None
    stream = torch.cuda.ExternalStream(ctx.stream, device=dev)
    n_kf = args.keyframes

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allsum(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    def allmax(x):
        if world == 1:
            return float(x)
        t = torch.tensor([float(x)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # keyframe clouds: pinned host copy (e2e) and a device-resident copy (value)
    xyz_pin, pin_owner = pinned_like(d["xyz"])
    xyz_dev = torch.empty(d["xyz"].shape, dtype=torch.float32, device=dev)
    xyz_dev.copy_(torch.from_numpy(xyz_pin), non_blocking=False)
    torch.cuda.synchronize()
    prm = pkg.hba_global_params(LOCAL, GBA, MAX_LAYER, args.total_max_iter)
    job_dev = pkg.HbaGlobalJob(None, d["off"], d["session"], d["start"], prm, xyz_device_ptr=xyz_dev.data_ptr())
    job_host = pkg.HbaGlobalJob(xyz_pin, d["off"], d["session"], d["start"], prm)

    # ---- resident timing -------------------------------------------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        out = job_dev.run(ctx, comm)
    vf_step = allsum(out["stats"]["voxel_frame_evals"])
    barrier()
    clocks = ClockSampler(local_rank); clocks.start()
    l0 = ctx.launch_count
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    phase = {}
    for _ in range(args.steps):
        out = job_dev.run(ctx, comm)
        for k, v in out["stats"].items():
            if k.startswith("ms_"):
                phase[k] = phase.get(k, 0.0) + v / args.steps
    e1.record(stream)
    barrier()
    ms_step = allmax(e0.elapsed_time(e1)) / args.steps
    launches = ctx.launch_count - l0
    value = vf_step / (ms_step * 1e-3)

    # ---- end to end: pinned host clouds in, edges + poses out (wall clock) -----------------------------------------------------
    for _ in range(2):
        out_h = job_host.run(ctx, comm)
    barrier()
    t0 = time.perf_counter()
    phase_h = {}
    for _ in range(args.steps):
        out_h = job_host.run(ctx, comm)
        for k, v in out_h["stats"].items():
            if k.startswith("ms_"):
                phase_h[k] = phase_h.get(k, 0.0) + v / args.steps
    barrier()
    t_e2e = allmax(time.perf_counter() - t0) / args.steps
    clk = clocks.stop()
    n_full, n_win = n_windows_of(n_kf)
    h2d = int(12 * d["off"][-1]) + 96 * 2 * n_kf      # every keyframe cloud once over all ranks + poses
    d2h = int(out_h["edges1"].nbytes + out_h["edges2"].nbytes + out_h["top_poses"].nbytes)
    e2e = {"value": vf_step / t_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": 1e3 * t_e2e, "steps": args.steps,
           "phases_ms_rank0": {k: round(v, 3) for k, v in phase_h.items()},
           "what": "vx_hba_global from pinned host keyframe clouds (each rank uploads the clouds of its own bottom windows), pose-graph edges of both "
                   "levels and the optimized submap poses back to the host, wall clock with a barrier on both sides, max over ranks"}
    same = bool(len(out_h["edges1"]) == len(out["edges1"]) and len(out_h["edges2"]) == len(out["edges2"]) and
                np.array_equal(out_h["top_poses"]["p"], out["top_poses"]["p"]))

    # ---- per-kernel timing of one job (rank-local), kept factors for the byte models ----------------------------------------------
    prm_keep = prm.copy(); prm_keep["keep_factors"] = 1
    job_prof = pkg.HbaGlobalJob(None, d["off"], d["session"], d["start"], prm_keep, xyz_device_ptr=xyz_dev.data_ptr())
    ctx.profile(True)
    outp = job_prof.run(ctx, comm)
    prof = ctx.profile_read()
    ctx.profile(False)
    fb = pkg.Factor.taken_from_hba_global(ctx, 0); ft = pkg.Factor.taken_from_hba_global(ctx, 1)
    rb = factor_ratios(fb) if fb is not None and fb.num_voxel_frames > 0 else None
    rt = factor_ratios(ft) if ft is not None and ft.num_voxel_frames > 0 else None
    stp = outp["stats"]
    # point instances this rank keys at the bottom level (every window re-references its 10 keyframe clouds)
    ws, wst, wsz = pkg.hba_schedule([n_kf])
    lo_w, hi_w = pkg.hba_window_range(wsz, 10, rank, world)
    sizes = np.diff(d["off"])
    inst_bottom = int(sum(sizes[wst[w]:wst[w] + wsz[w]].sum() for w in range(lo_w, hi_w)))
    n_groups = len(set(int(x) for x in wsz[lo_w:hi_w]))
    top_builds = max(prof.get("k_leaf_stats", (0.0, n_groups))[1] - n_groups, 0)
    models = kernel_models(stp, rb, rt, inst_bottom, stp["top_points"], top_builds)
    kern = {k: {"ms_per_step": v[0], "launches_per_step": v[1]} for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}
    kernel_ms_total = sum(v[0] for v in prof.values())
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak_hbm = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (burst copy)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    fp64 = {"dfma": ctx.fp64_peak(0), "dmma": ctx.fp64_peak(1)}
    rows = {}
    for k, (nbytes, nflops, what) in models.items():
        if k not in prof or prof[k][0] <= 0:
            continue
        t = prof[k][0] * 1e-3
        row = {"ms_per_step": prof[k][0], "launches_per_step": prof[k][1], "share_of_kernel_time": prof[k][0] / kernel_ms_total, "what": what}
        if nbytes:
            row.update({"hbm_gbs": nbytes / t / 1e9, "hbm_frac": nbytes / t / 1e9 / peak_hbm, "algorithmic_bytes_per_step": int(nbytes)})
        if nflops:
            row.update({"fp64_tflops": nflops / t / 1e12, "fp64_frac": nflops / t / 1e12 / fp64["dfma"], "algorithmic_flops_per_step": int(nflops)})
        if nbytes and nflops:
            row["intensity_flop_per_byte"] = nflops / nbytes
        rows[k] = row
    ridge = fp64["dfma"] * 1e12 / (peak_hbm * 1e9)
    top = next((k for k in kern if k in rows), None)
    roofline = None
    if top:
        r = rows[top]
        tensor_bound = ("fp64_frac" in r) and (("hbm_frac" not in r) or r.get("intensity_flop_per_byte", 0) > ridge)
        if tensor_bound:
            roofline = {"bound": "tensor", "achieved": r["fp64_tflops"], "peak": fp64["dfma"], "unit": "TFLOP/s", "frac": r["fp64_frac"],
                        "peak_source": "FP64 datapath peak measured in this run by vx_bench_fp64 (DFMA %.1f, DMMA m8n8k4 %.1f TFLOP/s: one shared datapath); "
                                       "MEASURED_PEAKS.json carries no FP64 number" % (fp64["dfma"], fp64["dmma"])}
        else:
            roofline = {"bound": "hbm", "achieved": r["hbm_gbs"], "peak": peak_hbm, "unit": "GB/s", "frac": r["hbm_frac"], "peak_source": peak_src}
        roofline.update({"kernel": top, "what": r["what"], "ms_per_launch": r["ms_per_step"] / max(r["launches_per_step"], 1),
                         "launches_per_step": r["launches_per_step"], "share_of_kernel_time": r["share_of_kernel_time"],
                         "traffic": traffic_of(top), "traffic_unit": "bytes of DRAM traffic per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
                         "traffic_source": "profiles/ncu_traffic.json: mean over the launches of one ncu --set full capture (scripts/final_profile.sh); null when this kernel has no capture",
                         "ridge_flop_per_byte": ridge, "rows": rows,
                         "library_kernels_not_listed": "cub::DeviceRadixSort / DeviceScan launches of the map build are not bracketed by the library's profiler",
                         "latency_kernels_without_a_model": "k_ldlt_dnext (ONE CTA: the 32 x 32 pivot chain of the next diagonal block, ~45 kflop per launch) runs on a second "
                                                            "branch of the captured sweep beside k_ldlt_trailing; serialised under a profiler its share (about that of "
                                                            "k_ldlt_trailing) overstates it, in the replayed graph it is hidden; k_ldlt_panel / k_ldlt_back are "
                                                            "dependent 32-step chains as well (DESIGN section 5)",
                         "kernels_ms_per_step": {k: round(v["ms_per_step"], 3) for k, v in list(kern.items())[:16]}})
    extra = {"phases_ms_rank0": {k: round(v, 3) for k, v in phase.items()}, "voxel_frames_per_step": int(vf_step),
             "job": {"edges1": int(len(out["edges1"])), "edges2": int(len(out["edges2"])), "top_points": int(allsum(out["stats"]["top_points"])),
                     "vf_evals_bottom": int(allsum(out["stats"]["vf_evals_bottom"])), "vf_evals_top": int(allsum(out["stats"]["vf_evals_top"])),
                     "top_lm_iterations": out["stats"]["top_res_passes"], "e2e_equals_resident": same},
             "factors": {"bottom": rb, "top": rt}, "fp64_peak_tflops": fp64}
    if comm is not None:
        extra["comm"] = comm.stats(); extra["comm"]["nccl_version"] = int(pkg.load_library().vx_comm_nccl_version())
        # every rank's phases: the wait inside `gather` is the skew between the ranks' bottom + merge times, not transfer time
        allp = [None] * world
        dist.all_gather_object(allp, {k: round(v, 3) for k, v in phase.items()})
        extra["phases_ms_all_ranks"] = allp
    if fb is not None:
        fb.free()
    if ft is not None:
        ft.free()

    # ---- cpu baseline (rank 0, N = 1 only): the oracle on a bounded sample of the same bottom windows -----------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        cores = os.cpu_count() or 1
        items = cpu_windows(args, d, args.cpu_sample_windows or 40)
        cpu_pass(d, items[:max(2, len(items) // 4)], cores)           # warm-up
        vf_c, t_c, (t_b, t_t) = cpu_pass(d, items, cores)
        cpu = {"value": vf_c / t_c, "unit": UNIT, "cores": cores, "kind": "port", "sample": cpu_sample_text(args, len(items), cores) +
               "; %.1f s (bottom %.1f s, top %.1f s), %d voxel-frames through the oracle's acc_evaluate2" % (t_c, t_b, t_t, vf_c)}
    if rank == 0 and world == 1 and not args.no_extras:
        extra["other_rows"] = extra_benches(pkg, ctx, peak_hbm, fp64)
    if rank == 0:
        res = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
               "data": "synthetic", "config": cfg, "clocks": clk, "e2e": e2e, "gpu_launches": int(launches),
               "roofline": roofline, "cpu_baseline": cpu}
        res.update(extra)
        print(json.dumps(res))
    if comm is not None:
        comm.free()
    ctx.close()
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
