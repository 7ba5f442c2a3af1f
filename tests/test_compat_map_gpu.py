"""The map side of the C++ adaptor (include/voxelba_compat_map.hpp) end to end: tests/cpp/compat_map.cpp holds the reference's own
call lines of the local-mapping loop (voxelslam.cpp:1905-1991: cut_voxel_multi, multi_recut, LI_BA_OptimizerGravity::damping_iter,
multi_margi), of match / lio_state_estimation (:1027, :962) and of HBA_add_edge's body (:2887-2899) plus HBA_add_edge itself;
linked against libvoxelba.so it must reproduce the oracle session fed with the same inputs."""
import os
import subprocess

import numpy as np
import pytest

import c3_scene
import scene
import synth

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOCAL = (1.0, 0.0025, [1.0] * 4)
GBA = (1.0, 0.01, [0.5] * 4)


def test_reference_call_lines_on_the_device_map(pkg, ob, tmp_path):
    W, thread_num = 10, 5
    c = c3_scene.build(pkg, W=W + 1, n_az=120)
    exe = str(tmp_path / "compat_map")
    libdir = os.path.join(ROOT, "voxel-slam_b200")
    r = subprocess.run(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "compat_map.cpp"), "-o", exe,
                        "-L", libdir, "-lvoxelba", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    # ---- the oracle session, collecting what the program needs
    om = ob.LocalMap(win_size=W, dept_err=0.01, beam_err=0.01, voxel_size=LOCAL[0], min_eigen_value=LOCAL[1], thre=LOCAL[2], max_layer=2, thread_num=thread_num)
    cov = ob.default_cov()
    start = c["start"][:W].copy(); imus = c["imus"][:W - 1].copy()
    Rs, ps = synth.states_poses(start)
    fin = str(tmp_path / "in.bin"); fout = str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(np.array([W, thread_num], dtype=np.int32).tobytes())
        f.write(np.array([LOCAL[0], LOCAL[1]] + LOCAL[2] + [GBA[0], GBA[1]] + GBA[2], dtype=np.float64).tobytes())
        f.write(np.ascontiguousarray(start).tobytes()); f.write(np.ascontiguousarray(imus).tobytes())
        for i in range(W):
            pv, pw = om.insert(c["scans"][i], (Rs[i:i + 1], ps[i:i + 1]), cov); om.recut()
            f.write(np.array([pv.shape[0]], dtype=np.int32).tobytes()); f.write(np.ascontiguousarray(pv).tobytes()); f.write(np.ascontiguousarray(pw).tobytes())
        of = om.factor(W)
        n_vox = of.num_voxels
        b = of.ba_li(start, imus, imu_coef=1e-4, gravity=True, max_iter=5)
        om.set_poses(synth.states_poses(b["states"]))
        assert om.margi() == 1
        # a fresh sweep for match() / lio_state_estimation against the map the session left
        q = c["scans"][W]
        pvq = ob.var_init(q, 0.01, 0.01)
        Rq, pq = synth.states_poses(c["truth"][W:W + 1])
        R0, p0 = synth.perturb_poses((Rq, pq), 3, 0.003, 0.02, fix_first=False)
        s0 = np.zeros(1, dtype=pkg.imu_state_dtype); s0["R"] = R0[0].T.reshape(9); s0["p"] = p0[0]
        f.write(np.array([pvq.shape[0]], dtype=np.int32).tobytes()); f.write(np.ascontiguousarray(pvq).tobytes()); f.write(s0.tobytes())
        f.write(np.ascontiguousarray(cov.T).tobytes())
        zero = np.zeros((15, 15))
        sub = np.arange(0, pvq.shape[0], 50)
        pvw = pvq[sub].copy()
        for k in range(pvw.shape[0]):                      # the program hands match() var_wld = R var R^T
            V = pvw[k, 3:].reshape(3, 3).T
            pvw[k, 3:] = (R0[0] @ V @ R0[0].T).T.reshape(9)
        lio = om.lio_state_estimation(pvq, R0[0], p0[0], cov)
        # ---- part B input: 8 keyframe clouds of config 1
        clouds, (Rk, pk) = scene.config1(W=8, pts_per_scan=6000)
        kstart = synth.perturb_poses((Rk, pk), 3, np.deg2rad(0.5), 0.05)
        kp = pkg.as_poses(kstart)
        f.write(np.array([8], dtype=np.int32).tobytes())
        for i in range(8):
            f.write(kp[i:i + 1].tobytes()); f.write(np.array([clouds[i].shape[0]], dtype=np.int32).tobytes()); f.write(np.ascontiguousarray(clouds[i], dtype=np.float32).tobytes())
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    o = np.fromfile(fout, dtype=np.float64)
    k = 0
    assert int(o[k]) == n_vox > 100; k += 1
    sd = pkg.imu_state_dtype
    ns = sd.itemsize // 8
    got = o[k:k + ns * W].view(sd); k += ns * W
    for name in ("R", "p", "v", "bg", "ba", "g"):
        assert np.abs(got[name] - b["states"][name]).max() <= 1e-8, name
    assert np.abs(o[k:k + 2] - b["resis"]).max() <= 1e-9 * np.abs(b["resis"]).max(); k += 2
    v6 = 1.0 / np.abs(np.diag(b["hess"][0:6, 15:21]))
    assert np.abs(o[k:k + 6] - v6).max() <= 1e-6 * np.abs(v6).max(); k += 6
    matched, sig_sum, ok = o[k], o[k + 1], o[k + 2]; k += 3
    st = o[k:k + ns].view(sd)[0]; k += ns
    assert matched > 50 and sig_sum > 0
    assert bool(ok) == lio["ok"]
    assert np.abs(st["R"].reshape(3, 3).T - lio["R"]).max() < 1e-9 and np.abs(st["p"] - lio["p"]).max() < 1e-9
    # ---- part B: OctreeGBA::cut_voxel + OctreeGBA_multi_recut + Lidar_BA_Optimizer::damping_iter(up = 4)
    og = ob.gba_build(clouds, kstart, GBA[0], GBA[1], GBA[2], 2)
    rb = og.ba_lidar(kstart, max_iter=4, thd_num=2)
    assert int(o[k]) == og.num_voxels; k += 1
    assert np.abs(o[k:k + 2] - rb["resis"]).max() <= 1e-9 * np.abs(rb["resis"]).max(); k += 2
    assert bool(o[k]) == (rb["rc"] == 1); k += 1
    pg = o[k:k + 12 * 8].reshape(8, 12); k += 96
    assert np.abs(pg[:, :9] - rb["poses"]["R"]).max() < 1e-8 and np.abs(pg[:, 9:] - rb["poses"]["p"]).max() < 1e-8
    # ---- HBA_add_edge(max_iter 1, thread_num 2, plptr)
    gba_hba = (GBA[0], GBA[1], GBA[2])
    rh = ob.hba_add_edge(clouds, kstart, LOCAL, gba_hba, 2, 1, thread_num=2, ids=np.arange(100, 108), mps=np.zeros(8), want_cloud=True)
    ne, nc = int(o[k]), int(o[k + 1]); k += 2
    assert ne == rh["n_edges"] and nc == rh["merged"].shape[0] > 1000
    e = o[k:k + 11 * ne].reshape(ne, 11); k += 11 * ne
    assert np.array_equal(e[:, 0].astype(int), rh["edges"]["id1"]) and np.array_equal(e[:, 1].astype(int), rh["edges"]["id2"])
    assert np.abs(e[:, 2:5] - rh["edges"]["tra"]).max() < 1e-8
    assert (np.abs(e[:, 5:] - rh["edges"]["cov"]) / np.abs(rh["edges"]["cov"])).max() < 1e-6
    ph = o[k:k + 96].reshape(8, 12)
    assert np.abs(ph[:, :9] - rh["poses"]["R"]).max() < 1e-8 and np.abs(ph[:, 9:] - rh["poses"]["p"]).max() < 1e-8
