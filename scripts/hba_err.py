#!/usr/bin/env python
"""Prints the pose error of the batched bottom-level HBA against the oracle (tests/test_hba_gpu.py) for the current build."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import __graft_entry__ as ge
import scene, synth, oracle_bind as ob
from test_hba_gpu import LOCAL, GBA, pose_err
ge.build_oracle(); pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
clouds, (R, p) = scene.config1(W=25, pts_per_scan=5000)
W = 10; starts = [0, 5, 10, 15]; NW = 4
Rs, ps_ = synth.perturb_poses((R, p), 4, 0.004, 0.03)
fc = np.concatenate([np.arange(s, s + W) for s in starts]).astype(np.int32)
out = pkg.hba_add_edge(ctx, clouds, (Rs[fc], ps_[fc]), W, NW, LOCAL, GBA, 2, 1, frame_cloud=fc, ids=fc, mps=np.zeros(fc.size))
for b, s in enumerate(starts):
    ref = ob.hba_add_edge([clouds[i] for i in range(s, s + W)], (Rs[s:s + W], ps_[s:s + W]), LOCAL, GBA, 2, 1, ids=np.arange(s, s + W), mps=np.zeros(W))
    print("window %d pose_err %.3e" % (b, pose_err(out["poses"][b * W:(b + 1) * W], ref["poses"])))
