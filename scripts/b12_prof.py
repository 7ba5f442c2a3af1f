#!/usr/bin/env python
"""Times B1 (acc_evaluate2) and B2 (evaluate_only_residual) on the round-1 workload (synthetic batched factor, 399 windows x 1000
voxels, W = 10, occupancy 0.6): CUDA events on the library stream around `steps` back-to-back passes; prints cache checksums so
that kernel variants (environment switches) can be compared across processes.  usage: python scripts/b12_prof.py [steps] [b1|b2|both]"""
import hashlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import __graft_entry__ as ge
import synth
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 50
which = sys.argv[2] if len(sys.argv) > 2 else "both"
pkg = ge.load_package(); pkg.load_library()
fac = synth.batched_factor_csr(20241008 + 3, 399, 1000, 10, occupancy=0.6, pts_max=12)
nnz = int(fac["offsets"][-1]); V = int(fac["sig"].shape[0]); k = np.diff(fac["offsets"])
ctx = pkg.Context(0)
stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
f = pkg.Factor(ctx, 10, 399); f.upload_csr(fac); f.set_poses(fac["poses"])
f.residual_async(); ctx.sync()
def timed(fn):
    for _ in range(5): fn()
    ctx.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(steps): fn()
    b.record(stream); ctx.sync()
    return a.elapsed_time(b) / steps
def kern(fn):
    ctx.profile(True)
    for _ in range(10): fn()
    prof = ctx.profile_read(); ctx.profile(False)
    return {kk: round(v[0] / 10, 4) for kk, v in prof.items()}
peak = 6551.4
if which in ("b2", "both"):
    ms = timed(f.residual_async)
    b2 = 76 * nnz + 304 * V
    print("B2 pass %.4f ms  %.0f GB/s  frac %.3f  kernels %s" % (ms, b2 / ms / 1e6, b2 / ms / 1e6 / peak, kern(f.residual_async)))
    ev, evec, mrg = f.download_cache()
    h = hashlib.sha1(); [h.update(np.ascontiguousarray(a).tobytes()) for a in (ev, evec, mrg["P"], mrg["v"], mrg["N"])]
    print("B2 cache sha1", h.hexdigest(), "cost %.17g" % np.sum(f.residual(fac["poses"])))
if which in ("b1", "both"):
    ms = timed(f.hessian_async)
    fl = int(900 * k.sum() + 216 * np.sum(k * (k - 1) // 2))
    print("B1 pass %.4f ms  %.2f TFLOP/s  kernels %s" % (ms, fl / ms / 1e9, kern(f.hessian_async)))
    H, J, c = f.hessian(fac["poses"])
    h = hashlib.sha1(); h.update(np.ascontiguousarray(H).tobytes()); h.update(np.ascontiguousarray(J).tobytes())
    print("B1 H/J sha1", h.hexdigest(), "c %.17g" % np.sum(c))
