"""The C++ adaptor (include/voxelba_compat.hpp) end to end: reference-style code (tests/cpp/compat_smoke.cpp) linked against
libvoxelba.so reproduces the committed fixture tests/golden/factor.npz."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def test_reference_style_calls_through_the_adaptor(pkg, tmp_path):
    z = np.load(os.path.join(ROOT, "tests", "golden", "factor.npz"))
    W = int(z["W"]); V = z["sig"].shape[0]
    exe = str(tmp_path / "compat_smoke")
    libdir = os.path.join(ROOT, "voxel-slam_b200")
    r = subprocess.run(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "compat_smoke.cpp"), "-o", exe,
                        "-L", libdir, "-lvoxelba", "-Wl,-rpath," + libdir], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as f:
        f.write(np.array([V, W], dtype=np.int32).tobytes())
        for name in ("plvec", "sig", "coe", "eigval", "eigvec", "merged"):
            f.write(np.ascontiguousarray(z[name]).tobytes())
        f.write(pkg.as_poses((z["poses_R"], z["poses_p"])).tobytes())
        f.write(pkg.as_poses((z["start_R"], z["start_p"])).tobytes())
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True)
    assert r.returncode == 0, (r.stdout, r.stderr)
    o = np.fromfile(fout, dtype=np.float64)
    jl = 6 * W
    k = 0
    residual, cost = o[0], o[1]; k = 2
    H = o[k:k + jl * jl].reshape(jl, jl).T; k += jl * jl
    J = o[k:k + jl]; k += jl
    ev0 = o[k:k + 3]; k += 3
    assert abs(residual - float(z["residual"])) <= 1e-9 * abs(float(z["residual"]))
    assert abs(cost - float(z["cost"])) <= 1e-9 * abs(float(z["cost"])) and rel(H, z["H"]) < 1e-9 and rel(J, z["J"]) < 1e-9
    scale = np.abs(z["cache_merged"]["P"][0]).max() / z["cache_merged"]["N"][0]
    assert np.abs(ev0 - z["cache_eigval"][0]).max() <= 1e-12 * scale
    ps = o[k:k + 12 * W].reshape(W, 12); k += 12 * W
    resis = o[k:k + 2]; k += 2
    conv = o[k]; k += 1
    hess = o[k:k + jl * jl].reshape(jl, jl).T
    assert np.abs(ps[:, :9] - z["ba_poses"]["R"]).max() < 1e-8 and np.abs(ps[:, 9:] - z["ba_poses"]["p"]).max() < 1e-8
    assert rel(resis, z["ba_resis"]) < 1e-9 and conv == float(int(z["ba_rc"]) == 1) and rel(hess, z["ba_hess"]) < 1e-9
