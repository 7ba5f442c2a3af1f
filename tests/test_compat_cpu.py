"""include/voxelba_compat.hpp: the reference-style call sites of tests/cpp/compat_smoke.cpp must compile against the adaptor
(no GPU, no linking: syntax + type check only)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_compat_header_compiles():
    r = subprocess.run(["g++", "-std=c++14", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "cpp", "compat_smoke.cpp")],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
