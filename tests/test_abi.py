"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/*.h declares, the POD layouts match, and compute calls fail loudly without a GPU."""
import ctypes as C
import glob
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = []
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names += re.findall(r"\b(vx_[a-z0-9_]+)\s*\(", src)
    return sorted(set(names))


def test_library_exports_every_declared_symbol(pkg):
    lib = pkg.load_library()
    syms = declared_symbols()
    assert len(syms) >= 15
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing


def test_pod_layouts(pkg):
    assert pkg.cluster_dtype.itemsize == 104          # tools.hpp:304-310 {Matrix3d, Vector3d, int}
    assert pkg.cluster_dtype.fields["N"][1] == 96
    assert pkg.pose_dtype.itemsize == 96


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.VxError) as ei:
        pkg.Context(0)
    assert ei.value.code == -1
    assert "no CPU fallback" in str(ei.value)


def test_product_does_not_touch_oracle():
    # the product path must never import / link / include anything under oracle/
    for path in glob.glob(os.path.join(ROOT, "voxel-slam_b200", "**", "*"), recursive=True):
        if os.path.isfile(path) and path.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
            assert "oracle" not in open(path, errors="replace").read().replace("no oracle", ""), path
