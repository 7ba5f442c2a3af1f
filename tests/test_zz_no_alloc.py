"""vx_keys_static — the allocation-free variant of vx_keys (no counterpart in the reference)
— against the existing entry point, through the C ABI.  Integer work: the bar is bit-exact, as for vx_keys itself
(tests/test_voxelize_gpu.py::test_keys_bit_exact).  Two calls in a row share the static scratch: the second must not see the
first's contents.  Collected last (file name), so a failure here hides no other test under pytest -x."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_keys_static_matches_keys_two_calls_in_a_row(pkg, ctx):
    rng = np.random.Generator(np.random.PCG64(20241008 + 99))
    # call 1: doubles, the maximum size; call 2: fewer float points at another voxel size (stale keys of call 1 stay behind it)
    a = rng.uniform(-250.0, 250.0, (pkg.KEYS_STATIC_MAX_POINTS, 3))
    a[:64] = np.array([-1.0, -0.5, 0.0])                      # cell faces and the (-1, 0) band of the key rule
    b = rng.uniform(-40.0, 40.0, (70001, 3)).astype(np.float32)
    b[:8] = np.float32(-0.0)
    for xyz, vs in ((a, 0.5), (b, 1.0), (a[:1000], 0.25)):
        want = pkg.keys(ctx, xyz, vs)
        got = pkg.keys_static(ctx, xyz, vs)
        assert got.dtype == np.int64 and got.shape == want.shape
        assert np.array_equal(got, want)


def test_keys_static_rejects_oversize(pkg, ctx):
    xyz = np.zeros((pkg.KEYS_STATIC_MAX_POINTS + 1, 3))
    with pytest.raises(pkg.VxError) as e:
        pkg.keys_static(ctx, xyz, 1.0)
    assert e.value.code == -2 and "VX_KEYS_STATIC_MAX_POINTS" in str(e.value)
    # the failed call leaves the entry point usable
    ok = np.array([[0.3, -0.3, 7.9]])
    assert np.array_equal(pkg.keys_static(ctx, ok, 1.0), pkg.keys(ctx, ok, 1.0))
