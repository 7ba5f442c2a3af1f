"""C2h window schedule (voxelslam.cpp:3037-3139): vx_hba_schedule (host code in libvoxelba.so, no device needed) against the
literal walk of the reference's state machine in tests/oracle_global.py."""
import numpy as np
import pytest

import oracle_global as og


def lit(sizes, wd, mg):
    w = og.schedule_literal(list(sizes), wd, mg)
    return [(s, ids[0], len(ids)) for s, ids in w]


@pytest.mark.parametrize("sizes", [[2000], [35], [10], [7], [12], [14], [15], [20, 20], [20, 23, 9], [10, 10, 10], [3, 40], [11, 2, 26], [500] * 5])
@pytest.mark.parametrize("wd,mg", [(10, 5), (10, 3), (4, 4)])
def test_schedule_matches_literal(pkg, sizes, wd, mg):
    a, b, c = pkg.hba_schedule(sizes, wd, mg)
    got = list(zip(a.tolist(), b.tolist(), c.tolist()))
    assert got == lit(sizes, wd, mg)


def test_schedule_c4_counts(pkg):
    a, b, c = pkg.hba_schedule([2000])
    # 399 full windows (voxelslam.cpp:3137-3138: stride 5) + the queue left when the global pass fires (:3054-3058)
    assert a.size == 400 and np.all(c[:399] == 10) and c[399] == 5 and b[399] == 1995 and np.all(np.diff(b[:399]) == 5)
    a, b, c = pkg.hba_schedule([500] * 5)
    assert a.size == 5 * 99 + 1      # aligned sessions switch right after their last full window (:3127-3133); only the last leaves a tail


@pytest.mark.parametrize("sizes", [[2000], [35], [10], [20, 23, 9], [500] * 5, [11, 2, 26]])
@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
def test_window_partition_is_a_contiguous_cover_balanced_by_cost(pkg, sizes, world):
    # vx_hba_window_range: every window of the schedule on exactly one rank, runs contiguous and in rank order; a window shorter than
    # wdsize (a batch of its own in vx_hba_global) weighs 16 full ones, and no rank carries more than its share + one window
    ws, wst, wsz = pkg.hba_schedule(sizes)
    r = [pkg.hba_window_range(wsz, 10, k, world) for k in range(world)]
    assert r[0][0] == 0 or all(a == b for a, b in r[:1])
    at = 0
    for a, b in r:
        assert a <= b
        if b > a:
            assert a == at
            at = b
    assert at == ws.size
    wt = np.where(wsz == 10, 1, 16)
    share = wt.sum() / world
    for a, b in r:
        assert wt[a:b].sum() <= share + 16


def test_window_partition_c4(pkg):
    ws, wst, wsz = pkg.hba_schedule([2000])
    r = [pkg.hba_window_range(wsz, 10, k, 8) for k in range(8)]
    assert r[-1] == (363, 400) and r[0] == (0, 52)          # the rank with the tail window takes 36 full windows, the others 51-52
