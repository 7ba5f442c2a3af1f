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
