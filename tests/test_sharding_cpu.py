"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: block distribution of windows / voxels and the
all-reduce of [Hess | JacT | cost].  The per-rank evaluation is the oracle's acc_evaluate2 on
[head, end) — the reference's own partial-sum interface (voxel_map.hpp:150, :376-384)."""
import os
import socket

import numpy as np
import pytest

import synth


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, q):
    import sys
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    import oracle_bind as ob
    import importlib.util
    spec = importlib.util.spec_from_file_location("vx_sharding", os.path.join(ge.PKG_DIR, "sharding.py"))
    sh = importlib.util.module_from_spec(spec); spec.loader.exec_module(sh)
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    V, W = 501, 6
    fac = synth.random_factor(77, V, W)
    of = ob.Factor(W); of.set_dense(fac)
    head, end = sh.voxel_shard(V, rank, world)
    H, J, c = of.hessian(fac["poses"], head, end)
    ts = [torch.from_numpy(np.ascontiguousarray(H)), torch.from_numpy(J), torch.tensor([c], dtype=torch.float64)]
    sh.allreduce_sum_(ts)
    Hf, Jf, cf = of.hessian(fac["poses"])
    ok = (np.max(np.abs(ts[0].numpy() - Hf)) <= 1e-12 * np.max(np.abs(Hf)) and np.max(np.abs(ts[1].numpy() - Jf)) <= 1e-12 * np.max(np.abs(Jf))
          and abs(ts[2].item() - cf) <= 1e-12 * abs(cf))
    # every rank ends with the same bits (replicated LM solve stays in lock step)
    g = [torch.zeros_like(ts[0]) for _ in range(world)]
    dist.all_gather(g, ts[0])
    same = all(torch.equal(g[0], x) for x in g)
    q.put((rank, bool(ok), bool(same), (head, end)))
    dist.destroy_process_group()


def test_block_range_covers_everything():
    import importlib.util
    import __graft_entry__ as ge
    spec = importlib.util.spec_from_file_location("vx_sharding", os.path.join(ge.PKG_DIR, "sharding.py"))
    sh = importlib.util.module_from_spec(spec); spec.loader.exec_module(sh)
    for n in (0, 1, 7, 399, 2000):
        for world in (1, 2, 4, 8):
            r = [sh.block_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    assert sh.window_shard(399, 3, 8) == sh.block_range(399, 3, 8)


@pytest.mark.timeout(300)
def test_voxel_sharded_hessian_allreduce_gloo(ob):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=240) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    assert all(r[1] for r in res), res
    assert all(r[2] for r in res), res
    rng = sorted(r[3] for r in res)
    assert rng[0][0] == 0 and rng[0][1] == rng[1][0] and rng[1][1] == 501
