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


def _route_worker(rank, world, port, q):
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
    # every rank holds a different slice of a 4-frame point set; owner = (a hash of the root cell) mod world, cells from the
    # oracle's key rule (the device uses mix64 of the packed key; any deterministic function of the CELL gives the invariant)
    F, n = 4, 3000
    rng = np.random.Generator(np.random.PCG64(1234))
    all_xyz = rng.uniform(-20, 20, (world * n, 3)).astype(np.float32)
    all_frame = rng.integers(0, F, world * n).astype(np.int32)
    xyz = all_xyz[rank * n:(rank + 1) * n]; frame = all_frame[rank * n:(rank + 1) * n]
    cells = ob.keys_f32(xyz, 1.0)
    own = ((cells[:, 0] * 73856093) ^ (cells[:, 1] * 19349663) ^ (cells[:, 2] * 83492791)) % world
    ids = np.arange(rank * n, (rank + 1) * n, dtype=np.int64)
    (rx, rf, rid), rc = sh.route_by_owner([torch.from_numpy(xyz), torch.from_numpy(frame), torch.from_numpy(ids)], torch.from_numpy(own.astype(np.int32)), world)
    # 1. everything I received is mine, 2. arrival order is (source rank, original order)
    rcells = ob.keys_f32(rx.numpy(), 1.0)
    mine = ((rcells[:, 0] * 73856093) ^ (rcells[:, 1] * 19349663) ^ (rcells[:, 2] * 83492791)) % world
    ok_owner = bool((mine == rank).all())
    ok_order = bool((np.diff(rid.numpy()) > 0).all())
    ok_payload = bool(np.array_equal(rx.numpy(), all_xyz[rid.numpy()]) and np.array_equal(rf.numpy(), all_frame[rid.numpy()]))
    sx, off = sh.clouds_by_frame(rx, rf, F)
    ok_frames = bool(off[-1] == rx.shape[0] and all(np.array_equal(sx.numpy()[off[g]:off[g + 1]], rx.numpy()[rf.numpy() == g]) for g in range(F)))
    # 3. nothing lost: the counts over ranks add up
    tot = torch.tensor([rx.shape[0]], dtype=torch.int64)
    dist.all_reduce(tot)
    q.put((rank, ok_owner, ok_order, ok_payload, ok_frames, int(tot.item()) == world * n, int(rc.sum())))
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_route_points_to_root_voxel_owner_gloo(ob):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_route_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=240) for _ in ps]
    for p in ps:
        p.join(timeout=60)
    assert sorted(r[0] for r in res) == [0, 1]
    for r in res:
        assert all(r[1:6]), r
    assert sum(r[6] for r in res) == 2 * 3000
