"""Host-side logic of the N>1 path on CPU: world_size-2 gloo processes."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from aircraft_b200.sharding import _orderable, gather_varlen, global_argmin, local_min_or_sentinel, shard_range


def test_shard_range_partitions():
    for total in (0, 1, 7, 8, 1000003, 409600):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(total, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for a, b in zip(spans, spans[1:]):
                assert a[1] == b[0]
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(10, 2, 2)


def test_orderable_key_sorts_like_float():
    v = torch.tensor([-float("inf"), -3.5, -0.0, 0.0, 1e-30, 2.0, float("inf"), float("nan")])
    k = _orderable(v)
    assert torch.all(k[:-1][1:] >= k[:-1][:-1])
    assert k[-1] == 0xFFFFFFFF


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(5)
        cost = torch.rand(total, generator=g)          # same on every rank
        cost[total // 2 + 3] = cost.min() - 1.0        # unique winner in rank 1's shard
        a, b = shard_range(total, rank, world)
        loc = cost[a:b]
        lmin, lidx = loc.min(), loc.argmin()
        gmin, gidx, owner = global_argmin(lmin, lidx + a)
        allc = gather_varlen(loc, total, rank, world)
        # tie: identical minimum in both shards -> lower global index wins
        tie = torch.ones(total); tie[1] = -2.0; tie[total - 1] = -2.0
        tl = tie[a:b]
        tmin, tidx, towner = global_argmin(tl.min(), tl.argmin() + a)
        q.put((rank, float(gmin), gidx, owner, bool(torch.equal(allc, cost)), tidx, towner,
               float(cost.min()), int(cost.argmin())))
    finally:
        dist.destroy_process_group()


def test_global_argmin_and_gather_gloo_world2():
    world, total = 2, 1001
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    for rank, gmin, gidx, owner, gathered_ok, tidx, towner, ref_min, ref_idx in res:
        assert gidx == ref_idx and abs(gmin - ref_min) < 1e-7 and owner == 1
        assert gathered_ok
        assert tidx == 1 and towner == 0


class _TorchArgmin:
    """stands in for Aircraft.argmin (the kernel refuses B = 0, as airc_argmin does)"""

    def argmin(self, cost):
        assert cost.numel() > 0, "AIRC_EINVAL"
        c = torch.where(torch.isnan(cost), torch.full_like(cost, float("inf")), cost)
        return c.min().reshape(1), c.argmin().reshape(1)


def _worker_empty(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        total = 1                                        # fewer rollouts than ranks: rank 1's shard is empty
        cost = torch.tensor([3.25])
        a, b = shard_range(total, rank, world)
        lmin, lg = local_min_or_sentinel(_TorchArgmin(), cost[a:b], a)
        gmin, gidx, owner = global_argmin(lmin, lg)
        q.put((rank, b - a, float(gmin), gidx, owner))
    finally:
        dist.destroy_process_group()


def test_empty_shard_does_not_hang_or_win_gloo_world2():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_empty, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    assert [r[1] for r in res] == [1, 0]
    for _, _, gmin, gidx, owner in res:
        assert gmin == 3.25 and gidx == 0 and owner == 0
