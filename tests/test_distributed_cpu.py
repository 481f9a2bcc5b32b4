"""world_size-2 gloo tests of the N>1 host logic: sharding and the final gather of coefficient arrays."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from openflash_b200.distributed import gather_padded, plan_shards, shard_bounds, shard_by_cost


def test_shard_bounds_cover_everything():
    for n, w in [(4096, 8), (10, 3), (7, 8), (100000, 8)]:
        b = shard_bounds(n, w)
        assert b[0][0] == 0 and b[-1][1] == n
        assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        sizes = [e - s for s, e in b]
        assert max(sizes) - min(sizes) <= 1
    assert plan_shards(1, 4096, 8)[0] == "frequency"
    assert plan_shards(100000, 64, 8)[0] == "geometry"


def test_cost_balanced_sharding():
    rng = np.random.default_rng(0)
    M = rng.integers(2, 9, 5000)
    cost = (32.0 * M - 16) ** 3
    parts = shard_by_cost(cost, 8)
    assert sorted(np.concatenate(parts).tolist()) == list(range(5000))
    loads = np.array([cost[p].sum() for p in parts])
    assert loads.max() / loads.mean() < 1.01


def _worker(rank, world, port, tmp):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    F = 37
    full = torch.arange(F * 3, dtype=torch.float64)       # pretend [F, 3] coefficient rows
    (s, e) = shard_bounds(F, world)[rank]
    local = full.view(F, 3)[s:e].reshape(-1).clone()
    parts = gather_padded(local)
    got = torch.cat(parts)
    ok = torch.equal(got, full)
    t = torch.tensor([1.0 + rank])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)               # max-over-ranks timing reduction used by bench.py
    ok = ok and t.item() == float(world)
    with open(os.path.join(tmp, f"ok{rank}"), "w") as f:
        f.write("1" if ok else "0")
    dist.destroy_process_group()


def test_gather_world2_gloo(tmp_path):
    world = 2
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        assert open(tmp_path / f"ok{r}").read() == "1"
