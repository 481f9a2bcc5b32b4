"""Multi-GPU sharding of a batch: one process per GPU, independent shards, one gather at the end.

Every (geometry, frequency) system is independent, so there is no data-path collective; the only
inter-GPU traffic is the final gather of the coefficient arrays (NCCL all_gather over NVLink on
GPUs; the same code runs on gloo/CPU tensors for the host-logic tests).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def shard_bounds(n_units: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous shards whose sizes differ by at most one unit."""
    base, rem = divmod(n_units, world)
    out, start = [], 0
    for r in range(world):
        n = base + (1 if r < rem else 0)
        out.append((start, start + n))
        start += n
    return out


def shard_by_cost(costs: Sequence[float], world: int) -> List[np.ndarray]:
    """Greedy longest-processing-time partition: unit indices per rank with balanced total cost
    (heterogeneous geometry sweeps: cost ~ p^3 + F K^3)."""
    order = np.argsort(-np.asarray(costs, dtype=float), kind="stable")
    loads = np.zeros(world)
    parts: List[List[int]] = [[] for _ in range(world)]
    for u in order:
        r = int(np.argmin(loads))
        parts[r].append(int(u))
        loads[r] += costs[u]
    return [np.array(sorted(p), dtype=np.int64) for p in parts]


def plan_shards(n_geom: int, n_freq: int, world: int):
    """Shard by geometry when there are at least `world` of them, else by contiguous frequency chunks
    (each rank then repeats the small per-geometry stage) — SURVEY §8(e)."""
    if n_geom >= world:
        return "geometry", shard_bounds(n_geom, world)
    return "frequency", shard_bounds(n_freq, world)


def gather_padded(local, group=None):
    """all_gather of 1-D tensors whose length may differ by rank; returns the list of per-rank tensors."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    n = torch.tensor([local.numel()], dtype=torch.int64, device=local.device)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    mx = max(sizes)
    buf = torch.zeros(mx, dtype=local.dtype, device=local.device)
    buf[:local.numel()] = local
    out = torch.empty(mx * world, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    return [out[r * mx:r * mx + sizes[r]] for r in range(world)]
