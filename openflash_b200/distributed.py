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
    costs = np.asarray(costs, dtype=float)
    order = np.argsort(-costs, kind="stable")
    if len(costs) >= 256 * world:
        # many units: deal the cost-sorted list to the ranks in boustrophedon order (0..w-1, w-1..0, ...) - vectorised, and
        # within a fraction of a percent of the greedy partition when every rank gets hundreds of units
        pos = np.arange(len(order))
        lap, col = pos // world, pos % world
        rank_of = np.where(lap % 2 == 0, col, world - 1 - col)
        return [np.sort(order[rank_of == r]).astype(np.int64) for r in range(world)]
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


def geometry_cost(geo, n_freq: int) -> float:
    """Relative device cost of one geometry with n_freq frequencies: window elimination of the geometry stage
    (~25 N^3 per interior boundary) plus the per-frequency Schur stage (N^2 K + N K^2 + K^3 / 3)."""
    nmk = np.asarray(geo.NMK, dtype=float)
    N, K = nmk[-2], nmk[-1]
    geom = float(np.sum(25.0 * nmk[:-1] ** 3))
    return geom + n_freq * (N * N * K + N * K * K + K ** 3 / 3.0)


class ShardedPlan:
    """A batch spread over the ranks of a process group (one process per GPU): the product-level multi-GPU entry.

    Geometry-first sharding with cost-balanced parts when there are at least as many geometries as ranks, contiguous
    frequency chunks otherwise.  Every rank solves its shard with an ordinary ``BatchPlan``; the ONLY
    exchange is one all_gather of the packed output buffers (NCCL over NVLink on GPUs; gloo in the CPU tests, where
    ``plan_factory`` supplies a stand-in for the device plan).  Every rank can compute every other rank's output layout,
    so no sizes are exchanged.
    """

    def __init__(self, geometries, freqs, group=None, plan_factory=None, want_c=False, want_x=False, gather="all", root=0,
                 **plan_kw):
        import torch
        import torch.distributed as dist
        from .batch import BatchPlan, GeometryBatch, output_layout
        self.torch, self.dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.geometries = geometries if isinstance(geometries, GeometryBatch) else GeometryBatch.from_list(geometries)
        self.freqs = np.ascontiguousarray(np.asarray(freqs, dtype=np.float64).ravel())
        self.want_c, self.want_x = bool(want_c), bool(want_x)
        if gather not in ("all", "root"):
            raise ValueError("gather must be 'all' (every rank receives every result) or 'root' (only rank `root` does)")
        self.gather_mode, self.root = gather, int(root)
        gb = self.geometries
        G, F, W = len(gb), len(self.freqs), self.world
        self.mode, _ = plan_shards(G, F, W)
        if self.mode == "geometry":
            self.geom_parts = shard_by_cost(gb.costs(F), W)
            self.freq_parts = [(0, F)] * W
        else:
            self.geom_parts = [np.arange(G, dtype=np.int64)] * W
            self.freq_parts = shard_bounds(F, W)
        self.layouts = []
        for r in range(W):
            idx = self.geom_parts[r]
            Fr = self.freq_parts[r][1] - self.freq_parts[r][0]
            meta = (Fr, gb.R[idx], gb.NT[idx], gb.K[idx])
            self.layouts.append((meta, output_layout(*meta, self.want_c, self.want_x)))
        self.slot_bytes = max(lay[1] for _, lay in self.layouts)
        mine = self.geom_parts[self.rank]
        f0, f1 = self.freq_parts[self.rank]
        factory = plan_factory or BatchPlan
        my_geos = gb if (W == 1 or self.mode == "frequency") else gb.take(mine)
        self.local = factory(my_geos, self.freqs[f0:f1], want_c=self.want_c, want_x=self.want_x,
                             out_bytes_min=self.slot_bytes, **plan_kw) if (len(mine) and f1 > f0) else None
        dev = self.local.d_out.device if self.local is not None else torch.device("cpu")
        holds_all = W > 1 and (gather == "all" or self.rank == self.root)
        self.gathered = torch.empty(self.slot_bytes * W, dtype=torch.uint8, device=dev) if holds_all else None
        self.h_gathered = (torch.empty(self.slot_bytes * W, dtype=torch.uint8, pin_memory=dev.type == "cuda") if holds_all else None)
        self._empty = torch.zeros(self.slot_bytes, dtype=torch.uint8, device=dev) if self.local is None else None

    def upload(self):
        if self.local is not None:
            self.local.upload()

    def solve(self):
        """Enqueue this rank's kernels and the gather of every rank's packed outputs (asynchronous on GPUs)."""
        if self.local is not None:
            self.local.solve()
        if self.world > 1:
            src = self.local.d_out if self.local is not None else self._empty
            if self.gather_mode == "all":
                self.dist.all_gather_into_tensor(self.gathered, src, group=self.group)
            else:       # one receiver: 1 / world of the all-gather's traffic per link
                dst = self.dist.get_global_rank(self.group, self.root) if self.group is not None else self.root
                parts = list(self.gathered.view(self.world, self.slot_bytes).unbind(0)) if self.rank == self.root else None
                self.dist.gather(src, parts, dst=dst, group=self.group)

    def download_packed(self):
        """ONE device->host copy of every rank's packed outputs into pinned memory; returns the uint8 host buffer
        (rank r's slot at r * slot_bytes, laid out by ``layouts[r]``).  `download` turns it into per-geometry dicts."""
        if self.gathered is None and self.world > 1:
            return None                               # gather="root": only the root rank holds the results
        if self.world == 1:
            self.local.h_out.copy_(self.local.d_out, non_blocking=True)
            self.torch.cuda.current_stream(self.local.device).synchronize()
            return self.local.h_out.numpy()
        self.h_gathered.copy_(self.gathered, non_blocking=True)
        if self.gathered.is_cuda:
            self.torch.cuda.current_stream(self.gathered.device).synchronize()
        return self.h_gathered.numpy()

    def download_local_packed(self):
        """ONE device->host copy of THIS rank's packed outputs (its shard only) into its pinned buffer: with one process
        per GPU the shards reach host memory over all PCIe links in parallel, instead of every rank pulling all 8 shards
        through its own link.  Layout: ``layouts[rank]``; None for a rank without work."""
        if self.local is None:
            return None
        self.local.h_out.copy_(self.local.d_out, non_blocking=True)
        self.torch.cuda.current_stream(self.local.device).synchronize()
        return self.local.h_out.numpy()

    def download(self):
        """Per-geometry result dicts in the caller's geometry order, frequencies re-joined (every rank holds them all)."""
        from .batch import split_outputs, unpack_outputs
        if self.world == 1:
            return self.local.split(self.local.download())
        host = self.download_packed()
        if host is None:
            return None
        G, F = len(self.geometries), len(self.freqs)
        per_geom = [[None] * self.world for _ in range(G)]
        for r in range(self.world):
            (Fr, R, NT, K), (seg, _) = self.layouts[r]
            if Fr == 0 or len(R) == 0:
                continue
            flat = unpack_outputs(host[r * self.slot_bytes:(r + 1) * self.slot_bytes], seg, len(R), Fr, self.want_c, self.want_x)
            for gi, res in zip(self.geom_parts[r], split_outputs(flat, Fr, R, NT, K, self.want_c, self.want_x)):
                per_geom[int(gi)][r] = res
        out = []
        for g in range(G):
            parts = [x for x in per_geom[g] if x is not None]
            if len(parts) == 1:
                out.append(parts[0])
            else:       # frequency chunks of one geometry, in rank order
                merged = {k: np.concatenate([q[k] for q in parts], axis=0) for k in parts[0] if k != "c_vector"}
                if self.want_c:
                    merged["c_vector"] = parts[0]["c_vector"]
                out.append(merged)
        return out

    def download_stacked(self):
        """Design sweeps (geometry sharding, the same number of modes R for every geometry): the outputs of ALL geometries in
        the caller's order as arrays with a leading geometry axis - added_mass / damping [G, F, R, R], excitation_force /
        excitation_phase [G, F, R], m0 / status [G, F] - assembled with array indexing, no Python loop over geometries."""
        from .batch import unpack_outputs
        gb = self.geometries
        R = int(gb.R[0])
        if np.any(gb.R != R) or (self.world > 1 and self.mode != "geometry"):
            raise ValueError("download_stacked needs geometry sharding and the same number of modes for every geometry")
        if self.world == 1:
            return self.local.download_stacked(copy=True)
        host = self.download_packed()
        if host is None:
            return None
        G, F = len(gb), len(self.freqs)
        out = dict(added_mass=np.empty((G, F, R, R)), damping=np.empty((G, F, R, R)), excitation_force=np.empty((G, F, R)),
                   excitation_phase=np.empty((G, F, R)), m0=np.empty((G, F)), status=np.empty((G, F), dtype=np.int32))
        for r in range(self.world):
            (Fr, Rr, NT, K), (seg, _) = self.layouts[r]
            idx = self.geom_parts[r]
            if len(idx) == 0:
                continue
            flat = unpack_outputs(host[r * self.slot_bytes:(r + 1) * self.slot_bytes], seg, len(idx), Fr, False, False, copy=False)
            for k, v in out.items():
                v[idx] = flat[k].reshape((len(idx),) + v.shape[1:])
        return out

    def run(self):
        self.upload()
        self.solve()
        return self.download()


def solve_batch_distributed(geometries, freqs, group=None, **kw):
    """`solve_batch` over every GPU of the process group: shard, solve, one all_gather, results on every rank."""
    return ShardedPlan(geometries, freqs, group=group, **kw).run()
