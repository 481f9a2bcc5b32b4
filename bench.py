#!/usr/bin/env python
"""bench.py — MEEM systems solved per second on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--algo schur|dense] [--impl reference]

One "step" = one pass of the hot path over one batch: roots -> per-geometry stage -> per-(geometry,
frequency) solves -> hydrodynamic coefficients (+ for N > 1 the NCCL all-gather of the packed output
buffers, the path's only exchange).  The batch goes through the product's multi-GPU entry
(openflash_b200.distributed.ShardedPlan); the headline workload gives every rank the same-size shard
(weak scaling); `value` is whole-job systems/s with inputs resident in HBM; `e2e` times the public
host-buffer call per step (ShardedPlan.upload + solve + download: H2D of the packed inputs from pinned
memory, kernels, gather, ONE D2H of every output into pinned memory).  `extra.sweep_1e6` is the north
star's 10^6-system design sweep as a FIXED total spread over the N GPUs (strong scaling), `extra.workloads`
the other BASELINE.json configurations at N = 1.  `--impl reference` times the UNTOUCHED reference
(oracle/_ref, staged by oracle/build_ref.py; the numpy/scipy oracle port when it is absent) through its
public API in a pool of one process per host core on a bounded sample of the same workload, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MEEM systems solved/sec (hydro coeffs)"
UNIT = "systems/s"


# ------------------------------------------------------------------------------------------ workloads
def c4_geometries(n, seed=20260101, nmk=16):
    """SURVEY §8(d) C4 generator: random N-cylinder geometries, M in 2..8, h = 50."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        M = int(rng.integers(2, 9))
        h = 50.0
        a = np.cumsum(rng.uniform(1.0, 4.0, M))
        while True:
            d = np.sort(rng.uniform(0.04, 0.6, M) * h)[::-1]
            if np.min(-np.diff(d)) >= 0.01 * h:
                break
        if rng.uniform() < 0.1:
            d = rng.permutation(d)
        out.append(dict(h=h, a=a, d=d, NMK=[nmk] * (M + 1), body_map=[0] * M, modes=[0]))
    return out


def workload(name, rank, world):
    """Returns (geometry dicts, omega grid of THIS rank, description)."""
    if name == "big_tricylinder":       # BASELINE.json configs[2] / SURVEY C3
        F = 4096
        om = np.linspace(0.1, 3.0, F * world)[rank * F:(rank + 1) * F]
        g = dict(h=25.0, a=[10, 15, 20], d=[20, 15, 10], NMK=[200] * 4, body_map=[0, 0, 0], modes=[0])
        return [g], om, "big_tricylinder h=25 a=[10,15,20] d=[20,15,10], 200 modes/region (N_T=1200), 4096 omega per GPU in [0.1,3], heave R=1"
    if name == "small_tricylinder":     # configs[1] / C2
        F = 256
        om = np.linspace(0.1, 3.0, F * world)[rank * F:(rank + 1) * F]
        g = dict(h=20.0, a=[5, 10, 15], d=[15, 10, 5], NMK=[50] * 4, body_map=[0, 1, 1], modes=[0, 1])
        return [g], om, "small_tricylinder h=20 a=[5,10,15] d=[15,10,5], 50 modes/region (N_T=300), 256 omega per GPU, two heave bodies R=2"
    if name == "mini_bicylinder":       # configs[0] / C1 as a replica sweep
        F = 65536
        om = np.linspace(0.1, 3.0, F * world)[rank * F:(rank + 1) * F]
        g = dict(h=1.001, a=[0.125, 0.25], d=[0.25, 0.125], NMK=[10] * 3, body_map=[0, 0], modes=[0])
        return [g], om, "mini_bicylinder 10 modes/region (N_T=40), 65536 omega per GPU"
    if name == "geometry_sweep":        # configs[3] / C4: 12 500 geometries x 64 omega per GPU
        n = 12500
        gs = c4_geometries(n * world)[rank * n:(rank + 1) * n]
        return gs, np.linspace(0.3, 2.5, 64), "random N-cylinder (M=2..8, 16 modes/region) design sweep, 12500 geometries x 64 omega per GPU"
    if name == "convergence_sweep":     # configs[4] / C5: mode-count convergence, 12 mode counts x 2 geometries x 8 omega
        ns = [10, 15, 20, 30, 40, 50, 75, 100, 150, 200, 300, 400]
        gs = [dict(h=1.001, a=[0.5, 1.0], d=[0.5, 0.25], NMK=[n] * 3, body_map=[0, 0], modes=[0]) for n in ns]
        gs += [dict(h=20.0, a=[5.0, 10.0, 15.0], d=[15.0, 10.0, 5.0], NMK=[n] * 4, body_map=[0, 0, 0], modes=[0]) for n in ns]
        gs = gs[rank::world] if world > 1 else gs
        return gs, np.linspace(0.5, 3.0, 8), ("mode-count convergence sweep 10..400 modes/region on small_bicylinder and small_tricylinder, 8 omega "
                                               "(the reference's exp() overflows at the largest mode counts: those systems are NaN there and here)")
    if name == "sweep_1e6":             # north star: 10^6 systems = first 15 625 C4 geometries x 64 omega, FIXED total
        return c4_geometries(15625), np.linspace(0.3, 2.5, 64), "first 15625 C4 geometries x 64 omega = 10^6 systems in total (strong scaling)"
    raise SystemExit(f"unknown workload {name}")


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [x.strip() for x in r.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0]))
                mx = float(p[1])
            except ValueError:
                continue
            for nme, val in zip(names, p[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ CPU arm
def global_workload(name, world):
    """The whole job of `world` GPUs: (geometry dicts, omega grid, description, scaling)."""
    if name == "sweep_1e6":
        gs, om, desc = workload(name, 0, 1)
        return gs, om, desc, "strong"
    parts = [workload(name, r, world) for r in range(world)]
    desc = parts[0][2]
    if len(parts[0][0]) == 1:           # one geometry: the ranks' omega chunks, concatenated
        return parts[0][0], np.concatenate([p[1] for p in parts]), desc, "weak"
    return [g for p in parts for g in p[0]], parts[0][1], desc, "weak"


def cpu_sample_tasks(kind, geoms, omegas, target_systems, cores):
    """A bounded sample of the workload: (geometry, omega-chunk) tasks covering ~target_systems systems.  One geometry:
    the sampled omegas are dealt to one task per core, so the per-geometry set-up (the reference's build_problem_cache)
    is paid once per worker as in a real omega-chunked sweep."""
    tasks = []
    if len(geoms) == 1:
        idx = np.linspace(0, len(omegas) - 1, min(target_systems, len(omegas))).astype(int)
        nt = max(1, min(cores, len(idx)))
        for t in range(nt):
            tasks.append((kind, geoms[0], [float(omegas[i]) for i in idx[t::nt]]))
        n_sys = len(idx)
    else:
        ng = max(1, min(len(geoms), target_systems // len(omegas)))
        for g in geoms[:ng]:
            tasks.append((kind, g, [float(w) for w in omegas]))
        n_sys = ng * len(omegas)
    return tasks, n_sys


def run_cpu(kind, geoms, omegas, target_systems, pool, cores, min_seconds=0.0, max_seconds=30.0):
    """Time the CPU implementation on a sample of the workload.  With min_seconds > 0 the sample is repeated (same
    tasks) until that much wall time has been spent, so that the baseline rests on 10-30 s of CPU work."""
    from oracle.ref_runner import cpu_chunk
    tasks, n_sys = cpu_sample_tasks(kind, geoms, omegas, target_systems, cores)
    t0 = time.perf_counter()
    done = 0
    while True:
        list(pool.map(cpu_chunk, tasks, chunksize=1))
        done += n_sys
        el = time.perf_counter() - t0
        if el >= min_seconds or el * 2 > max_seconds:
            break
    return done, time.perf_counter() - t0


def cpu_target(name, kind):
    """Systems per timed pass.  The reference rebuilds its per-geometry cache in every task (meem_engine.py:100-305:
    0.3 s for big_tricylinder), exactly as a user's run_and_store_results call does."""
    t = {"big_tricylinder": 512, "small_tricylinder": 256, "mini_bicylinder": 2048, "geometry_sweep": 64, "sweep_1e6": 64,
         "convergence_sweep": 48}[name]
    return t if kind == "port" else max(t // 4, 32)


def cpu_kind():
    from oracle import ref_runner
    return "reference" if ref_runner.reference_available() else "port"


def cpu_description(kind, cores):
    what = ("the untouched reference package (oracle/_ref: MEEMEngine.run_and_store_results, numpy/scipy)" if kind == "reference"
            else "oracle port (numpy/scipy restatement of the reference path; oracle/_ref absent)")
    return f"{what}, {cores} worker processes x 1 BLAS thread"


# ------------------------------------------------------------------------------------------ GPU arm helpers
def band_flops(g):
    """Closed-form real FP64 flops of the window elimination of one geometry (meem_geom_band.cuh): per window the
    partial-pivot elimination of n0 columns of an nrows x ncols window, plus the back-substitution of the last window."""
    nmk = [int(x) for x in g.NMK]
    M, R = len(nmk) - 1, int(g.R)
    nun = [nmk[0]] + [2 * nmk[i] for i in range(1, M)]
    fl, nc = 0.0, 0
    for bd in range(M):
        last = bd == M - 1
        n0 = nun[bd]
        n1 = nmk[M - 1] if last else nun[bd + 1]
        nsearch = nc + (nmk[bd] if last else nmk[bd] + nmk[bd + 1])
        nrows, ncols = nsearch + R, n0 + n1 + R
        k = np.arange(n0, dtype=float)
        fl += float(np.sum(2.0 * (nrows - 1 - k) * (ncols - 1 - k)))
        if last:
            fl += float(n0) * n0 * (n1 + R)
        nc = nsearch - n0
    return fl


def coop2_flops(g):
    """Closed-form real FP64 flops of the static-pivot window elimination (k_geom_coop2): per window one rank-N_bd product
    for the C1 block, the partial-pivot elimination of the C2 block, and the back-substitution of the last window."""
    nmk = [int(x) for x in g.NMK]
    M, R = len(nmk) - 1, int(g.R)
    fl, nc = 0.0, 0
    for bd in range(M):
        last = bd == M - 1
        Nb, n2 = nmk[bd], (nmk[bd] if bd > 0 else 0)
        n1 = nmk[M - 1] if last else 2 * nmk[bd + 1]
        hv = 0 if last else nmk[bd + 1]
        nsub, ncs = nc + hv + R, n2 + n1 + R
        fl += 2.0 * nsub * Nb * ncs
        k = np.arange(n2, dtype=float)
        fl += float(np.sum(2.0 * (nsub - 1 - k) * (ncs - 1 - k)))
        if last:
            fl += float(n2) * n2 * (n1 + R)
        nc = hv
    return fl


def kernel_flops(flat, F_, algo, factorised, geom_kernel):
    """Closed-form real FP64 flops per kernel family for one step over geometries `flat` x F_ frequencies."""
    kf = {}
    for g in flat:
        N, K, R, p = float(g.NMK[-2]), float(g.NMK[-1]), float(g.R), float(g.size - g.NMK[-1])
        NT = float(g.size)
        if algo == "dense":
            d = {"frequency": F_ * (8 / 3 * NT ** 3 + 8 * NT * NT * R)}
        else:
            geom = (band_flops(g) if geom_kernel.startswith("k_geom_band") else coop2_flops(g) if geom_kernel == "k_geom_coop2"
                    else 2 / 3 * p ** 3 + 2 * p * p * (N + R))
            if algo == "schur":
                gemm = (N * N * K + 4 * R * N * K + N * K * K) if factorised else (2 * N * N * K + N * K * K)
                d = {"geometry": geom + (N ** 3 / 3 if factorised else 0.0), "freq_gemm": F_ * gemm,
                     "freq_solve": F_ * (K ** 3 / 3 + 2 * K * K * R + 2 * K * R * R)}
            else:
                d = {"geometry": geom, "frequency": F_ * (2 * N * N * K + 2 * N * K * K + 2 / 3 * K ** 3 + 2 * K * K * (1 + 2 * R) + 4 * N * K * R)}
        for k, v in d.items():
            kf[k] = kf.get(k, 0.0) + v
    return kf


class Runner:
    """One workload on the process group through the product entry (ShardedPlan): timed steps, e2e, stage times."""

    def __init__(self, ofb, name, world, rank, dev, algo, flush, gather="all"):
        import torch
        from openflash_b200.distributed import ShardedPlan
        self.torch, self.ofb, self.name, self.world, self.rank, self.dev, self.algo, self.flush = torch, ofb, name, world, rank, dev, algo, flush
        geoms, omegas, self.desc, self.scaling = global_workload(name, world)
        self.geoms, self.omegas = geoms, omegas
        t0 = time.perf_counter()
        batch = ofb.GeometryBatch.from_dicts(geoms)          # packed arrays (no Python object per geometry)
        algo_id = {"schur": ofb.ALGO_SCHUR, "schur_lu": ofb.ALGO_SCHUR_LU, "dense": ofb.ALGO_DENSE}[algo]
        self.sp = ShardedPlan(batch, omegas, algo=algo_id, device=dev.index, gather=gather)
        self.plan_ms = 1e3 * (time.perf_counter() - t0)
        self.flat = [SimpleNamespace(NMK=g["NMK"], R=len(g["modes"]), size=int(g["NMK"][0] + g["NMK"][-1] + 2 * sum(g["NMK"][1:-1])))
                     for g in geoms]
        self.total_sys = len(geoms) * len(omegas)

    def step(self):
        self.sp.solve()

    def timed(self, steps, warmup):
        torch, dist = self.torch, self.torch.distributed
        self.sp.upload()
        for _ in range(warmup):
            self.step()
        torch.cuda.synchronize()
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        stage = []
        for s in range(steps):
            self.flush.fill_(s & 0xFF)
            ev[s][0].record()
            self.step()
            ev[s][1].record()
            torch.cuda.synchronize()
            if self.sp.local is not None:
                stage.append(self.sp.local.stage_ms())
        if self.world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        self.stage = stage
        return sum(a.elapsed_time(b) for a, b in ev)

    def e2e(self, steps):
        """H2D of the packed inputs + kernels + gather + ONE D2H of every output, per step, wall clock with syncs."""
        torch = self.torch
        ms = 0.0
        for s in range(steps):
            self.flush.fill_(s)
            torch.cuda.synchronize()
            if self.world > 1:
                torch.distributed.barrier()
            t0 = time.perf_counter()
            self.sp.upload()
            self.sp.solve()
            # every output of this rank's shard into pinned host memory, one copy per rank (N > 1: all PCIe links in parallel;
            # the gathered copy of every shard stays on every GPU)
            packed = self.sp.download_local_packed()
            checksum = int(packed[::4096].sum()) if packed is not None else 0      # the host really reads the result buffer
            torch.cuda.synchronize()
            ms += 1e3 * (time.perf_counter() - t0)
        self.res = self.sp.download() or []              # per-geometry dicts (host post-processing, not timed); None off-root
        self.checksum = checksum
        return ms

    def reduce_max(self, vals):
        torch = self.torch
        tt = torch.tensor(vals, dtype=torch.float64, device=self.dev)
        if self.world > 1:
            torch.distributed.all_reduce(tt, op=torch.distributed.ReduceOp.MAX)
        return tt.tolist()

    def gather_all(self, val):
        torch = self.torch
        tt = torch.tensor([val], dtype=torch.float64, device=self.dev)
        if self.world == 1:
            return [val]
        out = torch.empty(self.world, dtype=torch.float64, device=self.dev)
        torch.distributed.all_gather_into_tensor(out, tt)
        return out.tolist()

    def roofline(self, peak, peak_txt):
        """Dominant kernel of this rank's shard: closed-form flops / CUDA-event time, against the measured FP64 peak."""
        plan = self.sp.local
        st = {k: float(np.mean([x[k] for x in self.stage])) for k in self.stage[0]}
        factorised = self.algo == "schur" and plan.factorised_geometries() == plan.G
        mine = [self.flat[i] for i in self.sp.geom_parts[self.rank]]
        kf = kernel_flops(mine, plan.F, self.algo, factorised, plan.geom_kernel)
        if self.algo == "schur" and (plan.freq_kernel == "k_freq_small" or st.get("freq_solve", 0.0) == 0.0):
            kf = {"geometry": kf["geometry"], "frequency": kf["freq_gemm"] + kf["freq_solve"]}      # one kernel / several waves: combined stage
        dom = max(kf, key=lambda k: st[k])
        kname = {"geometry": plan.geom_kernel, "freq_gemm": "k_freq_gemm", "freq_solve": "k_freq_solve",
                 "frequency": plan.freq_kernel}[dom]
        achieved = kf[dom] / (st[dom] * 1e-3)
        traffic = None
        try:     # DRAM bytes per launch of the dominant kernel from the committed ncu capture of this workload
            with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as fh:
                traffic = json.load(fh).get(self.name, {}).get(kname.split("<")[0], {}).get("dram_bytes_per_launch")
        except Exception:
            pass
        if traffic is not None and not np.isfinite(traffic):
            traffic = None
        return {"bound": "tensor", "pipe": "FP64 DMMA (mma.sync m8n8k4 f64; tcgen05 has no FP64 kind)", "kernel": kname,
                "achieved": achieved / 1e12, "peak": peak / 1e12, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_txt, "flops_per_launch": kf[dom], "kernel_ms": st[dom], "stage_ms": st,
                "step_fp64_frac": sum(kf.values()) / (sum(st[k] for k in ("roots", "geometry", "frequency")) * 1e-3) / peak}


def solve_host_ms(ofb, geoms, omegas, reps=3):
    """The C-ABI host-buffer call INTEGRATION.md's reference-side stub makes (meemb200_solve_host: stream-ordered device
    allocation, H2D, kernels, D2H into caller-owned host arrays), timed through raw ctypes on pinned host outputs."""
    import ctypes as C
    import torch
    from openflash_b200 import _lib
    flat = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in geoms]
    plan = ofb.BatchPlan(flat, omegas)             # reuses the planner; its device buffers stay unused below
    I = plan.info
    host = {k: torch.empty(n, dtype=dt, pin_memory=True) for k, n, dt in
            (("am", int(I.n_rr), torch.float64), ("dp", int(I.n_rr), torch.float64), ("f", int(I.n_r), torch.float64),
             ("p", int(I.n_r), torch.float64), ("m0", int(I.n_sys), torch.float64), ("st", int(I.n_sys), torch.int32))}
    o = _lib.MeemOutputs(host["am"].data_ptr(), host["dp"].data_ptr(), host["f"].data_ptr(), host["p"].data_ptr(),
                         host["m0"].data_ptr(), host["st"].data_ptr(), None, None)
    lib = _lib.load()
    best = None
    for _ in range(reps + 1):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _lib.check(lib.meemb200_solve_host(plan.plan, C.byref(o), None))
        dt = 1e3 * (time.perf_counter() - t0)
        best = dt if best is None else min(best, dt)
    assert int(np.count_nonzero(host["st"].numpy())) == 0
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="big_tricylinder")
    ap.add_argument("--algo", default="schur", choices=["schur", "schur_lu", "dense"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip extra.sweep_1e6 / extra.workloads")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        args.gpus = world
    geoms, omegas, desc = workload(args.workload, rank, world)
    n_sys_rank = len(geoms) * len(omegas)
    config = {"workload": desc, "systems_per_gpu": n_sys_rank if args.workload != "sweep_1e6" else None,
              "l2": "256 MiB L2 flush between timed steps (GPU arm)"}
    cores = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return
        import multiprocessing as mp
        kind = cpu_kind()
        W = max(args.warmup, 1)
        tgt = cpu_target(args.workload, kind)
        with mp.get_context("fork").Pool(cores) as pool:
            for _ in range(W):
                run_cpu(kind, geoms, omegas, min(cores, tgt), pool, cores)       # pool + BLAS warm-up
            tot_sys, tot_t = 0, 0.0
            for _ in range(args.steps):
                n, t = run_cpu(kind, geoms, omegas, tgt, pool, cores, min_seconds=10.0 / max(args.steps, 1))
                tot_sys += n
                tot_t += t
        val = tot_sys / tot_t
        sample = f"{tot_sys // args.steps} systems per step sampled evenly from the workload, {cpu_description(kind, cores)}"
        details = {"timing": "time.perf_counter around each step of the process pool", "algo": "reference: dense complex LU per system (scipy.linalg.lu_factor)"}
        line = {"impl": "reference", "details": details, "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": W, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
                "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import torch.distributed as dist
    import openflash_b200 as ofb
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    details = {"timing": "CUDA events per step on the launching stream, max over ranks"}
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    W = max(args.warmup, 3)
    run = Runner(ofb, args.workload, world, rank, dev, args.algo, flush)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_ms = run.timed(args.steps, W)
    e2e_steps = max(2, min(args.steps, 5))
    e2e_ms = run.e2e(e2e_steps)
    clocks = sampler.stop() if rank == 0 else None
    t_ms, e2e_ms = run.reduce_max([t_ms, e2e_ms])
    bad = int(sum(np.count_nonzero(r["status"]) for r in run.res))
    plan = run.sp.local

    extra = {}
    if not args.no_extra and args.algo == "schur":
        # north star: the 10^6-system sweep as a fixed total over the N GPUs (cost-balanced geometry shards + one gather)
        sw = Runner(ofb, "sweep_1e6", world, rank, dev, "schur", flush, gather="all")
        ksw = max(args.steps, 3)
        sw_ms = sw.timed(ksw, 3)
        local_ms = float(np.mean([s["roots"] + s["geometry"] + s["frequency"] for s in sw.stage]))
        sw_e2e_ms = sw.e2e(2)
        per_rank = sw.gather_all(local_ms)
        sw_ms, sw_e2e_ms = sw.reduce_max([sw_ms, sw_e2e_ms])
        failed = sw.reduce_max([float(sum(np.count_nonzero(r["status"]) for r in sw.res))])[0]
        extra["sweep_1e6"] = {"systems": sw.total_sys, "scaling": "strong", "value": sw.total_sys * ksw / (sw_ms * 1e-3), "unit": UNIT,
                              "ms_per_step": sw_ms / ksw, "per_rank_kernel_ms": [round(x, 4) for x in per_rank],
                              "gather_ms": round(sw_ms / ksw - max(per_rank), 4),
                              "gather": "NCCL all_gather of the packed output buffers (the path's only exchange): every rank ends with every result",
                              "e2e": {"value": sw.total_sys * 2 / (sw_e2e_ms * 1e-3), "ms_per_step": sw_e2e_ms / 2,
                                      "includes": "H2D inputs, kernels, all_gather, one pinned D2H of each rank's own shard (all PCIe links in parallel)"},
                              "plan_ms_once": round(sw.plan_ms, 1), "workload": sw.desc, "failed_systems": int(failed)}
        if world > 1:       # the same sweep with the results delivered to rank 0 only (NCCL gather: serialised receives, not faster)
            sw2 = Runner(ofb, "sweep_1e6", world, rank, dev, "schur", flush, gather="root")
            ms2 = sw2.reduce_max([sw2.timed(ksw, 3)])[0]
            extra["sweep_1e6"]["value_gather_to_root"] = sw2.total_sys * ksw / (ms2 * 1e-3)
            del sw2
        del sw

    if rank == 0:
        total_sys = n_sys_rank * world if args.workload != "sweep_1e6" else run.total_sys
        value = total_sys * args.steps / (t_ms * 1e-3)
        e2e_val = total_sys * e2e_steps / (e2e_ms * 1e-3)
        peak_dfma = ofb.fp64_peak(0, local_rank)
        peak_dmma = ofb.fp64_peak(1, local_rank)
        peak = max(peak_dfma, peak_dmma)
        peak_txt = ("measured in this run: max(DFMA register loop %.1f, DMMA m8n8k4 %.1f) TFLOP/s; MEASURED_PEAKS.json has no FP64 entry"
                    % (peak_dfma / 1e12, peak_dmma / 1e12))
        roofline = run.roofline(peak, peak_txt)
        roofline["phase_share"] = ({k: (round(v, 4) if isinstance(v, float) else v) for k, v in plan.phase_profile().items()}
                                   if args.algo == "schur" else None)
        roofline["dense_equivalent_tflops"] = value * (8.0 / 3.0 * run.flat[0].size ** 3 + 8.0 * run.flat[0].size ** 2 * run.flat[0].R) / 1e12
        e2e = {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": plan.h2d_bytes * world, "d2h_bytes_per_step": int(plan.d2h_bytes) * world,
               "ms_per_step": e2e_ms / e2e_steps,
               "api": "ShardedPlan.upload + solve + download_local_packed on every rank (pinned H2D of the packed inputs, kernels, all_gather, one pinned D2H of the rank's own shard)"}
        if world == 1 and not args.no_extra and args.algo == "schur":
            ms = solve_host_ms(ofb, geoms, omegas)
            e2e["solve_host"] = {"ms_per_call": ms, "value": total_sys / (ms * 1e-3),
                                 "api": "meemb200_solve_host via raw ctypes (INTEGRATION.md stub): H2D + kernels + D2H into caller-owned host arrays"}
            wl = {}
            for name in ("mini_bicylinder", "small_tricylinder", "geometry_sweep", "convergence_sweep"):
                if name == args.workload:
                    continue
                r2 = Runner(ofb, name, 1, 0, dev, "schur", flush)
                ms2 = r2.timed(3, 3)
                e2 = r2.e2e(2)
                rf = r2.roofline(peak, peak_txt)
                wl[name] = {"value": r2.total_sys * 3 / (ms2 * 1e-3), "ms_per_step": ms2 / 3, "e2e": r2.total_sys * 2 / (e2 * 1e-3),
                            "plan_ms_once": round(r2.plan_ms, 1), "kernel": rf["kernel"], "roofline_frac": rf["frac"],
                            "step_fp64_frac": rf["step_fp64_frac"], "stage_ms": rf["stage_ms"], "workload": r2.desc,
                            "failed_systems": int(sum(np.count_nonzero(r["status"]) for r in r2.res))}
                del r2
            extra["workloads"] = wl
            # BASELINE.json configs[0] exactly as the reference's CPU artifact: ONE problem, ONE frequency, through the class API
            geo = ofb.BasicRegionGeometry.from_vectors(a=np.array([0.125, 0.25]), d=np.array([0.25, 0.125]), h=1.001, NMK=[10] * 3,
                                                       body_map=[0, 0], heaving_map=[True])
            prob = ofb.MEEMProblem(geo)
            prob.set_frequencies(np.array([float(np.sqrt(np.tanh(1.001) * 9.81))]))
            eng = ofb.MEEMEngine([prob], device=local_rank)
            lat = []
            for _ in range(25):
                t0 = time.perf_counter()
                ds = eng.run_and_store_results(0).dataset
                lat.append(1e3 * (time.perf_counter() - t0))
            extra["latency"] = {"mini_bicylinder_single_frequency_ms": float(np.median(lat[5:])),
                                "api": "MEEMEngine.run_and_store_results(0): plan + H2D + kernels + D2H + Results packaging, median of 20 calls",
                                "added_mass": float(np.asarray(ds["added_mass"].values)[0, 0, 0])}
        cpu = None
        if not args.no_cpu and world == 1:
            import multiprocessing as mp
            kind = cpu_kind()
            tgt = cpu_target(args.workload, kind)
            with mp.get_context("fork").Pool(cores) as pool:
                run_cpu(kind, geoms, omegas, min(cores, tgt), pool, cores)
                n, t = run_cpu(kind, geoms, omegas, tgt, pool, cores, min_seconds=10.0)
            cpu = {"value": n / t, "unit": UNIT, "cores": cores, "kind": kind,
                   "sample": f"{n} systems sampled evenly from the workload, {cpu_description(kind, cores)}, {t:.1f} s"}
            from oracle import ref_runner
            cpu["legs"] = ref_runner.time_legs(geoms, omegas, cores)      # (1) one process, (3) compiled
        details["algo"] = args.algo
        details["geometry_kernel"] = plan.geom_kernel
        line = {"details": details, "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": W,
                "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": run.scaling, "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "clocks": clocks, "e2e": e2e,
                "gpu_launches": (int(plan.info.n_launches)) * args.steps, "roofline": roofline, "cpu_baseline": cpu,
                "failed_systems": bad, "extra": extra}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
