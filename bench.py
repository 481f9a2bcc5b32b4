#!/usr/bin/env python
"""bench.py — MEEM systems solved per second on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--workload NAME] [--algo schur|dense] [--impl reference]

One "step" = one pass of the hot path over one batch: roots -> per-geometry stage -> per-(geometry,
frequency) solves -> hydrodynamic coefficients (+ for N > 1 the NCCL all-gather of the coefficient
arrays, the path's only exchange).  Each rank runs the same-size shard (weak scaling); `value` is
whole-job systems/s with inputs resident in HBM; `e2e` times the public host-buffer call
(BatchPlan.upload + solve + download: H2D of the packed inputs from pinned memory, D2H of every
output) per step.  `--impl reference` times the CPU oracle port of the reference path (numpy/scipy,
all host cores) on a bounded sample of the same workload, rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

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
def _cpu_chunk(args):
    g, omegas = args
    try:
        from threadpoolctl import threadpool_limits
        ctx = threadpool_limits(limits=1)
    except Exception:
        import contextlib
        ctx = contextlib.nullcontext()
    import warnings
    from oracle import meem_oracle as mo
    with ctx, warnings.catch_warnings():
        warnings.simplefilter("ignore")
        og = mo.FlatGeometry(g["h"], g["a"], g["d"], g["NMK"])
        out = mo.run_sweep(og, omegas, g["body_map"], g["modes"])
    return float(np.nansum(out["added_mass"]))


def cpu_sample_tasks(geoms, omegas, target_systems):
    """A bounded sample of the workload: (geometry, omega-chunk) tasks covering ~target_systems systems."""
    tasks = []
    if len(geoms) == 1:
        idx = np.linspace(0, len(omegas) - 1, min(target_systems, len(omegas))).astype(int)
        for i in idx:
            tasks.append((geoms[0], [float(omegas[i])]))
        n_sys = len(idx)
    else:
        ng = max(1, min(len(geoms), target_systems // len(omegas)))
        for g in geoms[:ng]:
            tasks.append((g, [float(w) for w in omegas]))
        n_sys = ng * len(omegas)
    return tasks, n_sys


def run_cpu(geoms, omegas, target_systems, pool, min_seconds=0.0, max_seconds=30.0):
    """Time the oracle on a sample of the workload.  With min_seconds > 0 the sample is repeated (same tasks) until
    that much wall time has been spent, so that the baseline rests on 10-30 s of CPU work."""
    tasks, n_sys = cpu_sample_tasks(geoms, omegas, target_systems)
    t0 = time.perf_counter()
    done = 0
    while True:
        list(pool.map(_cpu_chunk, tasks, chunksize=1))
        done += n_sys
        el = time.perf_counter() - t0
        if el >= min_seconds or el * 2 > max_seconds:
            break
    return done, time.perf_counter() - t0


def cpu_target(name):
    return {"big_tricylinder": 512, "small_tricylinder": 256, "mini_bicylinder": 2048, "geometry_sweep": 64}[name]


# ------------------------------------------------------------------------------------------ main
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="big_tricylinder")
    ap.add_argument("--algo", default="schur", choices=["schur", "schur_lu", "dense"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        args.gpus = world
    geoms, omegas, desc = workload(args.workload, rank, world)
    n_sys_rank = len(geoms) * len(omegas)
    config = {"workload": desc, "algo": args.algo if args.impl == "b200" else "reference dense complex LU per system",
              "systems_per_gpu": n_sys_rank, "timing": "CUDA events per step on the launching stream; 256 MiB L2 flush between timed steps"}
    cores = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return
        import multiprocessing as mp
        W = max(args.warmup, 1)
        with mp.get_context("fork").Pool(cores) as pool:
            run_cpu(geoms, omegas, min(cores, cpu_target(args.workload)), pool)        # pool + BLAS warm-up
            for _ in range(W - 1):
                run_cpu(geoms, omegas, min(cores, cpu_target(args.workload)), pool)
            tot_sys, tot_t = 0, 0.0
            for _ in range(args.steps):
                n, t = run_cpu(geoms, omegas, cpu_target(args.workload), pool, min_seconds=10.0 / max(args.steps, 1))
                tot_sys += n
                tot_t += t
        val = tot_sys / tot_t
        sample = f"{tot_sys // args.steps} systems per step sampled evenly from the workload, {cores} worker processes x 1 BLAS thread"
        line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                "warmup": W, "ms_per_step": 1e3 * tot_t / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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
    flat = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in geoms]
    plan = ofb.BatchPlan(flat, omegas, algo={"schur": ofb.ALGO_SCHUR, "schur_lu": ofb.ALGO_SCHUR_LU, "dense": ofb.ALGO_DENSE}[args.algo], device=local_rank)
    n_out = int(2 * plan.info.n_rr + 2 * plan.info.n_r)
    packed = torch.empty(n_out, dtype=torch.float64, device=dev)
    gathered = torch.empty(n_out * world, dtype=torch.float64, device=dev) if world > 1 else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def step():
        plan.solve()
        if world > 1:   # the path's only exchange: gather the coefficient arrays (north_star)
            torch.cat([plan.d_am, plan.d_dp, plan.d_force, plan.d_phase], out=packed)
            dist.all_gather_into_tensor(gathered, packed)

    plan.upload()
    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    stage = []
    for s in range(args.steps):
        flush.fill_(s & 0xFF)
        ev[s][0].record()
        step()
        ev[s][1].record()
        torch.cuda.synchronize()
        stage.append(plan.stage_ms())
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_ms = sum(a.elapsed_time(b) for a, b in ev)
    # end-to-end through the host-buffer API: H2D inputs + kernels + D2H outputs, every step
    e2e_ms = 0.0
    for s in range(max(2, min(args.steps, 5))):
        flush.fill_(s)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        plan.upload()
        plan.solve()
        res = plan.download()
        torch.cuda.synchronize()
        e2e_ms += 1e3 * (time.perf_counter() - t0)
    e2e_steps = max(2, min(args.steps, 5))
    clocks = sampler.stop() if rank == 0 else None
    tt = torch.tensor([t_ms, e2e_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    t_ms, e2e_ms = tt.tolist()
    bad = int(np.count_nonzero(res["status"]))
    if rank == 0:
        total_sys = n_sys_rank * world
        value = total_sys * args.steps / (t_ms * 1e-3)
        e2e_val = total_sys * e2e_steps / (e2e_ms * 1e-3)
        # roofline of the dominant kernel
        peak_dfma = ofb.fp64_peak(0, local_rank)
        peak_dmma = ofb.fp64_peak(1, local_rank)
        peak = max(peak_dfma, peak_dmma)
        st = {k: float(np.mean([x[k] for x in stage])) for k in stage[0]}
        phase_share = {k: (round(v, 4) if isinstance(v, float) else v) for k, v in plan.phase_profile().items()} if args.algo == "schur" else None
        F_ = len(omegas)

        factorised = args.algo == "schur" and plan.factorised_geometries() == len(flat)

        def kernel_flops(g):   # closed-form real FP64 flops per kernel for geometry g (kernel table)
            N, K, R, p = float(g.NMK[-2]), float(g.NMK[-1]), float(g.R), float(g.size - g.NMK[-1])
            NT = float(g.size)
            if args.algo == "dense":
                return {"frequency": F_ * (8 / 3 * NT ** 3 + 8 * NT * NT * R)}
            geom = 2 / 3 * p ** 3 + 2 * p * p * (N + R)
            if args.algo == "schur":
                # factorised-H form (k_hchol): W = L^T Imk triangular (N^2 K) + right-hand-side rows (4 R N K) + W^T W lower (N K^2);
                # H itself when it did not factorise: 2 N^2 K + N K^2
                gemm = (N * N * K + 4 * R * N * K + N * K * K) if factorised else (2 * N * N * K + N * K * K)
                return {"geometry": geom + (N ** 3 / 3 if factorised else 0.0), "freq_gemm": F_ * gemm,
                        "freq_solve": F_ * (K ** 3 / 3 + 2 * K * K * R + 2 * K * R * R)}     # Cholesky + 2R forward solves + dots
            return {"geometry": geom, "frequency": F_ * (2 * N * N * K + 2 * N * K * K + 2 / 3 * K ** 3 + 2 * K * K * (1 + 2 * R) + 4 * N * K * R)}
        kf = {}
        for g in flat:
            for k, v in kernel_flops(g).items():
                kf[k] = kf.get(k, 0.0) + v
        if args.algo == "schur" and st.get("freq_solve", 0.0) == 0.0:      # several waves: only the combined stage was timed
            kf = {"geometry": kf["geometry"], "frequency": kf["freq_gemm"] + kf["freq_solve"]}
        dom = max(kf, key=lambda k: st[k])
        kname = {"geometry": "k_geom_coop/k_geom", "freq_gemm": "k_freq_gemm", "freq_solve": "k_freq_solve",
                 "frequency": "k_dense" if args.algo == "dense" else "k_freq_*"}[dom]
        kflops = kf[dom]
        achieved = kflops / (st[dom] * 1e-3)
        traffic = None
        try:     # DRAM bytes per launch of the dominant kernel from the committed ncu capture of this workload
            with open(os.path.join(ROOT, "profiles", "ncu_summary.json")) as fh:
                traffic = json.load(fh).get(args.workload, {}).get(kname, {}).get("dram_bytes_per_launch")
        except Exception:
            pass
        if traffic is not None and not np.isfinite(traffic):
            traffic = None
        roofline = {"bound": "tensor", "pipe": "FP64 DMMA (mma.sync m8n8k4 f64; tcgen05 has no FP64 kind)", "kernel": kname,
                    "achieved": achieved / 1e12, "peak": peak / 1e12, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": traffic,
                    "peak_source": "measured in this run: max(DFMA register loop %.1f, DMMA m8n8k4 %.1f) TFLOP/s; MEASURED_PEAKS.json has no FP64 entry"
                                   % (peak_dfma / 1e12, peak_dmma / 1e12),
                    "flops_per_launch": kflops, "kernel_ms": st[dom], "stage_ms": st, "phase_share": phase_share,
                    "step_fp64_frac": sum(kf.values()) / (sum(st[k] for k in ("roots", "geometry", "frequency")) * 1e-3) / peak,
                    "dense_equivalent_tflops": value * (8.0 / 3.0 * flat[0].size ** 3 + 8.0 * flat[0].size ** 2 * flat[0].R) / 1e12}
        cpu = None
        if not args.no_cpu and world == 1:
            import multiprocessing as mp
            with mp.get_context("fork").Pool(cores) as pool:
                run_cpu(geoms, omegas, min(cores, cpu_target(args.workload)), pool)
                n, t = run_cpu(geoms, omegas, cpu_target(args.workload), pool, min_seconds=10.0)
            cpu = {"value": n / t, "unit": UNIT, "cores": cores, "kind": "port",
                   "sample": f"{n} systems sampled evenly from the workload, oracle (numpy/scipy restatement of the reference path), {cores} processes x 1 BLAS thread, {t:.1f} s"}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": t_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config, "clocks": clocks,
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": plan.h2d_bytes, "d2h_bytes_per_step": plan.d2h_bytes,
                        "ms_per_step": e2e_ms / e2e_steps},
                "gpu_launches": int(plan.info.n_launches) * args.steps, "roofline": roofline, "cpu_baseline": cpu,
                "failed_systems": bad}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
