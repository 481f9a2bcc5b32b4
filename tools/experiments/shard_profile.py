"""Stage times of ONE rank's share of the 10^6-system sweep at 8 GPUs (1 953 C4 geometries x 64 omega) on one GPU:
what the fixed costs (roots, tables, launches) weigh when the per-rank step is ~2.3 ms.
    python tools/experiments/shard_profile.py [n_geom]"""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
import bench
import openflash_b200 as ofb

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1953
gs = bench.c4_geometries(15625)[:n]
flat = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
plan = ofb.BatchPlan(flat, np.linspace(0.3, 2.5, 64))
plan.upload()
for _ in range(3):
    plan.solve()
torch.cuda.synchronize()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ms, st = [], []
for _ in range(5):
    ev[0].record(); plan.solve(); ev[1].record(); torch.cuda.synchronize()
    ms.append(ev[0].elapsed_time(ev[1])); st.append(plan.stage_ms())
print("step ms", np.round(ms, 4), {k: round(float(np.mean([s[k] for s in st])), 4) for k in st[0]})
