"""Cycle split of the grid-cooperative geometry kernel on big_tricylinder (measurement only)."""
import sys; sys.path.insert(0, '.')
import torch, bench
import openflash_b200.batch as ofb
geoms, om, d = bench.workload("big_tricylinder", 0, 1)
flat = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in geoms]
b = ofb.BatchPlan(flat, om, device=0)
b.upload(); b.solve(); torch.cuda.synchronize()
b.phase_profile()
b.solve(); torch.cuda.synchronize()
pp = b.phase_profile()
c = pp["coop_geom_cycles"]
print("cycles: factor(owner)=%.0f factor+bar=%.0f update=%.0f bar2=%.0f backsolve=%.0f backsolve+bar=%.0f extra=%s" % (c[0], c[1], c[2], c[3], c[4], c[5], c[6:]))
print({k: round(v, 3) for k, v in pp.items() if k != "coop_geom_cycles"})
