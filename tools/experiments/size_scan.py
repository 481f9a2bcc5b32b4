"""Timing scan over system sizes (no parity check: tests do that): tricylinder with n modes per region, 512 omega, and a
64-geometry batch of the same - looks for performance cliffs between the kernel families.
    python tools/experiments/size_scan.py"""
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
import openflash_b200 as ofb

om = np.linspace(0.3, 2.8, 512)
for n in (8, 16, 17, 24, 32, 48, 64, 100, 128, 160, 256):
    for G in (1, 64):
        geo = ofb.FlatGeometry.from_body_map(20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [n] * 4, [0, 1, 1], [0, 1])
        plan = ofb.BatchPlan([geo] * G, om)
        plan.upload()
        for _ in range(2):
            plan.solve()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            plan.solve()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        st = plan.stage_ms()
        bad = int(np.count_nonzero(plan.download()["status"]))
        print(f"n={n:4d} G={G:3d} {plan.geom_kernel:26s} {plan.freq_kernel:24s} {ms:9.3f} ms  {G * 512 / ms * 1e3:12.0f} sys/s  geom {st['geometry']:.3f} freq {st['frequency']:.3f} bad {bad}")
