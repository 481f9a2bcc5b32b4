"""A/B of k_freq_gemm variants selected by environment switches read at plan creation (MEEMB200_NO_STREAM, ...):
outputs of the variant against the baseline plan on the same inputs, and stage times of both.
    python tools/experiments/stream_check.py [n_freq] [modes] [ENV=1 ...]   (ENV switches define the BASELINE plan)"""
import os
import sys
import numpy as np
import torch
sys.path.insert(0, '.')
import openflash_b200 as ofb

args = [a for a in sys.argv[1:] if "=" not in a]
envs = [a.split("=") for a in sys.argv[1:] if "=" in a] or [["MEEMB200_NO_STREAM", "1"]]
F = int(args[0]) if len(args) > 0 else 4096
nm = int(args[1]) if len(args) > 1 else 200
geo = ofb.FlatGeometry.from_body_map(25.0, [10.0, 15.0, 20.0], [20.0, 15.0, 10.0], [nm] * 4, [0, 1, 2], [0])
om = np.linspace(0.1, 3.0, F)


def run(env):
    for k, v in env: os.environ[k] = v
    plan = ofb.BatchPlan([geo], om)
    for k, v in env: del os.environ[k]
    plan.upload()
    for _ in range(3): plan.solve()
    torch.cuda.synchronize()
    st = []
    for _ in range(5):
        plan.solve(); torch.cuda.synchronize(); st.append(plan.stage_ms())
    out = plan.download(copy=True)
    return out, {k: round(float(np.mean([s[k] for s in st])), 4) for k in st[0]}


new, t_new = run([])
old, t_old = run(envs)
print("variant ", t_new)
print("baseline", envs, t_old)
for k in ("added_mass", "damping", "excitation_force", "excitation_phase"):
    a, b = np.asarray(new[0][k] if isinstance(new, list) else new[k]), np.asarray(old[0][k] if isinstance(old, list) else old[k])
    print(k, "max rel diff", float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))), "nan", int(np.isnan(a).sum()), int(np.isnan(b).sum()))
st = np.asarray(new[0]["status"] if isinstance(new, list) else new["status"])
print("status nonzero:", int((st != 0).sum()))
