"""Product multi-GPU entry through NCCL: MEEMEngine.run_batch(problems, group=...) and solve_batch_distributed on every
rank of a torchrun job, checked against the oracle and against the single-GPU result.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/experiments/mgpu_engine_check.py"""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist
sys.path.insert(0, '.')
import bench
import openflash_b200 as ofb
from openflash_b200.distributed import ShardedPlan, solve_batch_distributed
from oracle import meem_oracle as mo

rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
world = dist.get_world_size()
ok = True
# 1. geometry sharding: 37 C4 geometries (uneven shards), results on every rank in caller order
gs = bench.c4_geometries(37)
om = np.linspace(0.3, 2.5, 5)
geos = [ofb.FlatGeometry.from_body_map(g["h"], g["a"], g["d"], g["NMK"], g["body_map"], g["modes"]) for g in gs]
res = solve_batch_distributed(geos, om, device=local)
single = ofb.solve_batch(geos, om, device=local)
for g, r, s in zip(gs, res, single):
    for v in ("added_mass", "damping", "excitation_force", "excitation_phase", "m0", "status"):
        ok &= bool(np.array_equal(r[v], s[v]))
ref = mo.run_sweep(mo.FlatGeometry(gs[5]["h"], gs[5]["a"], gs[5]["d"], gs[5]["NMK"]), om, gs[5]["body_map"], gs[5]["modes"])
ok &= bool(np.max(np.abs(res[5]["added_mass"] - ref["added_mass"]) / np.abs(ref["added_mass"])) < 1e-10)
st = ShardedPlan(ofb.GeometryBatch.from_dicts(gs), om, device=local)
st.upload(); st.solve()
stk = st.download_stacked()
ok &= all(bool(np.array_equal(stk["added_mass"][i], single[i]["added_mass"])) for i in range(37))
# 2. frequency sharding + engine API with potentials and c vectors (R = 2)
geo = ofb.BasicRegionGeometry.from_vectors(a=np.array([5.0, 10.0, 15.0]), d=np.array([15.0, 10.0, 5.0]), h=20.0, NMK=[20] * 4,
                                           body_map=[0, 1, 1], heaving_map=[False, False])
prob = ofb.MEEMProblem(geo)
for b in prob.geometry.body_arrangement.bodies:
    b.heaving = True
omegas = np.linspace(0.2, 2.9, 11)
prob.set_frequencies(omegas)
eng = ofb.MEEMEngine([prob], device=local)
ds = eng.run_batch([prob], group=dist.group.WORLD)[0].dataset
ds1 = eng.run_batch([prob])[0].dataset
for v in ("added_mass", "damping", "excitation_force", "excitation_phase", "c_vector", "potentials_real", "potentials_imag"):
    ok &= bool(np.array_equal(np.asarray(ds[v].values), np.asarray(ds1[v].values), equal_nan=True))
ref = mo.run_sweep(mo.FlatGeometry(20.0, [5.0, 10.0, 15.0], [15.0, 10.0, 5.0], [20] * 4), omegas, [0, 1, 1], [0, 1])
ok &= bool(np.max(np.abs(np.asarray(ds["added_mass"].values) - ref["added_mass"]) / np.abs(ref["added_mass"])) < 1e-10)
# 3. gather to one rank
r3 = ShardedPlan(geos, om, device=local, gather="root", root=world - 1).run()
ok &= (r3 is None) if rank != world - 1 else all(bool(np.array_equal(a["damping"], b["damping"])) for a, b in zip(r3, single))
t = torch.tensor([1.0 if ok else 0.0], device=f"cuda:{local}")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print("mgpu_engine_check:", "OK" if t.item() == 1.0 else "FAILED", "world", world)
dist.destroy_process_group()
sys.exit(0 if t.item() == 1.0 else 1)
