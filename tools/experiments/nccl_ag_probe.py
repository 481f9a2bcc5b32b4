"""Time of the one exchange of the 10^6-system sweep by itself: all_gather_into_tensor of 5.5 MB per rank (the packed
outputs of 125 000 systems), for NCCL environment settings given on the command line of the launcher.
    NCCL_ALGO=Ring python -m torch.distributed.run --nproc-per-node 8 ... tools/experiments/nccl_ag_probe.py"""
import os
import torch
import torch.distributed as dist
lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
w = dist.get_world_size()
for nbytes in (5_500_160, 1_375_232):
    src = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
    dst = torch.empty(nbytes * w, dtype=torch.uint8, device="cuda")
    for _ in range(10): dist.all_gather_into_tensor(dst, src)
    torch.cuda.synchronize(); dist.barrier()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ts = []
    for _ in range(30):
        dist.barrier(); torch.cuda.synchronize()
        ev[0].record(); dist.all_gather_into_tensor(dst, src); ev[1].record(); torch.cuda.synchronize()
        ts.append(ev[0].elapsed_time(ev[1]))
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if dist.get_rank() == 0:
        print({k: v for k, v in os.environ.items() if k.startswith("NCCL_")}, f"{nbytes} B/rank: median {float(t):.4f} ms", flush=True)
dist.destroy_process_group()
