#!/usr/bin/env bash
# 1/2/4/8-GPU runs of the default bench line (weak big_tricylinder + extra.sweep_1e6 strong scaling) on one 8-GPU box.
set -u
mkdir -p gpurun_out
for n in 8 4 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/scale_${n}gpu.json 2> gpurun_out/scale_${n}gpu.err || echo "N=$n failed"
done
python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu > gpurun_out/scale_1gpu.json 2> gpurun_out/scale_1gpu.err || echo "N=1 failed"
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/scale_{n}gpu.json").read().strip().splitlines()[-1])
        s = d["extra"]["sweep_1e6"]
        print(n, "weak C3:", round(d["value"]), "e2e", round(d["e2e"]["value"]), "| strong 1e6:", round(s["value"]), "ms", round(s["ms_per_step"], 3),
              "per-rank", s["per_rank_kernel_ms"], "gather", s["gather_ms"], "e2e", round(s["e2e"]["value"]))
    except Exception as e:
        print(n, "unreadable:", e)
PY
