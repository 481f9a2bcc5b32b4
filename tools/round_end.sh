#!/usr/bin/env bash
# Round-end validation on one B200 (run through gpurun from the repo root):
#   parity tests, smoke, the default bench line (with extra.sweep_1e6 / extra.workloads / e2e.solve_host / cpu_baseline),
#   the reference arm, the ncu launch lists and one full capture per workload family.
#   Everything lands in gpurun_out/; tools/summarize_ncu.py turns the captures into profiles/.
#   gpurun copies back at most 64 MiB: `round_end.sh sweep` runs only the geometry_sweep captures (a second call).
set -u
mkdir -p gpurun_out
if [ "${1:-}" = sweep ]; then
    python bench.py --workload geometry_sweep --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/bench_geometry_sweep.json 2> gpurun_out/bench_gs.err &&
    timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_sweep.csv \
        python bench.py --workload geometry_sweep --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/ncu_gs_l.log 2>&1
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_geom|k_hchol|k_freq|k_tables" -c 5 -f -o gpurun_out/prof_sweep \
        python bench.py --workload geometry_sweep --steps 1 --warmup 0 --no-cpu --no-extra > gpurun_out/ncu_gs.log 2>&1
    echo "ncu sweep rc=$?"
    exit 0
fi
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
rc=0
timeout 900 python bench.py > gpurun_out/bench_big.json 2> gpurun_out/bench_big.err || rc=1
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err || rc=1
echo "bench rc=$rc"
rc=0
python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > /dev/null 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/ncu_l.log 2>&1 || rc=1
python bench.py --steps 1 --warmup 0 --no-cpu --no-extra > /dev/null 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_geom|k_hchol|k_freq" -c 5 -f -o gpurun_out/prof_big \
    python bench.py --steps 1 --warmup 0 --no-cpu --no-extra > gpurun_out/ncu_f.log 2>&1 || rc=1
echo "ncu rc=$rc"
python - <<'PY'
import json
try:
    d = json.loads(open("gpurun_out/bench_big.json").read().strip().splitlines()[-1])
    print("big", round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"], 4), d["roofline"]["stage_ms"], d.get("failed_systems"))
    for k, w in d["extra"].get("workloads", {}).items():
        print(k, round(w["value"]), round(w["e2e"]), w["kernel"], round(w["roofline_frac"], 4), w["stage_ms"])
    print("sweep_1e6", round(d["extra"]["sweep_1e6"]["value"]), d["extra"]["sweep_1e6"]["ms_per_step"])
    r = json.loads(open("gpurun_out/bench_ref.json").read().strip().splitlines()[-1])
    print("reference arm", r["value"], r["cpu_baseline"]["kind"], r["cpu_baseline"]["cores"])
except Exception as e:
    print("unreadable:", e)
PY
