#!/usr/bin/env bash
# Round-end validation on one B200 (run through gpurun from the repo root):
#   parity tests, smoke, the four bench workloads, the reference arm, the ncu launch list and one full capture per
#   workload family.  Everything lands in gpurun_out/; tools/summarize_ncu.py turns the captures into profiles/.
#   gpurun copies back at most 64 MiB: `round_end.sh sweep` runs only the geometry_sweep capture (a second call).
set -u
mkdir -p gpurun_out
if [ "${1:-}" = sweep ]; then
    timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_geom|k_hchol|k_freq" -c 4 -f -o gpurun_out/prof_sweep \
        python bench.py --workload geometry_sweep --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_gs.log 2>&1
    echo "ncu sweep rc=$?"
    exit 0
fi
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
rc=0
for w in big_tricylinder mini_bicylinder small_tricylinder geometry_sweep; do
    o=$w; [ $w = big_tricylinder ] && o=big
    timeout 600 python bench.py --workload $w > gpurun_out/bench_$o.json 2> gpurun_out/bench_$o.err || rc=1
done
timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err || rc=1
echo "bench rc=$rc"
rc=0
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu > gpurun_out/ncu_l.log 2>&1 || rc=1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"k_geom|k_hchol|k_freq" -c 5 -f -o gpurun_out/prof_big \
    python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_f.log 2>&1 || rc=1
echo "ncu rc=$rc"
python - <<'PY'
import json
for w in ("big", "mini_bicylinder", "small_tricylinder", "geometry_sweep"):
    try:
        d = json.loads(open(f"gpurun_out/bench_{w}.json").read().strip().splitlines()[-1])
        print(w, round(d["value"]), round(d["e2e"]["value"]), round(d["roofline"]["frac"], 4), d["roofline"].get("stage_ms"), d.get("failed_systems"))
    except Exception as e:
        print(w, "unreadable:", e)
PY
