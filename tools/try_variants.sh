#!/usr/bin/env bash
# Scratch A/B runs on one B200: geometry-stage variants on C2 / C3 (stage times from bench.py --no-extra --no-cpu).
for env in "" "MEEMB200_NO_COOP=1"; do
  for w in small_tricylinder big_tricylinder; do
    echo "== $w $env"
    env $env timeout 300 python bench.py --workload $w --no-extra --no-cpu --steps 3 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), d['details']['geometry_kernel'], {k: round(v, 3) for k, v in d['roofline']['stage_ms'].items()})
    elif 'rror' in l: print(l.strip())
"
  done
done
