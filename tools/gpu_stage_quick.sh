#!/bin/bash
# Stage times of both bench workloads (no CPU baseline, one e2e step) - for A/B runs of a rebuilt library.
for wl in 1843_RadVel 6843_RadVel_ods_20Hz; do
  echo "== $wl"
  python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['stage_ms_per_step'])"
done
