#!/bin/bash
# Range-angle quick check: parity tests touching the kernel, then stage times of both bench workloads.
python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -3
for wl in 1843_RadVel 6843_RadVel_ods_20Hz; do
  echo "== $wl"
  python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['stage_ms_per_step'])"
done
