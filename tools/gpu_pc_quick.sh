#!/bin/bash
# Point-cloud kernel check: the full GPU parity suite, then stage times of both bench workloads with the thread-per-detection
# kernel (default) and the warp-per-detection kernel (RADAR_B200_PC=1).
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in 1843_RadVel 6843_RadVel_ods_20Hz; do
  for mode in 0 1; do
    echo "== $wl RADAR_B200_PC=$mode"
    RADAR_B200_PC=$mode python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
      | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['stage_ms_per_step'])"
  done
done
