#!/bin/bash
# OS-CFAR first-phase depth experiment: stage times for RADAR_B200_CFAR_PH1 = 0..3 on both bench workloads.
for wl in 1843_RadVel 6843_RadVel_ods_20Hz; do
  for ph in 0 1 2 3; do
    echo "== $wl CFAR_PH1=$ph"
    RADAR_B200_CFAR_PH1=$ph python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
      | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['stage_ms_per_step']['cfar'], d['points_per_step_rank0'])"
  done
done
