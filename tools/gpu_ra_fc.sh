#!/bin/bash
# Range-angle pass-size experiment: stage times for forced frames-per-pass values and the cost model's pick.
for wl in 1843_RadVel 6843_RadVel_ods_20Hz; do
  for fc in 0 1 2 4; do
    echo "== $wl RA_FC=$fc"
    RADAR_B200_RA_FC=$fc python bench.py --workload $wl --steps 30 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
      | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print(d['value'], d['stage_ms_per_step'])"
  done
done
python -m pytest tests/test_gpu_parity.py -x -q -k "range_angle or ra" 2>&1 | tail -3
