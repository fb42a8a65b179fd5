#!/bin/bash
# Range-Doppler streaming-kernel variants (RADAR_B200_STREAM_VARIANT) of one workload: kernel time and roofline fraction.
WL=${1:-6843_RadVel_ods_20Hz}; shift
for v in ${@:-0 1 2 3}; do
  RADAR_B200_STREAM_VARIANT=$v python bench.py --workload $WL --steps 20 --warmup 5 --no-cpu-baseline --e2e-steps 1 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); r=d['roofline']; print('variant $v', round(r['kernel_ms_per_launch'],4), round(r['frac'],3), round(d['value']))"
done
