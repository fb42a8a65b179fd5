#!/bin/bash
# CFAR tests + stage times with and without the level-ladder screen. usage (under gpurun): bash tools/gpu_cfar_ab.sh <tag>
TAG=${1:-cfar}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q -k "cfar or parity or dropin or smoke" > $OUT/pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest.log
for WL in 6843_RadVel_ods_20Hz 1843_RadVel; do for LAD in 1 0; do
  RADAR_B200_CFAR_LADDER=$LAD timeout 300 python bench.py --workload $WL --frames 4096 --steps 30 --warmup 5 --no-cpu-baseline --no-extra --e2e-steps 1 2>$OUT/bench_${WL}_lad$LAD.err > $OUT/bench_${WL}_lad$LAD.json
  echo "$WL ladder=$LAD rc=$?"; python -c "import sys,json; d=json.loads(open('$OUT/bench_${WL}_lad$LAD.json').readlines()[-1]); print(d['value'], d['stage_ms_per_step'], d.get('points_per_step_rank0'))" || tail -5 $OUT/bench_${WL}_lad$LAD.err
done; done
