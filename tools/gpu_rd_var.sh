#!/bin/bash
# Range-Doppler kernel variants on one workload: RADAR_B200_RD_REG in $VARS (0 = shared-memory streaming kernel).
# usage (under gpurun): VARS="1 2 3 0" bash tools/gpu_rd_var.sh <tag> [workload] [frames]
TAG=${1:-var}; WL=${2:-6843_RadVel_ods_20Hz}; FR=${3:-4096}
OUT=gpurun_out/$TAG; mkdir -p $OUT
for REG in ${VARS:-1 0}; do
  RADAR_B200_RD_REG=$REG timeout 600 python -m pytest tests -m gpu -x -q -k "range_doppler or smoke" > $OUT/pytest_$REG.log 2>&1; echo "reg=$REG pytest rc=$? $(tail -1 $OUT/pytest_$REG.log)"
  RADAR_B200_RD_REG=$REG timeout 300 python bench.py --workload $WL --frames $FR --steps 30 --warmup 5 --no-cpu-baseline --no-extra --e2e-steps 1 2>$OUT/bench_reg$REG.err > $OUT/bench_reg$REG.json
  echo "reg=$REG rc=$?"; python -c "import sys,json; d=json.loads(open('$OUT/bench_reg$REG.json').readlines()[-1]); print(d['value'], d['stage_ms_per_step'], d['roofline']['frac'])" || tail -5 $OUT/bench_reg$REG.err
done
