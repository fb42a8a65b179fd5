#!/bin/bash
# Stage times and rooflines of the other BASELINE shapes. usage (under gpurun): bash tools/gpu_shapes.sh <tag> [workloads...]
TAG=${1:-shapes}; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
for WL in ${@:-RadSAR_1843_official RadCloud RadarHD 1843_RadVel}; do
  timeout 300 python bench.py --workload $WL --frames ${FRAMES:-4096} --steps 30 --warmup 5 --no-cpu-baseline --no-extra --e2e-steps 1 2>$OUT/bench_$WL.err > $OUT/bench_$WL.json
  echo "$WL rc=$?"; python -c "import sys,json; d=json.loads(open('$OUT/bench_$WL.json').readlines()[-1]); print(d['value'], d['stage_ms_per_step'], [ (k['kernel'], round(k['frac'],3)) for k in d['roofline']['kernels']])" || tail -5 $OUT/bench_$WL.err
done
