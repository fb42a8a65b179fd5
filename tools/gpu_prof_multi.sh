#!/bin/bash
# ncu full capture of several kernels (one launch each) on a small bench run.
# usage: bash tools/gpu_prof_multi.sh <tag> <workload> <frames> <regex1> [regex2 ...]
TAG=$1; WL=$2; FR=$3; shift 3
OUT=gpurun_out/$TAG; mkdir -p $OUT
SMALL="python bench.py --workload $WL --steps 2 --warmup 3 --frames $FR --e2e-steps 1 --no-cpu-baseline"
$SMALL > $OUT/plain_$WL.log 2>&1 || { echo "plain run failed"; tail -5 $OUT/plain_$WL.log; exit 1; }
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o $OUT/prof_${K}_${WL} $SMALL > $OUT/ncu_${K}_$WL.log 2>&1
  echo "ncu $K rc=$?"
done
