#!/bin/bash
# ncu full capture of one kernel on a small bench run.
# usage: bash tools/gpu_prof.sh <tag> <kernel-regex> <workload> [frames] [stream-variant]
TAG=$1; KREGEX=$2; WL=$3; FR=${4:-2048}; export RADAR_B200_STREAM_VARIANT=${5:-0}
OUT=gpurun_out/$TAG; mkdir -p $OUT
SMALL="python bench.py --workload $WL --steps 2 --warmup 3 --frames $FR --e2e-steps 1 --no-cpu-baseline"
$SMALL > $OUT/plain_$WL.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 3 -c 1 -f -o $OUT/prof_${KREGEX}_${WL}_v$RADAR_B200_STREAM_VARIANT $SMALL > $OUT/ncu_full_$WL.log 2>&1
echo "ncu full rc=$?"; tail -2 $OUT/ncu_full_$WL.log
