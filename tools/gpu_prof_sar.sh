#!/bin/bash
# ncu full capture of the tensor-core SAR kernel on the throughput grid. usage: bash tools/gpu_prof_sar.sh <tag>
TAG=${1:-sar}; OUT=gpurun_out/$TAG; mkdir -p $OUT
python tools/bench_sar.py > $OUT/bench_sar_plain.json 2>&1 || { echo "plain run failed"; tail -5 $OUT/bench_sar_plain.json; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:sar_tc_kernel -s 30 -c 1 -f -o $OUT/prof_sar_tc python tools/bench_sar.py > $OUT/ncu_sar_tc.log 2>&1
echo "ncu rc=$?"; tail -3 $OUT/ncu_sar_tc.log
