#!/bin/bash
# One GPU-box session: parity tests, bench lines, ncu launch list and a full capture of the top kernel.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [kernel-regex]
TAG=${1:-r1}
KREGEX=${2:-rd_stream}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem,power.limit --format=csv > $OUT/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/pytest_gpu.log
for WL in 1843_RadVel 6843_RadVel_ods_20Hz; do
  timeout 600 python bench.py --workload $WL > $OUT/bench_$WL.json 2> $OUT/bench_$WL.err; echo "bench $WL rc=$?"; cat $OUT/bench_$WL.json; tail -3 $OUT/bench_$WL.err
done
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/bench_reference.json 2>&1; echo "ref rc=$?"; cat $OUT/bench_reference.json
SMALL="python bench.py --steps 2 --warmup 3 --frames 2048 --e2e-steps 1 --no-cpu-baseline"
$SMALL > $OUT/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/launches.csv $SMALL > $OUT/ncu_launch.log 2>&1
echo "ncu launches rc=$?"
$SMALL > $OUT/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$KREGEX -s 3 -c 2 -f -o $OUT/prof_$KREGEX $SMALL > $OUT/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 $OUT/ncu_full.log
ls -la $OUT
