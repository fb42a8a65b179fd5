#!/bin/bash
# SAR tensor-core path: parity tests, then timing of the three modes on the throughput grid. usage: bash tools/gpu_sar.sh <tag>
TAG=${1:-sar}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_sar_tc.py -m gpu -q > $OUT/pytest_sar_tc.log 2>&1; echo "pytest rc=$?"; tail -30 $OUT/pytest_sar_tc.log
timeout 300 python tools/bench_sar.py > $OUT/bench_sar.json 2> $OUT/bench_sar.err; echo "bench_sar rc=$?"; cat $OUT/bench_sar.json; tail -5 $OUT/bench_sar.err
