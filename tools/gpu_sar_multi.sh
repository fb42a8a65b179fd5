#!/bin/bash
# Voxel-sharded SAR image at N ranks: 2-GPU equality test + bench lines. usage (under gpurun --gpus N): bash tools/gpu_sar_multi.sh <tag> <n...>
TAG=${1:-sarm}; shift; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_multi.py -m gpu -x -q -k sar > $OUT/pytest_sar.log 2>&1; echo "pytest rc=$? $(tail -1 $OUT/pytest_sar.log)"
for N in $@; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --workload sar > $OUT/bench_sar_n$N.json 2> $OUT/bench_sar_n$N.err; echo "sar N=$N rc=$?"
  python -c "import json; d=json.loads(open('$OUT/bench_sar_n$N.json').readlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['e2e']['value'])" || tail -5 $OUT/bench_sar_n$N.err
done
