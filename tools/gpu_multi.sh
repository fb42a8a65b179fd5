#!/bin/bash
# Multi-GPU bench: peer-memory exchange (default) against NCCL all-gathers. usage: bash tools/gpu_multi.sh <n_gpus> [tag]
N=${1:-2}; TAG=${2:-multi}
OUT=gpurun_out/$TAG; mkdir -p $OUT
for ex in ${EXCH:-peer copy nccl}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540 \
    bench.py --gpus $N --steps 50 --warmup 5 --exchange $ex > $OUT/bench_n${N}_$ex.json 2> $OUT/bench_n${N}_$ex.err
  echo "== N=$N exchange=$ex rc=$?"
  python -c "import sys,json; d=json.loads(open('$OUT/bench_n${N}_$ex.json').readlines()[-1]); print(d['value'], d['ms_per_step'], d['stage_ms_per_step'])" || tail -5 $OUT/bench_n${N}_$ex.err
done
