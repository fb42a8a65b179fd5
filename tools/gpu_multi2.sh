#!/bin/bash
# N-GPU session: the multi-GPU tests, the frame-sharded 6843 stream and the voxel-sharded SAR image at N ranks.
# usage (under gpurun --gpus N): bash tools/gpu_multi2.sh <n_gpus> <tag>
N=${1:-2}; TAG=${2:-multi}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $OUT/pytest_multi_n$N.log 2>&1; echo "pytest rc=$? $(tail -1 $OUT/pytest_multi_n$N.log)"
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540"
timeout 400 $TR bench.py --gpus $N > $OUT/bench_stream_n$N.json 2> $OUT/bench_stream_n$N.err; echo "stream rc=$?"
python -c "import json; d=json.loads(open('$OUT/bench_stream_n$N.json').readlines()[-1]); print(d['value'], d['ms_per_step'], d['stage_ms_per_step'], d['e2e']['value'], d['e2e'].get('h2d_ceiling'))" || tail -5 $OUT/bench_stream_n$N.err
timeout 300 $TR bench.py --gpus $N --workload sar > $OUT/bench_sar_n$N.json 2> $OUT/bench_sar_n$N.err; echo "sar rc=$?"
python -c "import json; d=json.loads(open('$OUT/bench_sar_n$N.json').readlines()[-1]); print(d['value'], d['ms_per_step'], d['roofline'], d['e2e'])" || tail -5 $OUT/bench_sar_n$N.err
