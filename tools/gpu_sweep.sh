#!/bin/bash
# Variant sweep of the streaming range-Doppler kernel. usage (under gpurun): bash tools/gpu_sweep.sh <tag> "<variants 1843>" "<variants 6843>"
TAG=${1:-sweep}; V1=${2:-"0 1 2 3 4 5"}; V2=${3:-"0 1 2 3"}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/pytest_gpu.log
run() { # workload variant generic
  RADAR_B200_STREAM_VARIANT=$2 RADAR_B200_GENERIC=$3 timeout 300 python bench.py --workload $1 --steps 20 --warmup 3 --no-cpu-baseline --e2e-steps 1 > $OUT/b_$1_$2_$3.json 2> $OUT/b_$1_$2_$3.err
  python - $OUT/b_$1_$2_$3.json $1 $2 $3 <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d["roofline"]; s=d["stage_ms_per_step"]
    print(f"{sys.argv[2]:22s} var={sys.argv[3]} generic={sys.argv[4]} rd_ms={r['kernel_ms_per_launch']:.4f} frac={r['frac']:.3f} value={d['value']:.0f} stages={ {k: round(v,3) for k,v in s.items()} }")
except Exception as e:
    print(sys.argv[2:], "FAILED", e)
PY
}
run 1843_RadVel 0 2
for v in $V1; do run 1843_RadVel $v 0; done
run 6843_RadVel_ods_20Hz 0 2
for v in $V2; do run 6843_RadVel_ods_20Hz $v 0; done
