#!/bin/bash
# compute-sanitizer memcheck + racecheck over the kernels where they earn their keep: the register-resident range-Doppler
# kernel (mbarrier ring, TMA), the OS-CFAR ladder / work lists, the tensor-core SAR pipeline and the point-cloud mirrors.
# usage (under gpurun): bash tools/sanitize.sh <tag>     -> gpurun_out/<tag>/sanitize_*.log
TAG=${1:-sanitize}; OUT=gpurun_out/$TAG; mkdir -p $OUT
SEL='os_cfar_ladder or (range_doppler_both_layouts and 6843) or tensor_core_path_matches or point_mirrors or frame_file_to_point or cfar_edge_cases'
for TOOL in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 9 --print-limit 20 \
    python -m pytest tests -m gpu -x -q -k "$SEL" > $OUT/sanitize_$TOOL.log 2>&1
  echo "$TOOL rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" $OUT/sanitize_$TOOL.log | tail -3
done
