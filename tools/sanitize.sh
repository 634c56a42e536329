#!/bin/bash
# compute-sanitizer over the kernel parity tests (SURVEY.md §5 "race detection / sanitizers").
#   bash tools/sanitize.sh [memcheck|racecheck|initcheck|synccheck ...]     (default: memcheck racecheck initcheck)
# Runs on a GPU box (gpurun); writes one log per tool to gpurun_out/ and prints their summaries.  The test subset
# is the one that touches every kernel with small shapes -- the sanitizer slows kernels 10-100x, so the bench-size
# tests (tests/test_gpu_shapes.py) are left out.
set -u
mkdir -p gpurun_out
TOOLS=${*:-"memcheck racecheck initcheck"}
SEL='nms_golden or nms_cfg2 or decode_golden or decode_cfg2 or dilate_golden or crop_golden or build_crops or gemm_plain or gemm_epilogues or gemm_fp16 or folded_layernorm or test_attention or test_layernorm or vit_tiny or letterbox'
rc_all=0
for tool in $TOOLS; do
  log=gpurun_out/sanitize_${tool}.log
  timeout ${SAN_TIMEOUT:-700} compute-sanitizer --tool $tool --error-exitcode 9 --launch-timeout 600 \
      python -m pytest tests/test_gpu_ops.py tests/test_gpu_gemm.py tests/test_gpu_vit.py -m gpu -q -x -k "$SEL" > $log 2>&1
  rc=$?
  [ $rc -ne 0 ] && rc_all=$rc
  echo "== compute-sanitizer --tool $tool: exit $rc"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error|hazard" $log | sort | uniq -c | sort -rn | head -12
done
exit $rc_all
