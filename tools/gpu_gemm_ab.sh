#!/bin/bash
# A/B of the GEMM epilogue warp count on one box: per-shape timings + the whole step
for w in 8 16; do
  echo "== YV_GEMM_EPI_WARPS=$w"
  YV_GEMM_EPI_WARPS=$w timeout 300 python tools/gemm_bench.py 2>&1 | grep -E "^(qkv|proj|fc1|fc2) " | cut -c1-70
done
for w in 8 16 0 8 16 0; do
  YV_GEMM_EPI_WARPS=$w timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('warps=$w', round(d['value'],1), d['ms_per_step'], d['stages_ms_per_step']['gemm_tcgen05'], d['e2e']['value'], d['clocks']['sm_mhz'])"
done
