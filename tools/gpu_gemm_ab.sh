#!/bin/bash
# A/B of GEMM epilogue variants on one box: per-shape timings + the whole step.  usage: gpu_gemm_ab.sh "ENV=1" "ENV2=1" ...
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg timeout 300 python tools/gemm_bench.py 2>&1 | grep -E "^(qkv|proj|fc1|fc2) " | cut -c1-70
done
for rep in 1 2; do
for cfg in "$@"; do
  env $cfg timeout 600 python bench.py --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$cfg', round(d['value'],1), round(d['ms_per_step'],2), d['stages_ms_per_step']['gemm_tcgen05'], round(d['e2e']['value'],1), d['clocks']['sm_mhz'])"
done
done
