#!/bin/bash
# whole-step A/B on one box: ViT pass size of the resident path (YV_VIT_CHUNK) and of the host-buffer loop
for rep in 1 2; do
for cfg in "256 256" "2048 0"; do
  set -- $cfg
  YV_VIT_CHUNK=$1 timeout 600 python bench.py --no-cpu-baseline --e2e-vit-chunk $2 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
st=d['stages_ms_per_step']
print('chunk=$1 e2e_chunk=$2', round(d['value'],1), round(d['ms_per_step'],2), st['gemm_tcgen05'], st['layernorm'], round(d['e2e']['value'],1), d['gpu_launches'], d['clocks']['sm_mhz'])"
done
done
