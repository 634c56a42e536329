#!/bin/bash
# A/B on one box: LayerNorm with a persistent grid (YV_LN_PERSIST = CTAs per SM; 0 = one CTA per 8 rows), alternating runs
mkdir -p gpurun_out
YV_LN_PERSIST=4 timeout 300 python -m pytest tests/test_gpu_vit.py -q -m gpu --timeout=300 -p no:cacheprovider -x 2>&1 | tail -n 3
for i in 1 2; do
  for v in 0 4 8; do
    YV_LN_PERSIST=$v timeout 300 python bench.py --no-cpu-baseline --steps 8 > gpurun_out/ln_persist_${v}_$i.json 2> gpurun_out/ln_persist.err
    python - <<PY
import json
d=json.loads(open('gpurun_out/ln_persist_${v}_$i.json').read().strip().splitlines()[-1])
print('YV_LN_PERSIST=$v run $i', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ln', d['stages_ms_per_step']['layernorm'], 'gemm', d['stages_ms_per_step']['gemm_tcgen05'], 'attn', d['stages_ms_per_step']['attention'], d['clocks']['sm_mhz'])
PY
  done
done
tail -n 3 gpurun_out/ln_persist.err
