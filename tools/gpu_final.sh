#!/bin/bash
# end-of-round pass: full GPU pass on the committed build + ViT-B/8 timing + LayerNorm grid sweep
bash tools/gpu_full.sh
timeout 300 python tools/vit_b8_bench.py > gpurun_out/vit_b8_bench.json 2> gpurun_out/vit_b8_bench.err; echo "b8 exit $?"; cat gpurun_out/vit_b8_bench.json; tail -n 3 gpurun_out/vit_b8_bench.err
for v in 0 2 3 4 6; do
  YV_LN_PERSIST=$v timeout 300 python bench.py --no-cpu-baseline --steps 8 > gpurun_out/ln_sweep_$v.json 2> gpurun_out/ln_sweep.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/ln_sweep_$v.json').read().strip().splitlines()[-1])
print('YV_LN_PERSIST=$v', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ln', d['stages_ms_per_step']['layernorm'], 'gemm', d['stages_ms_per_step']['gemm_tcgen05'], d['clocks']['sm_mhz'])
PY
done
