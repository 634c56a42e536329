#!/bin/bash
# strip (default) vs quarters form of the ViT-B/16 attention: all ViT / shape / canary tests on the default, attention
# tests on the other forms, whole bench step both ways twice
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_vit.py tests/test_gpu_shapes.py tests/test_gpu_canary.py tests/test_gpu_ref.py -m gpu -q --timeout 900 2>&1 | tail -3
for m in quarters split; do YV_ATTN_B16=$m timeout 600 python -m pytest tests/test_gpu_vit.py -m gpu -q --timeout 300 -k "attention" 2>&1 | tail -1; done
for m in strip quarters strip quarters; do
  YV_ATTN_B16=$m timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/attn_ab_$m.json 2> gpurun_out/attn_ab_$m.err
  python - <<PY
import json
d=json.load(open("gpurun_out/attn_ab_$m.json"))
print("$m", round(d["value"]), round(d["ms_per_step"],2), "attention", d["stages_ms_per_step"]["attention"], "gemm", d["stages_ms_per_step"]["gemm_tcgen05"], "e2e", round(d["e2e"]["value"]), "verified", d["verified"], d["verification"]["logit_err_rel_max"])
PY
done
