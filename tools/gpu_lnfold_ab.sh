#!/bin/bash
# A/B of the folded LayerNorm (default) against the LayerNorm kernels (YV_VIT_LNFOLD=0): GEMM + ViT tests, bench.
mkdir -p gpurun_out
: > gpurun_out/lnf_pytest.log
for f in tests/test_gpu_gemm.py tests/test_gpu_vit.py tests/test_gpu_shapes.py; do
  echo "=== $f" >> gpurun_out/lnf_pytest.log
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 --maxfail=8 >> gpurun_out/lnf_pytest.log 2>&1
  echo "rc=$?" >> gpurun_out/lnf_pytest.log
done
grep -E "^===|passed|failed|rc=|^FAILED|Error|error" gpurun_out/lnf_pytest.log | head -40
for mode in 1 0 1 0; do
  YV_VIT_LNFOLD=$mode timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/lnf_bench_$mode.json 2> gpurun_out/lnf_bench_$mode.err
  echo "fold=$mode rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/lnf_bench_$mode.json"))
    print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], "e2e", round(d["e2e"]["value"]), "verified", d.get("verified"), d.get("verification"), d["clocks"])
except Exception as e: print("no json", e)
PY
done
