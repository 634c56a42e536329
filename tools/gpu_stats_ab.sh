#!/bin/bash
# A/B of the row statistics fused into the proj / fc2 epilogues (default) against the statistics kernel
# (YV_VIT_STATS=kernel): GEMM + ViT + canary tests, bench both ways twice.
mkdir -p gpurun_out
: > gpurun_out/sf_pytest.log
for f in tests/test_gpu_gemm.py tests/test_gpu_vit.py tests/test_gpu_canary.py tests/test_gpu_shapes.py; do
  echo "=== $f" >> gpurun_out/sf_pytest.log
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 --maxfail=8 >> gpurun_out/sf_pytest.log 2>&1
  echo "rc=$?" >> gpurun_out/sf_pytest.log
done
grep -E "^===|passed|failed|rc=|^FAILED|^E  " gpurun_out/sf_pytest.log | head -40
for mode in fused kernel fused kernel; do
  YV_VIT_STATS=$mode timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/sf_bench_$mode.json 2> gpurun_out/sf_bench_$mode.err
  echo "stats=$mode rc=$?"
  python - <<PY
import json
try:
    d=json.load(open("gpurun_out/sf_bench_$mode.json"))
    print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], "e2e", round(d["e2e"]["value"]), "verified", d.get("verified"), d["verification"]["logit_err_rel_max"], d["clocks"])
except Exception as e: print("no json", e)
PY
done
