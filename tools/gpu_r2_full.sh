#!/bin/bash
# round 2 verification pass on one B200: build, every GPU test file in its own process, smoke, bench (with extras +
# cpu baseline + self-check), attention A/B, ViT-B/8 bench, stage bench.  Outputs under gpurun_out/r2f_*.
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2f_build.log 2>&1
: > gpurun_out/r2f_pytest.log
for f in tests/test_gpu_canary.py tests/test_gpu_gemm.py tests/test_gpu_ops.py tests/test_gpu_ref.py tests/test_gpu_vit.py tests/test_gpu_shapes.py tests/test_wrapper.py tests/test_zz_no_alloc.py; do
  echo "=== $f" >> gpurun_out/r2f_pytest.log
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 --maxfail=6 >> gpurun_out/r2f_pytest.log 2>&1
  echo "rc=$?" >> gpurun_out/r2f_pytest.log
done
grep -E "^===|passed|failed|rc=|^FAILED" gpurun_out/r2f_pytest.log | head -40
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2f_smoke.log 2>&1; tail -1 gpurun_out/r2f_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2f_bench.json 2> gpurun_out/r2f_bench.err
echo "bench rc=$?"; tail -2 gpurun_out/r2f_bench.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2f_bench.json"))
    print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], "e2e", round(d["e2e"]["value"]), d["e2e"].get("records_equal_resident"), "verified", d.get("verified"), d.get("verification"))
    print("roofline", d["roofline"]["achieved"], d["roofline"]["frac"], "extras", json.dumps(d.get("extras"))[:1500]); print(d.get("cpu_baseline"))
except Exception as e: print("no json", e)
PY
timeout 300 python tools/attn_bench.py > gpurun_out/r2f_attn_default.json 2>&1; tail -1 gpurun_out/r2f_attn_default.json
YV_ATTN_B16=fa timeout 300 python tools/attn_bench.py > gpurun_out/r2f_attn_b16fa.json 2>&1; tail -1 gpurun_out/r2f_attn_b16fa.json
timeout 600 python tools/vit_b8_bench.py > gpurun_out/r2f_vit_b8.json 2>&1; tail -1 gpurun_out/r2f_vit_b8.json
