#!/bin/bash
# round 2, GPU call B: flash attention kernel (tests, A/B vs the round-1 kernels), fp16 LayerNorm A/B, B/8 bench
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2b_build.log 2>&1
: > gpurun_out/r2b_pytest.log
echo "=== attention tests (fa)" >> gpurun_out/r2b_pytest.log
timeout 900 python -m pytest tests/test_gpu_vit.py -m gpu -q --timeout 600 -k "attention" >> gpurun_out/r2b_pytest.log 2>&1
echo "rc=$?" >> gpurun_out/r2b_pytest.log
for f in tests/test_gpu_vit.py tests/test_gpu_shapes.py tests/test_gpu_ref.py tests/test_wrapper.py; do
  echo "=== $f" >> gpurun_out/r2b_pytest.log
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 --maxfail=6 -k "not attention" >> gpurun_out/r2b_pytest.log 2>&1
  echo "rc=$?" >> gpurun_out/r2b_pytest.log
done
grep -E "^===|passed|failed|rc=|Error|error|assert" gpurun_out/r2b_pytest.log | head -60
for v in "fa 2" "legacy 2" "legacy 0" "legacy 1"; do set -- $v
  YV_ATTN=$1 YV_LN_F16_CTAS=$2 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2b_bench_$1_ln$2.json 2> gpurun_out/r2b_bench_$1_ln$2.err
  echo "bench attn=$1 ln_ctas=$2 rc=$?"; python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2b_bench_$1_ln$2.json")); print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], round(d["e2e"]["value"]))
except Exception as e: print("no json", e); print(open("gpurun_out/r2b_bench_$1_ln$2.err").read()[-1500:])
PY
done
timeout 600 python tools/vit_b8_bench.py > gpurun_out/r2b_vit_b8_fa.json 2>&1; tail -1 gpurun_out/r2b_vit_b8_fa.json
YV_ATTN=legacy timeout 600 python tools/vit_b8_bench.py > gpurun_out/r2b_vit_b8_legacy.json 2>&1; tail -1 gpurun_out/r2b_vit_b8_legacy.json
