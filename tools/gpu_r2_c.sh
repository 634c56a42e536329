#!/bin/bash
# round 2, GPU call C: attention kernel: tests, race diagnostic, isolated timing fa vs legacy, bench, ncu of the fa kernel
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2c_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_vit.py -m gpu -q --timeout 600 -k "attention or b8" > gpurun_out/r2c_pytest.log 2>&1
echo "rc=$?" >> gpurun_out/r2c_pytest.log; tail -3 gpurun_out/r2c_pytest.log
timeout 600 python tools/attn_diag.py > gpurun_out/r2c_attn_diag.log 2>&1; grep -c "bad 0, differs from run0: 0" gpurun_out/r2c_attn_diag.log; grep -v "bad 0, differs from run0: 0" gpurun_out/r2c_attn_diag.log | head -20 | cut -c1-300
timeout 300 python tools/attn_bench.py > gpurun_out/r2c_attn_fa.json 2>&1; tail -1 gpurun_out/r2c_attn_fa.json
YV_ATTN=legacy timeout 300 python tools/attn_bench.py > gpurun_out/r2c_attn_legacy.json 2>&1; tail -1 gpurun_out/r2c_attn_legacy.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2c_bench_fa.json 2> gpurun_out/r2c_bench_fa.err
python - <<PY
import json
try:
    d=json.load(open("gpurun_out/r2c_bench_fa.json")); print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], round(d["e2e"]["value"]))
except Exception as e: print("no json", e); print(open("gpurun_out/r2c_bench_fa.err").read()[-1500:])
PY
NCU=1 N16=268 N8=64 timeout 900 ncu --set full --clock-control none --import-source on -k regex:attention_fa -c 2 -o gpurun_out/r2c_attn_fa_ncu -f python tools/attn_bench.py > gpurun_out/r2c_ncu.log 2>&1
tail -2 gpurun_out/r2c_ncu.log
ncu -i gpurun_out/r2c_attn_fa_ncu.ncu-rep --page raw --csv > gpurun_out/r2c_attn_fa_ncu_raw.csv 2>/dev/null
