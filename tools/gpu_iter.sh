#!/bin/bash
# quick iteration: GEMM + ViT parity, per-shape GEMM timing, bench line, ncu full capture of every GEMM shape
mkdir -p gpurun_out
for f in test_gpu_gemm test_gpu_vit; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu --timeout=300 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?"; tail -n 6 gpurun_out/$f.log
done
timeout 300 python tools/gemm_bench.py 2>&1 | tee gpurun_out/gemm_bench.log | tail -n 8
python bench.py --no-cpu-baseline > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_iter.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','stages_ms_per_step','roofline','e2e','clocks'): print(k, d.get(k))
PY
tail -n 5 gpurun_out/bench_iter.err
if [ -n "$DO_NCU" ]; then
NCU=1 python tools/gemm_bench.py > gpurun_out/plain_gemm.log 2>&1 &&
NCU=1 ncu --set full --clock-control none --import-source on -k regex:gemm_bf16 -c 10 -o gpurun_out/prof_gemm_shapes python tools/gemm_bench.py > gpurun_out/ncu_gemm_shapes.log 2>&1
echo "ncu exit $?"
fi
