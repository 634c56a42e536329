#!/bin/bash
# round 2, GPU call A: full GPU test suite (one process per file, so a trapped kernel cannot poison the rest)
# + bench (fp16 stream default, fp32 stream A/B) + GEMM shapes alone
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2a_build.log 2>&1
: > gpurun_out/r2a_pytest.log
for f in tests/test_gpu_gemm.py tests/test_gpu_ops.py tests/test_gpu_ref.py tests/test_gpu_vit.py tests/test_gpu_shapes.py tests/test_wrapper.py tests/test_abi.py tests/test_dist_gloo.py; do
  echo "=== $f" >> gpurun_out/r2a_pytest.log
  timeout 1500 python -m pytest $f -m gpu -q --timeout 900 --maxfail=6 >> gpurun_out/r2a_pytest.log 2>&1
  echo "rc=$?" >> gpurun_out/r2a_pytest.log
done
grep -E "^===|passed|failed|rc=|Error|error" gpurun_out/r2a_pytest.log | head -60
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; tail -2 gpurun_out/r2a_smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?"; cut -c1-1800 gpurun_out/r2a_bench.json; tail -3 gpurun_out/r2a_bench.err
YV_VIT_STREAM=f32 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline > gpurun_out/r2a_bench_f32.json 2> gpurun_out/r2a_bench_f32.err
echo "bench f32 rc=$?"; cut -c1-600 gpurun_out/r2a_bench_f32.json
timeout 300 python tools/gemm_bench.py > gpurun_out/r2a_gemm_bench.log 2>&1
tail -3 gpurun_out/r2a_gemm_bench.log
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_tmem tools/ubench_tmem.cu && timeout 120 /tmp/ubench_tmem > gpurun_out/r2a_ubench_tmem.json 2>&1; cat gpurun_out/r2a_ubench_tmem.json
YV_NMS_CLUSTER=0 timeout 300 python tools/stage_bench.py > gpurun_out/r2a_stage_bench_c1.json 2>&1; tail -1 gpurun_out/r2a_stage_bench_c1.json | cut -c1-1200
