#!/bin/bash
# full pass: GPU parity tests, smoke, both bench arms, ncu launch list, full captures of the GEMM and attention kernels
mkdir -p gpurun_out
for f in test_gpu_ops test_gpu_gemm test_gpu_vit test_wrapper; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu --timeout=300 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?"; tail -n 4 gpurun_out/$f.log
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?"; tail -n 2 gpurun_out/smoke.log
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?"
tail -c 3500 gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err; echo "ref exit $?"
tail -c 700 gpurun_out/bench_ref.json
timeout 300 python tools/gemm_bench.py 2>&1 | tee gpurun_out/gemm_bench.log | tail -n 8
timeout 300 python tools/attn_bench.py 2>&1 | tee gpurun_out/attn_bench.log | tail -n 3
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|attention" -s 40 -c 8 -o gpurun_out/prof_vit $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la gpurun_out
