#!/bin/bash
# first GPU pass: each test file in its own process (a CUDA fault in one must not hide the others)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
for f in test_gpu_ops test_gpu_gemm test_gpu_vit; do
  timeout 600 python -m pytest tests/$f.py -q -m gpu --timeout=300 -p no:cacheprovider > gpurun_out/$f.log 2>&1
  echo "$f exit $?" >> gpurun_out/summary.txt
done
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
tail -n 30 gpurun_out/test_gpu_ops.log gpurun_out/test_gpu_gemm.log gpurun_out/test_gpu_vit.log gpurun_out/smoke.log
