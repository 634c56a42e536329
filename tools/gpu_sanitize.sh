#!/bin/bash
# compute-sanitizer memcheck over the op-level parity tests and the small-ViT / pipeline tests
mkdir -p gpurun_out
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
for t in "tests/test_gpu_ops.py" "tests/test_gpu_gemm.py" "tests/test_gpu_vit.py -k 'layernorm or attention or tiny or pipeline_cfg1 or no_detections or launch_ahead'"; do
  name=$(echo $t | sed 's/[^a-z_]//g' | cut -c1-24)
  eval timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 --log-file gpurun_out/sanitize_$name.log \
    python -m pytest $t -q -m gpu -x --timeout=1400 -p no:cacheprovider > gpurun_out/sanitize_$name.out 2>&1
  echo "$t exit $?"; tail -n 3 gpurun_out/sanitize_$name.out; grep -E "ERROR SUMMARY|Invalid|out of bounds" gpurun_out/sanitize_$name.log | head -5
done
