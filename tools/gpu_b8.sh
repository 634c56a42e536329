#!/bin/bash
# long-sequence attention (ViT-B/8): parity tests + timing
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_vit.py -q -m gpu --timeout=300 -p no:cacheprovider 2>&1 | tail -n 6
timeout 300 python tools/vit_b8_bench.py > gpurun_out/vit_b8_bench_v7.json 2> gpurun_out/vit_b8_bench.err; echo "b8 exit $?"; cat gpurun_out/vit_b8_bench_v7.json; tail -n 3 gpurun_out/vit_b8_bench.err
N_CROPS=1024 timeout 300 python tools/vit_b8_bench.py > gpurun_out/vit_b8_bench_v7_1024.json 2>> gpurun_out/vit_b8_bench.err; echo "b8 exit $?"; cat gpurun_out/vit_b8_bench_v7_1024.json
