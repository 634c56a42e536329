#!/bin/bash
# clusters of 4 / 6 with W multicast (YV_GEMM_CLUSTER) against CTA pairs: GEMM parity, sustained power-capped
# throughput per shape (usage: gpu_cl4.sh [6|4])
C=${1:-6}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 -k cluster 2>&1 | tail -3
YV_GEMM_VERBOSE=1 YV_GEMM_CLUSTER=$C timeout 600 python tools/gemm_power.py 2>&1 | grep -E "yv8vit|^\{\"YV" > gpurun_out/gemm_power_cl$C.json; cat gpurun_out/gemm_power_cl$C.json
timeout 600 python tools/gemm_power.py 2>&1 | tail -1 > gpurun_out/gemm_power_cl2b.json; cat gpurun_out/gemm_power_cl2b.json
