#!/bin/bash
# clusters of 4 with W multicast (YV_GEMM_CLUSTER=4) against CTA pairs: GEMM parity tests, sustained power-capped
# throughput per shape, whole bench step both ways.
mkdir -p gpurun_out
YV_GEMM_VERBOSE=1 YV_GEMM_CLUSTER=4 timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x --timeout 300 2>&1 | tail -8
YV_GEMM_CLUSTER=4 timeout 600 python tools/gemm_power.py 2>&1 | tail -1 > gpurun_out/gemm_power_cl4.json; cat gpurun_out/gemm_power_cl4.json
timeout 600 python tools/gemm_power.py 2>&1 | tail -1 > gpurun_out/gemm_power_cl2.json; cat gpurun_out/gemm_power_cl2.json
