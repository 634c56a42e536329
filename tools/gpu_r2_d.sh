#!/bin/bash
bash tools/gpu_r2_c.sh
YV_ATTN_WIDE=0 timeout 300 python tools/attn_bench.py > gpurun_out/r2c_attn_fa_nowide.json 2>&1; tail -1 gpurun_out/r2c_attn_fa_nowide.json
