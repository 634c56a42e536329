#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_vit.py -q -m gpu --timeout=300 -p no:cacheprovider -k "attention or vit_b16_golden" 2>&1 | tail -n 5
python tools/attn_bench.py 2>&1 | tail -n 2
if [ -n "$DO_NCU" ]; then
NCU=1 python tools/attn_bench.py > gpurun_out/plain_attn.log 2>&1 &&
NCU=1 ncu --set full --clock-control none --import-source on -k regex:attention -c 2 -o gpurun_out/prof_attn python tools/attn_bench.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu exit $?"
fi
