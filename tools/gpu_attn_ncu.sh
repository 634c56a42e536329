#!/bin/bash
# ncu --set full (source-level stall sampling) of the ViT-B/16 attention alone; read with
#   ncu -i gpurun_out/prof_attn.ncu-rep --page source --csv > src.csv; python tools/ncu_stalls.py src.csv 60
mkdir -p gpurun_out
python -m yolov8_vit_b200.build > /dev/null 2>&1
NCU=1 N8=0 timeout 300 python tools/attn_bench.py > gpurun_out/plain_attn.log 2>&1 &&
NCU=1 N8=0 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention -c 2 -f -o gpurun_out/prof_attn python tools/attn_bench.py > gpurun_out/ncu_attn.log 2>&1
echo "ncu rc=$?"
