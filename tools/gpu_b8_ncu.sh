#!/bin/bash
mkdir -p gpurun_out
N_CROPS=64 timeout 300 python tools/vit_b8_bench.py > gpurun_out/plain_b8.log 2>&1 &&
N_CROPS=64 timeout 600 ncu --set full --clock-control none --import-source on -k regex:"attention_long" -s 12 -c 2 -o gpurun_out/prof_attn_long python tools/vit_b8_bench.py > gpurun_out/ncu_b8.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_attn_long.ncu-rep --page raw --csv > gpurun_out/prof_attn_long_raw.csv 2>/dev/null
ncu -i gpurun_out/prof_attn_long.ncu-rep --page details > gpurun_out/prof_attn_long_details.txt 2>/dev/null
grep -E "Duration|Registers|Theoretical Occupancy|Achieved Occupancy|Executed Ipc|Issue Slots Busy|Compute \(SM\) Throughput|Memory Throughput|Shared Memory Configuration|Block Limit" gpurun_out/prof_attn_long_details.txt | head -30
