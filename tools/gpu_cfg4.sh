#!/bin/bash
# BASELINE.json configs[3] (YOLOv8s-1280, 512 images) on one GPU: the whole batch and the 64-image shard an
# 8-GPU run gives each rank; then ncu --set full captures of the row-f1 / f2 kernels (raw-head decode, letterbox)
mkdir -p gpurun_out
timeout 900 python bench.py --workload cfg4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg4_b512.json 2> gpurun_out/bench_cfg4.err
echo "cfg4 b512 exit $?"; tail -c 2500 gpurun_out/bench_cfg4_b512.json; tail -n 5 gpurun_out/bench_cfg4.err
timeout 600 python bench.py --workload cfg4 --batch 64 --steps 5 --warmup 3 > gpurun_out/bench_cfg4_b64.json 2>> gpurun_out/bench_cfg4.err
echo "cfg4 b64 exit $?"; tail -c 2500 gpurun_out/bench_cfg4_b64.json; tail -n 5 gpurun_out/bench_cfg4.err
NCU=1 timeout 300 python tools/stage_bench.py > gpurun_out/plain_stage.log 2>&1 &&
NCU=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"decode_filter|letterbox" -o gpurun_out/prof_f1f2 python tools/stage_bench.py > gpurun_out/ncu_f1f2.log 2>&1
echo "ncu exit $?"
ncu -i gpurun_out/prof_f1f2.ncu-rep --page raw --csv --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,launch__registers_per_thread,launch__grid_size,sm__warps_active.avg.pct_of_peak_sustained_active > gpurun_out/prof_f1f2_raw.csv 2>/dev/null
head -c 3000 gpurun_out/prof_f1f2_raw.csv
