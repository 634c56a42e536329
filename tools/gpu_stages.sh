#!/bin/bash
# HBM-bound stages: per-config timings + one ncu --set full capture of each kernel
mkdir -p gpurun_out
timeout 600 python tools/stage_bench.py > gpurun_out/stage_bench.json 2> gpurun_out/stage_bench.err; echo "stage exit $?"
cat gpurun_out/stage_bench.json; tail -n 3 gpurun_out/stage_bench.err
NCU=1 timeout 300 python tools/stage_bench.py > gpurun_out/plain_stage.log 2>&1 &&
NCU=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"decode_filter|nms_kernel|crop_resize" -o gpurun_out/prof_stages python tools/stage_bench.py > gpurun_out/ncu_stages.log 2>&1
echo "ncu exit $?"
timeout 300 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/plain3.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"layernorm|build_crops|wrapper_head" -s 20 -c 4 -o gpurun_out/prof_ln python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_ln.log 2>&1
echo "ncu ln exit $?"
