#!/bin/bash
# final round-2 profiles on one B200: ncu launch list of the bench command (first four steps are enough for the per-step
# table of tools/summarise_launches.py) and one --set full capture of the crop kernel + the ViT-B/16 attention.
# Keep gpurun_out small: it is only merged back below 64 MiB.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/fp_plain.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 420 --csv --log-file gpurun_out/fp_launches.csv $CMD > gpurun_out/fp_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"attention_pipe|crop_resize" -s 13 -c 2 -f -o gpurun_out/fp_crop_attn $CMD > gpurun_out/fp_ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out | grep fp_; du -sh gpurun_out
