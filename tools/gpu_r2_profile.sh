#!/bin/bash
# round 2 profiles on one B200: bench line, ncu launch list of the same command, --set full captures of the top
# kernels (GEMM with its three epilogues, ViT-B/16 attention, row statistics) and of the ViT-B/8 attention.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-extras --no-cpu-baseline"
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/r2p_bench.json")); print(round(d["value"]), round(d["ms_per_step"],2), d["stages_ms_per_step"], "e2e", round(d["e2e"]["value"]), d["roofline"], d["clocks"])
PY
timeout 300 $CMD > gpurun_out/r2p_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2p_launches.csv $CMD > gpurun_out/r2p_ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|attention|row_stats|stats_finalize" -s 30 -c 14 -o gpurun_out/r2p_vit $CMD > gpurun_out/r2p_ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 600 python tools/vit_b8_bench.py > gpurun_out/r2p_vit_b8.json 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"attention_fa" -s 2 -c 2 -o gpurun_out/r2p_b8_attn python tools/vit_b8_bench.py > gpurun_out/r2p_ncu_b8.log 2>&1
echo "ncu b8 rc=$?"; tail -1 gpurun_out/r2p_vit_b8.json
NCU=1 timeout 600 python tools/stage_bench.py > gpurun_out/r2p_stage_plain.log 2>&1 && \
NCU=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:"nms_kernel|decode_filter" -c 8 -o gpurun_out/r2p_stages python tools/stage_bench.py > gpurun_out/r2p_ncu_stages.log 2>&1
echo "ncu stages rc=$?"
ls -la gpurun_out | grep r2p
