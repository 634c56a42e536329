#!/bin/bash
# usage: gpu_r2_n.sh N -- round 2 multi-GPU check: north-star line on N GPUs (weak; with the cfg4_strong extra, per-rank
# times and the gather timing) and the reference arm launched the same way (rank 0 alone works).
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_n${N}.json 2> gpurun_out/r2_bench_n${N}.err
echo "north-star n$N exit $?"; tail -n 5 gpurun_out/r2_bench_n${N}.err; wc -c gpurun_out/r2_bench_n${N}.json; head -c 200 gpurun_out/r2_bench_n${N}.json; echo
python - <<PY
import json
d=json.loads(open("gpurun_out/r2_bench_n${N}.json").read().strip().splitlines()[-1])
print(round(d["value"]), round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), d.get("per_rank_ms_per_step"), "gather", d.get("gather_records_ms"), "cfg4", d.get("cfg4_strong"))
print(json.dumps(d.get("extras"))[:1200])
PY
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29543 \
  bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r2_bench_ref_n${N}.json 2> gpurun_out/r2_bench_ref_n${N}.err
echo "reference n$N exit $?"; head -c 700 gpurun_out/r2_bench_ref_n${N}.json
