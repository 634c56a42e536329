#!/bin/bash
# usage: gpu_scale_n.sh N -- north-star workload (weak) and BASELINE.json configs[3] (cfg4, strong) on N GPUs of one box
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n${N}_v7.json 2> gpurun_out/bench_n${N}_v7.err
echo "north-star n$N exit $?"; tail -c 1200 gpurun_out/bench_n${N}_v7.json; tail -n 3 gpurun_out/bench_n${N}_v7.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 \
  bench.py --gpus $N --workload cfg4 --steps 5 --warmup 3 > gpurun_out/bench_cfg4_n${N}.json 2> gpurun_out/bench_cfg4_n${N}.err
echo "cfg4 n$N exit $?"; tail -c 1200 gpurun_out/bench_cfg4_n${N}.json; tail -n 3 gpurun_out/bench_cfg4_n${N}.err
