#!/bin/bash
# BASELINE.json configs[3] on 2 GPUs: 512 images 1280x1280 batch-sharded (256 per rank), final NCCL gather of the records
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
  bench.py --gpus 2 --workload cfg4 --steps 3 --warmup 3 > gpurun_out/bench_cfg4_n2.json 2> gpurun_out/bench_cfg4_n2.err
echo "cfg4 n2 exit $?"; tail -c 2500 gpurun_out/bench_cfg4_n2.json; tail -n 5 gpurun_out/bench_cfg4_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 \
  bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2_v7.json 2> gpurun_out/bench_n2_v7.err
echo "north-star n2 exit $?"; tail -c 2500 gpurun_out/bench_n2_v7.json; tail -n 5 gpurun_out/bench_n2_v7.err
