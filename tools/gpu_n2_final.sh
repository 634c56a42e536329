#!/bin/bash
# final 2-GPU record of the round-2 build + a look at the host's NUMA layout (for the N = 8 end-to-end upload bound)
mkdir -p gpurun_out
ls /sys/devices/system/node/ | grep -c "^node" ; nvidia-smi topo -m 2>&1 | head -20
for d in /sys/bus/pci/devices/*; do if [ "$(cat $d/class 2>/dev/null)" = "0x030200" ]; then echo "$d numa=$(cat $d/numa_node) cpus=$(cat $d/local_cpulist)"; fi; done
nproc; python -c "import os; print('affinity', len(os.sched_getaffinity(0)))"
bash tools/gpu_r2_n.sh 2
