#!/bin/bash
# fused-LayerNorm pipeline: parity, bench line, per-launch durations of one full chunk under ncu
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vit.py -q -m gpu --timeout=300 -p no:cacheprovider -x 2>&1 | tail -n 3
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','stages_ms_per_step','e2e','clocks'): print(k, d.get(k))
PY
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 1100 -c 300 --csv --log-file gpurun_out/launches_ln.csv $CMD > gpurun_out/ncu_list.log 2>&1
python - <<'PY'
import csv
lines=[l for l in open('gpurun_out/launches_ln.csv') if not l.startswith('==')]
rows=list(csv.DictReader(lines))
g=[(r['Kernel Name'].split('(')[0][-28:], float(r['Metric Value'])/1e3) for r in rows if any(k in r['Kernel Name'] for k in ('gemm_bf16','attention','row_stats','layernorm'))]
for i,(n,t) in enumerate(g):
    if '0, 1, 1' in n and t>80 and i+8 < len(g):
        print(' | '.join(f"{n2.split('::')[-1]} {t2:.1f}" for n2,t2 in g[i:i+8]))
        break
PY
