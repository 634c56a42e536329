#!/bin/bash
# final-build check: every GPU test file, smoke, one bench line (no ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu --timeout=300 -p no:cacheprovider 2>&1 | tail -n 4
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -n 1
timeout 600 python bench.py > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_final.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','stages_ms_per_step','e2e','clocks','cpu_baseline','gpu_launches'): print(k, d.get(k))
print('roofline', {k: d['roofline'][k] for k in ('achieved','peak','frac','traffic')})
PY
tail -n 3 gpurun_out/bench_final.err
