#!/bin/bash
# quick pass: ViT/pipeline parity tests + bench line (no ncu)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_vit.py -q -m gpu --timeout=300 -p no:cacheprovider -x 2>&1 | tail -n 15
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err; echo "bench exit $?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_quick.json').read().strip().splitlines()[-1])
for k in ('value','ms_per_step','stages_ms_per_step','roofline','e2e','clocks','nms_us_per_img'): print(k, d.get(k))
PY
tail -n 5 gpurun_out/bench_quick.err
