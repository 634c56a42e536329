#!/bin/bash
# crop + resize kernel: thread order inside a crop, (patch, row, segment) against (output row, segment) -- YV_CROP_MAP=row
mkdir -p gpurun_out
for m in patch row; do
  echo -n "map=$m tests: "; YV_CROP_MAP=$m timeout 600 python -m pytest tests/test_gpu_ops.py tests/test_gpu_ref.py -q -m gpu --timeout=300 -p no:cacheprovider -k "crop or pipeline or letterbox" 2>&1 | tail -n 1
done
for m in patch row patch row; do
  YV_CROP_MAP=$m timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/crop_ab_$m.json 2> gpurun_out/crop_ab_$m.err
  python - <<PY
import json
d=json.load(open("gpurun_out/crop_ab_$m.json"))
print("map=$m", round(d["value"]), round(d["ms_per_step"],2), "crop_resize", d["stages_ms_per_step"]["crop_resize"], "e2e", round(d["e2e"]["value"]), "verified", d["verified"], "cfg3", d.get("extras",{}).get("cfg3"))
PY
done | tee gpurun_out/crop_ab.txt
