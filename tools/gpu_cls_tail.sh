#!/bin/bash
# last ViT block on the class-token rows only (opt-in): bit-identity test + the bench line with its extra
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_vit.py -q -m gpu --timeout=600 -p no:cacheprovider -k "last_block or chunking or b16_golden" 2>&1 | tail -n 3
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/cls_tail_bench.json 2> gpurun_out/cls_tail_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/cls_tail_bench.json"))
print(round(d["value"]), round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "verified", d["verified"], d["extras"]["last_block_cls_rows"], d["extras"]["cfg3_crop_vit"])
PY
tail -3 gpurun_out/cls_tail_bench.err
