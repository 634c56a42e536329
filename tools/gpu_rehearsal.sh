#!/bin/bash
# what the driver runs at round end, in its order: GPU tests (one pytest process, -x), smoke, reference arm, our arm
mkdir -p gpurun_out
timeout 1800 python -m pytest tests/ -x -q -m gpu > gpurun_out/reh_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/reh_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/reh_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/reh_smoke.log
timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 3 > gpurun_out/reh_ref.json 2> gpurun_out/reh_ref.err; echo "ref rc=$?"; head -c 300 gpurun_out/reh_ref.json; echo
timeout 900 python bench.py --gpus 1 --steps 20 --warmup 3 > gpurun_out/reh_bench.json 2> gpurun_out/reh_bench.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/reh_bench.json")); r=json.load(open("gpurun_out/reh_ref.json"))
print(round(d["value"]), round(d["ms_per_step"],2), "e2e", round(d["e2e"]["value"]), "ref", round(r["value"],2), "ratio e2e", round(d["e2e"]["value"]/r["value"],1), "verified", d["verified"], d["stages_ms_per_step"], d["clocks"], "launches", d["gpu_launches"])
PY
