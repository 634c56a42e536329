#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2d_build.log 2>&1
timeout 300 python - > gpurun_out/r2d_dbg.log 2>&1 <<PY
import torch, sys
sys.path.insert(0, '.')
from yolov8_vit_b200 import ops
for n, T in [(1, 300), (2, 785)]:
    qkv = (torch.randn(n, T, 3 * 12 * 64) * 0.7).to(torch.bfloat16).cuda()
    try:
        o = ops.attention_bf16(qkv, 12); torch.cuda.synchronize(); print("ok", n, T, float(o.float().abs().max()))
    except Exception as e:
        print("FAIL", n, T, str(e)[:100]); break
PY
head -40 gpurun_out/r2d_dbg.log | cut -c1-200
bash tools/gpu_r2_c.sh
