#!/bin/bash
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.build()" > gpurun_out/r2d_build.log 2>&1
timeout 600 python tools/attn_diag.py > gpurun_out/r2d_attn_diag.log 2>&1; cat gpurun_out/r2d_attn_diag.log | cut -c1-400
