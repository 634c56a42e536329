"""Attention alone: CUDA-event timing of ops.attention_bf16 at the two shapes the ViTs use (ViT-B/16: T = 197, ViT-B/8:
T = 785), L2 flushed between reps.  YV_ATTN_B16=fa runs the ViT-B/16 shape on the flash-style kernel too.  NCU=1: one launch per shape.
Reports ms, TFLOP/s (4 T^2 dh flop per (crop, head)) and the fraction of the MUFU floor (16 ex2 per clock per SM at the
SM clock sampled by nvidia-smi is not known here, so the floor is given in clocks per SM instead)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops

dev = "cuda:0"
NCU = bool(os.environ.get("NCU"))
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
out = {"b16_kernel": "attention_fa (WIDE)" if os.environ.get("YV_ATTN_B16", "")[:1] == "f" else "attention_pipe", "b8_kernel": "attention_fa"}
shapes = [("b16", int(os.environ.get("N16", "1072")), 197), ("b8", int(os.environ.get("N8", "256")), 785)]
shapes = [s for s in shapes if s[1] > 0]            # N8=0 skips the ViT-B/8 shape
for name, n, T in shapes:
    heads, dh = 12, 64
    qkv = (torch.randn(n, T, 3 * heads * dh, device=dev) * 0.7).to(torch.bfloat16)
    for _ in range(1 if NCU else 3):
        o = ops.attention_bf16(qkv, heads)
    ts = []
    for _ in range(1 if NCU else 10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.attention_bf16(qkv, heads); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    t16 = (T + 15) // 16 * 16
    live_quarters = sum(1 for r0 in range(0, (T + 127) // 128 * 128, 32) if r0 < T)       # 32-row warps with a live row
    exps_per_sm = n * heads * live_quarters * 32 * t16 / 148
    out[name] = {"crops": n, "T": T, "ms": ms, "tflops": 4.0 * n * heads * T * T * dh / ms / 1e9,
                 "mufu_floor_clocks_per_sm": exps_per_sm / 16, "clocks_per_sm_at_1p5GHz": ms * 1.5e6}
    del qkv, o
print(json.dumps(out))
