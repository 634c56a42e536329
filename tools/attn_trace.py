"""Per-tile timeline of CTA 0 of the pipelined attention kernel (clock64 stamps, debug hook)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops, _lib
lib = _lib.load()
lib.yv_debug_attn_trace.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
n, T, H = 268, 197, 12
qkv = (torch.randn(n, T, 3 * H * 64, device="cuda:0") * 0.7).to(torch.bfloat16)
for _ in range(3):
    ops.attention_bf16(qkv, H)
lib.yv_debug_attn_trace(1, None, 0)
ops.attention_bf16(qkv, H)
buf = (ctypes.c_ulonglong * 512)()
lib.yv_debug_attn_trace(0, buf, 512)
t0 = min(v for v in buf if v)
names = ["mma:p_full", "mma:issued", "-", "-", "smx:s_full", "smx:p_arrive", "out:o_full", "out:o_free"]
print("tile " + " ".join(f"{x:>12s}" for x in names))
for t in range(44):
    print(f"{t:4d} " + " ".join(f"{(buf[t * 8 + e] - t0) if buf[t * 8 + e] else -1:12d}" for e in range(8)))
