"""Times the attention kernel alone at the bench shape (268 crops x 12 heads, T=197)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops
dev = "cuda:0"
n, T, H = int(os.environ.get('N', 268)), 197, 12
qkv = (torch.randn(n, T, 3 * H * 64, device=dev) * 0.7).to(torch.bfloat16)
NCU = bool(os.environ.get("NCU"))
for _ in range(1 if NCU else 3):
    o = ops.attention_bf16(qkv, H)
ts = []
for _ in range(1 if NCU else 10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.attention_bf16(qkv, H); e1.record(); torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ts.sort()
ms = ts[len(ts) // 2]
fl = 4.0 * n * H * T * T * 64
print(f"attention n={n} T={T}: {ms*1e3:.1f} us, {fl/ms/1e9:.1f} TFLOP/s, {ms*1e3/(n*H/148):.2f} us per item per SM")

