"""Diagnostic: run attention twice on the same input, locate elements that differ between the runs and vs the fp32 reference."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops
dev = "cuda:0"
heads, dh = 12, 64
for n, T in [(60, 197), (148, 197), (300, 197), (40, 785)]:
    qkv = (torch.randn(n, T, 3 * heads * dh, generator=torch.Generator().manual_seed(3)) * 0.7).to(torch.bfloat16)
    q, k, v = qkv.float().reshape(n, T, 3, heads, dh).permute(2, 0, 3, 1, 4)
    ref = (torch.softmax(q @ k.transpose(-1, -2) * dh ** -0.5, -1) @ v).transpose(1, 2).reshape(n, T, heads * dh)
    d = qkv.to(dev)
    outs = [ops.attention_bf16(d, heads).cpu().float() for _ in range(4)]
    tol = 2e-2 * float(ref.abs().max())
    for r, o in enumerate(outs):
        err = (o - ref).abs()
        bad = err > tol
        print(f"n={n} T={T} run{r}: max err {float(err.max()):.4g} (tol {tol:.4g}), bad {int(bad.sum())}, differs from run0: {int((o != outs[0]).sum())}")
        diff = (o != outs[0]) if r else bad
        if int(diff.sum()):
            idx = diff.nonzero()
            crops = sorted(set(idx[:, 0].tolist()))[:12]
            rows = sorted(set(idx[:, 1].tolist()))
            hd = sorted(set((idx[:, 2] // 64).tolist()))
            halves = sorted(set(((idx[:, 2] % 64) // 32).tolist()))
            print("   crops", crops, "rows min/max", rows[0], rows[-1], "n_rows", len(rows), "heads", hd, "col halves", halves,
                  "max |delta| vs run0", float((o - outs[0]).abs().max()))
            # which CTA processed it: item = crop * heads + head (npair = 1 for T = 197), CTA = item % grid
            items = sorted(set((idx[:, 0] * heads + idx[:, 2] // 64).tolist()))[:16]
            grid = min(148, n * heads * (1 if T <= 256 else ((T + 127) // 128 + 1) // 2))
            print("   items", items, "-> (cta, local index)", [(it % grid, it // grid) for it in items])
