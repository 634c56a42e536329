"""Seeded synthetic inputs for the BASELINE.json configs.

Pure torch-CPU generators shared by tests, bench.py (both arms) and the golden
writer; nothing here computes the hot path.  Seed convention: 20240410 + cfg_id.
"""
from __future__ import annotations

import torch

SEED0 = 20240410


def distinct_scores(n: int, total: int, g: torch.Generator, lo: float = 0.35, span: float = 0.6):
    """n pairwise-distinct fp32 scores from a random permutation of the grid
    lo + span*(k+0.5)/total (distinct scores make the
    reference's unstable argsort order well defined)."""
    k = torch.randperm(total, generator=g)[:n].to(torch.float64)
    return (lo + span * (k + 0.5) / total).to(torch.float32)


def make_head(B: int, nc: int = 80, A: int = 8400, n_obj: int = 32, n_jit: int = 24,
              img: int = 640, seed: int = SEED0 + 2, bg_max: float = 0.30):
    """Head tensor [B, 4+nc, A] fp32: rows cx,cy,w,h (input pixels) then class
    probabilities.  Each image gets n_obj objects (centre U[40,img-40], w,h
    U[24,200]) x n_jit jittered copies (+-6 px, size x U[.85,1.15]) placed on
    random anchors with one class prob from the distinct-score grid; all other
    probabilities are U[0,bg_max) (below conf=0.35)."""
    g = torch.Generator().manual_seed(seed)
    head = torch.empty(B, 4 + nc, A, dtype=torch.float32)
    head[:, 0:2] = torch.rand(B, 2, A, generator=g) * img
    head[:, 2:4] = torch.rand(B, 2, A, generator=g) * 100 + 8
    head[:, 4:] = torch.rand(B, nc, A, generator=g) * bg_max
    K = n_obj * n_jit
    for b in range(B):
        cxy = torch.rand(n_obj, 2, generator=g) * (img - 80) + 40
        wh = torch.rand(n_obj, 2, generator=g) * 176 + 24
        cls = torch.randint(0, nc, (n_obj,), generator=g)
        jit = (torch.rand(n_obj, n_jit, 2, generator=g) * 12 - 6)
        sc = torch.rand(n_obj, n_jit, 2, generator=g) * 0.30 + 0.85
        c = (cxy[:, None, :] + jit).reshape(K, 2)
        s = (wh[:, None, :] * sc).reshape(K, 2)
        pos = torch.randperm(A, generator=g)[:K]
        head[b, 0, pos], head[b, 1, pos] = c[:, 0], c[:, 1]
        head[b, 2, pos], head[b, 3, pos] = s[:, 0], s[:, 1]
        head[b, 4 + cls.repeat_interleave(n_jit), pos] = distinct_scores(K, A, g)
    return head


def make_images(B: int, H: int = 640, W: int = 640, seed: int = SEED0 + 3, cell: int = 32):
    """uint8 HWC RGB images [B,H,W,3]: random low-resolution colour blocks plus
    +-24 noise, so different crops give clearly different ViT inputs."""
    g = torch.Generator().manual_seed(seed)
    gh, gw = (H + cell - 1) // cell, (W + cell - 1) // cell
    low = torch.randint(0, 256, (B, gh, gw, 3), generator=g, dtype=torch.int16)
    up = low.repeat_interleave(cell, 1).repeat_interleave(cell, 2)[:, :H, :W]
    noise = torch.randint(-24, 25, (B, H, W, 3), generator=g, dtype=torch.int16)
    return (up + noise).clamp_(0, 255).to(torch.uint8)


def make_nms_stress(B: int = 32, N: int = 30000, seed: int = SEED0 + 5, field: float = 256.0):
    """cfg5: dense-overlap candidates, centres U[0,field]^2, w,h U[32,96],
    pairwise-distinct scores.  Returns boxes [B,N,4] xyxy f32, scores [B,N] f32."""
    g = torch.Generator().manual_seed(seed)
    c = torch.rand(B, N, 2, generator=g) * field
    wh = torch.rand(B, N, 2, generator=g) * 64 + 32
    boxes = torch.cat([c - 0.5 * wh, c + 0.5 * wh], dim=2).to(torch.float32)
    scores = torch.stack([distinct_scores(N, N, g, 0.0, 1.0) for _ in range(B)])
    return boxes, scores


def make_boxes_int(C: int, H: int = 640, W: int = 640, seed: int = SEED0 + 3,
                   lo: int = 32, hi: int = 320):
    """cfg3: integer boxes with w,h U[lo,hi] inside (or slightly over) the image."""
    g = torch.Generator().manual_seed(seed + 1000)
    w = torch.randint(lo, hi + 1, (C,), generator=g)
    h = torch.randint(lo, hi + 1, (C,), generator=g)
    x0 = (torch.rand(C, generator=g) * (W - w + 20) - 10).round().to(torch.int64).clamp_(0, W - 1)
    y0 = (torch.rand(C, generator=g) * (H - h + 20) - 10).round().to(torch.int64).clamp_(0, H - 1)
    return torch.stack([x0, y0, x0 + w, y0 + h], dim=1).to(torch.int32)


LEVELS_640 = [(80, 80, 8.0), (40, 40, 16.0), (20, 20, 32.0)]       # YOLOv8 @ 640: A = 8400


def make_raw_head(B: int, nc: int = 80, levels=LEVELS_640, n_obj: int = 8, seed: int = SEED0 + 6):
    """Row f1 input: raw Detect-head output [B, 64+nc, A] f32 -- DFL bin logits ~ N(0, 2) and class logits
    ~ N(-6, 1) (sigmoid << conf), with `n_obj` random anchors per image given one confident class logit
    (pairwise-distinct values in [0, 4], away from logit(0.35))."""
    g = torch.Generator().manual_seed(seed)
    A = sum(h * w for h, w, _ in levels)
    raw = torch.empty(B, 64 + nc, A, dtype=torch.float32)
    raw[:, :64] = torch.randn(B, 64, A, generator=g) * 2.0
    raw[:, 64:] = torch.randn(B, nc, A, generator=g) - 6.0
    for b in range(B):
        idx = torch.randperm(A, generator=g)[:n_obj]
        cls = torch.randint(0, nc, (n_obj,), generator=g)
        val = (torch.randperm(n_obj, generator=g).to(torch.float32) + 0.5) * (4.0 / n_obj)
        raw[b, 64 + cls, idx] = val
    return raw


def make_vit_weights(cfg=None, seed: int = 20240410, fc_gain: float = 8.0):
    """Seeded synthetic ``Network_Wrapper.state_dict()`` (keys ``model.*`` in
    timm naming + ``fc.1.*``/``fc.3.*``, utils/utils.py:59-69).  trunc-normal
    0.02 like timm's init; ``fc`` scaled by ``fc_gain`` so top-1 margins are
    well above bf16 noise (no trained weights exist in the tree).  Values are
    rounded to bf16-representable fp32 so oracle and GPU see identical weights."""
    if cfg is None:
        from .vit import ViTConfig
        cfg = ViTConfig()
    g = torch.Generator().manual_seed(seed)
    D, P = cfg.dim, cfg.patch

    def tn(*shape, std=0.02):
        t = torch.empty(*shape)
        torch.nn.init.trunc_normal_(t, std=std, a=-2 * std, b=2 * std, generator=g)
        return t.to(torch.bfloat16).to(torch.float32)

    def small(*shape, std=0.02):
        return (torch.randn(*shape, generator=g) * std).to(torch.bfloat16).to(torch.float32)

    sd = {
        "model.cls_token": tn(1, 1, D),
        "model.pos_embed": tn(1, cfg.tokens, D),
        "model.patch_embed.proj.weight": tn(D, 3, P, P),
        "model.patch_embed.proj.bias": small(D),
    }
    for i in range(cfg.depth):
        p = f"model.blocks.{i}."
        sd[p + "norm1.weight"] = (1.0 + small(D, std=0.1)).to(torch.bfloat16).to(torch.float32)
        sd[p + "norm1.bias"] = small(D)
        sd[p + "attn.qkv.weight"] = tn(3 * D, D, std=0.04)
        sd[p + "attn.qkv.bias"] = small(3 * D)
        sd[p + "attn.proj.weight"] = tn(D, D)
        sd[p + "attn.proj.bias"] = small(D)
        sd[p + "norm2.weight"] = (1.0 + small(D, std=0.1)).to(torch.bfloat16).to(torch.float32)
        sd[p + "norm2.bias"] = small(D)
        sd[p + "mlp.fc1.weight"] = tn(cfg.mlp, D)
        sd[p + "mlp.fc1.bias"] = small(cfg.mlp)
        sd[p + "mlp.fc2.weight"] = tn(D, cfg.mlp)
        sd[p + "mlp.fc2.bias"] = small(D)
    sd["model.norm.weight"] = (1.0 + small(D, std=0.1)).to(torch.bfloat16).to(torch.float32)
    sd["model.norm.bias"] = small(D)
    sd["model.head.weight"] = tn(cfg.head_out, D, std=0.04)
    sd["model.head.bias"] = small(cfg.head_out, std=0.2)
    sd["fc.1.weight"] = tn(cfg.hidden, cfg.head_out, std=0.02 * fc_gain)
    sd["fc.1.bias"] = small(cfg.hidden, std=0.1)
    sd["fc.3.weight"] = tn(cfg.num_classes, cfg.hidden, std=0.02 * fc_gain)
    sd["fc.3.bias"] = small(cfg.num_classes, std=0.1)
    return sd
