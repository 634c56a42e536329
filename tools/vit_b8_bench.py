"""Row f4 measurement: ViT-B/8 (785 tokens, the reference-literal `vit_base_patch8_224` of utils/class_config.py:21)
forward on N crops, CUDA-event-timed (inputs 473 MB of patches > L2), with the per-kernel breakdown of the library's
own launch profiler.  Weights: seeded random values in the Network_Wrapper.state_dict() layout (timing only here)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ViTConfig, ViTModel, _lib


def random_state_dict(cfg, seed=9):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g) * 0.02
    D, P, H = cfg.dim, cfg.patch, cfg.mlp
    sd = {"model.cls_token": r(1, 1, D), "model.pos_embed": r(1, cfg.tokens, D),
          "model.patch_embed.proj.weight": r(D, 3, P, P), "model.patch_embed.proj.bias": r(D),
          "model.norm.weight": 1 + r(D), "model.norm.bias": r(D),
          "model.head.weight": r(cfg.head_out, D), "model.head.bias": r(cfg.head_out),
          "fc.1.weight": r(cfg.hidden, cfg.head_out), "fc.1.bias": r(cfg.hidden),
          "fc.3.weight": r(cfg.num_classes, cfg.hidden), "fc.3.bias": r(cfg.num_classes)}
    for i in range(cfg.depth):
        p = f"model.blocks.{i}."
        sd.update({p + "norm1.weight": 1 + r(D), p + "norm1.bias": r(D), p + "attn.qkv.weight": r(3 * D, D),
                   p + "attn.qkv.bias": r(3 * D), p + "attn.proj.weight": r(D, D), p + "attn.proj.bias": r(D),
                   p + "norm2.weight": 1 + r(D), p + "norm2.bias": r(D), p + "mlp.fc1.weight": r(H, D),
                   p + "mlp.fc1.bias": r(H), p + "mlp.fc2.weight": r(D, H), p + "mlp.fc2.bias": r(D)})
    return sd


N = int(os.environ.get("N_CROPS", "256"))
dev = "cuda:0"
cfg = ViTConfig(patch=8)
m = ViTModel(cfg, dev).load_state_dict(random_state_dict(cfg))
T = cfg.tokens
patches = (torch.rand(N * (T - 1), 3 * 64, device=dev) * 2 - 1).to(torch.bfloat16)
for _ in range(2):
    m.forward_patches(patches, N)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
K = 3
e0.record()
for _ in range(K):
    m.forward_patches(patches, N)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / K
_lib.prof_enable(True)
m.forward_patches(patches, N)
torch.cuda.synchronize()
prof = _lib.prof_read(); _lib.prof_enable(False)
flop = 156.297e9 * N
print(json.dumps({"model": "ViT-B/8 @224 (785 tokens)", "crops": N, "ms": ms, "crops_per_s": N / ms * 1e3,
                  "tflops_effective": flop / ms / 1e9,
                  "stages_ms": {k: round(v["ms"], 3) for k, v in prof.items() if v["launches"]}}))
