"""ViT-B/{16,8} + Network_Wrapper head on the C-ABI (yv_vit_*).

Mirrors utils/utils.py:59-87 of the reference: ``Network_Wrapper(model,
num_class)`` whose ``forward(x)`` returns ``fc(model(x))`` logits, and
``build_model(CFG, modelName, pretrained_path=...)``.  The arithmetic runs in
libyv8vit.so (tcgen05 GEMMs, bf16 operands, fp32 accumulate).
"""
from __future__ import annotations

import ctypes
import re
from dataclasses import dataclass

import torch

from . import ops
from ._lib import VitConfigC, check, load, ptr, require_device, stream_ptr


@dataclass(frozen=True)
class ViTConfig:
    """timm ``vit_base_patch{16,8}_224`` shape + the wrapper head (utils/utils.py:64-69)."""
    img: int = 224
    patch: int = 16
    dim: int = 768
    depth: int = 12
    heads: int = 12
    mlp: int = 3072
    head_out: int = 1000
    hidden: int = 128
    num_classes: int = 5
    eps: float = 1e-6

    @property
    def tokens(self) -> int:
        return (self.img // self.patch) ** 2 + 1

    @staticmethod
    def from_name(model_name: str, num_classes: int = 5) -> "ViTConfig":
        """Parses timm names like ``vit_base_patch8_224.augreg_in21k`` (utils/class_config.py:21)."""
        m = re.match(r"vit_base_patch(\d+)_(\d+)", model_name)
        if not m:
            raise ValueError(f"unsupported modelName {model_name!r}: only vit_base_patch{{8,16}}_224 variants")
        return ViTConfig(img=int(m.group(2)), patch=int(m.group(1)), num_classes=num_classes)


class ViTModel:
    """Device-resident packed weights + forward.  Not an nn.Module: inference only."""

    def __init__(self, cfg: ViTConfig = ViTConfig(), device="cuda:0"):
        self.cfg = cfg
        self.device = torch.device(device)
        require_device(torch.empty(1, device=self.device))
        c = VitConfigC(cfg.img, cfg.patch, cfg.dim, cfg.depth, cfg.heads, cfg.mlp, cfg.head_out, cfg.hidden,
                       cfg.num_classes, cfg.eps)
        h = ctypes.c_void_p()
        with torch.cuda.device(self.device):
            check(load().yv_vit_create(ctypes.byref(c), ctypes.byref(h)), "yv_vit_create")
        self._h = h
        self._ws = {}           # activation workspace per CUDA stream (threads on different streams never share one)

    def __del__(self):
        h = getattr(self, "_h", None)
        if h:
            try:
                load().yv_vit_destroy(h)
            except Exception:
                pass
            self._h = None

    def load_state_dict(self, sd: dict, strict: bool = True):
        """``sd``: Network_Wrapper.state_dict() layout (``model.*`` timm names + ``fc.1.*``/``fc.3.*``), or the
        same wrapper around ``torchvision.models.vit_b_16`` (``model.encoder.layers.encoder_layer_N...``,
        SURVEY.md §5 'checkpoint')."""
        lib = load()
        sd = torchvision_to_timm(sd)
        with torch.cuda.device(self.device):
            for k, v in sd.items():
                t = v.detach().to(device=self.device, dtype=torch.float32).contiguous()
                rc = lib.yv_vit_set_param(self._h, k.encode(), ptr(t), t.numel(), stream_ptr(self.device))
                if rc != 0 and strict:
                    check(rc, f"yv_vit_set_param({k})")
        missing = lib.yv_vit_missing_params(self._h)
        if strict and missing:
            raise KeyError(f"{missing} ViT parameters missing from the state dict")
        return self

    def last_block_cls_rows(self, enable: bool = True) -> "ViTModel":
        """Classification reads only the class-token row of a crop and no operation after the last attention mixes
        rows, so proj / norm2 / MLP of the LAST block can run on the class rows alone (``yv_vit_last_block_cls_rows``):
        logits and top-1 are bit-identical, 5.9 % of the FLOPs are not executed.  Off by default."""
        check(load().yv_vit_last_block_cls_rows(self._h, 1 if enable else 0), "yv_vit_last_block_cls_rows")
        return self

    def _workspace(self, n: int) -> torch.Tensor:
        """Workspace of the CURRENT stream: forwards on different streams (one per serving thread) run
        concurrently on separate buffers; on one stream they are ordered.  A buffer that is outgrown is
        released to the allocator of the stream it was allocated (and only ever used) on."""
        nb = load().yv_vit_workspace_bytes(self._h, n)          # sized for min(n, the library's chunk)
        key = torch.cuda.current_stream(self.device).cuda_stream
        ws = self._ws.get(key)
        if ws is None or ws.numel() < nb:
            ws = self._ws[key] = torch.empty(nb, dtype=torch.uint8, device=self.device)
        return ws

    def forward_patches(self, patches: torch.Tensor, n: int, return_hidden: bool = False,
                        out_logits: torch.Tensor | None = None, out_top1: torch.Tensor | None = None):
        """patches: layout-0 output of ops.crop_resize for n crops -> (logits [n,C] f32, top1 [n] i32).
        ``out_logits`` / ``out_top1`` (contiguous, >= n rows) receive the results in place."""
        cfg = self.cfg
        logits = torch.empty(n, cfg.num_classes, dtype=torch.float32, device=self.device) \
            if out_logits is None else out_logits
        top1 = torch.empty(n, dtype=torch.int32, device=self.device) if out_top1 is None else out_top1
        hidden = torch.empty(n, cfg.head_out, dtype=torch.float32, device=self.device) if return_hidden else None
        if n == 0:
            return (logits, top1, hidden) if return_hidden else (logits, top1)
        assert patches.dtype == torch.bfloat16 and patches.is_contiguous()
        ws = self._workspace(n)
        with torch.cuda.device(self.device):
            check(load().yv_vit_forward(self._h, ptr(patches), n, ptr(logits), ptr(top1), ptr(hidden), ptr(ws),
                                        ws.numel(), stream_ptr(self.device)), "yv_vit_forward")
        return (logits, top1, hidden) if return_hidden else (logits, top1)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: f32 [n,3,img,img] normalised crops (the reference's transform output) -> logits [n,C]."""
        require_device(x)
        p = patchify_bf16(x.to(self.device, torch.float32), self.cfg.patch)
        return self.forward_patches(p, x.shape[0])[0]

    __call__ = forward


_TV_TOP = {"model.class_token": "model.cls_token", "model.encoder.pos_embedding": "model.pos_embed",
           "model.conv_proj.weight": "model.patch_embed.proj.weight", "model.conv_proj.bias": "model.patch_embed.proj.bias",
           "model.encoder.ln.weight": "model.norm.weight", "model.encoder.ln.bias": "model.norm.bias",
           "model.heads.head.weight": "model.head.weight", "model.heads.head.bias": "model.head.bias"}
_TV_BLOCK = {"ln_1.weight": "norm1.weight", "ln_1.bias": "norm1.bias", "ln_2.weight": "norm2.weight",
             "ln_2.bias": "norm2.bias", "self_attention.in_proj_weight": "attn.qkv.weight",
             "self_attention.in_proj_bias": "attn.qkv.bias", "self_attention.out_proj.weight": "attn.proj.weight",
             "self_attention.out_proj.bias": "attn.proj.bias", "mlp.0.weight": "mlp.fc1.weight",
             "mlp.0.bias": "mlp.fc1.bias", "mlp.3.weight": "mlp.fc2.weight", "mlp.3.bias": "mlp.fc2.bias",
             # torchvision checkpoints older than 0.12 name the MLP layers linear_1 / linear_2
             "mlp.linear_1.weight": "mlp.fc1.weight", "mlp.linear_1.bias": "mlp.fc1.bias",
             "mlp.linear_2.weight": "mlp.fc2.weight", "mlp.linear_2.bias": "mlp.fc2.bias"}


def torchvision_to_timm(sd: dict) -> dict:
    """Re-keys a ``Network_Wrapper(torchvision vit_b_16)`` state dict to the timm names the packed model uses;
    a dict that is already in timm naming is returned unchanged.  A bare torchvision ViT state dict (no
    ``model.`` prefix, no ``fc.*``) is accepted too."""
    if not any(k.startswith(("model.encoder.", "encoder.", "model.conv_proj", "conv_proj")) for k in sd):
        return sd
    out = {}
    for k, v in sd.items():
        kk = k if k.startswith(("model.", "fc.")) else "model." + k
        if kk in _TV_TOP:
            out[_TV_TOP[kk]] = v
        elif kk.startswith("model.encoder.layers.encoder_layer_"):
            i, rest = kk[len("model.encoder.layers.encoder_layer_"):].split(".", 1)
            if rest not in _TV_BLOCK:
                raise KeyError(f"unknown torchvision ViT parameter {k!r}")
            out[f"model.blocks.{i}.{_TV_BLOCK[rest]}"] = v
        else:
            out[kk] = v
    return out


def patchify_bf16(x: torch.Tensor, patch: int) -> torch.Tensor:
    """[n,3,S,S] f32 -> patch-major bf16 [n*(S/p)^2, 3*p*p] (layout glue for callers that bring
    their own CHW tensors; the pipeline's crop kernel writes this layout directly)."""
    n, ch, S, _ = x.shape
    g = S // patch
    x = x.reshape(n, ch, g, patch, g, patch).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(n * g * g, ch * patch * patch).to(torch.bfloat16).contiguous()
