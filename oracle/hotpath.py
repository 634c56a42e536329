"""CPU restatement of the reference hot path (TEST INFRASTRUCTURE, see oracle/__init__.py).

Every function cites the reference location (relative to the upstream repo
Voyager0587/yolov8-vit) whose behaviour it restates.  The reference has no tests or
golden vectors for any of these stages; what pins this file instead is the reference's
OWN TEXT, extracted and executed by oracle/make_ref.py into oracle/_ref (custom_nms,
crop_image, Network_Wrapper, make_anchors + the DFL fragment) and compared with the
functions below in tests/test_ref.py, plus third-party libraries for the pieces the
reference delegates (cv2 nearest resize, torch bilinear interpolate, torchvision ViT).

Third-party arithmetic that is not in the reference tree:
  * torchvision.ops.box_iou (unpinned in requirements.txt:7; 0.26 used here)
  * cv2.resize INTER_NEAREST (unpinned; 4.13 used here)
  * timm vit_base_patch{16,8}_224 (unpinned requirements.txt:5, NOT installed):
    restated below from its published architecture (pre-norm blocks, LayerNorm
    eps 1e-6, erf-GELU, class token, learned position embedding, final norm,
    class-token pooling, linear head).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch
from torchvision.ops import box_iou

# --------------------------------------------------------------------------- a2
def custom_nms(boxes: torch.Tensor, scores: torch.Tensor, iou_threshold: float = 0.45):
    """Greedy score-sorted NMS.  Verbatim restatement of README.md:62-83
    (= tech.md:72-93) with one pinned change: ``stable=True`` on the argsort
    (ties -> lower original index first; identical to the
    snippet whenever scores are pairwise distinct).  Returns list[int] of
    original indices in descending-score order."""
    sorted_indices = torch.argsort(scores, descending=True, stable=True)
    keep = []
    while sorted_indices.size(0) > 0:
        i = sorted_indices[0].item()
        keep.append(i)
        if sorted_indices.size(0) == 1:
            break
        ious = box_iou(boxes[i:i + 1], boxes[sorted_indices[1:]])
        mask = ious.squeeze(0) < iou_threshold   # .squeeze() in the snippet; squeeze(0) keeps the 1-survivor case 1-D
        sorted_indices = sorted_indices[1:][mask]
    return keep


# --------------------------------------------------------------------------- a1
def decode_head(head: torch.Tensor, conf: float = 0.35, ratio: float = 1.0,
                dwdh=(0.0, 0.0)):
    """Decode one image's head ``[4+nc, A]`` (rows cx,cy,w,h then class probs).

    Follows docs/YOLO_TensorRT_Technical.md:243-267 (per-anchor max over class
    scores, first max wins, cxcywh -> xyxy) and the wrapper's
    ``if score < 0.35: continue`` (YOLOTensorRT_yolodet_py_解读.md:93-94, D4:
    keep iff not(score < conf)); optional un-letterbox ``(v - d) / ratio``
    (解读.md:89-92).  All arithmetic fp32.  Returns boxes[K,4] f32, scores[K]
    f32, labels[K] i32, anchor_idx[K] i32 in ascending anchor order."""
    head = head.to(torch.float32)
    cls = head[4:]
    scores, labels = cls.max(dim=0)          # torch.max returns the first max index on CPU
    keep = ~(scores < torch.tensor(conf, dtype=torch.float32))
    idx = torch.nonzero(keep).squeeze(1)
    cx, cy, w, h = head[0, idx], head[1, idx], head[2, idx], head[3, idx]
    half = torch.tensor(0.5, dtype=torch.float32)
    x0, y0 = cx - half * w, cy - half * h
    x1, y1 = cx + half * w, cy + half * h
    if ratio != 1.0 or dwdh[0] != 0.0 or dwdh[1] != 0.0:
        r = torch.tensor(ratio, dtype=torch.float32)
        dw = torch.tensor(dwdh[0], dtype=torch.float32)
        dh = torch.tensor(dwdh[1], dtype=torch.float32)
        x0, x1 = (x0 - dw) / r, (x1 - dw) / r
        y0, y1 = (y0 - dh) / r, (y1 - dh) / r
    boxes = torch.stack([x0, y0, x1, y1], dim=1)
    return boxes, scores[idx], labels[idx].to(torch.int32), idx.to(torch.int32)


# --------------------------------------------------------------------------- f1
def make_anchors(levels, grid_cell_offset: float = 0.5):
    """docs/YOLO_TensorRT_Technical.md:14-29 (make_anchors) for levels [(h, w, stride)]:
    anchor points (x + .5, y + .5) row-major per level and the stride of every anchor."""
    pts, strides = [], []
    for h, w, stride in levels:
        sx = torch.arange(w, dtype=torch.float32) + grid_cell_offset
        sy = torch.arange(h, dtype=torch.float32) + grid_cell_offset
        sy, sx = torch.meshgrid(sy, sx, indexing="ij")
        pts.append(torch.stack((sx, sy), -1).view(-1, 2))
        strides.append(torch.full((h * w, 1), float(stride), dtype=torch.float32))
    return torch.cat(pts), torch.cat(strides)


def decode_raw_head(raw: torch.Tensor, levels, reg_max: int = 16) -> torch.Tensor:
    """Detect-head decode of ONE image's raw output ``[4*reg_max+nc, A]`` -> ``[4+nc, A]`` with rows
    x0,y0,x1,y1 (pixels) then sigmoid class probabilities.  Follows the reference's PyTorch form
    (docs/YOLO_TensorRT_Technical.md:71-77): ``boxes.view(4, reg_max, A).softmax(bins) @ arange``,
    ``anchors.repeat(2) + cat(-lt, rb)``, ``* strides``; class scores ``sigmoid``
    (YOLOTensorRT_yolodet_py_解读.md:64-92).  fp32."""
    raw = raw.to(torch.float32)
    A = raw.shape[1]
    anchors, strides = make_anchors(levels)
    boxes = raw[:4 * reg_max].view(4, reg_max, A).permute(0, 2, 1)             # [4, A, reg_max]
    boxes = boxes.softmax(-1) @ torch.arange(reg_max, dtype=torch.float32)     # [4, A]
    boxes0, boxes1 = -boxes[:2], boxes[2:]
    boxes = anchors.t().repeat(2, 1) + torch.cat([boxes0, boxes1], 0)          # x0,y0,x1,y1 in cells
    boxes = boxes * strides.t()
    return torch.cat([boxes, raw[4 * reg_max:].sigmoid()], 0)


def decode_raw_filter(raw: torch.Tensor, levels, conf: float = 0.35):
    """f1 + the a1 filter on xyxy boxes: returns boxes[K,4], scores[K], labels[K] i32, anchor_idx[K] i32."""
    dec = decode_raw_head(raw, levels)
    scores, labels = dec[4:].max(dim=0)
    keep = ~(scores < torch.tensor(conf, dtype=torch.float32))
    idx = torch.nonzero(keep).squeeze(1)
    return dec[:4, idx].t().contiguous(), scores[idx], labels[idx].to(torch.int32), idx.to(torch.int32)


# --------------------------------------------------------------------------- f2
def letterbox_blob(bgr: np.ndarray, new_shape=(640, 640), color=(114, 114, 114)):
    """The wrapper's detector-input preprocessing of one BGR image (YOLOTensorRT_yolodet_py_解读.md:67-78;
    letterbox as in the upstream YOLOv8-TensorRT ``models/utils.py``): resize keeping the aspect ratio with
    cv2.INTER_LINEAR, pad to ``new_shape`` (w, h) with 114, BGR->RGB, /255, CHW.  Returns
    (blob [1,3,H,W] f32, ratio, (dw, dh)).  cv2 itself is the arithmetic (third-party, installed 4.13)."""
    import cv2
    shape = bgr.shape[:2]
    r = min(new_shape[1] / shape[0], new_shape[0] / shape[1])
    unpad = int(round(shape[1] * r)), int(round(shape[0] * r))
    dw, dh = (new_shape[0] - unpad[0]) / 2, (new_shape[1] - unpad[1]) / 2
    im = bgr
    if shape[::-1] != unpad:
        im = cv2.resize(im, unpad, interpolation=cv2.INTER_LINEAR)
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    im = cv2.copyMakeBorder(im, top, bottom, left, right, cv2.BORDER_CONSTANT, value=color)
    rgb = cv2.cvtColor(im, cv2.COLOR_BGR2RGB)
    blob = np.ascontiguousarray(rgb.transpose(2, 0, 1)[None].astype(np.float32) / 255.0)
    return blob, r, (dw, dh)


def round_boxes(boxes: torch.Tensor) -> torch.Tensor:
    """D6: float xyxy -> round half-to-even -> int32 (解读.md:97 ``bbox.round().int()``)."""
    return torch.round(boxes).to(torch.int32)


# --------------------------------------------------------------------------- a3
def dilate_box(x_min, y_min, x_max, y_max, W, H, num=1, den=10, mode="full"):
    """Integer box dilation + image clamp.  README.md:86-124 (alpha = num/den,
    clamp to [0,W]x[0,H]) in the integer form of utils/trainClass.py:76-77
    (``dis = (max - min) // 10``).  ``mode='infer_half'`` reproduces
    ``crop_image(training=False)`` (utils/trainClass.py:85-91: ``dis // 2``).
    Returns (x0, y0, x1, y1, valid)."""
    dx = ((x_max - x_min) * num) // den
    dy = ((y_max - y_min) * num) // den
    if mode == "infer_half":
        dx, dy = dx // 2, dy // 2
    x1 = min(W, x_max + dx)
    x0 = max(0, x_min - dx)
    y1 = min(H, y_max + dy)
    y0 = max(0, y_min - dy)
    return x0, y0, x1, y1, bool(x1 > x0 and y1 > y0)


# --------------------------------------------------------------------------- a4
def nearest_index_table(src: int, dst: int = 224) -> np.ndarray:
    """cv2 INTER_NEAREST source index for every destination index, the rule
    behind ``A.Resize(224,224,interpolation=cv2.INTER_NEAREST)`` (app.py:40):
    ``sx = min(floor(x * (1.0 / (dst / src))), src - 1)`` evaluated in fp64."""
    fx = float(dst) / float(src)
    ifx = 1.0 / fx
    x = np.arange(dst, dtype=np.float64)
    return np.minimum(np.floor(x * ifx).astype(np.int64), src - 1).astype(np.int32)


_INV_STD = np.float32(1.0 / 127.5)
_MEAN = np.float32(127.5)


def crop_resize_norm(image: np.ndarray, box, out: int = 224, interp: str = "nearest") -> np.ndarray:
    """Crop ``image[y0:y1, x0:x1]`` (PIL crop, utils/trainClass.py:92), resize to
    ``out x out`` (app.py:40 NEAREST; 'bilinear' = half-pixel centres, the
    north-star mode), normalise ``(p - 127.5) * (1/127.5)`` in fp32
    (A.Normalize(mean=.5,std=.5), app.py:41) and go HWC -> CHW
    (utils/trainClass.py:266).  ``image`` uint8 [H,W,3]; returns f32 [3,out,out]."""
    x0, y0, x1, y1 = [int(v) for v in box]
    crop = image[y0:y1, x0:x1]
    h, w = crop.shape[:2]
    if interp == "nearest":
        ys = nearest_index_table(h, out)
        xs = nearest_index_table(w, out)
        res = crop[ys][:, xs].astype(np.float32)
    elif interp == "bilinear":
        f32 = np.float32
        def axis(n):
            scale = f32(n) / f32(out)
            d = np.arange(out, dtype=np.float32)
            s = (d + f32(0.5)) * scale - f32(0.5)
            s = np.maximum(s, f32(0.0)).astype(np.float32)
            i0 = np.floor(s).astype(np.int32)
            frac = (s - i0.astype(np.float32)).astype(np.float32)
            i0 = np.minimum(i0, n - 1)
            i1 = np.minimum(i0 + 1, n - 1)
            return i0, i1, frac
        yi0, yi1, fy = axis(h)
        xi0, xi1, fx = axis(w)
        c = crop.astype(np.float32)
        fxv = fx[None, :, None]
        fyv = fy[:, None, None]
        r0, r1 = c[yi0], c[yi1]
        top = (r0[:, xi0] + fxv * (r0[:, xi1] - r0[:, xi0])).astype(np.float32)
        bot = (r1[:, xi0] + fxv * (r1[:, xi1] - r1[:, xi0])).astype(np.float32)
        res = (top + fyv * (bot - top)).astype(np.float32)
    else:
        raise ValueError(interp)
    res = (res - _MEAN) * _INV_STD
    return np.ascontiguousarray(res.transpose(2, 0, 1)).astype(np.float32)


def patchify(x: torch.Tensor, patch: int = 16) -> torch.Tensor:
    """[C,3,S,S] -> patch-major [C*(S/p)^2, 3*p*p], column = c*p*p + iy*p + ix
    (the flattening of the timm patch-embed conv weight [768,3,p,p])."""
    C, ch, S, _ = x.shape
    g = S // patch
    x = x.reshape(C, ch, g, patch, g, patch).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(C * g * g, ch * patch * patch)


# --------------------------------------------------------------------------- a5
@dataclass(frozen=True)
class ViTConfig:
    """timm ``vit_base_patch{16,8}_224`` shape (utils/class_config.py:21 names
    patch8; the north star pins patch16) + the wrapper head of
    utils/utils.py:59-72 (1000 -> 128 -> num_classes)."""
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
    def tokens(self):
        return (self.img // self.patch) ** 2 + 1


def make_vit_weights(cfg: ViTConfig = ViTConfig(), seed: int = 20240410, fc_gain: float = 8.0):
    """Seeded synthetic ``Network_Wrapper.state_dict()``: the shared generator lives with the other synthetic
    inputs (yolov8_vit_b200/synth.py) so that the product-side processes never import ``oracle/``."""
    from yolov8_vit_b200 import synth
    return synth.make_vit_weights(cfg, seed, fc_gain)


@torch.no_grad()
def vit_forward(sd, x: torch.Tensor, cfg: ViTConfig = ViTConfig(), return_hidden: bool = False):
    """fp32 forward of ``Network_Wrapper(timm ViT)`` (utils/utils.py:71-72:
    ``self.fc(self.model(x))``).  ``x`` f32 [C,3,224,224].  Returns logits
    [C,num_classes] (and the 1000-way backbone output if ``return_hidden``)."""
    F = torch.nn.functional
    C = x.shape[0]
    D, Hh = cfg.dim, cfg.heads
    dh = D // Hh
    w = sd["model.patch_embed.proj.weight"].reshape(D, -1)
    t = patchify(x.to(torch.float32), cfg.patch) @ w.t() + sd["model.patch_embed.proj.bias"]
    t = t.reshape(C, cfg.tokens - 1, D)
    t = torch.cat([sd["model.cls_token"].expand(C, -1, -1), t], dim=1) + sd["model.pos_embed"]
    for i in range(cfg.depth):
        p = f"model.blocks.{i}."
        y = F.layer_norm(t, (D,), sd[p + "norm1.weight"], sd[p + "norm1.bias"], cfg.eps)
        qkv = y @ sd[p + "attn.qkv.weight"].t() + sd[p + "attn.qkv.bias"]
        qkv = qkv.reshape(C, cfg.tokens, 3, Hh, dh).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        att = torch.softmax((q @ k.transpose(-1, -2)) * (dh ** -0.5), dim=-1)
        y = (att @ v).transpose(1, 2).reshape(C, cfg.tokens, D)
        t = t + (y @ sd[p + "attn.proj.weight"].t() + sd[p + "attn.proj.bias"])
        y = F.layer_norm(t, (D,), sd[p + "norm2.weight"], sd[p + "norm2.bias"], cfg.eps)
        y = F.gelu(y @ sd[p + "mlp.fc1.weight"].t() + sd[p + "mlp.fc1.bias"])
        t = t + (y @ sd[p + "mlp.fc2.weight"].t() + sd[p + "mlp.fc2.bias"])
    t = F.layer_norm(t, (D,), sd["model.norm.weight"], sd["model.norm.bias"], cfg.eps)
    hidden = t[:, 0] @ sd["model.head.weight"].t() + sd["model.head.bias"]
    y = torch.relu(hidden) @ sd["fc.1.weight"].t() + sd["fc.1.bias"]
    logits = torch.relu(y) @ sd["fc.3.weight"].t() + sd["fc.3.bias"]
    return (logits, hidden) if return_hidden else logits


def timm_to_torchvision(sd) -> dict:
    """Re-keys a ``Network_Wrapper.state_dict()`` in timm naming (``model.blocks.N...``) to the layout the same
    wrapper has around ``torchvision.models.vit_b_16`` (``model.encoder.layers.encoder_layer_N...``); timm is
    not installed, torchvision's ViT is the independent implementation of the same published architecture the
    reference's own ``Network_Wrapper`` text (oracle/_ref) is executed around.  ``fc.*`` keys pass through."""
    out = {}
    top = {"model.cls_token": "model.class_token", "model.pos_embed": "model.encoder.pos_embedding",
           "model.patch_embed.proj.weight": "model.conv_proj.weight", "model.patch_embed.proj.bias": "model.conv_proj.bias",
           "model.norm.weight": "model.encoder.ln.weight", "model.norm.bias": "model.encoder.ln.bias",
           "model.head.weight": "model.heads.head.weight", "model.head.bias": "model.heads.head.bias"}
    blk = {"norm1.weight": "ln_1.weight", "norm1.bias": "ln_1.bias", "norm2.weight": "ln_2.weight", "norm2.bias": "ln_2.bias",
           "attn.qkv.weight": "self_attention.in_proj_weight", "attn.qkv.bias": "self_attention.in_proj_bias",
           "attn.proj.weight": "self_attention.out_proj.weight", "attn.proj.bias": "self_attention.out_proj.bias",
           "mlp.fc1.weight": "mlp.0.weight", "mlp.fc1.bias": "mlp.0.bias", "mlp.fc2.weight": "mlp.3.weight", "mlp.fc2.bias": "mlp.3.bias"}
    for k, v in sd.items():
        if k in top:
            out[top[k]] = v
        elif k.startswith("model.blocks."):
            _, _, i, rest = k.split(".", 3)
            out[f"model.encoder.layers.encoder_layer_{i}.{blk[rest]}"] = v
        else:
            out[k] = v
    return out


# --------------------------------------------------------------------------- whole path
@torch.no_grad()
def pipeline_image(head, image, sd=None, cfg: ViTConfig = ViTConfig(), conf=0.35, iou=0.45,
                   alpha=(1, 10), mode="full", interp="bilinear", max_det=100,
                   ratio=1.0, dwdh=(0.0, 0.0), nms_fn=None):
    """One image through the whole second stage, in the order of the wrapper's
    per-image loop (YOLOTensorRT_yolodet_py_解读.md:64-100 + the authors' bracketed
    additions): decode + conf filter -> custom_nms -> round (D6)
    -> dilate + clamp -> crop/resize/normalise -> ViT -> argmax.  ``max_det``
    truncates the keep list in score order (engine top-100, test.ipynb:22).
    ``nms_fn`` substitutes another implementation of ``custom_nms`` (the reference's own text from oracle/_ref).
    Returns a dict of per-detection arrays (keep order = descending score)."""
    boxes, scores, labels, anchors = decode_head(head, conf, ratio, dwdh)
    keep = (nms_fn or custom_nms)(boxes, scores, iou)[:max_det]
    kb = round_boxes(boxes[keep]) if keep else torch.zeros(0, 4, dtype=torch.int32)
    H, W = image.shape[:2]
    crops, cboxes, valid = [], [], []
    for b in kb.tolist():
        x0, y0, x1, y1, ok = dilate_box(b[0], b[1], b[2], b[3], W, H, alpha[0], alpha[1], mode)
        cboxes.append([x0, y0, x1, y1])
        valid.append(ok)
        if ok:
            crops.append(crop_resize_norm(image, (x0, y0, x1, y1), cfg.img, interp))
    out = {
        "keep": np.asarray(keep, dtype=np.int32),
        "anchor": anchors[keep].numpy() if keep else np.zeros(0, np.int32),
        "score": scores[keep].numpy() if keep else np.zeros(0, np.float32),
        "label": labels[keep].numpy() if keep else np.zeros(0, np.int32),
        "box": kb.numpy(),
        "crop_box": np.asarray(cboxes, dtype=np.int32).reshape(-1, 4),
        "valid": np.asarray(valid, dtype=bool),
    }
    if sd is not None and crops:
        x = torch.from_numpy(np.stack(crops))
        logits = vit_forward(sd, x, cfg)
        out["logits"] = logits.numpy()
        out["cls"] = logits.argmax(dim=1).numpy().astype(np.int32)
    return out
