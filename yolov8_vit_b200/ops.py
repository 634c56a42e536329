"""Tensor-level wrappers of the C-ABI calls (one function per entry point).

Inputs are CUDA tensors on a B200; outputs are freshly allocated CUDA tensors.
Everything is enqueued on the current torch stream; nothing here synchronises
unless stated.  No CPU fallback.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np
import torch

from . import _lib
from ._lib import check, load, ptr, require_device, stream_ptr

INTERP = {"nearest": 0, "bilinear": 1}
DILATE_MODE = {"full": 0, "infer_half": 1}


def _ws(nbytes: int, device) -> torch.Tensor:
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


# ------------------------------------------------------------------------- a1
def decode_filter(head: torch.Tensor, conf: float = 0.35, ratio: float = 1.0, dwdh=(0.0, 0.0), cap: int | None = None):
    """head [B,4+nc,A] f32 -> (boxes [B,cap,4], scores [B,cap], labels, anchors, counts [B])."""
    require_device(head)
    assert head.dtype == torch.float32 and head.dim() == 3
    head = head.contiguous()
    B, C, A = head.shape
    nc = C - 4
    cap = A if cap is None else int(cap)
    dev = head.device
    boxes = torch.empty(B, cap, 4, dtype=torch.float32, device=dev)
    scores = torch.empty(B, cap, dtype=torch.float32, device=dev)
    labels = torch.empty(B, cap, dtype=torch.int32, device=dev)
    anchors = torch.empty(B, cap, dtype=torch.int32, device=dev)
    counts = torch.empty(B, dtype=torch.int32, device=dev)
    lib = load()
    nb = lib.yv_decode_filter_workspace_bytes(B, nc, A)
    ws = _ws(nb, dev)
    unl = int(ratio != 1.0 or dwdh[0] != 0.0 or dwdh[1] != 0.0)
    check(lib.yv_decode_filter(ptr(head), B, nc, A, float(np.float32(conf)), unl, float(np.float32(ratio)),
                               float(np.float32(dwdh[0])), float(np.float32(dwdh[1])), cap, ptr(boxes), ptr(scores),
                               ptr(labels), ptr(anchors), ptr(counts), ptr(ws), ws.numel(), stream_ptr(dev)),
          "yv_decode_filter")
    return boxes, scores, labels, anchors, counts


def decode_filter_raw(raw: torch.Tensor, levels, conf: float = 0.35, ratio: float = 1.0, dwdh=(0.0, 0.0),
                      cap: int | None = None):
    """Row f1: raw Detect-head output [B, 64+nc, A] f32 (DFL bin logits + class logits) ->
    the same outputs as :func:`decode_filter`.  ``levels`` = [(h, w, stride)] x 3."""
    import ctypes
    require_device(raw)
    assert raw.dtype == torch.float32 and raw.dim() == 3 and len(levels) == 3
    raw = raw.contiguous()
    B, C, A = raw.shape
    nc = C - 64
    assert nc >= 1 and A == sum(int(h) * int(w) for h, w, _ in levels), "levels do not match the anchor count"
    cap = A if cap is None else int(cap)
    dev = raw.device
    boxes = torch.empty(B, cap, 4, dtype=torch.float32, device=dev)
    scores = torch.empty(B, cap, dtype=torch.float32, device=dev)
    labels = torch.empty(B, cap, dtype=torch.int32, device=dev)
    anchors = torch.empty(B, cap, dtype=torch.int32, device=dev)
    counts = torch.empty(B, dtype=torch.int32, device=dev)
    lib = load()
    ws = _ws(lib.yv_decode_filter_workspace_bytes(B, nc, A), dev)
    hw = (ctypes.c_int32 * 6)(*[int(v) for h, w, _ in levels for v in (h, w)])
    st = (ctypes.c_float * 3)(*[float(s) for _, _, s in levels])
    unl = int(ratio != 1.0 or dwdh[0] != 0.0 or dwdh[1] != 0.0)
    check(lib.yv_decode_filter_raw(ptr(raw), B, nc, hw, st, float(np.float32(conf)), unl, float(np.float32(ratio)),
                                   float(np.float32(dwdh[0])), float(np.float32(dwdh[1])), cap, ptr(boxes),
                                   ptr(scores), ptr(labels), ptr(anchors), ptr(counts), ptr(ws), ws.numel(),
                                   stream_ptr(dev)), "yv_decode_filter_raw")
    return boxes, scores, labels, anchors, counts


def letterbox_geometry(H: int, W: int, out_hw=(640, 640)):
    """(ratio, (dw, dh), (unpad_w, unpad_h), (left, top)) of the wrapper's letterbox (host arithmetic)."""
    import ctypes
    r, dw, dh = ctypes.c_double(), ctypes.c_double(), ctypes.c_double()
    un, lt = (ctypes.c_int32 * 2)(), (ctypes.c_int32 * 2)()
    check(load().yv_letterbox_geometry(int(H), int(W), int(out_hw[0]), int(out_hw[1]), ctypes.byref(r),
                                       ctypes.byref(dw), ctypes.byref(dh), un, lt), "yv_letterbox_geometry")
    return r.value, (dw.value, dh.value), (un[0], un[1]), (lt[0], lt[1])


def letterbox_blob(images: torch.Tensor, out_hw=(640, 640), bgr: bool = True):
    """Row f2: images [B,H,W,3] u8 -> detector input blob [B,3,out_h,out_w] f32 (RGB, /255), letterboxed
    exactly like cv2.resize(INTER_LINEAR) + copyMakeBorder(114).  Returns (blob, ratio, (dw, dh))."""
    require_device(images)
    assert images.dtype == torch.uint8 and images.dim() == 4 and images.shape[3] == 3
    images = images.contiguous()
    B, H, W, _ = images.shape
    blob = torch.empty(B, 3, int(out_hw[0]), int(out_hw[1]), dtype=torch.float32, device=images.device)
    check(load().yv_letterbox_blob(ptr(images), B, H, W, 1 if bgr else 0, int(out_hw[0]), int(out_hw[1]), ptr(blob),
                                   stream_ptr(images.device)), "yv_letterbox_blob")
    ratio, dwdh, _, _ = letterbox_geometry(H, W, out_hw)
    return blob, ratio, dwdh


# ------------------------------------------------------------------------- a2
def nms_batched(boxes: torch.Tensor, scores: torch.Tensor, counts: torch.Tensor | None = None,
                iou_threshold: float = 0.45, max_keep: int = 0, rule: str = "custom"):
    """boxes [B,S,4] f32, scores [B,S] f32, counts [B] i32 (or None = all S) ->
    keep [B,S] i32 (original indices, descending score), keep_counts [B] i32."""
    require_device(boxes)
    assert boxes.dtype == torch.float32 and scores.dtype == torch.float32
    boxes, scores = boxes.contiguous(), scores.contiguous()
    B, S = scores.shape
    dev = boxes.device
    keep = torch.empty(B, max(S, 1), dtype=torch.int32, device=dev)
    keep_counts = torch.zeros(B, dtype=torch.int32, device=dev)
    if S == 0 or B == 0:
        return keep[:, :0], keep_counts
    lib = load()
    ws = _ws(lib.yv_nms_workspace_bytes(B, S), dev)
    if counts is not None:
        assert counts.dtype == torch.int32 and counts.is_cuda
    check(lib.yv_nms_batched_rule(ptr(boxes), ptr(scores), ptr(counts), B, S, S, float(np.float32(iou_threshold)),
                                  int(max_keep), {"custom": 0, "standard": 1}[rule], ptr(keep), ptr(keep_counts),
                                  ptr(ws), ws.numel(), stream_ptr(dev)), "yv_nms_batched")
    return keep, keep_counts


NMS_NOALLOC_BYTES = 4 << 20          # include/yv8vit.h: YV_NMS_NOALLOC_BYTES


def nms_batched_noalloc(boxes: torch.Tensor, scores: torch.Tensor, counts: torch.Tensor | None, keep: torch.Tensor,
                        keep_counts: torch.Tensor, iou_threshold: float = 0.45, max_keep: int = 0):
    """``nms_batched`` (custom_nms rule) through ``yv_nms_batched_noalloc``: the CALLER brings ``keep`` [B,S] i32 and
    ``keep_counts`` [B] i32, the scratch is the library's built-in 4 MiB device array, and nothing is allocated here
    or in the library (no counterpart in the reference; for callers that cannot allocate).  Calls on one device share
    that array: do not overlap them on different streams.  B x S x 56 B must fit the array."""
    require_device(boxes)
    assert boxes.dtype == torch.float32 and scores.dtype == torch.float32
    assert boxes.is_contiguous() and scores.is_contiguous() and keep.is_contiguous() and keep_counts.is_contiguous()
    B, S = scores.shape
    assert keep.dtype == torch.int32 and keep.shape == (B, S) and keep_counts.dtype == torch.int32 and keep_counts.shape == (B,)
    if counts is not None:
        assert counts.dtype == torch.int32 and counts.is_cuda
    check(load().yv_nms_batched_noalloc(ptr(boxes), ptr(scores), ptr(counts), B, S, S, float(np.float32(iou_threshold)),
                                        int(max_keep), ptr(keep), ptr(keep_counts), stream_ptr(boxes.device)),
          "yv_nms_batched_noalloc")
    return keep, keep_counts


def efficient_nms(boxes: torch.Tensor, scores: torch.Tensor, labels: torch.Tensor, counts: torch.Tensor | None = None,
                  iou_threshold: float = 0.65, max_det: int = 100):
    """Row f4: the engine's class-aware EfficientNMS stage (iou .65, top-100;
    docs/YOLO_TensorRT_Technical.md:108-150) = torchvision.ops.batched_nms per image: boxes of different
    classes never suppress each other (coordinate offset label * (max + 1), as torchvision does), survivor
    rule ``iou > thr`` suppresses.  boxes [B,S,4], scores [B,S], labels [B,S] -> keep [B,S] i32, counts [B]."""
    require_device(boxes)
    mx = boxes.amax(dim=(1, 2), keepdim=True) if counts is None else boxes.masked_fill(
        (torch.arange(boxes.shape[1], device=boxes.device)[None] >= counts[:, None])[..., None], float("-inf")
    ).amax(dim=(1, 2), keepdim=True)
    off = labels.to(boxes.dtype)[..., None] * (mx + 1)
    return nms_batched(boxes + off, scores, counts, iou_threshold, max_det, rule="standard")


# ------------------------------------------------------------------------- a3
def dilate_boxes(boxes_i32: torch.Tensor, W: int, H: int, alpha=(1, 10), mode: str = "full"):
    require_device(boxes_i32)
    assert boxes_i32.dtype == torch.int32
    b = boxes_i32.contiguous().view(-1, 4)
    out = torch.empty_like(b)
    valid = torch.empty(b.shape[0], dtype=torch.int32, device=b.device)
    check(load().yv_dilate_boxes(ptr(b), b.shape[0], int(W), int(H), int(alpha[0]), int(alpha[1]), DILATE_MODE[mode],
                                 ptr(out), ptr(valid), stream_ptr(b.device)), "yv_dilate_boxes")
    return out, valid


@dataclass
class CropList:
    det_box: torch.Tensor        # [B,max_det,4] i32 rounded boxes
    det_crop_box: torch.Tensor   # [B,max_det,4] i32 dilated + clamped
    det_score: torch.Tensor      # [B,max_det] f32
    det_label: torch.Tensor      # [B,max_det] i32
    det_anchor: torch.Tensor     # [B,max_det] i32
    det_count: torch.Tensor      # [B] i32
    det_crop: torch.Tensor       # [B,max_det] i32 crop index or -1
    crop_box: torch.Tensor       # [cap,4] i32
    crop_img: torch.Tensor       # [cap] i32
    crop_det: torch.Tensor       # [cap] i32
    n_crops: torch.Tensor        # [1] i32 (device)


def build_crops(boxes, scores, labels, anchors, keep, keep_counts, max_det: int, W: int, H: int,
                alpha=(1, 10), mode: str = "full", crop_cap: int | None = None) -> CropList:
    require_device(boxes)
    B, S = scores.shape
    dev = boxes.device
    cap = B * max_det if crop_cap is None else int(crop_cap)
    i32 = dict(dtype=torch.int32, device=dev)
    cl = CropList(
        det_box=torch.empty(B, max_det, 4, **i32), det_crop_box=torch.empty(B, max_det, 4, **i32),
        det_score=torch.empty(B, max_det, dtype=torch.float32, device=dev),
        det_label=torch.empty(B, max_det, **i32), det_anchor=torch.empty(B, max_det, **i32),
        det_count=torch.empty(B, **i32), det_crop=torch.empty(B, max_det, **i32),
        crop_box=torch.empty(cap, 4, **i32), crop_img=torch.empty(cap, **i32), crop_det=torch.empty(cap, **i32),
        n_crops=torch.empty(1, **i32))
    check(load().yv_build_crops(ptr(boxes), ptr(scores), ptr(labels), ptr(anchors), ptr(keep), ptr(keep_counts),
                                B, S, int(max_det), int(W), int(H), int(alpha[0]), int(alpha[1]), DILATE_MODE[mode],
                                ptr(cl.det_box), ptr(cl.det_crop_box), ptr(cl.det_score), ptr(cl.det_label),
                                ptr(cl.det_anchor), ptr(cl.det_count), ptr(cl.det_crop), ptr(cl.crop_box),
                                ptr(cl.crop_img), ptr(cl.crop_det), cap, ptr(cl.n_crops), stream_ptr(dev)),
          "yv_build_crops")
    return cl


# ------------------------------------------------------------------------- a4
def crop_resize(images: torch.Tensor, crop_box: torch.Tensor, crop_img: torch.Tensor, n: int | None = None,
                interp: str = "bilinear", out_size: int = 224, patch: int = 16, layout: str = "patch",
                n_dev: torch.Tensor | None = None, out: torch.Tensor | None = None):
    """images [B,H,W,3] u8 -> patch-major bf16 [n*(out/patch)^2, 3*patch^2] (layout='patch')
    or the reference's f32 [n,3,out,out] (layout='chw')."""
    require_device(images)
    assert images.dtype == torch.uint8 and images.dim() == 4 and images.shape[3] == 3
    images = images.contiguous()
    B, H, W, _ = images.shape
    n = crop_box.shape[0] if n is None else int(n)
    dev = images.device
    if out is None:
        if layout == "patch":
            g = out_size // patch
            out = torch.empty(n * g * g, 3 * patch * patch, dtype=torch.bfloat16, device=dev)
        else:
            out = torch.empty(n, 3, out_size, out_size, dtype=torch.float32, device=dev)
    check(load().yv_crop_resize(ptr(images), B, H, W, ptr(crop_box.contiguous()), ptr(crop_img.contiguous()), n,
                                ptr(n_dev), INTERP[interp], out_size, patch, 0 if layout == "patch" else 1, ptr(out),
                                stream_ptr(dev)), "yv_crop_resize")
    return out


# ------------------------------------------------------------------------- a5 building blocks
def gemm_bf16(a: torch.Tensor, w: torch.Tensor, bias=None, residual=None, gelu: bool = False, out_f32: bool = False,
              out: torch.Tensor | None = None):
    """a [M,K] bf16, w [N,K] bf16 -> a @ w.T (+bias) (GELU) (+residual).  Output bf16, or f32 (``out_f32``), or --
    when ``out`` is an f16 tensor -- f16 with an f16 ``residual`` that may be ``out`` itself (the ViT's in-place
    residual-stream update, TMA in both directions)."""
    require_device(a)
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16
    a, w = a.contiguous(), w.contiguous()
    M, K = a.shape
    N = w.shape[0]
    if out is not None:
        assert out.dtype == torch.float16 and out.is_contiguous() and out.shape == (M, N)
        assert residual is None or (residual.dtype == torch.float16 and residual.is_contiguous())
        d, out_type = out, 2
    else:
        assert residual is None or residual.dtype == torch.bfloat16
        d = torch.empty(M, N, dtype=torch.float32 if out_f32 else torch.bfloat16, device=a.device)
        out_type = 1 if out_f32 else 0
    check(load().yv_gemm_bf16(ptr(a), ptr(w), ptr(bias), ptr(residual), ptr(d), M, N, K, 1 if gelu else 0,
                              out_type, stream_ptr(a.device)), "yv_gemm_bf16")
    return d


def linear_ln_f16(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor,
                  eps: float = 1e-6, gelu: bool = False):
    """[GELU](Linear(LayerNorm(x))): x [M,K] f16, w [N,K] bf16 -> bf16 [M,N], the LayerNorm folded into the GEMM."""
    require_device(x)
    assert x.dtype == torch.float16 and w.dtype == torch.bfloat16
    x, w = x.contiguous(), w.contiguous()
    M, K = x.shape
    N = w.shape[0]
    lib = load()
    ws = torch.empty(int(lib.yv_linear_ln_f16_ws_bytes(M, N, K)), dtype=torch.uint8, device=x.device)
    d = torch.empty(M, N, dtype=torch.bfloat16, device=x.device)
    check(lib.yv_linear_ln_f16(ptr(x), ptr(w), ptr(bias), ptr(gamma), ptr(beta), ptr(d), M, N, K, float(eps),
                               1 if gelu else 0, ptr(ws), ws.numel(), stream_ptr(x.device)), "yv_linear_ln_f16")
    return d


def residual_linear_f16(a: torch.Tensor, w: torch.Tensor, bias: torch.Tensor | None, x: torch.Tensor,
                        want_stats: bool = True, eps: float = 1e-6):
    """x += a @ w.T + bias IN PLACE on the f16 stream x [M,N] (a bf16 [M,K], w bf16 [N,K]); returns the (mean, rstd)
    rows [M,2] f32 of the updated x computed inside the GEMM epilogue (or None)."""
    require_device(a)
    assert a.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and x.dtype == torch.float16 and x.is_contiguous()
    a, w = a.contiguous(), w.contiguous()
    M, K = a.shape
    N = w.shape[0]
    assert x.shape == (M, N)
    lib = load()
    stats = torch.empty(M, 2, dtype=torch.float32, device=a.device) if want_stats else None
    ws = _ws(lib.yv_residual_linear_f16_ws_bytes(M, N), a.device) if want_stats else None
    check(lib.yv_residual_linear_f16(ptr(a), ptr(w), ptr(bias), ptr(x), M, N, K, float(eps), ptr(stats), ptr(ws),
                                     ws.numel() if ws is not None else 0, stream_ptr(a.device)), "yv_residual_linear_f16")
    return stats


def layernorm_bf16(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-6):
    require_device(x)
    assert x.dtype == torch.bfloat16
    x = x.contiguous()
    rows, dim = x.shape
    y = torch.empty_like(x)
    check(load().yv_layernorm_bf16(ptr(x), ptr(gamma), ptr(beta), ptr(y), rows, dim, float(eps),
                                   stream_ptr(x.device)), "yv_layernorm_bf16")
    return y


def attention_bf16(qkv: torch.Tensor, heads: int):
    """qkv [n,T,3*heads*dh] bf16 -> [n,T,heads*dh] bf16."""
    require_device(qkv)
    assert qkv.dtype == torch.bfloat16
    qkv = qkv.contiguous()
    n, T, D3 = qkv.shape
    D = D3 // 3
    out = torch.empty(n, T, D, dtype=torch.bfloat16, device=qkv.device)
    check(load().yv_attention_bf16(ptr(qkv), ptr(out), n, T, heads, D // heads, stream_ptr(qkv.device)),
          "yv_attention_bf16")
    return out


def pack_results(cl: CropList, top1: torch.Tensor | None, out: torch.Tensor | None = None):
    """top1 is indexed by det_crop (crop index inside `cl`); `out` [B,max_det,8] f32 is written in place."""
    B, max_det = cl.det_score.shape
    rec = torch.empty(B, max_det, 8, dtype=torch.float32, device=cl.det_score.device) if out is None else out
    assert rec.is_contiguous() and rec.shape == (B, max_det, 8)
    check(load().yv_pack_results(ptr(cl.det_box), ptr(cl.det_score), ptr(cl.det_anchor), ptr(cl.det_count),
                                 ptr(cl.det_crop), ptr(top1), B, max_det, ptr(rec), stream_ptr(rec.device)),
          "yv_pack_results")
    return rec
