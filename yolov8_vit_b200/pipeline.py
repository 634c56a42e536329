"""The batched second-stage hot path on one B200:

    head [B,4+nc,A] -> decode + conf filter (K1) -> custom_nms (K2/K3) -> round + dilate + clamp
    -> crop list (a3) -> crop + resize + normalise, patch-major bf16 (K4) -> ViT-B/16 + wrapper
    head + argmax (K5-K8) -> per-image result records (K9).

Order of operations = the detector wrapper's per-image loop
(YOLOTensorRT_yolodet_py_解读.md:64-100 + the authors' NMS / dilation / classify additions,
app.py:61), batched over images.  Every stage is a C-ABI call into libyv8vit.so.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import torch

from . import ops
from .vit import ViTConfig, ViTModel


@dataclass(frozen=True)
class HotPathConfig:
    """The reference's hard-coded constants (SURVEY.md §5 'config'): conf 0.35 (解读.md:94), iou 0.45
    (README.md:62), alpha 0.1 (README.md:106) as a rational, 224 output (app.py:40), top-100
    detections (test.ipynb:22)."""
    conf: float = 0.35
    iou: float = 0.45
    alpha: tuple = (1, 10)
    dilate_mode: str = "full"          # 'infer_half' = crop_image(training=False), utils/trainClass.py:85-91
    interp: str = "bilinear"           # 'nearest' = the reference-literal A.Resize(INTER_NEAREST), app.py:40
    out_size: int = 224
    max_det: int = 100
    ratio: float = 1.0                 # un-letterbox (解读.md:89-92)
    dwdh: tuple = (0.0, 0.0)
    cand_cap: int | None = None        # candidates kept per image after the conf filter (None = all anchors)


@dataclass
class HotPathResult:
    records: torch.Tensor              # [B,max_det,8] f32: x0,y0,x1,y1,cls,score,anchor,valid
    det_count: torch.Tensor            # [B] i32
    n_crops: int
    logits: torch.Tensor | None = None  # [n_crops, num_classes] f32
    top1: torch.Tensor | None = None    # [n_crops] i32
    crops: ops.CropList | None = None
    keep: torch.Tensor | None = None
    keep_counts: torch.Tensor | None = None
    cand: tuple | None = None


class HotPath:
    """Stateless apart from the (optional) classifier; safe to call from several threads as long
    as each uses its own CUDA stream (the reference shares module-global models across Flask
    request threads, app.py:24-37)."""

    def __init__(self, model: ViTModel | None = None, cfg: HotPathConfig = HotPathConfig()):
        self.model = model
        self.cfg = cfg
        self._pin = None

    def detect(self, head: torch.Tensor):
        """K1..K3: candidates and keep lists (device tensors, no sync)."""
        c = self.cfg
        cand = ops.decode_filter(head, c.conf, c.ratio, c.dwdh, cap=c.cand_cap)
        boxes, scores, labels, anchors, counts = cand
        keep, keep_counts = ops.nms_batched(boxes, scores, counts, c.iou, max_keep=c.max_det)
        return cand, keep, keep_counts

    def __call__(self, head: torch.Tensor, images: torch.Tensor, keep_intermediates: bool = False) -> HotPathResult:
        """head [B,4+nc,A] f32, images [B,H,W,3] u8 (both on the GPU)."""
        c = self.cfg
        B, H, W, _ = images.shape
        cand, keep, keep_counts = self.detect(head)
        boxes, scores, labels, anchors, _ = cand
        cl = ops.build_crops(boxes, scores, labels, anchors, keep, keep_counts, c.max_det, W, H, c.alpha,
                             c.dilate_mode)
        logits = top1 = None
        n = 0
        if self.model is not None:
            # The crop count sizes the ViT launch: one 4-byte read-back (the only host sync of the path).
            if self._pin is None:
                self._pin = torch.empty(1, dtype=torch.int32).pin_memory()
            self._pin.copy_(cl.n_crops, non_blocking=True)
            torch.cuda.current_stream(head.device).synchronize()
            n = min(int(self._pin.item()), cl.crop_box.shape[0])
            patches = ops.crop_resize(images, cl.crop_box, cl.crop_img, n, c.interp, c.out_size,
                                      self.model.cfg.patch, "patch")
            logits, top1 = self.model.forward_patches(patches, n)
        rec = ops.pack_results(cl, top1)
        r = HotPathResult(records=rec, det_count=cl.det_count, n_crops=n, logits=logits, top1=top1)
        if keep_intermediates:
            r.crops, r.keep, r.keep_counts, r.cand = cl, keep, keep_counts, cand
        return r


def records_to_tuples(records: torch.Tensor, names=None):
    """[(image_name, class_id, conf, xmin, ymin, xmax, ymax)] sorted by image name -- the return
    value of the wrapper's main() (YOLOTensorRT_yolodet_py_解读.md:103-116).  Host side."""
    rec = records.detach().cpu()
    B = rec.shape[0]
    names = [str(i) for i in range(B)] if names is None else list(names)
    out = []
    for b in range(B):
        for r in rec[b]:
            if r[7] <= 0:
                continue
            out.append((names[b], int(r[4]), float(r[5]), int(r[0]), int(r[1]), int(r[2]), int(r[3])))
    out.sort(key=lambda t: t[0])
    return out
