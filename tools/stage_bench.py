"""Times the HBM-bound stages alone at BASELINE.json's configs (CUDA events, median of 10, L2 flushed):
cfg2 decode + custom_nms (64 x 8400 x 80), cfg5 NMS stress (32 x 30 000 boxes, thr sweep), cfg3 crop-resize
(256 crops).  Prints one JSON line; NCU=1 runs every kernel once (for ncu captures)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops, synth

dev = "cuda:0"
NCU = bool(os.environ.get("NCU"))
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timed(f, reps=10):
    for _ in range(1 if NCU else 3):
        f()
    ts = []
    for _ in range(1 if NCU else reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


out = {}
# cfg2: decode + NMS, batch 64
head = synth.make_head(64, 80, 8400, n_obj=32, n_jit=24, seed=synth.SEED0 + 2).to(dev)
ms = timed(lambda: ops.decode_filter(head, 0.35))
nbytes = head.numel() * 4
out["cfg2_decode"] = {"ms": ms, "GBs": nbytes / ms / 1e6, "bytes": nbytes}
cand = ops.decode_filter(head, 0.35)
ms = timed(lambda: ops.nms_batched(cand[0], cand[1], cand[4], 0.45))
out["cfg2_nms"] = {"ms": ms, "us_per_img": ms * 1e3 / 64, "mean_candidates": float(cand[4].float().mean())}
# bench-shape decode (batch 256)
head256 = synth.make_head(256, 80, 8400, n_obj=4, n_jit=24, seed=7).to(dev)
ms = timed(lambda: ops.decode_filter(head256, 0.35))
out["b256_decode"] = {"ms": ms, "GBs": head256.numel() * 4 / ms / 1e6}
del head256
# cfg5: NMS stress
boxes, scores = synth.make_nms_stress(32, 30000, seed=synth.SEED0 + 5)
boxes, scores = boxes.to(dev), scores.to(dev)
for thr in ([0.5] if NCU else [0.3, 0.4, 0.5, 0.6, 0.7]):
    ms = timed(lambda: ops.nms_batched(boxes, scores, None, thr), reps=5)
    keep, kc = ops.nms_batched(boxes, scores, None, thr)
    out[f"cfg5_nms_thr{thr}"] = {"ms": ms, "us_per_img": ms * 1e3 / 32, "mean_keep": float(kc.float().mean())}
# cfg3: crop + resize, 256 crops from 64 images
imgs = synth.make_images(64, 640, 640, seed=3).to(dev)
cb = synth.make_boxes_int(256, 640, 640, seed=4)
cb = cb if isinstance(cb, torch.Tensor) else cb[0]
cb = cb.to(torch.int32)
cb[:, 2].clamp_(max=640); cb[:, 3].clamp_(max=640)
cb = cb.to(dev)
ci = (torch.arange(256, device=dev, dtype=torch.int32) % 64).contiguous()
for interp in ("bilinear", "nearest"):
    ms = timed(lambda: ops.crop_resize(imgs, cb, ci, 256, interp, 224, 16, "patch"))
    wb = 256 * 224 * 224 * 3 * 2
    out[f"cfg3_crop_{interp}"] = {"ms": ms, "write_GBs": wb / ms / 1e6}
del imgs
# row f1: raw Detect-head decode (DFL + anchors fused into the filter), batch 128: the 80 class rows of every
# anchor are streamed (algorithmic bytes), the 64 DFL rows are touched for kept anchors only
raw = synth.make_raw_head(128, 80, synth.LEVELS_640, n_obj=4, seed=11).to(dev)
ms = timed(lambda: ops.decode_filter_raw(raw, synth.LEVELS_640, 0.35))
out["f1_raw_decode_b128"] = {"ms": ms, "GBs": 128 * 80 * 8400 * 4 / ms / 1e6, "tensor_GB": raw.numel() * 4 / 1e9}
del raw
# row f2: letterbox + BGR->RGB + /255 + CHW blob of the detector input, 64 frames 1080x1920 -> 640x640
frames = synth.make_images(64, 1080, 1920, seed=12).to(dev)
ms = timed(lambda: ops.letterbox_blob(frames, (640, 640), bgr=True))
wb = 64 * 3 * 640 * 640 * 4
out["f2_letterbox_64x1080p"] = {"ms": ms, "read_GB": frames.numel() / 1e9, "write_GB": wb / 1e9,
                                "GBs": (frames.numel() + wb) / ms / 1e6}
print(json.dumps(out))
