"""``YOLOTensorRT.inferdet`` of the reference (call sites app.py:16,61 / test.py:5,28; walk-through in
YOLOTensorRT_yolodet_py_解读.md:35-116), batched:

    main(Engine, imgs, device, model_list, transform, aliyunoss, func)
        -> [(image_name, class_id, conf, xmin, ymin, xmax, ymax)] sorted by image name

Per image the reference does letterbox -> blob -> Engine -> det_postprocess -> un-letterbox ->
``score < 0.35`` filter -> custom_nms -> dilation -> crop -> transform -> model_list -> argmax, one Python
iteration per detection.  Here the images of a call are grouped by size and every group goes through the
CUDA hot path (:class:`yolov8_vit_b200.HotPath`) as ONE batch, including the detector-input letterbox/blob
(row f2).  Host-side work kept as upstream: image decoding.  ``letterbox`` / ``blob`` below are the host
helpers of the upstream wrapper, kept for callers that use them directly.
"""
from __future__ import annotations

import os

import numpy as np
import torch

from .. import ops
from ..pipeline import HotPath, HotPathConfig

CLASSES = ["good", "broke", "lose", "uncovered", "circle"]          # utils/utils.py:174-181 order
# one BGR colour per class (the upstream wrapper's config.COLORS is in the un-vendored submodule: any fixed
# palette indexed like CLASSES serves, 解读.md:25)
COLORS = [(0, 200, 0), (0, 0, 230), (0, 165, 255), (255, 128, 0), (200, 0, 200)]
_SUFFIXES = (".bmp", ".dng", ".jpeg", ".jpg", ".mpo", ".png", ".tif", ".tiff", ".webp", ".pfm")


def path_to_list(images_path):
    """Single file, directory or list of files -> sorted list of image paths (解读.md:31)."""
    if isinstance(images_path, (list, tuple)):
        return list(images_path)
    if os.path.isdir(images_path):
        return sorted(os.path.join(images_path, f) for f in os.listdir(images_path) if f.lower().endswith(_SUFFIXES))
    return [images_path]


def letterbox(im: np.ndarray, new_shape=(640, 640), color=(114, 114, 114)):
    """Resize + pad keeping the aspect ratio (解读.md:67-72); returns (image, ratio, (dw, dh))."""
    import cv2
    shape = im.shape[:2]                                   # h, w
    new_w, new_h = int(new_shape[0]), int(new_shape[1])    # the wrapper passes (W, H)
    r = min(new_h / shape[0], new_w / shape[1])
    unpad = int(round(shape[1] * r)), int(round(shape[0] * r))
    dw, dh = (new_w - unpad[0]) / 2, (new_h - unpad[1]) / 2
    if shape[::-1] != unpad:
        im = cv2.resize(im, unpad, interpolation=cv2.INTER_LINEAR)
    top, bottom = int(round(dh - 0.1)), int(round(dh + 0.1))
    left, right = int(round(dw - 0.1)), int(round(dw + 0.1))
    im = cv2.copyMakeBorder(im, top, bottom, left, right, cv2.BORDER_CONSTANT, value=color)
    return im, r, (dw, dh)


def blob(im: np.ndarray) -> np.ndarray:
    """HWC uint8 RGB -> 1x3xHxW float32 in [0, 1] (解读.md:75-78)."""
    return np.ascontiguousarray(im.transpose(2, 0, 1)[None].astype(np.float32) / 255.0)


def det_postprocess(data):
    """Engine outputs ``(num_dets [1,1], bboxes [1,100,4], scores [1,100], labels [1,100])``
    (test.ipynb:20-24) -> ``(bboxes[:n], scores[:n], labels[:n])`` of the first image (解读.md:27,85-87)."""
    assert len(data) == 4
    num_dets, bboxes, scores, labels = (d[0] for d in data)
    n = int(num_dets.item()) if num_dets.numel() == 1 else int(num_dets[0].item())
    return bboxes[:n], scores[:n], labels[:n]


def draw_image(image: np.ndarray, box, cls):
    """``draw_image(image, box, cls)`` of the wrapper (imported by app.py:16; described in
    YOLOTensorRT_yolodet_py_解读.md:35-45): draws the box ``[xmin, ymin, xmax, ymax]`` with ``cv2.rectangle`` in the
    class colour and the label ``CLASSES[cls]`` followed by the fixed ``:1`` the upstream code prints, with
    ``cv2.putText`` at the box's top-left corner, IN PLACE on the BGR image.  Host-side visualisation (cv2),
    not part of the GPU path; returns the image for convenience."""
    import cv2
    cls = int(cls)
    color = COLORS[cls % len(COLORS)]
    name = CLASSES[cls] if 0 <= cls < len(CLASSES) else str(cls)
    x0, y0, x1, y1 = (int(v) for v in box)
    cv2.rectangle(image, (x0, y0), (x1, y1), color, 2)
    cv2.putText(image, f"{name}:1", (x0, max(y0 - 2, 0)), cv2.FONT_HERSHEY_SIMPLEX, 0.75, color, thickness=2)
    return image


def _model_of(net):
    """Network_Wrapper / ViTModel -> ViTModel."""
    return getattr(net, "model", net)


def main(Engine, imgs, device, model_list=None, transform=None, aliyunoss=None, func=None, *,
         conf: float = 0.35, iou: float = 0.45, names=None):
    """The wrapper's ``main`` (app.py:61, test.py:28).  ``imgs``: directory / file / list of files, or a list
    of ``(name, BGR uint8 array)`` pairs.  ``Engine``: a :class:`..models.HeadEngine`-like callable (``kind``
    'head' / 'raw' / 'nms', ``inp_info[0].shape`` = model input).  ``transform``: the reference passes an
    albumentations Compose (app.py:39-42); only its interpolation is read (INTER_NEAREST = 0 -> 'nearest';
    None -> the north star's bilinear).  ``aliyunoss`` uploads are outside the hot path and ignored;
    ``func`` (``generate_annotation``) is called per image as ``func("", name, name, objects)`` (app.py:77).
    Returns the flattened tuples sorted by image name (解读.md:103-116)."""
    import cv2
    dev = torch.device(device)
    H_in, W_in = Engine.inp_info[0].shape[-2:]
    items = []
    for it in path_to_list(imgs):
        if isinstance(it, (tuple, list)):
            name, bgr = it
        else:
            name, bgr = os.path.basename(str(it)), cv2.imread(str(it))
            if bgr is None:
                raise FileNotFoundError(f"cannot read image {it}")
        items.append((name, bgr))
    interp = "bilinear"
    if transform is not None:
        t = transform.get("valid_test", transform) if isinstance(transform, dict) else transform
        for op in getattr(t, "transforms", []):
            if getattr(op, "interpolation", None) == 0:
                interp = "nearest"
    kind = getattr(Engine, "kind", "head")
    models = [_model_of(m) for m in (model_list or [])]
    out = []
    groups = {}
    for idx, (_, bgr) in enumerate(items):
        groups.setdefault(bgr.shape[:2], []).append(idx)
    for (h0, w0), idxs in groups.items():
        # one upload of the group's BGR frames; detector input (letterbox -> RGB -> /255 -> CHW) on the GPU
        # (row f2, bit-exact against the cv2 path the wrapper runs on the host)
        bgr_d = torch.from_numpy(np.stack([items[i][1] for i in idxs])).to(dev)
        x, ratio, dwdh = ops.letterbox_blob(bgr_d, (H_in, W_in), bgr=True)
        rgb = bgr_d.flip(-1).contiguous()                # crop source: the reference crops the RGB PIL image
        hp = HotPath(models[0] if models else None,
                     HotPathConfig(conf=conf, iou=iou, interp=interp, ratio=float(ratio),
                                   dwdh=(float(dwdh[0]), float(dwdh[1]))))
        hp.extra_models = tuple(models[1:])
        res = Engine(x)
        if kind == "nms":                               # engine with the EfficientNMS plugin: <= 100 boxes / image
            num, bb, sc, lb_ = res
            bb = (bb.to(dev, torch.float32) - torch.tensor([dwdh[0], dwdh[1]] * 2, device=dev)) / float(ratio)
            sc = sc.to(dev, torch.float32)
            ok = ~(sc < conf) & (torch.arange(sc.shape[1], device=dev)[None] < num.to(dev).view(-1, 1))
            order = torch.argsort((~ok).to(torch.int8), dim=1, stable=True)       # kept entries first, order kept
            bb = torch.gather(bb, 1, order[..., None].expand(-1, -1, 4)).contiguous()
            sc = torch.gather(sc, 1, order).contiguous()
            lb_ = torch.gather(lb_.to(dev, torch.int32), 1, order.to(torch.int64)).to(torch.int32).contiguous()
            r = hp.from_candidates(bb, sc, lb_, ok.sum(1).to(torch.int32), rgb)
        elif kind == "raw":
            r = hp(res.to(dev, torch.float32), rgb, levels=Engine.levels)
        else:
            r = hp(res.to(dev, torch.float32), rgb)
        rec = r.records.cpu()
        for j, i in enumerate(idxs):
            name = items[i][0]
            objects = []
            for d in rec[j]:
                if d[7] <= 0:
                    continue
                cls_id = int(d[4])
                box = [int(d[0]), int(d[1]), int(d[2]), int(d[3])]
                out.append((name, cls_id, float(d[5]), *box))
                cname = (names or CLASSES)[cls_id] if 0 <= cls_id < len(names or CLASSES) else str(cls_id)
                objects.append({"sort": cname, "xmin": box[0], "ymin": box[1], "xmax": box[2], "ymax": box[3]})
            if func is not None:
                func("", name, name, objects)
    out.sort(key=lambda t: t[0])
    return out
