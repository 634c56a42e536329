"""Reference call surface of the hot path, same names / arguments / return conventions:

  custom_nms(boxes, scores, iou_threshold)      README.md:62-83 (= tech.md:72-93)
  crop_image(image, x_min, y_min, x_max, y_max) utils/trainClass.py:70-93 (+ README.md:86-124)
  Network_Wrapper(model, num_class)             utils/utils.py:59-72
  build_model(CFG, modelName, pretrained_path)  utils/utils.py:75-87

Each call runs the CUDA kernels of libyv8vit.so; CPU tensors are rejected (no fallback), except
that host inputs (PIL / numpy images, CPU box tensors) are copied to the configured device.
"""
from __future__ import annotations

import numpy as np
import torch

from .. import ops
from ..vit import ViTConfig, ViTModel

_DEFAULT_DEVICE = "cuda:0"


def _to_dev(t, dtype, device=None):
    t = torch.as_tensor(t)
    if not t.is_cuda:
        t = t.to(device or _DEFAULT_DEVICE)
    return t.to(dtype).contiguous()


def custom_nms(boxes, scores, iou_threshold=0.45):
    """Greedy score-sorted NMS; returns ``list[int]`` of original indices in descending-score
    order, exactly like the reference snippet (README.md:62-83)."""
    boxes = _to_dev(boxes, torch.float32)
    scores = _to_dev(scores, torch.float32, boxes.device)
    n = int(scores.shape[0])
    if n == 0:
        return []
    keep, cnt = ops.nms_batched(boxes.view(1, n, 4), scores.view(1, n), None, float(iou_threshold), 0)
    k = int(cnt.item())
    return keep[0, :k].tolist()


def box_dilate(x_min, y_min, x_max, y_max, width, height, alpha=(1, 10), mode="full", device=None):
    """README.md:86-124 / utils/trainClass.py:76-91 on the GPU; returns (x0, y0, x1, y1, valid)."""
    b = torch.tensor([[x_min, y_min, x_max, y_max]], dtype=torch.int32, device=device or _DEFAULT_DEVICE)
    out, valid = ops.dilate_boxes(b, int(width), int(height), alpha, mode)
    o = out[0].tolist()
    return o[0], o[1], o[2], o[3], bool(valid[0].item())


def _image_tensor(image, device):
    """PIL image / path / numpy HWC uint8 / torch tensor -> [1,H,W,3] u8 on the device."""
    if isinstance(image, str):
        from PIL import Image
        image = Image.open(image).convert("RGB")
    if not isinstance(image, (np.ndarray, torch.Tensor)):
        image = np.asarray(image.convert("RGB"))          # PIL
    t = torch.as_tensor(image)
    if t.dim() == 3:
        t = t.unsqueeze(0)
    if not t.is_cuda:
        t = t.to(device or _DEFAULT_DEVICE)
    return t.to(torch.uint8).contiguous()


def crop_image(image_path, x_min, y_min, x_max, y_max, training=False, *, interp="nearest", out_size=224,
               alpha=(1, 10), device=None):
    """``crop_image`` of utils/trainClass.py:70-93 fused with the inference transform of
    app.py:39-42: dilate (validation/test rule: ``dis // 2`` per side) + clamp + crop + resize +
    normalise.  Returns the f32 [3,out,out] CHW tensor the reference feeds to the classifier.
    ``training=True`` (random expansion, utils/trainClass.py:79-84) is out of scope."""
    if training:
        raise NotImplementedError("crop_image(training=True) is a training-time augmentation (out of scope)")
    img = _image_tensor(image_path, device)
    _, H, W, _ = img.shape
    b = torch.tensor([[x_min, y_min, x_max, y_max]], dtype=torch.int32, device=img.device)
    box, valid = ops.dilate_boxes(b, W, H, alpha, "infer_half")
    if not bool(valid[0].item()):
        raise ValueError("crop_image: empty crop after dilation + clamp")
    idx = torch.zeros(1, dtype=torch.int32, device=img.device)
    return ops.crop_resize(img, box, idx, 1, interp, out_size, 16, "chw")[0]


class Network_Wrapper:
    """``Network_Wrapper(model, num_class)``: ``forward(x) = fc(model(x))`` (utils/utils.py:59-72).
    ``model`` is a :class:`ViTModel` (or a ViTConfig / timm model name to build one)."""

    def __init__(self, model, num_class=5, device=_DEFAULT_DEVICE):
        if isinstance(model, str):
            model = ViTConfig.from_name(model, num_class)
        if isinstance(model, ViTConfig):
            model = ViTModel(model, device)
        assert model.cfg.num_classes == num_class, "num_class does not match the packed head"
        self.model = model

    def load_state_dict(self, sd, strict=True):
        self.model.load_state_dict(sd, strict)
        return self

    def to(self, device):      # the reference calls .to(CFG.device) / .eval() (app.py:36-37)
        return self

    def eval(self):
        return self

    def forward(self, x):
        return self.model.forward(x)

    __call__ = forward


def build_model(CFG, modelName=None, pretrained_path=None, pretrained=None, state_dict=None):
    """utils/utils.py:75-87.  Accepts both keyword spellings the reference's call sites use
    (``pretrained_path=`` in utils/utils.py:75, ``pretrained=`` in app.py:35 / trainClass.py:341);
    ``state_dict`` lets callers pass weights that are already in memory."""
    name = modelName or getattr(CFG, "modelName")
    cfg = ViTConfig.from_name(name, getattr(CFG, "num_classes", 5))
    net = Network_Wrapper(cfg, cfg.num_classes, getattr(CFG, "device", _DEFAULT_DEVICE))
    path = pretrained_path if pretrained_path is not None else pretrained
    if state_dict is None and path is not None:
        state_dict = torch.load(path, map_location="cpu")
    if state_dict is not None:
        net.load_state_dict(state_dict)
    return net


def classify_crops(model_list, x):
    """``for net in model_list: net(x)`` then argmax (app.py:30-37 ensembling; logits are summed)."""
    logits = None
    for net in model_list:
        l = net(x)
        logits = l if logits is None else logits + l
    return logits.argmax(dim=1), logits


def generate_annotation(folder_name, image_filename, image_path, objects_data, save_dir="train/new/"):
    """VOC-style XML writer with the reference's tag set (utils/utils.py:133-228): <annotation> with
    folder / filename / path / source.database / size (0, 0, 3 as upstream) / segmented, and one <object>
    per detection carrying <sort> (label string mapped to its id), pose, truncated, difficult, bndbox.
    Host-side result formatting (row f3).  Returns the path written (raises on I/O errors)."""
    import os
    import xml.etree.ElementTree as ET
    root = ET.Element("annotation")
    ET.SubElement(root, "folder").text = folder_name
    ET.SubElement(root, "filename").text = image_filename
    ET.SubElement(root, "path").text = image_path
    ET.SubElement(ET.SubElement(root, "source"), "database").text = "Unknown"
    size = ET.SubElement(root, "size")
    for tag, val in (("width", "0"), ("height", "0"), ("depth", "3")):
        ET.SubElement(size, tag).text = val
    ET.SubElement(root, "segmented").text = "0"
    mapping = {"good": "0", "broke": "1", "lose": "2", "loss": "2", "uncovered": "3", "circle": "4"}
    for obj in objects_data:
        node = ET.SubElement(root, "object")
        v = obj["sort"]
        ET.SubElement(node, "sort").text = str(v) if isinstance(v, int) else (
            mapping.get(v, str(v)) if isinstance(v, str) else "unknown")
        ET.SubElement(node, "pose").text = "Unspecified"
        ET.SubElement(node, "truncated").text = "0"
        ET.SubElement(node, "difficult").text = "0"
        bb = ET.SubElement(node, "bndbox")
        for tag in ("xmin", "ymin", "xmax", "ymax"):
            ET.SubElement(bb, tag).text = str(obj[tag])
    ET.indent(root, space="  ")
    os.makedirs(save_dir, exist_ok=True)
    path = os.path.join(save_dir, f"{os.path.splitext(image_filename)[0]}.xml")
    ET.ElementTree(root).write(path, encoding="utf-8", xml_declaration=False)
    return path
