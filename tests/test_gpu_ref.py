"""GPU path (through the C-ABI) against the REFERENCE'S OWN TEXT, executed from oracle/_ref (extracted byte for
byte by oracle/make_ref.py; prebuilt files travel to the GPU box, /root/reference is not read here):

  custom_nms       README.md:62-83              keep lists bit-exact                          (row a2)
  crop_image       utils/trainClass.py:70-93    crop box + pixels bit-exact (nearest resize)  (rows a3, a4)
  Network_Wrapper  utils/utils.py:59-72         logits <= 1e-2, identical top-1               (row a5)
  make_anchors/DFL docs/YOLO_TensorRT_Technical.md:14-29,72-76   raw-head decode 1e-5         (row f1)
"""
import warnings

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import hotpath as O          # checker only
from oracle import make_ref
from yolov8_vit_b200 import ViTConfig, ViTModel, ops, synth
from yolov8_vit_b200.utils import utils as U

if not make_ref.ensure():
    pytest.skip("oracle/_ref is absent (build it with `python -m oracle.make_ref` where /root/reference exists)",
                allow_module_level=True)
from oracle import _ref as R  # noqa: E402

DEV = "cuda:0"


def test_custom_nms_vs_reference_text(golden):
    g = golden("nms")
    n_checked = 0
    for name in g["names"]:
        b, s, thr = g[f"{name}__boxes"], g[f"{name}__scores"], float(g[f"{name}__thr"])
        if len(np.unique(s)) != len(s):
            continue          # tied scores: the snippet's unstable argsort is implementation-defined (D5 pins stable)
        assert U.custom_nms(torch.from_numpy(b), torch.from_numpy(s), thr) == \
            R.custom_nms(torch.from_numpy(b), torch.from_numpy(s), thr), name
        n_checked += 1
    assert n_checked >= 9
    # cfg2 candidates of three images and a 3000-box dense case, three thresholds
    head = synth.make_head(3, nc=80, A=8400, n_obj=32, n_jit=24, seed=synth.SEED0 + 2)
    for b in range(3):
        boxes, scores, _, _ = O.decode_head(head[b], 0.35)
        assert U.custom_nms(boxes, scores, 0.45) == R.custom_nms(boxes, scores, 0.45)
    bx, sc = synth.make_nms_stress(1, 3000, seed=synth.SEED0 + 55)
    for thr in (0.3, 0.5, 0.7):
        assert U.custom_nms(bx[0], sc[0], thr) == R.custom_nms(bx[0], sc[0], thr)


def test_crop_image_vs_reference_text(tmp_path):
    """The reference's crop_image(training=False) on a file + its inference transform (A.Resize NEAREST = cv2,
    Normalize(.5, .5), HWC -> CHW; app.py:39-42, utils/trainClass.py:266) == the fused GPU crop_image."""
    import cv2
    from PIL import Image
    W, H = 300, 200
    img = synth.make_images(1, H, W, seed=5)[0].numpy()
    path = str(tmp_path / "img.png")
    Image.fromarray(img).save(path)
    for box in [(50, 40, 150, 120), (0, 0, 300, 200), (290, 190, 300, 200), (3, 5, 44, 199), (100, 100, 101, 130)]:
        ref = np.asarray(R.crop_image(path, *box, training=False))
        ref = cv2.resize(ref, (224, 224), interpolation=cv2.INTER_NEAREST).astype(np.float32)
        ref = ((ref - np.float32(127.5)) * np.float32(1.0 / 127.5)).transpose(2, 0, 1)
        got = U.crop_image(path, *box).cpu().numpy()
        assert np.array_equal(got, ref), box


def test_network_wrapper_vs_reference_text():
    """The reference's Network_Wrapper class around torchvision's vit_b_16 (fp32, CPU) vs the GPU model loaded
    from that wrapper's OWN state dict (torchvision key layout): logits within 1e-2 relative, identical top-1."""
    from torchvision.models import vit_b_16
    cfg = O.ViTConfig()
    sd = O.make_vit_weights(cfg, seed=synth.SEED0 + 41)
    net = R.Network_Wrapper(vit_b_16(weights=None), 5).eval()
    net.load_state_dict(O.timm_to_torchvision(sd), strict=True)
    m = U.Network_Wrapper(ViTModel(ViTConfig(), DEV), 5).load_state_dict(net.state_dict())
    imgs = synth.make_images(2, 640, 640, seed=6)
    boxes = synth.make_boxes_int(6, seed=7)
    x = torch.from_numpy(np.stack([O.crop_resize_norm(imgs[i % 2].numpy(), O.dilate_box(*boxes[i].tolist(), 640, 640)[:4],
                                                      224, "bilinear") for i in range(6)]))
    with torch.no_grad():
        ref = net(x)
    got = m(x.to(DEV)).cpu()
    err = float((got - ref).abs().max() / ref.abs().max())
    assert err <= 1e-2, err
    assert got.argmax(1).tolist() == ref.argmax(1).tolist()


def test_raw_head_decode_vs_reference_text():
    from types import SimpleNamespace
    levels = synth.LEVELS_640
    raw = synth.make_raw_head(3, 80, levels, n_obj=9, seed=79)
    feats = [torch.zeros(1, 1, h, w) for h, w, _ in levels]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        pts, strides = R.make_anchors(feats, torch.tensor([s for _, _, s in levels]))
    self = SimpleNamespace(reg_max=16, anchors=pts.t()[None], strides=strides.t()[None])
    ref_boxes = R.dfl_decode(self, raw[:, :64], 3)                 # [3, 4, A]
    ref_scores, ref_labels = raw[:, 64:].sigmoid().max(dim=1)
    boxes, scores, labels, anchors, counts = [t.cpu() for t in ops.decode_filter_raw(raw.to(DEV), levels, 0.35)]
    for b in range(3):
        idx = torch.nonzero(~(ref_scores[b] < 0.35)).squeeze(1)
        n = int(counts[b])
        assert anchors[b, :n].tolist() == idx.tolist() and n >= 5
        assert labels[b, :n].tolist() == ref_labels[b, idx].tolist()
        assert torch.allclose(scores[b, :n], ref_scores[b, idx], rtol=1e-5, atol=1e-6)
        assert torch.allclose(boxes[b, :n], ref_boxes[b][:, idx].t(), rtol=1e-5, atol=2e-3)
