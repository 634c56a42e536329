"""Detector-wrapper mirror (yolov8_vit_b200/YOLOTensorRT) and result formatting: host logic on CPU,
``main`` end to end on the GPU against the oracle pipeline."""
import os

import numpy as np
import pytest
import torch

from oracle import hotpath as O          # checker only
from yolov8_vit_b200 import synth
from yolov8_vit_b200.YOLOTensorRT import inferdet
from yolov8_vit_b200.YOLOTensorRT.models import HeadEngine, TRTModule
from yolov8_vit_b200.utils import generate_annotation


def test_letterbox_blob_and_postprocess():
    im = np.full((480, 640, 3), 200, np.uint8)
    lb, r, (dw, dh) = inferdet.letterbox(im, (640, 640))
    assert lb.shape == (640, 640, 3) and r == 1.0 and (dw, dh) == (0.0, 80.0)
    assert (lb[:80] == 114).all() and (lb[80:560] == 200).all() and (lb[560:] == 114).all()
    b = inferdet.blob(lb)
    assert b.shape == (1, 3, 640, 640) and b.dtype == np.float32 and abs(float(b[0, 0, 100, 0]) - 200 / 255) < 1e-7
    lb2, r2, (dw2, dh2) = inferdet.letterbox(np.zeros((1280, 960, 3), np.uint8), (640, 640))
    assert lb2.shape == (640, 640, 3) and r2 == 0.5 and (dw2, dh2) == (80.0, 0.0)
    data = (torch.tensor([[2]]), torch.arange(400.).view(1, 100, 4), torch.rand(1, 100), torch.zeros(1, 100, dtype=torch.int32))
    bb, sc, lab = inferdet.det_postprocess(data)
    assert bb.shape == (2, 4) and sc.shape == (2,) and lab.shape == (2,)


def test_path_to_list_and_trtmodule(tmp_path):
    for n in ("b.jpg", "a.png", "note.txt"):
        (tmp_path / n).write_bytes(b"x")
    got = inferdet.path_to_list(str(tmp_path))
    assert [os.path.basename(p) for p in got] == ["a.png", "b.jpg"]
    assert inferdet.path_to_list("x.jpg") == ["x.jpg"] and inferdet.path_to_list(["p", "q"]) == ["p", "q"]
    with pytest.raises((RuntimeError, NotImplementedError)):
        TRTModule("best.engine", "cuda:0")


def test_draw_image_draws_in_place():
    """draw_image(image, box, cls) (app.py:16 imports it; 解读.md:35-45): rectangle + label in the class colour."""
    from yolov8_vit_b200.YOLOTensorRT import draw_image, COLORS
    img = np.zeros((120, 160, 3), np.uint8)
    out = draw_image(img, [20, 40, 100, 90], 1)
    assert out is img
    assert tuple(img[40, 60]) == COLORS[1] and tuple(img[90, 60]) == COLORS[1]       # top / bottom edges
    assert tuple(img[60, 20]) == COLORS[1] and tuple(img[60, 100]) == COLORS[1]      # left / right edges
    assert (img[44:86, 24:96] == 0).all()                                            # interior untouched
    assert (img[15:39, 20:120] == np.array(COLORS[1])).all(-1).any()                 # label text above the box


def test_torchvision_state_dict_keys_map_to_timm_names():
    from yolov8_vit_b200.vit import torchvision_to_timm
    sd = O.make_vit_weights(O.ViTConfig(depth=2), seed=3)
    tv = O.timm_to_torchvision(sd)
    assert "model.encoder.layers.encoder_layer_1.self_attention.in_proj_weight" in tv and "fc.3.bias" in tv
    back = torchvision_to_timm(tv)
    assert set(back) == set(sd) and all(back[k] is sd[k] for k in sd)
    assert torchvision_to_timm(sd) is sd                                             # already timm naming
    bare = {k[len("model."):]: v for k, v in tv.items() if k.startswith("model.")}   # a plain torchvision ViT
    assert set(torchvision_to_timm(bare)) == {k for k in sd if k.startswith("model.")}


def test_generate_annotation_tags(tmp_path):
    p = generate_annotation("", "img1.jpg", "img1.jpg", [{"sort": "lose", "xmin": 1, "ymin": 2, "xmax": 30, "ymax": 40},
                                                         {"sort": 4, "xmin": 5, "ymin": 6, "xmax": 7, "ymax": 8}], str(tmp_path))
    import xml.etree.ElementTree as ET
    root = ET.parse(p).getroot()
    assert os.path.basename(p) == "img1.xml" and root.tag == "annotation"
    assert [o.find("sort").text for o in root.findall("object")] == ["2", "4"]
    assert root.find("object/bndbox/xmax").text == "30" and root.find("size/depth").text == "3"


@pytest.mark.gpu
def test_main_matches_oracle_pipeline(tmp_path):
    from yolov8_vit_b200 import ViTConfig, ViTModel
    dev = "cuda:0"
    cfg = O.ViTConfig(depth=2)
    sd = O.make_vit_weights(cfg, seed=5)
    model = ViTModel(ViTConfig(depth=2), dev).load_state_dict(sd)
    head = synth.make_head(3, 80, 8400, n_obj=3, n_jit=12, seed=91)
    rgb = synth.make_images(3, 640, 640, seed=92)
    imgs = [(f"im{2 - i}.jpg", np.ascontiguousarray(rgb[i].numpy()[..., ::-1])) for i in range(3)]   # BGR, names reversed
    eng = HeadEngine(lambda blob: head.to(blob.device), (640, 640), "head")
    eng.set_desired(["num_dets", "bboxes", "scores", "labels"])
    seen = []
    out = inferdet.main(eng, imgs, torch.device(dev), [model], None, None,
                        lambda folder, name, path, objs: seen.append((name, len(objs))))
    want = []
    for i in range(3):
        ref = O.pipeline_image(head[i], rgb[i].numpy(), sd, cfg)
        for k in range(len(ref["keep"])):
            want.append((f"im{2 - i}.jpg", int(ref["cls"][k]), *[int(v) for v in ref["box"][k]]))
    want.sort(key=lambda t: t[0])
    assert [(t[0], t[1], *t[3:]) for t in out] == want
    assert all(0.35 <= t[2] <= 1.0 for t in out) and sorted(seen) == sorted((f"im{i}.jpg", sum(w[0] == f"im{i}.jpg" for w in want)) for i in range(3))
    # EfficientNMS-style engine output (<= 100 boxes / image) through det_postprocess-compatible tensors
    cand = O.decode_head(head[0], 0.25)
    n = min(100, len(cand[1]))
    order = torch.argsort(cand[1], descending=True, stable=True)[:n]
    bb = torch.zeros(1, 100, 4); sc = torch.zeros(1, 100); lab = torch.zeros(1, 100, dtype=torch.int32)
    bb[0, :n], sc[0, :n], lab[0, :n] = cand[0][order], cand[1][order], cand[2][order]
    eng2 = HeadEngine(lambda blob: (torch.tensor([[n]], dtype=torch.int32), bb, sc, lab), (640, 640), "nms")
    out2 = inferdet.main(eng2, imgs[:1], torch.device(dev), [model], None, None, None)
    keepm = ~(sc[0, :n] < 0.35)
    kb, ks = bb[0, :n][keepm], sc[0, :n][keepm]
    ref_keep = O.custom_nms(kb, ks, 0.45)
    assert [t[3:] for t in out2] == [tuple(int(v) for v in O.round_boxes(kb[k:k + 1])[0]) for k in ref_keep]
