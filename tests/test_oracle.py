"""CPU tests: the oracle against the committed golden vectors, the C NMS oracle against the
verbatim torch snippet, and the ViT restatement against torchvision's vit_b_16 (an independent
implementation of the same published architecture).  PARITY UNPINNED by the reference itself
(no tests / vectors upstream): these goldens pin the oracle."""
import numpy as np
import pytest
import torch

from oracle import hotpath as O
from oracle import nms_c
from yolov8_vit_b200 import synth


def test_nms_golden(golden):
    g = golden("nms")
    for name in g["names"]:
        b, s, thr = g[f"{name}__boxes"], g[f"{name}__scores"], float(g[f"{name}__thr"])
        keep = O.custom_nms(torch.from_numpy(b), torch.from_numpy(s), thr)
        assert keep == g[f"{name}__keep"].tolist(), name
        assert nms_c.nms(b, s, thr) == keep, f"C oracle differs on {name}"


def test_nms_known_answers():
    f = torch.float32
    # strict '<': IoU == thr is suppressed (torchvision.ops.nms would keep both)
    assert O.custom_nms(torch.tensor([[0, 0, 10, 10], [0, 0, 10, 5]], dtype=f), torch.tensor([.9, .8]), 0.5) == [0]
    # NaN IoU (two zero-area boxes) is suppressed
    assert O.custom_nms(torch.tensor([[5, 5, 5, 5], [5, 5, 5, 5]], dtype=f), torch.tensor([.9, .8]), 0.45) == [0]
    # the last survivor is always kept; order is descending score
    assert O.custom_nms(torch.tensor([[0, 0, 1, 1], [5, 5, 6, 6], [9, 9, 10, 10]], dtype=f),
                        torch.tensor([.1, .3, .2]), 0.45) == [1, 2, 0]


def test_nms_matches_unstable_snippet_when_scores_distinct():
    """README.md:64 uses the unstable argsort; with pairwise-distinct scores it equals stable=True."""
    from torchvision.ops import box_iou
    b, s = synth.make_nms_stress(1, 500, seed=5)
    b, s = b[0], s[0]
    si = torch.argsort(s, descending=True)
    keep = []
    while si.size(0) > 0:
        i = si[0].item()
        keep.append(i)
        if si.size(0) == 1:
            break
        m = box_iou(b[i:i + 1], b[si[1:]]).squeeze(0) < 0.45
        si = si[1:][m]
    assert keep == O.custom_nms(b, s, 0.45)


@pytest.mark.parametrize("n,thr", [(3000, 0.3), (3000, 0.6)])
def test_c_oracle_vs_torch_snippet(n, thr):
    b, s = synth.make_nms_stress(1, n, seed=77 + n)
    assert nms_c.nms(b[0].numpy(), s[0].numpy(), thr) == O.custom_nms(b[0], s[0], thr)


def test_decode_golden(golden):
    g = golden("decode")
    head = torch.from_numpy(g["head"])
    for b in range(3):
        for tag, kw in (("plain", {}), ("unl", dict(ratio=0.8, dwdh=(12.0, 4.0)))):
            bx, sc, lb, an = O.decode_head(head[b], 0.35, **kw)
            assert np.array_equal(bx.numpy(), g[f"{tag}{b}__boxes"])
            assert np.array_equal(sc.numpy(), g[f"{tag}{b}__scores"])
            assert np.array_equal(lb.numpy(), g[f"{tag}{b}__labels"])
            assert np.array_equal(an.numpy(), g[f"{tag}{b}__anchors"])
    # D4: score == conf is kept; first max wins
    an0 = g["plain0__anchors"].tolist()
    assert 17 in an0 and g["plain0__labels"][an0.index(17)] == 3
    an1 = g["plain1__anchors"].tolist()
    assert g["plain1__labels"][an1.index(5)] == 2


def test_dilate_golden_and_readme_formula(golden):
    g = golden("dilate")
    W, H = int(g["W"]), int(g["H"])
    for mode in ("full", "infer_half"):
        for b, o, v in zip(g["boxes"], g[f"{mode}__out"], g[f"{mode}__valid"]):
            r = O.dilate_box(*[int(t) for t in b], W, H, 1, 10, mode)
            assert list(r[:4]) == o.tolist() and int(r[4]) == int(v)
    # README.md:91-121 with alpha = 0.1 on a box whose sides are multiples of 10
    assert O.dilate_box(100, 200, 200, 300, 640, 480)[:4] == (90, 190, 210, 310)
    assert O.dilate_box(0, 0, 640, 480, 640, 480)[:4] == (0, 0, 640, 480)


def test_crop_golden_and_cv2(golden):
    import cv2
    g = golden("crop")
    imgs, boxes, idx = g["images"], g["boxes"], g["img_idx"]
    for interp in ("nearest", "bilinear"):
        for out in (32, 224):
            ref = g[f"{interp}_{out}"]
            for i in range(ref.shape[0]):
                got = O.crop_resize_norm(imgs[idx[i]], boxes[i], out, interp)
                assert np.array_equal(got, ref[i]), (interp, out, i)
    # the nearest rule is cv2's (the library behind A.Resize(..., INTER_NEAREST), app.py:40)
    for i in range(len(boxes)):
        x0, y0, x1, y1 = boxes[i]
        crop = imgs[idx[i]][y0:y1, x0:x1]
        ref = cv2.resize(crop, (224, 224), interpolation=cv2.INTER_NEAREST).astype(np.float32)
        ref = ((ref - np.float32(127.5)) * np.float32(1.0 / 127.5)).transpose(2, 0, 1)
        assert np.array_equal(O.crop_resize_norm(imgs[idx[i]], boxes[i], 224, "nearest"), ref)


def test_vit_golden_tiny(golden):
    g = golden("vit")
    tiny = O.ViTConfig(img=32, patch=16, dim=256, depth=2, heads=4, mlp=512, head_out=64, hidden=16, num_classes=5)
    sd = O.make_vit_weights(tiny, seed=synth.SEED0 + 13)
    logits, hidden = O.vit_forward(sd, torch.from_numpy(g["tiny_x"]), tiny, return_hidden=True)
    assert np.allclose(logits.numpy(), g["tiny_logits"], rtol=1e-4, atol=1e-5)
    assert np.allclose(hidden.numpy(), g["tiny_hidden"], rtol=1e-4, atol=1e-5)


def test_vit_restatement_vs_torchvision():
    """oracle.vit_forward (restated timm ViT-B/16) == torchvision.models.vit_b_16 with the same weights."""
    from torchvision.models import vit_b_16
    torch.manual_seed(1)
    tv = vit_b_16(weights=None).eval()
    cfg = O.ViTConfig()
    sd = O.make_vit_weights(cfg, seed=3)
    with torch.no_grad():
        tv.conv_proj.weight.copy_(sd["model.patch_embed.proj.weight"])
        tv.conv_proj.bias.copy_(sd["model.patch_embed.proj.bias"])
        tv.class_token.copy_(sd["model.cls_token"])
        tv.encoder.pos_embedding.copy_(sd["model.pos_embed"])
        for i, blk in enumerate(tv.encoder.layers):
            p = f"model.blocks.{i}."
            blk.ln_1.weight.copy_(sd[p + "norm1.weight"]); blk.ln_1.bias.copy_(sd[p + "norm1.bias"])
            blk.self_attention.in_proj_weight.copy_(sd[p + "attn.qkv.weight"])
            blk.self_attention.in_proj_bias.copy_(sd[p + "attn.qkv.bias"])
            blk.self_attention.out_proj.weight.copy_(sd[p + "attn.proj.weight"])
            blk.self_attention.out_proj.bias.copy_(sd[p + "attn.proj.bias"])
            blk.ln_2.weight.copy_(sd[p + "norm2.weight"]); blk.ln_2.bias.copy_(sd[p + "norm2.bias"])
            blk.mlp[0].weight.copy_(sd[p + "mlp.fc1.weight"]); blk.mlp[0].bias.copy_(sd[p + "mlp.fc1.bias"])
            blk.mlp[3].weight.copy_(sd[p + "mlp.fc2.weight"]); blk.mlp[3].bias.copy_(sd[p + "mlp.fc2.bias"])
        tv.encoder.ln.weight.copy_(sd["model.norm.weight"]); tv.encoder.ln.bias.copy_(sd["model.norm.bias"])
        tv.heads.head.weight.copy_(sd["model.head.weight"]); tv.heads.head.bias.copy_(sd["model.head.bias"])
        x = torch.rand(2, 3, 224, 224) * 2 - 1
        ref = tv(x)
        _, hidden = O.vit_forward(sd, x, cfg, return_hidden=True)
    assert torch.allclose(hidden, ref, rtol=1e-3, atol=1e-3), float((hidden - ref).abs().max())


def test_pipeline_cfg1_golden(golden):
    g = golden("pipeline_cfg1")
    head = synth.make_head(1, nc=80, A=8400, n_obj=int(g["n_obj"]), n_jit=int(g["n_jit"]), seed=int(g["seed"]))
    img = synth.make_images(1, 640, 640, seed=int(g["seed"])).numpy()[0]
    r = O.pipeline_image(head[0], img, None)
    for k in ("keep", "anchor", "label", "box", "crop_box", "valid"):
        assert np.array_equal(r[k], g[k]), k
    assert len(r["keep"]) >= 6


def test_raw_head_decode_oracle_properties():
    """f1 oracle: uniform DFL logits -> every distance is the mean bin (7.5 cells); a one-hot bin gives that
    bin; boxes scale with the level stride; anchors follow make_anchors (row-major, +0.5)."""
    levels = [(4, 6, 8.0), (2, 3, 16.0), (1, 2, 32.0)]
    A = 24 + 6 + 2
    raw = torch.zeros(64 + 3, A)
    raw[64:] = -5.0
    dec = O.decode_raw_head(raw, levels)
    pts, st = O.make_anchors(levels)
    assert pts.shape == (A, 2) and pts[0].tolist() == [0.5, 0.5] and pts[7].tolist() == [1.5, 1.5]
    assert torch.allclose(dec[0], (pts[:, 0] - 7.5) * st[:, 0]) and torch.allclose(dec[3], (pts[:, 1] + 7.5) * st[:, 0])
    raw[:64] = -30.0
    for k, bin_ in enumerate((2, 0, 5, 15)):
        raw[k * 16 + bin_] = 30.0
    dec = O.decode_raw_head(raw, levels)
    a = 24 + 4                                    # level 1, cell (1, 1)
    assert torch.allclose(dec[:4, a], torch.tensor([(1.5 - 2) * 16, (1.5 - 0) * 16, (1.5 + 5) * 16, (1.5 + 15) * 16]))
    assert torch.allclose(dec[4:], torch.sigmoid(torch.tensor(-5.0)).expand(3, A))
