"""Writes tests/golden/*.npz from the CPU oracle with fixed seeds.

PARITY UNPINNED by the reference (it ships no tests / vectors): these goldens pin
the ORACLE (oracle/hotpath.py, which restates the reference snippets verbatim and calls the real
torchvision.ops.box_iou / the cv2 rule) so that the CUDA path and later oracle edits are both
checked against the same committed numbers.  Run from the repo root:

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import hotpath as O          # noqa: E402
from yolov8_vit_b200 import synth        # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def nms_cases():
    """Edge cases of SURVEY.md §8c + seeded random cases.  Returns list of (name, boxes, scores, thr)."""
    f = np.float32
    cases = []
    cases.append(("empty", np.zeros((0, 4), f), np.zeros((0,), f), 0.45))
    cases.append(("one", np.array([[0, 0, 10, 10]], f), np.array([0.9], f), 0.45))
    cases.append(("two_disjoint", np.array([[0, 0, 10, 10], [20, 20, 30, 30]], f), np.array([0.5, 0.9], f), 0.45))
    cases.append(("two_overlap", np.array([[0, 0, 10, 10], [1, 1, 11, 11]], f), np.array([0.5, 0.9], f), 0.45))
    cases.append(("duplicates", np.array([[0, 0, 10, 10]] * 4 + [[50, 50, 60, 60]] * 3, f),
                  np.array([0.3, 0.9, 0.5, 0.7, 0.2, 0.8, 0.6], f), 0.45))
    # IoU exactly == thr (0.5): suppressed (strict <), unlike torchvision.ops.nms
    cases.append(("iou_eq_thr", np.array([[0, 0, 10, 10], [0, 0, 10, 5]], f), np.array([0.9, 0.8], f), 0.5))
    # zero-area pair: 0/0 = NaN -> suppressed
    cases.append(("zero_area_nan", np.array([[5, 5, 5, 5], [5, 5, 5, 5], [0, 0, 4, 4]], f),
                  np.array([0.9, 0.8, 0.7], f), 0.45))
    # tied scores: stable rule, lower index first
    rng = np.random.default_rng(7)
    c = rng.uniform(0, 100, (64, 2)).astype(f)
    wh = rng.uniform(10, 40, (64, 2)).astype(f)
    b = np.concatenate([c - wh / 2, c + wh / 2], 1).astype(f)
    cases.append(("tied_scores", b, np.repeat(np.array([0.9, 0.8, 0.7, 0.6], f), 16), 0.45))
    cases.append(("all_tied", b, np.full((64,), 0.5, f), 0.3))
    cases.append(("neg_zero_scores", b[:8], np.array([0.0, -0.0, 0.0, -0.0, 0.5, -0.5, 0.0, -0.0], f), 0.45))
    for i, (n, thr) in enumerate([(65, 0.45), (300, 0.45), (1000, 0.3), (1000, 0.7), (2500, 0.5)]):
        g = torch.Generator().manual_seed(synth.SEED0 + 100 + i)
        bx, sc = synth.make_nms_stress(1, n, seed=synth.SEED0 + 100 + i, field=256.0 if i % 2 else 640.0)
        cases.append((f"rand_{n}_{thr}", bx[0].numpy(), sc[0].numpy(), thr))
    return cases


def main():
    torch.manual_seed(0)
    # ---------------------------------------------------------------- a2 custom_nms
    d = {}
    names = []
    for name, b, s, thr in nms_cases():
        keep = O.custom_nms(torch.from_numpy(b), torch.from_numpy(s), thr)
        d[f"{name}__boxes"], d[f"{name}__scores"] = b, s
        d[f"{name}__thr"] = np.float64(thr)
        d[f"{name}__keep"] = np.asarray(keep, np.int32)
        names.append(name)
    d["names"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "nms.npz"), **d)

    # ---------------------------------------------------------------- a1 decode (+ conf equality)
    head = synth.make_head(3, nc=12, A=640, n_obj=4, n_jit=6, img=160, seed=synth.SEED0 + 11)
    head[0, 4 + 3, 17] = 0.35          # score == conf (as f32) -> kept (D4)
    head[0, 4:, 17][head[0, 4:, 17] > 0.35] = 0.35
    head[1, 4 + 2, 5] = head[1, 4 + 7, 5] = 0.5   # tie between classes -> first max wins
    d = {"head": head.numpy()}
    for b in range(3):
        for tag, kw in (("plain", {}), ("unl", dict(ratio=0.8, dwdh=(12.0, 4.0)))):
            bx, sc, lb, an = O.decode_head(head[b], 0.35, **kw)
            d[f"{tag}{b}__boxes"], d[f"{tag}{b}__scores"] = bx.numpy(), sc.numpy()
            d[f"{tag}{b}__labels"], d[f"{tag}{b}__anchors"] = lb.numpy(), an.numpy()
    np.savez_compressed(os.path.join(OUT, "decode.npz"), **d)

    # ---------------------------------------------------------------- a3 dilation
    rng = np.random.default_rng(3)
    W, H = 640, 480
    x0 = rng.integers(-5, W, 200); y0 = rng.integers(-5, H, 200)
    bx = np.stack([x0, y0, x0 + rng.integers(-3, 300, 200), y0 + rng.integers(-3, 300, 200)], 1).astype(np.int32)
    d = {"boxes": bx, "W": W, "H": H}
    for mode in ("full", "infer_half"):
        o = [O.dilate_box(*[int(v) for v in b], W, H, 1, 10, mode) for b in bx]
        d[f"{mode}__out"] = np.asarray([r[:4] for r in o], np.int32)
        d[f"{mode}__valid"] = np.asarray([r[4] for r in o], np.int32)
    np.savez_compressed(os.path.join(OUT, "dilate.npz"), **d)

    # ---------------------------------------------------------------- a4 crop/resize/normalise
    imgs = synth.make_images(2, 120, 160, seed=synth.SEED0 + 12, cell=8).numpy()
    boxes = np.array([[0, 0, 160, 120], [10, 20, 43, 97], [100, 5, 160, 60], [7, 7, 8, 8], [30, 40, 130, 41],
                      [0, 0, 23, 117]], np.int32)
    img_idx = np.array([0, 1, 0, 1, 0, 1], np.int32)
    d = {"images": imgs, "boxes": boxes, "img_idx": img_idx}
    for interp in ("nearest", "bilinear"):
        for out in (32, 224):
            n = len(boxes) if out == 32 else 2
            d[f"{interp}_{out}"] = np.stack([O.crop_resize_norm(imgs[img_idx[i]], boxes[i], out, interp)
                                             for i in range(n)])
    np.savez_compressed(os.path.join(OUT, "crop.npz"), **d)

    # ---------------------------------------------------------------- a5 ViT (tiny + full B/16)
    tiny = O.ViTConfig(img=32, patch=16, dim=256, depth=2, heads=4, mlp=512, head_out=64, hidden=16, num_classes=5)
    sd = O.make_vit_weights(tiny, seed=synth.SEED0 + 13)
    g = torch.Generator().manual_seed(synth.SEED0 + 14)
    x = (torch.rand(5, 3, 32, 32, generator=g) * 2 - 1).to(torch.bfloat16).to(torch.float32)
    logits, hidden = O.vit_forward(sd, x, tiny, return_hidden=True)
    d = {"tiny_x": x.numpy(), "tiny_logits": logits.numpy(), "tiny_hidden": hidden.numpy()}
    full = O.ViTConfig()
    sd = O.make_vit_weights(full, seed=synth.SEED0 + 3)
    x = (torch.rand(3, 3, 224, 224, generator=g) * 2 - 1).to(torch.bfloat16).to(torch.float32)
    logits, hidden = O.vit_forward(sd, x, full, return_hidden=True)
    d.update({"full_x_seed": synth.SEED0 + 14, "full_logits": logits.numpy(), "full_hidden": hidden.numpy()})
    np.savez_compressed(os.path.join(OUT, "vit.npz"), **d)

    # ---------------------------------------------------------------- whole path, cfg1 (batch 1)
    head = synth.make_head(1, nc=80, A=8400, n_obj=6, n_jit=24, seed=synth.SEED0 + 1)
    img = synth.make_images(1, 640, 640, seed=synth.SEED0 + 1).numpy()[0]
    r = O.pipeline_image(head[0], img, sd, full, interp="bilinear")
    np.savez_compressed(os.path.join(OUT, "pipeline_cfg1.npz"),
                        **{k: v for k, v in r.items()}, n_obj=6, n_jit=24, seed=synth.SEED0 + 1)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
