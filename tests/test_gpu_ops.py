"""GPU parity tests (through the C-ABI) for decode / NMS / dilation / crop against the CPU oracle:
bit-exact for every integer, index and pre-bf16 fp32 result."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import hotpath as O          # checker only
from oracle import nms_c
from yolov8_vit_b200 import ops, synth
from yolov8_vit_b200.utils import utils as U

DEV = "cuda:0"


def _decode_check(head, conf=0.35, **kw):
    bx, sc, lb, an, cnt = ops.decode_filter(head.to(DEV), conf, **kw)
    cnt = cnt.cpu()
    for b in range(head.shape[0]):
        rb, rs, rl, ra = O.decode_head(head[b], conf, **kw)
        k = int(cnt[b])
        assert k == rb.shape[0], (b, k, rb.shape[0])
        assert torch.equal(bx[b, :k].cpu(), rb) and torch.equal(sc[b, :k].cpu(), rs)
        assert torch.equal(lb[b, :k].cpu(), rl) and torch.equal(an[b, :k].cpu(), ra)


def test_decode_golden(golden):
    g = golden("decode")
    head = torch.from_numpy(g["head"])
    bx, sc, lb, an, cnt = ops.decode_filter(head.to(DEV), 0.35)
    bu, *_ = ops.decode_filter(head.to(DEV), 0.35, ratio=0.8, dwdh=(12.0, 4.0))
    for b in range(3):
        k = int(cnt[b])
        assert k == len(g[f"plain{b}__anchors"])
        assert np.array_equal(bx[b, :k].cpu().numpy(), g[f"plain{b}__boxes"])
        assert np.array_equal(sc[b, :k].cpu().numpy(), g[f"plain{b}__scores"])
        assert np.array_equal(lb[b, :k].cpu().numpy(), g[f"plain{b}__labels"])
        assert np.array_equal(an[b, :k].cpu().numpy(), g[f"plain{b}__anchors"])
        assert np.array_equal(bu[b, :k].cpu().numpy(), g[f"unl{b}__boxes"])


def test_decode_cfg2_and_ragged():
    _decode_check(synth.make_head(8, 80, 8400, seed=synth.SEED0 + 2))
    _decode_check(synth.make_head(2, 5, 8400, n_obj=3, seed=9))            # nc=5 (the real detector)
    _decode_check(synth.make_head(2, 7, 1001, n_obj=3, n_jit=5, seed=10))  # A % 4 != 0: scalar path
    _decode_check(synth.make_head(1, 80, 33600, n_obj=4, img=1280, seed=11))
    h = synth.make_head(2, 80, 8400, seed=12)
    _decode_check(h, conf=2.0)     # nothing passes
    _decode_check(h, conf=0.0)     # everything passes (count == A)


def test_decode_workspace_reuse_is_idempotent():
    h = synth.make_head(4, 80, 8400, seed=13).to(DEV)
    a = ops.decode_filter(h)
    b = ops.decode_filter(h)
    assert torch.equal(a[4], b[4])
    for i in range(4):           # entries beyond the count are unspecified
        k = int(a[4][i])
        assert all(torch.equal(x[i, :k], y[i, :k]) for x, y in zip(a[:4], b[:4]))


def test_nms_golden(golden):
    g = golden("nms")
    for name in g["names"]:
        b, s, thr = g[f"{name}__boxes"], g[f"{name}__scores"], float(g[f"{name}__thr"])
        keep = U.custom_nms(torch.from_numpy(b), torch.from_numpy(s), thr)
        assert keep == g[f"{name}__keep"].tolist(), name


def test_nms_cfg2_batched_bit_exact():
    head = synth.make_head(64, 80, 8400, seed=synth.SEED0 + 2)
    bx, sc, lb, an, cnt = ops.decode_filter(head.to(DEV))
    keep, kc = ops.nms_batched(bx, sc, cnt, 0.45)
    keep, kc, cnt_c, bxc, scc = keep.cpu(), kc.cpu(), cnt.cpu(), bx.cpu(), sc.cpu()
    for b in range(64):
        n = int(cnt_c[b])
        ref = nms_c.nms(bxc[b, :n].numpy(), scc[b, :n].numpy(), 0.45)
        assert keep[b, :int(kc[b])].tolist() == ref, b
    for b in (0, 17):      # and the verbatim torch snippet on two images
        n = int(cnt_c[b])
        assert keep[b, :int(kc[b])].tolist() == O.custom_nms(bxc[b, :n], scc[b, :n], 0.45)
    # max_keep truncates in score order
    k10, c10 = ops.nms_batched(bx, sc, cnt, 0.45, max_keep=10)
    for b in range(64):
        assert k10[b, :int(c10[b])].tolist() == keep[b, :min(10, int(kc[b]))].tolist()


def test_nms_properties_full_size():
    """Size-independent properties at cfg5's full batch: idempotence (NMS of the keeps keeps all),
    keeps are pairwise below the threshold, scores descend."""
    from torchvision.ops import box_iou
    boxes, scores = synth.make_nms_stress(32, 30000, seed=synth.SEED0 + 5)
    bd, sd = boxes.to(DEV), scores.to(DEV)
    keep, kc = ops.nms_batched(bd, sd, None, 0.5)
    for b in (0, 31):
        k = keep[b, :int(kc[b])].long()
        kb, ks = bd[b, k], sd[b, k]
        assert bool((ks[:-1] >= ks[1:]).all())
        iou = box_iou(kb, kb).fill_diagonal_(0)
        assert float(iou.max()) < 0.5
        k2, c2 = ops.nms_batched(kb[None].contiguous(), ks[None].contiguous(), None, 0.5)
        assert int(c2[0]) == k.numel() and k2[0, :int(c2[0])].tolist() == list(range(k.numel()))


def test_dilate_golden(golden):
    g = golden("dilate")
    b = torch.from_numpy(g["boxes"]).to(DEV)
    for mode in ("full", "infer_half"):
        out, valid = ops.dilate_boxes(b, int(g["W"]), int(g["H"]), (1, 10), mode)
        assert np.array_equal(out.cpu().numpy(), g[f"{mode}__out"])
        assert np.array_equal(valid.cpu().numpy(), g[f"{mode}__valid"])
    assert U.box_dilate(100, 200, 200, 300, 640, 480) == (90, 190, 210, 310, True)


def test_crop_golden_both_interps_and_layouts(golden):
    g = golden("crop")
    imgs = torch.from_numpy(g["images"]).to(DEV)
    boxes = torch.from_numpy(g["boxes"]).to(DEV)
    idx = torch.from_numpy(g["img_idx"]).to(DEV)
    for interp in ("nearest", "bilinear"):
        for out, patch in ((32, 16), (32, 8), (224, 16)):
            ref = torch.from_numpy(g[f"{interp}_{out}"])
            n = ref.shape[0]
            chw = ops.crop_resize(imgs, boxes, idx, n, interp, out, patch, "chw").cpu()
            assert torch.equal(chw, ref), (interp, out)
            pm = ops.crop_resize(imgs, boxes, idx, n, interp, out, patch, "patch").cpu()
            assert torch.equal(pm, O.patchify(ref, patch).to(torch.bfloat16)), (interp, out, patch)


def test_crop_cfg3_random_boxes_bit_exact():
    imgs = synth.make_images(4, 640, 640, seed=synth.SEED0 + 3)
    boxes = synth.make_boxes_int(32, seed=synth.SEED0 + 3)
    dil, valid = ops.dilate_boxes(boxes.to(DEV), 640, 640)
    assert bool(valid.all())
    idx = (torch.arange(32) % 4).to(torch.int32)
    for interp in ("bilinear", "nearest"):
        got = ops.crop_resize(imgs.to(DEV), dil, idx.to(DEV), 32, interp, 224, 16, "chw").cpu()
        for i in range(32):
            ref = O.crop_resize_norm(imgs[idx[i]].numpy(), dil[i].cpu().numpy(), 224, interp)
            assert np.array_equal(got[i].numpy(), ref), (interp, i)


def test_crop_patch_order_mode_is_bit_identical():
    """YV_CROP_MAP=patch keeps the round-1 thread order of the crop kernel (by patch instead of by output row) for A/B
    timing; only the order of the work changes, so both layouts and both interpolations must equal the oracle there too.
    Child process: the switch is read once."""
    import os, subprocess, sys
    code = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r)
from oracle import hotpath as O
from yolov8_vit_b200 import ops, synth
imgs = synth.make_images(4, 640, 640, seed=synth.SEED0 + 3)
boxes = synth.make_boxes_int(24, seed=synth.SEED0 + 7)
dil, _ = ops.dilate_boxes(boxes.cuda(), 640, 640)
idx = (torch.arange(24) %% 4).to(torch.int32)
for interp in ("bilinear", "nearest"):
    chw = ops.crop_resize(imgs.cuda(), dil, idx.cuda(), 24, interp, 224, 16, "chw").cpu()
    pm = ops.crop_resize(imgs.cuda(), dil, idx.cuda(), 24, interp, 224, 16, "patch").cpu()
    for i in range(24):
        ref = O.crop_resize_norm(imgs[idx[i]].numpy(), dil[i].cpu().numpy(), 224, interp)
        assert np.array_equal(chw[i].numpy(), ref), (interp, i)
        assert torch.equal(pm[i * 196:(i + 1) * 196], O.patchify(torch.from_numpy(ref)[None], 16).to(torch.bfloat16)), (interp, i)
print('patch order ok')
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, YV_CROP_MAP="patch"), capture_output=True,
                       text=True, timeout=280)
    assert r.returncode == 0 and "patch order ok" in r.stdout, r.stdout + r.stderr


def test_crop_image_reference_surface():
    img = synth.make_images(1, 200, 300, seed=5)[0].numpy()
    t = U.crop_image(img, 50, 40, 150, 120)
    x0, y0, x1, y1, ok = O.dilate_box(50, 40, 150, 120, 300, 200, 1, 10, "infer_half")
    assert np.array_equal(t.cpu().numpy(), O.crop_resize_norm(img, (x0, y0, x1, y1), 224, "nearest"))


def test_decode_raw_head_against_oracle():
    """Row f1: DFL + anchor decode fused into the filter kernel vs the torch fp32 restatement.
    Integer outputs (labels, anchors, counts) exact; scores / boxes to 1e-5 relative (expf)."""
    levels = synth.LEVELS_640
    raw = synth.make_raw_head(5, 80, levels, n_obj=9, seed=77)
    boxes, scores, labels, anchors, counts = [t.cpu() for t in ops.decode_filter_raw(raw.to(DEV), levels, 0.35)]
    for b in range(5):
        rb, rs, rl, ra = O.decode_raw_filter(raw[b], levels, 0.35)
        n = int(counts[b])
        assert n == len(ra) and n >= 5
        assert anchors[b, :n].tolist() == ra.tolist() and labels[b, :n].tolist() == rl.tolist()
        assert torch.allclose(scores[b, :n], rs, rtol=1e-5, atol=1e-6)
        assert torch.allclose(boxes[b, :n], rb, rtol=1e-5, atol=2e-3)


def test_decode_raw_feeds_the_same_pipeline():
    """raw head -> f1 decode -> custom_nms gives the keep lists of the oracle chain on the same input."""
    levels = [(8, 8, 8.0), (4, 4, 16.0), (2, 2, 32.0)]
    raw = synth.make_raw_head(3, 5, levels, n_obj=20, seed=78)
    cand = ops.decode_filter_raw(raw.to(DEV), levels, 0.35)
    keep, kc = ops.nms_batched(cand[0], cand[1], cand[4], 0.45)
    for b in range(3):
        rb, rs, _, _ = O.decode_raw_filter(raw[b], levels, 0.35)
        assert keep[b, :int(kc[b])].cpu().tolist() == O.custom_nms(rb, rs, 0.45)


@pytest.mark.parametrize("H,W", [(640, 640), (480, 640), (1280, 960), (333, 517), (100, 100), (1080, 1920), (97, 211)])
def test_letterbox_blob_bit_exact_vs_cv2(H, W):
    """Row f2: GPU letterbox + BGR->RGB + /255 + CHW equals the cv2-based oracle bit for bit (8-bit INTER_LINEAR
    fixed-point resize, 114 padding), and returns the same ratio / (dw, dh)."""
    g = torch.Generator().manual_seed(H * 7 + W)
    imgs = torch.randint(0, 256, (3, H, W, 3), generator=g, dtype=torch.uint8)
    blob, ratio, dwdh = ops.letterbox_blob(imgs.to(DEV), (640, 640), bgr=True)
    blob = blob.cpu().numpy()
    for b in range(3):
        ref, r, (dw, dh) = O.letterbox_blob(imgs[b].numpy(), (640, 640))
        assert r == ratio and (dw, dh) == dwdh
        assert np.array_equal(blob[b], ref[0]), (H, W, b, int((blob[b] != ref[0]).sum()))
    rgb_blob, _, _ = ops.letterbox_blob(imgs.to(DEV), (640, 640), bgr=False)
    assert np.array_equal(rgb_blob.cpu().numpy()[:, [2, 1, 0]], blob)


@pytest.mark.parametrize("thr", [0.65, 0.5])
def test_efficient_nms_equals_torchvision_batched_nms(thr):
    """Row f4: class-aware standard NMS (iou > thr suppresses, top-100) vs torchvision.ops.batched_nms."""
    import torchvision
    g = torch.Generator().manual_seed(11)
    B, N = 6, 900
    c = torch.rand(B, N, 2, generator=g) * 300
    wh = torch.rand(B, N, 2, generator=g) * 80 + 10
    boxes = torch.cat([c - wh / 2, c + wh / 2], 2)
    scores = torch.stack([synth.distinct_scores(N, N, g, 0.0, 1.0) for _ in range(B)])
    labels = torch.randint(0, 5, (B, N), generator=g, dtype=torch.int32)
    keep, kc = ops.efficient_nms(boxes.to(DEV), scores.to(DEV), labels.to(DEV), None, thr, 100)
    for b in range(B):
        ref = torchvision.ops.batched_nms(boxes[b], scores[b], labels[b].long(), thr)[:100]
        assert keep[b, :int(kc[b])].cpu().tolist() == ref.tolist(), b
    # class-agnostic standard rule = torchvision.ops.nms, including an IoU == thr pair (kept under this rule)
    bx = torch.tensor([[0., 0, 2, 2], [0, 0, 2, 1], [5, 5, 6, 6]]); sc = torch.tensor([.9, .8, .7])
    k, n = ops.nms_batched(bx[None].to(DEV), sc[None].to(DEV), None, 0.5, 0, rule="standard")
    assert k[0, :int(n[0])].cpu().tolist() == torchvision.ops.nms(bx, sc, 0.5).tolist() == [0, 1, 2]
    k, n = ops.nms_batched(bx[None].to(DEV), sc[None].to(DEV), None, 0.5, 0)          # custom_nms suppresses equality
    assert k[0, :int(n[0])].cpu().tolist() == [0, 2]


@pytest.mark.parametrize("B,S,max_det,cap_frac", [(3, 64, 100, 1.0), (37, 200, 100, 1.0), (700, 24, 10, 1.0),
                                                   (5, 400, 300, 1.0), (16, 128, 100, 0.4)])
def test_build_crops_vs_oracle(B, S, max_det, cap_frac):
    """Crop list (round -> dilate -> clamp -> valid-crop compaction in (image, rank) order) against the
    oracle's round_boxes + dilate_box: degenerate boxes (invalid crops), more keeps than max_det, more than
    one 512-image scan block, max_det above one 128-slot sweep, and an overflowing crop_cap."""
    g = torch.Generator().manual_seed(B * 1000 + S)
    W, H = 640, 480
    xy = torch.rand(B, S, 2, generator=g) * torch.tensor([W, H]) * 1.1 - 20.0
    wh = torch.rand(B, S, 2, generator=g) * 120.0
    wh[torch.rand(B, S, generator=g) < 0.15] = 0.2                    # rounds to an empty box -> invalid crop
    boxes = torch.cat([xy, xy + wh], -1).contiguous()
    scores = torch.rand(B, S, generator=g)
    labels = torch.randint(0, 80, (B, S), generator=g, dtype=torch.int32)
    anchors = torch.randint(0, 8400, (B, S), generator=g, dtype=torch.int32)
    keep_counts = torch.randint(0, min(S, max_det + 40) + 1, (B,), generator=g, dtype=torch.int32)
    keep = torch.stack([torch.randperm(S, generator=g) for _ in range(B)]).to(torch.int32)
    ref_total = 0
    ref = []
    for b in range(B):
        nk = min(int(keep_counts[b]), max_det)
        for k in range(nk):
            i = int(keep[b, k])
            q = O.round_boxes(boxes[b, i]).tolist()
            x0, y0, x1, y1, ok = O.dilate_box(*q, W, H)
            ref.append((b, k, q, [x0, y0, x1, y1], ok, i))
            ref_total += int(ok)
    cap = max(1, int(ref_total * cap_frac))
    cl = ops.build_crops(boxes.to(DEV), scores.to(DEV), labels.to(DEV), anchors.to(DEV), keep.to(DEV),
                         keep_counts.to(DEV), max_det, W, H, crop_cap=cap)
    assert int(cl.n_crops) == ref_total
    assert torch.equal(cl.det_count.cpu(), keep_counts.clamp(max=max_det))
    det_box, det_cb, det_crop = cl.det_box.cpu(), cl.det_crop_box.cpu(), cl.det_crop.cpu()
    det_score, det_label, det_anchor = cl.det_score.cpu(), cl.det_label.cpu(), cl.det_anchor.cpu()
    crop_box, crop_img, crop_det = cl.crop_box.cpu(), cl.crop_img.cpu(), cl.crop_det.cpu()
    rank = 0
    filled = torch.zeros(B, max_det, dtype=torch.bool)
    for b, k, q, c, ok, i in ref:
        filled[b, k] = True
        assert det_box[b, k].tolist() == q and det_cb[b, k].tolist() == c
        assert det_score[b, k] == scores[b, i] and det_label[b, k] == labels[b, i] and det_anchor[b, k] == anchors[b, i]
        if ok and rank < cap:
            assert int(det_crop[b, k]) == rank
            assert crop_box[rank].tolist() == c and int(crop_img[rank]) == b and int(crop_det[rank]) == b * max_det + k
        else:
            assert int(det_crop[b, k]) == -1
        rank += int(ok)
    assert bool((det_crop[~filled] == -1).all()) and bool((det_label[~filled] == -1).all())
    assert bool((det_box[~filled] == 0).all()) and bool((det_score[~filled] == 0).all())


def test_crop_image_edges_and_unaligned_buffers():
    """The bilinear gather reads two RGB pixels as three aligned 32-bit words: crops that end on the last
    pixel of the last image (buffer tail -> byte path), odd image sizes (row starts at every alignment, total
    size not a multiple of 4), 1-pixel-wide / 1-pixel-high crops (both taps the same pixel), and an image
    tensor whose base is not 4-byte aligned (byte-load kernel) all stay bit-exact against the oracle."""
    H, W = 37, 53
    imgs = synth.make_images(3, H, W, seed=91)
    boxes = torch.tensor([[0, 0, W, H], [W - 1, H - 1, W, H], [W - 2, H - 3, W, H], [0, H - 1, W, H],
                          [W - 1, 0, W, H], [5, 7, 6, 30], [11, 13, 40, 14], [1, 1, 2, 2], [W - 7, H - 5, W - 1, H - 1]],
                         dtype=torch.int32)
    idx = torch.tensor([2, 2, 2, 2, 2, 0, 1, 1, 2], dtype=torch.int32)
    n = boxes.shape[0]
    raw = torch.zeros(imgs.numel() + 8, dtype=torch.uint8, device=DEV)
    for off in (0, 1, 2, 3):                                   # off = 0: word loads; 1..3: unaligned base
        view = raw[off:off + imgs.numel()].view(imgs.shape)
        view.copy_(imgs)
        for interp in ("bilinear", "nearest"):
            got = ops.crop_resize(view, boxes.to(DEV), idx.to(DEV), n, interp, 32, 16, "chw").cpu()
            for i in range(n):
                ref = O.crop_resize_norm(imgs[idx[i]].numpy(), boxes[i].numpy(), 32, interp)
                assert np.array_equal(got[i].numpy(), ref), (off, interp, i)
