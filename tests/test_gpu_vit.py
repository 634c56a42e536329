"""GPU parity tests for LayerNorm, attention, the ViT forward and the whole path.
ViT tolerance (north star): max|Δlogits| / max|ref logits| <= 1e-2 (bf16 operands, fp32 accumulate)
with identical top-1."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import hotpath as O          # checker only
from yolov8_vit_b200 import HotPath, HotPathConfig, ViTConfig, ViTModel, ops, synth
from yolov8_vit_b200.utils import utils as U

DEV = "cuda:0"
REL_TOL = 1e-2


def rel_err(got, ref):
    return float((got - ref).abs().max() / ref.abs().max())


def test_layernorm():
    x = (torch.randn(1000, 768) * 2 + 0.5).to(torch.bfloat16)
    g, b = torch.rand(768) + 0.5, torch.randn(768) * 0.1
    ref = torch.nn.functional.layer_norm(x.float(), (768,), g, b, 1e-6)
    got = ops.layernorm_bf16(x.to(DEV), g.to(DEV), b.to(DEV), 1e-6).cpu().float()
    assert float((got - ref).abs().max()) <= 2 ** -7 * float(ref.abs().max())


def _attention_ref(qkv, heads, dh):
    n, T, _ = qkv.shape
    q, k, v = qkv.float().reshape(n, T, 3, heads, dh).permute(2, 0, 3, 1, 4)
    return (torch.softmax(q @ k.transpose(-1, -2) * dh ** -0.5, -1) @ v).transpose(1, 2).reshape(n, T, heads * dh)


# T covers every shape of the key-block / query-tile decomposition: one block (<= 128 keys), two blocks (197 = 112 +
# 96, 256 = 128 + 128, 129 = 80 + 64), many blocks with a short last one (785 = 6 x 128 + 32, 300, 400), one / two /
# seven query tiles, an odd tile count (the second slot of the last pair does not exist), T = 1
@pytest.mark.parametrize("n,T", [(3, 197), (2, 64), (2, 5), (1, 256), (2, 785), (1, 300), (2, 1), (3, 16), (2, 17),
                                 (2, 128), (2, 129), (1, 400), (13, 197)])
def test_attention(n, T):
    heads, dh = 12, 64
    qkv = (torch.randn(n, T, 3 * heads * dh, generator=torch.Generator().manual_seed(T)) * 0.7).to(torch.bfloat16)
    ref = _attention_ref(qkv, heads, dh)
    got = ops.attention_bf16(qkv.to(DEV), heads).cpu().float()
    assert bool(torch.isfinite(got).all())
    assert float((got - ref).abs().max()) <= 2e-2 * float(ref.abs().max())


@pytest.mark.parametrize("T", [197, 785])
def test_attention_online_softmax_rescale(T):
    """Scores whose running maximum jumps by far more than the 2^8 rescale threshold from one key block to the next
    (a few keys with a 30x larger norm placed in LATER blocks), peaked rows, and large-magnitude logits: the O
    accumulator in tensor memory is rescaled in place and the result still matches the fp32 softmax."""
    heads, dh = 12, 64
    g = torch.Generator().manual_seed(7 + T)
    qkv = (torch.randn(2, T, 3 * heads * dh, generator=g) * 0.5)
    D = heads * dh
    hot = [T - 1, T - 40, 130] if T < 300 else [T - 1, 300, 520, 700, 130]
    for t in hot:
        qkv[:, t, D:2 * D] *= 30.0                  # keys with a huge norm: the block they sit in raises the row maximum
    qkv[1, :, :D] *= 6.0                            # large logits on every row of the second sample
    qkv = qkv.to(torch.bfloat16)
    ref = _attention_ref(qkv, heads, dh)
    got = ops.attention_bf16(qkv.to(DEV), heads).cpu().float()
    assert bool(torch.isfinite(got).all())
    assert float((got - ref).abs().max()) <= 2e-2 * float(ref.abs().max())


def test_attention_many_items_persistent_phases():
    """More work items than SMs x 4, so every CTA wraps its Q / K-V rings, the three rotating score buffers and
    both O accumulators several times; deterministic across launches."""
    heads, dh, n, T = 12, 64, 60, 197
    qkv = (torch.randn(n, T, 3 * heads * dh, generator=torch.Generator().manual_seed(3)) * 0.7).to(torch.bfloat16)
    ref = _attention_ref(qkv, heads, dh)
    a = ops.attention_bf16(qkv.to(DEV), heads)
    b = ops.attention_bf16(qkv.to(DEV), heads)
    assert torch.equal(a, b)
    assert float((a.cpu().float() - ref).abs().max()) <= 2e-2 * float(ref.abs().max())


def test_vit_tiny_golden(golden):
    g = golden("vit")
    tiny = O.ViTConfig(img=32, patch=16, dim=256, depth=2, heads=4, mlp=512, head_out=64, hidden=16, num_classes=5)
    sd = O.make_vit_weights(tiny, seed=synth.SEED0 + 13)
    m = ViTModel(ViTConfig(**{k: getattr(tiny, k) for k in tiny.__dataclass_fields__}), DEV).load_state_dict(sd)
    x = torch.from_numpy(g["tiny_x"])
    logits, top1, hidden = m.forward_patches(O.patchify(x, 16).to(torch.bfloat16).to(DEV), 5, return_hidden=True)
    assert rel_err(hidden.cpu(), torch.from_numpy(g["tiny_hidden"])) <= REL_TOL
    assert rel_err(logits.cpu(), torch.from_numpy(g["tiny_logits"])) <= REL_TOL
    assert top1.cpu().tolist() == g["tiny_logits"].argmax(1).tolist()


@pytest.fixture(scope="module")
def full_model():
    sd = O.make_vit_weights(O.ViTConfig(), seed=synth.SEED0 + 3)
    return ViTModel(ViTConfig(), DEV).load_state_dict(sd), sd


def test_vit_b16_golden(golden, full_model):
    g = golden("vit")
    m, _ = full_model
    gen = torch.Generator().manual_seed(int(g["full_x_seed"]))
    _ = torch.rand(5, 3, 32, 32, generator=gen)              # the tiny case consumed the generator first
    x = (torch.rand(3, 3, 224, 224, generator=gen) * 2 - 1).to(torch.bfloat16).to(torch.float32)
    logits = U.Network_Wrapper(m, 5)(x.to(DEV)).cpu()
    ref = torch.from_numpy(g["full_logits"])
    assert rel_err(logits, ref) <= REL_TOL, rel_err(logits, ref)
    assert logits.argmax(1).tolist() == ref.argmax(1).tolist()


def test_vit_b16_chunking_and_odd_counts(full_model):
    """300 crops (> one 256-crop chunk, M not a multiple of 128): chunked == unchunked rows."""
    m, sd = full_model
    imgs = synth.make_images(4, 640, 640, seed=3).to(DEV)
    boxes = synth.make_boxes_int(300, seed=4).to(DEV)
    dil, _ = ops.dilate_boxes(boxes, 640, 640)
    idx = (torch.arange(300, device=DEV) % 4).to(torch.int32)
    p = ops.crop_resize(imgs, dil, idx, 300, "bilinear", 224, 16, "patch")
    la, ta = m.forward_patches(p, 300)
    lb, tb = m.forward_patches(p[: 37 * 196].contiguous(), 37)
    assert torch.allclose(la[:37], lb, rtol=1e-3, atol=1e-3)
    # oracle on a few of them
    chw = ops.crop_resize(imgs, dil, idx, 300, "bilinear", 224, 16, "chw")
    sel = [0, 123, 255, 256, 299]
    ref = O.vit_forward(sd, chw[sel].cpu())
    assert rel_err(la[sel].cpu(), ref) <= REL_TOL
    assert ta[sel].cpu().tolist() == ref.argmax(1).tolist()


@pytest.mark.parametrize("n", [1, 37, 300])
def test_last_block_on_class_rows_is_bit_identical(full_model, n):
    """yv_vit_last_block_cls_rows: proj / norm2 / MLP of the last block on the class-token rows alone.  No operation
    after the last attention mixes rows and every GEMM row is computed independently, so logits, backbone features
    and top-1 equal the full computation BIT FOR BIT (1, 37 and 300 crops: partial row tiles of every size)."""
    m, _ = full_model
    imgs = synth.make_images(4, 640, 640, seed=5).to(DEV)
    boxes = synth.make_boxes_int(n, seed=6).to(DEV)
    dil, _ = ops.dilate_boxes(boxes, 640, 640)
    idx = (torch.arange(n, device=DEV) % 4).to(torch.int32)
    p = ops.crop_resize(imgs, dil, idx, n, "bilinear", 224, 16, "patch")
    la, ta, ha = m.forward_patches(p, n, return_hidden=True)
    try:
        m.last_block_cls_rows(True)
        lb, tb, hb = m.forward_patches(p, n, return_hidden=True)
    finally:
        m.last_block_cls_rows(False)
    lc, tc = m.forward_patches(p, n)
    assert torch.equal(la, lb) and torch.equal(ta, tb) and torch.equal(ha, hb)
    assert torch.equal(la, lc) and torch.equal(ta, tc)          # and the switch goes back


def test_pipeline_cfg1_golden(golden, full_model):
    g = golden("pipeline_cfg1")
    m, _ = full_model
    head = synth.make_head(1, nc=80, A=8400, n_obj=int(g["n_obj"]), n_jit=int(g["n_jit"]), seed=int(g["seed"]))
    img = synth.make_images(1, 640, 640, seed=int(g["seed"]))
    r = HotPath(m, HotPathConfig())(head.to(DEV), img.to(DEV), keep_intermediates=True)
    k = int(r.keep_counts[0])
    assert r.keep[0, :k].cpu().tolist() == g["keep"].tolist()
    n = int(r.det_count[0])
    assert np.array_equal(r.crops.det_box[0, :n].cpu().numpy(), g["box"])
    assert np.array_equal(r.crops.det_crop_box[0, :n].cpu().numpy(), g["crop_box"])
    assert np.array_equal(r.crops.det_anchor[0, :n].cpu().numpy(), g["anchor"])
    assert np.array_equal(r.crops.det_label[0, :n].cpu().numpy(), g["label"])
    assert r.n_crops == int(g["valid"].sum())
    assert rel_err(r.logits.cpu(), torch.from_numpy(g["logits"])) <= REL_TOL
    assert r.top1.cpu().tolist() == g["cls"].tolist()
    rec = r.records[0, :n].cpu()
    assert rec[:, 4].int().tolist() == g["cls"].tolist() and bool((rec[:, 7] == 1).all())


def test_pipeline_batch_against_oracle(full_model):
    """8 images end to end: integer outputs bit-exact per image, logits within 1e-2, classes identical."""
    m, sd = full_model
    head = synth.make_head(8, 80, 8400, n_obj=3, n_jit=12, seed=21)
    imgs = synth.make_images(8, 640, 640, seed=22)
    r = HotPath(m, HotPathConfig())(head.to(DEV), imgs.to(DEV), keep_intermediates=True)
    rec = r.records.cpu()
    det_crop, logits = r.crops.det_crop.cpu(), r.logits.cpu()
    for b in range(8):
        ref = O.pipeline_image(head[b], imgs[b].numpy(), sd)
        n = int(r.det_count[b])
        assert n == len(ref["keep"])
        assert np.array_equal(r.crops.det_box[b, :n].cpu().numpy(), ref["box"])
        assert np.array_equal(r.crops.det_crop_box[b, :n].cpu().numpy(), ref["crop_box"])
        assert rec[b, :n, 4].int().tolist() == ref["cls"].tolist(), b
        assert bool(ref["valid"].all())
        got = logits[det_crop[b, :n].long()]
        assert rel_err(got, torch.from_numpy(ref["logits"])) <= REL_TOL, b


@pytest.mark.parametrize("slices,chunk", [(1, 256), (3, 256), (8, 16), (13, 7)])
def test_run_host_equals_resident_path(full_model, slices, chunk):
    """The pipelined host-buffer entry point (sliced upload on a second stream, crops gathered into
    fixed ViT chunks) returns exactly the records / logits of the resident call on the same inputs."""
    m, _ = full_model
    head = synth.make_head(13, 80, 8400, n_obj=3, n_jit=12, seed=31)
    imgs = synth.make_images(13, 640, 640, seed=32)
    hp = HotPath(m, HotPathConfig())
    ref = hp(head.to(DEV), imgs.to(DEV))
    rec_h = torch.empty(13, 100, 8).pin_memory()
    for _ in range(2):                      # second call reuses the staging buffers
        r = hp.run_host(head.pin_memory(), imgs.pin_memory(), rec_h, device=DEV, slices=slices, vit_chunk=chunk)
    assert r.n_crops == ref.n_crops
    assert torch.equal(r.det_count.cpu(), ref.det_count.cpu())
    assert torch.equal(rec_h, ref.records.cpu())
    assert torch.equal(r.top1.cpu(), ref.top1.cpu())
    # chunk boundaries differ from the resident call (different GEMM M), the arithmetic per row does not
    assert torch.equal(r.logits.cpu(), ref.logits.cpu())


def test_run_host_no_detections(full_model):
    m, _ = full_model
    head = torch.zeros(4, 84, 8400)
    imgs = synth.make_images(4, 640, 640, seed=33)
    r = HotPath(m, HotPathConfig()).run_host(head, imgs, device=DEV, slices=2)
    assert r.n_crops == 0 and float(r.records[..., 7].sum()) == 0.0


@pytest.mark.parametrize("depth", [2, 12])
def test_vit_b8_reference_literal_model(depth):
    """Row f4: the reference-literal vit_base_patch8_224 shape (785 tokens, utils/class_config.py:21) runs end to
    end (patch-8 crop layout, K = 192 patch-embed GEMM, 7-key-block flash attention) at its full depth of 12."""
    cfg = O.ViTConfig(patch=8, depth=depth)
    sd = O.make_vit_weights(cfg, seed=9)
    m = ViTModel(ViTConfig(patch=8, depth=depth), DEV).load_state_dict(sd)
    x = torch.randn(3, 3, 224, 224, generator=torch.Generator().manual_seed(3)).clamp_(-1, 1)
    ref = O.vit_forward(sd, x, cfg)
    got = m(x.to(DEV)).cpu()
    assert rel_err(got, ref) <= REL_TOL
    assert got.argmax(1).tolist() == ref.argmax(1).tolist()
    try:                                            # the class-rows-only last block at 785 tokens: bit-identical
        got2 = m.last_block_cls_rows(True)(x.to(DEV)).cpu()
    finally:
        m.last_block_cls_rows(False)
    assert torch.equal(got, got2)


def test_fp32_residual_stream_mode_parity():
    """YV_VIT_STREAM=f32 keeps the round-1 layout (fp32 residual stream, bf16 deltas folded in by the LayerNorms) as
    an A/B and high-precision mode; it is held to the same bar as the default fp16 stream.  The switch is read
    once per process, so this runs in a child process."""
    import os, subprocess, sys
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from oracle import hotpath as O
from yolov8_vit_b200 import ViTConfig, ViTModel
cfg = O.ViTConfig(depth=12)
sd = O.make_vit_weights(cfg, seed=12)
m = ViTModel(ViTConfig(depth=12), 'cuda:0').load_state_dict(sd)
x = torch.randn(37, 3, 224, 224, generator=torch.Generator().manual_seed(4)).clamp_(-1, 1)
ref = O.vit_forward(sd, x, cfg)
got = m(x.cuda()).cpu()
err = float((got - ref).abs().max() / ref.abs().max())
assert err <= 1e-2, err
top2 = ref.topk(2, dim=1).values
clear = (top2[:, 0] - top2[:, 1]) > 2e-2 * float(ref.abs().max())      # near-ties may flip inside the tolerance
assert int(clear.sum()) >= 30 and torch.equal(got.argmax(1)[clear], ref.argmax(1)[clear])
print('f32 stream ok', err)
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, YV_VIT_STREAM="f32")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=280)
    assert r.returncode == 0 and "f32 stream ok" in r.stdout, r.stdout + r.stderr


def test_layernorm_kernel_mode_parity():
    """YV_VIT_LNFOLD=0 runs the LayerNorm kernels (fp16 stream -> bf16 rows -> bf16 GEMM) instead of the default
    folded form; both are held to the same bar (full depth: tools/vit_err_modes.py measured 3.3e-3 .. 3.9e-3 folded,
    4.4e-3 .. 5.0e-3 with the kernels, 4.8e-3 .. 5.5e-3 with the fp32 stream -- the folded form feeds the tensor cores
    11-bit fp16 rows instead of 8-bit bf16 ones).  Child process: the switch is read once."""
    import os, subprocess, sys
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from oracle import hotpath as O
from yolov8_vit_b200 import ViTConfig, ViTModel
cfg = O.ViTConfig(depth=12)
sd = O.make_vit_weights(cfg, seed=12)
m = ViTModel(ViTConfig(depth=12), 'cuda:0').load_state_dict(sd)
x = torch.randn(37, 3, 224, 224, generator=torch.Generator().manual_seed(4)).clamp_(-1, 1)
ref = O.vit_forward(sd, x, cfg)
got = m(x.cuda()).cpu()
err = float((got - ref).abs().max() / ref.abs().max())
assert err <= 1e-2, err
top2 = ref.topk(2, dim=1).values
clear = (top2[:, 0] - top2[:, 1]) > 2e-2 * float(ref.abs().max())      # near-ties may flip inside the tolerance
assert int(clear.sum()) >= 30 and torch.equal(got.argmax(1)[clear], ref.argmax(1)[clear])
print('ln kernels ok', err)
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, YV_VIT_LNFOLD="0")
    r = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, timeout=280)
    assert r.returncode == 0 and "ln kernels ok" in r.stdout, r.stdout + r.stderr


def test_statistics_kernel_mode_parity():
    """YV_VIT_STATS=kernel: the row statistics of the folded LayerNorms come from the stream-reading kernel instead
    of the proj / fc2 epilogues (the default); same bar.  The two modes differ only in the last bits of the statistics
    (the epilogue sums the values before their fp16 rounding), but twelve blocks of bf16 roundings amplify any such
    perturbation to the level of each mode's own error against the oracle (measured 4.5e-3 of max|logit| between the
    modes, 3.3e-3 .. 3.9e-3 each against the oracle), so the modes are only required to agree within the same 1e-2.
    Child process: the switch is read once."""
    import os, subprocess, sys
    code = r"""
import sys, torch
sys.path.insert(0, %r)
from oracle import hotpath as O
from yolov8_vit_b200 import ViTConfig, ViTModel
cfg = O.ViTConfig(depth=12)
sd = O.make_vit_weights(cfg, seed=12)
m = ViTModel(ViTConfig(depth=12), 'cuda:0').load_state_dict(sd)
x = torch.randn(37, 3, 224, 224, generator=torch.Generator().manual_seed(4)).clamp_(-1, 1)
ref = O.vit_forward(sd, x, cfg)
got = m(x.cuda()).cpu()
err = float((got - ref).abs().max() / ref.abs().max())
assert err <= 1e-2, err
top2 = ref.topk(2, dim=1).values
clear = (top2[:, 0] - top2[:, 1]) > 2e-2 * float(ref.abs().max())      # near-ties may flip inside the tolerance
assert int(clear.sum()) >= 30 and torch.equal(got.argmax(1)[clear], ref.argmax(1)[clear])
torch.save(got, sys.argv[1])
print('stats mode ok', err)
""" % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    import tempfile
    outs = []
    with tempfile.TemporaryDirectory() as d:
        for mode in ("kernel", "fused"):
            env = dict(os.environ, YV_VIT_STATS=mode)
            f = os.path.join(d, mode + ".pt")
            r = subprocess.run([sys.executable, "-c", code, f], env=env, capture_output=True, text=True, timeout=280)
            assert r.returncode == 0 and "stats mode ok" in r.stdout, r.stdout + r.stderr
            outs.append(torch.load(f))
    assert float((outs[0] - outs[1]).abs().max() / outs[0].abs().max()) <= 1e-2


def test_submit_host_pipelined_batches(full_model):
    """Two batches in flight (submit k+1 before result k, alternating staging slots) give each batch exactly the
    records of its own resident run; a third submit reuses slot 0 safely."""
    m, _ = full_model
    hp = HotPath(m, HotPathConfig())
    batches = [(synth.make_head(6, 80, 8400, n_obj=3, n_jit=12, seed=40 + i).pin_memory(),
                synth.make_images(6, 640, 640, seed=50 + i).pin_memory()) for i in range(3)]
    refs = [hp(h.to(DEV), im.to(DEV)).records.cpu() for h, im in batches]
    outs = [torch.empty(6, 100, 8).pin_memory() for _ in range(3)]
    pend = hp.submit_host(*batches[0], outs[0], device=DEV, slices=3)
    for k in range(3):
        nxt = hp.submit_host(*batches[k + 1], outs[k + 1], device=DEV, slices=3) if k < 2 else None
        r = pend.result()
        assert torch.equal(outs[k], refs[k]) and torch.equal(r.records.cpu(), refs[k]), k
        pend = nxt


def test_submit_host_launch_ahead(full_model):
    """Serving-loop order of bench.py's e2e: batch k+1 is submitted AND launched (its ViT queued behind batch
    k's) before result(k); every batch still gets exactly the records, logits and top-1 of its resident run."""
    m, _ = full_model
    hp = HotPath(m, HotPathConfig())
    batches = [(synth.make_head(5, 80, 8400, n_obj=2 + i % 2, n_jit=10, seed=60 + i).pin_memory(),
                synth.make_images(5, 640, 640, seed=70 + i).pin_memory()) for i in range(4)]
    refs = [hp(h.to(DEV), im.to(DEV)) for h, im in batches]
    outs = [torch.empty(5, 100, 8).pin_memory() for _ in range(4)]
    pend = hp.submit_host(*batches[0], outs[0], device=DEV, slices=2).launch()
    for k in range(4):
        nxt = hp.submit_host(*batches[k + 1], outs[k + 1], device=DEV, slices=2).launch() if k < 3 else None
        r = pend.result()
        assert r is pend.result()                                  # idempotent
        assert torch.equal(outs[k], refs[k].records.cpu()) and torch.equal(r.records, refs[k].records), k
        assert r.n_crops == refs[k].n_crops and torch.equal(r.top1, refs[k].top1) and torch.equal(r.logits, refs[k].logits)
        pend = nxt
