"""Out-of-bounds WRITE check with guard bands (compute-sanitizer is closed on this GPU pool -- its wrapper exits 86 --
so this is the repo's own bounds check; tools/sanitize.sh stays for pools where it runs).

Every device buffer the Python layer allocates for a kernel (outputs AND workspaces: `torch.empty` / `zeros` /
`empty_like` inside yolov8_vit_b200.ops / .vit) is replaced by the interior of a larger allocation whose 4 KB bands on
both sides hold a byte pattern; after the kernels ran -- at awkward sizes: row / column counts that are not multiples
of any tile, single rows, token counts on both sides of a block boundary -- every band must be intact.  Results are
compared with the unguarded run, so the guard itself is shown not to change anything.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu

from yolov8_vit_b200 import ViTConfig, ViTModel, ops, synth
from yolov8_vit_b200 import vit as vit_mod

DEV = "cuda:0"
BAND = 4096
PATTERN = 0xA5


class Guard:
    """torch stand-in for one module: allocation calls return guarded interiors, everything else is torch's."""

    def __init__(self):
        self.allocs = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _alloc(self, shape, dtype, device, zero=False):
        shape = tuple(int(s) for s in (shape[0] if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)) else shape))
        numel = 1
        for s in shape:
            numel *= s
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        pad = (-nbytes) % 512                            # keep the rear band's start 512-byte aligned like the front
        raw = torch.full((BAND + nbytes + pad + BAND,), PATTERN, dtype=torch.uint8, device=device)
        inner = raw[BAND:BAND + nbytes]
        if zero:
            inner.zero_()
        self.allocs.append((raw, nbytes))
        return inner.view(dtype).view(shape)

    def empty(self, *shape, dtype=torch.float32, device=None, **kw):
        return self._alloc(shape, dtype, device)

    def zeros(self, *shape, dtype=torch.float32, device=None, **kw):
        return self._alloc(shape, dtype, device, zero=True)

    def empty_like(self, t, **kw):
        return self._alloc(tuple(t.shape), t.dtype, t.device)

    def check(self):
        torch.cuda.synchronize()
        assert self.allocs
        for k, (raw, nbytes) in enumerate(self.allocs):
            front, rear = raw[:BAND], raw[BAND + nbytes:]
            assert bool((front == PATTERN).all()), f"allocation {k} ({nbytes} B): write BEFORE the buffer"
            assert bool((rear == PATTERN).all()), f"allocation {k} ({nbytes} B): write PAST the buffer"
        n = len(self.allocs)
        self.allocs.clear()
        return n


@pytest.fixture()
def guard(monkeypatch):
    g = Guard()
    monkeypatch.setattr(ops, "torch", g)
    monkeypatch.setattr(vit_mod, "torch", g)
    return g


def test_front_stage_kernels_stay_inside_their_buffers(guard):
    head = synth.make_head(3, nc=80, A=8400, n_obj=5, n_jit=9, seed=1).to(DEV)
    imgs = synth.make_images(3, 333, 517, seed=2).to(DEV)                 # odd image size
    cand = ops.decode_filter(head, 0.35)
    cut = ops.decode_filter(head, 0.35, cap=7)                            # more candidates than capacity
    assert int(cut[4].max()) == 7
    keep, kc = ops.nms_batched(cand[0], cand[1], cand[4], 0.45)
    k3, c3 = ops.nms_batched(cand[0], cand[1], cand[4], 0.45, max_keep=3)
    cl = ops.build_crops(cand[0], cand[1], cand[2], cand[3], keep, kc, 100, 517, 333)
    n = int(cl.n_crops.item())
    assert n > 0
    for interp in ("nearest", "bilinear"):
        for layout in ("chw", "patch"):
            ops.crop_resize(imgs, cl.crop_box, cl.crop_img, n, interp, 224, 16, layout)
    ops.pack_results(cl, None)
    ops.dilate_boxes(synth.make_boxes_int(37, seed=3).to(DEV), 640, 640)
    ops.letterbox_blob(synth.make_images(2, 270, 481, seed=4).to(DEV), (640, 640), bgr=True)
    levels = synth.LEVELS_640
    ops.decode_filter_raw(synth.make_raw_head(2, 80, levels, n_obj=4, seed=5).to(DEV), levels, 0.35)
    # long lists: the 4-CTA cluster path (>= 4096 boxes, no counts) and the single-CTA path just below it
    for nbox in (4095, 4097, 6001):
        bx, sc = synth.make_nms_stress(2, nbox, seed=6)
        ops.nms_batched(bx.to(DEV), sc.to(DEV), None, 0.6)
    assert guard.check() > 30


def test_vit_kernels_stay_inside_their_buffers(guard):
    g = torch.Generator().manual_seed(7)
    for M, N, K in ((129, 264, 128), (1, 256, 64), (300, 1000, 768), (257, 776, 192)):
        a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
        w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
        b = torch.randn(N, generator=g).to(DEV)
        ops.gemm_bf16(a, w, b)
        ops.gemm_bf16(a, w, b, gelu=True)
        ops.gemm_bf16(a, w, b, out_f32=True)
    for M, N, K in ((333, 272, 512), (1, 256, 768), (130, 2304, 768)):
        x = torch.randn(M, K, generator=g).to(torch.float16).to(DEV)
        w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
        ops.linear_ln_f16(x, w, torch.randn(N, generator=g).to(DEV), torch.ones(K, device=DEV), torch.zeros(K, device=DEV), 1e-6, True)
    for M, N, K in ((130, 768, 64), (1, 256, 128), (259, 1024, 192)):
        a = (torch.randn(M, K, generator=g) * 0.5).to(torch.bfloat16).to(DEV)
        w = (torch.randn(N, K, generator=g) * 0.05).to(torch.bfloat16).to(DEV)
        ops.residual_linear_f16(a, w, None, torch.randn(M, N, generator=g).to(torch.float16).to(DEV))
    for rows, dim in ((1, 768), (33, 256), (1001, 1024)):
        ops.layernorm_bf16(torch.randn(rows, dim, generator=g).to(torch.bfloat16).to(DEV), torch.ones(dim, device=DEV), torch.zeros(dim, device=DEV))
    for n, T in ((3, 197), (2, 5), (1, 1), (2, 128), (2, 129), (1, 300), (1, 785)):
        ops.attention_bf16((torch.randn(n, T, 3 * 128, generator=g) * 0.5).to(torch.bfloat16).to(DEV), 2)
    assert guard.check() >= 25


@pytest.mark.parametrize("patch,n", [(16, 5), (16, 1), (8, 2)])
def test_vit_forward_stays_inside_workspace_and_outputs(guard, monkeypatch, patch, n):
    """The whole forward inside a guarded workspace (sized by yv_vit_workspace_bytes) and guarded logits / top-1 /
    hidden buffers; logits equal to the unguarded model's bit for bit."""
    cfg = ViTConfig(patch=patch, depth=2)
    sd = synth.make_vit_weights(cfg, seed=8)
    x = torch.randn(n, 3, 224, 224, generator=torch.Generator().manual_seed(9)).clamp_(-1, 1).to(DEV)
    m = ViTModel(cfg, DEV).load_state_dict(sd)
    got = m(x).clone()
    p = vit_mod.patchify_bf16(x, cfg.patch)
    logits, top1, hidden = m.forward_patches(p, n, return_hidden=True)
    assert torch.equal(logits, got) and torch.equal(top1.long(), got.argmax(1))
    assert guard.check() >= 6                           # workspace, logits, top-1, hidden, patches ...
    monkeypatch.setattr(ops, "torch", torch)            # the comparison run allocates plainly
    monkeypatch.setattr(vit_mod, "torch", torch)
    ref = ViTModel(cfg, DEV).load_state_dict(sd)(x)
    assert torch.equal(got, ref)
