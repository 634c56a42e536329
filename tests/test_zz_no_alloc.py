"""yv_nms_batched_noalloc (no counterpart in the reference): the NMS
kernels with their scratch in the library's built-in 4 MiB __device__ array instead of a caller workspace, for callers
that cannot allocate.  Checked through the C-ABI against the existing entry point (yv_nms_batched) -- keep lists are
integer work, so the bar is bit-exact like every other NMS test -- on two calls in a row whose shapes differ (the
second call finds the first call's contents in the array), with the device's free memory unchanged across the calls,
and the size limit is rejected with a clear error.  This file sorts last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from yolov8_vit_b200 import ops, synth
from yolov8_vit_b200._lib import YvError, load, ptr, stream_ptr

DEV = "cuda:0"


def _same_as_workspace_entry(boxes, scores, counts, thr, max_keep=0):
    B, S = scores.shape
    ref_keep, ref_n = ops.nms_batched(boxes, scores, counts, thr, max_keep)
    keep = torch.full((B, S), -7, dtype=torch.int32, device=DEV)          # outputs are the caller's, pre-filled with junk
    n = torch.full((B,), -7, dtype=torch.int32, device=DEV)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info(0)[0]
    alloc0 = torch.cuda.memory_allocated(0)
    ops.nms_batched_noalloc(boxes, scores, counts, keep, n, thr, max_keep)
    torch.cuda.synchronize()
    used = (free0 - torch.cuda.mem_get_info(0)[0], torch.cuda.memory_allocated(0) - alloc0)
    assert torch.equal(n, ref_n)
    for b in range(B):
        k = int(n[b])
        assert torch.equal(keep[b, :k], ref_keep[b, :k]), b
    return used, int(n.sum())


def test_noalloc_nms_matches_workspace_entry_on_two_calls_in_a_row():
    # call 1: the detector shape (cfg2: 768 candidates per image, per-image counts on the device), one CTA per image
    head = synth.make_head(16, nc=80, A=8400, n_obj=32, n_jit=24, seed=synth.SEED0 + 2).to(DEV)
    bx, sc, _, _, cnt = ops.decode_filter(head, 0.35, cap=1024)      # candidate capacity 1024: 16 x 1024 x 56 B of scratch
    assert 500 < int(cnt.max()) < 1024
    _, kept1 = _same_as_workspace_entry(bx, sc, cnt, 0.45, max_keep=100)
    # call 2, straight after, another shape: long dense lists without counts (the 4-CTA cluster path), lists that
    # reach further into the array than call 1 did and find its contents at the front
    b2, s2 = synth.make_nms_stress(2, 9000, seed=synth.SEED0 + 7)
    b2, s2 = b2.to(DEV), s2.to(DEV)
    used2, kept2 = _same_as_workspace_entry(b2, s2, None, 0.6)
    # call 3: call 1's inputs again, after the long lists have overwritten the array
    used3, kept3 = _same_as_workspace_entry(bx, sc, cnt, 0.45, max_keep=100)
    assert kept1 == kept3 and kept1 > 0 and kept2 > 0
    # nothing was allocated by calls 2 and 3 (call 1 may have loaded the CUDA module lazily): the device's free
    # memory and torch's own allocator are where they were
    assert used2 == (0, 0) and used3 == (0, 0), (used2, used3)


def test_noalloc_nms_rejects_sizes_beyond_the_array():
    B, S = 4, 30000                                   # 4 x 30 000 x 56 B = 6.7 MB > 4 MiB
    assert load().yv_nms_workspace_bytes(B, S) > ops.NMS_NOALLOC_BYTES
    bx = torch.zeros(B, S, 4, device=DEV)
    sc = torch.zeros(B, S, device=DEV)
    keep = torch.zeros(B, S, dtype=torch.int32, device=DEV)
    n = torch.full((B,), -1, dtype=torch.int32, device=DEV)
    with pytest.raises(YvError, match="built-in array holds"):
        ops.nms_batched_noalloc(bx, sc, None, keep, n, 0.5)
    torch.cuda.synchronize()
    assert n.tolist() == [-1] * B                     # nothing ran
    # the largest shape that fits still runs
    S_ok = 18000
    assert load().yv_nms_workspace_bytes(B, S_ok) <= ops.NMS_NOALLOC_BYTES
    b3, s3 = synth.make_nms_stress(B, S_ok, seed=synth.SEED0 + 8)
    _same_as_workspace_entry(b3.to(DEV), s3.to(DEV), None, 0.5)
