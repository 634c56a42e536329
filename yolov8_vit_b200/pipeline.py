"""The batched second-stage hot path on one B200:

    head [B,4+nc,A] -> decode + conf filter (K1) -> custom_nms (K2/K3) -> round + dilate + clamp
    -> crop list (a3) -> crop + resize + normalise, patch-major bf16 (K4) -> ViT-B/16 + wrapper
    head + argmax (K5-K8) -> per-image result records (K9).

Order of operations = the detector wrapper's per-image loop
(YOLOTensorRT_yolodet_py_解读.md:64-100 + the authors' NMS / dilation / classify additions,
app.py:61), batched over images.  Every stage is a C-ABI call into libyv8vit.so.
"""
from __future__ import annotations

import dataclasses
import threading
from contextlib import contextmanager
from dataclasses import dataclass, field

import torch

from . import ops
from .vit import ViTConfig, ViTModel


@dataclass(frozen=True)
class HotPathConfig:
    """The reference's hard-coded constants (SURVEY.md §5 'config'): conf 0.35 (解读.md:94), iou 0.45
    (README.md:62), alpha 0.1 (README.md:106) as a rational, 224 output (app.py:40), top-100
    detections (test.ipynb:22)."""
    conf: float = 0.35
    iou: float = 0.45
    alpha: tuple = (1, 10)
    dilate_mode: str = "full"          # 'infer_half' = crop_image(training=False), utils/trainClass.py:85-91
    interp: str = "bilinear"           # 'nearest' = the reference-literal A.Resize(INTER_NEAREST), app.py:40
    out_size: int = 224
    max_det: int = 100
    ratio: float = 1.0                 # un-letterbox (解读.md:89-92)
    dwdh: tuple = (0.0, 0.0)
    cand_cap: int | None = None        # candidates kept per image after the conf filter (None = all anchors)


@dataclass
class HotPathResult:
    records: torch.Tensor              # [B,max_det,8] f32: x0,y0,x1,y1,cls,score,anchor,valid
    det_count: torch.Tensor            # [B] i32
    n_crops: int
    logits: torch.Tensor | None = None  # [n_crops, num_classes] f32
    top1: torch.Tensor | None = None    # [n_crops] i32
    crops: ops.CropList | None = None
    keep: torch.Tensor | None = None
    keep_counts: torch.Tensor | None = None
    cand: tuple | None = None


@contextmanager
def _nvtx(name: str):
    """NVTX range per stage (visible in nsys / ncu --nvtx; a no-op otherwise)."""
    torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        torch.cuda.nvtx.range_pop()


class HotPath:
    """Stateless apart from the (optional) classifier.  The resident entry points (``__call__``, ``detect``,
    ``classify``, ``from_candidates``) may be called from several threads that share one instance and one
    model, as the reference's Flask request threads share ``model_list`` (app.py:24-37), PROVIDED each thread
    works on its own CUDA stream: the 4-byte pinned read-back buffer is per thread and the ViT activation
    workspace is per stream (``ViTModel._workspace``); two threads on the SAME stream are ordered by the
    stream.  The host-buffer entry points (``submit_host`` / ``run_host``) own staging slots and are not
    re-entrant: one HotPath per serving thread."""

    def __init__(self, model: ViTModel | None = None, cfg: HotPathConfig = HotPathConfig()):
        self.model = model
        self.cfg = cfg
        self.extra_models = ()         # further classifiers of the reference's model_list (logits are summed)
        self._tls = threading.local()  # per-thread pinned crop-count buffer

    def detect(self, head: torch.Tensor):
        """K1..K3: candidates and keep lists (device tensors, no sync)."""
        c = self.cfg
        with _nvtx("yv8vit.decode_filter"):
            cand = ops.decode_filter(head, c.conf, c.ratio, c.dwdh, cap=c.cand_cap)
        boxes, scores, labels, anchors, counts = cand
        with _nvtx("yv8vit.custom_nms"):
            keep, keep_counts = ops.nms_batched(boxes, scores, counts, c.iou, max_keep=c.max_det)
        return cand, keep, keep_counts

    def __call__(self, head: torch.Tensor, images: torch.Tensor, keep_intermediates: bool = False,
                 levels=None) -> HotPathResult:
        """head [B,4+nc,A] f32 (or, with ``levels`` = [(h,w,stride)]x3, the Detect head's RAW output
        [B,64+nc,A]: row f1), images [B,H,W,3] u8 (both on the GPU)."""
        c = self.cfg
        if levels is None:
            cand, keep, keep_counts = self.detect(head)
        else:
            cand = ops.decode_filter_raw(head, levels, c.conf, c.ratio, c.dwdh, cap=c.cand_cap)
            keep, keep_counts = ops.nms_batched(cand[0], cand[1], cand[4], c.iou, max_keep=c.max_det)
        return self.classify(cand, keep, keep_counts, images, keep_intermediates)

    def from_candidates(self, boxes: torch.Tensor, scores: torch.Tensor, labels: torch.Tensor,
                        counts: torch.Tensor, images: torch.Tensor, keep_intermediates: bool = False):
        """Entry for detectors that already emit per-image candidate lists (the reference's engine with its
        EfficientNMS plugin, test.ipynb:20-24): boxes [B,S,4] f32 xyxy in image pixels, scores [B,S] f32,
        labels [B,S] i32, counts [B] i32 -> custom_nms -> dilation -> crop -> ViT as usual."""
        anchors = torch.arange(scores.shape[1], dtype=torch.int32, device=scores.device).repeat(scores.shape[0], 1)
        keep, keep_counts = ops.nms_batched(boxes, scores, counts, self.cfg.iou, max_keep=self.cfg.max_det)
        cand = (boxes.contiguous(), scores.contiguous(), labels.contiguous(), anchors, counts)
        return self.classify(cand, keep, keep_counts, images, keep_intermediates)

    def classify(self, cand, keep, keep_counts, images: torch.Tensor, keep_intermediates: bool = False):
        """a3..K9: crop list -> crop + resize -> ViT (ensemble: logits of ``extra_models`` are added, the
        reference's ``for net in model_list`` loop, app.py:30-37) -> records."""
        c = self.cfg
        B, H, W, _ = images.shape
        boxes, scores, labels, anchors, _ = cand
        with _nvtx("yv8vit.build_crops"):
            cl = ops.build_crops(boxes, scores, labels, anchors, keep, keep_counts, c.max_det, W, H, c.alpha,
                                 c.dilate_mode)
        logits = top1 = None
        n = 0
        if self.model is not None:
            # The crop count sizes the ViT launches (GEMM grids, TMA maps): one 4-byte read-back, the only host
            # sync of the path (~20 us of a 40 ms step; the GPU has the whole ViT queued again right after it).
            pin = getattr(self._tls, "pin", None)
            if pin is None:
                pin = self._tls.pin = torch.empty(1, dtype=torch.int32).pin_memory()
            pin.copy_(cl.n_crops, non_blocking=True)
            torch.cuda.current_stream(images.device).synchronize()
            n = min(int(pin.item()), cl.crop_box.shape[0])
            with _nvtx("yv8vit.crop_resize"):
                patches = ops.crop_resize(images, cl.crop_box, cl.crop_img, n, c.interp, c.out_size,
                                          self.model.cfg.patch, "patch")
            with _nvtx("yv8vit.vit_forward"):
                logits, top1 = self.model.forward_patches(patches, n)
            for m in getattr(self, "extra_models", ()):
                logits = logits + m.forward_patches(patches, n)[0]
            if getattr(self, "extra_models", ()) and n:
                top1 = logits.argmax(dim=1).to(torch.int32)
        rec = ops.pack_results(cl, top1)
        r = HotPathResult(records=rec, det_count=cl.det_count, n_crops=n, logits=logits, top1=top1)
        if keep_intermediates:
            r.crops, r.keep, r.keep_counts, r.cand = cl, keep, keep_counts, cand
        return r

    # ------------------------------------------------------------------ host-buffer entry points
    def _host_state(self, head_h: torch.Tensor, imgs_h: torch.Tensor, dev, slices: int, slot: int):
        key = (tuple(head_h.shape), tuple(imgs_h.shape), str(dev), slices)
        sets = getattr(self, "_hs", None)
        if sets is None or sets["key"] != key:
            sets = {"key": key, "front": torch.cuda.Stream(device=dev, priority=-1), "slots": {}, "next": 0}
            self._hs = sets
        st = sets["slots"].get(slot)
        if st is None:
            st = {"head": torch.empty(head_h.shape, dtype=torch.float32, device=dev),
                  "imgs": torch.empty(imgs_h.shape, dtype=torch.uint8, device=dev),
                  "pin": torch.empty(slices, dtype=torch.int32).pin_memory(),
                  "events": [torch.cuda.Event() for _ in range(slices)],
                  "free": None,                # recorded on the caller's stream when the slot's buffers are no longer read
                  "patches": None, "top1": None, "logits": None, "cap": 0}
            sets["slots"][slot] = st
        return sets, st

    def _grow(self, st, need: int, used: int, dev):
        """Crop-indexed buffers (patches, logits, top1) with geometric growth; `used` crops are preserved."""
        if need <= st["cap"]:
            return
        cap = max(need, 2 * st["cap"], 1024)
        m = self.model.cfg
        g = (self.cfg.out_size // m.patch) ** 2
        new = (torch.empty(cap * g, 3 * m.patch * m.patch, dtype=torch.bfloat16, device=dev),
               torch.empty(cap, m.num_classes, dtype=torch.float32, device=dev),
               torch.empty(cap, dtype=torch.int32, device=dev))
        if used and st["patches"] is not None:
            new[0][:used * g].copy_(st["patches"][:used * g])
            new[1][:used].copy_(st["logits"][:used])
            new[2][:used].copy_(st["top1"][:used])
        st["patches"], st["logits"], st["top1"], st["cap"] = new[0], new[1], new[2], cap

    def submit_host(self, head_h: torch.Tensor, imgs_h: torch.Tensor, out_records_h: torch.Tensor | None = None,
                    device="cuda:0", slices: int = 8, vit_chunk: int | None = None, ramp=()) -> "PendingHost":
        """Asynchronous half of :meth:`run_host`: enqueues the sliced upload and the per-slice decode ->
        custom_nms -> crop list on the high-priority front stream and returns at once.  ``.result()`` of the
        returned handle runs the ViT part and delivers the records.  Two staging slots alternate, so a server
        loop ``nxt = submit_host(batch i+1).launch(); cur.result(); cur = nxt`` uploads batch i+1 while batch i is
        in the ViT and has batch i+1's ViT queued behind it -- the end-to-end rate then approaches the resident
        rate (bench.py's `e2e`).  At most two batches may be in flight (two slots)."""
        if self.model is None:
            raise ValueError("the host-buffer path needs a classifier")
        if head_h.is_cuda or imgs_h.is_cuda:
            raise ValueError("submit_host / run_host take host tensors; call the instance for device-resident inputs")
        c = self.cfg
        dev = torch.device(device)
        B, H, W, _ = imgs_h.shape
        slices = max(1, min(slices, B))
        sets = getattr(self, "_hs", None)
        slot = 0 if sets is None else sets.get("next", 0)
        sets, st = self._host_state(head_h, imgs_h, dev, slices, slot)
        sets["next"] = slot ^ 1
        front = sets["front"]
        bounds = [(B * i) // slices for i in range(slices + 1)]
        if st["free"] is not None:
            front.wait_event(st["free"])       # the previous user of this slot no longer reads its buffers
        else:
            front.wait_stream(torch.cuda.current_stream(dev))
        cls = []
        with torch.cuda.stream(front):
            for i in range(slices):
                a, b = bounds[i], bounds[i + 1]
                hd, im = st["head"][a:b], st["imgs"][a:b]
                hd.copy_(head_h[a:b], non_blocking=True)
                im.copy_(imgs_h[a:b], non_blocking=True)
                cand, keep, keep_counts = self.detect(hd)
                boxes, scores, labels, anchors, _ = cand
                cl = ops.build_crops(boxes, scores, labels, anchors, keep, keep_counts, c.max_det, W, H, c.alpha,
                                     c.dilate_mode)
                st["pin"][i:i + 1].copy_(cl.n_crops, non_blocking=True)
                st["events"][i].record(front)
                cls.append(cl)
        return PendingHost(self, st, cls, bounds, dev, out_records_h, vit_chunk, ramp, front)

    def run_host(self, head_h: torch.Tensor, imgs_h: torch.Tensor, out_records_h: torch.Tensor | None = None,
                 device="cuda:0", slices: int = 8, vit_chunk: int | None = None, ramp=()) -> HotPathResult:
        """The same path from HOST buffers (pinned memory recommended): head [B,4+nc,A] f32 and images
        [B,H,W,3] u8 on the CPU -> records on the CPU (`out_records_h`, pinned, optional).

        The batch is cut into `slices` image slices.  A high-priority "front" stream uploads slice
        after slice and runs decode -> custom_nms -> crop list on each as soon as it has landed; the
        caller's stream gathers the crops of every finished slice into one patch buffer and runs
        the ViT over `vit_chunk`-crop chunks as soon as enough crops exist (None = auto, see launch()), so the upload of
        slice i+1.. overlaps the ViT of slice ..i (`ramp` = smaller first chunks, e.g. (64, 128); measured
        not to pay off: small chunks run the GEMMs at half efficiency).  Host syncs: one 4-byte crop count per
        slice (it sizes the ViT launches) and the final record download.  = submit_host(...).result().
        Not re-entrant per instance (it owns staging buffers); use one HotPath per thread.
        """
        return self.submit_host(head_h, imgs_h, out_records_h, device, slices, vit_chunk, ramp).result()


class PendingHost:
    """Handle of :meth:`HotPath.submit_host`: the upload + detection of the batch are in flight."""

    def __init__(self, hp, st, cls, bounds, dev, out_records_h, vit_chunk, ramp, front):
        self.hp, self.st, self.cls, self.bounds, self.dev = hp, st, cls, bounds, dev
        self.out_records_h, self.vit_chunk, self.ramp, self.front = out_records_h, vit_chunk, ramp, front
        self._res = None
        self._launched = False
        self._done = None

    def launch(self) -> "PendingHost":
        """Enqueues the rest of the batch on the caller's stream -- per slice: wait for its crop count (a 4-byte
        host sync on work that ran on the front stream), crop-resize into the patch buffer, ViT over fixed chunks;
        then the records and their download -- and returns without waiting for the GPU.  A serving loop calls
        ``submit_host(k+1).launch()`` BEFORE ``result()`` of batch k, so the GPU already holds batch k+1's work
        when batch k finishes (no idle gap while the host prepares the next batch)."""
        if self._launched:
            return self
        hp, st, cls, bounds, dev = self.hp, self.st, self.cls, self.bounds, self.dev
        c, model, vit_chunk = hp.cfg, hp.model, self.vit_chunk
        if vit_chunk is None:
            # auto: while an earlier batch is still in the ViT this batch's upload is hidden anyway, so its ViT
            # runs as ONE pass once all slices are in (fewest launches, no remainder pass: 43.5 vs 46.7 ms per
            # batch); on an idle GPU the ViT starts on 256-crop chunks as soon as the first slices have landed
            prev = getattr(hp, "_last_done", None)
            vit_chunk = (1 << 30) if (prev is not None and not prev.query()) else 256
        slices, B = len(cls), bounds[-1]
        main = torch.cuda.current_stream(dev)
        total = done = 0
        offs = []
        g = (c.out_size // model.cfg.patch) ** 2
        sizes = [int(r) for r in self.ramp if 0 < int(r) < vit_chunk]
        launched = 0

        def next_chunk():
            return sizes[launched] if launched < len(sizes) else vit_chunk

        def vit(upto):
            nonlocal done, launched
            while done < upto:
                n = min(next_chunk(), upto - done)
                launched += 1
                model.forward_patches(st["patches"][done * g:(done + n) * g], n,
                                      out_logits=st["logits"][done:done + n], out_top1=st["top1"][done:done + n])
                done += n

        for i in range(slices):
            a, b = bounds[i], bounds[i + 1]
            st["events"][i].synchronize()
            n_i = min(int(st["pin"][i].item()), cls[i].crop_box.shape[0])
            main.wait_event(st["events"][i])
            hp._grow(st, total + n_i, total, dev)
            if n_i:
                ops.crop_resize(st["imgs"][a:b], cls[i].crop_box, cls[i].crop_img, n_i, c.interp, c.out_size,
                                model.cfg.patch, "patch", out=st["patches"][total * g:(total + n_i) * g])
            offs.append(total)
            total += n_i
            while total - done >= next_chunk():                        # only full chunks while slices remain
                vit(done + next_chunk())
        vit(total)
        rec = torch.empty(B, c.max_det, 8, dtype=torch.float32, device=dev)
        for i in range(slices):
            a, b = bounds[i], bounds[i + 1]
            ops.pack_results(cls[i], st["top1"][offs[i]:] if total else None, out=rec[a:b])
        if self.out_records_h is not None:
            self.out_records_h.copy_(rec, non_blocking=True)
        # results are copied out of the slot's crop-indexed buffers (the next-but-one submit reuses them)
        self._res = HotPathResult(records=rec, det_count=torch.cat([cl.det_count for cl in cls]), n_crops=total,
                                  logits=st["logits"][:total].clone() if total else None,
                                  top1=st["top1"][:total].clone() if total else None)
        # The crop lists were allocated on the FRONT stream and are read here on the caller's stream (crop_resize,
        # and pack_results after the whole ViT): tell the caching allocator, or their blocks would go back to the
        # front-stream pool as soon as the last reference drops and the next submit_host -- which only waits for
        # the OTHER slot's `free` event -- would overwrite them while this batch's pack_results is still queued.
        for cl in cls:
            for f in dataclasses.fields(cl):
                getattr(cl, f.name).record_stream(main)
        if st["free"] is None:
            st["free"] = torch.cuda.Event()
        st["free"].record(main)                # the slot's staging buffers may be overwritten after this point
        self._done = torch.cuda.Event()
        self._done.record(main)
        hp._last_done = self._done
        self._launched = True                  # (self.cls stays referenced until result() has seen the batch finish)
        return self

    def result(self) -> HotPathResult:
        """Waits for this batch only (an event, not the whole stream: a later batch may already be queued)."""
        self.launch()
        if self._done is not None:
            self._done.synchronize()
            self._done = None
            self.cls = None
        return self._res


def records_to_tuples(records: torch.Tensor, names=None):
    """[(image_name, class_id, conf, xmin, ymin, xmax, ymax)] sorted by image name -- the return
    value of the wrapper's main() (YOLOTensorRT_yolodet_py_解读.md:103-116).  Host side."""
    rec = records.detach().cpu()
    B = rec.shape[0]
    names = [str(i) for i in range(B)] if names is None else list(names)
    out = []
    for b in range(B):
        for r in rec[b]:
            if r[7] <= 0:
                continue
            out.append((names[b], int(r[4]), float(r[5]), int(r[0]), int(r[1]), int(r[2]), int(r[3])))
    out.sort(key=lambda t: t[0])
    return out
