"""bench.py -- images/sec of the second-stage hot path (BASELINE.json metric) on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A "step" is one pass of the hot path over one batch of synthetic input on every rank:
  head [B,84,8400] f32 -> conf 0.35 -> custom_nms 0.45 -> alpha=0.1 dilation -> 224x224 bilinear
  crop -> ViT-B/16 + wrapper head -> per-image records (+ final NCCL all-gather when N > 1).
Workload = the north star's single-GPU target "YOLOv8n-640 -> ViT-B/16 at batch 256": B = 256
images per GPU, OBJ objects x 24 candidate boxes per image (cfg4's 4 objects/image).  Images are
batch-sharded (weak scaling: per-GPU work fixed), no collective on the data path.

`value`   : whole-job images/s, inputs resident in HBM, CUDA events, max over ranks.
`e2e`     : same metric through the public HotPath API from pinned HOST buffers (H2D of head +
            images and D2H of the records of EVERY step inside the timed region; the serving loop submits
            batch k+1 before collecting batch k, so uploads overlap the previous batch's ViT).
`roofline`: the dominant kernel (tcgen05 GEMM): flops per launch / event-timed launch duration.
`cpu_baseline` / `--impl reference`: the oracle (a verbatim restatement of the reference's torch
            CPU path; the reference itself is not importable) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU = 256
N_OBJ = 4
N_JIT = 24
NC = 80
ANCHORS = 8400
IMG = 640
SEED = 20240410 + 1
FLOP_PER_CROP = 35.128e9           # SURVEY.md §8d (ViT-B/16 + heads)
# ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch, mean of the four block GEMMs of the
# one-pass ViT at M = 1072 crops x 197 = 211 184 rows (profiles/r01_vit_ncu_v7.md): (1624 + 1250 + 610 + 1576) / 4 MB
# = 5.99 KB per token row; scaled by the run's own row count below (the driver's workload gives exactly 1072 crops)
GEMM_DRAM_BYTES_PER_ROW = 1265e6 / 211184


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="images per GPU (0 = the workload's own)")
    ap.add_argument("--workload", default="north_star", choices=["north_star", "cfg4"],
                    help="north_star: YOLOv8n-640, 256 images per GPU (weak scaling; the driver's line). "
                         "cfg4: BASELINE.json configs[3], YOLOv8s-1280 head [512,84,33600] + 512 images "
                         "1280x1280 sharded over the ranks (strong scaling: 512 / N images per GPU)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-vit-chunk", type=int, default=0,
                    help="crops per ViT pass in the host-buffer loop (0 = the API's default: auto)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], d.get("bf16_tflops_sustained", d["bf16_tflops"]), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def set_workload(args):
    """Select the synthetic workload; returns the scaling kind ("weak" / "strong")."""
    global ANCHORS, IMG
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.workload == "cfg4":
        ANCHORS, IMG = 33600, 1280                 # 160^2 + 80^2 + 40^2 anchors of a 1280x1280 input
        args.batch = args.batch or 512 // world
        return "strong"
    args.batch = args.batch or B_PER_GPU
    return "weak"


def workload_name(batch):
    if IMG == 1280:
        return (f"cfg4: YOLOv8s-1280 head [{batch},{4 + NC},{ANCHORS}] f32 + {batch} images {IMG}x{IMG} u8 per GPU "
                f"(512 images over the ranks), {N_OBJ} objects x {N_JIT} candidates/img -> conf .35 -> custom_nms "
                f".45 -> dilate a=0.1 -> 224 bilinear crop -> ViT-B/16 (+wrapper head)")
    return (f"YOLOv8n-640 head [{batch},{4 + NC},{ANCHORS}] f32 + {batch} images {IMG}x{IMG} u8 per GPU, "
            f"{N_OBJ} objects x {N_JIT} candidates/img -> conf .35 -> custom_nms .45 -> dilate a=0.1 -> "
            f"224 bilinear crop -> ViT-B/16 (+wrapper head)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.monotonic(), [c.strip() for c in line.split(",")]))

    def stop(self, windows=()):
        """`windows`: [(t0, t1)] (time.monotonic) of the timed regions; only samples taken inside them count
        (the sampler is started early because nvidia-smi needs ~1 s to come up on an 8-GPU box)."""
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        rows = [r for t, r in self.rows if not windows or any(a <= t <= b + 0.11 for a, b in windows)]
        sm = sorted(int(r[0]) for r in rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i] == "Active" for r in rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------------ CPU
def make_inputs(batch, seed):
    from yolov8_vit_b200 import synth
    head = synth.make_head(batch, nc=NC, A=ANCHORS, n_obj=N_OBJ, n_jit=N_JIT, img=IMG, seed=seed)
    imgs = synth.make_images(batch, IMG, IMG, seed=seed + 1)
    return head, imgs


def vit_weights():
    from oracle import hotpath as O      # only to GENERATE the seeded synthetic weights
    return O.make_vit_weights(O.ViTConfig(), seed=SEED)


def cpu_pipeline(head, imgs, sd, threads):
    """The reference's CPU path (oracle restatement) on `head`/`imgs`; returns (seconds, crops)."""
    import torch
    from oracle import hotpath as O
    torch.set_num_threads(threads)
    t0 = time.perf_counter()
    crops = 0
    for b in range(head.shape[0]):
        r = O.pipeline_image(head[b], imgs[b].numpy(), sd)
        crops += int(r["valid"].sum())
    return time.perf_counter() - t0, crops


def cpu_baseline(sd, budget_s=15.0):
    """Bounded sample of the same workload: images are added until ~budget_s of CPU work."""
    cores = os.cpu_count() or 1
    head, imgs = make_inputs(64, SEED)
    cpu_pipeline(head[:1], imgs[:1], sd, cores)                      # warm-up
    t, c = cpu_pipeline(head[1:3], imgs[1:3], sd, cores)
    n = int(max(2, min(61, budget_s / max(t / 2, 1e-3))))
    t, c = cpu_pipeline(head[3:3 + n], imgs[3:3 + n], sd, cores)
    return {"value": n / t, "unit": "images/s", "cores": cores, "kind": "port",
            "sample": f"{n} images ({c} crops) of the same synthetic workload, oracle pipeline fp32, "
                      f"torch {cores} threads, {t:.1f} s"}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port: custom_nms is a
    markdown snippet upstream and timm is absent, so the reference itself cannot be imported)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    sd = vit_weights()
    per_step = 6
    head, imgs = make_inputs(per_step * (args.steps + args.warmup), SEED)
    crops = 0
    for w in range(args.warmup):
        cpu_pipeline(head[w * per_step:(w + 1) * per_step], imgs[w * per_step:(w + 1) * per_step], sd, cores)
    t0 = time.perf_counter()
    for s in range(args.warmup, args.warmup + args.steps):
        _, c = cpu_pipeline(head[s * per_step:(s + 1) * per_step], imgs[s * per_step:(s + 1) * per_step], sd, cores)
        crops += c
    dt = time.perf_counter() - t0
    val = per_step * args.steps / dt
    sample = f"{per_step} images/step x {args.steps} steps ({crops} crops), oracle pipeline fp32, torch {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": "images/sec detect-post->NMS->crop->ViT-B/16", "value": val,
        "unit": "images/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(args.batch), "sample": sample},
        "cpu_baseline": {"value": val, "unit": "images/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# ------------------------------------------------------------------------------------------ GPU
def run_ours(args):
    import torch
    import torch.distributed as dist
    from yolov8_vit_b200 import HotPath, HotPathConfig, ViTConfig, ViTModel, _lib
    from yolov8_vit_b200.dist import gather_records

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B = args.batch
    clocks = ClockSampler(local)                   # started early: nvidia-smi needs ~1 s to come up
    clocks.start()
    sd = vit_weights()
    model = ViTModel(ViTConfig(), dev).load_state_dict(sd)
    hp = HotPath(model, HotPathConfig())
    head_h, imgs_h = make_inputs(B, SEED + 100 * rank)             # every rank owns its own shard
    head_h, imgs_h = head_h.pin_memory(), imgs_h.pin_memory()
    head_d, imgs_d = head_h.to(dev), imgs_h.to(dev)
    rec_h = torch.empty(B, hp.cfg.max_det, 8, dtype=torch.float32).pin_memory()
    rec_h2 = (rec_h, torch.empty_like(rec_h).pin_memory())         # two batches are in flight in the e2e loop
    gstream = torch.cuda.Stream(device=dev)

    def step_resident():
        r = hp(head_d, imgs_d)
        if world > 1:      # the single collective of the path: final gather of the per-image records
            gather_records(r.records, world * B)
        return r

    def run_e2e(steps):
        """K steps through the public host-buffer API, the way a serving loop uses it: submit_host(batch k+1)
        (pinned head + images -> sliced upload + detection on the front stream) is issued before
        result(batch k) (ViT + records -> pinned host buffer), so every step's 1.04 GB upload and its record
        download are inside the timed region but overlap the previous step's ViT."""
        pend = hp.submit_host(head_h, imgs_h, rec_h2[0], device=dev, vit_chunk=args.e2e_vit_chunk or None).launch()
        r = None
        for k in range(steps):
            # batch k+1: upload + detection on the front stream, its ViT queued behind batch k's
            nxt = hp.submit_host(head_h, imgs_h, rec_h2[(k + 1) & 1], device=dev,
                                 vit_chunk=args.e2e_vit_chunk or None).launch() if k + 1 < steps else None
            r = pend.result()
            if world > 1:          # the gather runs beside batch k+1's queued ViT, not behind it
                with torch.cuda.stream(gstream):
                    r.records.record_stream(gstream)
                    gather_records(r.records, world * B)
                gstream.synchronize()
            pend = nxt
        return r

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, prof=False):
        barrier()
        if prof:
            _lib.prof_enable(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            r = fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, r

    for _ in range(max(args.warmup, 3)):
        r = step_resident()
    w0 = time.monotonic()
    ms, r = timed(step_resident, args.steps)
    windows = [(w0, time.monotonic())]
    # second pass of the same K steps with every launch event-bracketed on its stream: per-kernel
    # device time for the roofline (kept out of the headline timing: 2 events per launch)
    ms_prof, _ = timed(step_resident, args.steps, prof=True)
    prof = _lib.prof_read()
    _lib.prof_enable(False)
    n_crops = r.n_crops
    run_e2e(2)
    w0 = time.monotonic()
    ms_e2e, _ = timed(lambda: run_e2e(args.steps), 1)
    windows.append((w0, time.monotonic()))
    clk = clocks.stop(windows)                     # samples taken during the two timed regions

    if rank == 0:
        hbm, tf_burst, tf_sus, which = peaks()
        g = prof["gemm_tcgen05"]
        launches = sum(v["launches"] for v in prof.values())
        per_launch_flops = g["work"] / max(g["launches"], 1)
        per_launch_ms = g["ms"] / max(g["launches"], 1)
        achieved = (per_launch_flops / (per_launch_ms * 1e-3)) / 1e12 if per_launch_ms > 0 else 0.0
        stage_ms = {k: round(v["ms"] / args.steps, 4) for k, v in prof.items() if v["launches"]}
        dec = prof["decode_filter"]
        out = {
            "metric": "images/sec detect-post->NMS->crop->ViT-B/16", "value": world * B * args.steps / (ms * 1e-3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic",
            "config": {"workload": workload_name(B), "images_per_gpu": B, "crops_per_gpu": n_crops,
                       "l2": f"inputs larger than L2 (head {head_h.numel() * 4 / 1e6:.0f} MB + images "
                             f"{imgs_h.numel() / 1e6:.0f} MB per step)",
                       "parallelism": f"batch-sharded x{world}, final NCCL all-gather of records"},
            "roofline": {"bound": "tensor", "kernel": "gemm_bf16_kernel (tcgen05)", "achieved": achieved,
                         "peak": tf_sus, "unit": "TFLOP/s", "frac": achieved / tf_sus,
                         "traffic": GEMM_DRAM_BYTES_PER_ROW * n_crops * 197,
                         "traffic_note": "mean dram read+write bytes of the qkv/proj/fc1/fc2 launches of the one-pass "
                                         "ViT (M = crops x 197 rows), ncu --set full (profiles/r01_vit_ncu_v7.md: "
                                         "1265 MB at 1072 crops = the algorithmic bytes)",
                         "peak_source": f"{which} bf16_tflops_sustained (kernel timed inside a long step)",
                         "launches_per_step": g["launches"] / args.steps,
                         "flops_per_launch": per_launch_flops, "ms_per_launch": per_launch_ms,
                         "share_of_step": g["ms"] / ms_prof},
            "stages_ms_per_step": stage_ms,
            "decode_hbm": {"achieved_gbs": dec["work"] / max(dec["ms"], 1e-9) / 1e6, "peak_gbs": hbm,
                           "frac": dec["work"] / max(dec["ms"], 1e-9) / 1e6 / hbm},
            "nms_us_per_img": prof["nms"]["ms"] * 1e3 / (B * args.steps),
            "crops_per_s": world * n_crops * args.steps / (ms * 1e-3),
            "vit_tflops_effective": world * n_crops * FLOP_PER_CROP * args.steps / (ms * 1e-3) / 1e12,
            "e2e": {"value": world * B * args.steps / (ms_e2e * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": head_h.numel() * 4 + imgs_h.numel(),
                    "d2h_bytes_per_step": rec_h.numel() * 4 + 4, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clk,
        }
        if not args.no_cpu_baseline and world == 1:
            out["cpu_baseline"] = cpu_baseline(sd)
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    a.scaling = set_workload(a)
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
