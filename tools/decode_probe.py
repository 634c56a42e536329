"""Where does the small-batch decode lose its time?  Kernel-only device time (library launch profiler, L2 flushed) of
yv_decode_filter over batch size x candidate density x threshold (conf 2.0 keeps nothing: no staging, no gather).
Prints one JSON line per case: us, TB/s, and the (a, b) of us = a + b MB over the batch sizes."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import _lib, ops, synth

dev = "cuda:0"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def kernel_us(fn, reps=7):
    fn()
    _lib.prof_enable(True)
    for _ in range(reps):
        flush.zero_()
        fn()
    torch.cuda.synchronize()
    p = _lib.prof_read()["decode_filter"]
    _lib.prof_enable(False)
    return 1e3 * p["ms"] / max(p["launches"], 1)


for n_obj, conf in ((32, 0.35), (32, 2.0), (4, 0.35)):
    pts = []
    for B in (16, 32, 64, 128, 256):
        head = synth.make_head(B, 80, 8400, n_obj=n_obj, n_jit=24, seed=3).to(dev)
        us = kernel_us(lambda: ops.decode_filter(head, conf))
        mb = head.numel() * 4 / 1e6
        pts.append((mb, us))
        print(json.dumps({"n_obj": n_obj, "conf": conf, "B": B, "us": round(us, 2), "TBs": round(mb / us, 3)}))
        del head
    n = len(pts); sx = sum(p[0] for p in pts); sy = sum(p[1] for p in pts)
    sxx = sum(p[0] ** 2 for p in pts); sxy = sum(p[0] * p[1] for p in pts)
    b = (n * sxy - sx * sy) / (n * sxx - sx * sx); a = (sy - b * sx) / n
    print(json.dumps({"n_obj": n_obj, "conf": conf, "fit_us": {"fixed": round(a, 2), "per_MB": round(b, 4), "marginal_TBs": round(1 / b, 2)}}))
