"""Sustained (power-capped) throughput of the tcgen05 GEMM against cuBLAS on the same operands: each case loops
back to back for SECS seconds (CUDA events around the whole loop) while nvidia-smi samples power and SM clock.
YV_GEMM_DBG=1 (no global stores) / 2 (no epilogue work at all) in the environment show what the epilogue costs in
POWER: under the 1000 W cap a cheaper epilogue buys clock, not idle time."""
import json, os, subprocess, sys, threading, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops

dev = "cuda:0"
SECS = float(os.environ.get("SECS", "2.5"))


class Smi:
    def __init__(self):
        self.rows, self.stop = [], False
    def run(self):
        while not self.stop:
            try:
                o = subprocess.run(["nvidia-smi", "-i", "0", "--query-gpu=power.draw,clocks.sm", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.rows.append((time.monotonic(), float(o[0]), float(o[1])))
            except Exception:
                pass
            time.sleep(0.1)


def sustained(f, flops):
    for _ in range(3):
        f()
    torch.cuda.synchronize()
    smi = Smi()
    th = threading.Thread(target=smi.run); th.start()
    t0 = time.monotonic()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 0
    e0.record()
    while time.monotonic() - t0 < SECS:
        for _ in range(50):
            f()
        n += 50
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    smi.stop = True; th.join()
    ms = e0.elapsed_time(e1)
    rows = [r for r in smi.rows if r[0] - t0 > 0.8]           # after the power limiter has settled
    pw = sorted(r[1] for r in rows); ck = sorted(r[2] for r in rows)
    return {"tflops": round(flops * n / ms / 1e9, 1), "ms_per_launch": round(ms / n, 4), "launches": n, "power_w_median": pw[len(pw) // 2] if pw else None,
            "sm_mhz_median": ck[len(ck) // 2] if ck else None}


out = {"YV_GEMM_DBG": os.environ.get("YV_GEMM_DBG", "0")}
for name, M, N, K in (("plain_8192", 8192, 8192, 8192), ("fc1_shape", 211184, 3072, 768), ("fc2_shape", 211184, 768, 3072)):
    a = (torch.randn(M, K, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    fl = 2.0 * M * N * K
    out[name] = {"ours": sustained(lambda: ops.gemm_bf16(a, w), fl)}
    if not os.environ.get("YV_GEMM_DBG"):
        c = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
        wt = w.t()
        out[name]["cublas"] = sustained(lambda: torch.matmul(a, wt, out=c), fl)
    del a, w
    print(json.dumps({name: out[name]}), flush=True)
print(json.dumps(out))
