"""Can attention hide behind the GEMMs?  The step is power-capped: the GEMMs run at whatever clock 1000 W leave and
lose little when they get fewer SMs, while attention (exp-bound, far below the cap) runs at the GEMMs' depressed clock.
This probe runs one ViT block's GEMM work of HALF a bench batch (qkv, proj, fc1, fc2 at M = 536 x 197) on one stream
and the attention of the other half on a second stream, with disjoint SM budgets (yv_tune_sm_budget), back to back
for SECS seconds, against the same work serialised on one stream with all SMs."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops
from yolov8_vit_b200._lib import load, check

dev = "cuda:0"
SECS = float(os.environ.get("SECS", "2.5"))
n = 536
M, D, H = n * 197, 768, 3072
g = torch.Generator(device=dev).manual_seed(1)
rb = lambda *s: (torch.randn(*s, device=dev, generator=g) * 0.05).to(torch.bfloat16)
y = (torch.randn(M, D, device=dev, generator=g) * 0.5).to(torch.bfloat16)
x16 = torch.randn(M, D, device=dev, generator=g).to(torch.float16)
hbuf = (torch.randn(M, H, device=dev, generator=g) * 0.5).to(torch.bfloat16)
wq, wp, w1, w2 = rb(3 * D, D), rb(D, D), rb(H, D), rb(D, H)
bq, bp, b1, b2 = (torch.randn(k, device=dev, generator=g) * 0.1 for k in (3 * D, D, H, D))
ones, zeros = torch.ones(D, device=dev), torch.zeros(D, device=dev)
qkv = (torch.randn(n, 197, 3 * D, device=dev, generator=g) * 0.7).to(torch.bfloat16)


def gemms():
    ops.gemm_bf16(y, wq, bq)                    # qkv
    ops.gemm_bf16(y, wp, bp, residual=x16, out=x16)     # proj (fp16 stream in place)
    ops.gemm_bf16(y, w1, b1, gelu=True)         # fc1
    ops.gemm_bf16(hbuf, w2, b2, residual=x16, out=x16)  # fc2


def attention():
    ops.attention_bf16(qkv, 12)


def run(mode, gemm_sms=0, attn_sms=0):
    check(load().yv_tune_sm_budget(gemm_sms, attn_sms), "tune")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(3):
        gemms(); attention()
    torch.cuda.synchronize()
    t0 = time.monotonic()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    it = 0
    e0.record()
    while time.monotonic() - t0 < SECS:
        for _ in range(20):
            if mode == "serial":
                gemms(); attention()
            else:
                with torch.cuda.stream(s1):
                    gemms()
                with torch.cuda.stream(s2):
                    attention()
        it += 20
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    check(load().yv_tune_sm_budget(0, 0), "tune")
    return {"mode": mode, "gemm_sms": gemm_sms, "attn_sms": attn_sms, "ms_per_iteration": round(e0.elapsed_time(e1) / it, 4), "iterations": it}


print(json.dumps(run("serial")))
print(json.dumps(run("gemms_only_probe", 0, 0) if False else {"note": "iteration = qkv + proj + fc1 + fc2 at M = 105 592 rows, + attention of 536 crops"}))
for gs in (132, 124, 116, 108):
    print(json.dumps(run("overlap", gs, 148 - gs)), flush=True)
print(json.dumps(run("serial")))
