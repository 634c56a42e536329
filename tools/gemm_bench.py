"""Times the ViT's GEMM shapes one by one (CUDA events, L2 flushed between reps) -> TFLOP/s per shape."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from yolov8_vit_b200 import ops

dev = "cuda:0"
M = int(os.environ.get('CROPS', '1072')) * 197
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
res = {}
for name, N, K, kw in [("qkv", 2304, 768, dict(bias=True)), ("proj", 768, 768, dict(bias=True, res=True)),
                       ("fc1", 3072, 768, dict(bias=True, gelu=True)), ("fc2", 768, 3072, dict(bias=True, res=True)),
                       ("plain_8192", 8192, 8192, dict())]:
    m = 8192 if name == "plain_8192" else M
    a = (torch.randn(m, K, device=dev) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
    bias = torch.randn(N, device=dev) if kw.get("bias") else None
    r = (torch.randn(m, N, device=dev)).to(torch.float16) if kw.get("res") else None     # fp16 stream, updated in place
    f = (lambda: ops.gemm_bf16(a, w, bias=bias, residual=r, out=r)) if kw.get("res") else \
        (lambda: ops.gemm_bf16(a, w, bias=bias, gelu=kw.get("gelu", False)))
    NCU = bool(os.environ.get('NCU'))
    for _ in range(1 if NCU else 3):
        f()
    ts = []
    for _ in range(1 if NCU else 10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); f(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    med = ts[len(ts) // 2]
    res[name] = {"ms": med, "tflops": 2.0 * m * N * K / med / 1e9}
    ref = torch.matmul(a, w.t())
    print(name, res[name], "ref_ms(torch.matmul, warm)", end=" ")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush.zero_(); e0.record(); torch.matmul(a, w.t()); e1.record(); torch.cuda.synchronize()
    print(round(e0.elapsed_time(e1), 4))
print(json.dumps(res))
