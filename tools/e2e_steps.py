"""Per-step wall clock of the host-buffer serving loop (bench.py's e2e): is the gap to the resident rate the
pipeline fill of batch 0 or a steady-state cost?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from yolov8_vit_b200 import HotPath, HotPathConfig, ViTConfig, ViTModel

dev = torch.device("cuda:0")
model = ViTModel(ViTConfig(), dev).load_state_dict(bench.vit_weights())
hp = HotPath(model, HotPathConfig())
head_h, imgs_h = bench.make_inputs(256, bench.SEED)
head_h, imgs_h = head_h.pin_memory(), imgs_h.pin_memory()
head_d, imgs_d = head_h.to(dev), imgs_h.to(dev)
rec = [torch.empty(256, 100, 8).pin_memory() for _ in range(2)]
for _ in range(3):
    hp(head_d, imgs_d)
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    hp(head_d, imgs_d)
torch.cuda.synchronize()
print(f"resident: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms/step")
for chunk, slices in ((256, 8), (4096, 8), (4096, 4), (4096, 16)):
    for _ in range(2):
        hp.run_host(head_h, imgs_h, rec[0], device=dev, slices=slices, vit_chunk=chunk)
    torch.cuda.synchronize()
    K = 10
    t = [time.perf_counter()]
    pend = hp.submit_host(head_h, imgs_h, rec[0], device=dev, slices=slices, vit_chunk=chunk).launch()
    for k in range(K):
        nxt = hp.submit_host(head_h, imgs_h, rec[(k + 1) & 1], device=dev, slices=slices, vit_chunk=chunk).launch() if k + 1 < K else None
        pend.result()
        t.append(time.perf_counter())
        pend = nxt
    d = [1e3 * (b - a) for a, b in zip(t, t[1:])]
    print(f"vit_chunk={chunk} slices={slices}: " + " ".join(f"{x:.1f}" for x in d) + f" | mean {sum(d) / K:.2f} steady {sum(d[2:]) / (K - 2):.2f}")
torch.cuda.synchronize(); t0 = time.perf_counter()
a = head_h.to(dev, non_blocking=True); b = imgs_h.to(dev, non_blocking=True); torch.cuda.synchronize()
print(f"plain H2D of head + images: {1e3 * (time.perf_counter() - t0):.2f} ms")
