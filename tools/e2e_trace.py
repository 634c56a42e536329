"""Where does run_host's time go?  GPU events + CPU timestamps around one pipelined host-buffer call."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from yolov8_vit_b200 import HotPath, HotPathConfig, ViTConfig, ViTModel

dev = torch.device("cuda:0")
sd = bench.vit_weights()
model = ViTModel(ViTConfig(), dev).load_state_dict(sd)
hp = HotPath(model, HotPathConfig())
head_h, imgs_h = bench.make_inputs(256, bench.SEED)
head_h, imgs_h = head_h.pin_memory(), imgs_h.pin_memory()
rec_h = torch.empty(256, 100, 8).pin_memory()
for _ in range(3):
    hp.run_host(head_h, imgs_h, rec_h, device=dev)
marks = []
orig = model.forward_patches
def fp(*a, **k):
    e = torch.cuda.Event(enable_timing=True); e.record(); marks.append(("vit_launch", time.perf_counter(), e, a[1]))
    return orig(*a, **k)
model.forward_patches = fp
for slices, ramp in ((16, (64, 128)), (8, ()), (4, ()), (16, (32, 64, 128))):
    marks.clear()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    hp.run_host(head_h, imgs_h, rec_h, device=dev, slices=slices, ramp=ramp)
    e1.record(); t1 = time.perf_counter(); torch.cuda.synchronize()
    print(f"slices={slices} ramp={ramp}: total {e0.elapsed_time(e1):.2f} ms gpu, {1e3 * (t1 - t0):.2f} ms cpu wall")
    print("   chunk launches (crops, cpu ms since start, gpu ms since start): " +
          " ".join(f"({n},{1e3 * (t - t0):.1f},{e0.elapsed_time(e):.1f})" for _, t, e, n in marks))
# upload alone
torch.cuda.synchronize(); t0 = time.perf_counter()
a = head_h.to(dev, non_blocking=True); b = imgs_h.to(dev, non_blocking=True); torch.cuda.synchronize()
print(f"plain H2D of head + images: {1e3 * (time.perf_counter() - t0):.2f} ms")
