"""Logit error of the ViT against the fp32 oracle for the three stream modes (child process per mode: the switches are
read once).  python tools/vit_err_modes.py [depth]"""
import json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CODE = r"""
import sys, json, torch
sys.path.insert(0, %r)
from oracle import hotpath as O
from yolov8_vit_b200 import ViTConfig, ViTModel
depth = int(sys.argv[1])
out = []
for seed in (12, 13, 14):
    cfg = O.ViTConfig(depth=depth)
    sd = O.make_vit_weights(cfg, seed=seed)
    m = ViTModel(ViTConfig(depth=depth), 'cuda:0').load_state_dict(sd)
    x = torch.randn(37, 3, 224, 224, generator=torch.Generator().manual_seed(seed - 8)).clamp_(-1, 1)
    ref = O.vit_forward(sd, x, cfg)
    got = m(x.cuda()).cpu()
    out.append(round(float((got - ref).abs().max() / ref.abs().max()), 5))
    out.append(round(float(((got - ref).pow(2).mean().sqrt()) / ref.abs().max()), 5))
print(json.dumps(out))
""" % ROOT

if __name__ == "__main__":
    depth = sys.argv[1] if len(sys.argv) > 1 else "3"
    for name, env in (("fold", {}), ("ln_kernels", {"YV_VIT_LNFOLD": "0"}), ("f32_stream", {"YV_VIT_STREAM": "f32"})):
        r = subprocess.run([sys.executable, "-c", CODE, depth], env=dict(os.environ, **env), capture_output=True, text=True)
        print(json.dumps({"mode": name, "depth": int(depth), "max_rel,rms_rel per seed": r.stdout.strip() or r.stderr[-400:]}))
