#!/usr/bin/env python
"""Where the tensor-core conv time of one 1080p DAT forward goes: ms and TFLOP/s per (resolution, Cin -> Cout, residual
mode, upscale), from CUDA events around every vfi_conv3x3_f32 call of an eager pass."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200 import _lib
from videoframeinterpolation_b200.workloads import init_dat_weights

dev = torch.device("cuda", 0)
torch.backends.cudnn.allow_tf32 = True
torch.backends.cudnn.benchmark = True
torch.manual_seed(0)
net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval().to(memory_format=torch.channels_last)
g = torch.Generator().manual_seed(0)
x0 = torch.rand(1, 3, 1088, 1920, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
x1 = torch.rand(1, 3, 1088, 1920, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
t = torch.full((1, 1, 1, 1), 0.5, device=dev)
with torch.no_grad():
    for _ in range(3):
        net.inference(x0, x1, t)
    torch.cuda.synchronize()
    N = 5
    tm = _lib.OpTimer()
    tm.detailed = True
    with tm:
        for _ in range(N):
            net.inference(x0, x1, t)
    torch.cuda.synchronize()
rows = []
for name, (calls, ms) in tm.summary().items():
    if not name.startswith("conv3x3 "):
        continue
    _, hw, cc, res, up, ps = name.split()[:6]
    H, W = (int(a) for a in hw.split("x"))
    cin, cout = (int(a) for a in cc.split("->"))
    parts = name.split()
    fl = 2.0 * 9 * cin * cout * H * W * (int(parts[6][1:]) if len(parts) > 6 else 1)      # " B2": a batch of two images
    rows.append((ms / N, calls // N, name, fl * (calls // N) / (ms / N) / 1e9, ms / calls * 1e3))
rows.sort(reverse=True)
tot = sum(r[0] for r in rows)
print(f"{'ms/step':>8} {'calls':>5} {'us/call':>8} {'TFLOP/s':>8}  shape")
for ms, calls, name, tf, us in rows:
    print(f"{ms:8.3f} {calls:5d} {us:8.1f} {tf:8.1f}  {name}")
print(f"total {tot:.3f} ms over {sum(r[1] for r in rows)} launches")
other = {k: (c // N, m / N) for k, (c, m) in tm.summary().items() if not k.startswith("conv3x3 ")}
print("other vfi ops (calls, ms/step):", {k: (c, round(m, 3)) for k, (c, m) in other.items()})
