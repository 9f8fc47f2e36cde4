"""Where the time of one DCNDATv1 1080p forward goes: vfi ops by OpTimer, the rest by torch.profiler."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200 import _lib
from videoframeinterpolation_b200.workloads import init_dat_weights
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True; torch.backends.cuda.matmul.allow_tf32 = True; torch.backends.cudnn.benchmark = True
net = init_dat_weights(v.DCNDATv1(v.dcndat_default_args())).to(dev).eval().to(memory_format=torch.channels_last)
g = torch.Generator().manual_seed(0)
x0 = torch.rand(1, 3, 1088, 1920, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
x1 = torch.rand(1, 3, 1088, 1920, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
t = torch.full((1, 1, 1, 1), 0.5, device=dev)
with torch.no_grad():
    for _ in range(3):
        net.inference(x0, x1, t)
    torch.cuda.synchronize()
    tm = _lib.OpTimer(); tm.detailed = True
    with tm:
        for _ in range(3):
            net.inference(x0, x1, t)
    torch.cuda.synchronize()
    rows = sorted(((ms / 3, c // 3, k) for k, (c, ms) in tm.summary().items()), reverse=True)
    print("vfi ops (ms per forward, calls):")
    for ms, c, k in rows[:22]:
        print(f"{ms:8.3f} {c:4d}  {k}")
    print("vfi total", round(sum(r[0] for r in rows), 3))
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        net.inference(x0, x1, t); torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CPU and e.kernels:
        for k in e.kernels:
            if "vfi::" in k.name: continue
            agg[(e.name, k.name[:50])][0] += 1; agg[(e.name, k.name[:50])][1] += k.duration
print("non-vfi kernels:", sum(a[0] for a in agg.values()), "launches", round(sum(a[1] for a in agg.values()) / 1e3, 3), "ms")
for (op, kn), (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:14]:
    print(f"{us:8.1f} us {n:3d}x {op:30s} {kn}")
