"""Time every stride-2 / non-tcgen05 conv module of the DAT forward in place (eager, CUDA events around the module)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200.workloads import init_dat_weights
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval().to(memory_format=torch.channels_last)
x0, x1 = torch.rand(1, 3, 1088, 1920, device=dev), torch.rand(1, 3, 1088, 1920, device=dev)
t = torch.full((1, 1, 1, 1), 0.5, device=dev)
rows = {}
def pre(m, inp):
    m._e0 = torch.cuda.Event(enable_timing=True); m._e0.record()
def post(m, inp, out):
    e1 = torch.cuda.Event(enable_timing=True); e1.record()
    rows.setdefault(m._nm, []).append((m._e0, e1, tuple(inp[0].shape), inp[0].stride(), m.weight.stride()))
for nm, m in net.named_modules():
    if isinstance(m, (nn.Conv2d, nn.ConvTranspose2d)) and (m.stride != (1, 1) or m.kernel_size != (3, 3)):
        m._nm = f"{nm} {type(m).__name__} k{m.kernel_size} s{m.stride} {m.in_channels}->{m.out_channels}"
        m.register_forward_pre_hook(pre); m.register_forward_hook(post)
import torch.nn.functional as F
_orig = F.conv2d
frows = []
def timed_conv2d(x, w, b=None, stride=1, padding=0, dilation=1, groups=1):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); y = _orig(x, w, b, stride, padding, dilation, groups); e1.record()
    frows.append((e0, e1, tuple(x.shape), x.stride(), tuple(w.shape), w.stride(), stride, b is not None, y.stride()))
    return y
F.conv2d = timed_conv2d
with torch.no_grad():
    for i in range(4):
        if i == 3: rows.clear(); frows.clear()
        net.interpolate(x0, x1, t)
torch.cuda.synchronize()
for nm, lst in rows.items():
    for e0, e1, shp, st, wst in lst:
        print(f"{e0.elapsed_time(e1)*1000:8.1f} us  {nm}  in {shp} strides {st} w strides {wst}")
print("F.conv2d calls:")
for e0, e1, shp, st, wshp, wst, stride, hb, yst in frows:
    print(f"{e0.elapsed_time(e1)*1000:8.1f} us  in {shp} strides {st} w {wshp} strides {wst} stride {stride} bias {hb} out strides {yst}")
