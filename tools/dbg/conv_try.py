"""Bring-up check of vfi_conv3x3_f32 against TF32-emulated fp64 references (tools/dbg, not a test)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn as nn, torch.nn.functional as F
from videoframeinterpolation_b200 import conv as vc

def trunc(x): return (x.view(torch.int32) & ~0x1FFF).view(torch.float32)
def rna(x): return ((x.view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)

dev = torch.device("cuda")
torch.set_grad_enabled(False)
torch.manual_seed(0)
def run(B, Cin, Cout, H, W, prelu=False, res=0, ocl=True):
    conv = nn.Conv2d(Cin, Cout, 3, 1, 1).to(dev)
    x = torch.randn(B, Cin, H, W, device=dev).contiguous(memory_format=torch.channels_last)
    slope = torch.rand(Cout, device=dev) if prelu else None
    r = torch.randn(B, Cout, H, W, device=dev).contiguous(memory_format=torch.channels_last) if res else None
    assert vc.eligible(conv, x), "not eligible"
    y = vc.conv3x3(conv, x, slope, r, residual_after_act=(res == 2), out_channels_last=ocl)
    torch.cuda.synchronize()
    outs = {}
    for name, fx in (("trunc", trunc), ("rna", rna), ("fp32", lambda t: t)):
        wq = rna(conv.weight.detach()) if name != "fp32" else conv.weight.detach()
        ref = F.conv2d(fx(x).double(), wq.double(), conv.bias.double(), padding=1)
        if res == 1: ref = r.double() + ref
        if prelu: ref = torch.where(ref > 0, ref, slope.double().view(1, -1, 1, 1) * ref)
        if res == 2: ref = r.double() + ref
        outs[name] = float((y.double() - ref).abs().max() / ref.abs().max())
    print(f"B{B} Cin{Cin} Cout{Cout} {H}x{W} prelu={prelu} res={res} ocl={ocl}: " + " ".join(f"{k}={v:.2e}" for k, v in outs.items()), flush=True)

run(1, 72, 72, 16, 8)
run(1, 72, 72, 32, 24, prelu=True, res=1)
run(2, 72, 72, 45, 37, prelu=True, res=2)
run(1, 72, 72, 35, 29, ocl=False)
run(1, 72, 512, 33, 40)
run(1, 72, 288, 32, 32)
run(1, 72, 3, 32, 32, ocl=False)
run(1, 32, 40, 32, 32)
# timing at the DAT 1/2-res shape
conv = nn.Conv2d(72, 72, 3, 1, 1).to(dev)
x = torch.randn(1, 72, 544, 960, device=dev).contiguous(memory_format=torch.channels_last)
for f, name in ((lambda: vc.conv3x3(conv, x), "vfi tc"), (lambda: conv(x), "cudnn tf32")):
    torch.backends.cudnn.benchmark = True
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): f()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    print(f"{name}: {ms*1e3:.1f} us  {2*72*72*9*544*960/ms/1e9:.1f} TFLOP/s", flush=True)
