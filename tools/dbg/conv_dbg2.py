import os, sys
R = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import torch, torch.nn as nn
import oracle
from videoframeinterpolation_b200 import conv as vc
torch.set_grad_enabled(False)
def case(B, Cin, Cout, H, W, prelu, res, slice_of=None, scale=0.1, tag=""):
    g = torch.Generator().manual_seed(0)
    conv = nn.Conv2d(Cin, Cout, 3, 1, 1)
    conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * scale)
    conv.bias.copy_(torch.randn(Cout, generator=g))
    conv = conv.cuda()
    ctot = slice_of or Cin
    xfull = torch.randn(B, ctot, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    x = xfull[:, ctot - Cin:]
    slope = torch.rand(Cout, generator=g).cuda() if prelu else None
    r = torch.randn(B, Cout, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last) if res else None
    for rep in range(int(os.environ.get('REPS','3'))):
        y = vc.conv3x3(conv, x, slope, r, residual_after_act=(res == 2))
        torch.cuda.synchronize()
        ref = oracle.conv3x3_tf32(x.cpu(), conv.weight.cpu(), conv.bias.cpu(), None if slope is None else slope.cpu(),
                                  None if r is None else r.cpu(), res == 2)
        d = (y.double().cpu() - ref).abs()
        idx = torch.nonzero(d > 1e-4 * ref.abs().max())
        print(tag, f"B{B} {Cin}->{Cout} {H}x{W} prelu={prelu} res={res} rep{rep}: max {float(d.max()):.3e} nbad {len(idx)}",
              (idx[:3].tolist(), idx[-3:].tolist()) if len(idx) else "", flush=True)
import subprocess
which = os.environ.get("CASES", "abcdef")
if "a" in which: case(2, 72, 72, 45, 37, False, 0, tag="a")
if "b" in which: case(2, 72, 72, 45, 37, True, 0, tag="b")
if "c" in which: case(2, 72, 72, 45, 37, False, 1, tag="c")
if "d" in which: case(2, 72, 72, 45, 37, True, 1, tag="d")
if "e" in which: case(1, 72, 72, 33, 20, True, 2, tag="e")
if "f" in which: case(1, 36, 36, 32, 24, True, 0, slice_of=72, tag="f")
if "z" in which:
    g = torch.Generator().manual_seed(0)
    conv = nn.Conv2d(72, 72, 3, 1, 1).cuda()
    x = torch.randn(1, 72, 32, 24, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    r = torch.randn(1, 72, 32, 24, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    r2 = torch.randn(1, 72, 32, 24, device="cuda").contiguous(memory_format=torch.channels_last)
    y0 = vc.conv3x3(conv, x)
    for name, rr in (("cpu-made r", r), ("gpu-made r", r2)):
        y1 = vc.conv3x3(conv, x, None, rr)
        torch.cuda.synchronize()
        print(name, "strides", rr.stride(), "ptr%16", rr.data_ptr() % 16, "y1-y0 vs r:", float((y1 - y0 - rr).abs().max()),
              "y1-y0 absmax", float((y1 - y0).abs().max()), flush=True)
