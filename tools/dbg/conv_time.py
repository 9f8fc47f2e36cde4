import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn as nn
from videoframeinterpolation_b200 import conv as vc
torch.set_grad_enabled(False)
dev = torch.device("cuda")
COUT = int(os.environ.get('COUT', '72')); CIN = int(os.environ.get('CIN', '72'))
conv = nn.Conv2d(CIN, COUT, 3, 1, 1).to(dev)
x = torch.randn(1, CIN, 544, 960, device=dev).contiguous(memory_format=torch.channels_last)
RES = int(os.environ.get('RES', '0'))
r = torch.randn(1, COUT, 544, 960, device=dev).contiguous(memory_format=torch.channels_last) if RES else None
slope = torch.rand(COUT, device=dev)
f = lambda: vc.conv3x3(conv, x, slope, r)
for _ in range(3): f()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(20): f()
b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b) / 20
print(f"res={RES} cin={CIN} cout={COUT} dbg={os.environ.get('VFI_CONV_DBG','0')} ngran={os.environ.get('VFI_CONV_NGRAN','16')}: {ms*1e3:.1f} us  {2*CIN*COUT*9*544*960/ms/1e9:.1f} TFLOP/s", flush=True)
