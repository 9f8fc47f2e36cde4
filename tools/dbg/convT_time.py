import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, torch.nn as nn
from videoframeinterpolation_b200 import conv as vc
torch.set_grad_enabled(False)
dev = torch.device("cuda")
def tm(f, n=20):
    for _ in range(3): f()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): f()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for (cin, cout, H, W, ocl) in [(72, 72, 272, 480, True), (72, 2, 272, 480, False), (72, 72, 136, 240, True), (72, 2, 136, 240, False), (76, 76, 68, 120, True)]:
    ct = nn.ConvTranspose2d(cin, cout, 4, 2, 1).to(dev).to(memory_format=torch.channels_last)
    x = torch.randn(1, cin, H, W, device=dev).contiguous(memory_format=torch.channels_last)
    torch.backends.cudnn.benchmark = True
    t_ref = tm(lambda: ct(x))
    t_tc = tm(lambda: vc.conv_transpose_up2(ct, x, out_channels_last=ocl))
    print(f"convT {cin}->{cout} @{H}x{W}: cudnn {t_ref:.1f} us, tc {t_tc:.1f} us", flush=True)
