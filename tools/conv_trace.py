"""One traced launch of the tensor-core conv (variants/libvfi_trace.so, built with -DVFI_CONV_TRACE): role timelines of CTA 0."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn as nn
from videoframeinterpolation_b200 import conv as C
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True
cin, cout, H, W = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
res = len(sys.argv) > 5 and sys.argv[5] == "res"
m = nn.Conv2d(cin, cout, 3, 1, 1).to(dev)
x = torch.randn(1, H, W, cin, device=dev).permute(0, 3, 1, 2)
r = torch.randn(1, H, W, cout, device=dev).permute(0, 3, 1, 2) if res else None
with torch.no_grad():
    for i in range(3):
        os.environ["VFI_TRACE_QUIET"] = "1"
        y = C.conv3x3(m, x, None, r)
    torch.cuda.synchronize()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    ts = []
    for i in range(3):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); y = C.conv3x3(m, x, None, r); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1000)
    print("event-timed launches (us):", [round(t, 1) for t in ts])
