"""Which PyTorch (ATen / cuDNN / cuBLAS) kernels one eager 1080p DAT forward still launches, and from which line of this
package: torch.profiler with stacks, kernels joined to their launching op by correlation id."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200.workloads import init_dat_weights
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval().to(memory_format=torch.channels_last)
x0, x1 = torch.rand(1, 3, 1080, 1920, device=dev), torch.rand(1, 3, 1080, 1920, device=dev)
t = torch.full((1, 1, 1, 1), 0.5, device=dev)
with torch.no_grad():
    for _ in range(3):
        net.interpolate(x0, x1, t)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True, record_shapes=True) as prof:
        net.interpolate(x0, x1, t)
        torch.cuda.synchronize()
evs = prof.events()
cpu_by_corr = {}
def pkg_frame(e):
    for fr in (e.stack or []):
        if "videoframeinterpolation_b200" in fr and "_lib.py" not in fr:
            return fr.split("videoframeinterpolation_b200/")[-1]
    return "?"
# kernels: device_type CUDA events; their launching CPU op = the innermost CPU event enclosing the launch (same correlation via e.kernels)
rows = collections.defaultdict(lambda: [0, 0.0])
for e in evs:
    if e.device_type == torch.autograd.DeviceType.CPU and e.kernels:
        # only leaf-most ops carry kernels in this API
        where = pkg_frame(e)
        for k in e.kernels:
            if k.name.startswith("void vfi::") or "vfi::" in k.name:
                continue
            key = (where, e.name, k.name[:60])
            rows[key][0] += 1
            rows[key][1] += k.duration
tot = sum(r[1] for r in rows.values())
print(f"non-vfi kernels of one forward: {sum(r[0] for r in rows.values())} launches, {tot / 1e3:.3f} ms")
for (where, op, kn), (n, us) in sorted(rows.items(), key=lambda kv: -kv[1][1])[:70]:
    print(f"{us:8.1f} us {n:3d}x  {op:32s} {kn:60s} {where}")

print("largest elementwise glue launches (op, input shapes, us):")
glue = []
for e in evs:
    if e.device_type == torch.autograd.DeviceType.CPU and e.kernels and e.name in ("aten::copy_", "aten::add_", "aten::cat", "aten::fill_", "aten::add", "aten::gelu"):
        glue.append((sum(k.duration for k in e.kernels), e.name, str(e.input_shapes)[:110]))
for us, nm, shp in sorted(glue, reverse=True)[:28]:
    print(f"{us:8.1f} us  {nm:12s} {shp}")
