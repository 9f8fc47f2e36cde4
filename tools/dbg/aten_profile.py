"""Which ATen ops are left in one eager 1080p forward (torch profiler, grouped by input shape)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200.workloads import init_dat_weights
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda")
torch.manual_seed(0)
if len(sys.argv) > 1 and sys.argv[1] == "dcndat":
    net = init_dat_weights(v.DCNDATv1(v.dcndat_default_args())).to(dev).eval().to(memory_format=torch.channels_last)
else:
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval().to(memory_format=torch.channels_last)
x0, x1 = (torch.rand(1, 3, 1088, 1920, device=dev).contiguous(memory_format=torch.channels_last) for _ in range(2))
t = torch.full((1, 1, 1, 1), 0.5, device=dev)
for _ in range(3):
    net.inference(x0, x1, t)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True, with_stack=False) as prof:
    net.inference(x0, x1, t)
    torch.cuda.synchronize()
print("==== kernels ====")
kern = {}
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        k = kern.setdefault(e.name[:90], [0, 0.0]); k[0] += 1; k[1] += e.device_time
for name, (n, us) in sorted(kern.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"{us:10.1f} us  n={n:4d}  {name}")
print(prof.key_averages(group_by_input_shape=True).table(sort_by="cuda_time_total", row_limit=45, max_name_column_width=40, max_shapes_column_width=70))
print("==== ATen ops that launch kernels, by shape ====")
rows = [e for e in prof.key_averages(group_by_input_shape=True) if e.key.startswith("aten::") and e.device_time_total > 0
        and not any(k in e.key for k in ("conv", "addmm", "gelu", "cudnn"))]
for e in sorted(rows, key=lambda e: -e.device_time_total)[:60]:
    print(f"{e.key:28s} n={e.count:3d} cuda_us={e.device_time_total:9.1f} shapes={str(e.input_shapes)[:110]}")
print("==== elementwise_kernel launches by op ====")
for e in sorted([e for e in prof.key_averages(group_by_input_shape=True) if e.device_time_total > 0 and e.key.startswith("aten::")
                 and any(k in e.key for k in ("copy_", "fill_", "add", "mul", "sub", "cat", "contiguous", "tanh", "sigmoid", "clone"))],
                key=lambda e: -e.device_time_total)[:25]:
    print(f"{e.key:24s} n={e.count:3d} cuda_us={e.device_time_total:9.1f} shapes={str(e.input_shapes)[:120]}")
