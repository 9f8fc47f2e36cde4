"""One training step (forward + backward, fp32, grad wrt every parameter) of the DAT model through the CUDA drop-ins:
peak device memory and time per frame size, fused-attention training route against the kv-materialising one."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import videoframeinterpolation_b200 as v
from videoframeinterpolation_b200.workloads import init_dat_weights
dev = torch.device("cuda:0")
torch.backends.cudnn.allow_tf32 = True
torch.backends.cuda.matmul.allow_tf32 = True
net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).train()
cases = [((544, 960), "fused"), ((544, 960), "materialize"), ((1088, 1920), "fused")]
if len(sys.argv) > 1 and sys.argv[1] == "all":
    cases.append(((1088, 1920), "materialize"))
for (H, W), mode in cases:
    os.environ["VFI_DATTN_TRAIN"] = mode
    g = torch.Generator().manual_seed(0)
    x0, x1, gt = (torch.rand(1, 3, H, W, generator=g).to(dev) for _ in range(3))
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    try:
        for it in range(2):
            net.zero_grad(set_to_none=True)
            torch.cuda.synchronize(); torch.cuda.reset_peak_memory_stats()
            t0 = time.perf_counter()
            loss = (net(x0, x1, t) - gt).abs().mean()
            loss.backward()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        gn = sum(float(p.grad.square().sum()) for p in net.parameters() if p.grad is not None) ** 0.5
        print(f"{W}x{H} {mode:12s}: step {dt * 1e3:8.1f} ms, peak {torch.cuda.max_memory_allocated() / 2**30:6.1f} GiB, loss {float(loss):.5f}, |grad| {gn:.4e}", flush=True)
    except torch.OutOfMemoryError as e:
        print(f"{W}x{H} {mode:12s}: out of memory ({str(e)[:120]})", flush=True)
    del x0, x1, gt
    torch.cuda.empty_cache()
