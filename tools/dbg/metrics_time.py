"""1080p timing of the validators: vfi_frame_metrics_f32 (one launch: PSNR + SSIM) next to the reference's composition
(evaluate.py:29-84 restated in oracle/torch_oracle.py) run with PyTorch CUDA ops on the same GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import oracle
import videoframeinterpolation_b200 as v

dev = torch.device("cuda")
a = torch.rand(1, 3, 1080, 1920, device=dev)
b = (a + 0.03 * torch.randn_like(a)).clamp(0, 1)


def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    return sorted(ts)[len(ts) // 2]


ms = timeit(lambda: v.frame_metrics(a, b))
ms_ref = timeit(lambda: (oracle.calculate_ssim(a, b, val_range=1.0), oracle.calculate_psnr(a, b)), 5)
print({"frame": "1080x1920x3", "vfi_frame_metrics_ms": ms, "pytorch_cuda_composition_ms": ms_ref, "ratio": ms_ref / ms,
       "algorithmic_GBps": 2 * a.numel() * 4 / ms / 1e6})
