"""On-device PSNR / SSIM validators with the reference's signatures.

    calculate_psnr(img1, img2)                                   evaluate.py:82-84
    calculate_ssim(img1, img2, window_size=11, window=None,
                   size_average=True, full=False, val_range=None)  evaluate.py:29-79

Both run in one pass of vfi_frame_metrics_f32 (csrc/metrics.cu): the 11x11x11 Gaussian window of the reference is
separable, so the five conv3d calls over replicate-padded copies become a channel mix + two 11-tap passes in shared
memory.  `frame_metrics` returns per-sample values for validation loops that want both numbers from one launch.
No CPU fallback: CPU tensors raise, like the reference's `.cuda()` window (evaluate.py:24).
"""
from __future__ import annotations

import torch

from . import _lib

__all__ = ["calculate_psnr", "calculate_ssim", "frame_metrics"]


def _val_range(img1: torch.Tensor) -> float:
    """evaluate.py:30-40: dynamic range guessed from img1."""
    max_val = 255 if float(img1.max()) > 128 else 1
    min_val = -1 if float(img1.min()) < -0.5 else 0
    return float(max_val - min_val)


def _sums(img1: torch.Tensor, img2: torch.Tensor, val_range: float) -> torch.Tensor:
    """(B, 3) fp64: per-sample sum of squared differences, sum of the SSIM map, sum of the cs map."""
    dev = _lib.require_cuda(img1, img2)
    assert img1.dim() == 4 and img1.shape == img2.shape, "img1 / img2 must be (B, C, H, W) of the same shape"
    a, b = _lib.f32c(img1), _lib.f32c(img2)
    B, C, H, W = a.shape
    out = torch.empty((B, 3), dtype=torch.float64, device=dev)
    nbytes = int(_lib.lib().vfi_frame_metrics_workspace_bytes(B, C, H, W))
    ws = torch.empty((max(nbytes // 8, 1),), dtype=torch.float64, device=dev)
    _lib.call("vfi_frame_metrics_f32", dev, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, C, H, W, float(val_range),
              ws.data_ptr(), ws.numel() * 8, _lib.stream_ptr(dev))
    return out


def frame_metrics(pred: torch.Tensor, gt: torch.Tensor, val_range: float = 1.0):
    """Per-sample (psnr, ssim) of (B, C, H, W) frames, fp64 tensors of shape (B,), from one launch."""
    s = _sums(pred, gt, val_range)
    n = max(pred.shape[1] * pred.shape[2] * pred.shape[3], 1)
    return -10.0 * torch.log10(s[:, 0] / n), s[:, 1] / n


def calculate_psnr(img1: torch.Tensor, img2: torch.Tensor) -> torch.Tensor:
    """-10 log10(mean((img1 - img2)^2)) over the whole tensor; 3-D (C, H, W) inputs are accepted as the reference's
    per-sample calls pass them (evaluate.py:108)."""
    if img1.dim() == 3:
        img1, img2 = img1[None], img2[None]
    s = _sums(img1, img2, 1.0)
    return (-10.0 * torch.log10(s[:, 0].sum() / img1.numel())).to(torch.float32)


def calculate_ssim(img1, img2, window_size=11, window=None, size_average=True, full=False, val_range=None):
    assert window_size == 11 and window is None, "the kernel implements the reference's default 11-tap sigma-1.5 window"
    assert size_average, "size_average=False (the reference reduces it to a (B, W) tensor) is not implemented"
    L = _val_range(img1) if val_range is None else float(val_range)
    s = _sums(img1, img2, L)
    n = img1.numel()
    ret = (s[:, 1].sum() / n).to(torch.float32)
    if full:
        return ret, (s[:, 2].sum() / n).to(torch.float32)
    return ret
