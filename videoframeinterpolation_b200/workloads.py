"""Synthetic workloads of BASELINE.json (shapes, seeds, algorithmic bytes) shared by bench.py, smoke() and tests.

Inputs follow SURVEY.md 8(d): frames U[0,1), features N(0,1), flows N(0, sigma_l^2) with sigma_l = 8 / 2^l px at
pyramid level l plus 2 % of pixels leaving the frame, metric N(0,1).  ``smooth`` flows (default for the bench)
are spatially correlated like optical flow (coarse Gaussian field upsampled bicubically, same sigma) with a
0.25 px i.i.d. jitter; its correlation length is 32 px, i.e. a strain of ~0.35 px/px, rougher than camera or
object motion; ``gentle`` is the same field with a 256 px correlation length (strain ~0.05 px/px, closer to
real optical flow); ``iid`` flows are independent per pixel (SURVEY.md 8(d)'s literal wording, the adversarial
case, used by tests and reported by the micro-benchmark as a separate line).
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


def seed_for(cfg: int, idx: int) -> int:
    return 20260000 + 100 * cfg + idx


def make_flow(B: int, H: int, W: int, sigma: float, gen: torch.Generator, kind: str = "smooth",
              outlier_frac: float = 0.02, device="cpu") -> torch.Tensor:
    if kind == "iid":
        f = torch.randn(B, 2, H, W, generator=gen) * sigma
    else:
        cell = 256 if kind == "gentle" else 32          # correlation length of the field in pixels
        ch, cw = max(2, H // cell + 2), max(2, W // cell + 2)
        coarse = torch.randn(B, 2, ch, cw, generator=gen)
        f = F.interpolate(coarse, size=(H, W), mode="bicubic", align_corners=True)
        f = f / f.std().clamp_min(1e-6) * sigma + torch.randn(B, 2, H, W, generator=gen) * 0.25
    if outlier_frac > 0:
        m = torch.rand(B, 1, H, W, generator=gen) < outlier_frac
        big = (torch.rand(B, 2, H, W, generator=gen) * 2 - 1) * W
        f = torch.where(m, big, f)
    return f.contiguous().to(device)


def init_dat_weights(model: torch.nn.Module, seed: int = 0) -> torch.nn.Module:
    """Constructor init under a fixed seed; the heads the reference zero-initialises (modules/dcnv2.py:35,
    modules/deformable_attn.py:64) are re-randomised N(0, 0.02^2) so that residual offsets are non-zero."""
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in model.named_parameters():
            if "conv_res_offset" in name or "conv_offset_mask" in name:
                if p.ndim > 1 or name.endswith("bias"):
                    p.copy_(torch.randn(p.shape, generator=g) * 0.02)
    return model


# ---- algorithmic bytes per launch (SURVEY.md 8d / BASELINE.md 5), fp32 ------------------------------
def bytes_bwarp(B, C, H, W):
    return 4 * (2 * C + 2) * B * H * W


def bytes_splat(B, C, H, W, mode="soft"):
    return 4 * (2 * C + (3 if mode.split("-")[0] in ("soft", "linear") else 2)) * B * H * W


def bytes_dattn(B, C, H, W, nG, nS):
    return 4 * (4 * C + 4 * nG * nS + 4) * B * H * W


def bytes_dcn(B, Cin, Cout, H, W, OG, K=9):
    return 4 * (Cin + 3 * OG * K + Cout) * B * H * W


def bytes_dsample(B, C, H, W, nG, nS):
    return 4 * (C + 2 * nG * nS + C * nS) * B * H * W
