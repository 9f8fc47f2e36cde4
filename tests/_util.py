"""Parity metrics: rel = max|a-b| / max|b|; elementwise = max |a-b| / (|b| + 1e-3 max|b|)."""
import torch


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    assert a.shape == b.shape, (a.shape, b.shape)
    denom = b.abs().max().clamp_min(1e-30)
    return float((a - b).abs().max() / denom)


def elem_err(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    m = b.abs().max().clamp_min(1e-30)
    return float(((a - b).abs() / (b.abs() + 1e-3 * m)).max())


def assert_close(a, b, tol, what=""):
    assert torch.isfinite(a.detach().float().cpu()).all() or not torch.isfinite(b.detach().float().cpu()).all(), f"{what}: non-finite output"
    r = rel_err(a, b)
    assert r <= tol, f"{what}: rel err {r:.3e} > {tol:.1e}"
    return r
