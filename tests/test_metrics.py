"""PSNR / SSIM validators (evaluate.py:29-84).

CPU: the oracle restatement against values computed by the reference's own functions (tests/golden/metrics.npz).
GPU: vfi_frame_metrics_f32 through the reference-named wrappers against the fixtures, against the oracle on further
sizes (tiles with ragged edges, 1080p), and the properties SSIM(x, x) = 1, symmetry and run-to-run reproducibility.
"""
import pytest
import torch

import _det as det
import golden_cases as gc
import oracle

TOL = 1e-5


def _close(a, b, tol, what):
    a, b = float(a), float(b)
    # SSIM lives on a unit scale (it is a mean of mixed-sign map values, ~0 for unrelated frames): absolute below 1
    assert abs(a - b) <= tol * max(abs(b), 1.0), f"{what}: {a} vs {b}"


def test_metrics_oracle_matches_reference(golden):
    g = golden("metrics")
    for key, (a, b, L) in gc.metrics_cases().items():
        ssim, cs = oracle.calculate_ssim(a, b, val_range=L, full=True)
        _close(oracle.calculate_psnr(a, b), g[key][0], 1e-6, f"psnr[{key}]")
        _close(ssim, g[key][1], 1e-6, f"ssim[{key}]")
        _close(cs, g[key][2], 1e-6, f"cs[{key}]")


def test_metrics_refuse_cpu_tensors():
    import videoframeinterpolation_b200 as vfi
    a = torch.rand(1, 3, 16, 16)
    with pytest.raises(Exception):
        vfi.calculate_psnr(a, a)


@pytest.mark.gpu
def test_metrics_golden(golden):
    import videoframeinterpolation_b200 as vfi
    g = golden("metrics")
    for key, (a, b, L) in gc.metrics_cases().items():
        ssim, cs = vfi.calculate_ssim(a.cuda(), b.cuda(), full=True, val_range=L)
        _close(vfi.calculate_psnr(a.cuda(), b.cuda()), g[key][0], TOL, f"psnr[{key}]")
        _close(ssim, g[key][1], TOL, f"ssim[{key}]")
        _close(cs, g[key][2], TOL, f"cs[{key}]")
    a, b, _ = gc.metrics_cases()["unit"]
    _close(vfi.calculate_psnr(a[0].cuda(), b[0].cuda()), oracle.calculate_psnr(a[0], b[0]), TOL, "psnr of a (C,H,W) pair")


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(2, 3, 97, 131), (1, 3, 256, 448), (3, 3, 32, 33), (1, 3, 1080, 1920)])
def test_metrics_vs_oracle_and_properties(shape):
    import videoframeinterpolation_b200 as vfi
    a = det.uniform("mtg.a", shape)
    b = (a + 0.03 * det.normal("mtg.n", shape)).clamp(0, 1)
    ca, cb = a.cuda(), b.cuda()
    ssim, cs = vfi.calculate_ssim(ca, cb, full=True)
    o_ssim, o_cs = oracle.calculate_ssim(a, b, full=True)
    _close(ssim, o_ssim, TOL, "ssim")
    _close(cs, o_cs, TOL, "cs")
    _close(vfi.calculate_psnr(ca, cb), oracle.calculate_psnr(a, b), TOL, "psnr")
    # per-sample values of one launch == the reference functions called sample by sample
    psnr_b, ssim_b = vfi.frame_metrics(ca, cb)
    for i in range(shape[0]):
        _close(psnr_b[i], oracle.calculate_psnr(a[i], b[i]), TOL, f"psnr[{i}]")
        _close(ssim_b[i], oracle.calculate_ssim(a[i:i + 1], b[i:i + 1], val_range=1.0), TOL, f"ssim[{i}]")
    # properties: identical frames, symmetry, bitwise reproducibility
    _close(vfi.calculate_ssim(ca, ca), 1.0, 1e-6, "ssim(x, x)")
    assert torch.isinf(vfi.calculate_psnr(ca, ca))
    _close(vfi.calculate_ssim(cb, ca, val_range=1.0), vfi.calculate_ssim(ca, cb, val_range=1.0), 1e-6, "symmetry")
    assert torch.equal(vfi.calculate_ssim(ca, cb), ssim)


@pytest.mark.gpu
def test_metrics_edge_cases():
    import videoframeinterpolation_b200 as vfi
    from videoframeinterpolation_b200._lib import VfiError
    with pytest.raises(VfiError):                       # window larger than the frame: refused, not approximated
        vfi.calculate_ssim(torch.rand(1, 3, 8, 40).cuda(), torch.rand(1, 3, 8, 40).cuda())
    empty = torch.empty(0, 3, 16, 16).cuda()
    p, s = vfi.frame_metrics(empty, empty)
    assert p.numel() == 0 and s.numel() == 0
    a = torch.rand(1, 3, 24, 24).cuda()                 # non-contiguous views are accepted
    b = torch.rand(1, 3, 24, 48).cuda()[..., ::2]
    _close(vfi.calculate_ssim(a, b, val_range=1.0), oracle.calculate_ssim(a.cpu(), b.cpu().contiguous(), val_range=1.0), TOL, "strided")
