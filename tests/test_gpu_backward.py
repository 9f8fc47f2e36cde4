"""GPU parity of the backward kernels through autograd: the gradients of the
CUDA drop-ins against torch autograd run through the oracle's differentiable restatement of the same forward
(oracle.splat_sum is built from index_add_, oracle.bwarp from F.grid_sample: their autograd is the analytic derivative
the reference's softsplat_ingrad / softsplat_flowgrad kernels and grid_sample's backward compute)."""
import pytest
import torch

import _det as det
import oracle
from _util import rel_err

pytestmark = pytest.mark.gpu


def _grads(fn, *tensors):
    leaves = [t.clone().requires_grad_(True) if t is not None else None for t in tensors]
    out = fn(*leaves)
    w = det.normal("bwd.cotangent", tuple(out.shape)).to(out.device)
    (out * w).sum().backward()
    return out.detach(), [None if t is None else t.grad.detach() for t in leaves]


@pytest.mark.parametrize("shape", [(2, 5, 19, 23), (1, 64, 32, 40)])
def test_bwarp_backward_matches_grid_sample_autograd(shape):
    import videoframeinterpolation_b200 as v
    B, C, H, W = shape
    img = det.normal(f"bwd.img.{shape}", shape)
    flow = det.flow_field(f"bwd.flow.{shape}", B, H, W, sigma=2.5, outlier_frac=0.05)
    with oracle.library_ops():
        ref_out, (ref_gi, ref_gf) = _grads(oracle.bwarp, img, flow)
    out, (gi, gf) = _grads(v.bwarp, img.cuda(), flow.cuda())
    assert rel_err(out, ref_out) <= 1e-5
    assert rel_err(gi, ref_gi) <= 1e-5
    assert rel_err(gf, ref_gf) <= 1e-4        # C-term sums of products of differences
    # image-only and flow-only gradients
    x = img.cuda().requires_grad_(True)
    v.bwarp(x, flow.cuda()).sum().backward()
    ones_gi = _grads(lambda a: v.bwarp(a, flow.cuda()), img.cuda())[1][0]
    assert x.grad is not None and x.grad.shape == ones_gi.shape


@pytest.mark.parametrize("shape", [(2, 4, 17, 21), (1, 16, 24, 36)])
def test_splat_sum_backward_matches_autograd(shape):
    from videoframeinterpolation_b200.softsplat import softsplat_func
    B, C, H, W = shape
    x = det.normal(f"bwd.sin.{shape}", shape)
    flow = det.flow_field(f"bwd.sflow.{shape}", B, H, W, sigma=2.0, outlier_frac=0.05, nonfinite_frac=0.01)
    ref_out, (ref_gx, ref_gf) = _grads(oracle.splat_sum, x, flow)
    out, (gx, gf) = _grads(softsplat_func.apply, x.cuda(), flow.cuda())
    assert rel_err(out, ref_out) <= 1e-4
    assert rel_err(gx, ref_gx) <= 1e-5
    fin = torch.isfinite(ref_gf)              # autograd through a NaN flow gives NaN; the kernels give 0 there (as the reference)
    assert rel_err(torch.where(fin, gf.cpu(), torch.zeros_like(ref_gf)), torch.where(fin, ref_gf, torch.zeros_like(ref_gf))) <= 1e-4
    assert bool((gf.cpu()[~fin] == 0).all())


@pytest.mark.parametrize("mode", ["avg", "linear", "soft", "soft-zeroeps", "linear-clipeps"])
def test_softsplat_backward_all_modes(mode):
    import videoframeinterpolation_b200 as v
    B, C, H, W = 2, 6, 18, 22
    x = det.normal(f"bwd.ss.{mode}.x", (B, C, H, W))
    flow = det.flow_field(f"bwd.ss.{mode}.f", B, H, W, sigma=2.0, outlier_frac=0.03)
    head = mode.split("-")[0]
    metric = None if head == "avg" else (det.uniform(f"bwd.ss.{mode}.m", (B, 1, H, W), lo=0.2, hi=1.5) if head == "linear"
                                         else det.normal(f"bwd.ss.{mode}.m", (B, 1, H, W), std=0.7))
    ref_out, ref_g = _grads(lambda a, f, m: oracle.fwarp(a, f, m, mode), x, flow, metric)
    cu = lambda t: None if t is None else t.cuda()
    out, g = _grads(lambda a, f, m: v.softsplat(a, f, m, mode), cu(x), cu(flow), cu(metric))
    assert rel_err(out, ref_out) <= 1e-4
    for name, a, b in zip(("in", "flow", "metric"), g, ref_g):
        if b is not None:
            assert rel_err(a, b) <= 2e-4, name


def test_inference_path_unchanged_without_grad():
    import videoframeinterpolation_b200 as v
    x = det.normal("bwd.nograd.x", (1, 16, 20, 24)).cuda()
    flow = det.flow_field("bwd.nograd.f", 1, 20, 24).cuda()
    m = det.normal("bwd.nograd.m", (1, 1, 20, 24)).cuda()
    with torch.no_grad():
        a = v.softsplat(x.requires_grad_(True), flow, m, "soft")
    assert not a.requires_grad
