"""Drop-in for the reference's ``modules/warp.py``: same names, argument order and return types.

    bwarp(img, flow)                                                       modules/warp.py:7-15
    fwarp(tenIn, tenFlow, tenMetric, strMode='soft')                       modules/warp.py:18-59
    fwarp_using_two_frames(in1, flow1, t1, in2, flow2, t2, m1, m2)         modules/warp.py:68-83
    fwarp_mframes(in1, flow1, t1, in2, flow2, t2, m1, m2)                  modules/warp.py:86-114

All arithmetic runs in libvfi_b200.so (csrc/bwarp.cu, csrc/splat.cu, csrc/splat_fused.cu).
"""
from __future__ import annotations

import torch

from . import _lib
from .softsplat import _SplatSumFn, softsplat, splat_accumulate, splat_last, splat_normalize

__all__ = ["bwarp", "fwarp", "fwarp_using_two_frames", "fwarp_mframes"]


class _BwarpFn(torch.autograd.Function):
    """bwarp with its backward (vfi_bwarp_bwd_f32): the gradients F.grid_sample's autograd gives the reference."""

    @staticmethod
    def forward(ctx, img, flow):
        img_c, flow_c = _lib.f32c(img), _lib.f32c(flow)
        ctx.save_for_backward(img_c, flow_c)
        with torch.no_grad():
            return _bwarp_forward(img_c, flow_c)

    @staticmethod
    def backward(ctx, grad_out):
        img, flow = ctx.saved_tensors
        dev = img.device
        B, C, H, W = img.shape
        g = _lib.f32c(grad_out)
        g_img = torch.empty_like(img) if ctx.needs_input_grad[0] else None
        g_flow = torch.empty_like(flow) if ctx.needs_input_grad[1] else None
        _lib.call("vfi_bwarp_bwd_f32", dev, img.data_ptr(), flow.data_ptr(), g.data_ptr(), _lib.linspace_m11(W, dev).data_ptr(),
                  _lib.linspace_m11(H, dev).data_ptr(), _lib.ptr(g_img), _lib.ptr(g_flow), B, C, H, W, _lib.stream_ptr(dev))
        return g_img, g_flow


def bwarp(img: torch.Tensor, flow: torch.Tensor) -> torch.Tensor:
    """Backward bilinear warp, border clamp.  img (B,C,H,W), flow (B,2,H,W) [ch0 = dx, ch1 = dy, pixels].

    NCHW-contiguous and channels-last inputs both run vectorised kernels; the output has the input's
    memory format.  Flow may be a non-contiguous channel slice (e.g. ``pred_scale_3[:, nf:nf+2]``,
    models/DAT.py:82): strides are passed through, nothing is copied.  Differentiable wrt both arguments.
    """
    if torch.is_grad_enabled() and (img.requires_grad or flow.requires_grad):
        _lib.require_cuda(img, flow)
        return _BwarpFn.apply(img, flow)
    return _bwarp_forward(img, flow)


def _bwarp_forward(img: torch.Tensor, flow: torch.Tensor, out: torch.Tensor = None, flowcat: bool = False) -> torch.Tensor:
    """``out``: optional (B,C,H,W) fp32 view to write into (any strides, e.g. a channel slice of a wider buffer).
    ``flowcat``: out is the first C channels of a channels-last buffer with >= C + 4 channels per pixel; the kernel also
    writes (flow_x, flow_y, 0, 0) into channels [C, C + 4) (vfi_bwarp_flowcat_f32)."""
    dev = _lib.require_cuda(img, flow)
    B, _, H, W = flow.shape
    assert img.shape[0] == B and tuple(img.shape[2:]) == (H, W), "img / flow shape mismatch"
    img, flow = _lib.f32(img), _lib.f32(flow)
    C = img.shape[1]
    if out is not None:
        assert tuple(out.shape) == (B, C, H, W) and out.dtype == torch.float32 and out.device == dev, "bwarp: bad out"
    elif img.stride(1) == 1 and C > 1:
        out = torch.empty_like(img, memory_format=torch.channels_last)
    else:
        out = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    ext, lin_x, lin_y = _lib.frame_ext(H, W, dev)
    _lib.call("vfi_bwarp_flowcat_f32" if flowcat else "vfi_bwarp_f32", dev, img.data_ptr(), flow.data_ptr(), lin_x.data_ptr(),
              lin_y.data_ptr(), out.data_ptr(), B, C, H, W, _lib.strides4(img), _lib.strides4(flow), _lib.strides4(out), ext,
              _lib.stream_ptr(dev))
    return out


def fwarp(tenIn: torch.Tensor, tenFlow: torch.Tensor, tenMetric: torch.Tensor, strMode: str = 'soft'):
    """Forward splatting front-end -- identical contract to ``softsplat`` (the reference keeps two copies)."""
    return softsplat(tenIn, tenFlow, tenMetric, strMode)


def _two_directions(num, den, tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1, tenMetric2, first: bool, last_eps=None):
    """Both directions of one branch into num / den.  ``last_eps``: this is the final branch -- the second direction
    also divides by the total weight + last_eps (vfi_splat_last_f32) and (out, mask) is returned when the gather route
    applies; None otherwise (the caller normalises)."""
    B = tenIn1.shape[0]
    splat_accumulate(tenIn1, tenFlow1, tenMetric1, t1.reshape(B), num, den, "soft_clip_t", accumulate=not first)
    if last_eps is not None:
        done = splat_last(tenIn2, tenFlow2, tenMetric2, t2.reshape(B), num, den, "soft_clip_t", "add", last_eps, 0.00001)
        if done is not None:
            return done
    splat_accumulate(tenIn2, tenFlow2, tenMetric2, t2.reshape(B), num, den, "soft_clip_t", accumulate=True)
    return None


def _needs_grad(*tensors) -> bool:
    return torch.is_grad_enabled() and any(t is not None and torch.is_tensor(t) and t.requires_grad for t in tensors)


def _one_direction_autograd(tenIn, tenFlow, td, tenMetric):
    """modules/warp.py:69-73 / 91-96 over the differentiable summation splat (training route)."""
    e = td * tenMetric.clip(-20.0, 20.0).exp()
    out = _SplatSumFn.apply(torch.cat([tenIn * td * tenMetric.clip(-20.0, 20.0).exp(), e], 1), tenFlow)
    return out[:, :-1], out[:, -1:] + 0.0000001


def fwarp_using_two_frames(tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1=None, tenMetric2=None):
    """Two-direction soft splat blended by t; returns (num / den, den < 1e-5).

    Both directions accumulate into ONE numerator / denominator pair (2 kernel launches instead of the
    reference's 2 splats + ~10 elementwise passes); the two +1e-7 terms of modules/warp.py:73 become
    one +2e-7 in the divide.
    """
    assert tenMetric1 is not None and tenMetric2 is not None, "fwarp_using_two_frames needs both metrics"
    dev = _lib.require_cuda(tenIn1, tenFlow1, tenIn2, tenFlow2, tenMetric1, tenMetric2)
    if _needs_grad(tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1, tenMetric2):
        # training route: the reference's composition (modules/warp.py:68-83), gradients through vfi_splat_bwd_f32
        of, nf = _one_direction_autograd(tenIn1, tenFlow1, t1, tenMetric1)
        ob, nb = _one_direction_autograd(tenIn2, tenFlow2, t2, tenMetric2)
        den_ = nf + nb
        return (of + ob) / den_, den_ < 0.00001
    B, C, H, W = tenIn1.shape
    num = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    den = torch.empty((B, 1, H, W), dtype=torch.float32, device=dev)
    done = _two_directions(num, den, tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1, tenMetric2, first=True,
                           last_eps=2 * 0.0000001)
    if done is not None:
        return done
    return splat_normalize(num, den, "add", 2 * 0.0000001, want_mask=True, blank_threshold=0.00001)


def fwarp_mframes(tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1, tenMetric2):
    """Many-to-many splatting: flows (n_branch,B,2,H,W), t (n_branch,B,1,1,1); inputs shared by branches."""
    dev = _lib.require_cuda(tenIn1, tenFlow1, tenIn2, tenFlow2, tenMetric1, tenMetric2)
    n_branch = tenFlow1.shape[0]
    if _needs_grad(tenIn1, tenFlow1, t1, tenIn2, tenFlow2, t2, tenMetric1, tenMetric2):
        # training route: modules/warp.py:86-114 over the differentiable summation splat
        num_, den_ = 0, 0
        for idx in range(n_branch):
            of, nf = _one_direction_autograd(tenIn1, tenFlow1[idx], t1[idx], tenMetric1)
            ob, nb = _one_direction_autograd(tenIn2, tenFlow2[idx], t2[idx], tenMetric2)
            num_, den_ = num_ + (of + ob), den_ + (nf + nb)
        return num_ / den_, den_ < 0.00001
    B, C, H, W = tenIn1.shape
    num = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    den = torch.empty((B, 1, H, W), dtype=torch.float32, device=dev)
    for idx in range(n_branch):
        done = _two_directions(num, den, tenIn1, tenFlow1[idx], t1[idx], tenIn2, tenFlow2[idx], t2[idx], tenMetric1, tenMetric2,
                               first=idx == 0, last_eps=2 * n_branch * 0.0000001 if idx == n_branch - 1 else None)
        if done is not None:
            return done
    return splat_normalize(num, den, "add", 2 * n_branch * 0.0000001, want_mask=True, blank_threshold=0.00001)
