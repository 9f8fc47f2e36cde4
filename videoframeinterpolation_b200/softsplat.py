"""Drop-in for the reference's ``modules/softsplat.py`` (forward): ``softsplat`` and ``softsplat_func``.

    softsplat(tenIn, tenFlow, tenMetric, strMode)          modules/softsplat.py:237-279
    softsplat_func.apply(tenIn, tenFlow)                   modules/softsplat.py:282-360 (forward)

The reference JIT-compiles one NVRTC kernel per tensor shape and wraps it in cat / exp / multiply /
zero-fill / slice / divide passes; here one C-ABI call does pre-scale + splat (csrc/splat.cu) and one
does the divide -- or a single fused call (csrc/splat_fused.cu) when the flow statistics allow it.
Backward (modules/softsplat.py:364-533): ``vfi_splat_bwd_f32`` behind ``_SplatSumFn``; with grad enabled and an input
that requires grad, ``softsplat`` / ``fwarp`` and the two-frame / multi-branch wrappers of ``warp.py`` run the reference's
own composition around that differentiable summation splat, so gradients reach tenIn, tenFlow and tenMetric.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch

from . import _lib

__all__ = ["softsplat", "softsplat_func", "splat_accumulate", "splat_normalize"]

_WEIGHT_MODES = {"none": 0, "ones": 1, "linear": 2, "soft": 3, "soft_clip_t": 4}
_EPS_MODES = {"add": 0, "zero": 1, "clip": 2}


def splat_accumulate(tenIn, tenFlow, tenMetric, t, num, den, weight: str, accumulate: bool) -> None:
    """num (+)= splat(tenIn * w), den (+)= splat(w) with w chosen by ``weight`` (vfi_splat_f32)."""
    dev = _lib.require_cuda(tenIn, tenFlow, tenMetric, num, den)
    tenIn, tenFlow, tenMetric, t = _lib.f32c(tenIn), _lib.f32c(tenFlow), _lib.f32c(tenMetric), _lib.f32c(t)
    B, C, H, W = tenIn.shape
    assert tuple(tenFlow.shape) == (B, 2, H, W), "flow must be (B,2,H,W)"
    assert tenMetric is None or tuple(tenMetric.shape) == (B, 1, H, W), "metric must be (B,1,H,W)"
    assert num.is_contiguous() and (den is None or den.is_contiguous())
    # a workspace selects the gather form (no floating-point atomics).  Its sums are read-modify-written once per
    # source, so the crossover is higher than for the one-call splat: measured at 1080p, two sources
    # (profiles/r2d_op_bench.txt): C = 16 0.55 vs 0.59 ms (scatter), C = 64 1.22 vs 2.23 ms
    ws = _workspace(dev, int(_lib.lib().vfi_splat_workspace_bytes(B, C, H, W))) if _fused_ok(B, C, H, W, min_c=16) else None
    _lib.call("vfi_splat_f32", dev, tenIn.data_ptr(), tenFlow.data_ptr(), _lib.ptr(tenMetric), _lib.ptr(t),
                                        num.data_ptr(), _lib.ptr(den), B, C, H, W, _WEIGHT_MODES[weight],
                                        1 if accumulate else 0, _lib.ptr(ws), ws.numel() if ws is not None else 0,
                                        _lib.stream_ptr(dev))


def splat_normalize(num, den, eps_mode: str, eps_value: float, want_mask: bool = False,
                    blank_threshold: float = 0.0):
    """out = num / f(den) in place of ``num`` (vfi_splat_normalize_f32); optionally the blank-pixel mask."""
    dev = _lib.require_cuda(num, den)
    B, C, H, W = num.shape
    mask = torch.empty((B, 1, H, W), dtype=torch.uint8, device=dev) if want_mask else None
    _lib.call("vfi_splat_normalize_f32", dev, num.data_ptr(), den.data_ptr(), num.data_ptr(), _lib.ptr(mask), B, C, H, W,
                                                  _EPS_MODES[eps_mode], eps_value, blank_threshold,
                                                  _lib.stream_ptr(dev))
    if want_mask:
        return num, mask.view(torch.bool)
    return num


def splat_last(tenIn, tenFlow, tenMetric, t, num, den, weight: str, eps_mode: str, eps_value: float, blank_threshold: float):
    """Last source of an accumulate sequence on the gather route (vfi_splat_last_f32): adds this source and divides by
    the total weight in the same pass.  Returns (num normalised in place, bool mask), or None when the gather route does
    not apply to this shape (the caller then accumulates and calls ``splat_normalize``)."""
    B, C, H, W = tenIn.shape
    if not _fused_ok(B, C, H, W, min_c=16):
        return None
    dev = _lib.require_cuda(tenIn, tenFlow, tenMetric, num, den)
    tenIn, tenFlow, tenMetric, t = _lib.f32c(tenIn), _lib.f32c(tenFlow), _lib.f32c(tenMetric), _lib.f32c(t)
    ws = _workspace(dev, int(_lib.lib().vfi_splat_workspace_bytes(B, C, H, W)))
    mask = torch.empty((B, 1, H, W), dtype=torch.uint8, device=dev)
    _lib.call("vfi_splat_last_f32", dev, tenIn.data_ptr(), tenFlow.data_ptr(), _lib.ptr(tenMetric), _lib.ptr(t), num.data_ptr(),
              den.data_ptr(), mask.data_ptr(), B, C, H, W, _WEIGHT_MODES[weight], _EPS_MODES[eps_mode], eps_value, blank_threshold,
              ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev))
    return num, mask.view(torch.bool)


def _workspace(dev, nbytes: int) -> torch.Tensor:
    """Scratch for the gather-form splat (claim slots, list heads, nodes), allocated per call: the caching allocator makes
    that nearly free and stream-ordered, so two streams / threads splatting on one device never share a buffer, and a
    CUDA graph that captured a call keeps its own buffer alive (a cached global buffer could be replaced under it)."""
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=dev)


def splat_fused(tenIn, tenFlow, tenMetric, weight: str, eps_mode: str, eps_value: float) -> torch.Tensor:
    """One-call splat + divide without floating-point atomics (vfi_splat_fused_f32, csrc/splat_fused.cu)."""
    dev = _lib.require_cuda(tenIn, tenFlow, tenMetric)
    tenIn, tenFlow, tenMetric = _lib.f32c(tenIn), _lib.f32c(tenFlow), _lib.f32c(tenMetric)
    B, C, H, W = tenIn.shape
    assert tuple(tenFlow.shape) == (B, 2, H, W), "flow must be (B,2,H,W)"
    assert tenMetric is None or tuple(tenMetric.shape) == (B, 1, H, W), "metric must be (B,1,H,W)"
    out = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    nbytes = int(_lib.lib().vfi_splat_workspace_bytes(B, C, H, W))
    ws = _workspace(dev, nbytes)
    _lib.call("vfi_splat_fused_f32", dev, tenIn.data_ptr(), tenFlow.data_ptr(), _lib.ptr(tenMetric), out.data_ptr(), B, C, H, W,
              _WEIGHT_MODES[weight], _EPS_MODES[eps_mode], eps_value, ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev))
    return out


def _fused_ok(B, C, H, W, min_c: int = 8) -> bool:
    """Gather form (no fp atomics) pays a fixed ~0.1 ms/Mpx for building the lists: it wins from ~8 channels up;
    below that the scatter kernel's 4*C atomics per pixel are cheaper.  VFI_SPLAT_ALGO=gather|scatter forces one."""
    algo = os.environ.get("VFI_SPLAT_ALGO", "auto")
    if algo == "scatter" or B * H * W * 4 >= 2 ** 31 or C == 0:
        return False
    return algo == "gather" or C >= min_c


def _check_mode(tenMetric, strMode: str) -> Tuple[str, Optional[str]]:
    # same assertions as the reference (modules/softsplat.py:238-243)
    parts = strMode.split('-')
    assert(parts[0] in ['sum', 'avg', 'linear', 'soft'])
    if strMode == 'sum': assert(tenMetric is None)
    if strMode == 'avg': assert(tenMetric is None)
    if parts[0] == 'linear': assert(tenMetric is not None)
    if parts[0] == 'soft': assert(tenMetric is not None)
    return parts[0], (parts[1] if len(parts) > 1 else None)


def softsplat(tenIn: torch.Tensor, tenFlow: torch.Tensor, tenMetric: torch.Tensor, strMode: str):
    """Forward splatting: 'sum' | 'avg' | 'linear[-eps]' | 'soft[-eps]', eps in addeps / zeroeps / clipeps."""
    head, eps = _check_mode(tenMetric, strMode)
    dev = _lib.require_cuda(tenIn, tenFlow, tenMetric)
    if torch.is_grad_enabled() and (tenIn.requires_grad or tenFlow.requires_grad
                                    or (tenMetric is not None and tenMetric.requires_grad)):
        return _softsplat_autograd(tenIn, tenFlow, tenMetric, head, eps)
    B, C, H, W = tenIn.shape
    weight = {"sum": "none", "avg": "ones", "linear": "linear", "soft": "soft"}[head]
    if _fused_ok(B, C, H, W):
        if head == 'sum':
            return splat_fused(tenIn, tenFlow, None, "none", "add", 0.0)
        eps_mode, eps_value = {None: ("add", 0.0000001), 'addeps': ("add", 0.0000001), 'zeroeps': ("zero", 0.0),
                               'clipeps': ("clip", 0.0000001)}.get(eps, ("add", 0.0))
        return splat_fused(tenIn, tenFlow, tenMetric, weight, eps_mode, eps_value)
    num = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
    if head == 'sum':
        splat_accumulate(tenIn, tenFlow, None, None, num, None, "none", accumulate=False)
        return num
    den = torch.empty((B, 1, H, W), dtype=torch.float32, device=dev)
    splat_accumulate(tenIn, tenFlow, tenMetric, None, num, den, weight, accumulate=False)
    if eps is None or eps == 'addeps':
        return splat_normalize(num, den, "add", 0.0000001)
    if eps == 'zeroeps':
        return splat_normalize(num, den, "zero", 0.0)
    if eps == 'clipeps':
        return splat_normalize(num, den, "clip", 0.0000001)
    # the reference leaves tenNormalize un-adjusted for an unknown suffix (modules/softsplat.py:262-274)
    return splat_normalize(num, den, "add", 0.0)


class _SplatSumFn(torch.autograd.Function):
    """Summation splat with the backward of softsplat_func (modules/softsplat.py:364-533): vfi_splat_bwd_f32."""

    @staticmethod
    def forward(ctx, tenIn, tenFlow):
        tin, tfl = _lib.f32c(tenIn), _lib.f32c(tenFlow)
        ctx.save_for_backward(tin, tfl)
        with torch.no_grad():
            return softsplat_func.apply(tin, tfl)

    @staticmethod
    def backward(ctx, grad_out):
        tin, tfl = ctx.saved_tensors
        dev = tin.device
        B, C, H, W = tin.shape
        g = _lib.f32c(grad_out)
        g_in = torch.empty_like(tin) if ctx.needs_input_grad[0] else None
        g_flow = torch.empty_like(tfl) if ctx.needs_input_grad[1] else None
        _lib.call("vfi_splat_bwd_f32", dev, tin.data_ptr(), tfl.data_ptr(), g.data_ptr(), _lib.ptr(g_in), _lib.ptr(g_flow),
                  B, C, H, W, _lib.stream_ptr(dev))
        return g_in, g_flow


def _softsplat_autograd(tenIn, tenFlow, tenMetric, head: str, eps: Optional[str]):
    """Training path: the reference's own composition (modules/softsplat.py:245-276) around the differentiable
    summation splat, so that autograd reaches tenIn, tenFlow and tenMetric."""
    if head == 'avg':
        tenIn = torch.cat([tenIn, tenIn.new_ones([tenIn.shape[0], 1, tenIn.shape[2], tenIn.shape[3]])], 1)
    elif head == 'linear':
        tenIn = torch.cat([tenIn * tenMetric, tenMetric], 1)
    elif head == 'soft':
        e = tenMetric.exp()
        tenIn = torch.cat([tenIn * e, e], 1)
    out = _SplatSumFn.apply(tenIn, tenFlow)
    if head == 'sum':
        return out
    norm = out[:, -1:, :, :]
    if eps is None or eps == 'addeps':
        norm = norm + 0.0000001
    elif eps == 'zeroeps':
        norm = torch.where(norm == 0.0, torch.ones_like(norm), norm)
    elif eps == 'clipeps':
        norm = norm.clip(0.0000001, None)
    return out[:, :-1, :, :] / norm


class softsplat_func:
    """``softsplat_func.apply(tenIn, tenFlow)``: raw summation splat of every channel (differentiable through
    ``softsplat`` / ``_SplatSumFn``; this entry itself is the plain forward)."""

    @staticmethod
    def apply(tenIn: torch.Tensor, tenFlow: torch.Tensor) -> torch.Tensor:
        if torch.is_grad_enabled() and (tenIn.requires_grad or tenFlow.requires_grad):
            return _SplatSumFn.apply(tenIn, tenFlow)
        dev = _lib.require_cuda(tenIn, tenFlow)          # CPU tensors: the reference asserts False (352-353)
        B, C, H, W = tenIn.shape
        if _fused_ok(B, C, H, W):
            return splat_fused(tenIn, tenFlow, None, "none", "add", 0.0)
        num = torch.empty((B, C, H, W), dtype=torch.float32, device=dev)
        splat_accumulate(tenIn, tenFlow, None, None, num, None, "none", accumulate=False)
        return num

    forward = apply
