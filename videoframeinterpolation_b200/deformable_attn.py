"""Drop-in for the reference's ``modules/deformable_attn.py``.

    Attn(in_c, out_c, n_samples, n_heads).forward(q, kv)                          lines 10-40
    CrossDeformableAttentionBlockwFlow(...).forward(feat_t, feat0, feat1, ft0, ft1)   lines 43-124
        ._get_movement_feats / ._get_ref_coords / ._get_ref_feats                 lines 77-101

``forward`` never materialises the (B, C, 2nS, H*W) sample tensor: one fused kernel (csrc/dattn.cu) takes the
raw ``conv_res_offset`` outputs and does tanh/scale/+flow, sampling of both frames, q/k/v projections, softmax
and the weighted sum.  The piecewise methods keep the reference's signatures and go through the materialising
kernels (csrc/dsample.cu, csrc/attn.cu).
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from . import conv as _conv
from .blocks import FeadForward, HalfChannelConv5ResBlock, conv_bias_act, conv_prelu, flow_up2_add, resize
from .warp import bwarp, _bwarp_forward

__all__ = ["Attn", "CrossDeformableAttentionBlockwFlow", "grouped_sample", "attention_over_samples", "fused_deformable_attention",
           "fused_attention_train"]


class _GroupedSampleFn(torch.autograd.Function):
    """grouped_sample with its backward (vfi_dsample_bwd_f32): the gradients autograd gives the reference's
    ``_get_ref_feats`` (F.grid_sample + the coordinate arithmetic of modules/deformable_attn.py:88-101)."""

    @staticmethod
    def forward(ctx, feat, flow, n_groups, n_samples):
        feat_c, flow_c = _lib.f32c(feat), _lib.f32c(flow)
        ctx.save_for_backward(feat_c, flow_c)
        ctx.ng, ctx.ns = n_groups, n_samples
        with torch.no_grad():
            return grouped_sample(feat_c, flow_c, n_groups, n_samples)

    @staticmethod
    def backward(ctx, grad_out):
        feat, flow = ctx.saved_tensors
        dev = feat.device
        b, c, fh, fw = feat.shape
        g = _lib.f32c(grad_out)
        g_feat = torch.empty_like(feat) if ctx.needs_input_grad[0] else None
        g_flow = torch.empty_like(flow) if ctx.needs_input_grad[1] else None
        _lib.call("vfi_dsample_bwd_f32", dev, feat.data_ptr(), flow.data_ptr(), g.data_ptr(), _lib.linspace_m11(fw, dev).data_ptr(),
                  _lib.linspace_m11(fh, dev).data_ptr(), _lib.ptr(g_feat), _lib.ptr(g_flow), b, c, fh, fw, ctx.ng, ctx.ns,
                  _lib.stream_ptr(dev))
        return g_feat, g_flow, None, None


def grouped_sample(feat: torch.Tensor, flow: torch.Tensor, n_groups: int, n_samples: int, want_grid: bool = False):
    """feat (b,c,h,w), flow (b, nG*nS, 2, h, w) absolute displacement -> samples (b, c, nS, h*w) [, grid (b,nG*nS,h*w,2)].
    Differentiable wrt feat and flow (the grid output is not: the reference only logs it)."""
    if torch.is_grad_enabled() and (feat.requires_grad or flow.requires_grad) and not want_grid:
        return _GroupedSampleFn.apply(feat, flow, n_groups, n_samples)
    dev = _lib.require_cuda(feat, flow)
    feat, flow = _lib.f32c(feat), _lib.f32c(flow)
    b, c, fh, fw = feat.shape
    assert tuple(flow.shape) == (b, n_groups * n_samples, 2, fh, fw), f"flow shape {tuple(flow.shape)}"
    out = torch.empty((b, c, n_samples, fh * fw), dtype=torch.float32, device=dev)
    grid = torch.empty((b, n_groups * n_samples, fh * fw, 2), dtype=torch.float32, device=dev) if want_grid else None
    ext, lin_x, lin_y = _lib.frame_ext(fh, fw, dev)
    _lib.call("vfi_dsample_f32", dev, feat.data_ptr(), flow.data_ptr(), lin_x.data_ptr(), lin_y.data_ptr(), out.data_ptr(),
              _lib.ptr(grid), b, c, fh, fw, n_groups, n_samples, ext, _lib.stream_ptr(dev))
    return (out, grid) if want_grid else out


def _dense_nhwc(t: torch.Tensor) -> bool:
    """(B,C,H,W) fp32 view of a dense (B,H,W,C) buffer, 16-byte aligned."""
    if not (t.dim() == 4 and t.dtype == torch.float32 and t.shape[1] > 1 and t.data_ptr() % 16 == 0):
        return False
    B, C, H, W = t.shape
    return tuple(t.stride()) == (H * W * C, 1, W * C, C)


def _w2d(w: torch.Tensor) -> torch.Tensor:
    return _lib.f32c(w.reshape(w.shape[0], -1))


def attention_over_samples(q, kv, wq, bq, wk, bk, wv, bv, n_heads: int, want_scores: bool = False):
    """Per-pixel multi-head attention of q (b,c,h,w) over kv (b,c,S,h*w); weights (c,c[,1,1]) / biases (c)."""
    dev = _lib.require_cuda(q, kv)
    q, kv = _lib.f32c(q), _lib.f32c(kv)
    b, c, fh, fw = q.shape
    S = kv.shape[2]
    assert tuple(kv.shape) == (b, c, S, fh * fw), f"kv shape {tuple(kv.shape)}"
    out = torch.empty((b, c, fh, fw), dtype=torch.float32, device=dev)
    s0 = s1 = None
    if want_scores:
        s0 = torch.empty((b * n_heads, S // 2, fh * fw), dtype=torch.float32, device=dev)
        s1 = torch.empty_like(s0)
    _lib.call("vfi_attn_f32", dev, q.data_ptr(), kv.data_ptr(), _w2d(wq).data_ptr(), _lib.f32c(bq).data_ptr(),
                                       _w2d(wk).data_ptr(), _lib.f32c(bk).data_ptr(), _w2d(wv).data_ptr(),
                                       _lib.f32c(bv).data_ptr(), out.data_ptr(), _lib.ptr(s0), _lib.ptr(s1), b, c, fh, fw, S,
                                       n_heads, _lib.stream_ptr(dev))
    return (out, s0, s1) if want_scores else out


# (C / nG, nG) pairs vfi_dattn_f32 is instantiated for (csrc/dattn.cu, VFI_DATTN_CASE); other shapes take the
# materialising kernels (vfi_dsample_f32 + vfi_attn_f32)
_FUSED_INSTANCES = {(9, 8), (18, 4), (8, 8), (16, 4), (4, 4), (2, 8), (8, 4), (3, 8), (6, 4)}


def fused_supported(channels: int, n_groups: int, n_heads: int) -> bool:
    return n_groups == n_heads and channels % n_groups == 0 and (channels // n_groups, n_groups) in _FUSED_INSTANCES


def fused_deformable_attention(q, feat0, feat1, offs0, offs1, ft0: Optional[torch.Tensor], ft1: Optional[torch.Tensor],
                               wq, bq, wk, bk, wv, bv, n_groups: int, n_samples: int, n_heads: int, offset_scale: float,
                               coords_final: bool = False, want_scores: bool = False, offs_bias=None,
                               out_channels_last: Optional[bool] = None):
    """Sampler + attention over both frames in one kernel (vfi_dattn_f32).

    offs* (b, nG*nS*2, h, w): raw conv_res_offset outputs (``coords_final=False``: the kernel applies
    offset_scale*tanh(.) + ft*) or final displacements laid out (b, nG*nS, 2, h, w) (``coords_final=True``).
    """
    dev = _lib.require_cuda(q, feat0, feat1, offs0, offs1)
    # q is read once per pixel and out written once: both are taken / produced channels-last when q arrives that way
    # (the conv stacks around the block run NHWC); feat0 / feat1 are gathered and must be planar
    q_cl = (q.dim() == 4 and q.dtype == torch.float32 and q.shape[1] > 1 and not q.is_contiguous()
            and q.is_contiguous(memory_format=torch.channels_last))
    if not q_cl:
        q = _lib.f32c(q)
    # offsets are read in the layout the conv wrote them: channels-last (B,H,W,nG*nS*2) needs no transpose
    offs_cl = (offs0.dim() == 4 and offs0.dtype == torch.float32 and not offs0.is_contiguous()
               and offs0.is_contiguous(memory_format=torch.channels_last)
               and offs1.dtype == torch.float32 and not offs1.is_contiguous()
               and offs1.is_contiguous(memory_format=torch.channels_last))
    if not offs_cl:
        offs0, offs1 = _lib.f32c(offs0), _lib.f32c(offs1)
    offs_bias = _lib.f32c(offs_bias)
    b, c, fh, fw = q.shape
    # features: dense channels-last maps (what the encoder convolutions write) go to the shared-memory window kernel
    # as they are when it covers the shape; everything else is gathered from planar maps
    feat_cl = (_dense_nhwc(feat0) and _dense_nhwc(feat1) and not want_scores
               and os.environ.get("VFI_DATTN_ALGO", "win") != "gather"
               and bool(_lib.lib().vfi_dattn_win_covers(c, n_groups, n_samples, float(offset_scale), 1 if coords_final else 0,
                                                        1 if offs_cl else 0)))
    if not feat_cl:
        feat0, feat1 = _lib.planar(feat0), _lib.planar(feat1)
    assert offs0.numel() == b * n_groups * n_samples * 2 * fh * fw and offs1.numel() == offs0.numel(), "offset shape"
    if not coords_final:
        ft0, ft1 = _lib.f32(ft0), _lib.f32(ft1)
        assert ft0.stride() == ft1.stride(), "ft0 / ft1 must share strides"
        ft_strides = _lib.strides4(ft0)
    else:
        ft_strides = None
    out_cl = q_cl if out_channels_last is None else bool(out_channels_last)   # the kernel writes either layout
    if out_cl:
        out = torch.empty((b, fh, fw, c), dtype=torch.float32, device=dev).permute(0, 3, 1, 2)
    else:
        out = torch.empty((b, c, fh, fw), dtype=torch.float32, device=dev)
    s0 = s1 = None
    if want_scores:
        s0 = torch.empty((b * n_heads, n_samples, fh * fw), dtype=torch.float32, device=dev)
        s1 = torch.empty_like(s0)
    ext, lin_x, lin_y = _lib.frame_ext(fh, fw, dev)
    _lib.call("vfi_dattn_f32", dev, q.data_ptr(), feat0.data_ptr(), feat1.data_ptr(), offs0.data_ptr(), offs1.data_ptr(),
              _lib.ptr(ft0) if not coords_final else None, _lib.ptr(ft1) if not coords_final else None, ft_strides,
              _w2d(wq).data_ptr(), _lib.f32c(bq).data_ptr(), _w2d(wk).data_ptr(), _lib.f32c(bk).data_ptr(), _w2d(wv).data_ptr(),
              _lib.f32c(bv).data_ptr(), lin_x.data_ptr(), lin_y.data_ptr(), out.data_ptr(), _lib.ptr(s0), _lib.ptr(s1),
              b, c, fh, fw, n_groups, n_samples, n_heads, float(offset_scale), 1 if coords_final else 0, _lib.ptr(offs_bias),
              1 if offs_cl else 0, 1 if q_cl else 0, 1 if out_cl else 0, 1 if feat_cl else 0, ext, _lib.stream_ptr(dev))
    return (out, s0, s1) if want_scores else out


def _attn_math(q, kv, wq, bq, wk, bk, wv, bv, n_heads: int):
    """Attn.forward of the reference (modules/deformable_attn.py:26-40) as library ops: q (b,c,h,w), kv (b,c,S,h*w)."""
    b, c, fh, fw = q.shape
    S, hc = kv.shape[2], c // n_heads
    qp = F.conv2d(q, wq.reshape(c, c, 1, 1), bq).reshape(b * n_heads, hc, fh * fw)
    k = F.conv2d(kv, wk.reshape(c, c, 1, 1), bk).reshape(b * n_heads, hc, S, fh * fw)
    v = F.conv2d(kv, wv.reshape(c, c, 1, 1), bv).reshape(b * n_heads, hc, S, fh * fw)
    w = F.softmax(torch.einsum("bcd,bcsd->bsd", qp, k) * hc ** -0.5, dim=1)
    return torch.einsum("bsd,bcsd->bcd", w, v).reshape(b, c, fh, fw)


def _sample_rows(feat, coords_rows, n_groups, n_samples, row0, lin_x, lin_y):
    """grouped_sample for the output rows [row0, row0 + rows): coords_rows (b, nG*nS, 2, rows, w) -> (b, c, nS, rows*w)."""
    b, c, fh, fw = feat.shape
    rows = coords_rows.shape[3]
    out = torch.empty((b, c, n_samples, rows * fw), dtype=torch.float32, device=feat.device)
    _lib.call("vfi_dsample_rows_f32", feat.device, feat.data_ptr(), coords_rows.data_ptr(), lin_x.data_ptr(), lin_y.data_ptr(),
              out.data_ptr(), None, b, c, fh, fw, n_groups, n_samples, row0, rows, None, _lib.stream_ptr(feat.device))
    return out


class _FusedAttnTrainFn(torch.autograd.Function):
    """Training form of the fused sampler + attention.

    forward: vfi_dattn_f32 on final coordinates -- the (b, c, 2 nS, h*w) kv tensor of modules/deformable_attn.py:33-34
    is never formed and nothing but the inputs is kept for the backward.
    backward: the output rows are walked in strips; per strip the samples are RECOMPUTED (vfi_dsample_rows_f32), the
    attention of modules/deformable_attn.py:26-40 is re-run on the strip's kv under autograd (library ops, exactly what the
    reference differentiates), and the strip's d kv goes through the sampler's backward kernel (vfi_dsample_rows_bwd_f32:
    atomics into the whole-frame feature gradients, per-pixel coordinate gradients).  Peak extra memory is ~8x one strip of
    kv (``strip_bytes``, default 64 MB; the block reads ``VFI_DATTN_TRAIN_STRIP_MB``) instead of ~8x the 9.6 GB kv tensor
    of a 1080p pair at 1/2 resolution."""

    @staticmethod
    def forward(ctx, q, feat0, feat1, coords0, coords1, wq, bq, wk, bk, wv, bv, n_groups, n_samples, n_heads, strip_bytes):
        q_c, f0, f1 = _lib.f32c(q), _lib.f32c(feat0), _lib.f32c(feat1)
        c0, c1 = _lib.f32c(coords0), _lib.f32c(coords1)
        ws = tuple(_lib.f32c(t) for t in (wq, bq, wk, bk, wv, bv))
        ctx.save_for_backward(q_c, f0, f1, c0, c1, *ws)
        ctx.conf = (n_groups, n_samples, n_heads, int(strip_bytes))
        with torch.no_grad():
            return fused_deformable_attention(q_c, f0, f1, c0, c1, None, None, *ws, n_groups=n_groups, n_samples=n_samples,
                                              n_heads=n_heads, offset_scale=0.0, coords_final=True, out_channels_last=False)

    @staticmethod
    def backward(ctx, grad_out):
        q, f0, f1, c0, c1, wq, bq, wk, bk, wv, bv = ctx.saved_tensors
        nG, nS, nH, strip_bytes = ctx.conf
        dev = q.device
        b, c, fh, fw = q.shape
        g = _lib.f32c(grad_out)
        need = ctx.needs_input_grad
        lin_x, lin_y = _lib.linspace_m11(fw, dev), _lib.linspace_m11(fh, dev)
        g_q = torch.empty_like(q) if need[0] else None
        g_f0 = torch.zeros_like(f0) if need[1] else None
        g_f1 = torch.zeros_like(f1) if need[2] else None
        g_c0 = torch.empty_like(c0) if need[3] else None
        g_c1 = torch.empty_like(c1) if need[4] else None
        w_leaf = [t.detach().requires_grad_(need[5 + i]) for i, t in enumerate((wq, bq, wk, bk, wv, bv))]
        g_w = [torch.zeros_like(t) if need[5 + i] else None for i, t in enumerate(w_leaf)]
        rows_per = max(1, min(fh, strip_bytes // max(1, b * c * 2 * nS * fw * 4)))
        c0v, c1v = c0.view(b, nG * nS, 2, fh, fw), c1.view(b, nG * nS, 2, fh, fw)
        st = _lib.stream_ptr(dev)
        for r0 in range(0, fh, rows_per):
            rows = min(rows_per, fh - r0)
            cs0, cs1 = c0v[:, :, :, r0:r0 + rows].contiguous(), c1v[:, :, :, r0:r0 + rows].contiguous()
            kv = torch.cat((_sample_rows(f0, cs0, nG, nS, r0, lin_x, lin_y), _sample_rows(f1, cs1, nG, nS, r0, lin_x, lin_y)), dim=2)
            need_kv = need[1] or need[2] or need[3] or need[4]
            kv.requires_grad_(need_kv)
            q_s = q[:, :, r0:r0 + rows].detach().requires_grad_(need[0])
            inputs = [t for t in [q_s, kv, *w_leaf] if t.requires_grad]
            if not inputs:
                break
            with torch.enable_grad():
                att = _attn_math(q_s, kv, *w_leaf, n_heads=nH)
            grads = list(torch.autograd.grad(att, inputs, g[:, :, r0:r0 + rows]))
            if need[0]:
                g_q[:, :, r0:r0 + rows] = grads.pop(0)
            g_kv = grads.pop(0) if need_kv else None
            for i in range(6):
                if need[5 + i]:
                    g_w[i] += grads.pop(0)
            if need_kv:
                for (f, cs, g_f, g_c, lo) in ((f0, cs0, g_f0, g_c0, 0), (f1, cs1, g_f1, g_c1, nS)):
                    if g_f is None and g_c is None:
                        continue
                    gk = g_kv[:, :, lo:lo + nS].contiguous()
                    g_cs = torch.empty_like(cs) if g_c is not None else None
                    _lib.call("vfi_dsample_rows_bwd_f32", dev, f.data_ptr(), cs.data_ptr(), gk.data_ptr(), lin_x.data_ptr(),
                              lin_y.data_ptr(), _lib.ptr(g_f), _lib.ptr(g_cs), b, c, fh, fw, nG, nS, r0, rows, 1, st)
                    if g_c is not None:
                        g_c.view(b, nG * nS, 2, fh, fw)[:, :, :, r0:r0 + rows] = g_cs
            del kv, g_kv, grads
        return (g_q, g_f0, g_f1, g_c0, g_c1, *g_w, None, None, None, None)


def fused_attention_train(q, feat0, feat1, coords0, coords1, wq, bq, wk, bk, wv, bv, n_groups: int, n_samples: int, n_heads: int,
                          strip_bytes: int = 64 << 20):
    """Differentiable fused sampler + attention: q (b,c,h,w), feat0 / feat1 (b,c,h,w), coords0 / coords1 (b, nG*nS, 2, h, w)
    final displacements -> (b, c, h, w).  See :class:`_FusedAttnTrainFn`."""
    return _FusedAttnTrainFn.apply(q, feat0, feat1, coords0, coords1, wq, bq, wk, bk, wv, bv, n_groups, n_samples, n_heads,
                                   strip_bytes)


class Attn(nn.Module):
    def __init__(self, in_c, out_c, n_samples, n_heads):
        super().__init__()
        assert in_c == out_c, "the fused projections assume in_c == out_c (true for every reference block)"
        self.in_c, self.out_c = in_c, out_c
        self.q_proj = nn.Conv2d(in_c, out_c, 1, 1, 0)
        self.k_proj = nn.Conv2d(in_c, out_c, 1, 1, 0)
        self.v_proj = nn.Conv2d(in_c, out_c, 1, 1, 0)
        self.n_samples = n_samples
        self.n_head = n_heads
        self.n_head_c = out_c // n_heads
        self.scale = self.n_head_c ** -0.5

    def _weights(self):
        return (self.q_proj.weight, self.q_proj.bias, self.k_proj.weight, self.k_proj.bias, self.v_proj.weight,
                self.v_proj.bias)

    def forward(self, q, kv):
        """q: b, c, fh, fw;  kv: b, c, nS, fh*fw  ->  b, out_c, fh, fw"""
        if torch.is_grad_enabled() and (q.requires_grad or kv.requires_grad or self.q_proj.weight.requires_grad):
            return self._forward_autograd(q, kv)
        return attention_over_samples(q, kv, *self._weights(), n_heads=self.n_head)

    def _forward_autograd(self, q, kv):
        """Training route: the projections and the two contractions as library ops (what modules/deformable_attn.py:26-40
        runs), so that autograd reaches the sampler's and the warp's backward kernels."""
        b, c, fh, fw = q.shape
        S = kv.shape[2]
        qp = self.q_proj(q).reshape(b * self.n_head, self.n_head_c, fh * fw)
        k = self.k_proj(kv).reshape(b * self.n_head, self.n_head_c, S, fh * fw)
        v = self.v_proj(kv).reshape(b * self.n_head, self.n_head_c, S, fh * fw)
        w = F.softmax(torch.einsum("bcd,bcsd->bsd", qp, k) * self.scale, dim=1)
        return torch.einsum("bsd,bcsd->bcd", w, v).reshape(b, self.out_c, fh, fw)


def movement_feats_pair(me, feat_t, feat0, feat1, ft0, ft1):
    """(movement features of frame 0, of frame 1): views of :func:`movement_feats_batched`'s result."""
    mv = movement_feats_batched(me, feat_t, feat0, feat1, ft0, ft1)
    b = feat_t.shape[0]
    return mv[:b], mv[b:]


def movement_feats_batched(me, feat_t, feat0, feat1, ft0, ft1):
    """Both movement features, as ONE tensor of 2b images, with the first two convs split over their input channels (a convolution is linear
    in them), all passes on the tensor-core kernel:
        conv0(cat(feat_t, warped, flow)) = conv0_a(feat_t) + conv0_b(cat(warped, flow))
    The feat_t term is shared by the two frames, the concat (and cuDNN's channel padding of its 146 channels)
    disappears, and conv1 (144 -> 72) runs as two 72-channel passes.  Same sums up to fp32 addition order."""
    c0, p0, c1, p1 = me[0][0], me[0][1], me[1][0], me[1][1]
    C = feat_t.shape[1]
    b, _, fh, fw = feat_t.shape
    shared = _conv.conv3x3(c0, feat_t, in_slice=(0, C, 0))                       # + bias, no activation
    if os.environ.get("VFI_DAT_BATCH_FRAMES", "1") == "0":                       # A/B switch: one pass per frame
        outs = []
        for feat_x, ftx in ((feat0, ft0), (feat1, ft1)):
            wf = torch.empty((b, fh, fw, C + 4), dtype=torch.float32, device=feat_t.device).permute(0, 3, 1, 2)
            _bwarp_forward(feat_x, ftx, out=wf[:, :C])
            wf[:, C:C + 2] = ftx
            wf[:, C + 2:] = 0.0
            y = _conv.conv3x3(c0, wf, p0.weight, residual=shared, use_bias=False, in_slice=(C, 2 * C + 2, 2))
            r = _conv.conv3x3(c1, y[:, :C], in_slice=(0, C, 0))
            z = _conv.conv3x3(c1, y[:, C:], p1.weight, residual=r, use_bias=False, in_slice=(C, 2 * C, 0))
            outs.append(me[2](z))
        return torch.cat(outs, dim=0)
    # both frames as one batch of 2b images (frame 0 first): half the launches, and the 1/4 / 1/8-resolution levels fill
    # the machine better; the shared term is broadcast by the conv kernel (residual image = b % batch)
    wf = torch.empty((2 * b, fh, fw, C + 4), dtype=torch.float32, device=feat_t.device).permute(0, 3, 1, 2)
    if feat0.stride(1) == 1 and feat1.stride(1) == 1 and C % 4 == 0 and feat0.data_ptr() % 16 == 0 and feat1.data_ptr() % 16 == 0:
        _bwarp_forward(feat0, ft0, out=wf[:b, :C], flowcat=True)                 # warp + flow channels + zero pad in one pass
        _bwarp_forward(feat1, ft1, out=wf[b:, :C], flowcat=True)
    else:
        _bwarp_forward(feat0, ft0, out=wf[:b, :C])
        _bwarp_forward(feat1, ft1, out=wf[b:, :C])
        wf[:b, C:C + 2] = ft0
        wf[b:, C:C + 2] = ft1
        wf[:, C + 2:] = 0.0
    y = _conv.conv3x3(c0, wf, p0.weight, residual=shared, use_bias=False, in_slice=(C, 2 * C + 2, 2))
    r = _conv.conv3x3(c1, y[:, :C], in_slice=(0, C, 0))
    z = _conv.conv3x3(c1, y[:, C:], p1.weight, residual=r, use_bias=False, in_slice=(C, 2 * C, 0))
    return me[2](z)                                                              # (2b, C, fh, fw): [:b] frame 0, [b:] frame 1


def movement_split_ok(me, feat_t, feat0):
    C = feat_t.shape[1]
    return (C % 4 == 0 and _conv.part_eligible(me[0][0], feat_t) and _conv.part_eligible(me[1][0], feat_t)
            and feat0.is_cuda and feat0.dtype == torch.float32
            and bool(_lib.lib().vfi_conv3x3_supported(C + 4, C + 4, me[0][0].out_channels, 1))
            and bool(_lib.lib().vfi_conv3x3_supported(C, 2 * C, me[1][0].out_channels, 1)))


class CrossDeformableAttentionBlockwFlow(nn.Module):
    def __init__(self, in_c, out_c, n_samples=9, n_groups=12, n_heads=12, mlp_ratio=2.0, offset_scale=2.0,
                 pred_res_flow=True):
        super().__init__()
        self.in_c, self.out_c = in_c, out_c
        self.n_groups, self.n_group_c = n_groups, in_c // n_groups
        self.n_heads, self.n_samples = n_heads, n_samples
        self.offset_scale, self.pred_res_flow = offset_scale, pred_res_flow
        self.movement_extractor = nn.Sequential(conv_prelu(in_c * 2 + 2, in_c * 2), conv_prelu(in_c * 2, in_c),
                                                HalfChannelConv5ResBlock(in_c, in_c // 2))
        self.conv_res_offset = nn.Conv2d(in_c, n_groups * n_samples * 2, 3, 1, 1)
        nn.init.zeros_(self.conv_res_offset.weight)       # modules/deformable_attn.py:64
        nn.init.zeros_(self.conv_res_offset.bias)
        if pred_res_flow:
            self.conv_res_flow = nn.ConvTranspose2d(in_c, 2, 4, 2, 1, bias=True)
        self.attn = Attn(in_c, out_c, n_samples * 2, n_heads)
        self.mlp = FeadForward(in_dim=out_c, hidden_dim=out_c * mlp_ratio, out_dim=out_c)

    # -- piecewise API of the reference ---------------------------------------------------------
    def _get_movement_feats(self, feat_t, feat_x, ftx):
        return self.movement_extractor(torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1))

    def _get_ref_coords(self, feat_t, ftx, movement_feat):
        b, c, fh, fw = feat_t.shape
        res = self.offset_scale * torch.tanh(self.conv_res_offset(movement_feat))
        return res.view(b, self.n_groups * self.n_samples, 2, fh, fw) + ftx.unsqueeze(1)

    def _get_ref_feats(self, feat, flow):
        return grouped_sample(feat, flow, self.n_groups, self.n_samples)        # b, c, nS, fh*fw

    def _movement_feats_pair(self, feat_t, feat0, feat1, ft0, ft1):
        return movement_feats_pair(self.movement_extractor, feat_t, feat0, feat1, ft0, ft1)

    def _split_ok(self, feat_t, feat0):
        return movement_split_ok(self.movement_extractor, feat_t, feat0)

    # -- fused forward ------------------------------------------------------------------------------
    def _forward_autograd(self, feat_t, feat0, feat1, ft0, ft1):
        """Training route (grad enabled): the reference's piecewise composition (modules/deformable_attn.py:103-133) over
        the differentiable drop-ins -- bwarp runs its CUDA kernels forward and backward, and the sampler + attention run
        fused forward with a strip-wise recomputing backward (:func:`fused_attention_train`; ``VFI_DATTN_TRAIN=materialize``
        restores the kv-materialising composition)."""
        mv0 = self._get_movement_feats(feat_t, feat0, ft0)
        mv1 = self._get_movement_feats(feat_t, feat1, ft1)
        c0, c1 = self._get_ref_coords(feat_t, ft0, mv0), self._get_ref_coords(feat_t, ft1, mv1)
        if fused_supported(self.in_c, self.n_groups, self.n_heads) and os.environ.get("VFI_DATTN_TRAIN", "fused") != "materialize":
            # fused forward, strip-wise recomputing backward: no (b, c, 2 nS, h*w) tensor is kept for autograd
            attended = fused_attention_train(feat_t, feat0, feat1, c0, c1, *self.attn._weights(), n_groups=self.n_groups,
                                             n_samples=self.n_samples, n_heads=self.n_heads,
                                             strip_bytes=int(os.environ.get("VFI_DATTN_TRAIN_STRIP_MB", "64")) << 20)
        else:
            kv = torch.cat((self._get_ref_feats(feat0, c0), self._get_ref_feats(feat1, c1)), dim=2)
            attended = self.attn(feat_t, kv)
        out = attended + self.mlp(attended)
        if not self.pred_res_flow:
            return out
        return (out, self.conv_res_flow(mv0) + 2.0 * resize(ft0, scale_factor=2.0),
                self.conv_res_flow(mv1) + 2.0 * resize(ft1, scale_factor=2.0))

    def _wants_grad(self, *tensors) -> bool:
        """The materialising training route is needed only when something can receive a gradient: with grad merely enabled
        (model.eval()(x0, x1, t) without no_grad) the fused kernel still runs -- the (B, C, 2nS, HW) kv tensors of the
        training route are ~40 GB per 1080p pair."""
        if not torch.is_grad_enabled():
            return False
        return any(t.requires_grad for t in tensors) or (self.training and any(p.requires_grad for p in self.parameters()))

    def forward(self, feat_t, feat0, feat1, ft0, ft1):
        if self._wants_grad(feat_t, feat0, feat1, ft0, ft1):
            return self._forward_autograd(feat_t, feat0, feat1, ft0, ft1)
        if torch.is_grad_enabled():
            with torch.no_grad():
                return self.forward(feat_t, feat0, feat1, ft0, ft1)
        nb = feat_t.shape[0]
        mvb = None                      # both frames' movement features as one batch of 2 nb images (frame 0 first)
        if self._split_ok(feat_t, feat0):
            mvb = movement_feats_batched(self.movement_extractor, feat_t, feat0, feat1, ft0, ft1)
            mv0, mv1 = mvb[:nb], mvb[nb:]
            if os.environ.get("VFI_DAT_BATCH_FRAMES", "1") == "0":       # A/B switch: per-frame launches after the extractor
                mvb = None
        else:
            mv0 = self._get_movement_feats(feat_t, feat0, ft0)
            mv1 = self._get_movement_feats(feat_t, feat1, ft1)
        if fused_supported(self.in_c, self.n_groups, self.n_heads):
            # bias-free conv: the fused kernel adds conv_res_offset.bias itself (same fp32 order: conv + bias, then tanh)
            co = self.conv_res_offset
            if mvb is not None and _conv.eligible(co, mvb):
                raw = _conv.conv3x3(co, mvb, use_bias=False)                    # one pass per Cout tile for both frames
                raw0, raw1 = raw[:nb], raw[nb:]
            elif _conv.eligible(co, mv0) and _conv.eligible(co, mv1):
                raw0, raw1 = _conv.conv3x3(co, mv0, use_bias=False), _conv.conv3x3(co, mv1, use_bias=False)
            else:
                raw0 = F.conv2d(mv0, co.weight, None, co.stride, co.padding)
                raw1 = F.conv2d(mv1, co.weight, None, co.stride, co.padding)
            attended = fused_deformable_attention(feat_t, feat0, feat1, raw0, raw1, ft0, ft1, *self.attn._weights(),
                                                  n_groups=self.n_groups, n_samples=self.n_samples, n_heads=self.n_heads,
                                                  offset_scale=self.offset_scale, offs_bias=co.bias)
        else:   # groups != heads: materialising kernels (still CUDA, never a CPU path)
            raw0, raw1 = conv_bias_act(self.conv_res_offset, mv0), conv_bias_act(self.conv_res_offset, mv1)
            b, _, fh, fw = feat_t.shape
            shp = (b, self.n_groups * self.n_samples, 2, fh, fw)
            c0 = (self.offset_scale * torch.tanh(raw0)).view(shp) + ft0.unsqueeze(1)
            c1 = (self.offset_scale * torch.tanh(raw1)).view(shp) + ft1.unsqueeze(1)
            kv = torch.cat((self._get_ref_feats(feat0, c0), self._get_ref_feats(feat1, c1)), dim=2)
            attended = self.attn(feat_t, kv)
        out = self.mlp.forward_residual(attended)
        if not self.pred_res_flow:
            return out
        if mvb is not None:
            rf = _conv.conv_transpose_up2(self.conv_res_flow, mvb, out_channels_last=False)
            return out, flow_up2_add(rf[:nb], ft0), flow_up2_add(rf[nb:], ft1)
        up_ft0 = flow_up2_add(_conv.conv_transpose_up2(self.conv_res_flow, mv0, out_channels_last=False), ft0)
        up_ft1 = flow_up2_add(_conv.conv_transpose_up2(self.conv_res_flow, mv1, out_channels_last=False), ft1)
        return out, up_ft0, up_ft1
