"""Drop-in for the reference's ``modules/dcnv2.py`` plus the op it calls.

    deform_conv2d(input, offset, weight, bias, stride, padding, dilation, mask)   torchvision.ops signature
    DeformableConv2d(in_channels, out_channels, ...).forward(x, movement_feat)    modules/dcnv2.py:8-67
    DeformableConv2dGivenOffset(in_c, out_c).forward(x, offset, mask)             models/archive/DCNDAT.py:57-78

Parameter names (``offset_flow_conv``, ``conv_offset_mask.{0..4}``, ``regular_conv``) match the reference
so its checkpoints load.  The dense 3x3 convs around the op stay cuDNN; bwarp and the deformable
convolution itself run in libvfi_b200.so (csrc/bwarp.cu, csrc/dcn.cu, csrc/dcn_tc.cu).
"""
from __future__ import annotations

from typing import Optional, Tuple, Union

import torch
import torch.nn as nn

from . import _lib
from .warp import bwarp

__all__ = ["deform_conv2d", "DeformableConv2d", "DeformableConv2dGivenOffset"]

_ALGOS = {"auto": 0, "simt": 1, "tcgen05": 2, "window": 3}


def _pair(v: Union[int, Tuple[int, int]]) -> int:
    if isinstance(v, (tuple, list)):
        assert v[0] == v[1], "vfi_b200 deform_conv2d: only square stride / padding / dilation"
        return int(v[0])
    return int(v)


class _DeformConvFn(torch.autograd.Function):
    """deform_conv2d with torchvision's backward (row f1): vfi_dcn_im2col_f32 / vfi_dcn_col2im_f32 around the two
    per-group GEMMs (cuBLAS through torch.einsum: a training-only path)."""

    @staticmethod
    def forward(ctx, x, offset, weight, bias, mask, s, p, d, algo):
        x, offset, weight = _lib.f32c(x), _lib.f32c(offset), _lib.f32c(weight)
        bias, mask = _lib.f32c(bias), _lib.f32c(mask)
        ctx.save_for_backward(x, offset, weight, mask)
        ctx.conf = (s, p, d, bias is not None)
        with torch.no_grad():
            return deform_conv2d(x, offset, weight, bias, s, p, d, mask, algo)

    @staticmethod
    def backward(ctx, grad_out):
        x, offset, weight, mask = ctx.saved_tensors
        s, p, d, has_bias = ctx.conf
        dev = x.device
        B, Cin, H, W = x.shape
        Cout, cin_g, kh, kw = weight.shape
        K, WG, OG = kh * kw, Cin // cin_g, offset.shape[1] // (2 * kh * kw)
        g = _lib.f32c(grad_out)
        Ho, Wo = g.shape[-2:]
        Po = Ho * Wo
        need_x, need_off, need_w, need_b, need_m = (ctx.needs_input_grad[i] for i in (0, 1, 2, 3, 4))
        st = _lib.stream_ptr(dev)
        g_w = None
        gg = g.view(B, WG, Cout // WG, Po)
        if need_w:
            cols = torch.empty((B, Cin, K, Po), dtype=torch.float32, device=dev)
            _lib.call("vfi_dcn_im2col_f32", dev, x.data_ptr(), offset.data_ptr(), _lib.ptr(mask), cols.data_ptr(), B, Cin, H, W,
                      kh, kw, s, p, d, OG, st)
            g_w = torch.einsum("bgop,bgkp->gok", gg, cols.view(B, WG, cin_g * K, Po)).reshape(Cout, cin_g, kh, kw)
            del cols
        g_x = g_off = g_m = None
        if need_x or need_off or (need_m and mask is not None):
            gcol = torch.einsum("gok,bgop->bgkp", weight.view(WG, Cout // WG, cin_g * K), gg).contiguous()
            g_x = torch.empty_like(x) if need_x else None
            g_off = torch.empty_like(offset) if need_off else None
            g_m = torch.empty_like(mask) if (need_m and mask is not None) else None
            _lib.call("vfi_dcn_col2im_f32", dev, x.data_ptr(), offset.data_ptr(), _lib.ptr(mask), gcol.data_ptr(), _lib.ptr(g_x),
                      _lib.ptr(g_off), _lib.ptr(g_m), B, Cin, H, W, kh, kw, s, p, d, OG, st)
        g_b = g.sum(dim=(0, 2, 3)) if (need_b and has_bias) else None
        return g_x, g_off, g_w, g_b, g_m, None, None, None, None


def deform_conv2d(input: torch.Tensor, offset: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor] = None,
                  stride=1, padding=0, dilation=1, mask: Optional[torch.Tensor] = None, algo: str = "auto") -> torch.Tensor:
    """Modulated deformable convolution with torchvision's exact semantics (groups inferred the same way).
    Differentiable wrt input, offset, weight, bias and mask."""
    if torch.is_grad_enabled() and any(t is not None and t.requires_grad for t in (input, offset, weight, bias, mask)):
        return _DeformConvFn.apply(input, offset, weight, bias, mask, _pair(stride), _pair(padding), _pair(dilation), algo)
    dev = _lib.require_cuda(input, offset, weight, bias, mask)
    x, offset, weight = _lib.planar(input), _lib.f32c(offset), _lib.f32c(weight)   # channels-last input: one transpose kernel
    bias, mask = _lib.f32c(bias), _lib.f32c(mask)
    B, Cin, H, W = x.shape
    Cout, cin_g, kh, kw = weight.shape
    s, p, d = _pair(stride), _pair(padding), _pair(dilation)
    assert Cin % cin_g == 0, "weight.shape[1] must divide the input channels"
    WG = Cin // cin_g
    assert offset.shape[1] % (2 * kh * kw) == 0, "offset.shape[1] must be a multiple of 2*kh*kw"
    OG = offset.shape[1] // (2 * kh * kw)
    Ho = (H + 2 * p - (d * (kh - 1) + 1)) // s + 1
    Wo = (W + 2 * p - (d * (kw - 1) + 1)) // s + 1
    assert tuple(offset.shape) == (B, OG * 2 * kh * kw, Ho, Wo), f"offset shape {tuple(offset.shape)}"
    if mask is not None:
        assert tuple(mask.shape) == (B, OG * kh * kw, Ho, Wo), f"mask shape {tuple(mask.shape)}"
    out = torch.empty((B, Cout, Ho, Wo), dtype=torch.float32, device=dev)
    L = _lib.lib()
    a = _ALGOS[algo]
    ws_bytes = int(L.vfi_dcn_workspace_bytes(B, Cin, Cout, H, W, kh, kw, WG, OG, a))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev) if ws_bytes else None
    _lib.call("vfi_dcn_f32", dev, x.data_ptr(), offset.data_ptr(), _lib.ptr(mask), weight.data_ptr(), _lib.ptr(bias),
                             out.data_ptr(), B, Cin, Cout, H, W, kh, kw, s, p, d, WG, OG, a, _lib.ptr(ws), ws_bytes,
                             _lib.stream_ptr(dev))
    return out


def window_or_auto(x: torch.Tensor, weight: torch.Tensor, stride: int = 1) -> str:
    """'window' (csrc/dcn_win.cu) for the network-predicted offsets of the DCN blocks when the shape is covered and large
    enough to pay for the window staging, else 'auto'."""
    if (stride == 1 and x.dim() == 4 and x.shape[2] * x.shape[3] >= 4096 and x.shape[3] % 4 == 0 and weight.shape[2] <= 3
            and weight.shape[3] <= 3 and weight.shape[1] <= 16 and weight.shape[0] * weight.shape[1] // max(x.shape[1], 1) <= 16):
        return "window"
    return "auto"


class DeformableConv2d(nn.Module):
    """DCNv2 block: flow head -> bwarp -> offset/mask head -> modulated deformable conv.  Returns (out, flow)."""

    def __init__(self, in_channels, out_channels, kernel_size=3, stride=1, padding=1, groups=8):
        super().__init__()
        self.kernel_size, self.padding, self.stride, self.group = kernel_size, padding, stride, groups
        self.offset_flow_conv = nn.Conv2d(in_channels, 2, 3, 1, 1)
        self.conv_offset_mask = nn.Sequential(
            nn.Conv2d(in_channels * 2 + 2, in_channels, 3, 1, 1), nn.PReLU(in_channels),
            nn.Conv2d(in_channels, in_channels, 3, 1, 1), nn.PReLU(in_channels),
            nn.Conv2d(in_channels, groups * 3 * kernel_size * kernel_size, 3, 1, 1))
        for m in self.conv_offset_mask:              # zero-initialised head, modules/dcnv2.py:35,43-47
            if isinstance(m, nn.Conv2d):
                nn.init.zeros_(m.weight)
                nn.init.zeros_(m.bias)
        self.regular_conv = nn.Conv2d(in_channels // groups, out_channels, kernel_size, stride=1, padding=padding)

    def forward(self, x, movement_feat):
        if torch.is_grad_enabled():
            flow = self.offset_flow_conv(movement_feat)                           # B, 2, h, w
            warped = bwarp(x, flow)
            head = self.conv_offset_mask(torch.cat((warped, movement_feat, flow), dim=1))
        else:       # inference: conv + one fused bias / PReLU epilogue per layer (tensor-core kernel where eligible)
            from .blocks import conv_bias_act, run_conv_chain
            flow = conv_bias_act(self.offset_flow_conv, movement_feat)
            warped = bwarp(x, flow)
            head = run_conv_chain(self.conv_offset_mask, torch.cat((warped, movement_feat, flow), dim=1))
        n_off = head.shape[1] // 3 * 2
        res_offset = 2.0 * torch.tanh(head[:, :n_off])
        # base offset: the flow flipped to (dy, dx) and tiled over (group, tap) -- modules/dcnv2.py:57
        offset = res_offset + flow.flip(1).repeat(1, n_off // 2, 1, 1)
        mask = torch.sigmoid(head[:, n_off:])
        out = deform_conv2d(x, offset, self.regular_conv.weight, self.regular_conv.bias, stride=self.stride,
                            padding=self.padding, mask=mask)
        return out, flow


class DeformableConv2dGivenOffset(nn.Module):
    """DCN whose offsets come from the DAT block: offset (B, nG*9, 2, h, w), mask (B, nG*9, h, w) pre-sigmoid."""

    def __init__(self, in_c, out_c, kernel_size=3, stride=1, padding=1, groups=8):
        super().__init__()
        self.kernel_size, self.padding, self.stride, self.group = kernel_size, padding, stride, groups
        self.regular_conv = nn.Conv2d(in_c // groups, out_c, kernel_size, stride=1, padding=padding)

    def forward(self, x, offset, mask):
        b, _, fh, fw = x.shape
        # same permute / flip / view as models/archive/DCNDAT.py:69 (channel order is the reference's, quirks included)
        offset = offset.permute(0, 2, 1, 3, 4).flip(1).contiguous().view(b, -1, fh, fw)
        return deform_conv2d(x, offset, self.regular_conv.weight, self.regular_conv.bias, stride=self.stride,
                             padding=self.padding, mask=torch.sigmoid(mask), algo=window_or_auto(x, self.regular_conv.weight))
