"""Dense conv stacks that surround the hot path (cuDNN through PyTorch; out of scope as kernels).

Only the parameter layout is shared with the reference (``modules/basic_blocks.py``, ``modules/cnn_encoders.py``,
``modules/generator.py``) so that its state dicts load unchanged.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


from . import _lib
from . import conv as _conv


def _fusable(x: torch.Tensor, conv: nn.Conv2d) -> bool:
    return (x.is_cuda and x.dtype == torch.float32 and not torch.is_grad_enabled() and conv.bias is not None
            and conv.padding_mode == "zeros")


def conv_bias_act(conv: nn.Conv2d, x: torch.Tensor, prelu: nn.PReLU = None, residual: torch.Tensor = None,
                  residual_after_act: bool = False) -> torch.Tensor:
    """cuDNN convolution (no bias) + ONE fused epilogue kernel: bias, optional residual, optional per-channel PReLU
    (vfi_bias_act_f32).  Same arithmetic order as ``prelu(residual + conv(x))`` / ``residual + conv(x)`` in PyTorch, which
    runs bias add, PReLU and residual add as three strided elementwise kernels.  Training (grad enabled) and CPU
    tensors take the plain PyTorch route."""
    if _conv.eligible(conv, x):      # tcgen05 implicit GEMM with the epilogue fused (TF32, as cuDNN's default)
        return _conv.conv3x3(conv, x, None if prelu is None else prelu.weight, residual, residual_after_act)
    cin = conv.in_channels
    if cin == 3 and x.dim() == 4 and x.shape[1] == 4 and x.is_cuda and not torch.is_grad_enabled():
        # RGB0 channels-last frames from vfi_frame_prep: already in the padded layout the kernel reads
        if _conv.part_eligible(conv, x):
            return _conv.conv3x3(conv, x, None if prelu is None else prelu.weight, residual, residual_after_act,
                                 in_slice=(0, 3, 1))
        x = x[:, :3]
    if cin < 8 and cin % 4 and x.dim() == 4 and x.is_cuda and not torch.is_grad_enabled():
        # RGB-like inputs (the encoder's first conv): zero-pad the channels to 4 in NHWC so that the same kernel runs
        pad = 4 - cin % 4
        B, _, H, W = x.shape
        xp = torch.zeros((B, H, W, cin + pad), dtype=x.dtype, device=x.device).permute(0, 3, 1, 2)
        if _conv.part_eligible(conv, xp):
            xp[:, :cin] = x
            return _conv.conv3x3(conv, xp, None if prelu is None else prelu.weight, residual, residual_after_act,
                                 in_slice=(0, cin, pad))
    if not _fusable(x, conv):
        y = conv(x)
        if residual is not None and not residual_after_act:
            y = residual + y
        if prelu is not None:
            y = prelu(y)
        if residual is not None and residual_after_act:
            y = residual + y
        return y
    y = F.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)
    B, C, H, W = y.shape
    cl = y.is_contiguous(memory_format=torch.channels_last) and not y.is_contiguous()
    if not cl and not y.is_contiguous():
        y = y.contiguous()
    res = None
    if residual is not None:
        res = residual.contiguous(memory_format=torch.channels_last) if cl else residual.contiguous()
    slope = None
    if prelu is not None:
        slope = prelu.weight if prelu.weight.numel() == C else prelu.weight.expand(C).contiguous()
    mode = 0 if res is None else (2 if residual_after_act else 1)
    dev = y.device
    _lib.call("vfi_bias_act_f32", dev, y.data_ptr(), conv.bias.data_ptr(), _lib.ptr(slope), _lib.ptr(res), y.data_ptr(), B, C, H, W,
              1 if cl else 0, mode, _lib.stream_ptr(dev))
    return y


class ConvPReLU(nn.Sequential):
    """Conv2d + PReLU with the reference's Sequential key layout ('0' = conv, '1' = prelu); fused epilogue at inference."""

    def forward(self, x):
        return conv_bias_act(self[0], x, self[1])


def conv_prelu(cin: int, cout: int, stride: int = 1) -> nn.Sequential:
    return ConvPReLU(nn.Conv2d(cin, cout, 3, stride, 1), nn.PReLU(cout))


def run_conv_chain(seq: nn.Sequential, x: torch.Tensor) -> torch.Tensor:
    """Run a Sequential, fusing every Conv2d -> PReLU pair (and trailing bare Conv2d's bias) into conv + one epilogue."""
    mods = list(seq)
    i = 0
    while i < len(mods):
        m = mods[i]
        if isinstance(m, nn.Conv2d) and type(m) is nn.Conv2d:
            if i + 1 < len(mods) and isinstance(mods[i + 1], nn.PReLU):
                x = conv_bias_act(m, x, mods[i + 1])
                i += 2
                continue
            x = conv_bias_act(m, x)
        else:
            x = m(x)
        i += 1
    return x


class ResBlocknoBN(nn.Module):
    """x + conv2(prelu(conv1(x)))  (state-dict keys conv1 / conv2 / prelu)."""

    def __init__(self, nf: int = 64):
        super().__init__()
        self.conv1 = nn.Conv2d(nf, nf, 3, 1, 1)
        self.conv2 = nn.Conv2d(nf, nf, 3, 1, 1)
        self.prelu = nn.PReLU(nf)
        for conv in (self.conv1, self.conv2):      # kaiming * 0.1, zero bias (modules/basic_blocks.py:27)
            nn.init.kaiming_normal_(conv.weight, a=0, mode="fan_in")
            conv.weight.data.mul_(0.1)
            nn.init.zeros_(conv.bias)

    def forward(self, x):
        return conv_bias_act(self.conv2, conv_bias_act(self.conv1, x, self.prelu), residual=x)


class HalfChannelConv5ResBlock(nn.Module):
    """IFRNet-style 5-conv residual block whose even convs touch only the last ``side_c`` channels."""

    def __init__(self, in_c: int, side_c: int):
        super().__init__()
        self.side_c = side_c
        self.conv1, self.conv2 = conv_prelu(in_c, in_c), conv_prelu(side_c, side_c)
        self.conv3, self.conv4 = conv_prelu(in_c, in_c), conv_prelu(side_c, side_c)
        self.conv5 = nn.Conv2d(in_c, in_c, 3, 1, 1)
        self.prelu = nn.PReLU(in_c)

    def _side(self, t, conv):
        keep = t.shape[1] - self.side_c
        return torch.cat((t[:, :keep], conv(t[:, keep:])), dim=1)

    def _fused_ok(self, x):
        c, s = self.conv1[0].in_channels, self.side_c
        if not (x.dim() == 4 and c % 8 == 0 and s % 4 == 0 and (c - s) % 4 == 0 and _conv.eligible(self.conv1[0], x)):
            return False
        ctot = (c + s + 7) // 8 * 8
        ok = _lib.lib().vfi_conv3x3_supported
        return bool(ok(s, ctot, s, 1)) and bool(ok(c, ctot, c, 1))

    def _forward_no_cat(self, x):
        """Same five convs without the two concatenations: the full convs write their channels ROLLED (side half first)
        into a wider channels-last buffer [side | keep | side conv output], the side conv reads the first slice and
        writes the third, and the next full conv reads [keep | side conv output] in place."""
        B, c, H, W = x.shape
        s, keep = self.side_c, c - self.side_c
        ctot = (c + s + 7) // 8 * 8                     # pixel rows stay 32-byte aligned (256-bit stores)
        t = x
        for full, side in ((self.conv1, self.conv2), (self.conv3, self.conv4)):
            buf = torch.empty((B, H, W, ctot), dtype=torch.float32, device=x.device).permute(0, 3, 1, 2)
            _conv.conv3x3(full[0], t, prelu=full[1], out=buf[:, :c], roll=keep)
            _conv.conv3x3(side[0], buf[:, :s], prelu=side[1], out=buf[:, c:c + s])
            t = buf[:, s:c + s]
        return _conv.conv3x3(self.conv5, t, self.prelu.weight, residual=x)

    def forward(self, x):
        if self._fused_ok(x):
            return self._forward_no_cat(x)
        t = self._side(self.conv1(x), self.conv2)
        t = self._side(self.conv3(t), self.conv4)
        return conv_bias_act(self.conv5, t, self.prelu, residual=x)


class FeadForward(nn.Module):
    """1x1-conv MLP with GELU (keys fc1 / fc2; the reference's spelling of the class name is kept)."""

    def __init__(self, in_dim, hidden_dim=None, out_dim=None):
        super().__init__()
        hidden_dim, out_dim = int(hidden_dim or in_dim), int(out_dim or in_dim)
        self.fc1 = nn.Conv2d(in_dim, hidden_dim, 1)
        self.act = nn.GELU()
        self.fc2 = nn.Conv2d(hidden_dim, out_dim, 1)

    def forward(self, x):
        return self.fc2(self.act(self.fc1(x)))

    def forward_residual(self, x):
        """``x + fc2(gelu(fc1(x)))`` (the tail of modules/deformable_attn.py:130).  On channels-last CUDA tensors at
        inference the two 1x1 convolutions are plain GEMMs over the (pixels, C) view with the bias in the GEMM epilogue
        (cuBLASLt): cuDNN runs them as a conv plus a generic strided bias pass that costs more than the GEMM itself
        (measured at 544x960: 298 + 175 us -> 4 kernels, profiles/r1x_aten_ops.txt).  Same math mode as the cuDNN
        convs around it (TF32 iff cudnn.allow_tf32)."""
        cl = (x.dim() == 4 and x.is_cuda and x.dtype == torch.float32 and not torch.is_grad_enabled()
              and x.permute(0, 2, 3, 1).is_contiguous() and self.fc1.bias is not None and self.fc2.bias is not None)
        if not cl:
            return x + self.forward(x)
        B, C, H, W = x.shape
        x2 = x.permute(0, 2, 3, 1).reshape(B * H * W, C)
        w1 = self.fc1.weight.reshape(self.fc1.out_channels, C)
        w2 = self.fc2.weight.reshape(self.fc2.out_channels, self.fc1.out_channels)
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = bool(torch.backends.cudnn.allow_tf32)
        try:
            h = F.gelu(torch.addmm(self.fc1.bias, x2, w1.t()))
            y = torch.addmm(self.fc2.bias, h, w2.t())
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        y += x2
        return y.view(B, H, W, -1).permute(0, 3, 1, 2)


class SameChannelResEncoder(nn.Module):
    """4-level feature pyramid at 1/2, 1/4, 1/8, 1/16 resolution, nf channels each."""

    def __init__(self, nf: int, n_res_block: int):
        super().__init__()
        self.projection = nn.Sequential(nn.Conv2d(3, nf, 3, 1, 1), nn.PReLU(nf), nn.Conv2d(nf, nf, 3, 2, 1), nn.PReLU(nf),
                                        *[ResBlocknoBN(nf) for _ in range(n_res_block)])

        def down():
            return nn.Sequential(nn.Conv2d(nf, nf, 3, 2, 1), nn.PReLU(nf), nn.Conv2d(nf, nf, 3, 1, 1), nn.PReLU(nf))
        self.fea_L2_conv, self.fea_L3_conv, self.fea_L4_conv = down(), down(), down()

    def forward(self, x):
        f1 = run_conv_chain(self.projection, x)
        f2 = run_conv_chain(self.fea_L2_conv, f1)
        f3 = run_conv_chain(self.fea_L3_conv, f2)
        return f1, f2, f3, run_conv_chain(self.fea_L4_conv, f3)


class BasicResPixelShuffleGenerator(nn.Module):
    """Residual trunk -> pixel-shuffle x2 -> RGB, clamp(out + mean, 0, 1)."""

    def __init__(self, nf: int, num_res_blocks: int):
        super().__init__()
        self.nf = nf
        self.reconstruction = nn.Sequential(*[ResBlocknoBN(nf) for _ in range(num_res_blocks)])
        self.upconv1 = nn.Conv2d(nf, nf * 4, 3, 1, 1)
        self.prelu1 = nn.PReLU(nf)
        self.pixel_shuffle = nn.PixelShuffle(2)
        self.HRconv = nn.Conv2d(nf, nf, 3, 1, 1)
        self.prelu2 = nn.PReLU(nf)
        self.conv_last = nn.Conv2d(nf, 3, 3, 1, 1)

    def forward(self, feat, m):
        """``m`` = the rgb mean to add back before the clamp; ``m is None`` returns the un-shifted, un-clamped frame
        (the fused frame edge, vfi_frame_finish, applies mean + clamp + crop in its own pass)."""
        r = self.reconstruction(feat)
        if _conv.eligible(self.upconv1, r, upscale=2):     # conv + PixelShuffle(2) + PReLU in one kernel
            t = _conv.conv3x3(self.upconv1, r, self.prelu1.weight, upscale=2)
        else:
            t = self.prelu1(self.pixel_shuffle(conv_bias_act(self.upconv1, r)))
        y = conv_bias_act(self.conv_last, conv_bias_act(self.HRconv, t, self.prelu2))
        return y if m is None else torch.clamp(y + m, 0, 1)


def resize(x, scale_factor):
    """``BaseModel.resize`` (models/BaseModel.py:19-22)."""
    return F.interpolate(x, scale_factor=scale_factor, recompute_scale_factor=False, mode="bilinear", align_corners=True)


def flow_up2_add(res: torch.Tensor, ft: torch.Tensor, mul: float = 2.0) -> torch.Tensor:
    """``res + mul * resize(ft, scale_factor=2.0)`` (modules/deformable_attn.py:127-131) in one pass (vfi_resize2x_f32);
    frame-extent aware, so a tile of a larger frame upsamples exactly as the whole frame would."""
    if not (ft.is_cuda and ft.dtype == torch.float32 and res.dtype == torch.float32 and not torch.is_grad_enabled()):
        return res + mul * resize(ft, scale_factor=2.0)
    B, C, H, W = ft.shape
    assert tuple(res.shape) == (B, C, 2 * H, 2 * W), "flow_up2_add: res must be (B, C, 2H, 2W)"
    dev = ft.device
    out = torch.empty((B, C, 2 * H, 2 * W), dtype=torch.float32, device=dev)
    ext, _, _ = _lib.frame_ext(H, W, dev)
    _lib.call("vfi_resize2x_f32", dev, ft.data_ptr(), _lib.strides4(ft), res.data_ptr(), _lib.strides4(res), float(mul),
              out.data_ptr(), B, C, H, W, ext, _lib.stream_ptr(dev))
    return out
