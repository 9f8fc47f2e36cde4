"""3x3 / stride-1 convolutions of the DAT conv stacks on the tcgen05 tensor cores (``csrc/conv3x3.cu``; SURVEY.md 8f
row f2).  Used by ``blocks.conv_bias_act`` for channels-last fp32 CUDA tensors at inference when PyTorch's own switch
``torch.backends.cudnn.allow_tf32`` permits TF32 convolutions (its default, and therefore the arithmetic the reference
runs on a GPU); everything else stays on cuDNN.  ``VFI_CONV_ALGO=cudnn`` disables it, ``=tc`` makes ineligible calls
raise instead of falling back to cuDNN.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn

from . import _lib


_flops = [0]     # multiply-add flops issued through conv3x3 since the last reset (bench.py's tensor-roofline object)


def flop_count(reset: bool = False) -> int:
    n = _flops[0]
    if reset:
        _flops[0] = 0
    return n


def _algo() -> str:
    return os.environ.get("VFI_CONV_ALGO", "auto")


def _nhwc_pixel_stride(x: torch.Tensor) -> Optional[int]:
    """Pixel stride (floats) when x (B,C,H,W) is NHWC-dense or a channel slice of an NHWC tensor, else None."""
    if x.dim() != 4:
        return None
    B, C, H, W = x.shape
    sb, sc, sh, sw = x.stride()
    if C > 1 and sc != 1:
        return None
    if sw < C or (H > 1 and sh != W * sw) or (B > 1 and sb != H * W * sw):
        return None
    return sw


def eligible(conv: nn.Conv2d, x: torch.Tensor, upscale: int = 1) -> bool:
    if _algo() == "cudnn" or not torch.backends.cudnn.allow_tf32:
        return False
    if not (type(conv) is nn.Conv2d and conv.kernel_size == (3, 3) and conv.stride == (1, 1) and conv.padding == (1, 1)
            and conv.dilation == (1, 1) and conv.groups == 1 and conv.padding_mode == "zeros"):
        return False
    if not (x.is_cuda and x.dtype == torch.float32 and not torch.is_grad_enabled() and x.data_ptr() % 16 == 0):
        return False
    ps = _nhwc_pixel_stride(x)
    if ps is None:
        return False
    return bool(_lib.lib().vfi_conv3x3_supported(conv.in_channels, ps, conv.out_channels, upscale))


def prepared_weight(conv: nn.Conv2d, upscale: int = 1, roll: int = 0) -> torch.Tensor:
    """TF32-rounded, UMMA-ordered copy of conv.weight, cached on the module and rebuilt when the weight changes.
    ``roll``: output channel j of the prepared convolution is channel (j + roll) % Cout of ``conv``."""
    w = conv.weight
    key = (w.data_ptr(), w._version, str(w.device), upscale, roll)
    cache = getattr(conv, "_vfi_wprep", None)
    if cache is None or cache[0] != key[:3]:
        cache = (key[:3], {})
        object.__setattr__(conv, "_vfi_wprep", cache)
    if key in cache[1]:
        return cache[1][key]
    cout, cin = w.shape[:2]
    nbytes = _lib.lib().vfi_conv3x3_prep_bytes(cout, cin, upscale)
    wp = torch.empty(nbytes // 4, dtype=torch.float32, device=w.device)
    wc = w.detach()
    if roll:
        wc = torch.roll(wc, -roll, 0)
    wc = wc.contiguous()                  # (Cout, Cin, 3, 3) row-major whatever the module's memory format
    _lib.call("vfi_conv3x3_prep_f32", w.device, wc.data_ptr(), wp.data_ptr(), cout, cin, upscale, _lib.stream_ptr(w.device))
    cache[1][key] = wp
    return wp


_CACHE_ATTRS = ("_vfi_wprep", "_vfi_wprep_slices", "_vfi_as_conv", "_vfi_bias_rolled", "_vfi_slope_rolled")


def invalidate_prepared(model: nn.Module) -> None:
    """Drop every prepared-weight cache under ``model`` (TF32-packed weights, ConvTranspose-as-conv modules, rolled bias /
    slope vectors).  The caches are keyed on (data_ptr, _version); ``param.copy_`` / ``load_state_dict`` / optimizer steps
    bump the version, but writes through ``.data`` (``p.data.copy_(ema)``, ``m.weight.data.zero_()`` as the reference's own
    ``_init_zero`` does, modules/deformable_attn.py:72-75) do NOT -- call this after such writes."""
    for m in model.modules():
        for a in _CACHE_ATTRS:
            if a in m.__dict__:
                object.__delattr__(m, a)


def install_invalidation_hooks(model: nn.Module) -> nn.Module:
    """Clear the prepared-weight caches whenever ``model.load_state_dict`` runs (belt and braces: it uses ``copy_``, which
    already bumps the versions the caches are keyed on)."""
    model.register_load_state_dict_post_hook(lambda module, incompatible: invalidate_prepared(module))
    return model


def _rolled(owner: nn.Module, name: str, v: torch.Tensor, roll: int) -> torch.Tensor:
    """torch.roll(v, -roll) cached on ``owner`` (per-channel bias / PReLU slope of a conv whose outputs are rolled)."""
    key = (v.data_ptr(), v._version, str(v.device), roll)
    cached = getattr(owner, name, None)
    if cached is not None and cached[0] == key:
        return cached[1]
    r = torch.roll(v.detach(), -roll, 0).contiguous()
    object.__setattr__(owner, name, (key, r))
    return r


def prepared_weight_slice(conv: nn.Conv2d, lo: int, hi: int, pad: int = 0) -> torch.Tensor:
    """Prepared weights of the sub-convolution that sees input channels [lo, hi) of ``conv`` (plus ``pad`` zero
    channels): a convolution is linear in its input channels, so conv(cat(a, b)) = conv_a(a) + conv_b(b)."""
    w = conv.weight
    key = (w.data_ptr(), w._version, str(w.device), lo, hi, pad)
    cache = getattr(conv, "_vfi_wprep_slices", None)
    if cache is None:
        cache = {}
        object.__setattr__(conv, "_vfi_wprep_slices", cache)
    if key in cache:
        return cache[key]
    ws = w.detach()[:, lo:hi]
    if pad:
        ws = torch.nn.functional.pad(ws, (0, 0, 0, 0, 0, pad))
    ws = ws.contiguous()
    cout, cin = ws.shape[:2]
    wp = torch.empty(_lib.lib().vfi_conv3x3_prep_bytes(cout, cin, 1) // 4, dtype=torch.float32, device=w.device)
    _lib.call("vfi_conv3x3_prep_f32", w.device, ws.data_ptr(), wp.data_ptr(), cout, cin, 1, _lib.stream_ptr(w.device))
    for k in [k for k in cache if k[:3] != key[:3]]:      # the weight changed: drop stale slices
        del cache[k]
    cache[key] = wp
    return wp


def part_eligible(conv: nn.Conv2d, x: torch.Tensor) -> bool:
    """:func:`eligible` for a channel-slice pass (x holds only some of conv's input channels)."""
    if _algo() == "cudnn" or not torch.backends.cudnn.allow_tf32:
        return False
    if not (type(conv) is nn.Conv2d and conv.kernel_size == (3, 3) and conv.stride == (1, 1) and conv.padding == (1, 1)
            and conv.dilation == (1, 1) and conv.groups == 1 and conv.padding_mode == "zeros"):
        return False
    if not (x.is_cuda and x.dtype == torch.float32 and not torch.is_grad_enabled() and x.data_ptr() % 16 == 0):
        return False
    ps = _nhwc_pixel_stride(x)
    return ps is not None and bool(_lib.lib().vfi_conv3x3_supported(x.shape[1], ps, conv.out_channels, 1))


def conv3x3(conv: nn.Conv2d, x: torch.Tensor, prelu_weight: Optional[torch.Tensor] = None,
            residual: Optional[torch.Tensor] = None, residual_after_act: bool = False, use_bias: bool = True,
            out_channels_last: bool = True, upscale: int = 1, in_slice=None, out: Optional[torch.Tensor] = None,
            roll: int = 0, prelu: Optional[nn.Module] = None) -> torch.Tensor:
    """``act(residual + conv(x) + bias)`` in one kernel; x must satisfy :func:`eligible`.  ``upscale=2`` returns
    ``act(pixel_shuffle(conv(x) + bias, 2))`` (prelu_weight then has Cout/4 entries).  ``in_slice=(lo, hi, pad)``: x
    holds input channels [lo, hi) of conv followed by ``pad`` zero-weighted channels (one term of a convolution split
    over its input channels; add the terms through ``residual``).  ``residual`` may have fewer images than x (B % Br == 0):
    image b adds residual image b % Br (a term shared by both frames of a pair batch).  ``out``: a (B, Cout, H, W) channels-last view (e.g.
    a channel slice of a wider buffer) to write into.  ``roll``: output channel j is channel (j + roll) % Cout of
    ``conv`` (bias and the slopes of ``prelu`` are rolled with it; no residual)."""
    dev = _lib.require_cuda(x, residual)
    B, Cin, H, W = x.shape
    Cout = conv.out_channels
    ps = _nhwc_pixel_stride(x)
    if ps is None:
        raise _lib.VfiError("conv3x3: x must be channels-last (or a channel slice of a channels-last tensor)")
    if in_slice is not None:
        lo, hi, pad = in_slice
        assert upscale == 1 and roll == 0 and Cin == hi - lo + pad, "in_slice does not match x"
        wp = prepared_weight_slice(conv, lo, hi, pad)
    else:
        wp = prepared_weight(conv, upscale, roll)
    Cs, Ho, Wo = Cout // (upscale * upscale), H * upscale, W * upscale
    out_c = Cs
    if out is not None:
        out_c = _nhwc_pixel_stride(out)
        assert out_channels_last and out_c is not None and tuple(out.shape) == (B, Cs, Ho, Wo) and out.dtype == torch.float32, \
            "out must be a channels-last float32 view of the result's shape"
    elif out_channels_last:
        out = torch.empty((B, Ho, Wo, Cs), dtype=torch.float32, device=dev).permute(0, 3, 1, 2)
    else:
        out = torch.empty((B, Cs, Ho, Wo), dtype=torch.float32, device=dev)
    res, res_c, mode = None, Cout, 0
    if residual is not None:
        if out_channels_last:
            rps = _nhwc_pixel_stride(residual)
            res = residual if rps is not None else residual.contiguous(memory_format=torch.channels_last)
            res_c = rps if rps is not None else Cout
        else:
            res = residual.contiguous()
        mode = 2 if residual_after_act else 1
        if residual.shape[0] != B:      # one residual shared by groups of the batch: image b adds residual image b % Br
            assert B % residual.shape[0] == 0 and tuple(residual.shape[1:]) == (Cout, H, W), "residual batch must divide B"
            mode |= residual.shape[0] << 8
    if prelu is not None:
        prelu_weight = prelu.weight
    slope = None
    if prelu_weight is not None:
        slope = prelu_weight if prelu_weight.numel() == Cs else prelu_weight.expand(Cs).contiguous()
    bias = conv.bias if use_bias else None
    if roll:
        assert residual is None and upscale == 1, "roll: no residual / pixel shuffle"
        if bias is not None:
            bias = _rolled(conv, "_vfi_bias_rolled", bias, roll)
        if slope is not None and prelu_weight.numel() == Cs:
            slope = _rolled(prelu if prelu is not None else conv, "_vfi_slope_rolled", slope, roll)
    _flops[0] += 2 * 9 * (Cin if in_slice is None else in_slice[1] - in_slice[0]) * Cout * B * H * W
    tag = None
    if _lib._timer is not None and getattr(_lib._timer, "detailed", False):     # per-shape rows for tools/conv_table.py
        tag = f"conv3x3 {H}x{W} {Cin}->{Cout} res{mode & 3} up{upscale} ps{ps}" + (f" B{B}" if B > 1 else "")
    _lib.call("vfi_conv3x3_f32", dev, x.data_ptr(), wp.data_ptr(), _lib.ptr(bias), _lib.ptr(slope), _lib.ptr(res), out.data_ptr(),
              B, H, W, Cin, ps, Cout, 1 if out_channels_last else 0, out_c, res_c, mode, upscale, _lib.stream_ptr(dev), tag=tag)
    return out


# ---- ConvTranspose2d(k=4, s=2, p=1) as a 3x3 convolution + PixelShuffle(2) ---------------------------------------------
# out[2Y + i] gets in[y] * w[ky] with ky = 2 (Y - y) + i + 1, i.e. sub-pixel i = 0 sees in[Y - 1] * w[3] + in[Y] * w[1] and
# i = 1 sees in[Y] * w[2] + in[Y + 1] * w[0]: a 3-tap window around Y with one zero tap per sub-pixel (same in x).
_CT_TAPS = {0: ((0, 3), (1, 1)), 1: ((1, 2), (2, 0))}      # sub-pixel -> ((conv tap, transposed-conv tap), ...)


def _is_up2_transpose(ct: nn.Module) -> bool:
    return (type(ct) is nn.ConvTranspose2d and ct.kernel_size == (4, 4) and ct.stride == (2, 2) and ct.padding == (1, 1)
            and ct.output_padding == (0, 0) and ct.dilation == (1, 1) and ct.groups == 1)


def transpose_as_conv(ct: nn.ConvTranspose2d) -> nn.Conv2d:
    """The Conv2d(Cin, 4 Cout, 3, 1, 1) whose PixelShuffle(2) equals ``ct`` (models/DAT.py:29-36, conv_res_flow of
    modules/deformable_attn.py:66); cached on ``ct`` and rebuilt when its parameters change."""
    w, b = ct.weight, ct.bias
    key = (w.data_ptr(), w._version, str(w.device), None if b is None else (b.data_ptr(), b._version))
    cached = getattr(ct, "_vfi_as_conv", None)
    if cached is not None and cached[0] == key:
        return cached[1]
    cin, cout = w.shape[:2]
    with torch.no_grad():
        w3 = torch.zeros((cout, 2, 2, cin, 3, 3), dtype=w.dtype, device=w.device)
        for i, ytaps in _CT_TAPS.items():
            for j, xtaps in _CT_TAPS.items():
                for ty, ky in ytaps:
                    for tx, kx in xtaps:
                        w3[:, i, j, :, ty, tx] = w[:, :, ky, kx].t()
        conv = nn.Conv2d(cin, 4 * cout, 3, 1, 1, bias=b is not None).to(device=w.device, dtype=w.dtype)
        conv.weight.copy_(w3.reshape(4 * cout, cin, 3, 3))
        if b is not None:
            conv.bias.copy_(b.repeat_interleave(4))
        conv.requires_grad_(False)
    object.__setattr__(ct, "_vfi_as_conv", (key, conv))
    return conv


def conv_transpose_up2(ct: nn.ConvTranspose2d, x: torch.Tensor, out_channels_last: bool = True) -> torch.Tensor:
    """``ct(x)`` on the tensor-core kernel (3x3 conv with the fused PixelShuffle) when eligible, else PyTorch's.
    Below ~64k input pixels the kernel's fixed cost (resident weights per launch, one launch per sub-pixel) loses to
    cuDNN (measured: 76 -> 76 at 68x120 75 vs 30 us, 72 -> 72 at 136x240 59 vs 65 us, at 272x480 122 vs 240-470 us)."""
    if _is_up2_transpose(ct) and _algo() != "cudnn" and x.dim() == 4 and x.shape[2] * x.shape[3] >= 65536:
        conv = transpose_as_conv(ct)
        if 4 * ct.out_channels <= 72 and eligible(conv, x):
            # few output channels (the 2-channel flow heads): all four sub-pixels in ONE pass, shuffle afterwards
            y = torch.nn.functional.pixel_shuffle(conv3x3(conv, x), 2)
            return y if out_channels_last else y.contiguous()
        if eligible(conv, x, upscale=2):
            return conv3x3(conv, x, out_channels_last=out_channels_last, upscale=2)
    return ct(x)
