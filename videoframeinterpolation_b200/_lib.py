"""ctypes binding of libvfi_b200.so (C ABI: include/vfi_b200.h).

There is NO fallback: if the shared library is missing or a tensor is not on a CUDA device the
call raises (the reference's splat likewise refuses CPU tensors, modules/softsplat.py:352-353).
PyTorch is used only for device memory and the current stream.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_float, c_int, c_int64, c_size_t, c_void_p
from pathlib import Path
from typing import Dict, Optional, Tuple

import torch

_PKG = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("VFI_B200_LIB", _PKG / "libvfi_b200.so"))

_I64x4 = c_int64 * 4
_lib: Optional[ctypes.CDLL] = None

# name -> (restype, argtypes); every entry is declared in include/vfi_b200.h
SIGNATURES = {
    "vfi_version": (c_int, []),
    "vfi_last_error_string": (ctypes.c_char_p, []),
    "vfi_device_supported": (c_int, []),
    "vfi_bwarp_f32": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p] * 3 + [c_void_p, c_void_p]),
    "vfi_bwarp_flowcat_f32": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p] * 3 + [c_void_p, c_void_p]),
    "vfi_splat_f32": (c_int, [c_void_p] * 6 + [c_int] * 4 + [c_int, c_int, c_void_p, c_size_t, c_void_p]),
    "vfi_splat_workspace_bytes": (c_size_t, [c_int] * 4),
    "vfi_splat_normalize_f32": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_int, c_float, c_float, c_void_p]),
    "vfi_splat_last_f32": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_int, c_int, c_float, c_float, c_void_p, c_size_t, c_void_p]),
    "vfi_splat_fused_f32": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_int, c_int, c_float, c_void_p, c_size_t, c_void_p]),
    "vfi_dsample_f32": (c_int, [c_void_p] * 6 + [c_int] * 6 + [c_void_p, c_void_p]),
    "vfi_attn_f32": (c_int, [c_void_p] * 11 + [c_int] * 6 + [c_void_p]),
    "vfi_dattn_f32": (c_int, [c_void_p] * 19 + [c_int] * 7 + [c_float, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "vfi_dattn_win_covers": (c_int, [c_int, c_int, c_int, c_float, c_int, c_int]),
    "vfi_dcn_f32": (c_int, [c_void_p] * 6 + [c_int] * 13 + [c_void_p, c_size_t, c_void_p]),
    "vfi_dcn_workspace_bytes": (c_size_t, [c_int] * 10),
    "vfi_bias_act_f32": (c_int, [c_void_p] * 5 + [c_int] * 6 + [c_void_p]),
    "vfi_nhwc_to_nchw_f32": (c_int, [c_void_p] * 2 + [c_int] * 4 + [c_int64, c_void_p]),
    "vfi_conv3x3_supported": (c_int, [c_int] * 4),
    "vfi_conv3x3_prep_bytes": (c_size_t, [c_int] * 3),
    "vfi_conv3x3_prep_f32": (c_int, [c_void_p] * 2 + [c_int] * 3 + [c_void_p]),
    "vfi_conv3x3_f32": (c_int, [c_void_p] * 6 + [c_int] * 11 + [c_void_p]),
    "vfi_splat_bwd_f32": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "vfi_bwarp_bwd_f32": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p]),
    "vfi_dsample_bwd_f32": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_void_p]),
    "vfi_dsample_rows_f32": (c_int, [c_void_p] * 6 + [c_int] * 8 + [c_void_p, c_void_p]),
    "vfi_dsample_rows_bwd_f32": (c_int, [c_void_p] * 7 + [c_int] * 9 + [c_void_p]),
    "vfi_frame_prep_workspace_bytes": (c_size_t, [c_int]),
    "vfi_frame_prep": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p] + [c_int] * 7 + [c_void_p, c_size_t, c_void_p]),
    "vfi_frame_finish": (c_int, [c_void_p] * 4 + [c_int] * 8 + [c_void_p]),
    "vfi_dcn_im2col_f32": (c_int, [c_void_p] * 4 + [c_int] * 10 + [c_void_p]),
    "vfi_dcn_col2im_f32": (c_int, [c_void_p] * 7 + [c_int] * 10 + [c_void_p]),
    "vfi_dcndat_coords_f32": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p]),
    "vfi_frame_metrics_workspace_bytes": (c_size_t, [c_int] * 4),
    "vfi_frame_metrics_f32": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_float, c_void_p, c_size_t, c_void_p]),
    "vfi_frame_metrics_static_max_bytes": (c_size_t, []),
    "vfi_frame_metrics_static_f32": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_float, c_void_p]),
    "vfi_resize2x_f32": (c_int, [c_void_p] * 4 + [c_float, c_void_p] + [c_int] * 4 + [c_void_p, c_void_p]),
    "vfi_launch_count": (c_int64, []),
    "vfi_launch_count_reset": (None, []),
}


class VfiError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load (once) and return the shared library; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise VfiError(f"{LIB_PATH} not found: build it with `python -m videoframeinterpolation_b200.build` "
                           "(there is no CPU or PyTorch fallback)")
        handle = ctypes.CDLL(str(LIB_PATH))
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(status: int, what: str) -> None:
    if status != 0:
        raise VfiError(f"{what} failed with status {status}: {lib().vfi_last_error_string().decode()}")


def require_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise VfiError("vfi_b200 ops need CUDA tensors (no CPU fallback)")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise VfiError(f"tensors on different devices: {dev} vs {t.device}")
    if dev is None:
        raise VfiError("no tensor given")
    return dev


def f32c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """fp32 + contiguous (the splat reference also forces fp32: modules/softsplat.py:284)."""
    if t is None:
        return None
    if t.dtype != torch.float32:
        t = t.float()
    return t if t.is_contiguous() else t.contiguous()


def planar(t: torch.Tensor) -> torch.Tensor:
    """fp32 planar NCHW copy of ``t``; a channels-last CUDA tensor (or a channel slice of one) goes through
    vfi_nhwc_to_nchw_f32 instead of PyTorch's generic strided copy."""
    if (t.dim() == 4 and t.is_cuda and t.dtype == torch.float32 and not t.is_contiguous() and not torch.is_grad_enabled()
            and t.shape[1] > 1 and t.shape[1] <= 383 and t.shape[0] <= 65535):
        B, C, H, W = t.shape
        sb, sc, sh, sw = t.stride()
        if sc == 1 and sw >= C and sh == W * sw and sb == H * W * sw:
            out = torch.empty((B, C, H, W), dtype=torch.float32, device=t.device)
            call("vfi_nhwc_to_nchw_f32", t.device, t.data_ptr(), out.data_ptr(), B, C, H, W, sw, stream_ptr(t.device))
            return out
    return f32c(t)


def f32(t: torch.Tensor) -> torch.Tensor:
    return t if t.dtype == torch.float32 else t.float()


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def strides4(t: torch.Tensor):
    return _I64x4(*t.stride())


def stream_ptr(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


_lin_cache: Dict[Tuple[int, str], torch.Tensor] = {}


def linspace_m11(n: int, dev: torch.device) -> torch.Tensor:
    """torch.linspace(-1, 1, n) built on the CPU exactly as the reference does (modules/warp.py:9-10),
    cached on the device so the per-call H2D copy of the reference disappears."""
    key = (n, str(dev))
    t = _lin_cache.get(key)
    if t is None:
        t = torch.linspace(-1.0, 1.0, n).to(dev)
        _lin_cache[key] = t
    return t


# ---- tiles of a larger frame (BASELINE configs[4]) ---------------------------------------------------
# While a crop context is active, every frame-normalised op (bwarp, the samplers, the flow resize) is told the extent
# of the WHOLE frame and where its tensors sit in it, so a tile samples exactly where the whole frame would.
_crop: Optional[Tuple[int, int, int, int, int, int]] = None     # (Hg, Wg, y0, x0, Hc, Wc) at full resolution
_I32x4 = c_int * 4


class crop_context:
    """``with crop_context(Hg, Wg, y0, x0, Hc, Wc): net._generate_frame(crop0, crop1, t, mean_=...)`` -- the tensors of
    the forward are the (Hc, Wc) crop at (y0, x0) of an (Hg, Wg) frame (all multiples of 16: every pyramid level then
    has integer extents)."""

    def __init__(self, Hg, Wg, y0, x0, Hc, Wc):
        for v in (Hg, Wg, y0, x0, Hc, Wc):
            assert v % 16 == 0, "crop contexts need frame, origin and crop sizes that are multiples of 16"
        assert 0 <= y0 and y0 + Hc <= Hg and 0 <= x0 and x0 + Wc <= Wg, "crop outside the frame"
        self.ctx = (Hg, Wg, y0, x0, Hc, Wc)

    def __enter__(self):
        global _crop
        self._prev, _crop = _crop, self.ctx
        return self

    def __exit__(self, *exc):
        global _crop
        _crop = self._prev
        return False


def frame_ext(H: int, W: int, dev: torch.device):
    """-> (frame_ext argument of the C ABI or None, lin_x, lin_y) for an (H, W) tensor of the current forward."""
    if _crop is None:
        return None, linspace_m11(W, dev), linspace_m11(H, dev)
    Hg, Wg, y0, x0, Hc, Wc = _crop
    s = Hc // H
    assert s >= 1 and H * s == Hc and W * s == Wc and s in (1, 2, 4, 8, 16), f"({H}, {W}) is not a level of the ({Hc}, {Wc}) crop"
    ext = _I32x4(Wg // s, Hg // s, x0 // s, y0 // s)
    return ext, linspace_m11(Wg // s, dev)[x0 // s:x0 // s + W], linspace_m11(Hg // s, dev)[y0 // s:y0 // s + H]


def launch_count() -> int:
    return int(lib().vfi_launch_count())


def reset_launch_count() -> None:
    lib().vfi_launch_count_reset()


class OpTimer:
    """Collects CUDA-event timings of every C-ABI call made while active (bench.py's per-kernel numbers).

        with OpTimer() as tm: ...   ;   torch.cuda.synchronize()   ;   tm.summary() -> {name: (calls, total_ms)}
    Events are recorded on the stream the call is launched on.
    """

    def __init__(self):
        self.records = {}

    def __enter__(self):
        global _timer
        self._prev, _timer = _timer, self
        return self

    def __exit__(self, *exc):
        global _timer
        _timer = self._prev
        return False

    def summary(self):
        out = {}
        for name, evs in self.records.items():
            out[name] = (len(evs), sum(a.elapsed_time(b) for a, b in evs))
        return out


_timer: Optional[OpTimer] = None


def call(name: str, dev: torch.device, *args, tag: Optional[str] = None) -> None:
    """Invoke one C-ABI entry point on the current stream of ``dev`` and raise on a non-zero status."""
    fn = getattr(lib(), name)
    if _timer is None:
        check(fn(*args), name)
        return
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stream = torch.cuda.current_stream(dev)
    start.record(stream)
    status = fn(*args)
    end.record(stream)
    check(status, name)
    _timer.records.setdefault(tag or name, []).append((start, end))
