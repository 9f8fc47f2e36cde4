"""The caller of the hot path: ``DATwConstantnCv1`` (reference: models/DAT.py:13-119, models/BaseModel.py).

Same constructor argument (an ``args`` object with nf / enc_res_blocks / dec_res_blocks / mlp_ratio), same
parameter names (reference state dicts load with strict=True), same ``inference(x0, x1, t)`` convention
(evaluate.py:104).  Loss / logging halves of the reference class are out of scope.
"""
from __future__ import annotations

from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

from .blocks import BasicResPixelShuffleGenerator, SameChannelResEncoder, run_conv_chain
from . import conv as _conv
from .conv import conv_transpose_up2
from .dcnv2 import DeformableConv2d
from .deformable_attn import CrossDeformableAttentionBlockwFlow

__all__ = ["DATwConstantnCv1", "DCNInterFeatBuilderwithT", "InputPadder", "default_args"]


def default_args(**over):
    """configs/DAT.yaml:4-11."""
    d = dict(nf=72, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2, distill_lambda=0.01)
    d.update(over)
    return SimpleNamespace(**d)


def _batch_pair(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """(2B, ...) batch of a then b: a view when b already follows a in memory with the same strides (the two halves of one
    allocation, as ``interpolate`` makes them), else a copy."""
    if (a.shape == b.shape and a.stride() == b.stride() and a.dtype == b.dtype and a.dim() == 4 and a.shape[0] >= 1
            and a.untyped_storage().data_ptr() == b.untyped_storage().data_ptr()
            and b.storage_offset() == a.storage_offset() + a.shape[0] * a.stride(0)):
        return a.as_strided((2 * a.shape[0],) + tuple(a.shape[1:]), a.stride(), a.storage_offset())
    return torch.cat((a, b), dim=0)


class DCNInterFeatBuilderwithT(nn.Module):
    """Coarsest-level query builder: two DCN blocks blended, t / 1-t planes appended (modules/query_builder.py:7-36)."""

    def __init__(self, nc):
        super().__init__()
        self.nc = nc
        self.convblock = nn.Sequential(nn.Conv2d(nc * 2 + 1, nc, 3, 1, 1), nn.PReLU(nc), nn.Conv2d(nc, nc, 3, 1, 1), nn.PReLU(nc))
        self.dcnt0 = DeformableConv2d(nc, nc)
        self.dcnt1 = DeformableConv2d(nc, nc)
        self.blendblock = nn.Sequential(nn.Conv2d(nc * 2, nc, 3, 1, 1), nn.PReLU(nc), nn.Conv2d(nc, nc, 3, 1, 1))

    def forward(self, feat0, feat1, t):
        _, _, fh, fw = feat0.shape
        tt = t.repeat(1, 1, fh, fw)
        m01 = run_conv_chain(self.convblock, torch.cat((feat0, feat1, tt), 1))
        m10 = run_conv_chain(self.convblock, torch.cat((feat1, feat0, 1 - tt), 1))
        from0, ft0 = self.dcnt0(feat0, m01)
        from1, ft1 = self.dcnt1(feat1, m10)
        return run_conv_chain(self.blendblock, torch.cat((from0, from1), 1)), ft0, ft1


class DATwConstantnCv1(nn.Module):
    def __init__(self, args):
        super().__init__()
        self.args = args
        nf = self.nf = args.nf
        self.feature_encoder = SameChannelResEncoder(nf, args.enc_res_blocks)
        self.coarse_query_builder = DCNInterFeatBuilderwithT(nf)
        self.lv4_to_lv3 = nn.ConvTranspose2d(nf + 4, nf + 4, 4, 2, 1)
        self.dat_lv3 = CrossDeformableAttentionBlockwFlow(nf, nf, n_samples=8, n_groups=4, n_heads=4, offset_scale=2.0,
                                                          mlp_ratio=args.mlp_ratio)
        self.lv3_to_lv2 = nn.ConvTranspose2d(nf, nf, 4, 2, 1)
        self.dat_lv2 = CrossDeformableAttentionBlockwFlow(nf, nf, n_samples=16, n_groups=8, n_heads=8, offset_scale=4.0,
                                                          mlp_ratio=args.mlp_ratio)
        self.lv2_to_lv1 = nn.ConvTranspose2d(nf, nf, 4, 2, 1)
        self.dat_lv1 = CrossDeformableAttentionBlockwFlow(nf, nf, n_samples=32, n_groups=8, n_heads=8, offset_scale=8.0,
                                                          mlp_ratio=args.mlp_ratio, pred_res_flow=False)
        self.pixel_geneartor = BasicResPixelShuffleGenerator(nf, args.dec_res_blocks)   # (sic) reference attribute name
        _conv.install_invalidation_hooks(self)

    @staticmethod
    def norm_w_rgb_mean(x0, x1):
        mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
        return x0 - mean_, x1 - mean_, mean_

    def _generate_frame(self, x0, x1, t, return_flows: bool = False, mean_=None, normalized: bool = False):
        nf = self.nf
        if normalized:      # vfi_frame_prep already subtracted the mean; the caller adds it back (vfi_frame_finish)
            pass
        elif mean_ is None:
            x0, x1, mean_ = self.norm_w_rgb_mean(x0, x1)
        else:       # tile of a larger frame: the mean is the GLOBAL one (models/BaseModel.py:14-17)
            x0, x1 = x0 - mean_, x1 - mean_
        if torch.is_grad_enabled() or x0.shape != x1.shape:
            f0 = self.feature_encoder(x0)
            f1 = self.feature_encoder(x1)
        else:
            # inference: both frames through the encoder as one batch (the same convolutions on twice the pixels: half
            # the launches, and the 1/4 .. 1/16-resolution levels fill the machine better); per-frame pyramids are views
            nb = x0.shape[0]
            fb = self.feature_encoder(_batch_pair(x0, x1))
            f0, f1 = [f[:nb] for f in fb], [f[nb:] for f in fb]
        q4, ft0_4, ft1_4 = self.coarse_query_builder(f0[3], f1[3], t)
        s3 = conv_transpose_up2(self.lv4_to_lv3, torch.cat((q4, ft0_4, ft1_4), dim=1))
        q3, ft0_3, ft1_3 = s3[:, :nf], s3[:, nf:nf + 2], s3[:, nf + 2:nf + 4]
        a3, ft0_2, ft1_2 = self.dat_lv3(q3, f0[2], f1[2], ft0_3, ft1_3)
        a2, ft0_1, ft1_1 = self.dat_lv2(conv_transpose_up2(self.lv3_to_lv2, a3), f0[1], f1[1], ft0_2, ft1_2)
        a1 = self.dat_lv1(conv_transpose_up2(self.lv2_to_lv1, a2), f0[0], f1[0], ft0_1, ft1_1)
        img = self.pixel_geneartor(a1, None if normalized else mean_)
        if return_flows:
            return img, dict(ft0=[ft0_1, ft0_2, ft0_3, ft0_4], ft1=[ft1_1, ft1_2, ft1_3, ft1_4])
        return img

    @torch.no_grad()
    def inference(self, x0, x1, t):
        """x0, x1 (B,3,H,W) in [0,1] with H, W multiples of 16; t (B,1,1,1) -> (B,3,H,W) in [0,1]."""
        return self._generate_frame(x0, x1, t)

    def forward(self, x0, x1, t):
        """Training mode: differentiable (the blocks take the reference's piecewise composition over the backward kernels).
        Eval mode with inputs that do not require grad is inference: run it without recording a graph even if the caller
        forgot ``torch.no_grad()`` -- the training route materialises ~40 GB of sampled kv tensors per 1080p pair."""
        if not self.training and not (x0.requires_grad or x1.requires_grad or t.requires_grad):
            with torch.no_grad():
                return self._generate_frame(x0, x1, t)
        return self._generate_frame(x0, x1, t)

    @torch.no_grad()
    def interpolate(self, img0, img1, t, out_uint8=None, divisor: int = 16):
        """Frames in, frame out: img0, img1 (B,3,H,W) of ANY size, float32 in [0,1] or uint8 as
        ``imread`` delivers them (evaluate.py:138-139).  Runs uint8 -> float / 255, ``InputPadder`` replicate padding,
        ``norm_w_rgb_mean``, the forward, ``clamp(out + mean, 0, 1)`` and the unpad with two fused passes around the
        forward (vfi_frame_prep / vfi_frame_finish) instead of ~10 full-frame PyTorch kernels, and returns (B,3,H,W) in the
        input's dtype (``out_uint8`` overrides; uint8 = round(255 * frame))."""
        from . import _lib
        dev = _lib.require_cuda(img0, img1)
        assert img0.shape == img1.shape and img0.dim() == 4 and img0.shape[1] == 3, "frames must be (B,3,H,W)"
        assert img0.dtype == img1.dtype and img0.dtype in (torch.float32, torch.uint8), "frames must be float32 or uint8"
        is_u8 = img0.dtype == torch.uint8
        out_u8 = is_u8 if out_uint8 is None else bool(out_uint8)
        img0, img1 = img0.contiguous(), img1.contiguous()
        B, _, H, W = img0.shape
        pad = InputPadder((H, W), divisor)._pad            # [left, right, top, bottom]
        Hp, Wp = H + pad[2] + pad[3], W + pad[0] + pad[1]
        x01 = torch.empty((2 * B, Hp, Wp, 4), dtype=torch.float32, device=dev)     # one buffer: the encoder takes both frames as a batch
        x0, x1 = x01[:B], x01[B:]
        mean = torch.empty((B,), dtype=torch.float32, device=dev)
        ws = torch.empty((B,), dtype=torch.float64, device=dev)
        st = _lib.stream_ptr(dev)
        _lib.call("vfi_frame_prep", dev, img0.data_ptr(), img1.data_ptr(), 1 if is_u8 else 0, x0.data_ptr(), x1.data_ptr(),
                  mean.data_ptr(), B, H, W, Hp, Wp, pad[2], pad[0], ws.data_ptr(), ws.numel() * 8, st)
        y = self._generate_frame(x0.permute(0, 3, 1, 2), x1.permute(0, 3, 1, 2), t, normalized=True)
        out = torch.empty((B, 3, H, W), dtype=torch.uint8 if out_u8 else torch.float32, device=dev)
        _lib.call("vfi_frame_finish", dev, y.data_ptr(), _lib.strides4(y), mean.data_ptr(), out.data_ptr(), 1 if out_u8 else 0,
                  B, H, W, Hp, Wp, pad[2], pad[0], st)
        return out

    @torch.no_grad()
    def inference_tiled(self, x0, x1, t, grid=(2, 4), halo=64, rank: int = 0, world: int = 1, out=None):
        """Spatially tiled inference (BASELINE configs[4]): every tile+halo crop runs the whole forward with the
        global rgb mean and, through ``_lib.crop_context``, the FRAME's extent in every frame-normalised op (bwarp, the
        samplers, the flow resize: SURVEY.md H5), so a crop samples exactly where the whole frame would and what is
        left of the tile-vs-whole difference is the receptive field of the conv stacks cut by the halo.  Interiors are
        written into ``out``.  ``rank``/``world`` select this process's tiles (no collective: each rank fills its own
        interiors; gather them however the caller likes)."""
        from . import _lib
        from .shard import tile_plan, tiles_of_rank
        B, _, H, W = x0.shape
        mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
        if out is None:
            out = torch.zeros((B, 3, H, W), dtype=x0.dtype, device=x0.device)
        for tl in tiles_of_rank(tile_plan(H, W, grid, halo), rank, world):
            ys, xs = tl.crop
            with _lib.crop_context(H, W, ys.start, xs.start, ys.stop - ys.start, xs.stop - xs.start):
                pred = self._generate_frame(x0[:, :, ys, xs].contiguous(), x1[:, :, ys, xs].contiguous(), t, mean_=mean_)
            iy, ix = tl.interior_in_crop
            oy, ox = tl.interior
            out[:, :, oy, ox] = pred[:, :, iy, ix]
        return out


class InputPadder:
    """Replicate-pad to a multiple of ``divisor`` and crop back (reference: data/__init__.py:6-20)."""

    def __init__(self, dims, divisor=16):
        self.ht, self.wd = dims[-2:]
        ph = (-self.ht) % divisor
        pw = (-self.wd) % divisor
        self._pad = [pw // 2, pw - pw // 2, ph // 2, ph - ph // 2]

    def pad(self, *inputs):
        return [F.pad(x, self._pad, mode='replicate') for x in inputs]

    def unpad(self, x):
        ht, wd = x.shape[-2:]
        return x[..., self._pad[2]:ht - self._pad[3], self._pad[0]:wd - self._pad[1]]
