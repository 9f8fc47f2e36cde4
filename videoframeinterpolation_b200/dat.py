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
from .conv import conv_transpose_up2
from .dcnv2 import DeformableConv2d
from .deformable_attn import CrossDeformableAttentionBlockwFlow

__all__ = ["DATwConstantnCv1", "DCNInterFeatBuilderwithT", "InputPadder", "default_args"]


def default_args(**over):
    """configs/DAT.yaml:4-11."""
    d = dict(nf=72, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2, distill_lambda=0.01)
    d.update(over)
    return SimpleNamespace(**d)


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

    @staticmethod
    def norm_w_rgb_mean(x0, x1):
        mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
        return x0 - mean_, x1 - mean_, mean_

    def _generate_frame(self, x0, x1, t, return_flows: bool = False, mean_=None):
        nf = self.nf
        if mean_ is None:
            x0, x1, mean_ = self.norm_w_rgb_mean(x0, x1)
        else:       # tile of a larger frame: the mean is the GLOBAL one (models/BaseModel.py:14-17)
            x0, x1 = x0 - mean_, x1 - mean_
        f0 = self.feature_encoder(x0)
        f1 = self.feature_encoder(x1)
        q4, ft0_4, ft1_4 = self.coarse_query_builder(f0[3], f1[3], t)
        s3 = conv_transpose_up2(self.lv4_to_lv3, torch.cat((q4, ft0_4, ft1_4), dim=1))
        q3, ft0_3, ft1_3 = s3[:, :nf], s3[:, nf:nf + 2], s3[:, nf + 2:nf + 4]
        a3, ft0_2, ft1_2 = self.dat_lv3(q3, f0[2], f1[2], ft0_3, ft1_3)
        a2, ft0_1, ft1_1 = self.dat_lv2(conv_transpose_up2(self.lv3_to_lv2, a3), f0[1], f1[1], ft0_2, ft1_2)
        a1 = self.dat_lv1(conv_transpose_up2(self.lv2_to_lv1, a2), f0[0], f1[0], ft0_1, ft1_1)
        img = self.pixel_geneartor(a1, mean_)
        if return_flows:
            return img, dict(ft0=[ft0_1, ft0_2, ft0_3, ft0_4], ft1=[ft1_1, ft1_2, ft1_3, ft1_4])
        return img

    @torch.no_grad()
    def inference(self, x0, x1, t):
        """x0, x1 (B,3,H,W) in [0,1] with H, W multiples of 16; t (B,1,1,1) -> (B,3,H,W) in [0,1]."""
        return self._generate_frame(x0, x1, t)

    def forward(self, x0, x1, t):
        return self._generate_frame(x0, x1, t)

    @torch.no_grad()
    def inference_tiled(self, x0, x1, t, grid=(2, 4), halo=64, rank: int = 0, world: int = 1, out=None):
        """Spatially tiled inference (BASELINE configs[4]): every tile+halo crop runs the whole forward with the
        global rgb mean; interiors are written into ``out``.  ``rank``/``world`` select this process's tiles
        (no collective: each rank fills its own interiors; gather them however the caller likes)."""
        from .shard import tile_plan, tiles_of_rank
        B, _, H, W = x0.shape
        mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
        if out is None:
            out = torch.zeros((B, 3, H, W), dtype=x0.dtype, device=x0.device)
        for tl in tiles_of_rank(tile_plan(H, W, grid, halo), rank, world):
            ys, xs = tl.crop
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
