"""Drop-in for the hot-path half of the reference's ``modules/madat.py`` (motion-aware DAT block).

    DeformAttnwMotion(in_c, out_c, n_samples, n_heads).forward(q, kv) -> (out, score0, score1)     lines 70-115
    DATwithMotionEstimation._get_movement_feats / ._get_ref_flow / ._get_ref_feats                 lines 150-174

Upstream, ``DATwithMotionEstimation.forward`` cannot run (``corr.coords_grid`` returns a 5-D, h/w-swapped grid and
the following rearrange raises; SURVEY.md note A), so the parity targets are the sub-paths above.  ``attend``
is the fused equivalent of lines 179-190 (sampler of both frames + attention + per-frame scores + grids).
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib
from .deformable_attn import attention_over_samples, fused_deformable_attention, grouped_sample
from .warp import bwarp

__all__ = ["DeformAttnwMotion", "DATwithMotionEstimation", "SmallResBlock", "DWConvMlp"]


def _cp(cin, cout):
    return nn.Sequential(nn.Conv2d(cin, cout, 3, 1, 1), nn.PReLU(cout))


class SmallResBlock(nn.Module):
    """3-conv residual block, middle conv on the last ``side_channels`` channels (modules/madat.py:47-67)."""

    def __init__(self, in_channels, side_channels, bias=True):
        super().__init__()
        self.side_channels = side_channels
        self.conv1, self.conv2, self.conv3 = _cp(in_channels, in_channels), _cp(side_channels, side_channels), \
            _cp(in_channels, in_channels)

    def forward(self, x):
        t = self.conv1(x)
        keep = t.shape[1] - self.side_channels
        t = torch.cat((t[:, :keep], self.conv2(t[:, keep:])), dim=1)
        return x + self.conv3(t)


class _DW(nn.Module):
    def __init__(self, dim):
        super().__init__()
        self.dwconv = nn.Conv2d(dim, dim, 3, 1, 1, bias=True, groups=dim)

    def forward(self, x, H, W):
        B, N, C = x.shape
        return self.dwconv(x.transpose(1, 2).reshape(B, C, H, W)).reshape(B, C, -1).transpose(1, 2)


class DWConvMlp(nn.Module):
    """Linear -> depthwise 3x3 -> GELU -> Linear on (B, HW, C) tokens (keys fc1 / dwconv.dwconv / fc2; modules/corr.py:28-44)."""

    def __init__(self, in_features, hidden_features=None, out_features=None):
        super().__init__()
        hidden_features, out_features = hidden_features or in_features, out_features or in_features
        self.fc1 = nn.Linear(in_features, hidden_features)
        self.dwconv = _DW(hidden_features)
        self.act = nn.GELU()
        self.fc2 = nn.Linear(hidden_features, out_features)

    def forward(self, x, H, W):
        return self.fc2(self.act(self.dwconv(self.fc1(x), H, W)))


class DeformAttnwMotion(nn.Module):
    def __init__(self, in_c, out_c, n_samples, n_heads):
        super().__init__()
        assert in_c == out_c
        self.in_c, self.out_c = in_c, out_c
        self.q_proj, self.k_proj, self.v_proj = nn.Linear(in_c, out_c), nn.Linear(in_c, out_c), nn.Linear(in_c, out_c)
        self.n_samples, self.n_head, self.n_head_c = n_samples, n_heads, out_c // n_heads
        self.scale = self.n_head_c ** -0.5

    def _weights(self):
        return (self.q_proj.weight, self.q_proj.bias, self.k_proj.weight, self.k_proj.bias, self.v_proj.weight,
                self.v_proj.bias)

    def forward(self, q, kv):
        """q: b,c,fh,fw; kv: b,c,nS,fh*fw -> (out, softmax(logits[:, :nS/2]), softmax(logits[:, nS/2:]))."""
        return attention_over_samples(q, kv, *self._weights(), n_heads=self.n_head, want_scores=True)


class DATwithMotionEstimation(nn.Module):
    def __init__(self, in_c, out_c, n_samples=9, n_groups=12, n_heads=12, mlp_ratio=2.0):
        super().__init__()
        assert n_groups == n_heads, "For now different number of n_heads and n_groups are not supported"
        self.in_c, self.out_c = in_c, out_c
        self.n_groups, self.n_group_c, self.n_heads, self.n_samples = n_groups, in_c // n_groups, n_heads, n_samples
        self.norm1, self.norm2 = nn.LayerNorm(in_c), nn.LayerNorm(out_c)
        self.mlp = DWConvMlp(out_c, int(out_c * mlp_ratio))
        self.coord_proj = nn.Linear(2, out_c)
        self.motion_proj = nn.Linear(out_c, out_c)
        self.conv_res_feat = nn.Sequential(nn.Conv2d(in_c * 2 + 2, in_c, 3, 1, 1), nn.PReLU(in_c),
                                           SmallResBlock(in_c, in_c // 2))
        self.conv_res_offset = nn.Conv2d(in_c, n_groups * n_samples * 2, 3, 1, 1)
        nn.init.zeros_(self.conv_res_offset.weight)
        nn.init.zeros_(self.conv_res_offset.bias)
        self.attn = DeformAttnwMotion(in_c, out_c, n_samples * 2, n_heads)

    def _norm_c(self, feat):
        return self.norm1(feat.permute(0, 2, 3, 1)).permute(0, 3, 1, 2).contiguous()

    def _get_movement_feats(self, feat_t, feat_x, ftx):
        return self.conv_res_feat(torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1))

    def _get_ref_flow(self, feat_t, ftx, movement_feat):
        b, c, fh, fw = feat_t.shape
        res = 2.0 * torch.tanh(self.conv_res_offset(movement_feat))
        return res.view(b, self.n_groups * self.n_samples, 2, fh, fw) + ftx.unsqueeze(1)

    def _get_ref_feats(self, feat, flow):
        """-> (samples (b,c,nS,fh*fw), normalised grid (b, nG*nS, fh*fw, 2)); LayerNorm over C first (line 163)."""
        return grouped_sample(self._norm_c(feat), flow, self.n_groups, self.n_samples, want_grid=True)

    def attend(self, feat_t, feat0, feat1, ft0, ft1):
        """Fused lines 179-190: returns (attended, score0, score1) without materialising the samples."""
        mv0 = self._get_movement_feats(feat_t, feat0, ft0)
        mv1 = self._get_movement_feats(feat_t, feat1, ft1)
        return fused_deformable_attention(self._norm_c(feat_t), self._norm_c(feat0), self._norm_c(feat1),
                                          self.conv_res_offset(mv0), self.conv_res_offset(mv1), ft0, ft1,
                                          *self.attn._weights(), n_groups=self.n_groups, n_samples=self.n_samples,
                                          n_heads=self.n_heads, offset_scale=2.0, want_scores=True)

    def forward(self, feat_t, feat0, feat1, ft0, ft1):
        raise NotImplementedError("DATwithMotionEstimation.forward does not run in the reference either (modules/corr.py:8-12 "
                                  "feeds a 5-D grid to the rearrange at modules/madat.py:196); use attend() / the sub-paths")
