"""DCNDATv1 (BASELINE configs[3]): the archive model whose blocks use the same offsets twice -- as the taps of a
deformable convolution that builds the query and as the sampling points of the deformable attention.

Reference: models/archive/DCNDAT.py (ResEncoder 18-54, DeformableConv2dGivenOffset 57-78, DCNInterFeatBuilderwithT
81-110, CrossDeformableAttentionBlockwFlow 113-199, DCNDATv1 202-251).  The class is unimportable upstream (SURVEY.md
Note B: it names `DeformAttn`, `Basemodel`, `Mlp`, `utils.resize`, `make_layer`, `generate_rgb_frame`, none of which
exist any more); the "repaired" reading used here and by `tests/golden/make_golden.py` maps them to `Attn`,
`BaseModel`, `FeadForward`, `BaseModel.resize`, `make_residual_layers` and a `BasicResPixelShuffleGenerator` held as
`self.generator` (body of `generate_rgb_frame`, models/archive/DCNTrans.py:118-123).  State-dict keys follow the
reference's attribute names (`cnn_encoder`, `dcn_feat_t_builder.dcn`, `dat_scale3.conv_res_offset_mask`, ...).

`inference(x0, x1, t)` follows DCNDAT.py:231-251 literally: the rgb mean is computed and added back at the end, but
NOT subtracted from the inputs (the subtraction is commented out at line 233).

Hot-path calls per level: 2x bwarp (vfi_bwarp_f32), 2x deformable conv with given offsets and offset groups =
n_groups != 8 weight groups (vfi_dcn_f32), the 9-point grouped sampler + attention over 18 samples in one kernel
(vfi_dattn_f32, `coords_final`: the kernel takes the final displacements because the DCN needs them materialised
anyway); 2x DCNv2 at 1/16 resolution.  Inference only (no autograd route).
"""
from __future__ import annotations

from types import SimpleNamespace

import torch
import torch.nn as nn

from . import _lib
from . import conv as _conv
from .blocks import (BasicResPixelShuffleGenerator, FeadForward, HalfChannelConv5ResBlock, SameChannelResEncoder,
                     conv_bias_act, conv_prelu, flow_up2_add, resize, run_conv_chain)
from .dcnv2 import DeformableConv2d, DeformableConv2dGivenOffset, deform_conv2d, window_or_auto
from .deformable_attn import (Attn, fused_deformable_attention, fused_supported, grouped_sample, movement_feats_pair,
                              movement_split_ok)
from .warp import bwarp

__all__ = ["DCNDATv1", "dcndat_default_args"]


def dcndat_default_args(**over):
    """configs/archive/DCNDAT.yaml:4-12."""
    d = dict(nf=64, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2, geo_lambda=0.01, distill_lambda=0.01)
    d.update(over)
    return SimpleNamespace(**d)


class SharedDCNInterFeatBuilderwithT(nn.Module):
    """Coarsest-level query builder with ONE DeformableConv2d shared by both directions (DCNDAT.py:81-110)."""

    def __init__(self, nc):
        super().__init__()
        self.nc = nc
        self.convblock = nn.Sequential(nn.Conv2d(nc * 2 + 1, nc, 3, 1, 1), nn.PReLU(nc), nn.Conv2d(nc, nc, 3, 1, 1), nn.PReLU(nc))
        self.dcn = DeformableConv2d(nc, nc)
        self.blendblock = nn.Sequential(nn.Conv2d(nc * 2, nc, 3, 1, 1), nn.PReLU(nc), nn.Conv2d(nc, nc, 3, 1, 1))

    def forward(self, feat0, feat1, t):
        _, _, fh, fw = feat0.shape
        tt = t.repeat(1, 1, fh, fw)
        m01 = run_conv_chain(self.convblock, torch.cat((feat0, feat1, tt), 1))
        m10 = run_conv_chain(self.convblock, torch.cat((feat1, feat0, 1 - tt), 1))
        from0, ft0 = self.dcn(feat0, m01)
        from1, ft1 = self.dcn(feat1, m10)
        return run_conv_chain(self.blendblock, torch.cat((from0, from1), 1)), ft0, ft1


class DCNCrossDeformableAttentionBlockwFlow(nn.Module):
    """DCNDAT.py:113-199.  One conv predicts, per (group, sample), a residual offset pair and a DCN mask; the offsets
    + the level's flow are both the DCN taps of the query enhancer and the attention's sampling points."""

    def __init__(self, in_c, out_c, n_samples=9, n_groups=12, n_heads=12, mlp_ratio=2.0, pred_res_flow=True):
        super().__init__()
        self.in_c, self.out_c = in_c, out_c
        self.n_groups, self.n_group_c = n_groups, in_c // n_groups
        self.n_heads, self.n_samples = n_heads, n_samples
        self.pred_res_flow = pred_res_flow
        self.conv_res_feat = nn.Sequential(conv_prelu(in_c * 2 + 2, in_c * 2), conv_prelu(in_c * 2, in_c),
                                           HalfChannelConv5ResBlock(in_c, in_c // 2))
        self.conv_res_offset_mask = nn.Conv2d(in_c, n_groups * n_samples * 3, 3, 1, 1)
        nn.init.zeros_(self.conv_res_offset_mask.weight)      # DCNDAT.py:131, 140-144
        nn.init.zeros_(self.conv_res_offset_mask.bias)
        if pred_res_flow:
            self.conv_res_flow = nn.ConvTranspose2d(in_c, 2, 4, 2, 1, bias=True)
        self.query_enhancer_dcn = DeformableConv2dGivenOffset(in_c, out_c)
        self.query_blender = nn.Conv2d(in_c * 3, in_c, 1, 1, 0)
        self.attn = Attn(in_c, out_c, n_samples * 2, n_heads)
        self.mlp = FeadForward(in_dim=out_c, hidden_dim=int(out_c * mlp_ratio), out_dim=out_c)

    def _get_movement_feats(self, feat_t, feat_x, ftx):
        return self.conv_res_feat(torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1))

    def _get_ref_coords_mask(self, feat_t, ftx, movement_feat):
        b, _, fh, fw = feat_t.shape
        out = conv_bias_act(self.conv_res_offset_mask, movement_feat)
        n2 = 2 * self.n_groups * self.n_samples
        res = 2.0 * torch.tanh(out[:, :n2])                       # chunks 0 and 1 of 3 (DCNDAT.py:153-154)
        coords = res.reshape(b, self.n_groups * self.n_samples, 2, fh, fw) + ftx.unsqueeze(1)
        return coords, out[:, n2:]

    def _get_ref_feats(self, feat, flow):
        return grouped_sample(feat, flow, self.n_groups, self.n_samples)

    def _coords_offsets_mask(self, ftx, movement_feat):
        """`_get_ref_coords_mask` + the permute / flip / sigmoid of the query enhancer (DCNDAT.py:69-70) in one pass
        (vfi_dcndat_coords_f32): coords (b, nG*nS, 2, h, w), DCN offsets (b, 2*nG*nS, h, w), sigmoid mask."""
        raw = _lib.f32(conv_bias_act(self.conv_res_offset_mask, movement_feat))     # read in whatever layout the conv wrote
        ftx = _lib.f32(ftx)
        dev = _lib.require_cuda(raw, ftx)
        b, _, fh, fw = raw.shape
        n = self.n_groups * self.n_samples
        coords = torch.empty((b, n, 2, fh, fw), dtype=torch.float32, device=dev)
        off = torch.empty((b, 2 * n, fh, fw), dtype=torch.float32, device=dev)
        mask = torch.empty((b, n, fh, fw), dtype=torch.float32, device=dev)
        _lib.call("vfi_dcndat_coords_f32", dev, raw.data_ptr(), _lib.strides4(raw), ftx.data_ptr(), _lib.strides4(ftx),
                  coords.data_ptr(), off.data_ptr(), mask.data_ptr(), b, n, fh, fw, _lib.stream_ptr(dev))
        return coords, off, mask

    def _blend_query(self, parts):
        """``query_blender(cat(parts))`` (DCNDAT.py:182): the 1x1 conv as three planar GEMMs W_i @ part_i accumulated in
        place -- no 192-channel concat and no transpose of it to the conv's layout (1.1 ms at 544x960; concatenating
        straight into a channels-last buffer was measured slower still).  Same math mode as the cuDNN convs around it."""
        qb = self.query_blender
        x0 = parts[0]
        if not (x0.is_cuda and x0.dtype == torch.float32 and not torch.is_grad_enabled() and qb.bias is not None):
            return qb(torch.cat(parts, dim=1))
        b, c, fh, fw = x0.shape
        w = qb.weight.reshape(qb.out_channels, 3, c)
        out = qb.bias.view(1, -1, 1).expand(b, qb.out_channels, fh * fw).contiguous()
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = bool(torch.backends.cudnn.allow_tf32)
        try:
            for i, part in enumerate(parts):
                out.baddbmm_(w[:, i].unsqueeze(0).expand(b, -1, -1), _lib.planar(part).view(b, c, fh * fw))
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        return out.view(b, qb.out_channels, fh, fw)

    def forward(self, feat_t, feat0, feat1, ft0, ft1):
        if movement_split_ok(self.conv_res_feat, feat_t, feat0):
            mv0, mv1 = movement_feats_pair(self.conv_res_feat, feat_t, feat0, feat1, ft0, ft1)
        else:
            mv0 = self._get_movement_feats(feat_t, feat0, ft0)
            mv1 = self._get_movement_feats(feat_t, feat1, ft1)
        c0, off0, m0 = self._coords_offsets_mask(ft0, mv0)
        c1, off1, m1 = self._coords_offsets_mask(ft1, mv1)
        rc = self.query_enhancer_dcn.regular_conv
        feat0, feat1 = _lib.planar(feat0), _lib.planar(feat1)      # gathered by the DCN and by the attention: one transpose each
        b, c, fh, fw = feat_t.shape
        algo = window_or_auto(feat0, rc.weight)       # offsets = 2 tanh(.) + a smooth flow: the staged window holds the taps
        parts = (deform_conv2d(feat0, off0, rc.weight, rc.bias, padding=1, mask=m0, algo=algo), feat_t,
                 deform_conv2d(feat1, off1, rc.weight, rc.bias, padding=1, mask=m1, algo=algo))
        q = self._blend_query(parts)
        if fused_supported(self.in_c, self.n_groups, self.n_heads):
            attended = fused_deformable_attention(q, feat0, feat1, c0.reshape(b, -1, fh, fw), c1.reshape(b, -1, fh, fw), None, None,
                                                  *self.attn._weights(), n_groups=self.n_groups, n_samples=self.n_samples,
                                                  n_heads=self.n_heads, offset_scale=1.0, coords_final=True,
                                                  # q is planar (GEMMs over planar DCN outputs); the conv stacks and the
                                                  # MLP GEMMs downstream run channels-last when the features do
                                                  out_channels_last=feat_t.is_contiguous(memory_format=torch.channels_last)
                                                  and not feat_t.is_contiguous())
        else:
            attended = self.attn(q, torch.cat((self._get_ref_feats(feat0, c0), self._get_ref_feats(feat1, c1)), dim=2))
        out = self.mlp.forward_residual(attended)
        if not self.pred_res_flow:
            return out
        up_ft0 = flow_up2_add(_conv.conv_transpose_up2(self.conv_res_flow, mv0, out_channels_last=False), ft0)
        up_ft1 = flow_up2_add(_conv.conv_transpose_up2(self.conv_res_flow, mv1, out_channels_last=False), ft1)
        return out, up_ft0, up_ft1


class DCNDATv1(nn.Module):
    def __init__(self, args):
        super().__init__()
        self.args = args
        nf = self.nf = args.nf
        self.cnn_encoder = SameChannelResEncoder(nf, args.enc_res_blocks)         # == ResEncoder, DCNDAT.py:18-54
        self.dcn_feat_t_builder = SharedDCNInterFeatBuilderwithT(nf)
        self.query_builder3 = nn.ConvTranspose2d(nf + 4, nf + 4, 4, 2, 1)
        self.dat_scale3 = DCNCrossDeformableAttentionBlockwFlow(nf, nf, n_samples=9, n_groups=8, n_heads=8, mlp_ratio=args.mlp_ratio)
        self.query_builder2 = nn.ConvTranspose2d(nf, nf, 4, 2, 1)
        self.dat_scale2 = DCNCrossDeformableAttentionBlockwFlow(nf, nf, n_samples=9, n_groups=4, n_heads=4, mlp_ratio=args.mlp_ratio)
        self.query_builder1 = nn.ConvTranspose2d(nf, nf, 4, 2, 1)
        self.dat_scale1 = DCNCrossDeformableAttentionBlockwFlow(nf, nf, n_samples=9, n_groups=4, n_heads=4, mlp_ratio=args.mlp_ratio,
                                                                pred_res_flow=False)
        self.generator = BasicResPixelShuffleGenerator(nf, args.dec_res_blocks)   # generate_rgb_frame (Note B)

    @torch.no_grad()
    def inference(self, x0, x1, t):
        nf = self.nf
        mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
        f0_1, f0_2, f0_3, f0_4 = self.cnn_encoder(x0)          # inputs NOT mean-shifted (DCNDAT.py:233)
        f1_1, f1_2, f1_3, f1_4 = self.cnn_encoder(x1)
        feat_t_4, ft0_4, ft1_4 = self.dcn_feat_t_builder(f0_4, f1_4, t)
        s3 = self.query_builder3(torch.cat((feat_t_4, ft0_4, ft1_4), dim=1))
        feat_t_3, ft0_3, ft1_3 = s3[:, :nf], s3[:, nf:nf + 2], s3[:, nf + 2:nf + 4]
        a3, ft0_2, ft1_2 = self.dat_scale3(feat_t_3, f0_3, f1_3, ft0_3, ft1_3)
        a2, ft0_1, ft1_1 = self.dat_scale2(self.query_builder2(a3), f0_2, f1_2, ft0_2, ft1_2)
        a1 = self.dat_scale1(self.query_builder1(a2), f0_1, f1_1, ft0_1, ft1_1)
        return self.generator(a1, mean_)

    def forward(self, x0, x1, t):
        """Inference only: this archive model has no training route here (its DCN-with-given-offsets query builder and the
        fused attention on final coordinates run forward kernels).  Asking for gradients raises instead of silently
        returning a detached tensor."""
        if torch.is_grad_enabled() and (self.training or x0.requires_grad or x1.requires_grad):
            raise _lib.VfiError("DCNDATv1 is inference-only in vfi_b200: call .eval() and .inference(x0, x1, t) (or wrap the call "
                                "in torch.no_grad()); gradients are not propagated through this model")
        return self.inference(x0, x1, t)
