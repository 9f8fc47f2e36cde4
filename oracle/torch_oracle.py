"""CPU oracle (fp32, torch tensor ops) for the warp / splat / DCNv2 / deformable-attention path.

TEST INFRASTRUCTURE ONLY -- see ``oracle/__init__.py``.  Every function restates one reference
function op by op (same fp32 rounding sequence for coordinates) and cites it.  Paths below are
relative to the reference checkout (``/root/reference`` in the build container; absent on the GPU box).

Third-party arithmetic restated here (un-vendored by the reference, versions = this image's):
  * ``torch.nn.functional.grid_sample`` (torch 2.11.0, ATen GridSampler, bilinear, align_corners=True,
    padding ``border`` / ``zeros``)            -> ``grid_sample_bilinear``
  * ``torchvision.ops.deform_conv2d`` (torchvision 0.26.0, deform_conv2d_kernel)  -> ``deform_conv2d``
With ``library_ops(True)`` the two restatements are swapped for the library calls the reference itself
makes (used for the timed CPU baseline, where the reference's speed is what is being measured).

The model half is written functionally over a ``state_dict`` (parameter names identical to the
reference, e.g. ``dat_lv3.attn.k_proj.weight``) so reference checkpoints drive it directly.
"""
from __future__ import annotations

import contextlib
import math
from typing import Dict, Optional, Tuple

import torch
import torch.nn.functional as F

__all__ = [
    "library_ops", "grid_sample_bilinear", "bwarp", "splat_sum", "fwarp", "softsplat",
    "fwarp_using_two_frames", "fwarp_mframes", "deform_conv2d", "dcn_block", "dcn_given_offset",
    "ref_coords", "dsample", "attn", "dat_block", "madat_ref_feats", "madat_attn",
    "layer_norm_c", "conv3x3_tf32", "tf32_truncate", "tf32_round", "calculate_ssim", "calculate_psnr", "dat_forward", "dcndat_forward", "dcndat_config", "dcndat_block", "dat_config", "dat_param_shapes", "input_pad", "input_unpad", "count_params", "resize",
]

_USE_LIBRARY = False


@contextlib.contextmanager
def library_ops(flag: bool = True):
    """Route the bilinear sampler / DCN through torch / torchvision CPU ops (what the reference calls)."""
    global _USE_LIBRARY
    old, _USE_LIBRARY = _USE_LIBRARY, flag
    try:
        yield
    finally:
        _USE_LIBRARY = old


# ----------------------------------------------------------------------------------------------
# bilinear sampling (ATen GridSampler restated) and bwarp
# ----------------------------------------------------------------------------------------------

def _take(plane: torch.Tensor, xi: torch.Tensor, yi: torch.Tensor, wgt: torch.Tensor) -> torch.Tensor:
    """plane (C,H,W); xi, yi, wgt (P,) float -> (C,P): wgt * plane[:, yi, xi], zero where (xi,yi) is outside."""
    C, H, W = plane.shape
    ok = (xi >= 0) & (xi <= W - 1) & (yi >= 0) & (yi <= H - 1)
    lin = (yi.clamp(0, H - 1) * W + xi.clamp(0, W - 1)).to(torch.int64)
    vals = plane.reshape(C, H * W).index_select(1, lin)
    return vals * (wgt * ok.to(wgt.dtype)).unsqueeze(0)


def grid_sample_bilinear(img: torch.Tensor, gx: torch.Tensor, gy: torch.Tensor, padding: str) -> torch.Tensor:
    """ATen ``grid_sampler_2d`` (bilinear, align_corners=True) restated.

    img (N,C,H,W); gx, gy (N,Ho,Wo) normalised to [-1,1] -> (N,C,Ho,Wo).
    Call sites in the reference: modules/warp.py:14 (border), modules/deformable_attn.py:99 and
    modules/madat.py:171 (zeros).
    """
    N, C, H, W = img.shape
    if _USE_LIBRARY:
        return F.grid_sample(img, torch.stack((gx, gy), dim=-1), mode="bilinear", padding_mode=padding,
                             align_corners=True)
    Ho, Wo = gx.shape[1:]
    ix = ((gx + 1.0) / 2.0) * (W - 1)          # unnormalise, align_corners=True
    iy = ((gy + 1.0) / 2.0) * (H - 1)
    if padding == "border":
        ix = ix.clamp(0, W - 1)
        iy = iy.clamp(0, H - 1)
    else:
        assert padding == "zeros"
        # keep the float->int conversions below finite; far-out points have no in-bounds corner anyway
        ix = ix.clamp(-4.0, W + 4.0)
        iy = iy.clamp(-4.0, H + 4.0)
    xw, yn = ix.floor(), iy.floor()
    tx, ty = ix - xw, iy - yn
    ex, ey = 1.0 - tx, 1.0 - ty
    out = img.new_empty(N, C, Ho * Wo)
    for n in range(N):
        x0, y0 = xw[n].reshape(-1), yn[n].reshape(-1)
        acc = _take(img[n], x0, y0, (ey[n] * ex[n]).reshape(-1))
        acc = acc + _take(img[n], x0 + 1, y0, (ey[n] * tx[n]).reshape(-1))
        acc = acc + _take(img[n], x0, y0 + 1, (ty[n] * ex[n]).reshape(-1))
        acc = acc + _take(img[n], x0 + 1, y0 + 1, (ty[n] * tx[n]).reshape(-1))
        out[n] = acc
    return out.view(N, C, Ho, Wo)


def bwarp(img: torch.Tensor, flow: torch.Tensor) -> torch.Tensor:
    """Backward bilinear warp with border clamp -- modules/warp.py:7-15.

    The normalise -> unnormalise round trip is kept in fp32 exactly as the reference performs it
    (SURVEY.md H1): lin[x] + flow_x / ((W-1)/2), then ((g+1)/2)*(W-1).
    """
    B, _, H, W = flow.shape
    lin_x = torch.linspace(-1.0, 1.0, W)
    lin_y = torch.linspace(-1.0, 1.0, H)
    gx = lin_x.view(1, 1, W).to(img) + flow[:, 0] / ((W - 1.0) / 2.0)
    gy = lin_y.view(1, H, 1).to(img) + flow[:, 1] / ((H - 1.0) / 2.0)
    return grid_sample_bilinear(img, gx, gy, "border")


# ----------------------------------------------------------------------------------------------
# forward splatting (the reference has only a CUDA kernel string for this)
# ----------------------------------------------------------------------------------------------

def splat_sum(ten_in: torch.Tensor, ten_flow: torch.Tensor) -> torch.Tensor:
    """Summation splat -- kernel text at modules/softsplat.py:289-341, restated with index_add_.

    fp32 throughout; weights formed as ((x0+1)-fx)*((y0+1)-fy) etc. in the kernel's order (lines
    320-323); non-finite targets skipped (306-307); each corner bounds-tested on its own (325-339).
    """
    ten_in = ten_in.float()
    ten_flow = ten_flow.float()
    N, C, H, W = ten_in.shape
    xs = torch.arange(W, dtype=torch.float32).view(1, 1, W)
    ys = torch.arange(H, dtype=torch.float32).view(1, H, 1)
    fx = xs + ten_flow[:, 0]
    fy = ys + ten_flow[:, 1]
    finite = torch.isfinite(fx) & torch.isfinite(fy)
    big = 1.0e8                                         # beyond any frame; keeps int conversion defined
    fx = torch.where(finite, fx, torch.zeros_like(fx)).clamp(-big, big)
    fy = torch.where(finite, fy, torch.zeros_like(fy)).clamp(-big, big)
    x0, y0 = fx.floor(), fy.floor()
    x1, y1 = x0 + 1.0, y0 + 1.0
    corners = (
        (x0, y0, (x1 - fx) * (y1 - fy)),                # north-west
        (x1, y0, (fx - x0) * (y1 - fy)),                # north-east
        (x0, y1, (x1 - fx) * (fy - y0)),                # south-west
        (x1, y1, (fx - x0) * (fy - y0)),                # south-east
    )
    out = ten_in.new_zeros(N, C, H * W)
    src = ten_in.reshape(N, C, H * W)
    for cx, cy, wgt in corners:
        ok = finite & (cx >= 0) & (cx < W) & (cy >= 0) & (cy < H)
        lin = (cy.clamp(0, H - 1) * W + cx.clamp(0, W - 1)).to(torch.int64).reshape(N, -1)
        w = (wgt * ok.to(wgt.dtype)).reshape(N, 1, -1)
        for n in range(N):
            out[n].index_add_(1, lin[n], src[n] * w[n])
    return out.view(N, C, H, W)


def _normalise(ten_out: torch.Tensor, mode: str) -> torch.Tensor:
    """modules/warp.py:39-56 (== modules/softsplat.py:259-276)."""
    norm = ten_out[:, -1:, :, :]
    parts = mode.split("-")
    if len(parts) == 1 or parts[1] == "addeps":
        norm = norm + 0.0000001
    elif parts[1] == "zeroeps":
        norm = torch.where(norm == 0.0, torch.ones_like(norm), norm)
    elif parts[1] == "clipeps":
        norm = norm.clip(0.0000001, None)
    return ten_out[:, :-1, :, :] / norm


def fwarp(ten_in, ten_flow, ten_metric, mode: str = "soft"):
    """modules/warp.py:18-59 and its twin modules/softsplat.py:237-279."""
    head = mode.split("-")[0]
    assert head in ("sum", "avg", "linear", "soft")
    if mode in ("sum", "avg"):
        assert ten_metric is None
    if head in ("linear", "soft"):
        assert ten_metric is not None
    if mode == "avg":
        ten_in = torch.cat([ten_in, ten_in.new_ones(ten_in.shape[0], 1, *ten_in.shape[2:])], 1)
    elif head == "linear":
        ten_in = torch.cat([ten_in * ten_metric, ten_metric], 1)
    elif head == "soft":
        e = ten_metric.exp()
        ten_in = torch.cat([ten_in * e, e], 1)
    out = splat_sum(ten_in, ten_flow)
    if head in ("avg", "linear", "soft"):
        out = _normalise(out, mode)
    return out


softsplat = fwarp


def _one_direction(ten_in, ten_flow, td, ten_metric):
    """modules/warp.py:69-73 / 91-96."""
    e = td * ten_metric.clip(-20.0, 20.0).exp()
    # reference order of multiplies: tenIn * td * exp(...)
    num_in = ten_in * td * ten_metric.clip(-20.0, 20.0).exp()
    out = splat_sum(torch.cat([num_in, e], 1), ten_flow)
    return out[:, :-1], out[:, -1:] + 0.0000001


def fwarp_using_two_frames(in1, flow1, t1, in2, flow2, t2, metric1=None, metric2=None):
    """modules/warp.py:68-83."""
    of, nf = _one_direction(in1, flow1, t1, metric1)
    ob, nb = _one_direction(in2, flow2, t2, metric2)
    num, den = of + ob, nf + nb
    return num / den, den < 0.00001


def fwarp_mframes(in1, flow1, t1, in2, flow2, t2, metric1, metric2):
    """modules/warp.py:86-114: flows (n_branch,B,2,H,W), t (n_branch,B,1,1,1); inputs shared by branches."""
    n_branch = flow1.shape[0]
    num, den = 0, 0
    for i in range(n_branch):
        of, nf = _one_direction(in1, flow1[i], t1[i], metric1)
        ob, nb = _one_direction(in2, flow2[i], t2[i], metric2)
        num = num + (of + ob)
        den = den + (nf + nb)
    return num / den, den < 0.00001


# ----------------------------------------------------------------------------------------------
# DCNv2 (torchvision.ops.deform_conv2d restated) and the reference block around it
# ----------------------------------------------------------------------------------------------

def _tv_bilinear(plane: torch.Tensor, h: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """torchvision deform_conv2d ``bilinear_interpolate``: plane (C,H,W); h, w (P,) -> (C,P).

    Zero if h <= -1 or h >= H or w <= -1 or w >= W; otherwise the four corners, each dropped when
    it lies outside the plane.
    """
    C, H, W = plane.shape
    inside = (h > -1) & (h < H) & (w > -1) & (w < W)
    hc = torch.where(inside, h, torch.zeros_like(h))
    wc = torch.where(inside, w, torch.zeros_like(w))
    h_lo, w_lo = hc.floor(), wc.floor()
    lh, lw = hc - h_lo, wc - w_lo
    hh, hw = 1.0 - lh, 1.0 - lw
    ins = inside.to(plane.dtype)
    acc = _take(plane, w_lo, h_lo, hh * hw * ins)
    acc = acc + _take(plane, w_lo + 1, h_lo, hh * lw * ins)
    acc = acc + _take(plane, w_lo, h_lo + 1, lh * hw * ins)
    acc = acc + _take(plane, w_lo + 1, h_lo + 1, lh * lw * ins)
    return acc


def deform_conv2d(x, offset, weight, bias=None, stride=1, padding=1, dilation=1, mask=None):
    """torchvision.ops.deform_conv2d semantics (call sites modules/dcnv2.py:60-66, models/archive/DCNDAT.py:71-77).

    offset (B, OG*2*kh*kw, Ho, Wo) interleaved (dy, dx) per (offset group, tap); mask (B, OG*kh*kw, Ho, Wo);
    weight (Cout, Cin/WG, kh, kw).  OG and WG are inferred exactly as torchvision infers them.
    """
    if _USE_LIBRARY:
        import torchvision.ops
        return torchvision.ops.deform_conv2d(x, offset, weight, bias, stride=stride, padding=padding,
                                             dilation=dilation, mask=mask)
    B, Cin, H, W = x.shape
    Cout, cin_g, kh, kw = weight.shape
    K = kh * kw
    WG = Cin // cin_g
    OG = offset.shape[1] // (2 * K)
    cpg = Cin // OG
    Ho = (H + 2 * padding - (dilation * (kh - 1) + 1)) // stride + 1
    Wo = (W + 2 * padding - (dilation * (kw - 1) + 1)) // stride + 1
    assert offset.shape[2:] == (Ho, Wo)
    base_h = (torch.arange(Ho, dtype=x.dtype) * stride - padding).view(Ho, 1).expand(Ho, Wo)
    base_w = (torch.arange(Wo, dtype=x.dtype) * stride - padding).view(1, Wo).expand(Ho, Wo)
    cols = x.new_empty(B, Cin, K, Ho * Wo)
    for b in range(B):
        for og in range(OG):
            plane = x[b, og * cpg:(og + 1) * cpg]
            for i in range(kh):
                for j in range(kw):
                    tap = i * kw + j
                    dy = offset[b, og * 2 * K + 2 * tap]
                    dx = offset[b, og * 2 * K + 2 * tap + 1]
                    hs = (base_h + i * dilation + dy).reshape(-1)
                    ws = (base_w + j * dilation + dx).reshape(-1)
                    val = _tv_bilinear(plane, hs, ws)
                    if mask is not None:
                        val = val * mask[b, og * K + tap].reshape(1, -1)
                    cols[b, og * cpg:(og + 1) * cpg, tap] = val
    cols = cols.view(B, WG, cin_g * K, Ho * Wo)
    wmat = weight.reshape(WG, Cout // WG, cin_g * K)
    out = torch.einsum("gok,bgkp->bgop", wmat, cols).reshape(B, Cout, Ho, Wo)
    if bias is not None:
        out = out + bias.view(1, -1, 1, 1)
    return out


class _P:
    """Prefix view of a flat state dict."""

    def __init__(self, sd: Dict[str, torch.Tensor], prefix: str = ""):
        self.sd, self.prefix = sd, prefix

    def sub(self, name: str) -> "_P":
        return _P(self.sd, f"{self.prefix}{name}.")

    def __getitem__(self, name: str) -> torch.Tensor:
        return self.sd[self.prefix + name]

    def has(self, name: str) -> bool:
        return (self.prefix + name) in self.sd


def _conv(p: _P, name: str, x, stride=1, padding=1):
    return F.conv2d(x, p[name + ".weight"], p[name + ".bias"], stride=stride, padding=padding)


def _prelu(p: _P, name: str, x):
    return F.prelu(x, p[name + ".weight"])


def _deconv(p: _P, name: str, x):
    return F.conv_transpose2d(x, p[name + ".weight"], p[name + ".bias"], stride=2, padding=1)


def dcn_block(p: _P, x, movement_feat):
    """``DeformableConv2d.forward`` -- modules/dcnv2.py:49-67. Returns (out, flow)."""
    flow = _conv(p, "offset_flow_conv", movement_feat)
    warped = bwarp(x, flow)
    h = torch.cat((warped, movement_feat, flow), dim=1)
    h = _prelu(p, "conv_offset_mask.1", _conv(p, "conv_offset_mask.0", h))
    h = _prelu(p, "conv_offset_mask.3", _conv(p, "conv_offset_mask.2", h))
    h = _conv(p, "conv_offset_mask.4", h)
    o1, o2, m = torch.chunk(h, 3, dim=1)
    res = 2.0 * torch.tanh(torch.cat((o1, o2), dim=1))
    offset = res + flow.flip(1).repeat(1, res.size(1) // 2, 1, 1)
    out = deform_conv2d(x, offset, p["regular_conv.weight"], p["regular_conv.bias"], stride=1, padding=1,
                        mask=torch.sigmoid(m))
    return out, flow


def dcn_given_offset(weight, bias, x, offset, mask):
    """``DeformableConv2dGivenOffset.forward`` -- models/archive/DCNDAT.py:67-78.

    offset (B, nG*9, 2, h, w) as produced by the DAT block; permuted/flipped/viewed exactly as line 69.
    """
    b, _, fh, fw = x.shape
    off = offset.permute(0, 2, 1, 3, 4).flip(1).contiguous().view(b, -1, fh, fw)
    return deform_conv2d(x, off, weight, bias, stride=1, padding=1, mask=torch.sigmoid(mask))


# ----------------------------------------------------------------------------------------------
# deformable attention (DAT flavour: modules/deformable_attn.py; MADAT flavour: modules/madat.py)
# ----------------------------------------------------------------------------------------------

def ref_coords(raw_offsets, ftx, n_groups: int, n_samples: int, offset_scale: float):
    """``_get_ref_coords`` -- modules/deformable_attn.py:81-86 (and madat.py:154-159 with scale 2)."""
    b, _, fh, fw = raw_offsets.shape
    res = offset_scale * torch.tanh(raw_offsets)
    res = res.view(b, n_groups * n_samples, 2, fh, fw)
    return res + ftx.unsqueeze(1)


def _sample_groups(feat, flow, n_groups: int, n_samples: int):
    b, c, fh, fw = feat.shape
    lin_x = torch.linspace(-1.0, 1.0, fw).to(feat)
    lin_y = torch.linspace(-1.0, 1.0, fh).to(feat)
    # the x displacement is divided by (fw-1) and THEN by 2 (reference quirk, line 94); y by ((fh-1)/2)
    gx = lin_x.view(1, 1, 1, fw) + flow[:, :, 0] / (fw - 1.0) / 2.0      # (b, nG*nS, fh, fw)
    gy = lin_y.view(1, 1, fh, 1) + flow[:, :, 1] / ((fh - 1.0) / 2.0)
    gxv = gx.reshape(b * n_groups, n_samples, fh * fw)
    gyv = gy.reshape(b * n_groups, n_samples, fh * fw)
    samples = grid_sample_bilinear(feat.reshape(b * n_groups, c // n_groups, fh, fw), gxv, gyv, "zeros")
    return samples.view(b, c, n_samples, fh * fw), gxv, gyv


def dsample(feat, flow, n_groups: int, n_samples: int):
    """``_get_ref_feats`` -- modules/deformable_attn.py:88-101: (b,c,h,w),(b,nG*nS,2,h,w) -> (b,c,nS,h*w)."""
    return _sample_groups(feat, flow, n_groups, n_samples)[0]


def attn(p: _P, q, kv, n_heads: int):
    """``Attn.forward`` -- modules/deformable_attn.py:26-40 (1x1-conv projections)."""
    b, c, fh, fw = q.shape
    S = kv.shape[2]
    out_c = p["q_proj.weight"].shape[0]
    hc = out_c // n_heads
    qp = _conv(p, "q_proj", q, padding=0).view(b * n_heads, hc, fh * fw)
    k = _conv(p, "k_proj", kv, padding=0).view(b * n_heads, hc, S, fh * fw)
    v = _conv(p, "v_proj", kv, padding=0).view(b * n_heads, hc, S, fh * fw)
    logits = torch.einsum("bcd,bcsd->bsd", qp, k).mul(hc ** -0.5)
    w = F.softmax(logits, dim=1)
    return torch.einsum("bsd,bcsd->bcd", w, v).contiguous().view(b, out_c, fh, fw)


def _res5(p: _P, x, side: int):
    """``HalfChannelConv5ResBlock`` -- modules/basic_blocks.py:44-73."""
    out = _prelu(p, "conv1.1", _conv(p, "conv1.0", x))
    out = torch.cat((out[:, :-side], _prelu(p, "conv2.1", _conv(p, "conv2.0", out[:, -side:]))), 1)
    out = _prelu(p, "conv3.1", _conv(p, "conv3.0", out))
    out = torch.cat((out[:, :-side], _prelu(p, "conv4.1", _conv(p, "conv4.0", out[:, -side:]))), 1)
    return _prelu(p, "prelu", x + _conv(p, "conv5", out))


def resize(x, scale_factor):
    """BaseModel.resize (models/BaseModel.py:19-22)."""
    return F.interpolate(x, scale_factor=scale_factor, recompute_scale_factor=False, mode="bilinear", align_corners=True)


def _resize2(x):
    return resize(x, 2.0)


def dat_block(p: _P, feat_t, feat0, feat1, ft0, ft1, n_samples, n_groups, n_heads, offset_scale,
              pred_res_flow=True):
    """``CrossDeformableAttentionBlockwFlow.forward`` -- modules/deformable_attn.py:103-124."""
    c = feat_t.shape[1]

    def movement(feat_x, ftx):                        # lines 77-79
        h = torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1)
        m = p.sub("movement_extractor")
        h = _prelu(m, "0.1", _conv(m, "0.0", h))
        h = _prelu(m, "1.1", _conv(m, "1.0", h))
        return _res5(m.sub("2"), h, c // 2)

    mv0 = movement(feat0, ft0)
    kv0 = dsample(feat0, ref_coords(_conv(p, "conv_res_offset", mv0), ft0, n_groups, n_samples, offset_scale),
                  n_groups, n_samples)
    mv1 = movement(feat1, ft1)
    kv1 = dsample(feat1, ref_coords(_conv(p, "conv_res_offset", mv1), ft1, n_groups, n_samples, offset_scale),
                  n_groups, n_samples)
    a = attn(p.sub("attn"), feat_t, torch.cat((kv0, kv1), dim=2), n_heads)
    mlp = p.sub("mlp")
    out = a + _conv(mlp, "fc2", F.gelu(_conv(mlp, "fc1", a, padding=0)), padding=0)
    if not pred_res_flow:
        return out
    up0 = _deconv(p, "conv_res_flow", mv0) + 2.0 * _resize2(ft0)
    up1 = _deconv(p, "conv_res_flow", mv1) + 2.0 * _resize2(ft1)
    return out, up0, up1


def layer_norm_c(p: _P, name: str, x):
    """LayerNorm over the channel dim of an NCHW map -- modules/madat.py:163."""
    w, b = p[name + ".weight"], p[name + ".bias"]
    return F.layer_norm(x.permute(0, 2, 3, 1), (x.shape[1],), w, b).permute(0, 3, 1, 2).contiguous()


def madat_ref_feats(p: _P, feat, flow, n_groups: int, n_samples: int):
    """``DATwithMotionEstimation._get_ref_feats`` -- modules/madat.py:161-174.

    Returns (samples (b,c,nS,hw), grid (b, nG*nS, hw, 2)).
    """
    b = feat.shape[0]
    feat = layer_norm_c(p, "norm1", feat)
    samples, gxv, gyv = _sample_groups(feat, flow, n_groups, n_samples)
    grid = torch.stack((gxv, gyv), dim=-1).view(b, n_groups * n_samples, -1, 2)
    return samples, grid


def madat_attn(p: _P, q, kv, n_heads: int):
    """``DeformAttnwMotion.forward`` -- modules/madat.py:86-115 (nn.Linear projections).

    Returns (out, softmax(logits[:, :S/2]), softmax(logits[:, S/2:])) -- per-frame softmaxes of the
    scaled logits.
    """
    b, c, fh, fw = q.shape
    S = kv.shape[2]
    out_c = p["q_proj.weight"].shape[0]
    hc = out_c // n_heads
    hw = fh * fw
    qf = F.linear(q.reshape(b, c, hw).transpose(1, 2), p["q_proj.weight"], p["q_proj.bias"])         # b, hw, C
    kvf = kv.permute(0, 2, 3, 1)                                                                    # b, S, hw, c
    k = F.linear(kvf, p["k_proj.weight"], p["k_proj.bias"])
    v = F.linear(kvf, p["v_proj.weight"], p["v_proj.bias"])
    qh = qf.view(b, hw, n_heads, hc).permute(0, 2, 3, 1).reshape(b * n_heads, hc, hw)
    kh = k.view(b, S, hw, n_heads, hc).permute(0, 3, 4, 1, 2).reshape(b * n_heads, hc, S, hw)
    vh = v.view(b, S, hw, n_heads, hc).permute(0, 3, 4, 1, 2).reshape(b * n_heads, hc, S, hw)
    logits = torch.einsum("bcd,bcsd->bsd", qh, kh).mul(hc ** -0.5)
    score = F.softmax(logits, dim=1)
    out = torch.einsum("bsd,bcsd->bcd", score, vh).contiguous().view(b, out_c, fh, fw)
    l0, l1 = torch.chunk(logits, 2, dim=1)
    return out, F.softmax(l0, dim=1), F.softmax(l1, dim=1)


# ----------------------------------------------------------------------------------------------
# the DAT model forward (models/DAT.py:74-119) as a function of a reference state dict
# ----------------------------------------------------------------------------------------------

def dat_config(nf=72, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2):
    """configs/DAT.yaml:4-8 plus the per-level constants hard-coded at models/DAT.py:22-31."""
    return dict(nf=nf, enc_res_blocks=enc_res_blocks, dec_res_blocks=dec_res_blocks, mlp_ratio=mlp_ratio,
                levels=dict(lv3=dict(n_samples=8, n_groups=4, n_heads=4, offset_scale=2.0),
                            lv2=dict(n_samples=16, n_groups=8, n_heads=8, offset_scale=4.0),
                            lv1=dict(n_samples=32, n_groups=8, n_heads=8, offset_scale=8.0)))


def _resblock(p: _P, x):
    """``ResBlocknoBN`` -- modules/basic_blocks.py:13-33."""
    return x + _conv(p, "conv2", _prelu(p, "prelu", _conv(p, "conv1", x)))


def _encoder(p: _P, x, n_res: int):
    """``SameChannelResEncoder`` -- modules/cnn_encoders.py:6-42."""
    pr = p.sub("projection")
    h = _prelu(pr, "1", _conv(pr, "0", x))
    h = _prelu(pr, "3", _conv(pr, "2", h, stride=2))
    for i in range(n_res):
        h = _resblock(pr.sub(str(4 + i)), h)
    feats = [h]
    for lvl in ("fea_L2_conv", "fea_L3_conv", "fea_L4_conv"):
        q = p.sub(lvl)
        h = _prelu(q, "1", _conv(q, "0", h, stride=2))
        h = _prelu(q, "3", _conv(q, "2", h))
        feats.append(h)
    return feats


def _query_builder(p: _P, feat0, feat1, t):
    """``DCNInterFeatBuilderwithT.forward`` -- modules/query_builder.py:28-36."""
    _, _, fh, fw = feat0.shape
    tt = t.repeat(1, 1, fh, fw)
    cb = p.sub("convblock")

    def motion(a, b_, plane):
        h = _prelu(cb, "1", _conv(cb, "0", torch.cat((a, b_, plane), 1)))
        return _prelu(cb, "3", _conv(cb, "2", h))

    m01 = motion(feat0, feat1, tt)
    m10 = motion(feat1, feat0, 1 - tt)
    f0, ft0 = dcn_block(p.sub("dcnt0"), feat0, m01)
    f1, ft1 = dcn_block(p.sub("dcnt1"), feat1, m10)
    bb = p.sub("blendblock")
    out = _conv(bb, "2", _prelu(bb, "1", _conv(bb, "0", torch.cat((f0, f1), 1))))
    return out, ft0, ft1


def _generator(p: _P, feat, mean_, n_res: int):
    """``BasicResPixelShuffleGenerator.forward`` -- modules/generator.py:19-24."""
    h = feat
    for i in range(n_res):
        h = _resblock(p.sub(f"reconstruction.{i}"), h)
    h = _prelu(p, "prelu1", F.pixel_shuffle(_conv(p, "upconv1", h), 2))
    h = _prelu(p, "prelu2", _conv(p, "HRconv", h))
    h = _conv(p, "conv_last", h)
    return torch.clamp(h + mean_, 0, 1)


def dat_forward(sd: Dict[str, torch.Tensor], x0, x1, t, cfg: Optional[dict] = None, return_flows: bool = False):
    """``DATwConstantnCv1.inference(x0, x1, t)`` -- models/DAT.py:74-92,118-119."""
    cfg = cfg or dat_config()
    nf = cfg["nf"]
    p = _P(sd)
    mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
    x0, x1 = x0 - mean_, x1 - mean_                                       # models/BaseModel.py:14-17
    f0 = _encoder(p.sub("feature_encoder"), x0, cfg["enc_res_blocks"])
    f1 = _encoder(p.sub("feature_encoder"), x1, cfg["enc_res_blocks"])
    q4, ft0_4, ft1_4 = _query_builder(p.sub("coarse_query_builder"), f0[3], f1[3], t)
    s3 = _deconv(p, "lv4_to_lv3", torch.cat((q4, ft0_4, ft1_4), dim=1))
    q3, ft0_3, ft1_3 = s3[:, :nf], s3[:, nf:nf + 2], s3[:, nf + 2:nf + 4]
    a3, ft0_2, ft1_2 = dat_block(p.sub("dat_lv3"), q3, f0[2], f1[2], ft0_3, ft1_3, **cfg["levels"]["lv3"])
    q2 = _deconv(p, "lv3_to_lv2", a3)
    a2, ft0_1, ft1_1 = dat_block(p.sub("dat_lv2"), q2, f0[1], f1[1], ft0_2, ft1_2, **cfg["levels"]["lv2"])
    q1 = _deconv(p, "lv2_to_lv1", a2)
    a1 = dat_block(p.sub("dat_lv1"), q1, f0[0], f1[0], ft0_1, ft1_1, pred_res_flow=False, **cfg["levels"]["lv1"])
    img = _generator(p.sub("pixel_geneartor"), a1, mean_, cfg["dec_res_blocks"])
    if return_flows:
        return img, dict(ft0=[ft0_1, ft0_2, ft0_3, ft0_4], ft1=[ft1_1, ft1_2, ft1_3, ft1_4])
    return img


# ----------------------------------------------------------------------------------------------
# DCNDATv1 (models/archive/DCNDAT.py), read as repaired by SURVEY.md Note B
# ----------------------------------------------------------------------------------------------

def dcndat_config(nf=64, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2):
    """configs/archive/DCNDAT.yaml:4-8 + the per-level constants of models/archive/DCNDAT.py:209-218."""
    return dict(nf=nf, enc_res_blocks=enc_res_blocks, dec_res_blocks=dec_res_blocks, mlp_ratio=mlp_ratio,
                levels=dict(scale3=dict(n_samples=9, n_groups=8, n_heads=8), scale2=dict(n_samples=9, n_groups=4, n_heads=4),
                            scale1=dict(n_samples=9, n_groups=4, n_heads=4)))


def dcndat_block(p: _P, feat_t, feat0, feat1, ft0, ft1, n_samples, n_groups, n_heads, pred_res_flow=True):
    """``CrossDeformableAttentionBlockwFlow.forward`` of the DCN flavour -- models/archive/DCNDAT.py:146-199."""
    b, c, fh, fw = feat_t.shape

    def movement(feat_x, ftx):                        # lines 146-148
        h = torch.cat((feat_t, bwarp(feat_x, ftx), ftx), dim=1)
        m = p.sub("conv_res_feat")
        h = _prelu(m, "0.1", _conv(m, "0.0", h))
        h = _prelu(m, "1.1", _conv(m, "1.0", h))
        return _res5(m.sub("2"), h, c // 2)

    def coords_mask(ftx, mv):                          # lines 150-157
        o1, o2, mask = torch.chunk(_conv(p, "conv_res_offset_mask", mv), 3, dim=1)
        res = 2.0 * torch.tanh(torch.cat((o1, o2), dim=1))
        return res.view(b, n_groups * n_samples, 2, fh, fw) + ftx.unsqueeze(1), mask

    mv0 = movement(feat0, ft0)
    c0, m0 = coords_mask(ft0, mv0)
    mv1 = movement(feat1, ft1)
    c1, m1 = coords_mask(ft1, mv1)
    w, bias = p["query_enhancer_dcn.regular_conv.weight"], p["query_enhancer_dcn.regular_conv.bias"]
    q = _conv(p, "query_blender", torch.cat((dcn_given_offset(w, bias, feat0, c0, m0), feat_t,
                                             dcn_given_offset(w, bias, feat1, c1, m1)), dim=1), padding=0)
    kv = torch.cat((dsample(feat0, c0, n_groups, n_samples), dsample(feat1, c1, n_groups, n_samples)), dim=2)
    a = attn(p.sub("attn"), q, kv, n_heads)
    mlp = p.sub("mlp")
    out = a + _conv(mlp, "fc2", F.gelu(_conv(mlp, "fc1", a, padding=0)), padding=0)
    if not pred_res_flow:
        return out
    return out, _deconv(p, "conv_res_flow", mv0) + 2.0 * _resize2(ft0), _deconv(p, "conv_res_flow", mv1) + 2.0 * _resize2(ft1)


def dcndat_forward(sd: Dict[str, torch.Tensor], x0, x1, t, cfg: Optional[dict] = None):
    """``DCNDATv1.inference(x0, x1, t)`` -- models/archive/DCNDAT.py:231-251; the rgb mean is added at the end but the
    inputs are NOT shifted by it (line 233 is commented out upstream)."""
    cfg = cfg or dcndat_config()
    nf = cfg["nf"]
    p = _P(sd)
    mean_ = torch.cat((x0, x1), dim=2).mean(1, keepdim=True).mean(2, keepdim=True).mean(3, keepdim=True)
    f0 = _encoder(p.sub("cnn_encoder"), x0, cfg["enc_res_blocks"])
    f1 = _encoder(p.sub("cnn_encoder"), x1, cfg["enc_res_blocks"])
    # DCNInterFeatBuilderwithT with ONE shared DeformableConv2d (lines 81-110)
    qb = p.sub("dcn_feat_t_builder")
    _, _, fh, fw = f0[3].shape
    tt = t.repeat(1, 1, fh, fw)
    cb = qb.sub("convblock")

    def motion(a, b_, plane):
        return _prelu(cb, "3", _conv(cb, "2", _prelu(cb, "1", _conv(cb, "0", torch.cat((a, b_, plane), 1)))))

    from0, ft0_4 = dcn_block(qb.sub("dcn"), f0[3], motion(f0[3], f1[3], tt))
    from1, ft1_4 = dcn_block(qb.sub("dcn"), f1[3], motion(f1[3], f0[3], 1 - tt))
    bb = qb.sub("blendblock")
    q4 = _conv(bb, "2", _prelu(bb, "1", _conv(bb, "0", torch.cat((from0, from1), 1))))
    s3 = _deconv(p, "query_builder3", torch.cat((q4, ft0_4, ft1_4), dim=1))
    q3, ft0_3, ft1_3 = s3[:, :nf], s3[:, nf:nf + 2], s3[:, nf + 2:nf + 4]
    a3, ft0_2, ft1_2 = dcndat_block(p.sub("dat_scale3"), q3, f0[2], f1[2], ft0_3, ft1_3, **cfg["levels"]["scale3"])
    a2, ft0_1, ft1_1 = dcndat_block(p.sub("dat_scale2"), _deconv(p, "query_builder2", a3), f0[1], f1[1], ft0_2, ft1_2,
                                    **cfg["levels"]["scale2"])
    a1 = dcndat_block(p.sub("dat_scale1"), _deconv(p, "query_builder1", a2), f0[0], f1[0], ft0_1, ft1_1,
                      pred_res_flow=False, **cfg["levels"]["scale1"])
    return _generator(p.sub("generator"), a1, mean_, cfg["dec_res_blocks"])


# ----------------------------------------------------------------------------------------------
# validators (evaluate.py:15-84)
# ----------------------------------------------------------------------------------------------

def _gaussian(window_size: int, sigma: float) -> torch.Tensor:
    """evaluate.py:15-17."""
    import math
    g = torch.Tensor([math.exp(-(x - window_size // 2) ** 2 / float(2 * sigma ** 2)) for x in range(window_size)])
    return g / g.sum()


def calculate_ssim(img1, img2, window_size: int = 11, val_range: Optional[float] = None, full: bool = False):
    """``calculate_ssim`` -- evaluate.py:29-79 (3-D window of evaluate.py:20-26 without the ``.cuda()``)."""
    if val_range is None:
        L = (255 if torch.max(img1) > 128 else 1) - (-1 if torch.min(img1) < -0.5 else 0)
    else:
        L = val_range
    _, _, height, width = img1.size()
    real = min(window_size, height, width)
    w1 = _gaussian(real, 1.5).unsqueeze(1)
    w2 = w1.mm(w1.t())
    window = (w2.unsqueeze(2) @ w1.t()).expand(1, 1, real, real, real).contiguous().to(img1)
    a, b = img1.unsqueeze(1), img2.unsqueeze(1)

    def blur(x):
        return F.conv3d(F.pad(x, (5, 5, 5, 5, 5, 5), mode="replicate"), window, padding=0, groups=1)

    mu1, mu2 = blur(a), blur(b)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1, s2, s12 = blur(a * a) - mu1_sq, blur(b * b) - mu2_sq, blur(a * b) - mu1_mu2
    c1, c2 = (0.01 * L) ** 2, (0.03 * L) ** 2
    v1, v2 = 2.0 * s12 + c2, s1 + s2 + c2
    ssim = (((2 * mu1_mu2 + c1) * v1) / ((mu1_sq + mu2_sq + c1) * v2)).mean()
    return (ssim, torch.mean(v1 / v2)) if full else ssim


def calculate_psnr(img1, img2):
    """``calculate_psnr`` -- evaluate.py:82-84."""
    return -10 * torch.log10(((img1 - img2) * (img1 - img2)).mean())


def input_pad(x, divisor: int = 16) -> Tuple[torch.Tensor, Tuple[int, int, int, int]]:
    """``InputPadder.pad`` -- data/__init__.py:6-15 (replicate padding to a multiple of ``divisor``)."""
    ht, wd = x.shape[-2:]
    ph = (((ht // divisor) + 1) * divisor - ht) % divisor
    pw = (((wd // divisor) + 1) * divisor - wd) % divisor
    pad = (pw // 2, pw - pw // 2, ph // 2, ph - ph // 2)
    return F.pad(x, pad, mode="replicate"), pad


def input_unpad(x, pad):
    """``InputPadder.unpad`` -- data/__init__.py:17-20."""
    ht, wd = x.shape[-2:]
    return x[..., pad[2]:ht - pad[3], pad[0]:wd - pad[1]]


def count_params(sd) -> int:
    return sum(int(v.numel()) for v in sd.values())


def dat_param_shapes(cfg: Optional[dict] = None) -> Dict[str, tuple]:
    """Names and shapes of ``DATwConstantnCv1.state_dict()`` (checked against the reference in
    tests/test_oracle_golden.py::test_dat_param_table)."""
    cfg = cfg or dat_config()
    nf, hid = cfg["nf"], int(cfg["nf"] * cfg["mlp_ratio"])
    s: Dict[str, tuple] = {}

    def conv(name, co, ci, k=3):
        s[name + ".weight"] = (co, ci, k, k)
        s[name + ".bias"] = (co,)

    def prelu(name, c):
        s[name + ".weight"] = (c,)

    def resblock(name):
        conv(name + ".conv1", nf, nf); conv(name + ".conv2", nf, nf); prelu(name + ".prelu", nf)

    e = "feature_encoder."
    conv(e + "projection.0", nf, 3); prelu(e + "projection.1", nf)
    conv(e + "projection.2", nf, nf); prelu(e + "projection.3", nf)
    for i in range(cfg["enc_res_blocks"]):
        resblock(e + f"projection.{4 + i}")
    for lvl in ("fea_L2_conv", "fea_L3_conv", "fea_L4_conv"):
        conv(e + lvl + ".0", nf, nf); prelu(e + lvl + ".1", nf); conv(e + lvl + ".2", nf, nf); prelu(e + lvl + ".3", nf)
    q = "coarse_query_builder."
    conv(q + "convblock.0", nf, 2 * nf + 1); prelu(q + "convblock.1", nf)
    conv(q + "convblock.2", nf, nf); prelu(q + "convblock.3", nf)
    for d in ("dcnt0.", "dcnt1."):
        conv(q + d + "offset_flow_conv", 2, nf)
        conv(q + d + "conv_offset_mask.0", nf, 2 * nf + 2); prelu(q + d + "conv_offset_mask.1", nf)
        conv(q + d + "conv_offset_mask.2", nf, nf); prelu(q + d + "conv_offset_mask.3", nf)
        conv(q + d + "conv_offset_mask.4", 8 * 27, nf)
        conv(q + d + "regular_conv", nf, nf // 8)
    conv(q + "blendblock.0", nf, 2 * nf); prelu(q + "blendblock.1", nf); conv(q + "blendblock.2", nf, nf)
    s["lv4_to_lv3.weight"] = (nf + 4, nf + 4, 4, 4); s["lv4_to_lv3.bias"] = (nf + 4,)
    for lv, up in (("lv3", "lv3_to_lv2"), ("lv2", "lv2_to_lv1"), ("lv1", None)):
        L = cfg["levels"][lv]
        b = f"dat_{lv}."
        m = b + "movement_extractor."
        conv(m + "0.0", 2 * nf, 2 * nf + 2); prelu(m + "0.1", 2 * nf)
        conv(m + "1.0", nf, 2 * nf); prelu(m + "1.1", nf)
        for i, ch in ((1, nf), (2, nf // 2), (3, nf), (4, nf // 2)):
            conv(m + f"2.conv{i}.0", ch, ch); prelu(m + f"2.conv{i}.1", ch)
        conv(m + "2.conv5", nf, nf); prelu(m + "2.prelu", nf)
        conv(b + "conv_res_offset", L["n_groups"] * L["n_samples"] * 2, nf)
        if up is not None:
            s[b + "conv_res_flow.weight"] = (nf, 2, 4, 4); s[b + "conv_res_flow.bias"] = (2,)
        for n in ("q_proj", "k_proj", "v_proj"):
            conv(b + "attn." + n, nf, nf, 1)
        conv(b + "mlp.fc1", hid, nf, 1); conv(b + "mlp.fc2", nf, hid, 1)
        if up is not None:
            s[up + ".weight"] = (nf, nf, 4, 4); s[up + ".bias"] = (nf,)
    g = "pixel_geneartor."
    for i in range(cfg["dec_res_blocks"]):
        resblock(g + f"reconstruction.{i}")
    conv(g + "upconv1", 4 * nf, nf); prelu(g + "prelu1", nf)
    conv(g + "HRconv", nf, nf); prelu(g + "prelu2", nf); conv(g + "conv_last", 3, nf)
    return s


# ---- f2: 3x3 convolutions of the conv stacks in TF32 -------------------------------------------
def tf32_truncate(x: torch.Tensor) -> torch.Tensor:
    """fp32 -> TF32 by dropping the low 13 mantissa bits (what the tensor core does to an fp32 operand)."""
    return (x.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def tf32_round(x: torch.Tensor) -> torch.Tensor:
    """fp32 -> TF32 round-to-nearest, ties away from zero (cvt.rna.tf32.f32)."""
    return ((x.contiguous().view(torch.int32) + 0x1000) & ~0x1FFF).view(torch.float32)


def conv3x3_tf32(x, weight, bias=None, slope=None, residual=None, residual_after_act=False):
    """nn.Conv2d(k=3, s=1, p=1) [+ residual] [+ per-channel PReLU] as the reference's GPU runs it under PyTorch's
    default ``cudnn.allow_tf32 = True`` (modules/basic_blocks.py:13-73, modules/cnn_encoders.py:6-42,
    modules/generator.py:7-24): operands reduced to TF32 (activations truncated, weights rounded), products and sums
    exact (float64 here; fp32 accumulation order is the implementation's).  Returns float64."""
    y = F.conv2d(tf32_truncate(x).double(), tf32_round(weight).double(), None if bias is None else bias.double(), padding=1)
    if residual is not None and not residual_after_act:
        y = residual.double() + y
    if slope is not None:
        y = torch.where(y > 0, y, slope.double().view(1, -1, 1, 1) * y)
    if residual is not None and residual_after_act:
        y = residual.double() + y
    return y
