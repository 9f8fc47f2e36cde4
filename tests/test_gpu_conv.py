"""GPU parity of the tcgen05 3x3 convolution through the C ABI.

The kernel computes in TF32 (activations truncated by the tensor core, weights rounded to nearest) with fp32
accumulation -- PyTorch's default conv arithmetic on a GPU.  It is checked (a) against the oracle's exact TF32
restatement at 1e-5, and (b) against the fp32 convolution at a TF32-sized tolerance, on the shapes and epilogues the
DAT model uses, including ragged tiles, channel-slice inputs, Cout tiling and both output layouts."""
import pytest
import torch
import torch.nn as nn

import oracle
from _util import rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture()
def tf32():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    yield
    torch.backends.cudnn.allow_tf32 = old


def _case(B, Cin, Cout, H, W, prelu=False, res=0, ocl=True, slice_of=None, bias=True, seed=0, res_slice_of=None):
    from videoframeinterpolation_b200 import conv as vc
    g = torch.Generator().manual_seed(seed)
    conv = nn.Conv2d(Cin, Cout, 3, 1, 1, bias=bias)
    with torch.no_grad():
        conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * 0.1)
        if bias:
            conv.bias.copy_(torch.randn(Cout, generator=g))
    conv = conv.cuda()
    ctot = slice_of or Cin
    xfull = torch.randn(B, ctot, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    x = xfull[:, ctot - Cin:]
    slope = torch.rand(Cout, generator=g).cuda() if prelu else None
    r = None
    if res:      # optionally the trailing channels of a wider channels-last tensor (pixel stride > Cout)
        rtot = res_slice_of or Cout
        r = torch.randn(B, rtot, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)[:, rtot - Cout:]
    with torch.no_grad():
        assert vc.eligible(conv, x)
        y = vc.conv3x3(conv, x, slope, r, residual_after_act=(res == 2), out_channels_last=ocl)
    assert y.shape == (B, Cout, H, W)
    assert y.is_contiguous(memory_format=torch.channels_last) if (ocl and Cout > 1) else y.is_contiguous()
    ref = oracle.conv3x3_tf32(x.cpu(), conv.weight.detach().cpu(), None if not bias else conv.bias.detach().cpu(),
                              None if slope is None else slope.cpu(), None if r is None else r.cpu(), res == 2)
    assert rel_err(y.double().cpu(), ref) <= 1e-5
    fp32 = torch.nn.functional.conv2d(x.double().cpu(), conv.weight.detach().double().cpu(),
                                      None if not bias else conv.bias.detach().double().cpu(), padding=1)
    if not prelu and not res:
        assert rel_err(y.double().cpu(), fp32) <= 3e-3


@pytest.mark.parametrize("cfg", [
    dict(B=1, Cin=72, Cout=72, H=16, W=8),                               # exactly one tile
    dict(B=2, Cin=72, Cout=72, H=45, W=37, prelu=True, res=1),           # ragged tiles, ResBlock tail
    dict(B=1, Cin=72, Cout=72, H=33, W=20, prelu=True, res=2),
    dict(B=1, Cin=72, Cout=72, H=35, W=29, ocl=False),                   # planar output (feeds the gather kernels)
    dict(B=1, Cin=72, Cout=512, H=33, W=40, bias=False),                 # conv_res_offset: 8 Cout tiles
    dict(B=1, Cin=72, Cout=288, H=32, W=32),                             # upconv1
    dict(B=1, Cin=72, Cout=3, H=40, W=24, ocl=False),                    # conv_last
    dict(B=1, Cin=72, Cout=2, H=24, W=24),                               # flow head
    dict(B=1, Cin=36, Cout=36, H=32, W=24, prelu=True, slice_of=72),     # HalfChannelConv5ResBlock side conv (half chunk)
    dict(B=3, Cin=64, Cout=64, H=17, W=9, prelu=True),
    # residual tiles staged through shared memory by TMA: odd tile count (dummy tile in the last CTA pair), Cout tile not
    # a multiple of 8, residual = channel slice of a wider tensor, several Cout tiles, many tiles per CTA
    dict(B=1, Cin=72, Cout=72, H=42, W=24, res=1),
    dict(B=1, Cin=40, Cout=36, H=30, W=19, prelu=True, res=1),
    dict(B=2, Cin=72, Cout=72, H=29, W=17, prelu=True, res=2, res_slice_of=144),
    dict(B=1, Cin=72, Cout=144, H=31, W=23, res=1),
    dict(B=1, Cin=16, Cout=16, H=300, W=333, prelu=True, res=1),
    dict(B=1, Cin=72, Cout=72, H=20, W=12, res=1, ocl=False),            # planar residual: direct loads
])
def test_conv3x3_vs_oracle(tf32, cfg):
    _case(**cfg)


def test_conv3x3_full_size_properties(tf32):
    """1/2-res 1080p shape: linearity in the input and agreement with cuDNN's TF32 convolution."""
    from videoframeinterpolation_b200 import conv as vc
    torch.manual_seed(1)
    conv = nn.Conv2d(72, 72, 3, 1, 1).cuda()
    x = torch.randn(1, 72, 544, 960, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        y = vc.conv3x3(conv, x)
        y2 = vc.conv3x3(conv, 2.0 * x)              # scaling by 2 is exact in TF32
        b = conv.bias.view(1, -1, 1, 1)
        assert torch.equal(y2 - b, 2.0 * (y - b)) or rel_err(y2 - b, 2.0 * (y - b)) <= 1e-6
        ref = conv(x)
        assert rel_err(y, ref) <= 2e-3


def test_conv3x3_not_used_when_tf32_is_off():
    from videoframeinterpolation_b200 import conv as vc
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    try:
        conv = nn.Conv2d(72, 72, 3, 1, 1).cuda()
        x = torch.randn(1, 72, 16, 16, device="cuda").contiguous(memory_format=torch.channels_last)
        with torch.no_grad():
            assert not vc.eligible(conv, x)
    finally:
        torch.backends.cudnn.allow_tf32 = old


def test_dat_forward_tc_convs_match_fp32_convs(tf32):
    """Whole DAT forward (channels-last) with the tensor-core convs vs the same model with fp32 cuDNN convs."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import init_dat_weights
    torch.manual_seed(0)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).cuda().eval().to(memory_format=torch.channels_last)
    g = torch.Generator().manual_seed(3)
    x0 = torch.rand(1, 3, 128, 192, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    x1 = torch.rand(1, 3, 128, 192, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    t = torch.full((1, 1, 1, 1), 0.5, device="cuda")
    from videoframeinterpolation_b200 import _lib
    with torch.no_grad():
        _lib.reset_launch_count()
        y_tc = net.inference(x0, x1, t)
        n_tc = _lib.launch_count()
        torch.backends.cudnn.allow_tf32 = False
        _lib.reset_launch_count()
        y_fp32 = net.inference(x0, x1, t)
        n_fp32 = _lib.launch_count()
    assert n_tc != n_fp32, "the tensor-core conv path did not run"
    assert float((y_tc - y_fp32).abs().max()) <= 2e-2
    assert float((y_tc - y_fp32).abs().mean()) <= 2e-3


def test_dat_forward_measured_mode_vs_reference_ops_with_tf32(tf32):
    """The configuration bench.py times -- tcgen05 3x3 convs, split movement convs, no-concat HalfChannel block, fused
    PixelShuffle, ConvTranspose on the tensor-core kernel (>= 65536 px at 1/4 res needs a >= 1 Mpx frame), channels-last
    tensors, window-form attention -- at 1024x1024 against the reference's op composition on the PyTorch-CUDA library ops
    WITH cuDNN TF32 (what the reference itself runs on a GPU), and against the fp32 oracle composition.
    Both sides round conv operands to TF32, in different places, so the bound is a PSNR / mean-error bound."""
    import math
    import oracle
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import _lib
    from videoframeinterpolation_b200.workloads import init_dat_weights
    torch.manual_seed(0)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).cuda().eval().to(memory_format=torch.channels_last)
    g = torch.Generator().manual_seed(4)
    x0 = torch.rand(1, 3, 1024, 1024, generator=g).cuda()
    x1 = torch.rand(1, 3, 1024, 1024, generator=g).cuda()
    t = torch.full((1, 1, 1, 1), 0.5, device="cuda")
    cl = lambda a: a.contiguous(memory_format=torch.channels_last)
    prev_mm = torch.backends.cuda.matmul.allow_tf32
    with torch.no_grad():
        with _lib.OpTimer() as tm:
            y = net.inference(cl(x0), cl(x1), t).float()
        calls = {k: len(e) for k, e in tm.records.items()}
        assert calls.get("vfi_conv3x3_f32", 0) >= 60 and calls.get("vfi_dattn_f32", 0) == 3, calls
        sd = {k: w.detach().clone() for k, w in net.state_dict().items()}
        with oracle.library_ops():
            ref_tf32 = oracle.dat_forward(sd, x0, x1, t)            # cuDNN TF32 convs (allow_tf32 is on in this fixture)
            torch.backends.cudnn.allow_tf32 = False
            torch.backends.cuda.matmul.allow_tf32 = False
            ref_fp32 = oracle.dat_forward(sd, x0, x1, t)
    torch.backends.cuda.matmul.allow_tf32 = prev_mm
    def stats(a, b):
        d = (a - b).abs()
        return float(d.max()), float(d.mean()), -10.0 * math.log10(max(float((d.double() ** 2).mean()), 1e-30))
    mx, mean, psnr = stats(y, ref_fp32)
    mx_r, mean_r, psnr_r = stats(ref_tf32, ref_fp32)
    print(f"measured mode vs fp32 reference ops: max {mx:.2e} mean {mean:.2e} psnr {psnr:.1f} dB; "
          f"reference ops with cuDNN TF32 vs fp32: max {mx_r:.2e} mean {mean_r:.2e} psnr {psnr_r:.1f} dB")
    assert mean <= 2e-3 and psnr >= 50.0, (mx, mean, psnr)
    # no worse than a small multiple of what the reference's own TF32 arithmetic does to its fp32 result
    assert mean <= 4.0 * mean_r + 1e-4, (mean, mean_r)


@pytest.mark.parametrize("shape", [(1, 72, 288, 20, 12), (2, 72, 288, 15, 9), (1, 32, 64, 16, 8)])
def test_conv3x3_fused_pixel_shuffle(tf32, shape):
    """upconv1 -> PixelShuffle(2) -> PReLU of the generator (modules/generator.py:14-22) in one kernel."""
    from videoframeinterpolation_b200 import conv as vc
    B, Cin, Cout, H, W = shape
    g = torch.Generator().manual_seed(5)
    conv = nn.Conv2d(Cin, Cout, 3, 1, 1)
    with torch.no_grad():
        conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * 0.1)
        conv.bias.copy_(torch.randn(Cout, generator=g))
    conv = conv.cuda()
    x = torch.randn(B, Cin, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    slope = torch.rand(Cout // 4, generator=g).cuda()
    with torch.no_grad():
        assert vc.eligible(conv, x, upscale=2)
        y = vc.conv3x3(conv, x, slope, upscale=2)
    assert y.shape == (B, Cout // 4, 2 * H, 2 * W) and y.is_contiguous(memory_format=torch.channels_last)
    ref = torch.nn.functional.pixel_shuffle(oracle.conv3x3_tf32(x.cpu(), conv.weight.detach().cpu(), conv.bias.detach().cpu()), 2)
    ref = torch.where(ref > 0, ref, slope.double().cpu().view(1, -1, 1, 1) * ref)
    assert rel_err(y.double().cpu(), ref) <= 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("shape", [(1, 72, 36, 40, 27), (2, 32, 16, 19, 33)])
def test_half_channel_block_without_concat(tf32, shape, monkeypatch):
    """HalfChannelConv5ResBlock: the rolled-output / slice-input route equals the route with the two torch.cat."""
    from videoframeinterpolation_b200.blocks import HalfChannelConv5ResBlock
    B, C, S, H, W = shape
    torch.manual_seed(3)
    blk = HalfChannelConv5ResBlock(C, S).cuda()
    with torch.no_grad():
        for m in blk.modules():
            if isinstance(m, nn.PReLU):
                m.weight.copy_(torch.rand_like(m.weight))
            if isinstance(m, nn.Conv2d):
                m.bias.copy_(torch.randn_like(m.bias))
    x = torch.randn(B, C, H, W, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        assert blk._fused_ok(x)
        y = blk(x)
        monkeypatch.setattr(HalfChannelConv5ResBlock, "_fused_ok", lambda self, t: False)
        ref = blk(x)
    assert y.shape == ref.shape
    assert torch.equal(y, ref) or rel_err(y, ref) <= 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [(1, 72, 72, 260, 256, True), (2, 72, 2, 256, 257, False), (1, 76, 76, 300, 220, True), (1, 72, 72, 20, 13, True)])
def test_conv_transpose_up2_on_tensor_cores(tf32, cfg):
    """ConvTranspose2d(4, 2, 1) through the 3x3 kernel with the fused PixelShuffle vs PyTorch's own (fp64 on the CPU)."""
    from videoframeinterpolation_b200 import conv as vc
    B, cin, cout, H, W, ocl = cfg
    torch.manual_seed(5)
    ct = nn.ConvTranspose2d(cin, cout, 4, 2, 1).cuda()
    x = torch.randn(B, cin, H, W, device="cuda").contiguous(memory_format=torch.channels_last)
    with torch.no_grad():
        assert vc.eligible(vc.transpose_as_conv(ct), x, upscale=2)
        y = vc.conv_transpose_up2(ct, x, out_channels_last=ocl)
        ref = nn.functional.conv_transpose2d(x.double().cpu(), ct.weight.double().cpu(), ct.bias.double().cpu(), 2, 1)
    assert y.shape == ref.shape
    assert y.is_contiguous() if not ocl else y.is_contiguous(memory_format=torch.channels_last)
    assert rel_err(y.double().cpu(), ref) <= 3e-3      # TF32 inputs


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(B=2, Cin=72, Cout=72, H=30, W=20, res=0), dict(B=2, Cin=72, Cout=72, H=31, W=21, res=1),
                                 dict(B=1, Cin=72, Cout=64, H=17, W=27, res=0), dict(B=3, Cin=32, Cout=8, H=15, W=9, res=0),
                                 dict(B=1, Cin=72, Cout=136, H=29, W=18, res=0)])
def test_conv3x3_tma_store_stays_inside_the_output_view(tf32, cfg):
    """Round-2 epilogue (output tile in shared memory -> one TMA store per tile; the shared-memory tiles are wider than the
    channel count, csrc/conv3x3.cu:conv_tile_pitch): the store must drop the padding channels, the rows / columns of ragged
    edge tiles and the dummy tile of an odd tile count.  The output is a channel slice in the middle of a wider
    channels-last buffer with guard bands before and after it; every element outside the view must keep its sentinel."""
    from videoframeinterpolation_b200 import conv as vc
    B, Cin, Cout, H, W, res = (cfg[k] for k in ("B", "Cin", "Cout", "H", "W", "res"))
    g = torch.Generator().manual_seed(11)
    conv = nn.Conv2d(Cin, Cout, 3, 1, 1)
    with torch.no_grad():
        conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * 0.1)
        conv.bias.copy_(torch.randn(Cout, generator=g))
    conv = conv.cuda()
    x = torch.randn(B, Cin, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    r = torch.randn(B, Cout, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last) if res else None
    lead, trail, ctot = 4, 8, Cout + 12
    guard = 4096
    SENT = 12345.0
    flat = torch.full((guard + B * H * W * ctot + guard,), SENT, device="cuda")
    buf = flat[guard:guard + B * H * W * ctot].view(B, H, W, ctot).permute(0, 3, 1, 2)
    out = buf[:, lead:lead + Cout]
    with torch.no_grad():
        y = vc.conv3x3(conv, x, None, r, out=out)
    torch.cuda.synchronize()
    ref = oracle.conv3x3_tf32(x.cpu(), conv.weight.detach().cpu(), conv.bias.detach().cpu(), None, None if r is None else r.cpu(), False)
    assert rel_err(y.double().cpu(), ref) <= 1e-5
    assert bool((flat[:guard] == SENT).all()) and bool((flat[-guard:] == SENT).all()), "wrote outside the buffer"
    assert bool((buf[:, :lead] == SENT).all()) and bool((buf[:, lead + Cout:] == SENT).all()), "wrote outside the channel slice"


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [dict(B=4, Br=2, C=72, H=30, W=20, after=False), dict(B=2, Br=1, C=72, H=17, W=27, after=True),
                                 dict(B=6, Br=3, C=32, H=15, W=9, after=False)])
def test_conv3x3_residual_shared_by_groups_of_the_batch(tf32, cfg):
    """residual_mode bits 8..: output image b adds residual image b % Br (the feat_t term of the movement extractor's first
    conv, shared by the two frames of a pair batch) -- staged (TMA) and direct residual paths, against the expanded residual."""
    from videoframeinterpolation_b200 import conv as vc
    B, Br, C, H, W = (cfg[k] for k in ("B", "Br", "C", "H", "W"))
    g = torch.Generator().manual_seed(5)
    conv = nn.Conv2d(C, C, 3, 1, 1)
    with torch.no_grad():
        conv.weight.copy_(torch.randn(conv.weight.shape, generator=g) * 0.1)
        conv.bias.copy_(torch.randn(C, generator=g))
    conv = conv.cuda()
    x = torch.randn(B, C, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    r = torch.randn(Br, C, H, W, generator=g).cuda().contiguous(memory_format=torch.channels_last)
    slope = torch.rand(C, generator=g).cuda()
    with torch.no_grad():
        y = vc.conv3x3(conv, x, slope, r, residual_after_act=cfg["after"])
        full = vc.conv3x3(conv, x, slope, r.repeat(B // Br, 1, 1, 1).contiguous(memory_format=torch.channels_last),
                          residual_after_act=cfg["after"])
        yp = vc.conv3x3(conv, x, slope, r, residual_after_act=cfg["after"], out_channels_last=False)     # direct residual path
    assert torch.equal(y, full)
    assert rel_err(yp, full) <= 1e-6
