"""GPU parity: the CUDA path (through the C ABI) vs the golden fixtures and vs the oracle on seeded inputs."""
import pytest
import torch

import _det as det
import golden_cases as gc
import oracle
from oracle.torch_oracle import _P
from _util import assert_close, rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-5          # warp / DCN / sampler / attention (north_star)
TOL_SPLAT = 1e-4    # splat: summation order differs


@pytest.fixture(scope="module")
def vfi():
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import _lib
    assert _lib.lib().vfi_device_supported() == 1, "needs an sm_100 device"
    return v


def cu(*ts):
    r = tuple(None if t is None else t.cuda() for t in ts)
    return r if len(r) > 1 else r[0]


# ---- bwarp -----------------------------------------------------------------------------------
def test_bwarp_golden(vfi, golden):
    g = golden("bwarp")
    for key, (img, flow) in gc.bwarp_cases().items():
        assert_close(vfi.bwarp(*cu(img, flow)), g[key], TOL, f"bwarp[{key}]")


@pytest.mark.parametrize("shape", [(1, 3, 270, 480), (2, 64, 68, 120), (1, 7, 33, 65), (1, 1, 16, 16)])
def test_bwarp_vs_oracle(vfi, shape):
    B, C, H, W = shape
    img = det.normal(f"bwo.{shape}", shape)
    flow = det.flow_field(f"bwo.f.{shape}", B, H, W, sigma=4.0)
    assert_close(vfi.bwarp(*cu(img, flow)), oracle.bwarp(img, flow), TOL, f"bwarp{shape}")


def test_bwarp_channels_last_and_strided_flow(vfi):
    B, C, H, W = 2, 72, 34, 60
    img = det.normal("bwcl.img", (B, C, H, W))
    big = det.flow_field("bwcl.flow", B, H, W, sigma=3.0)
    packed = torch.cat((det.normal("bwcl.pad", (B, 5, H, W)), big), 1)        # flow as a channel slice (models/DAT.py:82)
    ref = oracle.bwarp(img, big)
    out_cl = vfi.bwarp(img.cuda().contiguous(memory_format=torch.channels_last), packed.cuda()[:, 5:7])
    assert out_cl.is_contiguous(memory_format=torch.channels_last)
    assert_close(out_cl, ref, TOL, "bwarp channels_last")
    assert_close(vfi.bwarp(img.cuda(), packed.cuda()[:, 5:7]), ref, TOL, "bwarp strided flow")


def test_bwarp_identity_and_idempotence_1080p(vfi):
    """Size-independent properties at BASELINE size: zero flow is the identity; integer shifts commute with slicing."""
    img = torch.rand(1, 16, 1080, 1920, device="cuda")
    zero = torch.zeros(1, 2, 1080, 1920, device="cuda")
    # NOT exactly the identity: the reference's fp32 normalise/unnormalise round trip moves coordinates by up to
    # ~2e-4 px at this width; the kernel must reproduce that, so compare with the oracle below too
    assert rel_err(vfi.bwarp(img, zero), img) <= 1e-3
    assert_close(vfi.bwarp(img[:, :2], zero), oracle.bwarp(img[:, :2].cpu(), zero.cpu()), TOL, "bwarp zero flow 1080p vs oracle")
    shift = zero.clone()
    shift[:, 0] = 3.0
    shift[:, 1] = -2.0
    out = vfi.bwarp(img, shift)
    # away from the borders an integer flow is an exact shift up to the fp32 round trip of the coordinates
    assert rel_err(out[..., 8:-8, 8:-8], img[..., 6:-10, 11:-5]) <= 2e-3
    lin = oracle.bwarp(img[:, :2].cpu(), shift.cpu())
    assert_close(out[:, :2], lin, TOL, "bwarp 1080p vs oracle (2 channels)")


@pytest.mark.parametrize("algo", ["tile", "gather"])
@pytest.mark.parametrize("C", [8, 16, 64])
@pytest.mark.parametrize("kind", ["smooth", "gentle", "iid", "edge"])
def test_bwarp_tile_and_gather_kernels_vs_oracle(vfi, monkeypatch, algo, C, kind):
    """Both NCHW kernels of vfi_bwarp_f32 forced by VFI_BWARP_ALGO at a size that really dispatches the TMA tile kernel
    (H*W >= 64K, W % 4 == 0, C >= 8; the 64x16 tiles are ragged in both directions at 270x480), on the bench's flow
    kinds WITH their 2 % outliers (taps outside the 96x32 window -> the tile kernel's global-gather tail pass), and on
    an 'edge' flow whose median shift pushes the window past each frame border (the window origin clamps + the
    16-byte rounding of the TMA x origin).  Reference: modules/warp.py:7-15 through the oracle."""
    from videoframeinterpolation_b200.workloads import make_flow
    monkeypatch.setenv("VFI_BWARP_ALGO", algo)
    B, H, W = 1, 270, 480
    gen = torch.Generator().manual_seed(100 + C)
    img = torch.randn(B, C, H, W, generator=gen)
    if kind == "edge":
        flow = make_flow(B, H, W, 1.0, gen, kind="smooth", outlier_frac=0.01)
        flow[:, 0, :, : W // 2] -= 300.0       # left half samples far left of the frame, right half far right
        flow[:, 0, :, W // 2:] += 300.0
        flow[:, 1, : H // 2] -= 200.0          # top half above, bottom half below
        flow[:, 1, H // 2:] += 200.0
        flow[:, :, 100:140, 200:280] = make_flow(B, 40, 80, 6.0, gen, kind="iid", outlier_frac=0.0)   # an interior patch
    else:
        flow = make_flow(B, H, W, 8.0, gen, kind=kind)
    if algo == "tile" and kind != "edge":
        # the out-of-window tail is provably hit: some pixel's displacement differs from every probe of its tile by
        # more than the window margin (2 % of the pixels have flows uniform in [-W, W])
        assert float((flow.abs() > 64).float().mean()) > 0.005
    assert_close(vfi.bwarp(*cu(img, flow)), oracle.bwarp(img, flow), TOL, f"bwarp {algo} kernel C={C} {kind}")


@pytest.mark.parametrize("layout", ["nchw_tile", "nchw_gather", "channels_last"])
def test_bwarp_on_a_crop_samples_like_the_whole_frame(vfi, monkeypatch, layout):
    """frame_ext (Wg, Hg, x0, y0): bwarp of a crop, told the frame's extent and the crop's origin, equals
    the crop of bwarp of the whole frame wherever the taps stay inside the crop -- normalisation (modules/warp.py:9-12)
    and border clamp are the frame's.  Without frame_ext the same call differs (the crop's own W-1 / H-1)."""
    from videoframeinterpolation_b200 import _lib
    monkeypatch.setenv("VFI_BWARP_ALGO", "tile" if layout == "nchw_tile" else "gather")
    Hg, Wg, C = 544, 960, 16
    y0, x0, Hc, Wc = 128, 320, 288, 512
    img = det.normal("crop.img", (1, C, Hg, Wg))
    flow = det.flow_field("crop.flow", 1, Hg, Wg, sigma=5.0, outlier_frac=0.0)
    whole = oracle.bwarp(img, flow)[:, :, y0:y0 + Hc, x0:x0 + Wc]
    ci, cf = img[:, :, y0:y0 + Hc, x0:x0 + Wc].contiguous().cuda(), flow[:, :, y0:y0 + Hc, x0:x0 + Wc].contiguous().cuda()
    if layout == "channels_last":
        ci = ci.contiguous(memory_format=torch.channels_last)
    with _lib.crop_context(Hg, Wg, y0, x0, Hc, Wc):
        out = vfi.bwarp(ci, cf)
    m = 32       # |flow| < 6 sigma = 30 px: taps of interior pixels stay inside the crop
    assert_close(out[:, :, m:-m, m:-m], whole[:, :, m:-m, m:-m], TOL, f"bwarp crop vs whole ({layout})")
    plain = vfi.bwarp(ci, cf)
    assert rel_err(plain[:, :, m:-m, m:-m], whole[:, :, m:-m, m:-m]) > 1e-4, "the crop's own extent gives a different warp"


def test_sampler_on_a_crop_samples_like_the_whole_frame(vfi):
    """frame_ext for the grouped sampler (modules/deformable_attn.py:88-101 on the whole frame vs on a crop)."""
    from videoframeinterpolation_b200 import _lib
    from videoframeinterpolation_b200.deformable_attn import grouped_sample
    Hg, Wg, C, nG, nS = 96, 160, 16, 4, 3
    y0, x0, Hc, Wc = 32, 48, 48, 80
    feat = det.normal("cs.feat", (1, C, Hg, Wg))
    coords = det.normal("cs.coords", (1, nG * nS, 2, Hg, Wg), std=3.0)
    with oracle.library_ops():
        whole = oracle.dsample(feat, coords, nG, nS).view(1, C, nS, Hg, Wg)[..., y0:y0 + Hc, x0:x0 + Wc]
    with _lib.crop_context(Hg, Wg, y0, x0, Hc, Wc):
        out = grouped_sample(feat[..., y0:y0 + Hc, x0:x0 + Wc].contiguous().cuda(),
                             coords[..., y0:y0 + Hc, x0:x0 + Wc].contiguous().cuda(), nG, nS).view(1, C, nS, Hc, Wc)
    m = 20
    assert_close(out[..., m:-m, m:-m], whole[..., m:-m, m:-m], TOL, "dsample crop vs whole")


def test_flow_resize_matches_interpolate_and_crops(vfi):
    """vfi_resize2x_f32 = res + 2 * BaseModel.resize(ft, 2) (models/BaseModel.py:19-22, modules/deformable_attn.py:127-131),
    on a whole frame and on a crop with frame_ext."""
    from videoframeinterpolation_b200 import _lib
    from videoframeinterpolation_b200.blocks import flow_up2_add
    Hg, Wg = 64, 112
    ft = det.normal("rs.ft", (2, 2, Hg, Wg), std=4.0)
    res = det.normal("rs.res", (2, 2, 2 * Hg, 2 * Wg))
    ref = res + 2.0 * oracle.resize(ft, 2.0)
    with torch.no_grad():
        assert_close(flow_up2_add(res.cuda(), ft.cuda()), ref, 1e-6, "flow_up2_add")
        packed = torch.cat((det.normal("rs.pad", (2, 3, Hg, Wg)), ft), 1).cuda()          # ft as a channel slice
        assert_close(flow_up2_add(res.cuda(), packed[:, 3:5]), ref, 1e-6, "flow_up2_add, strided input")
        y0, x0, Hc, Wc = 16, 32, 32, 64
        with _lib.crop_context(16 * Hg, 16 * Wg, 16 * y0, 16 * x0, 16 * Hc, 16 * Wc):     # these tensors are the 1/16 level
            out = flow_up2_add(res[..., 2 * y0:2 * (y0 + Hc), 2 * x0:2 * (x0 + Wc)].contiguous().cuda(),
                               ft[..., y0:y0 + Hc, x0:x0 + Wc].contiguous().cuda())
    crop_ref = ref[..., 2 * y0:2 * (y0 + Hc), 2 * x0:2 * (x0 + Wc)]
    assert_close(out[..., 2:-2, 2:-2], crop_ref[..., 2:-2, 2:-2], 1e-6, "flow_up2_add on a crop")


# ---- splat -------------------------------------------------------------------------------------
def test_splat_golden(vfi, golden):
    g = golden("splat")
    for key, (x, fl) in gc.splat_sum_cases().items():
        assert_close(vfi.softsplat_func.apply(*cu(x, fl)), g[key], TOL_SPLAT, f"splat_sum[{key}]")
    for key, (x, fl, metric, mode) in gc.fwarp_cases().items():
        assert_close(vfi.fwarp(*cu(x, fl, metric), mode), g[key], TOL_SPLAT, f"fwarp[{key}]")
        assert_close(vfi.softsplat(*cu(x, fl, metric), mode), g[key], TOL_SPLAT, f"softsplat[{key}]")
    for key, args in gc.two_frame_cases().items():
        o, m = vfi.fwarp_using_two_frames(*cu(*args))
        assert_close(o, g[key + "_out"], TOL_SPLAT, key)
        assert m.dtype == torch.bool and (m.cpu().to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01
    for key, args in gc.mframes_cases().items():
        o, m = vfi.fwarp_mframes(*cu(*args))
        assert_close(o, g[key + "_out"], TOL_SPLAT, key)
        assert (m.cpu().to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01


@pytest.mark.parametrize("algo", ["scatter", "gather"])
def test_two_frame_and_mframes_both_routes(vfi, golden, algo, monkeypatch):
    """a4 / a5 through the scatter kernel (fp32 atomics) and through the gather route of vfi_splat_f32 (a workspace is
    passed: link pass + one owner thread per target pixel adds each direction / branch, no fp atomics): the golden
    cases (C = 3) forced down each route, then C = 16 and C = 64 with folding flows against the oracle."""
    monkeypatch.setenv("VFI_SPLAT_ALGO", algo)
    g = golden("splat")
    for key, args in gc.two_frame_cases().items():
        o, m = vfi.fwarp_using_two_frames(*cu(*args))
        assert_close(o, g[key + "_out"], TOL_SPLAT, f"{key} ({algo})")
        assert (m.cpu().to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01
    for key, args in gc.mframes_cases().items():
        o, m = vfi.fwarp_mframes(*cu(*args))
        assert_close(o, g[key + "_out"], TOL_SPLAT, f"{key} ({algo})")
        assert (m.cpu().to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01
    for (B, C, H, W) in ((2, 16, 67, 121), (1, 64, 136, 240)):
        tag = f"tfr.{B}.{C}"
        in1, in2 = det.normal(tag + ".a", (B, C, H, W)), det.normal(tag + ".b", (B, C, H, W))
        f1 = det.flow_field(tag + ".f1", B, H, W, sigma=4.0, nonfinite_frac=0.001)
        f2 = det.flow_field(tag + ".f2", B, H, W, sigma=4.0)
        t1 = det.uniform(tag + ".t", (B, 1, 1, 1), lo=0.1, hi=0.9)
        m1, m2 = det.normal(tag + ".m1", (B, 1, H, W), std=2.0), det.normal(tag + ".m2", (B, 1, H, W), std=2.0)
        ref, ref_mask = oracle.fwarp_using_two_frames(in1, f1, t1, in2, f2, 1 - t1, m1, m2)
        o, m = vfi.fwarp_using_two_frames(*cu(in1, f1, t1, in2, f2, 1 - t1, m1, m2))
        assert_close(o, ref, TOL_SPLAT, f"two-frame {C} ch ({algo})")
        assert (m.cpu() != ref_mask).float().mean() < 0.01
        nb = 3
        fb1 = torch.stack([det.flow_field(f"{tag}.g1.{i}", B, H, W, sigma=3.0) for i in range(nb)])
        fb2 = torch.stack([det.flow_field(f"{tag}.g2.{i}", B, H, W, sigma=3.0) for i in range(nb)])
        tb = det.uniform(tag + ".tb", (nb, B, 1, 1, 1), lo=0.1, hi=0.9)
        ref, ref_mask = oracle.fwarp_mframes(in1, fb1, tb, in2, fb2, 1 - tb, m1, m2)
        o, m = vfi.fwarp_mframes(*cu(in1, fb1, tb, in2, fb2, 1 - tb, m1, m2))
        assert_close(o, ref, TOL_SPLAT, f"mframes {C} ch ({algo})")
        assert (m.cpu() != ref_mask).float().mean() < 0.01


@pytest.mark.parametrize("mode", ["sum", "avg", "linear", "soft", "soft-zeroeps", "linear-clipeps"])
@pytest.mark.parametrize("shape", [(1, 3, 135, 240), (2, 16, 67, 121)])
def test_splat_vs_oracle(vfi, mode, shape):
    B, C, H, W = shape
    x = det.normal(f"spo.{shape}", shape)
    fl = det.flow_field(f"spo.f.{shape}", B, H, W, sigma=3.0, nonfinite_frac=0.001)
    head = mode.split("-")[0]
    metric = None if head in ("sum", "avg") else (det.normal(f"spo.m.{shape}", (B, 1, H, W)) if head == "soft"
                                                  else det.uniform(f"spo.m.{shape}", (B, 1, H, W), lo=0.01, hi=1.0))
    assert_close(vfi.fwarp(*cu(x, fl, metric), mode), oracle.fwarp(x, fl, metric, mode), TOL_SPLAT, f"fwarp {mode} {shape}")


@pytest.mark.parametrize("shape", [(1, 16, 272, 480), (2, 40, 264, 260)])
def test_splat_large_planes_with_folding_and_far_outliers(vfi, shape, monkeypatch):
    """The link + gather splat at plane sizes where every list feature is exercised at once: every weight / eps mode
    against the oracle with folding flows (overflow nodes beyond the 4 kept in registers: the re-walk path), 1 % far
    outliers, non-finite flows, ragged 32x8 tiles and a channel count that is not a multiple of the chunk; then the
    accumulate modes through the two-frame splat."""
    B, C, H, W = shape
    x = det.normal(f"spt.{shape}", shape)
    fl = det.flow_field(f"spt.f.{shape}", B, H, W, sigma=5.0, nonfinite_frac=0.001)
    far = det.uniform(f"spt.far.{shape}", (B, 1, H, W)) < 0.01
    fl = torch.where(far, (det.uniform(f"spt.big.{shape}", (B, 2, H, W)) * 2 - 1) * W, fl)
    fl[:, :, : H // 3] *= 3.0                        # strong folding in the upper third: lists longer than the registers hold
    for mode in ("sum", "avg", "linear", "soft", "soft-zeroeps", "linear-clipeps"):
        head = mode.split("-")[0]
        metric = None if head in ("sum", "avg") else (det.normal(f"spt.m.{shape}", (B, 1, H, W)) if head == "soft"
                                                      else det.uniform(f"spt.m.{shape}", (B, 1, H, W), lo=0.01, hi=1.0))
        assert_close(vfi.fwarp(*cu(x, fl, metric), mode), oracle.fwarp(x, fl, metric, mode), TOL_SPLAT, f"fwarp {mode} {shape}")
    monkeypatch.setenv("VFI_SPLAT_ALGO", "gather")
    tag = f"spt2.{shape}"
    in2 = det.normal(tag + ".b", shape)
    f2 = det.flow_field(tag + ".f2", B, H, W, sigma=4.0)
    t1 = det.uniform(tag + ".t", (B, 1, 1, 1), lo=0.1, hi=0.9)
    m1, m2 = det.normal(tag + ".m1", (B, 1, H, W), std=2.0), det.normal(tag + ".m2", (B, 1, H, W), std=2.0)
    ref, ref_mask = oracle.fwarp_using_two_frames(x, fl, t1, in2, f2, 1 - t1, m1, m2)
    o, m = vfi.fwarp_using_two_frames(*cu(x, fl, t1, in2, f2, 1 - t1, m1, m2))
    assert_close(o, ref, TOL_SPLAT, f"two-frame, gather route {shape}")
    assert (m.cpu() != ref_mask).float().mean() < 0.01


def test_splat_rejects_cpu_and_bad_modes(vfi):
    x, fl = torch.zeros(1, 1, 4, 4), torch.zeros(1, 2, 4, 4)
    with pytest.raises(Exception):
        vfi.softsplat_func.apply(x, fl)                      # CPU tensors: no fallback (reference asserts False)
    with pytest.raises(AssertionError):
        vfi.fwarp(x.cuda(), fl.cuda(), None, "soft")         # soft needs a metric
    with pytest.raises(AssertionError):
        vfi.fwarp(x.cuda(), fl.cuda(), x.cuda(), "avg")      # avg takes none
    with pytest.raises(AssertionError):
        vfi.fwarp(x.cuda(), fl.cuda(), None, "median")


def test_splat_mass_conservation_1080p(vfi):
    """Linearity / conservation at full size: with flows that stay inside the frame, sum(splat(x)) == sum(x)."""
    H, W = 1080, 1920
    x = torch.rand(1, 3, H, W, device="cuda")
    fl = torch.randn(1, 2, H, W, device="cuda") * 2.0
    fl[..., :8, :] = 0; fl[..., -8:, :] = 0; fl[..., :, :8] = 0; fl[..., :, -8:] = 0
    out = vfi.softsplat(x, fl, None, "sum")
    assert abs(float(out.double().sum() / x.double().sum()) - 1.0) < 1e-5
    out2 = vfi.softsplat(2.0 * x, fl, None, "sum")
    assert rel_err(out2, 2.0 * out) < 1e-5


# ---- conv epilogue (bias + PReLU + residual), SURVEY.md 8f row f2 ---------------------------------
@pytest.mark.parametrize("cl", [False, True])
@pytest.mark.parametrize("shape", [(2, 72, 17, 24), (1, 3, 9, 7), (1, 76, 5, 6)])
def test_conv_epilogue_bit_identical_to_torch_ops(vfi, cl, shape, monkeypatch):
    from videoframeinterpolation_b200.blocks import conv_bias_act
    monkeypatch.setenv("VFI_CONV_ALGO", "cudnn")      # this test is about the epilogue kernel behind a cuDNN conv
    B, C, H, W = shape
    torch.manual_seed(1)
    conv = torch.nn.Conv2d(C, C, 3, 1, 1).cuda()
    prelu = torch.nn.PReLU(C).cuda()
    with torch.no_grad():
        prelu.weight.copy_(torch.rand(C) - 0.3)
    x = torch.randn(B, C, H, W, device="cuda")
    if cl:
        x = x.contiguous(memory_format=torch.channels_last)
        conv = conv.to(memory_format=torch.channels_last)
    with torch.no_grad():
        y = torch.nn.functional.conv2d(x, conv.weight, None, 1, 1)          # the same cuDNN call both sides
        b = conv.bias.view(1, C, 1, 1)
        assert torch.equal(conv_bias_act(conv, x, prelu), prelu(y + b))
        assert torch.equal(conv_bias_act(conv, x, None, residual=x), x + (y + b))
        assert torch.equal(conv_bias_act(conv, x, prelu, residual=x), prelu(x + (y + b)))
        assert torch.equal(conv_bias_act(conv, x, prelu, residual=x, residual_after_act=True), x + prelu(y + b))
        assert torch.equal(conv_bias_act(conv, x), y + b)


@pytest.mark.parametrize("shape", [(2, 72, 37, 52), (1, 8, 19, 23)])
def test_bwarp_flowcat_equals_bwarp_plus_concat(vfi, shape):
    """vfi_bwarp_flowcat_f32 (warp + `torch.cat((.., bwarp(feat_x, ftx), ftx))` of modules/deformable_attn.py:69 + the zero
    padding of the 4-aligned NHWC layout in one pass): bit-identical to vfi_bwarp_f32 followed by the copies, nothing
    written past channel C + 4."""
    from videoframeinterpolation_b200.warp import _bwarp_forward
    B, C, H, W = shape
    img = det.normal(f"bfc.{shape}", shape).cuda().contiguous(memory_format=torch.channels_last)
    fl = det.flow_field(f"bfc.f.{shape}", B, H, W, sigma=4.0).cuda()
    buf = torch.full((B, H, W, C + 8), 777.0, device="cuda").permute(0, 3, 1, 2)
    _bwarp_forward(img, fl, out=buf[:, :C], flowcat=True)
    ref = _bwarp_forward(img, fl)
    assert torch.equal(buf[:, :C], ref)
    assert torch.equal(buf[:, C:C + 2], fl) and bool((buf[:, C + 2:C + 4] == 0).all()) and bool((buf[:, C + 4:] == 777.0).all())
    with pytest.raises(Exception):
        _bwarp_forward(img.contiguous(), fl, out=torch.empty(B, C, H, W, device="cuda"), flowcat=True)     # planar: unsupported
