"""GPU parity for DCNv2, the grouped sampler, attention (unfused and fused), and the DAT forward."""
import os

import pytest
import torch

import _det as det
import golden_cases as gc
import oracle
from oracle.torch_oracle import _P
from _util import assert_close, rel_err

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def vfi():
    import videoframeinterpolation_b200 as v
    torch.backends.cudnn.allow_tf32 = False          # dense convs in fp32 for model-level parity
    torch.backends.cuda.matmul.allow_tf32 = False
    return v


def cu(*ts):
    r = tuple(None if t is None else t.cuda() for t in ts)
    return r if len(r) > 1 else r[0]


def cu_kw(kw):
    return {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in kw.items()}


def load_sd(module, sd):
    module.load_state_dict({k: v.clone() for k, v in sd.items()}, strict=True)
    return module.cuda().eval()


# ---- DCNv2 -----------------------------------------------------------------------------------
@pytest.mark.parametrize("algo", ["simt"])
def test_dcn_op_golden(vfi, golden, algo):
    g = golden("dcn")
    for key, kw in gc.dcn_op_cases().items():
        k = cu_kw(kw)
        out = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=k["stride"], padding=k["padding"],
                                dilation=k["dilation"], mask=k["mask"], algo=algo)
        assert_close(out, g[key], TOL, f"dcn[{key}]")


def test_dcn_tcgen05_path(vfi, golden):
    """The tensor-core path (tcgen05.mma kind::tf32, error-compensated 3xTF32, accumulators in TMEM) against the same
    fixtures and against the oracle at dim 64 (cin/group 8: one k-step per tap) and dim 72 (9: two k-steps per tap)."""
    from videoframeinterpolation_b200._lib import VfiError
    g = golden("dcn")
    covered = 0
    for key, kw in gc.dcn_op_cases().items():
        k = cu_kw(kw)
        try:
            out = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=k["stride"], padding=k["padding"],
                                    dilation=k["dilation"], mask=k["mask"], algo="tcgen05")
        except VfiError as e:
            assert "does not cover" in str(e), str(e)
            continue
        assert_close(out, g[key], TOL, f"dcn tcgen05[{key}]")
        covered += 1
    assert covered >= 2, f"tcgen05 path covered only {covered} fixture cases"
    # W = 53: direct gathers only; W = 64 / 96: input window staged by TMA (rows must be 16-byte aligned), large sigma
    # pushes many taps out of the window and out of the frame
    for name, C, OG, H, W, sig in (("tc64", 64, 8, 37, 53, 1.5), ("tc72", 72, 8, 37, 53, 1.5), ("tc64og4", 64, 4, 37, 53, 1.5),
                                   ("tc64w", 64, 8, 38, 64, 1.5), ("tc72w", 72, 8, 21, 96, 6.0), ("tc64wog4", 64, 4, 38, 64, 4.0)):
        kw = gc._dcn_op(name, 2, C, C, H, W, WG=8, OG=OG, sigma=sig)        # partial tiles on both axes
        with oracle.library_ops():
            ref = oracle.deform_conv2d(**kw)
        k = cu_kw(kw)
        out = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"], algo="tcgen05")
        assert_close(out, ref, TOL, f"dcn tcgen05 {name}")
        simt = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"], algo="simt")
        assert_close(out, simt, TOL, f"dcn tcgen05 vs simt {name}")


def test_dcn_module_golden(vfi, golden):
    from test_oracle_golden import _module_sd_dcn
    g = golden("dcn")
    C = gc.DCN_MODULE_C
    mod = load_sd(vfi.DeformableConv2d(C, C), _module_sd_dcn())
    x, mv = gc.dcn_module_inputs()
    with torch.no_grad():
        out, flow = mod(*cu(x, mv))
    assert_close(flow, g["module_flow"], TOL, "dcn module flow")
    assert_close(out, g["module_out"], 2e-5, "dcn module out")
    kw = gc.dcn_given_offset_case()
    m2 = vfi.DeformableConv2dGivenOffset(16, 16)
    m2.load_state_dict({"regular_conv.weight": kw["weight"], "regular_conv.bias": kw["bias"]})
    with torch.no_grad():
        assert_close(m2.cuda()(*cu(kw["x"], kw["offset"], kw["mask"])), g["given_offset"], TOL, "dcn given offset")


def test_dcn_vs_oracle_dim64(vfi):
    kw = gc._dcn_op("big", 2, 64, 64, 24, 40, WG=8, OG=8)
    with oracle.library_ops():
        ref = oracle.deform_conv2d(**kw)
    k = cu_kw(kw)
    out = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"])
    assert_close(out, ref, TOL, "dcn dim64")
    # zero offsets + unit mask == ordinary grouped convolution
    z = torch.zeros_like(k["offset"])
    out0 = vfi.deform_conv2d(k["x"], z, k["weight"], k["bias"], stride=1, padding=1, mask=torch.ones_like(k["mask"]))
    ref0 = torch.nn.functional.conv2d(k["x"], k["weight"], k["bias"], padding=1, groups=8)
    assert_close(out0, ref0, TOL, "dcn zero offsets == grouped conv")


@pytest.mark.parametrize("cfg", [dict(C=64, WG=8, OG=8, H=72, W=96), dict(C=72, WG=8, OG=8, H=68, W=120), dict(C=64, WG=8, OG=4, H=64, W=64),
                                 dict(C=72, WG=8, OG=4, H=40, W=104), dict(C=64, WG=8, OG=8, H=72, W=96, far=0.6)])
def test_dcn_window_kernel_vs_oracle(vfi, cfg):
    """The TMA-window form of the DCN (csrc/dcn_win.cu; algo='window', auto's choice for these shapes): ragged tiles in
    both directions, offset groups == and != weight groups (row a7), offsets with 2 % of the taps far outside the staged
    window (cooperative gather list) and outside the frame, and with 60 % of them far (list overflow: in-thread global
    gathers), against torchvision's op through the oracle; the SIMT kernel on the same inputs for comparison."""
    from videoframeinterpolation_b200._lib import VfiError
    C, WG, OG, H, W = cfg["C"], cfg["WG"], cfg["OG"], cfg["H"], cfg["W"]
    kw = gc._dcn_op(f"win.{C}.{OG}", 2, C, C, H, W, WG=WG, OG=OG)
    g = torch.Generator().manual_seed(C + OG)
    far = torch.rand(kw["offset"].shape, generator=g) < cfg.get("far", 0.02)
    kw["offset"] = torch.where(far, (torch.rand(kw["offset"].shape, generator=g) * 2 - 1) * W, kw["offset"] * 2.0)
    with oracle.library_ops():
        ref = oracle.deform_conv2d(**kw)
    k = cu_kw(kw)
    out = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"], algo="window")
    assert_close(out, ref, TOL, f"dcn window {cfg}")
    auto = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"])
    assert torch.equal(auto, out) or H * W < 4096, "auto takes the window kernel for covered shapes of >= 4096 pixels"
    simt = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], k["bias"], stride=1, padding=1, mask=k["mask"], algo="simt")
    assert_close(simt, ref, TOL, f"dcn simt {cfg}")
    # no mask, no bias
    out2 = vfi.deform_conv2d(k["x"], k["offset"], k["weight"], None, stride=1, padding=1, mask=None, algo="window")
    with oracle.library_ops():
        ref2 = oracle.deform_conv2d(kw["x"], kw["offset"], kw["weight"], None, stride=1, padding=1, dilation=1, mask=None)
    assert_close(out2, ref2, TOL, f"dcn window, no mask / bias {cfg}")
    # shapes the window form does not cover are refused when asked for explicitly (stride 2)
    with pytest.raises(VfiError, match="does not cover"):
        vfi.deform_conv2d(k["x"], k["offset"][:, :, ::2, ::2].contiguous(), k["weight"], k["bias"], stride=2, padding=1,
                          mask=k["mask"][:, :, ::2, ::2].contiguous(), algo="window")


# ---- sampler / attention -------------------------------------------------------------------------
def test_dsample_golden(vfi, golden):
    from videoframeinterpolation_b200.deformable_attn import grouped_sample
    g = golden("dat_block")
    cfg = gc.DAT_BLOCK
    feat, flow = gc.dsample_case()
    assert_close(grouped_sample(*cu(feat, flow), cfg["n_groups"], cfg["n_samples"]), g["dsample"], TOL, "dsample")


def test_dat_block_golden(vfi, golden):
    from test_oracle_golden import dat_block_sd
    g = golden("dat_block")
    cfg = gc.DAT_BLOCK
    blk = load_sd(vfi.CrossDeformableAttentionBlockwFlow(cfg["c"], cfg["c"], n_samples=cfg["n_samples"], n_groups=cfg["n_groups"],
                                                         n_heads=cfg["n_heads"], mlp_ratio=2, offset_scale=cfg["offset_scale"]),
                  dat_block_sd(cfg["c"], cfg["n_samples"], cfg["n_groups"]))
    feat_t, feat0, feat1, ft0, ft1 = cu(*gc.dat_block_inputs())
    with torch.no_grad():
        mv0 = blk._get_movement_feats(feat_t, feat0, ft0)
        coords0 = blk._get_ref_coords(feat_t, ft0, mv0)
        kv0 = blk._get_ref_feats(feat0, coords0)
        mv1 = blk._get_movement_feats(feat_t, feat1, ft1)
        kv1 = blk._get_ref_feats(feat1, blk._get_ref_coords(feat_t, ft1, mv1))
        assert_close(coords0, g["coords0"], TOL, "coords0")
        assert_close(kv0, g["kv0"], TOL, "kv0")
        assert_close(blk.attn(feat_t, torch.cat((kv0, kv1), dim=2)), g["attn"], TOL, "Attn (materialised kv)")
        out, u0, u1 = blk(feat_t, feat0, feat1, ft0, ft1)          # fused path
    assert_close(out, g["block_out"], 2e-5, "block out (fused)")
    assert_close(u0, g["up_ft0"], TOL, "up_ft0")
    assert_close(u1, g["up_ft1"], TOL, "up_ft1")


@pytest.mark.parametrize("cfg", [dict(c=72, nG=8, nS=32, scale=8.0, H=20, W=28), dict(c=72, nG=4, nS=8, scale=2.0, H=17, W=23),
                                 dict(c=64, nG=8, nS=9, scale=2.0, H=12, W=16), dict(c=64, nG=4, nS=9, scale=2.0, H=12, W=16)])
def test_fused_dattn_vs_oracle(vfi, cfg):
    from videoframeinterpolation_b200.deformable_attn import fused_deformable_attention
    c, nG, nS, H, W, B = cfg["c"], cfg["nG"], cfg["nS"], cfg["H"], cfg["W"], 2
    tag = f"fd.{c}.{nG}.{nS}"
    q, f0, f1 = (det.normal(f"{tag}.{n}", (B, c, H, W)) for n in ("q", "f0", "f1"))
    raw0, raw1 = (det.normal(f"{tag}.{n}", (B, nG * nS * 2, H, W)) for n in ("r0", "r1"))
    ft0, ft1 = det.flow_field(f"{tag}.ft0", B, H, W, sigma=2.0), det.flow_field(f"{tag}.ft1", B, H, W, sigma=2.0)
    names = ("q_proj", "k_proj", "v_proj")
    sd = gc.fill({f"{n}.{k}": torch.empty(*s) for n in names for k, s in (("weight", (c, c, 1, 1)), ("bias", (c,)))})
    with oracle.library_ops():
        kv = torch.cat((oracle.dsample(f0, oracle.ref_coords(raw0, ft0, nG, nS, cfg["scale"]), nG, nS),
                        oracle.dsample(f1, oracle.ref_coords(raw1, ft1, nG, nS, cfg["scale"]), nG, nS)), dim=2)
    ref = oracle.attn(_P(sd), q, kv, nG)
    w = [sd[f"{n}.{k}"].cuda() for n in names for k in ("weight", "bias")]
    out = fused_deformable_attention(*cu(q, f0, f1, raw0, raw1, ft0, ft1), *w, n_groups=nG, n_samples=nS, n_heads=nG,
                                     offset_scale=cfg["scale"])
    assert_close(out, ref, TOL, f"fused dattn {cfg}")
    # same offsets delivered the way a bias-free channels-last conv writes them: (B,H,W,nG*nS*2) memory + separate bias
    bias = det.normal(f"{tag}.ob", (nG * nS * 2,), std=0.3)
    r0 = (raw0 - bias.view(1, -1, 1, 1)).cuda().contiguous(memory_format=torch.channels_last)
    r1 = (raw1 - bias.view(1, -1, 1, 1)).cuda().contiguous(memory_format=torch.channels_last)
    out_cl = fused_deformable_attention(*cu(q, f0, f1), r0, r1, *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS, n_heads=nG,
                                        offset_scale=cfg["scale"], offs_bias=bias.cuda())
    assert_close(out_cl, ref, 2e-5, f"fused dattn, channels-last offsets + bias {cfg}")   # (raw - b) + b re-rounds the offsets
    # channels-last q: read in place, and the output comes back channels-last with identical values
    out_qcl = fused_deformable_attention(q.cuda().contiguous(memory_format=torch.channels_last), *cu(f0, f1, raw0, raw1, ft0, ft1), *w,
                                         n_groups=nG, n_samples=nS, n_heads=nG, offset_scale=cfg["scale"])
    assert out_qcl.is_contiguous(memory_format=torch.channels_last) and not out_qcl.is_contiguous()
    assert torch.equal(out_qcl, out), "channels-last q / out must not change the arithmetic"
    # channels-last features + channels-last offsets: the shared-memory window kernel (csrc/dattn_win.cu) where it covers
    # the shape (C = 72, nG = 8), else the features are made planar and the gather kernel runs -- same result either way
    f0cl, f1cl = (f.cuda().contiguous(memory_format=torch.channels_last) for f in (f0, f1))
    out_win = fused_deformable_attention(q.cuda(), f0cl, f1cl, r0, r1, *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS, n_heads=nG,
                                         offset_scale=cfg["scale"], offs_bias=bias.cuda())
    assert_close(out_win, ref, 2e-5, f"fused dattn, channels-last features {cfg}")


def _dattn_case(tag, B, c, nG, nS, H, W, flow_kind, sigma, seed):
    from videoframeinterpolation_b200.workloads import make_flow
    gen = torch.Generator().manual_seed(seed)
    q, f0, f1 = (torch.randn(B, c, H, W, generator=gen) for _ in range(3))
    raw0, raw1 = (torch.randn(B, nG * nS * 2, H, W, generator=gen) * 1.5 for _ in range(2))
    ft0 = make_flow(B, H, W, sigma, gen, kind=flow_kind)
    ft1 = make_flow(B, H, W, sigma, gen, kind=flow_kind)
    names = ("q_proj", "k_proj", "v_proj")
    sd = gc.fill({f"{n}.{k}": torch.empty(*s) for n in names for k, s in (("weight", (c, c, 1, 1)), ("bias", (c,)))})
    return q, f0, f1, raw0, raw1, ft0, ft1, sd, names


def _dattn_ref(q, f0, f1, raw0, raw1, ft0, ft1, sd, nG, nS, scale, dtype=torch.float32):
    """modules/deformable_attn.py:81-101 + 26-40 through the oracle, in ``dtype``."""
    c = lambda t: t.to(dtype)
    with oracle.library_ops():
        kv = torch.cat((oracle.dsample(c(f0), oracle.ref_coords(c(raw0), c(ft0), nG, nS, scale), nG, nS),
                        oracle.dsample(c(f1), oracle.ref_coords(c(raw1), c(ft1), nG, nS, scale), nG, nS)), dim=2)
        return oracle.attn(_P({k: c(v) for k, v in sd.items()}), c(q), kv, nG)


@pytest.mark.parametrize("flow_kind", ["gentle", "smooth", "iid"])
def test_fused_dattn_mid_size_vs_oracle(vfi, flow_kind):
    """The benchmark configuration of the fused attention (C = 72, nG = nH = 8, nS = 32, offset_scale 8, raw channels-last
    offsets, channels-last q / out) at 136x240: 1020 tiles over 148 persistent CTAs (about 7 tiles per CTA, ragged last
    round), random Wq / Wk / Wv, flows with 2 % outliers (window misses -> the global-gather fallback of the window
    kernel) -- both kernels (window form on channels-last features, gather form on planar features) against the
    reference composition modules/deformable_attn.py:26-40, 81-101 through the oracle."""
    from videoframeinterpolation_b200.deformable_attn import fused_deformable_attention
    B, c, nG, nS, H, W, scale = 1, 72, 8, 32, 136, 240, 8.0
    q, f0, f1, raw0, raw1, ft0, ft1, sd, names = _dattn_case("mid", B, c, nG, nS, H, W, flow_kind, 4.0, 11)
    ref = _dattn_ref(q, f0, f1, raw0, raw1, ft0, ft1, sd, nG, nS, scale)
    w = [sd[f"{n}.{k}"].cuda() for n in names for k in ("weight", "bias")]
    cl = lambda t: t.cuda().contiguous(memory_format=torch.channels_last)
    out_win = fused_deformable_attention(cl(q), cl(f0), cl(f1), cl(raw0), cl(raw1), *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS,
                                         n_heads=nG, offset_scale=scale)
    assert out_win.is_contiguous(memory_format=torch.channels_last)
    e_win = assert_close(out_win, ref, TOL, f"fused dattn 136x240 window kernel, {flow_kind} flow")
    out_g = fused_deformable_attention(cl(q), *cu(f0, f1), cl(raw0), cl(raw1), *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS,
                                       n_heads=nG, offset_scale=scale)
    e_g = assert_close(out_g, ref, TOL, f"fused dattn 136x240 gather kernel, {flow_kind} flow")
    print(f"fused dattn 136x240 {flow_kind}: window kernel {e_win:.2e}, gather kernel {e_g:.2e}")


def test_fused_dattn_softmax_rebase_vs_fp64_oracle(vfi):
    """Same shape with k_proj scaled 3x: logits spread over about +-40, so the running softmax shift re-bases in the
    middle of the sample loop (the lazy re-base fires when the maximum grows by more than 30).  At that logit magnitude
    fp32 rounding of the 72-term dot products alone moves the softmax weights by ~1e-5, for the reference's own fp32
    evaluation as much as for the kernels', so the yardstick is the oracle evaluated in fp64: the kernels must be as
    close to it as the reference's fp32 evaluation is (factor 3), or within 1e-5."""
    from videoframeinterpolation_b200.deformable_attn import fused_deformable_attention
    B, c, nG, nS, H, W, scale = 1, 72, 8, 32, 136, 240, 8.0
    q, f0, f1, raw0, raw1, ft0, ft1, sd, names = _dattn_case("reb", B, c, nG, nS, H, W, "gentle", 4.0, 12)
    sd["k_proj.weight"] = sd["k_proj.weight"] * 3.0
    ref64 = _dattn_ref(q, f0, f1, raw0, raw1, ft0, ft1, sd, nG, nS, scale, torch.float64)
    ref32 = _dattn_ref(q, f0, f1, raw0, raw1, ft0, ft1, sd, nG, nS, scale)
    yard = rel_err(ref32, ref64)
    w = [sd[f"{n}.{k}"].cuda() for n in names for k in ("weight", "bias")]
    cl = lambda t: t.cuda().contiguous(memory_format=torch.channels_last)
    out_win = fused_deformable_attention(cl(q), cl(f0), cl(f1), cl(raw0), cl(raw1), *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS,
                                         n_heads=nG, offset_scale=scale)
    out_g = fused_deformable_attention(cl(q), *cu(f0, f1), cl(raw0), cl(raw1), *cu(ft0, ft1), *w, n_groups=nG, n_samples=nS,
                                       n_heads=nG, offset_scale=scale)
    e_win, e_g = rel_err(out_win, ref64), rel_err(out_g, ref64)
    print(f"re-basing softmax: reference fp32 vs fp64 {yard:.2e}; window kernel vs fp64 {e_win:.2e}; gather kernel vs fp64 {e_g:.2e}")
    bound = max(TOL, 3.0 * yard)
    assert e_win <= bound and e_g <= bound, (e_win, e_g, yard)


def test_madat_golden(vfi, golden):
    g = golden("madat")
    cfg = gc.MADAT
    m = vfi.DATwithMotionEstimation(cfg["c"], cfg["c"], n_samples=cfg["n_samples"], n_groups=cfg["n_groups"], n_heads=cfg["n_heads"])
    sd = gc.fill(m.state_dict())
    m = load_sd(m, sd)
    feat, flow, q, kv = cu(*gc.madat_inputs())
    with torch.no_grad():
        s, grid = m._get_ref_feats(feat, flow)
        a, s0, s1 = m.attn(q, kv)
    assert_close(s, g["samples"], TOL, "madat samples")
    assert_close(grid, g["grid"], 1e-6, "madat grid")
    assert_close(a, g["attn_out"], TOL, "madat attn")
    assert_close(s0, g["score0"], TOL, "madat score0")
    assert_close(s1, g["score1"], TOL, "madat score1")


def test_madat_fused_scores_match_unfused(vfi):
    cfg = gc.MADAT
    m = vfi.DATwithMotionEstimation(cfg["c"], cfg["c"], n_samples=cfg["n_samples"], n_groups=cfg["n_groups"], n_heads=cfg["n_heads"])
    m = load_sd(m, gc.fill(m.state_dict()))
    B, c, H, W = 2, cfg["c"], 8, 10
    feat_t, feat0, feat1 = cu(det.normal("mf.t", (B, c, H, W)), det.normal("mf.0", (B, c, H, W)), det.normal("mf.1", (B, c, H, W)))
    ft0, ft1 = cu(det.flow_field("mf.ft0", B, H, W, sigma=1.5), det.flow_field("mf.ft1", B, H, W, sigma=1.5))
    with torch.no_grad():
        a, s0, s1 = m.attend(feat_t, feat0, feat1, ft0, ft1)
        mv0, mv1 = m._get_movement_feats(feat_t, feat0, ft0), m._get_movement_feats(feat_t, feat1, ft1)
        kv0, _ = m._get_ref_feats(feat0, m._get_ref_flow(feat_t, ft0, mv0))
        kv1, _ = m._get_ref_feats(feat1, m._get_ref_flow(feat_t, ft1, mv1))
        a_ref, s0_ref, s1_ref = m.attn(m._norm_c(feat_t), torch.cat((kv0, kv1), dim=2))
    assert_close(a, a_ref, TOL, "madat fused out")
    assert_close(s0, s0_ref, TOL, "madat fused score0")
    assert_close(s1, s1_ref, TOL, "madat fused score1")


# ---- whole model -------------------------------------------------------------------------------
def test_dat_model_golden(vfi, golden):
    g = golden("dat_model")
    args = vfi.default_args(**gc.DAT_SMALL_ARGS)
    net = vfi.DATwConstantnCv1(args)
    net = load_sd(net, gc.fill(net.state_dict()))
    x0, x1, t = cu(*gc.dat_model_inputs())
    assert_close(net.inference(x0, x1, t), g["pred"], 5e-5, "DAT small vs reference")


def test_dat_full_width_golden_and_oracle(vfi, golden):
    g = golden("dat_model_full")
    net = vfi.DATwConstantnCv1(vfi.default_args())
    sd = gc.fill(net.state_dict(), seed=1)
    net = load_sd(net, sd)
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=1)
    assert_close(net.inference(*cu(x0, x1, t)), g["pred"], 5e-5, "DAT nf=72 vs reference")
    # configs[0]: one 448x256 Vimeo90K-shaped triplet against the oracle
    x0, x1, t = gc.dat_model_inputs(H=256, W=448, seed=2)
    with oracle.library_ops():
        ref = oracle.dat_forward({k: v.cpu() for k, v in sd.items()}, x0, x1, t)
    assert_close(net.inference(*cu(x0, x1, t)), ref, 5e-5, "DAT 448x256 vs oracle")


def test_dat_tiled_inference_matches_whole_frame(vfi):
    """BASELINE configs[4] mechanism at a CPU-checkable size.  The tile kernels get the FRAME's extent and the crop's
    origin (frame_ext of the C ABI), so bwarp / sampler / flow-resize sample exactly where the
    whole-frame forward does; what remains is the receptive field of the conv stacks cut by the halo.  Measured
    impulse response of these weights (test_dat_impulse_response_radius): the output moves by more than 1e-5 of full
    scale only within RADIUS px of a perturbation, so interior pixels farther than RADIUS from a cut that is not a
    frame border must agree with the whole-frame forward to 1e-5."""
    from videoframeinterpolation_b200.workloads import init_dat_weights
    torch.manual_seed(0)                      # the reference constructors' own init (what bench.py runs), SURVEY.md 8(d)
    net = init_dat_weights(vfi.DATwConstantnCv1(vfi.default_args(**gc.DAT_SMALL_ARGS))).cuda().eval()
    x0, x1, t = cu(*gc.dat_model_inputs(H=256, W=384, seed=3))
    whole = net.inference(x0, x1, t)
    halo = 64
    tiled = net.inference_tiled(x0, x1, t, grid=(2, 2), halo=halo)
    err = rel_err(tiled, whole)
    print(f"tiled vs whole, halo {halo}: rel err {err:.3e} (frame-extent aware kernels)")
    assert err <= TOL, f"tiled vs whole (halo {halo}): rel err {err:.3e}"
    # a halo shorter than the receptive field shows the truncation, i.e. the test can see a difference
    rough = net.inference_tiled(x0, x1, t, grid=(2, 2), halo=16)
    print(f"tiled vs whole, halo 16: rel err {rel_err(rough, whole):.3e}")
    # rank-sharded tiles fill disjoint interiors whose union is the frame
    parts = [net.inference_tiled(x0, x1, t, grid=(2, 2), halo=halo, rank=r, world=2) for r in range(2)]
    d = float((parts[0] + parts[1] - tiled).abs().max())
    assert d <= 1e-6, f"rank-sharded tiles differ from the single-process tiled run by {d}"


def test_dat_impulse_response_radius(vfi):
    """SURVEY.md H5's measurement repeated on the benchmarked weights (reference constructors' init under seed 0 with the
    zero-initialised heads re-randomised, nf = 72): perturb one pixel of x0 by 0.25 and record how far from it the
    output moves by more than 1e-5 / 1e-7 of full scale.  The halo of the 4K split is chosen from this radius."""
    from videoframeinterpolation_b200.workloads import init_dat_weights
    torch.manual_seed(0)
    net = init_dat_weights(vfi.DATwConstantnCv1(vfi.default_args())).cuda().eval()
    H, W = 384, 512
    x0, x1, t = cu(*gc.dat_model_inputs(H=H, W=W, seed=5))
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False          # fp32 convs: the perturbation must not drown in TF32 rounding
    try:
        base = net.inference(x0, x1, t)
        x0p = x0.clone()
        x0p[:, :, H // 2, W // 2] += 0.25
        # keep the global mean (models/BaseModel.py:14-17) out of the picture: it moves every pixel
        x0p -= (x0p.mean() - x0.mean())
        d = (net.inference(x0p, x1, t) - base).abs().amax(dim=(0, 1))
    finally:
        torch.backends.cudnn.allow_tf32 = prev
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    cheb = torch.maximum((yy - H // 2).abs(), (xx - W // 2).abs())
    r5 = int(cheb[d > 1e-5].max()) if bool((d > 1e-5).any()) else 0
    r7 = int(cheb[d > 1e-7].max()) if bool((d > 1e-7).any()) else 0
    print(f"impulse response radius: > 1e-5 within {r5} px, > 1e-7 within {r7} px")
    assert r5 <= 64, f"the bench halo (64 px) does not cover the 1e-5 radius ({r5} px)"


@pytest.mark.parametrize("hw", [(70, 100), (64, 96)])
def test_frame_edge_interpolate_matches_pad_inference_unpad(vfi, hw):
    """Row f3: uint8 -> float, InputPadder, norm_w_rgb_mean, clamp(out + mean) and unpad as the two fused passes of
    ``interpolate`` against the reference's composition (evaluate.py:187-191) through the oracle."""
    H, W = hw
    kw = dict(nf=16, enc_res_blocks=1, dec_res_blocks=1)
    net = vfi.DATwConstantnCv1(vfi.default_args(**kw))
    sd = det.fill_state_dict(net.state_dict())
    net.load_state_dict(sd)
    net = net.cuda().eval()
    u0 = (det.uniform(f"fe.u0.{hw}", (2, 3, H, W)) * 255).round().to(torch.uint8)
    u1 = (det.uniform(f"fe.u1.{hw}", (2, 3, H, W)) * 255).round().to(torch.uint8)
    f0, f1 = u0.float() / 255.0, u1.float() / 255.0
    t = torch.full((2, 1, 1, 1), 0.5)
    p0, pad = oracle.input_pad(f0)
    p1, _ = oracle.input_pad(f1)
    ref = oracle.input_unpad(oracle.dat_forward({k: w.cpu() for k, w in sd.items()}, p0, p1, t, oracle.dat_config(**kw)), pad)
    out_f = net.interpolate(f0.cuda(), f1.cuda(), t.cuda())
    assert out_f.shape == (2, 3, H, W) and out_f.dtype == torch.float32
    assert_close(out_f, ref, 5e-5, "interpolate, float frames")
    # the drop-in's own pad -> inference -> unpad composition gives the same frame
    padder = vfi.InputPadder((H, W))
    a, b = padder.pad(f0.cuda(), f1.cuda())
    assert_close(out_f, padder.unpad(net.inference(a, b, t.cuda())), 2e-5, "interpolate vs pad/inference/unpad")
    out_u = net.interpolate(u0.cuda(), u1.cuda(), t.cuda())
    assert out_u.dtype == torch.uint8
    q = (ref * 255.0).round()
    assert int((out_u.cpu().float() - q).abs().max()) <= 1, "uint8 frames: at most one code value from the rounded reference"


def test_pipelined_inference_matches_direct_calls(vfi):
    """graph.py:PipelinedInference (uploads / downloads on copy streams around the replayed graph): a stream of different
    pairs gives, in order, exactly the frames of direct interpolate() calls -- no slot is overwritten early."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import init_dat_weights
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args(nf=16, enc_res_blocks=1, dec_res_blocks=1))).cuda().eval()
    H, W, n = 72, 104, 7
    g = torch.Generator().manual_seed(3)
    frames = [(torch.rand(1, 3, H, W, generator=g) * 255).round().to(torch.uint8).pin_memory() for _ in range(n + 1)]
    t = torch.full((1, 1, 1, 1), 0.5, device="cuda")
    with torch.no_grad():
        want = [net.interpolate(frames[i].cuda(), frames[i + 1].cuda(), t).cpu() for i in range(n)]
    pipe = v.PipelinedInference(net, frames[0], frames[1], t, fn=net.interpolate)
    got = []
    for i in range(n):
        pipe.submit(frames[i], frames[i + 1])
        if pipe.pending() > 1:
            got.append(pipe.collect().clone())
    while pipe.pending():
        got.append(pipe.collect().clone())
    assert len(got) == n
    for i in range(n):
        assert got[i].dtype == torch.uint8 and (got[i].int() - want[i].int()).abs().max() <= 1, i
