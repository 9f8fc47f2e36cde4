"""GPU parity of the backward kernels through autograd: the gradients of the
CUDA drop-ins against torch autograd run through the oracle's differentiable restatement of the same forward
(oracle.splat_sum is built from index_add_, oracle.bwarp from F.grid_sample: their autograd is the analytic derivative
the reference's softsplat_ingrad / softsplat_flowgrad kernels and grid_sample's backward compute)."""
import pytest
import torch

import _det as det
import oracle
from _util import rel_err

pytestmark = pytest.mark.gpu


def _grads(fn, *tensors):
    leaves = [t.clone().requires_grad_(True) if t is not None else None for t in tensors]
    out = fn(*leaves)
    w = det.normal("bwd.cotangent", tuple(out.shape)).to(out.device)
    (out * w).sum().backward()
    return out.detach(), [None if t is None else t.grad.detach() for t in leaves]


@pytest.mark.parametrize("shape", [(2, 5, 19, 23), (1, 64, 32, 40)])
def test_bwarp_backward_matches_grid_sample_autograd(shape):
    import videoframeinterpolation_b200 as v
    B, C, H, W = shape
    img = det.normal(f"bwd.img.{shape}", shape)
    flow = det.flow_field(f"bwd.flow.{shape}", B, H, W, sigma=2.5, outlier_frac=0.05)
    with oracle.library_ops():
        ref_out, (ref_gi, ref_gf) = _grads(oracle.bwarp, img, flow)
    out, (gi, gf) = _grads(v.bwarp, img.cuda(), flow.cuda())
    assert rel_err(out, ref_out) <= 1e-5
    assert rel_err(gi, ref_gi) <= 1e-5
    assert rel_err(gf, ref_gf) <= 1e-4        # C-term sums of products of differences
    # image-only and flow-only gradients
    x = img.cuda().requires_grad_(True)
    v.bwarp(x, flow.cuda()).sum().backward()
    ones_gi = _grads(lambda a: v.bwarp(a, flow.cuda()), img.cuda())[1][0]
    assert x.grad is not None and x.grad.shape == ones_gi.shape


@pytest.mark.parametrize("shape", [(2, 4, 17, 21), (1, 16, 24, 36)])
def test_splat_sum_backward_matches_autograd(shape):
    from videoframeinterpolation_b200.softsplat import softsplat_func
    B, C, H, W = shape
    x = det.normal(f"bwd.sin.{shape}", shape)
    flow = det.flow_field(f"bwd.sflow.{shape}", B, H, W, sigma=2.0, outlier_frac=0.05, nonfinite_frac=0.01)
    ref_out, (ref_gx, ref_gf) = _grads(oracle.splat_sum, x, flow)
    out, (gx, gf) = _grads(softsplat_func.apply, x.cuda(), flow.cuda())
    assert rel_err(out, ref_out) <= 1e-4
    assert rel_err(gx, ref_gx) <= 1e-5
    fin = torch.isfinite(ref_gf)              # autograd through a NaN flow gives NaN; the kernels give 0 there (as the reference)
    assert rel_err(torch.where(fin, gf.cpu(), torch.zeros_like(ref_gf)), torch.where(fin, ref_gf, torch.zeros_like(ref_gf))) <= 1e-4
    assert bool((gf.cpu()[~fin] == 0).all())


@pytest.mark.parametrize("mode", ["avg", "linear", "soft", "soft-zeroeps", "linear-clipeps"])
def test_softsplat_backward_all_modes(mode):
    import videoframeinterpolation_b200 as v
    B, C, H, W = 2, 6, 18, 22
    x = det.normal(f"bwd.ss.{mode}.x", (B, C, H, W))
    flow = det.flow_field(f"bwd.ss.{mode}.f", B, H, W, sigma=2.0, outlier_frac=0.03)
    head = mode.split("-")[0]
    metric = None if head == "avg" else (det.uniform(f"bwd.ss.{mode}.m", (B, 1, H, W), lo=0.2, hi=1.5) if head == "linear"
                                         else det.normal(f"bwd.ss.{mode}.m", (B, 1, H, W), std=0.7))
    ref_out, ref_g = _grads(lambda a, f, m: oracle.fwarp(a, f, m, mode), x, flow, metric)
    cu = lambda t: None if t is None else t.cuda()
    out, g = _grads(lambda a, f, m: v.softsplat(a, f, m, mode), cu(x), cu(flow), cu(metric))
    assert rel_err(out, ref_out) <= 1e-4
    for name, a, b in zip(("in", "flow", "metric"), g, ref_g):
        if b is not None:
            assert rel_err(a, b) <= 2e-4, name


def test_inference_path_unchanged_without_grad():
    import videoframeinterpolation_b200 as v
    x = det.normal("bwd.nograd.x", (1, 16, 20, 24)).cuda()
    flow = det.flow_field("bwd.nograd.f", 1, 20, 24).cuda()
    m = det.normal("bwd.nograd.m", (1, 1, 20, 24)).cuda()
    with torch.no_grad():
        a = v.softsplat(x.requires_grad_(True), flow, m, "soft")
    assert not a.requires_grad


@pytest.mark.parametrize("cfg", [dict(c=24, nG=4, nS=3, H=13, W=17), dict(c=72, nG=8, nS=4, H=16, W=20)])
def test_grouped_sampler_backward_matches_grid_sample_autograd(cfg):
    """Row f1: gradients of `_get_ref_feats` (modules/deformable_attn.py:88-101) wrt features and sampling flow."""
    from videoframeinterpolation_b200.deformable_attn import grouped_sample
    c, nG, nS, H, W, B = cfg["c"], cfg["nG"], cfg["nS"], cfg["H"], cfg["W"], 2
    feat = det.normal(f"bwd.ds.feat.{cfg}", (B, c, H, W))
    flow = det.normal(f"bwd.ds.flow.{cfg}", (B, nG * nS, 2, H, W), std=3.0)
    flow[:, 0, :, :2] = 3.0 * W                 # samples that leave the frame: zero taps, zero gradient
    with oracle.library_ops():
        ref_out, (ref_gf, ref_gfl) = _grads(lambda a, f: oracle.dsample(a, f, nG, nS), feat, flow)
    out, (gf, gfl) = _grads(lambda a, f: grouped_sample(a, f, nG, nS), feat.cuda(), flow.cuda())
    assert rel_err(out, ref_out) <= 1e-5
    assert rel_err(gf, ref_gf) <= 1e-5
    assert rel_err(gfl, ref_gfl) <= 1e-4
    # feature-only gradient (flow detached) takes the same kernel with grad_coords = NULL
    x = feat.cuda().requires_grad_(True)
    grouped_sample(x, flow.cuda(), nG, nS).sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()


@pytest.mark.parametrize("cfg", [dict(B=2, Cin=16, Cout=24, H=11, W=13, WG=8, OG=8, k=3, pad=1, stride=1, dil=1),
                                 dict(B=1, Cin=16, Cout=16, H=12, W=10, WG=8, OG=4, k=3, pad=1, stride=1, dil=1),
                                 dict(B=1, Cin=8, Cout=8, H=13, W=14, WG=2, OG=2, k=3, pad=2, stride=2, dil=2)])
def test_deform_conv2d_backward_matches_torchvision_autograd(cfg):
    """Row f1: all five gradients of torchvision.ops.deform_conv2d (OG == WG as modules/dcnv2.py, OG != WG as
    models/archive/DCNDAT.py:69, and a strided / dilated case)."""
    import videoframeinterpolation_b200 as v
    B, Cin, Cout, H, W, WG, OG, k = (cfg[n] for n in ("B", "Cin", "Cout", "H", "W", "WG", "OG", "k"))
    pad, stride, dil = cfg["pad"], cfg["stride"], cfg["dil"]
    Ho = (H + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
    Wo = (W + 2 * pad - (dil * (k - 1) + 1)) // stride + 1
    tag = f"bwd.dcn.{Cin}.{OG}.{stride}"
    x = det.normal(tag + ".x", (B, Cin, H, W))
    off = det.normal(tag + ".off", (B, OG * 2 * k * k, Ho, Wo), std=1.5)
    off[:, 0, :2] = 40.0                                   # taps far outside the frame
    mask = torch.sigmoid(det.normal(tag + ".m", (B, OG * k * k, Ho, Wo)))
    wgt = det.normal(tag + ".w", (Cout, Cin // WG, k, k), std=0.2)
    bias = det.normal(tag + ".b", (Cout,))
    f_ref = lambda a, o, w_, b_, m: oracle.deform_conv2d(a, o, w_, b_, stride=stride, padding=pad, dilation=dil, mask=m)
    f_gpu = lambda a, o, w_, b_, m: v.deform_conv2d(a, o, w_, b_, stride=stride, padding=pad, dilation=dil, mask=m)
    with oracle.library_ops():
        ref_out, ref_g = _grads(f_ref, x, off, wgt, bias, mask)
    out, g = _grads(f_gpu, *(t.cuda() for t in (x, off, wgt, bias, mask)))
    assert rel_err(out, ref_out) <= 1e-5
    for name, a, b in zip(("input", "offset", "weight", "bias", "mask"), g, ref_g):
        assert rel_err(a, b) <= 1e-4, name
    # no mask (DCNv1 call form) and input-only gradient
    with oracle.library_ops():
        _, ref_g1 = _grads(lambda a: oracle.deform_conv2d(a, off, wgt, None, stride=stride, padding=pad, dilation=dil), x)
    _, g1 = _grads(lambda a: v.deform_conv2d(a, off.cuda(), wgt.cuda(), None, stride=stride, padding=pad, dilation=dil), x.cuda())
    assert rel_err(g1[0], ref_g1[0]) <= 1e-4


def test_dat_model_trains_through_the_cuda_kernels():
    """Row f1 end to end: loss gradients of a reduced DAT model wrt its parameters, CUDA drop-ins (bwarp, grouped
    sampler, DCN backward kernels behind autograd) against torch autograd through the oracle port on CPU."""
    import videoframeinterpolation_b200 as v
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    kw = dict(nf=16, enc_res_blocks=1, dec_res_blocks=1)
    net = v.DATwConstantnCv1(v.default_args(**kw))
    sd = det.fill_state_dict(net.state_dict())
    net.load_state_dict(sd)
    net = net.cuda().train()
    x0, x1 = det.uniform("bwd.dat.x0", (1, 3, 64, 96)), det.uniform("bwd.dat.x1", (1, 3, 64, 96))
    t = torch.full((1, 1, 1, 1), 0.5)
    target = det.uniform("bwd.dat.gt", (1, 3, 64, 96))
    loss = (net(x0.cuda(), x1.cuda(), t.cuda()) - target.cuda()).square().mean()
    loss.backward()
    ref_sd = {k: w.clone().requires_grad_(True) for k, w in sd.items()}
    with oracle.library_ops():
        ref_loss = (oracle.dat_forward(ref_sd, x0, x1, t, oracle.dat_config(**kw)) - target).square().mean()
    ref_loss.backward()
    assert abs(float(loss.detach()) - float(ref_loss.detach())) <= 1e-5 * max(1.0, abs(float(ref_loss.detach())))
    checked = 0
    for name, p in net.named_parameters():
        rg = ref_sd[name].grad
        if rg is None or float(rg.abs().max()) < 1e-10:     # e.g. k_proj.bias: cancels in the softmax, pure rounding noise
            continue
        assert p.grad is not None, f"no gradient reached {name}"
        # fp32 through ~60 layers; PReLU slopes sum over the sign of near-zero pre-activations, which a last-bit
        # difference flips for a handful of pixels
        # ... and the gradients that pass through sampling coordinates (offset heads, movement extractor) see the
        # derivative of bilinear sampling jump at pixel boundaries; the atomics of the backward kernels and cuDNN's
        # algorithm choice make the run-to-run spread of those ~1e-2.  Direction and size must still agree.
        ga, gb = p.grad.detach().cpu().double().flatten(), rg.double().flatten()
        cos = float((ga @ gb) / (ga.norm() * gb.norm()).clamp_min(1e-300))
        assert cos >= 0.9995, (name, cos)
        assert rel_err(p.grad, rg) <= 3e-2, name
        checked += 1
    assert checked >= 40, checked
    # gradients reach the sampling heads through the sampler's and the DCN's coordinate gradients
    for key in ("dat_lv1.conv_res_offset.weight", "dat_lv3.conv_res_flow.weight", "coarse_query_builder"):
        assert any(key in n and p.grad is not None and float(p.grad.abs().max()) > 0 for n, p in net.named_parameters()), key


def test_two_frame_and_mframes_splat_are_differentiable():
    """fwarp_using_two_frames / fwarp_mframes (modules/warp.py:68-114) with inputs that require grad: the training route
    (reference composition over the differentiable summation splat) against torch autograd through the oracle."""
    import videoframeinterpolation_b200 as v
    B, C, H, W, nb = 2, 5, 14, 18, 3
    in1, in2 = det.normal("tf.in1", (B, C, H, W)), det.normal("tf.in2", (B, C, H, W))
    f1, f2 = det.flow_field("tf.f1", B, H, W, sigma=2.0), det.flow_field("tf.f2", B, H, W, sigma=2.0)
    m1, m2 = det.normal("tf.m1", (B, 1, H, W)), det.normal("tf.m2", (B, 1, H, W))
    t1 = torch.full((B, 1, 1, 1), 0.3)
    ref_out, ref_g = _grads(lambda a, fa, b, fb, ma, mb: oracle.fwarp_using_two_frames(a, fa, t1, b, fb, 1 - t1, ma, mb)[0],
                            in1, f1, in2, f2, m1, m2)
    t1c = t1.cuda()
    out, g = _grads(lambda a, fa, b, fb, ma, mb: v.fwarp_using_two_frames(a, fa, t1c, b, fb, 1 - t1c, ma, mb)[0],
                    *(x.cuda() for x in (in1, f1, in2, f2, m1, m2)))
    assert out.requires_grad is False and rel_err(out, ref_out) <= 1e-4
    for a, b, name in zip(g, ref_g, ("in1", "flow1", "in2", "flow2", "metric1", "metric2")):
        assert rel_err(a, b) <= 2e-4, f"two-frame grad wrt {name}: {rel_err(a, b):.2e}"
    # many-to-many form
    fl1 = torch.stack([det.flow_field(f"mf.f1.{i}", B, H, W, sigma=2.0) for i in range(nb)])
    fl2 = torch.stack([det.flow_field(f"mf.f2.{i}", B, H, W, sigma=2.0) for i in range(nb)])
    tt = torch.linspace(0.2, 0.8, nb).view(nb, 1, 1, 1, 1).expand(nb, B, 1, 1, 1).contiguous()
    ref_out, ref_g = _grads(lambda a, b, ma, mb: oracle.fwarp_mframes(a, fl1, tt, b, fl2, 1 - tt, ma, mb)[0], in1, in2, m1, m2)
    out, g = _grads(lambda a, b, ma, mb: v.fwarp_mframes(a, fl1.cuda(), tt.cuda(), b, fl2.cuda(), 1 - tt.cuda(), ma, mb)[0],
                    *(x.cuda() for x in (in1, in2, m1, m2)))
    assert rel_err(out, ref_out) <= 1e-4
    for a, b, name in zip(g, ref_g, ("in1", "in2", "metric1", "metric2")):
        assert rel_err(a, b) <= 2e-4, f"mframes grad wrt {name}: {rel_err(a, b):.2e}"
    # without grad the fused inference kernels answer, and agree with the training route's values
    with torch.no_grad():
        o_inf = v.fwarp_mframes(in1.cuda(), fl1.cuda(), tt.cuda(), in2.cuda(), fl2.cuda(), 1 - tt.cuda(), m1.cuda(), m2.cuda())[0]
    assert rel_err(o_inf, ref_out) <= 1e-4


def test_eval_forward_without_no_grad_stays_on_the_fused_kernels_and_dcndat_refuses_training():
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import _lib
    net = v.DATwConstantnCv1(v.default_args(nf=16, enc_res_blocks=1, dec_res_blocks=1))
    net.load_state_dict(det.fill_state_dict(net.state_dict()))
    net = net.cuda().eval()
    x0, x1 = det.uniform("ev.x0", (1, 3, 64, 96)).cuda(), det.uniform("ev.x1", (1, 3, 64, 96)).cuda()
    t = torch.full((1, 1, 1, 1), 0.5, device="cuda")
    with _lib.OpTimer() as tm:
        y = net(x0, x1, t)                       # grad mode is on, parameters require grad, inputs do not
    assert not y.requires_grad
    assert "vfi_dattn_f32" in tm.records and "vfi_dsample_f32" not in tm.records, sorted(tm.records)
    with torch.no_grad():
        assert rel_err(y, net.inference(x0, x1, t)) <= 1e-6      # same kernels; cuDNN may pick another algorithm on a repeat call
    net2 = v.DCNDATv1(v.dcndat_default_args(nf=32, enc_res_blocks=1, dec_res_blocks=1)).cuda().train()
    with pytest.raises(_lib.VfiError, match="inference-only"):
        net2(x0, x1, t)


def test_prepared_weight_caches_are_invalidated_explicitly_after_data_writes():
    """writes through .data do not bump Parameter._version, so the TF32-packed weight copies would go stale;
    conv.invalidate_prepared(model) is the documented remedy."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import conv as vc
    prev = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    try:
        net = v.DATwConstantnCv1(v.default_args(nf=16, enc_res_blocks=1, dec_res_blocks=1))
        net.load_state_dict(det.fill_state_dict(net.state_dict()))
        net = net.cuda().eval().to(memory_format=torch.channels_last)
        x0 = det.uniform("iv.x0", (1, 3, 64, 96)).cuda().contiguous(memory_format=torch.channels_last)
        x1 = det.uniform("iv.x1", (1, 3, 64, 96)).cuda().contiguous(memory_format=torch.channels_last)
        t = torch.full((1, 1, 1, 1), 0.5, device="cuda")
        with torch.no_grad():
            y0 = net.inference(x0, x1, t).clone()
            for p in net.pixel_geneartor.parameters():
                p.data.mul_(0.5)                                   # in place through .data: version not bumped
            vc.invalidate_prepared(net)
            y1 = net.inference(x0, x1, t).clone()
            torch.backends.cudnn.allow_tf32 = False
            y1_fp32 = net.inference(x0, x1, t)                     # cuDNN reads the live weights
        assert float((y1 - y0).abs().max()) > 1e-3, "the write changed nothing?"
        assert float((y1 - y1_fp32).abs().max()) <= 2e-2, "prepared weights are stale after invalidate_prepared"
    finally:
        torch.backends.cudnn.allow_tf32 = prev


@pytest.mark.parametrize("cfg", [dict(c=72, nG=8, nS=8, H=22, W=36, strip=3), dict(c=16, nG=8, nS=4, H=17, W=24, strip=5)])
def test_fused_attention_train_matches_the_materialising_composition(cfg):
    """the fused sampler + attention with its strip-wise recomputing backward against autograd
    through the reference's own composition (grouped sampler -> Attn, modules/deformable_attn.py:26-40, 88-101, on the
    oracle's library ops): output and the gradients wrt query, both feature maps, both coordinate fields and the six
    projection tensors.  The strip size forces several strips with a ragged last one."""
    from videoframeinterpolation_b200.deformable_attn import fused_attention_train
    c, nG, nS, H, W, B = cfg["c"], cfg["nG"], cfg["nS"], cfg["H"], cfg["W"], 2
    tag = f"bwd.fat.{c}.{nS}"
    q, f0, f1 = (det.normal(f"{tag}.{n}", (B, c, H, W)) for n in ("q", "f0", "f1"))
    c0, c1 = (det.normal(f"{tag}.{n}", (B, nG * nS, 2, H, W), std=2.5) for n in ("c0", "c1"))
    c0[:, 1, :, :3] = 2.0 * W                   # samples that leave the frame
    ws = [det.normal(f"{tag}.w{i}", (c, c) if i % 2 == 0 else (c,), std=(c ** -0.5 if i % 2 == 0 else 0.1)) for i in range(6)]

    def ref(q, f0, f1, c0, c1, *w):
        kv = torch.cat((oracle.dsample(f0, c0, nG, nS), oracle.dsample(f1, c1, nG, nS)), dim=2)
        hc = c // nG
        qp = torch.nn.functional.conv2d(q, w[0].view(c, c, 1, 1), w[1]).reshape(B * nG, hc, H * W)
        k = torch.nn.functional.conv2d(kv, w[2].view(c, c, 1, 1), w[3]).reshape(B * nG, hc, 2 * nS, H * W)
        v = torch.nn.functional.conv2d(kv, w[4].view(c, c, 1, 1), w[5]).reshape(B * nG, hc, 2 * nS, H * W)
        a = torch.softmax(torch.einsum("bcd,bcsd->bsd", qp, k) * hc ** -0.5, dim=1)
        return torch.einsum("bsd,bcsd->bcd", a, v).reshape(B, c, H, W)

    with oracle.library_ops():
        ref_out, ref_g = _grads(ref, q, f0, f1, c0, c1, *ws)
    strip_bytes = cfg["strip"] * B * c * 2 * nS * W * 4
    out, got = _grads(lambda *a: fused_attention_train(*a, n_groups=nG, n_samples=nS, n_heads=nG, strip_bytes=strip_bytes),
                      *(t.cuda() for t in (q, f0, f1, c0, c1, *ws)))
    assert rel_err(out, ref_out) <= 1e-5
    names = ["q", "feat0", "feat1", "coords0", "coords1", "Wq", "bq", "Wk", "bk", "Wv", "bv"]
    for n, a, r in zip(names, got, ref_g):
        if n == "bk":                           # cancels in the softmax: pure rounding noise on both sides
            assert float(a.abs().max()) <= 1e-4 * float(ref_g[7].abs().max())
            continue
        assert rel_err(a, r) <= (2e-4 if n.startswith("coords") else 5e-5), (n, rel_err(a, r))
    # gradient wrt a subset only (frozen projections, detached coordinates) takes the same route
    x = f0.cuda().requires_grad_(True)
    fused_attention_train(q.cuda(), x, f1.cuda(), c0.cuda(), c1.cuda(), *(w.cuda() for w in ws), n_groups=nG, n_samples=nS,
                          n_heads=nG, strip_bytes=strip_bytes).sum().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all() and float(x.grad.abs().max()) > 0


def test_fused_attention_training_memory_is_bounded_by_the_strip():
    """The point of the recomputing backward: at the 1/8-resolution shape of a 1080p pair (136 x 240, C = 72, 2 x 8 samples)
    the kv-materialising composition keeps several (B, C, 16, HW) tensors alive for autograd (150 MB each); the fused route's
    peak must stay far below it -- at 1/2 resolution the same ratio is 40 GB against ~1.5 GB."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.deformable_attn import CrossDeformableAttentionBlockwFlow
    import os
    torch.manual_seed(0)
    blk = CrossDeformableAttentionBlockwFlow(72, 72, n_samples=8, n_groups=8, n_heads=8, offset_scale=8.0).cuda().train()
    with torch.no_grad():
        blk.conv_res_offset.weight.normal_(0, 0.02)
    B, H, W = 1, 136, 240
    ft, f0, f1 = (det.normal(f"bwd.mem.{n}", (B, 72, H, W)).cuda().requires_grad_(True) for n in ("t", "0", "1"))
    fl0, fl1 = (det.normal(f"bwd.mem.fl{n}", (B, 2, H, W), std=2.0).cuda() for n in "01")
    peaks, grads = {}, {}
    for mode in ("fused", "materialize"):
        os.environ["VFI_DATTN_TRAIN"] = mode
        os.environ["VFI_DATTN_TRAIN_STRIP_MB"] = "8"          # 18 strips of 8 rows at this shape
        try:
            for t in (ft, f0, f1):
                t.grad = None
            blk.zero_grad(set_to_none=True)
            torch.cuda.synchronize()
            torch.cuda.reset_peak_memory_stats()
            base = torch.cuda.memory_allocated()
            out = blk(ft, f0, f1, fl0, fl1)[0]
            out.square().mean().backward()
            torch.cuda.synchronize()
            peaks[mode] = torch.cuda.max_memory_allocated() - base
            grads[mode] = [ft.grad.clone(), f0.grad.clone(), blk.attn.k_proj.weight.grad.clone(), blk.conv_res_offset.weight.grad.clone()]
        finally:
            os.environ.pop("VFI_DATTN_TRAIN", None)
    print(f"peak training memory of one block at 136x240: fused {peaks['fused'] / 2**20:.0f} MiB, materialising "
          f"{peaks['materialize'] / 2**20:.0f} MiB")
    assert peaks["fused"] < 0.6 * peaks["materialize"], peaks
    for a, b_ in zip(grads["fused"], grads["materialize"]):
        assert rel_err(a, b_) <= 2e-3
