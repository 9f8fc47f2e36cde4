"""DCNDATv1 (BASELINE configs[3]; models/archive/DCNDAT.py read as repaired by SURVEY.md Note B).

CPU: the product module's parameter table and the oracle restatement against outputs of the reference's own class
(tests/golden/dcndat_model.npz, written by tests/golden/make_golden.py::import_repaired_dcndat).
GPU: the CUDA path (vfi_bwarp_f32, vfi_dcn_f32 with offset groups != weight groups, vfi_dattn_f32 on final
coordinates, vfi_conv3x3_f32) against the same fixtures and against the oracle at a larger frame.
"""
import pytest
import torch

import golden_cases as gc
import oracle
from _util import assert_close

N_PARAMS = 3677845


def _cfg(args: dict):
    return oracle.dcndat_config(**{k: v for k, v in args.items() if k not in ("geo_lambda", "distill_lambda")})


def test_dcndat_param_table(golden):
    import videoframeinterpolation_b200 as vfi
    g = golden("dcndat_model")
    sd = vfi.DCNDATv1(vfi.dcndat_default_args()).state_dict()
    names = [str(n) for n in g["names"]]
    assert sorted(sd) == names
    for n, shp in zip(names, g["shapes"]):
        assert str(tuple(sd[n].shape)) == str(shp), n
    assert sum(v.numel() for v in sd.values()) == int(g["n_params"]) == N_PARAMS


def test_dcndat_oracle_matches_reference(golden):
    import videoframeinterpolation_b200 as vfi
    g = golden("dcndat_model")
    sd = gc.fill(vfi.DCNDATv1(vfi.dcndat_default_args(**gc.DCNDAT_SMALL_ARGS)).state_dict(), seed=3)
    x0, x1, t = gc.dat_model_inputs(H=64, W=96, seed=3)
    assert_close(oracle.dcndat_forward(sd, x0, x1, t, _cfg(gc.DCNDAT_SMALL_ARGS)), g["pred_small"], 2e-5, "dcndat small")
    sd = gc.fill(vfi.DCNDATv1(vfi.dcndat_default_args()).state_dict(), seed=4)
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=4)
    with oracle.library_ops():
        assert_close(oracle.dcndat_forward(sd, x0, x1, t), g["pred_full"], 2e-5, "dcndat nf=64")


@pytest.fixture()
def vfi_fp32():
    import videoframeinterpolation_b200 as v
    old = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
    torch.backends.cudnn.allow_tf32 = False          # dense convs in fp32 for model-level parity
    torch.backends.cuda.matmul.allow_tf32 = False
    yield v
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _load(net, sd):
    net.load_state_dict({k: v.clone() for k, v in sd.items()}, strict=True)
    return net.cuda().eval()


@pytest.mark.gpu
def test_dcndat_golden(vfi_fp32, golden):
    vfi = vfi_fp32
    g = golden("dcndat_model")
    net = vfi.DCNDATv1(vfi.dcndat_default_args(**gc.DCNDAT_SMALL_ARGS))
    net = _load(net, gc.fill(net.state_dict(), seed=3))
    x0, x1, t = gc.dat_model_inputs(H=64, W=96, seed=3)
    assert_close(net.inference(x0.cuda(), x1.cuda(), t.cuda()), g["pred_small"], 5e-5, "DCNDATv1 nf=32 vs reference")
    net = vfi.DCNDATv1(vfi.dcndat_default_args())
    net = _load(net, gc.fill(net.state_dict(), seed=4))
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=4)
    assert_close(net.inference(x0.cuda(), x1.cuda(), t.cuda()), g["pred_full"], 5e-5, "DCNDATv1 nf=64 vs reference")


@pytest.mark.gpu
def test_dcndat_vs_oracle_larger_frame_and_batch(vfi_fp32):
    vfi = vfi_fp32
    net = vfi.DCNDATv1(vfi.dcndat_default_args())
    sd = gc.fill(net.state_dict(), seed=5)
    net = _load(net, sd)
    x0, x1, t = gc.dat_model_inputs(B=2, H=128, W=192, seed=5)
    t = torch.tensor([0.5, 0.25]).view(2, 1, 1, 1)
    with oracle.library_ops():
        ref = oracle.dcndat_forward({k: v.cpu() for k, v in sd.items()}, x0, x1, t)
    out = net.inference(x0.cuda(), x1.cuda(), t.cuda())
    assert_close(out, ref, 5e-5, "DCNDATv1 2x128x192 vs oracle")
    # frame pairs are independent: the batch equals its pairs run one by one
    one = net.inference(x0[1:].cuda(), x1[1:].cuda(), t[1:].cuda())
    assert_close(out[1:], one, 5e-5, "pair 1 alone")     # cuDNN picks other fp32 algorithms for the other batch size


@pytest.mark.gpu
def test_dcndat_graph_replay(vfi_fp32):
    vfi = vfi_fp32
    net = vfi.DCNDATv1(vfi.dcndat_default_args()).cuda().eval()
    x0, x1, t = (v.cuda() for v in gc.dat_model_inputs(H=96, W=128, seed=6))
    eager = net.inference(x0, x1, t).clone()
    g = vfi.GraphedInference(net, x0, x1, t)
    first = g(x0, x1, t).clone()
    assert_close(first, eager, 1e-5, "graph replay vs eager")
    assert_close(g.replay(), first, 1e-6, "replay vs replay")   # not bitwise: the library convs around the path may split-K with atomics


@pytest.mark.gpu
@pytest.mark.parametrize("channels_last", [False, True])
def test_dcndat_coords_kernel_matches_the_composition(vfi_fp32, channels_last):
    """vfi_dcndat_coords_f32 against DCNDAT.py:150-157 + 69-70 written out in torch ops (same rounding sequence)."""
    import _det as det
    vfi = vfi_fp32
    b, nG, nS, h, w = 2, 4, 9, 37, 53
    n = nG * nS
    blk = vfi.dcndat.DCNCrossDeformableAttentionBlockwFlow(64, 64, n_samples=nS, n_groups=nG, n_heads=nG, mlp_ratio=2).cuda().eval()
    raw = det.normal("dcg.raw", (b, 3 * n, h, w), std=1.5).cuda()
    if channels_last:
        raw = raw.contiguous(memory_format=torch.channels_last)
    ft = det.normal("dcg.ft", (b, 6, h, w), std=3.0).cuda()[:, 2:4]          # a channel slice, as in the model
    blk.conv_res_offset_mask = torch.nn.Identity()                             # feed raw straight through
    import videoframeinterpolation_b200.dcndat as dd
    old = dd.conv_bias_act
    dd.conv_bias_act = lambda conv, x: x
    try:
        coords, off, mask = blk._coords_offsets_mask(ft, raw)
    finally:
        dd.conv_bias_act = old
    o1, o2, m = torch.chunk(raw, 3, dim=1)
    ref = (2.0 * torch.tanh(torch.cat((o1, o2), dim=1))).reshape(b, n, 2, h, w) + ft.unsqueeze(1)
    assert_close(coords, ref, 1e-6, "coords")
    assert_close(off, ref.permute(0, 2, 1, 3, 4).flip(1).contiguous().view(b, -1, h, w), 1e-6, "dcn offsets")
    assert_close(mask, torch.sigmoid(m), 1e-6, "mask")
