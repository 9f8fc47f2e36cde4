"""The oracle pinned against outputs of the reference's own code (tests/golden/*.npz). CPU only."""
import torch

import golden_cases as gc
import oracle
from oracle import build_oracle
from oracle.torch_oracle import _P
from _util import assert_close, rel_err

TOL = 2e-6        # oracle vs reference, both fp32 on CPU


def test_bwarp_matches_reference(golden):
    g = golden("bwarp")
    for key, (img, flow) in gc.bwarp_cases().items():
        assert_close(oracle.bwarp(img, flow), g[key], TOL, f"bwarp[{key}]")
        with oracle.library_ops():
            assert_close(oracle.bwarp(img, flow), g[key], 1e-7, f"bwarp-lib[{key}]")


def test_splat_sum_matches_reference_kernel_text(golden):
    g = golden("splat")
    for key, (x, fl) in gc.splat_sum_cases().items():
        assert_close(oracle.splat_sum(x, fl), g[key], TOL, f"splat_sum[{key}]")
        assert_close(build_oracle.c_splat_sum(x, fl), g[key], TOL, f"c_splat_sum[{key}]")


def test_fwarp_modes_match_reference(golden):
    g = golden("splat")
    for key, (x, fl, metric, mode) in gc.fwarp_cases().items():
        assert_close(oracle.fwarp(x, fl, metric, mode), g[key], 5e-6, f"fwarp[{key}]")
    for key, args in gc.two_frame_cases().items():
        o, m = oracle.fwarp_using_two_frames(*args)
        assert_close(o, g[key + "_out"], 5e-6, key)
        assert (m.to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01
    for key, args in gc.mframes_cases().items():
        o, m = oracle.fwarp_mframes(*args)
        assert_close(o, g[key + "_out"], 5e-6, key)
        assert (m.to(torch.uint8) != g[key + "_mask"]).float().mean() < 0.01


def test_dcn_matches_torchvision_and_module(golden):
    g = golden("dcn")
    for key, kw in gc.dcn_op_cases().items():
        assert_close(oracle.deform_conv2d(**kw), g[key], TOL, f"dcn[{key}]")
    import torch.nn as nn
    sd = _module_sd_dcn()
    x, mv = gc.dcn_module_inputs()
    out, flow = oracle.dcn_block(_P(sd), x, mv)
    assert_close(flow, g["module_flow"], TOL, "dcn module flow")
    assert_close(out, g["module_out"], 5e-6, "dcn module out")
    kw = gc.dcn_given_offset_case()
    assert_close(oracle.dcn_given_offset(kw["weight"], kw["bias"], kw["x"], kw["offset"], kw["mask"]),
                 g["given_offset"], TOL, "dcn given offset")


def _module_sd_dcn():
    C = gc.DCN_MODULE_C
    shapes = {"offset_flow_conv.weight": (2, C, 3, 3), "offset_flow_conv.bias": (2,),
              "conv_offset_mask.0.weight": (C, 2 * C + 2, 3, 3), "conv_offset_mask.0.bias": (C,),
              "conv_offset_mask.1.weight": (C,),
              "conv_offset_mask.2.weight": (C, C, 3, 3), "conv_offset_mask.2.bias": (C,),
              "conv_offset_mask.3.weight": (C,),
              "conv_offset_mask.4.weight": (8 * 27, C, 3, 3), "conv_offset_mask.4.bias": (8 * 27,),
              "regular_conv.weight": (C, C // 8, 3, 3), "regular_conv.bias": (C,)}
    return gc.fill({k: torch.empty(v) for k, v in shapes.items()})


def dat_block_sd(c, nS, nG):
    h = c // 2
    shapes = {}

    def conv(name, co, ci, k=3):
        shapes[name + ".weight"] = (co, ci, k, k)
        shapes[name + ".bias"] = (co,)
    conv("movement_extractor.0.0", 2 * c, 2 * c + 2); shapes["movement_extractor.0.1.weight"] = (2 * c,)
    conv("movement_extractor.1.0", c, 2 * c); shapes["movement_extractor.1.1.weight"] = (c,)
    for i, ch in ((1, c), (2, h), (3, c), (4, h)):
        conv(f"movement_extractor.2.conv{i}.0", ch, ch); shapes[f"movement_extractor.2.conv{i}.1.weight"] = (ch,)
    conv("movement_extractor.2.conv5", c, c); shapes["movement_extractor.2.prelu.weight"] = (c,)
    conv("conv_res_offset", nG * nS * 2, c)
    shapes["conv_res_flow.weight"] = (c, 2, 4, 4); shapes["conv_res_flow.bias"] = (2,)
    for n in ("q_proj", "k_proj", "v_proj"):
        conv("attn." + n, c, c, 1)
    conv("mlp.fc1", 2 * c, c, 1); conv("mlp.fc2", c, 2 * c, 1)
    return gc.fill({k: torch.empty(v) for k, v in shapes.items()})


def test_dat_block_matches_reference(golden):
    g = golden("dat_block")
    cfg = gc.DAT_BLOCK
    sd = dat_block_sd(cfg["c"], cfg["n_samples"], cfg["n_groups"])
    feat_t, feat0, feat1, ft0, ft1 = gc.dat_block_inputs()
    out, u0, u1 = oracle.dat_block(_P(sd), feat_t, feat0, feat1, ft0, ft1, cfg["n_samples"], cfg["n_groups"],
                                   cfg["n_heads"], cfg["offset_scale"])
    assert_close(out, g["block_out"], 2e-5, "block out")
    assert_close(u0, g["up_ft0"], 5e-6, "up_ft0")
    assert_close(u1, g["up_ft1"], 5e-6, "up_ft1")
    feat, flow = gc.dsample_case()
    assert_close(oracle.dsample(feat, flow, cfg["n_groups"], cfg["n_samples"]), g["dsample"], TOL, "dsample")
    with oracle.library_ops():
        assert_close(oracle.dsample(feat, flow, cfg["n_groups"], cfg["n_samples"]), g["dsample"], 1e-7, "dsample-lib")
    kv0 = oracle.dsample(feat0, g["coords0"], cfg["n_groups"], cfg["n_samples"])
    assert_close(kv0, g["kv0"], TOL, "kv0")


def test_madat_matches_reference(golden):
    g = golden("madat")
    cfg = gc.MADAT
    c = cfg["c"]
    shapes = {"norm1.weight": (c,), "norm1.bias": (c,)}
    for n in ("q_proj", "k_proj", "v_proj"):
        shapes[f"attn.{n}.weight"] = (c, c); shapes[f"attn.{n}.bias"] = (c,)
    sd = gc.fill({k: torch.empty(v) for k, v in shapes.items()})
    feat, flow, q, kv = gc.madat_inputs()
    s, grid = oracle.madat_ref_feats(_P(sd), feat, flow, cfg["n_groups"], cfg["n_samples"])
    assert_close(s, g["samples"], 5e-6, "madat samples")
    assert_close(grid, g["grid"], 1e-7, "madat grid")
    a, s0, s1 = oracle.madat_attn(_P(sd).sub("attn"), q, kv, cfg["n_heads"])
    assert_close(a, g["attn_out"], 5e-6, "madat attn")
    assert_close(s0, g["score0"], 5e-6, "madat score0")
    assert_close(s1, g["score1"], 5e-6, "madat score1")


def test_dat_param_table(golden):
    g = golden("dat_model_full")
    shapes = oracle.dat_param_shapes()
    names = [str(n) for n in g["names"]]
    assert sorted(shapes) == names
    for n, shp in zip(names, g["shapes"]):
        assert str(tuple(shapes[n])) == str(shp), n
    assert sum(int(torch.tensor(v).prod()) for v in shapes.values()) == int(g["n_params"]) == 5081063


def test_dat_model_matches_reference(golden):
    g = golden("dat_model")
    cfg = oracle.dat_config(**{k: v for k, v in gc.DAT_SMALL_ARGS.items() if k != "distill_lambda"})
    sd = gc.fill({k: torch.empty(v) for k, v in oracle.dat_param_shapes(cfg).items()})
    x0, x1, t = gc.dat_model_inputs()
    assert_close(oracle.dat_forward(sd, x0, x1, t, cfg), g["pred"], 2e-5, "dat small")
    with oracle.library_ops():
        assert_close(oracle.dat_forward(sd, x0, x1, t, cfg), g["pred"], 2e-5, "dat small (library ops)")


def test_dat_full_width_matches_reference(golden):
    g = golden("dat_model_full")
    sd = gc.fill({k: torch.empty(v) for k, v in oracle.dat_param_shapes().items()}, seed=1)
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=1)
    with oracle.library_ops():
        assert_close(oracle.dat_forward(sd, x0, x1, t), g["pred"], 2e-5, "dat nf=72")
