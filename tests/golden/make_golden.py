"""Generate tests/golden/*.npz by running the REFERENCE's own code on deterministic inputs.

Runs only in the build container (needs the read-only checkout at /root/reference; torch 2.11.0,
torchvision 0.26.0 CPU ops).  Fixtures hold reference OUTPUTS; inputs and weights are rebuilt at test
time from ``tests/_det.py``, which is shared with this script (``cases.py`` lists the cases).

    python tests/golden/make_golden.py [fixture names...]
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

from oracle import build_oracle  # noqa: E402
import golden_cases as gc  # noqa: E402

OUT = Path(__file__).resolve().parent


ONLY = set(sys.argv[1:])     # e.g. `make_golden.py dcndat_model` rewrites just that fixture


def save(name, **arrays):
    if ONLY and name not in ONLY:
        return
    np.savez_compressed(OUT / f"{name}.npz", **{k: np.ascontiguousarray(v.detach().numpy() if torch.is_tensor(v) else v)
                                                for k, v in arrays.items()})
    print("wrote", name, {k: tuple(v.shape) for k, v in arrays.items()})


def main():
    torch.manual_seed(0)
    torch.set_grad_enabled(False)
    build_oracle.import_reference()
    import models as rmodels                      # first: modules.deformable_attn <-> models import cycle
    import modules.warp as rwarp
    import modules.dcnv2 as rdcn
    import modules.deformable_attn as rdat
    import modules.madat as rmadat
    import torchvision.ops

    # ---- bwarp (modules/warp.py:7-15) ---------------------------------------------------------
    out = {}
    for key, (img, flow) in gc.bwarp_cases().items():
        out[key] = rwarp.bwarp(img, flow)
    save("bwarp", **out)

    # ---- splat: reference kernel text compiled as C++ (modules/softsplat.py:289-341) ------------
    rwarp.softsplat_func = types.SimpleNamespace(apply=build_oracle.reference_splat_sum)
    out = {}
    for key, (ten_in, flow) in gc.splat_sum_cases().items():
        out[key] = build_oracle.reference_splat_sum(ten_in, flow)
    for key, (ten_in, flow, metric, mode) in gc.fwarp_cases().items():
        out[key] = rwarp.fwarp(ten_in, flow, metric, mode)
    for key, args in gc.two_frame_cases().items():
        o, m = rwarp.fwarp_using_two_frames(*args)
        out[key + "_out"], out[key + "_mask"] = o, m.to(torch.uint8)
    for key, args in gc.mframes_cases().items():
        o, m = rwarp.fwarp_mframes(*args)
        out[key + "_out"], out[key + "_mask"] = o, m.to(torch.uint8)
    save("splat", **out)

    # ---- DCNv2: torchvision op + the reference module (modules/dcnv2.py:49-67) ------------------
    out = {}
    for key, kw in gc.dcn_op_cases().items():
        out[key] = torchvision.ops.deform_conv2d(kw["x"], kw["offset"], kw["weight"], kw["bias"], stride=kw["stride"],
                                                 padding=kw["padding"], dilation=kw["dilation"], mask=kw["mask"])
    C = gc.DCN_MODULE_C
    mod = rdcn.DeformableConv2d(C, C).eval()
    gc.fill(mod.state_dict())
    x, mv = gc.dcn_module_inputs()
    o, fl = mod(x, mv)
    out["module_out"], out["module_flow"] = o, fl
    # DeformableConv2dGivenOffset body, models/archive/DCNDAT.py:67-78 (class is unimportable; 3 lines around the op)
    kw = gc.dcn_given_offset_case()
    off = kw["offset"].permute(0, 2, 1, 3, 4).flip(1).contiguous().view(kw["x"].shape[0], -1, *kw["x"].shape[2:])
    out["given_offset"] = torchvision.ops.deform_conv2d(kw["x"], off, kw["weight"], kw["bias"], padding=1,
                                                        mask=torch.sigmoid(kw["mask"]), stride=1)
    save("dcn", **out)

    # ---- DAT block (modules/deformable_attn.py) --------------------------------------------------
    out = {}
    cfg = gc.DAT_BLOCK
    blk = rdat.CrossDeformableAttentionBlockwFlow(cfg["c"], cfg["c"], n_samples=cfg["n_samples"], n_groups=cfg["n_groups"],
                                                  n_heads=cfg["n_heads"], mlp_ratio=2, offset_scale=cfg["offset_scale"]).eval()
    gc.fill(blk.state_dict())
    feat_t, feat0, feat1, ft0, ft1 = gc.dat_block_inputs()
    mv0 = blk._get_movement_feats(feat_t, feat0, ft0)
    coords0 = blk._get_ref_coords(feat_t, ft0, mv0)
    kv0 = blk._get_ref_feats(feat0, coords0)
    mv1 = blk._get_movement_feats(feat_t, feat1, ft1)
    coords1 = blk._get_ref_coords(feat_t, ft1, mv1)
    kv1 = blk._get_ref_feats(feat1, coords1)
    out["coords0"], out["kv0"] = coords0, kv0
    out["attn"] = blk.attn(feat_t, torch.cat((kv0, kv1), dim=2))
    o, u0, u1 = blk(feat_t, feat0, feat1, ft0, ft1)
    out["block_out"], out["up_ft0"], out["up_ft1"] = o, u0, u1
    # stand-alone sampler case with out-of-frame coordinates
    feat, flow = gc.dsample_case()
    out["dsample"] = blk._get_ref_feats(feat, flow)
    save("dat_block", **out)

    # ---- MADAT flavour (modules/madat.py:86-115, 161-174) ----------------------------------------
    out = {}
    cfg = gc.MADAT
    m = rmadat.DATwithMotionEstimation(cfg["c"], cfg["c"], n_samples=cfg["n_samples"], n_groups=cfg["n_groups"],
                                       n_heads=cfg["n_heads"]).eval()
    gc.fill(m.state_dict())
    feat, flow, q, kv = gc.madat_inputs()
    s, g = m._get_ref_feats(feat, flow)
    out["samples"], out["grid"] = s, g
    a, s0, s1 = m.attn(q, kv)
    out["attn_out"], out["score0"], out["score1"] = a, s0, s1
    save("madat", **out)

    # ---- whole model (models/DAT.py:74-119), reduced width --------------------------------------
    class NS(dict):
        __getattr__ = dict.get
    args = NS(gc.DAT_SMALL_ARGS)
    net = rmodels.DATwConstantnCv1(args).eval()
    gc.fill(net.state_dict())
    x0, x1, t = gc.dat_model_inputs()
    pred = net.inference(x0, x1, t)
    save("dat_model", pred=pred, n_params=np.array(sum(p.numel() for p in net.parameters())))

    # full-width model: parameter count + a checksum-sized output sample at 64x64
    args = NS(dict(nf=72, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2, distill_lambda=0.01))
    net = rmodels.DATwConstantnCv1(args).eval()
    gc.fill(net.state_dict(), seed=1)
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=1)
    pred = net.inference(x0, x1, t)
    names = np.array(sorted(net.state_dict().keys()))
    save("dat_model_full", pred=pred, n_params=np.array(sum(p.numel() for p in net.parameters())), names=names,
         shapes=np.array([str(tuple(net.state_dict()[k].shape)) for k in names]))

    # ---- DCNDATv1 (models/archive/DCNDAT.py), repaired per SURVEY.md Note B ----------------------
    RepairedDCNDATv1 = import_repaired_dcndat()
    args = NS(gc.DCNDAT_SMALL_ARGS)
    net = RepairedDCNDATv1(args).eval()
    gc.fill(net.state_dict(), seed=3)
    x0, x1, t = gc.dat_model_inputs(H=64, W=96, seed=3)
    pred_small = net.inference(x0, x1, t)
    args = NS(dict(nf=64, enc_res_blocks=5, dec_res_blocks=10, mlp_ratio=2, geo_lambda=0.01, distill_lambda=0.01))
    net = RepairedDCNDATv1(args).eval()
    gc.fill(net.state_dict(), seed=4)
    x0, x1, t = gc.dat_model_inputs(H=64, W=64, seed=4)
    names = np.array(sorted(net.state_dict().keys()))
    save("dcndat_model", pred_small=pred_small, pred_full=net.inference(x0, x1, t),
         n_params=np.array(sum(p.numel() for p in net.parameters())), names=names,
         shapes=np.array([str(tuple(net.state_dict()[k].shape)) for k in names]))

    # ---- validators: evaluate.py:29-84 (its window is created with .cuda(); identity here) --------
    sys.modules.setdefault("imageio", types.SimpleNamespace(imread=None))
    torch.Tensor.cuda = lambda self, *a, **k: self
    import evaluate as rev
    out = {}
    for key, (a, b, L) in gc.metrics_cases().items():
        ssim, cs = rev.calculate_ssim(a, b, full=True, val_range=L)
        out[key] = torch.stack((rev.calculate_psnr(a, b), ssim, cs))
    save("metrics", **out)


def import_repaired_dcndat():
    """models/archive/DCNDAT.py names six symbols that no longer exist (SURVEY.md Note B).  Bind each to what it was
    renamed to, import the file UNMODIFIED, and give the class the `generate_rgb_frame` body of
    models/archive/DCNTrans.py:118-123 (= modules/generator.py BasicResPixelShuffleGenerator) as `self.generator`."""
    import importlib
    import models.BaseModel as rbase
    import models.DAT as rdatm
    import modules.deformable_attn as rdat
    import modules.basic_blocks as rbb
    import modules.generator as rgen
    import utils as rutils
    rdat.DeformAttn = rdat.Attn
    base = types.ModuleType("models.base")
    base.Basemodel = rbase.BaseModel
    sys.modules["models.base"] = base
    rdatm.Mlp = lambda in_features, hidden_features, out_features: rbb.FeadForward(in_features, int(hidden_features), out_features)
    rutils.resize = rbase.BaseModel.resize
    mod = importlib.import_module("models.archive.DCNDAT")
    mod.make_layer = lambda nf, n_layers: rbb.make_residual_layers(nf=nf, n_layers=n_layers)

    class RepairedDCNDATv1(mod.DCNDATv1):
        def __init__(self, args):
            super().__init__(args)
            self.generator = rgen.BasicResPixelShuffleGenerator(args.nf, args.dec_res_blocks)

        def generate_rgb_frame(self, feat, mean_):
            return self.generator(feat, mean_)
    return RepairedDCNDATv1


if __name__ == "__main__":
    main()
