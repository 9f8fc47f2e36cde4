"""Deterministic inputs for the golden fixtures (shared by make_golden.py and the tests)."""
from __future__ import annotations

import torch

import _det as det

fill = det.fill_state_dict


# ---- bwarp ------------------------------------------------------------------------------------
def bwarp_cases():
    cases = {}
    cases["a"] = (det.normal("bw.a.img", (2, 5, 13, 17)), det.flow_field("bw.a.flow", 2, 13, 17, sigma=3.0))
    cases["b"] = (det.uniform("bw.b.img", (1, 3, 32, 56)), det.flow_field("bw.b.flow", 1, 32, 56, sigma=2.0))
    # exact-integer coordinates, border hits, far out-of-frame
    f = torch.zeros(1, 2, 8, 8)
    f[0, 0, :, :4] = 1.0
    f[0, 0, :, 4:] = -3.0
    f[0, 1, :4] = 7.0
    f[0, 1, 4:] = -100.0
    cases["c"] = (det.normal("bw.c.img", (1, 2, 8, 8)), f)
    return cases


# ---- splatting --------------------------------------------------------------------------------
def splat_sum_cases():
    return {
        "sum_a": (det.normal("sp.a.in", (2, 4, 12, 20)), det.flow_field("sp.a.flow", 2, 12, 20, sigma=2.5,
                                                                         nonfinite_frac=0.02)),
        "sum_b": (det.normal("sp.b.in", (1, 3, 16, 16)), det.flow_field("sp.b.flow", 1, 16, 16, sigma=6.0,
                                                                         outlier_frac=0.1)),
    }


def fwarp_cases():
    cases = {}
    H, W = 14, 18
    x = det.normal("fw.in", (2, 3, H, W))
    fl = det.flow_field("fw.flow", 2, H, W, sigma=2.0, nonfinite_frac=0.01)
    soft = det.normal("fw.soft", (2, 1, H, W))
    lin = det.uniform("fw.lin", (2, 1, H, W), lo=0.01, hi=1.0)
    cases["mode_sum"] = (x, fl, None, "sum")
    cases["mode_avg"] = (x, fl, None, "avg")
    cases["mode_linear"] = (x, fl, lin, "linear")
    cases["mode_soft"] = (x, fl, soft, "soft")
    cases["mode_soft_addeps"] = (x, fl, soft, "soft-addeps")
    cases["mode_soft_zeroeps"] = (x, fl, soft, "soft-zeroeps")
    cases["mode_linear_clipeps"] = (x, fl, lin, "linear-clipeps")
    # everything leaves the frame except a strip: exercises the eps variants on empty pixels
    fl2 = fl.clone()
    fl2[:, 0, :, : W // 2] = 1000.0
    cases["mode_soft_zeroeps_holes"] = (x, fl2, soft, "soft-zeroeps")
    cases["mode_avg_holes"] = (x, fl2, None, "avg")
    return cases


def two_frame_cases():
    H, W = 12, 16
    t1 = torch.full((2, 1, 1, 1), 0.5)
    t2 = torch.tensor([0.3, 0.8]).view(2, 1, 1, 1)
    return {"two": (det.normal("tf.in1", (2, 3, H, W)), det.flow_field("tf.f1", 2, H, W, sigma=2.0), t1,
                    det.normal("tf.in2", (2, 3, H, W)), det.flow_field("tf.f2", 2, H, W, sigma=2.0), t2,
                    det.normal("tf.m1", (2, 1, H, W), std=8.0), det.normal("tf.m2", (2, 1, H, W), std=30.0))}


def mframes_cases():
    H, W, nb, B = 10, 12, 3, 2
    f1 = torch.stack([det.flow_field(f"mf.f1.{i}", B, H, W, sigma=1.5) for i in range(nb)])
    f2 = torch.stack([det.flow_field(f"mf.f2.{i}", B, H, W, sigma=1.5) for i in range(nb)])
    t1 = det.uniform("mf.t1", (nb, B, 1, 1, 1), lo=0.1, hi=0.9)
    t2 = 1.0 - t1
    return {"mframes": (det.normal("mf.in1", (B, 3, H, W)), f1, t1, det.normal("mf.in2", (B, 3, H, W)), f2, t2,
                        det.normal("mf.m1", (B, 1, H, W)), det.normal("mf.m2", (B, 1, H, W)))}


# ---- DCNv2 ------------------------------------------------------------------------------------
def _dcn_op(tag, B, Cin, Cout, H, W, WG, OG, k=3, stride=1, padding=1, dilation=1, sigma=1.5, bias=True):
    K = k * k
    Ho = (H + 2 * padding - (dilation * (k - 1) + 1)) // stride + 1
    Wo = (W + 2 * padding - (dilation * (k - 1) + 1)) // stride + 1
    off = det.normal(f"dcn.{tag}.off", (B, OG * 2 * K, Ho, Wo), std=sigma)
    # a few far offsets so whole taps fall outside the plane
    far = det.uniform(f"dcn.{tag}.far", (B, OG * 2 * K, Ho, Wo)) < 0.03
    off = torch.where(far, off * 10.0, off)
    return dict(x=det.normal(f"dcn.{tag}.x", (B, Cin, H, W)), offset=off,
                mask=torch.sigmoid(det.normal(f"dcn.{tag}.mask", (B, OG * K, Ho, Wo))),
                weight=det.normal(f"dcn.{tag}.w", (Cout, Cin // WG, k, k), std=0.2),
                bias=det.normal(f"dcn.{tag}.b", (Cout,), std=0.1) if bias else None,
                stride=stride, padding=padding, dilation=dilation)


def dcn_op_cases():
    return {
        "op_g8": _dcn_op("g8", 2, 16, 16, 9, 11, WG=8, OG=8),
        "op_og4_wg8": _dcn_op("og4", 2, 16, 24, 9, 11, WG=8, OG=4),
        "op_c72": _dcn_op("c72", 1, 72, 72, 6, 7, WG=8, OG=8),
        "op_dense_s2": _dcn_op("s2", 1, 6, 5, 10, 9, WG=1, OG=1, stride=2, padding=2, dilation=2, bias=False),
        "op_c64_og4": _dcn_op("c64", 1, 64, 64, 5, 6, WG=8, OG=4),
    }


DCN_MODULE_C = 16


def dcn_module_inputs():
    return det.normal("dcnm.x", (2, DCN_MODULE_C, 10, 12)), det.normal("dcnm.mv", (2, DCN_MODULE_C, 10, 12))


def dcn_given_offset_case():
    B, C, H, W, nG = 1, 16, 7, 9, 4
    return dict(x=det.normal("dgo.x", (B, C, H, W)), offset=det.normal("dgo.off", (B, nG * 9, 2, H, W), std=1.5),
                mask=det.normal("dgo.mask", (B, nG * 9, H, W)), weight=det.normal("dgo.w", (C, C // 8, 3, 3), std=0.2),
                bias=det.normal("dgo.b", (C,), std=0.1))


# ---- deformable attention -------------------------------------------------------------------
DAT_BLOCK = dict(c=16, n_samples=4, n_groups=4, n_heads=4, offset_scale=2.0)


def dat_block_inputs(B=2, H=10, W=12):
    c = DAT_BLOCK["c"]
    return (det.normal("db.t", (B, c, H, W)), det.normal("db.f0", (B, c, H, W)), det.normal("db.f1", (B, c, H, W)),
            det.flow_field("db.ft0", B, H, W, sigma=1.5), det.flow_field("db.ft1", B, H, W, sigma=1.5))


def dsample_case(B=2, H=9, W=11):
    c, nG, nS = DAT_BLOCK["c"], DAT_BLOCK["n_groups"], DAT_BLOCK["n_samples"]
    flow = det.normal("ds.flow", (B, nG * nS, 2, H, W), std=4.0)
    far = det.uniform("ds.far", (B, nG * nS, 2, H, W)) < 0.05
    return det.normal("ds.feat", (B, c, H, W)), torch.where(far, flow * 20.0, flow)


MADAT = dict(c=16, n_samples=3, n_groups=4, n_heads=4)


def madat_inputs(B=2, H=8, W=10):
    c, nG, nS = MADAT["c"], MADAT["n_groups"], MADAT["n_samples"]
    return (det.normal("ma.feat", (B, c, H, W)), det.normal("ma.flow", (B, nG * nS, 2, H, W), std=3.0),
            det.normal("ma.q", (B, c, H, W)), det.normal("ma.kv", (B, c, 2 * nS, H * W)))


# ---- whole model ------------------------------------------------------------------------------
DAT_SMALL_ARGS = dict(nf=16, enc_res_blocks=1, dec_res_blocks=1, mlp_ratio=2, distill_lambda=0.01)
DCNDAT_SMALL_ARGS = dict(nf=32, enc_res_blocks=1, dec_res_blocks=1, mlp_ratio=2, geo_lambda=0.01, distill_lambda=0.01)


def dat_model_inputs(B=1, H=64, W=96, seed=0):
    return (det.uniform("dm.x0", (B, 3, H, W), seed), det.uniform("dm.x1", (B, 3, H, W), seed),
            torch.full((B, 1, 1, 1), 0.5))


# ---- validators (evaluate.py:29-84) ------------------------------------------------------------
def metrics_cases():
    """name -> (img1, img2, val_range or None).  Frames in [0,1], in [0,255], a ragged size, one channel, six channels."""
    out = {}
    a = det.uniform("mt.a", (2, 3, 40, 50))
    out["unit"] = (a, (a + 0.05 * det.normal("mt.n", (2, 3, 40, 50))).clamp(0, 1), None)
    a = det.uniform("mt.b", (1, 3, 67, 45)) * 255.0
    out["bytes"] = (a, (a + 4.0 * det.normal("mt.m", (1, 3, 67, 45))).clamp(0, 255), None)
    a = det.uniform("mt.c", (1, 1, 33, 64))
    out["gray"] = (a, det.uniform("mt.d", (1, 1, 33, 64)), 1.0)
    a = det.uniform("mt.e", (1, 6, 32, 32))
    out["six"] = (a, (a * 0.9 + 0.03), None)
    a = det.uniform("mt.f", (1, 3, 11, 11))
    out["tiny"] = (a, (a + 0.1 * det.normal("mt.g", (1, 3, 11, 11))).clamp(0, 1), None)
    return out
