#!/usr/bin/env python
"""Per-op timing of the vfi_b200 kernels on BASELINE-shaped synthetic inputs (CUDA events, L2 flushed between
iterations).  Used for kernel iteration and as the short command profiled under ncu.

    python tools/op_bench.py --op dattn --level 1 [--iters 10] [--pairs 1]
    python tools/op_bench.py --op bwarp --C 64 --level 0
    python tools/op_bench.py --op splat --C 64 --level 0 [--mode soft]
    python tools/op_bench.py --op dcn   (BASELINE configs[2]: x (32,64,256,256), OG=WG=8)
    python tools/op_bench.py --op all
Prints one JSON object per op: ms (median), GB/s of ALGORITHMIC bytes, fraction of the measured HBM peak.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import videoframeinterpolation_b200 as v  # noqa: E402
from videoframeinterpolation_b200 import workloads as wl  # noqa: E402
from videoframeinterpolation_b200.deformable_attn import fused_deformable_attention  # noqa: E402


def peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    return float(json.load(open(p))["hbm_gbs"]) if os.path.exists(p) else 6650.0


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(x.elapsed_time(y) for x, y in evs)
    return ts[len(ts) // 2], ts[0]


def report(name, cfg, ms, best, nbytes, extra=None):
    d = {"op": name, **cfg, "ms": round(ms, 4), "ms_best": round(best, 4), "alg_bytes": nbytes,
         "gbs": round(nbytes / ms / 1e6, 1), "frac_of_measured_hbm": round(nbytes / ms / 1e6 / peak(), 4)}
    if extra:
        d.update(extra)
    print(json.dumps(d), flush=True)


def run_dattn(a, dev, flush):
    lv = a.level                       # 1: 1/2 res, 2: 1/4, 3: 1/8
    H, W = 1088 >> lv, 1920 >> lv
    nG, nS, scale = {1: (8, 32, 8.0), 2: (8, 16, 4.0), 3: (4, 8, 2.0)}[lv]
    C, B = a.C or 72, a.pairs
    g = torch.Generator().manual_seed(wl.seed_for(4, 10 + lv))
    q = torch.randn(B, C, H, W, generator=g).to(dev)
    f0 = torch.randn(B, C, H, W, generator=g).to(dev)
    f1 = torch.randn(B, C, H, W, generator=g).to(dev)
    def raw_offsets():
        # conv_res_offset outputs are spatially smooth (a 3x3 conv of smooth features): correlated field + small noise;
        # --flow iid gives per-pixel independent offsets (adversarial for the gather)
        n = nG * nS * 2
        if a.flow == "iid":
            return (torch.randn(B, n, H, W, generator=g) * 0.5).to(dev)
        coarse = torch.randn(B, n, H // 16 + 2, W // 16 + 2, generator=g).to(dev)
        f = torch.nn.functional.interpolate(coarse, size=(H, W), mode="bilinear", align_corners=True) * 0.5
        return f + torch.randn(B, n, H, W, generator=g).to(dev) * 0.02
    raw0, raw1 = raw_offsets(), raw_offsets()
    ft0 = wl.make_flow(B, H, W, 8.0 / (1 << lv), g, device=dev)
    ft1 = wl.make_flow(B, H, W, 8.0 / (1 << lv), g, device=dev)
    ws = [torch.randn(C, C, generator=g).to(dev) * C ** -0.5 if i % 2 == 0 else torch.randn(C, generator=g).to(dev) * 0.1
          for i in range(6)]
    if a.layout == "cl":     # what the model hands the kernel: channels-last q / features / offsets (window kernel where covered)
        q, f0, f1, raw0, raw1 = (t.contiguous(memory_format=torch.channels_last) for t in (q, f0, f1, raw0, raw1))
    fn = lambda: fused_deformable_attention(q, f0, f1, raw0, raw1, ft0, ft1, *ws, n_groups=nG, n_samples=nS, n_heads=nG,
                                            offset_scale=scale)
    ms, best = timeit(fn, a.iters, flush)
    flops = 2.0 * B * H * W * (2 * C * C + C * C + 2 * (2 * nS) * nG * (C // nG) * nG + 4 * C * 2 * nS)
    report("dattn", dict(level=lv, B=B, C=C, H=H, W=W, nG=nG, nS=nS, offsets=a.flow, layout=a.layout,
                         algo=os.environ.get("VFI_DATTN_ALGO", "win")), ms, best, wl.bytes_dattn(B, C, H, W, nG, nS),
           {"tflops_fp32": round(flops / ms / 1e9, 2)})


def run_bwarp(a, dev, flush):
    H, W = 1080 >> a.level, 1920 >> a.level
    C, B = a.C or 64, a.pairs
    g = torch.Generator().manual_seed(wl.seed_for(2, a.level * 10 + C))
    img = torch.rand(B, C, H, W, generator=g).to(dev)
    flow = wl.make_flow(B, H, W, 8.0 / (1 << a.level), g, kind=a.flow, device=dev)
    ms, best = timeit(lambda: v.bwarp(img, flow), a.iters, flush)
    report("bwarp", dict(level=a.level, B=B, C=C, H=H, W=W, flow=a.flow), ms, best, wl.bytes_bwarp(B, C, H, W))


def run_splat(a, dev, flush):
    H, W = 1080 >> a.level, 1920 >> a.level
    C, B = a.C or 64, a.pairs
    g = torch.Generator().manual_seed(wl.seed_for(2, a.level * 10 + C))
    img = torch.rand(B, C, H, W, generator=g).to(dev)
    metric = torch.randn(B, 1, H, W, generator=g).to(dev)
    flow = wl.make_flow(B, H, W, 8.0 / (1 << a.level), g, kind=a.flow, device=dev)
    m = None if a.mode in ("sum", "avg") else metric
    ms, best = timeit(lambda: v.softsplat(img, flow, m, a.mode), a.iters, flush)
    report("splat", dict(level=a.level, B=B, C=C, H=H, W=W, flow=a.flow, mode=a.mode), ms, best,
           wl.bytes_splat(B, C, H, W, a.mode))


def run_twoframe(a, dev, flush):
    """fwarp_using_two_frames (a4): SURVEY.md 8d traffic 4*(2(C+3)+C+1) B/px."""
    H, W = 1080 >> a.level, 1920 >> a.level
    C, B = a.C or 64, a.pairs
    g = torch.Generator().manual_seed(wl.seed_for(2, 500 + a.level * 10 + C))
    in1, in2 = torch.rand(B, C, H, W, generator=g).to(dev), torch.rand(B, C, H, W, generator=g).to(dev)
    m1, m2 = torch.randn(B, 1, H, W, generator=g).to(dev), torch.randn(B, 1, H, W, generator=g).to(dev)
    f1 = wl.make_flow(B, H, W, 8.0 / (1 << a.level), g, kind=a.flow, device=dev)
    f2 = wl.make_flow(B, H, W, 8.0 / (1 << a.level), g, kind=a.flow, device=dev)
    t = torch.full((B, 1, 1, 1), 0.5, device=dev)
    ms, best = timeit(lambda: v.fwarp_using_two_frames(in1, f1, t, in2, f2, 1 - t, m1, m2), a.iters, flush)
    report("twoframe", dict(level=a.level, B=B, C=C, H=H, W=W, flow=a.flow, algo=os.environ.get("VFI_SPLAT_ALGO", "auto")), ms, best,
           4 * (2 * (C + 3) + C + 1) * B * H * W)


def run_dcn(a, dev, flush):
    B, C, H, W = a.pairs if a.pairs > 1 else 32, a.C or 64, 256, 256
    OG = WG = 8
    g = torch.Generator().manual_seed(wl.seed_for(3, 0))
    x = torch.randn(B, C, H, W, generator=g).to(dev)
    flow = wl.make_flow(B, H, W, 2.0, g, device="cpu", outlier_frac=a.outliers)
    off = (2 * torch.tanh(torch.randn(B, OG * 18, H, W, generator=g)) + flow.flip(1).repeat(1, OG * 9, 1, 1)).to(dev)
    mask = torch.sigmoid(torch.randn(B, OG * 9, H, W, generator=g)).to(dev)
    wgt = (torch.randn(C, C // WG, 3, 3, generator=g) * 0.1).to(dev)
    bias = torch.randn(C, generator=g).to(dev)
    algo = os.environ.get("VFI_DCN_ALGO", "auto")
    fn = lambda: v.deform_conv2d(x, off, wgt, bias, padding=1, mask=mask, algo=algo)
    ms, best = timeit(fn, a.iters, flush)
    flops = 2.0 * C * (C // WG) * 9 * B * H * W
    report("dcn", dict(B=B, C=C, H=H, W=W, OG=OG, WG=WG, algo=algo, outliers=a.outliers), ms, best, wl.bytes_dcn(B, C, C, H, W, OG),
           {"gemm_tflops": round(flops / ms / 1e9, 2)})


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--op", default="all", choices=["dattn", "bwarp", "splat", "twoframe", "dcn", "all"])
    ap.add_argument("--level", type=int, default=None)
    ap.add_argument("--C", type=int, default=None)
    ap.add_argument("--pairs", type=int, default=1)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--mode", default="soft")
    ap.add_argument("--flow", default="smooth", choices=["smooth", "gentle", "iid"])
    ap.add_argument("--layout", default="cl", choices=["cl", "planar"], help="dattn: channels-last (model) or planar tensors")
    ap.add_argument("--outliers", type=float, default=0.02, help="dcn: fraction of pixels whose flow leaves the frame (SURVEY.md 8d: 2 %)")
    ap.add_argument("--no-flush", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    flush = None if a.no_flush else torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    with torch.no_grad():
        if a.op in ("dattn", "all"):
            for lv in ([a.level] if a.level is not None and a.op == "dattn" else [1, 2, 3]):
                a.level = lv
                run_dattn(a, dev, flush)
        if a.op in ("bwarp", "all"):
            lv0 = a.level if a.op == "bwarp" and a.level is not None else 0
            for C in ([a.C] if a.C else [3, 64]):
                a.level, a.C = lv0, C
                run_bwarp(a, dev, flush)
            a.C = None
        if a.op in ("splat", "all"):
            lv0 = a.level if a.op == "splat" and a.level is not None else 0
            for C in ([a.C] if a.C else [3, 64]):
                a.level, a.C = lv0, C
                run_splat(a, dev, flush)
            a.C = None
        if a.op == "twoframe":
            run_twoframe(a, dev, flush)
        if a.op in ("dcn", "all"):
            run_dcn(a, dev, flush)


if __name__ == "__main__":
    main()
