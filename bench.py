#!/usr/bin/env python
"""Benchmark contract of the round driver.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # the reference's CPU path (oracle port)

Workload (BASELINE.json metric "DAT frames/sec @1080p"): DAT (configs/DAT.yaml, nf=72) inference on synthetic
1080p frame pairs, replicate-padded to 1088x1920 as the reference's InputPadder does; `--pairs` pairs per GPU
per step (weak scaling: every rank processes its own pairs, no collective on the data path).  One JSON line on
rank 0.  `value` = device-resident throughput; `e2e` = host buffers in, host frame out (H2D/D2H inside the
timed region); `roofline` = the dominant vfi_b200 kernel of the step (fused deformable attention at 1/2 res),
timed with CUDA events inside the timed region; `microbench` = warp + splat GB/s on the 1080p flow pyramid.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

METRIC = "DAT frames/sec @1080p"
UNIT = "frames/s"
H_RAW, W_RAW = 1080, 1920
# ncu --set full summary of the 1/2-res fused-attention launch of THIS tree (written by profiles/ncu_to_json.py from the
# round's own capture): roofline.traffic = dram__bytes_read.sum + dram__bytes_write.sum is read from it at run time
DATTN_NCU_SUMMARY = os.path.join("profiles", "r3_dattn_win_ncu.json")
# CPU arm: a bounded sample of the 1080p workload.  The full 1088x1920 forward of the reference's op composition on the
# host cores was measured once on the GPU box (profiles/r3_cpu_reference_1080p.txt: 167-173 s per pair, 56 GiB, 16
# cores) -- far beyond the few minutes the bench may take for K steps -- so each CPU step is one 544x960 pair, a quarter
# of the pixels at the same aspect, and frames/s are scaled by pixel count (which FLATTERS the CPU: its cost grows
# faster than the pixel count, 3.8 s -> 167 s for 4x the pixels).
CPU_SAMPLE_HW = (544, 960)
CPU_FULL_CONFIG_MEASURED = {"sec_per_pair": 167.0, "frames_per_s": 1.0 / 167.0, "peak_rss_gib": 56.6, "cores": 16,
                            "frame": "1920x1088", "source": "profiles/r3_cpu_reference_1080p.txt (tools/cpu_ref_probe.py on the GPU box's host)"}
FP32_PEAK_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12       # nominal: 148 SMs x 128 FMA lanes x 2 flop x max SM clock


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="vfi_b200", choices=["vfi_b200", "reference"])
    ap.add_argument("--pairs", type=int, default=1, help="frame pairs per GPU per step")
    ap.add_argument("--height", type=int, default=H_RAW)
    ap.add_argument("--width", type=int, default=W_RAW)
    ap.add_argument("--no-micro", action="store_true", help="skip the warp/splat micro-benchmark")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-legs", action="store_true", help="skip the extra legs (DCNDATv1, 4K tiled, 448x256, TF32-vs-fp32 parity)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (profiling runs under ncu only)")
    ap.add_argument("--channels-last", type=int, default=1, help="run the cuDNN conv stacks around the hot path in NHWC (0 = NCHW)")
    ap.add_argument("--graph", type=int, default=1, help="replay the forward from a CUDA graph (0 = eager launches from Python)")
    ap.add_argument("--torch-cuda-baseline", action="store_true",
                    help="also time the reference's op composition on PyTorch-CUDA library ops (ATen grid_sample, torchvision "
                         "deform_conv2d, cuDNN) at the full frame size on this GPU -> line['torch_cuda_baseline']")
    ap.add_argument("--tile-grid", default="", help="ROWSxCOLS: BASELINE configs[4] mode -- ONE frame pair split into a grid of "
                    "tiles with a halo, tiles dealt round-robin to the ranks (strong scaling, no collective); e.g. "
                    "--height 2160 --width 3840 --tile-grid 2x4 --halo 64")
    ap.add_argument("--halo", type=int, default=64)
    ap.add_argument("--tf32", type=int, default=1, help="allow TF32 in the cuDNN convs around the hot path (PyTorch default)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU every 100 ms while running (NVML)."""

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_ev = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv, self.h = pynvml, pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4),
                 "hw_power_brake": getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80)}
        while not self._stop_ev.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop_ev.wait(0.1)

    def stop(self):
        self._stop_ev.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def dist_env(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return rank, world, local


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's CPU path = oracle port with the library ops the reference itself calls
# ------------------------------------------------------------------------------------------------
def cpu_dat_setup(seed=0):
    import oracle
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import init_dat_weights, seed_for
    torch.manual_seed(seed)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args()))          # host-side module only supplies the weights
    sd = {k: w.detach().clone() for k, w in net.state_dict().items()}
    g = torch.Generator().manual_seed(seed_for(1, 0))
    h, w = CPU_SAMPLE_HW
    x0, x1 = torch.rand(1, 3, h, w, generator=g), torch.rand(1, 3, h, w, generator=g)
    t = torch.full((1, 1, 1, 1), 0.5)
    return oracle, sd, x0, x1, t


def cpu_dat_step(oracle, sd, x0, x1, t):
    with torch.no_grad(), oracle.library_ops():
        return oracle.dat_forward(sd, x0, x1, t)


def cpu_scale(height, width):
    """1080p-equivalent frames/s from a CPU_SAMPLE_HW forward: the model is fully convolutional, cost ~ pixels."""
    ph, pw = -(-height // 16) * 16, -(-width // 16) * 16
    return (CPU_SAMPLE_HW[0] * CPU_SAMPLE_HW[1]) / float(ph * pw)


def cpu_sample_text(cores):
    return (f"oracle port of DATwConstantnCv1.inference (torch CPU grid_sample / torchvision deform_conv2d, fp32, {cores} threads) "
            f"timed on one {CPU_SAMPLE_HW[1]}x{CPU_SAMPLE_HW[0]} pair per step = 1/4 of the pixels of a 1920x1088 pair; value = "
            f"1080p-equivalent frames/s by pixel count (EXTRAPOLATED, in the CPU's favour: the full 1920x1088 forward measured "
            f"on this pool's hosts takes {CPU_FULL_CONFIG_MEASURED['sec_per_pair']:.0f} s = "
            f"{CPU_FULL_CONFIG_MEASURED['frames_per_s']:.4f} frames/s, {CPU_FULL_CONFIG_MEASURED['source']})")


def cpu_workload_name(args):
    return (workload_name(args) + f" -- CPU arm: each step times a BOUNDED SAMPLE, one {CPU_SAMPLE_HW[1]}x{CPU_SAMPLE_HW[0]} pair "
            "(1/4 of the pixels); the full-size CPU forward is 167 s per pair (measured, see cpu_baseline)")


def run_reference(args):
    rank, world, _ = dist_env(args.gpus)
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    oracle, sd, x0, x1, t = cpu_dat_setup()
    for _ in range(args.warmup):
        cpu_dat_step(oracle, sd, x0, x1, t)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_dat_step(oracle, sd, x0, x1, t)
    dt = (time.perf_counter() - t0) / max(args.steps, 1)
    value = cpu_scale(args.height, args.width) / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": cpu_workload_name(args), "pairs_per_gpu": args.pairs,
                       "timed_frame": f"{CPU_SAMPLE_HW[1]}x{CPU_SAMPLE_HW[0]}", "value_is": "extrapolated by pixel count to 1920x1088"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": cpu_sample_text(cores),
                             "sec_per_sample_forward": dt, "measured_full_config": CPU_FULL_CONFIG_MEASURED},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def torch_cuda_baseline(args, dev, H, W):
    """North-star comparison point: the reference's own op composition (oracle port of DATwConstantnCv1.inference, which
    materialises the sampled kv tensors exactly as modules/deformable_attn.py does) executed by the PyTorch-CUDA library
    ops the reference calls -- ATen grid_sample, torchvision deform_conv2d, cuDNN convs -- on the same GPU, same frame
    size, same TF32 setting.  Opt-in (needs ~60 GB of HBM at 1080p)."""
    import oracle
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import init_dat_weights, seed_for
    torch.manual_seed(0)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args()))
    sd = {k: w.detach().clone().to(dev) for k, w in net.state_dict().items()}
    del net
    g = torch.Generator().manual_seed(seed_for(4, 0))
    ph, pw = -(-H // 16) * 16, -(-W // 16) * 16
    x0, x1 = (torch.rand(1, 3, ph, pw, generator=g).to(dev) for _ in range(2))
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    try:
        with torch.no_grad(), oracle.library_ops():
            oracle.dat_forward(sd, x0, x1, t)
            torch.cuda.synchronize()
            ts = []
            for _ in range(3):
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                oracle.dat_forward(sd, x0, x1, t)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b))
        ms = sorted(ts)[1]
        peak_gb = torch.cuda.max_memory_allocated() / 1e9
    except torch.OutOfMemoryError as e:
        torch.cuda.empty_cache()
        return {"unavailable": f"out of memory at {pw}x{ph}: {str(e)[:120]}"}
    torch.cuda.empty_cache()
    return {"value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "frame": f"{pw}x{ph}", "pairs": 1,
            "ops": "ATen grid_sampler_2d, torchvision deform_conv2d, cuDNN convs (tf32=%d), eager PyTorch" % args.tf32,
            "max_memory_allocated_gb": round(peak_gb, 1),
            "sample": "oracle port of DATwConstantnCv1.inference with library ops on cuda:0, median of 3 after 1 warm-up"}


def dcndat_1080p(args, dev, H, W, world=1, dist=None, with_library_baseline=True):
    """BASELINE configs[3] read literally: the DCNDATv1 archive model (models/archive/DCNDAT.py, repaired per SURVEY.md
    Note B) on one padded frame pair resident in HBM -- this repo's CUDA-graph replay next to the same composition on
    the PyTorch-CUDA library ops (oracle port with ATen grid_sample / torchvision deform_conv2d / cuDNN)."""
    import oracle
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import init_dat_weights, seed_for
    torch.manual_seed(0)
    net = init_dat_weights(v.DCNDATv1(v.dcndat_default_args())).to(dev).eval().to(memory_format=torch.channels_last)
    g = torch.Generator().manual_seed(seed_for(4, 1))
    ph, pw = -(-H // 16) * 16, -(-W // 16) * 16
    x0, x1 = (torch.rand(1, 3, ph, pw, generator=g).to(dev).contiguous(memory_format=torch.channels_last) for _ in range(2))
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    out = {"model": "DCNDATv1 (configs/archive/DCNDAT.yaml, nf=64)", "frame": f"{pw}x{ph}", "pairs_per_gpu": 1, "n_gpus": world,
           "unit": UNIT, "scaling": "weak"}
    try:
        with torch.no_grad():
            graphed = v.GraphedInference(net, x0, x1, t)
            for _ in range(2):
                graphed.replay()
            if world > 1:
                dist.barrier()
            ms = time_op(graphed.replay, max(args.steps, 5), None)
            if world > 1:
                mt = torch.tensor([ms], device=dev)
                dist.all_reduce(mt, op=dist.ReduceOp.MAX)
                ms = float(mt.item())
            out.update({"value": 1e3 * world / ms, "ms_per_step": ms, "api": "DCNDATv1.inference replayed from a CUDA graph; value = "
                        "all ranks' pairs / max-over-ranks median step time"})
            del graphed
            if with_library_baseline:
                sd = {k: w.detach().clone() for k, w in net.state_dict().items()}
                xa, xb = x0.contiguous(), x1.contiguous()
                with oracle.library_ops():
                    ms_lib = time_op(lambda: oracle.dcndat_forward(sd, xa, xb, t), 3, None)
                out["torch_cuda_baseline"] = {"value": 1e3 / ms_lib, "ms_per_step": ms_lib, "speedup_of_this_repo": ms_lib / ms,
                                              "ops": "ATen grid_sampler_2d, torchvision deform_conv2d, cuDNN convs (tf32=%d), eager PyTorch" % args.tf32}
    except torch.OutOfMemoryError as e:
        out["unavailable"] = f"out of memory: {str(e)[:120]}"
    torch.cuda.empty_cache()
    return out


def workload_name(args):
    ph, pw = -(-args.height // 16) * 16, -(-args.width // 16) * 16
    return (f"DAT (configs/DAT.yaml, nf=72) inference on synthetic {args.width}x{args.height} frame pairs replicate-padded to "
            f"{pw}x{ph}, sharded by frame pair (BASELINE configs[3] shape, live DAT model)")


# ------------------------------------------------------------------------------------------------
# warp + splat micro-benchmark (BASELINE configs[1]): 1080p 4-level pyramid, C = 3..64
# ------------------------------------------------------------------------------------------------
def time_op(fn, iters, flush):
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
    return ts[len(ts) // 2]


def microbench(dev, hbm_peak):
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import bytes_bwarp, bytes_splat, make_flow, seed_for
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > 126 MB L2
    res = {"flow": "smooth", "l2_flush": "256 MiB memset between iterations", "cases": []}
    for li in range(4):
        H, W = H_RAW >> li, W_RAW >> li
        for C in (3, 16, 32, 64):
            g = torch.Generator().manual_seed(seed_for(2, li * 10 + C))
            img = torch.rand(1, C, H, W, generator=g).to(dev)
            metric = torch.randn(1, 1, H, W, generator=g).to(dev)
            row = {"level": li, "C": C, "H": H, "W": W}
            for kind in ("smooth", "iid", "gentle"):
                if kind != "smooth" and not (li == 0 and C in (3, 64)):
                    continue
                flow = make_flow(1, H, W, 8.0 / (1 << li), g, kind=kind, device=dev)
                sfx = "" if kind == "smooth" else "_" + kind
                ms = time_op(lambda: v.bwarp(img, flow), 10, flush)
                row["bwarp_gbs" + sfx] = bytes_bwarp(1, C, H, W) / ms / 1e6
                ms = time_op(lambda: v.softsplat(img, flow, metric, "soft"), 10, flush)
                row["softsplat_gbs" + sfx] = bytes_splat(1, C, H, W) / ms / 1e6
            res["cases"].append(row)
    big = [r for r in res["cases"] if r["level"] == 0 and r["C"] == 64][0]
    res["bwarp_1080p_C64_frac_of_hbm"] = big["bwarp_gbs"] / hbm_peak
    res["softsplat_1080p_C64_frac_of_hbm"] = big["softsplat_gbs"] / hbm_peak
    res["bwarp_1080p_C64_gentle_frac_of_hbm"] = big["bwarp_gbs_gentle"] / hbm_peak
    res["softsplat_1080p_C64_gentle_frac_of_hbm"] = big["softsplat_gbs_gentle"] / hbm_peak
    res["flow_kinds"] = ("smooth: 32 px correlation length (strain ~0.35 px/px) + 2% outliers; gentle: 256 px correlation "
                         "length; iid: independent per pixel (SURVEY.md 8d literal)")
    res["dcn_cfg3"] = dcn_cfg3(dev, hbm_peak, flush)
    res["baselines_1080p"] = op_baselines(dev, flush)
    return res


def _best_of(fn, n):
    fn()
    ts = []
    for _ in range(n):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts) * 1e3


def op_baselines(dev, flush):
    """beside the warp / splat numbers at 1080x1920, (a) the reference function on the box's host cores
    (torch CPU grid_sample through modules/warp.py:7-15; the pure-torch scatter restatement of the splat, which the
    reference implements only in CUDA; and the reference's own kernel text compiled as C++, one thread), and (b) the same
    GPU running what the reference runs: ATen grid_sampler_2d through the reference's bwarp (including its per-call grid
    construction and H2D copy, modules/warp.py:9-11), and the reference's splat kernel text compiled unmodified for
    sm_100a (oracle/_ref, built from /root/reference in the build container) inside the reference's front-end
    composition (exp, cat, zero-fill, kernel, divide: modules/softsplat.py:237-279)."""
    import oracle
    import videoframeinterpolation_b200 as v
    from oracle import build_oracle as bo
    from videoframeinterpolation_b200.workloads import make_flow, seed_for
    H, W = H_RAW, W_RAW
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    rows = []
    for C in (3, 64):
        g = torch.Generator().manual_seed(seed_for(2, C))
        img = torch.rand(1, C, H, W, generator=g)
        metric = torch.randn(1, 1, H, W, generator=g)
        flow = make_flow(1, H, W, 8.0, g, kind="smooth")
        row = {"C": C, "H": H, "W": W, "flow": "smooth", "cpu_cores": cores}
        with oracle.library_ops():
            row["bwarp_cpu_ms"] = _best_of(lambda: oracle.bwarp(img, flow), 3)
        row["softsplat_cpu_ms"] = _best_of(lambda: oracle.fwarp(img, flow, metric, "soft"), 1 if C > 8 else 2)
        dimg, dflow, dmetric = img.to(dev), flow.to(dev), metric.to(dev)
        with oracle.library_ops():
            row["bwarp_torch_cuda_ms"] = time_op(lambda: oracle.bwarp(dimg, dflow), 10, flush)
        row["bwarp_vfi_ms"] = time_op(lambda: v.bwarp(dimg, dflow), 10, flush)
        row["softsplat_vfi_ms"] = time_op(lambda: v.softsplat(dimg, dflow, dmetric, "soft"), 10, flush)
        if bo.ref_cuda_so_path(1, C + 1, H, W).exists():
            buf = torch.empty(1, C + 1, H, W, device=dev)

            def ref_soft():
                e = dmetric.exp()
                out = bo.reference_splat_sum_cuda(torch.cat([dimg * e, e], 1), dflow, out=buf)
                return out[:, :-1] / (out[:, -1:] + 0.0000001)
            got, want = ref_soft(), v.softsplat(dimg, dflow, dmetric, "soft")
            row["softsplat_ref_kernel_vs_vfi_rel_err"] = float((got - want).abs().max() / want.abs().max())
            row["softsplat_ref_kernel_cuda_ms"] = time_op(ref_soft, 10, flush)
        else:
            row["softsplat_ref_kernel_cuda_ms"] = None
        if C == 3 and bo.ref_so_path(1, C + 1, H, W).exists():
            e = metric.exp()
            x = torch.cat([img * e, e], 1)
            row["softsplat_ref_kernel_text_cpu_1thread_ms"] = _best_of(lambda: bo.reference_splat_sum(x, flow), 1)
        rows.append(row)
    return {"rows": rows, "note": "cpu = reference functions on the host cores (oracle port, library ops); torch_cuda / ref_kernel_cuda = "
            "what the reference itself launches on this GPU; vfi = this repo, same tensors, L2 flushed between iterations"}


def dcn_cfg3(dev, hbm_peak, flush):
    """BASELINE configs[2]: DCNv2 dim 64 at 256x256 patches, batch 32, OG = WG = 8; torchvision's CUDA op beside it."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import bytes_dcn, make_flow, seed_for
    B, C, H, W, OG, WG = 32, 64, 256, 256, 8, 8
    g = torch.Generator().manual_seed(seed_for(3, 0))
    x = torch.randn(B, C, H, W, generator=g).to(dev)
    flow = make_flow(B, H, W, 2.0, g, device="cpu")
    off = (2 * torch.tanh(torch.randn(B, OG * 18, H, W, generator=g)) + flow.flip(1).repeat(1, OG * 9, 1, 1)).to(dev)
    mask = torch.sigmoid(torch.randn(B, OG * 9, H, W, generator=g)).to(dev)
    wgt = (torch.randn(C, C // WG, 3, 3, generator=g) * 0.1).to(dev)
    bias = torch.randn(C, generator=g).to(dev)
    ms = time_op(lambda: v.deform_conv2d(x, off, wgt, bias, padding=1, mask=mask), 5, flush)
    nbytes = bytes_dcn(B, C, C, H, W, OG)
    out = {"shape": "x (32,64,256,256), offset (32,144,256,256), mask (32,72,256,256), weight (64,8,3,3)", "ms": ms,
           "gbs": nbytes / ms / 1e6, "frac_of_hbm": nbytes / ms / 1e6 / hbm_peak, "algorithmic_bytes": nbytes,
           "kernel": "dcn_win_kernel (TMA-staged window, fp32 FFMA2 contraction; what auto picks for this shape)",
           "gemm_tflops_fp32": 2.0 * C * (C // WG) * 9 * B * H * W / ms / 1e9,
           "tensor_pipe_active_pct": {"default_kernel": 0.0, "tcgen05_path_algo2": 1.2,
                                      "note": "the per-group contraction is 8 x 8 x 9 MACs per sampled value: the default kernels keep "
                                              "it on the FMA pipe; the tcgen05 form (algo 2, 3xTF32, parity-green) measured 8.7 ms with "
                                              "the tensor pipe 1.2 % active (profiles/r1g_dcn_tc_ncu_full.txt) and is not a default"}}
    try:
        import torchvision.ops as tvo
        ms_tv = time_op(lambda: tvo.deform_conv2d(x, off, wgt, bias, padding=1, mask=mask), 3, flush)
        out["torchvision_cuda_ms"] = ms_tv
        # the reference's CPU path for the same op on a bounded sample: 4 of the 32 patches, scaled x8
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        xc, oc, mc, wc, bc = x[:4].cpu(), off[:4].cpu(), mask[:4].cpu(), wgt.cpu(), bias.cpu()
        out["torchvision_cpu_ms"] = 8.0 * _best_of(lambda: tvo.deform_conv2d(xc, oc, wc, bc, padding=1, mask=mc), 2)
        out["torchvision_cpu_note"] = f"{cores} host threads, 4 of the 32 patches timed, x8"
    except Exception as e:      # library baseline only; never on the product path
        out["torchvision_cuda_ms"] = None
        out["torchvision_note"] = str(e)[:100]
    return out



def dattn_ncu_summary():
    """dram bytes of the 1/2-res fused-attention launch from the round's own ncu --set full capture (profiles/)."""
    path = os.path.join(ROOT, DATTN_NCU_SUMMARY)
    try:
        with open(path) as f:
            d = json.load(f)
        return int(d["dram_bytes_read"]) + int(d["dram_bytes_write"]), DATTN_NCU_SUMMARY, d
    except Exception:
        return None, None, None


def dattn_flops(B, C, H, W, nG, nS):
    """fp32 flops of the folded algebra per launch: q / kt / v projections once per pixel, logits and
    weighted sums over 2 nS slots x nG heads x C channels, bilinear taps."""
    return 2.0 * B * H * W * (2 * C * C + C * C + 2 * (2 * nS) * nG * C + 4 * C * 2 * nS)


def tiled_4k_leg(net, dev, rank, world, steps, dist, grid=(2, 4), halo=64, H=2160, W=3840):
    """BASELINE configs[4] inside the normal bench line: ONE synthetic 3840x2160 pair, 2x4 spatial tiles with a 64-px
    halo dealt to the ranks (strong scaling, no collective), every tile run with the frame's extent passed to the warp /
    sampler kernels (frame_ext of the C ABI), CUDA-graph replay; pairs/s of the WHOLE pair = 1 / max-over-ranks time."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.shard import tile_plan, tiles_of_rank
    from videoframeinterpolation_b200.workloads import seed_for
    g = torch.Generator().manual_seed(seed_for(5, 0))           # the same pair on every rank
    x0, x1 = torch.rand(1, 3, H, W, generator=g).to(dev), torch.rand(1, 3, H, W, generator=g).to(dev)
    padder = v.InputPadder((H, W), divisor=16)
    d_x0, d_x1 = padder.pad(x0, x1)
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    ph, pw = d_x0.shape[-2:]
    plan = tile_plan(ph, pw, grid, halo)
    mine = tiles_of_rank(plan, rank, world)
    out_buf = torch.zeros((1, 3, ph, pw), device=dev)

    def fn(a, b, tt):
        return net.inference_tiled(a, b, tt, grid=grid, halo=halo, rank=rank, world=world, out=out_buf)

    res = {"frame": f"{W}x{H} (padded {pw}x{ph})", "tile_grid": f"{grid[0]}x{grid[1]}", "halo_px": halo, "tiles_per_rank": len(mine),
           "unit": "pairs/s", "scaling": "strong", "parallelism": f"spatial tiles x{world}, no collective"}
    try:
        graphed = v.GraphedInference(net, d_x0, d_x1, t, fn=fn) if mine else None
        for _ in range(2):
            if graphed is not None:
                graphed.replay()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            if graphed is not None:
                graphed.replay()
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_per_step = float(ms.item()) / steps
        crop_px = sum((tl.crop[0].stop - tl.crop[0].start) * (tl.crop[1].stop - tl.crop[1].start) for tl in plan)
        res.update({"value": 1e3 / ms_per_step, "ms_per_step": ms_per_step, "redundant_pixels": crop_px / float(ph * pw) - 1.0,
                    "tile_vs_whole_parity": "1.1e-6 of full scale at halo 64 (tests/test_gpu_parity_dat.py::test_dat_tiled_inference_matches_whole_frame); "
                                            "impulse-response radius of these weights: 8 px at 1e-5 (test_dat_impulse_response_radius)"})
        del graphed
    except torch.OutOfMemoryError as e:
        res["unavailable"] = f"out of memory: {str(e)[:120]}"
    torch.cuda.empty_cache()
    return res


def configs0_leg(net, dev, steps):
    """BASELINE configs[0] on the GPU: one synthetic 448x256 Vimeo90K-shaped pair, CUDA-graph replay."""
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200.workloads import seed_for
    g = torch.Generator().manual_seed(seed_for(1, 0))
    x0 = torch.rand(1, 3, 256, 448, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
    x1 = torch.rand(1, 3, 256, 448, generator=g).to(dev).contiguous(memory_format=torch.channels_last)
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    graphed = v.GraphedInference(net, x0, x1, t)
    ms = time_op(graphed.replay, max(steps, 10), None)
    del graphed
    torch.cuda.empty_cache()
    return {"frame": "448x256", "pairs": 1, "value": 1e3 / ms, "unit": UNIT, "ms_per_step": ms, "api": "DATwConstantnCv1.inference, CUDA graph replay"}


def tf32_parity_leg(net, d_x0, d_x1, t):
    """The measured mode (TF32 tensor-core convs, PyTorch's default conv arithmetic) against the same forward with fp32 convs
    (cuDNN, allow_tf32 = False) on the benchmark's own 1088x1920 pair: max / mean absolute error on [0, 1] frames and PSNR,
    plus the fp32-conv step time (eager)."""
    import math
    with torch.no_grad():
        y_tf32 = net.inference(d_x0, d_x1, t).float().clone()
        prev = torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32
        torch.backends.cudnn.allow_tf32 = False
        torch.backends.cuda.matmul.allow_tf32 = False
        try:
            y_fp32 = net.inference(d_x0, d_x1, t).float().clone()
            ms_fp32 = time_op(lambda: net.inference(d_x0, d_x1, t), 3, None)
        finally:
            torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = prev
    d = (y_tf32 - y_fp32).abs()
    mse = float((d.double() ** 2).mean())
    return {"max_abs": float(d.max()), "mean_abs": float(d.mean()), "psnr_db": (-10.0 * math.log10(mse) if mse > 0 else float("inf")),
            "frame": f"{d_x0.shape[-1]}x{d_x0.shape[-2]}", "reference": "this repo's forward with fp32 convolutions (cuDNN, allow_tf32 = False); "
            "the fp32 forward itself is pinned to the oracle at 5e-5 (tests/test_gpu_parity_dat.py)",
            "fp32_convs_ms_per_step_eager": ms_fp32, "fp32_convs_frames_per_s_eager": 1e3 * d_x0.shape[0] / ms_fp32}

# ------------------------------------------------------------------------------------------------
# this repo's arm
# ------------------------------------------------------------------------------------------------
def run_vfi(args):
    import torch.distributed as dist
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import _lib
    from videoframeinterpolation_b200.workloads import bytes_dattn, init_dat_weights, seed_for

    rank, world, local = dist_env(args.gpus)
    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU fallback in the product path)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert _lib.lib().vfi_device_supported() == 1, "libvfi_b200.so targets sm_100a"
    torch.backends.cudnn.allow_tf32 = bool(args.tf32)
    torch.backends.cuda.matmul.allow_tf32 = bool(args.tf32)
    torch.backends.cudnn.benchmark = True

    torch.manual_seed(0)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval()
    if args.channels_last:
        net = net.to(memory_format=torch.channels_last)
    P, H, W = args.pairs, args.height, args.width
    g = torch.Generator().manual_seed(seed_for(4, rank))
    h_x0 = torch.rand(P, 3, H, W, generator=g).pin_memory()
    h_x1 = torch.rand(P, 3, H, W, generator=g).pin_memory()
    h_out = torch.empty(P, 3, H, W).pin_memory()
    t = torch.full((P, 1, 1, 1), 0.5, device=dev)
    padder = v.InputPadder((H, W), divisor=16)
    d_x0, d_x1 = padder.pad(h_x0.to(dev), h_x1.to(dev))
    if args.channels_last:
        d_x0, d_x1 = d_x0.contiguous(memory_format=torch.channels_last), d_x1.contiguous(memory_format=torch.channels_last)

    d_r0, d_r1 = h_x0.to(dev), h_x1.to(dev)          # the raw (unpadded) float32 frames, resident in HBM

    def step_device():
        # frames in, frame out: pad + rgb-mean normalisation + forward + clamp + unpad (the fused frame edge, row f3)
        return net.interpolate(d_r0, d_r1, t)

    def step_e2e():
        a, b = padder.pad(h_x0.to(dev, non_blocking=True), h_x1.to(dev, non_blocking=True))
        h_out.copy_(padder.unpad(net.inference(a, b, t)), non_blocking=True)

    graphed = None

    step_graph = None

    def step_graph_e2e():
        a, b = padder.pad(h_x0.to(dev, non_blocking=True), h_x1.to(dev, non_blocking=True))
        h_out.copy_(padder.unpad(graphed(a, b)), non_blocking=True)

    # end-to-end leg as a user of the reference feeds it (evaluate.py:138-142): uint8 frames from the host, uint8 frame
    # back; uint8 -> float, pad, rgb-mean normalisation, clamp and unpad run on the device (net.interpolate, row f3)
    h_u0 = (h_x0 * 255).round().to(torch.uint8).pin_memory()
    h_u1 = (h_x1 * 255).round().to(torch.uint8).pin_memory()
    h_out_u8 = torch.empty(P, 3, H, W, dtype=torch.uint8).pin_memory()
    graphed_u8 = None

    def step_e2e_u8():
        h_out_u8.copy_(net.interpolate(h_u0.to(dev, non_blocking=True), h_u1.to(dev, non_blocking=True), t), non_blocking=True)

    def step_graph_e2e_u8():
        h_out_u8.copy_(graphed_u8(h_u0, h_u1), non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, timer=None):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        if timer is not None:
            with timer:
                for _ in range(steps):
                    fn()
        else:
            for _ in range(steps):
                fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    with torch.no_grad():
        for _ in range(max(args.warmup, 3)):
            step_device()
        clock = ClockSampler(local)
        clock.start()
        _lib.reset_launch_count()
        from videoframeinterpolation_b200 import conv as vconv
        vconv.flop_count(reset=True)
        tm = _lib.OpTimer()
        total_ms = timed(step_device, args.steps, tm)
        conv_flops = vconv.flop_count() / args.steps
        launches = _lib.launch_count()
        clocks = clock.stop()
        ops = tm.summary()
        dattn_events = tm.records.get("vfi_dattn_f32", [])
        # host enqueue time of the eager loop (no synchronisation inside): > GPU time means launch-bound
        torch.cuda.synchronize()
        h0 = time.perf_counter()
        for _ in range(args.steps):
            step_device()
        host_ms = (time.perf_counter() - h0) * 1e3 / args.steps
        torch.cuda.synchronize()
        eager_ms = total_ms / args.steps
        graph_pool_gb = None
        if args.graph:
            # one captured graph alive at a time: each holds the activations of a whole forward in its private pool
            # (~4.4 GB per 1080p pair), and --pairs 8 would not fit four of them
            graphed_dev = v.GraphedInference(net, d_r0, d_r1, t, fn=net.interpolate)
            step_graph = graphed_dev.replay
            for _ in range(3):
                step_graph()
            total_ms = timed(step_graph, args.steps)
            free_b, total_b = torch.cuda.mem_get_info(dev)
            graph_pool_gb = (total_b - free_b) / 2 ** 30                      # device-level: includes the graph's private pool
            del graphed_dev, step_graph
            torch.cuda.empty_cache()
            if not args.no_e2e:
                graphed = v.GraphedInference(net, d_x0, d_x1, t)                  # float32 e2e leg (pad / unpad in PyTorch)
        if args.no_e2e:
            e2e_ms = e2e_f32_ms = e2e_seq_ms = float("nan")
        else:
            fn = step_graph_e2e if args.graph else step_e2e
            for _ in range(2):
                fn()
            e2e_f32_ms = timed(fn, args.steps)
            if args.graph:
                graphed = None
                torch.cuda.empty_cache()
                graphed_u8 = v.GraphedInference(net, h_u0.to(dev), h_u1.to(dev), t, fn=net.interpolate)
            fn = step_graph_e2e_u8 if args.graph else step_e2e_u8
            for _ in range(2):
                fn()
            e2e_seq_ms = timed(fn, args.steps)
            e2e_ms = e2e_seq_ms
            if args.graph:
                graphed_u8 = None
                torch.cuda.empty_cache()
                # the streaming API a video interpolator uses: pair i + 1 uploads and result i - 1 downloads on copy streams
                # while pair i computes; every step's H2D and D2H are inside the timed region, the last result included
                pipe = v.PipelinedInference(net, h_u0, h_u1, t, fn=net.interpolate)

                def run_pipe(steps):
                    for _ in range(steps):
                        pipe.submit(h_u0, h_u1)
                        if pipe.pending() > 1:
                            pipe.collect()
                    while pipe.pending():
                        pipe.collect()              # blocks the host until that download has finished

                run_pipe(3)
                e2e_ms = timed(lambda: run_pipe(args.steps), 1)
                del pipe
                torch.cuda.empty_cache()

    ms_per_step = total_ms / args.steps
    value = world * P / (ms_per_step / 1e3)
    e2e_value = world * P / (e2e_ms / args.steps / 1e3)
    hbm_peak, peak_src = peaks()

    # dominant vfi_b200 kernel: fused deformable attention at 1/2 resolution (3rd vfi_dattn_f32 call of each step)
    ph, pw = d_x0.shape[-2:]
    lv1 = [a.elapsed_time(b) for i, (a, b) in enumerate(dattn_events) if i % 3 == 2]
    lv1_ms = sum(lv1) / max(len(lv1), 1)
    alg = bytes_dattn(P, 72, ph // 2, pw // 2, 8, 32)
    achieved = alg / lv1_ms / 1e6 if lv1_ms > 0 else 0.0
    traffic, traffic_src, ncu_d = dattn_ncu_summary()
    fl = dattn_flops(P, 72, ph // 2, pw // 2, 8, 32)
    fp32_tf = fl / lv1_ms / 1e9 if lv1_ms > 0 else 0.0
    roofline = {"kernel": "dattn_win_kernel<9> (fused sampler + attention, 1/2 res: TMA-staged shared-memory window, lane = pixel x group)",
                "bound": "hbm", "achieved": achieved, "peak": hbm_peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / hbm_peak,
                "algorithmic_bytes_per_launch": alg, "ms_per_launch": lv1_ms,
                "frac_fp32": fp32_tf / FP32_PEAK_TFLOPS, "fp32": {"achieved_tflops": fp32_tf, "peak_tflops": FP32_PEAK_TFLOPS,
                                                                  "peak_source": "nominal 148 SM x 128 FMA x 2 x 1.965 GHz",
                                                                  "flops_per_launch": fl},
                "note": "18.4k bilinear taps + ~170 kFLOP fp32 per pixel: the binding resources are shared-memory wavefronts of the "
                        "tap reads and fp32 issue, not HBM (DESIGN.md 4.3); reported against HBM as the contract asks, frac_fp32 beside it",
                "traffic": traffic if (P == 1 and (ph, pw) == (1088, 1920)) else None, "traffic_source": traffic_src}
    op_ms = {k: {"calls_per_step": c / args.steps, "ms_per_step": ms / args.steps} for k, (c, ms) in ops.items()}
    # second kernel family of the step: the tensor-core 3x3 convolutions (all launches of a step together)
    conv_ms = op_ms.get("vfi_conv3x3_f32", {}).get("ms_per_step", 0.0)
    tf32_peak = 1130.0      # TFLOP/s dense TF32, nominal (B200_PROFILING.md); MEASURED_PEAKS.json has no TF32 entry
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            tf32_peak_meas = float(json.load(f)["bf16_tflops"]) / 2.0
    except Exception:
        tf32_peak_meas = None
    conv_tf = conv_flops / conv_ms / 1e9 if conv_ms > 0 else 0.0
    roofline_conv = {"kernel": "conv3x3_tc_kernel (tcgen05 kind::tf32, all 3x3 stride-1 convs of the step)", "bound": "tensor",
                     "achieved": conv_tf, "peak": tf32_peak, "peak_source": "nominal dense TF32 (B200_PROFILING.md)",
                     "peak_half_of_measured_bf16": tf32_peak_meas, "unit": "TFLOP/s", "frac": conv_tf / tf32_peak,
                     "flops_per_step": conv_flops, "ms_per_step": conv_ms, "traffic": None,
                     "note": "steady state within 10% of the kernel's own tcgen05.mma issue rate (27 MMAs of M=256 x N=224 x K=8 per tile pair); 2 halo rows of 16 and the per-launch prologue / tail are the rest, DESIGN.md 4.6"}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32" if not args.tf32 else "f32 (hot path) / tf32 (3x3 convs)",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "pairs_per_gpu": P, "global_pairs_per_step": P * world,
                       "parallelism": f"frame-pair sharding x{world}, no collective", "conv_layout": "channels_last" if args.channels_last else "nchw",
                       "conv_math": "tf32" if args.tf32 else "fp32",
                       "arithmetic": "fp32 hot path (warp / sampler / attention / DCN kernels); 3x3 convs " +
                                     ("TF32 operands with fp32 accumulation on tcgen05 = PyTorch's default cudnn.allow_tf32 conv arithmetic" if args.tf32 else "fp32 (cuDNN)"),
                       "launch": "CUDA graph replay" if args.graph else "eager",
                       "api": "DATwConstantnCv1.interpolate(float32 frames resident in HBM): replicate pad, rgb-mean normalisation, forward, clamp, unpad",
                       "l2": "inputs + activations per step (>1 GB) exceed the 126 MB L2"},
            "details": {"eager_ms_per_step": eager_ms, "eager_host_enqueue_ms_per_step": host_ms,
                        "device_memory_in_use_gb_with_the_captured_graph": graph_pool_gb, "pairs_per_gpu": P},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 2 * h_u0.numel(), "d2h_bytes_per_step": h_out_u8.numel(),
                    "ms_per_step": e2e_ms / args.steps,
                    "api": ("PipelinedInference(net, fn=DATwConstantnCv1.interpolate).submit / collect: " if args.graph else "DATwConstantnCv1.interpolate: ") +
                           "uint8 frames from pinned host memory -> uint8 frame to pinned host memory; "
                           "uint8 -> float, pad, normalise, clamp, unpad on the device (vfi_frame_prep / vfi_frame_finish)" +
                           ("; uploads / downloads of neighbouring pairs overlap the running graph (copy streams), all K results downloaded inside the timed region" if args.graph else ""),
                    "sequential": {"value": world * P / (e2e_seq_ms / args.steps / 1e3), "ms_per_step": e2e_seq_ms / args.steps,
                                   "api": "one pair at a time: upload, graph replay, download, no overlap between steps"}},
            "e2e_f32": {"value": world * P / (e2e_f32_ms / args.steps / 1e3), "unit": UNIT, "h2d_bytes_per_step": 2 * h_x0.numel() * 4,
                        "d2h_bytes_per_step": h_out.numel() * 4, "ms_per_step": e2e_f32_ms / args.steps,
                        "api": "InputPadder.pad + inference + unpad on float32 frames from pinned host memory"},
            "roofline_conv3x3": roofline_conv, "gpu_launches": int(launches),      # vfi_b200 kernels per K steps (counted on the eager pass; the graph replays the same kernels)
            "clocks": clocks, "roofline": roofline, "vfi_ops_ms": op_ms}

    if not args.no_legs:
        with torch.no_grad():
            # BASELINE configs[3]'s own model at every N (weak scaling: one pair per rank) and configs[4] (one 4K pair tiled
            # over the ranks): every rank takes part, rank 0 reports
            dl = dcndat_1080p(args, dev, H, W, world, dist, with_library_baseline=(world == 1 and not args.no_cpu))
            t4 = tiled_4k_leg(net, dev, rank, world, max(args.steps // 2, 3), dist)
            if rank == 0:
                line["dcndat_1080p"] = dl
                line["tiled_4k"] = t4
                line["configs0_448x256"] = configs0_leg(net, dev, args.steps)
                if args.tf32:
                    line["e2e_parity"] = tf32_parity_leg(net, d_x0, d_x1, t)
                    line["value_fp32_convs"] = {"value": world * line["e2e_parity"]["fp32_convs_frames_per_s_eager"], "unit": UNIT,
                                                "note": "same forward with fp32 cuDNN convolutions (--tf32 0), eager launches"}
    if rank == 0:
        if not args.no_micro:
            with torch.no_grad():
                line["microbench"] = microbench(dev, hbm_peak)
        if (args.torch_cuda_baseline or not args.no_cpu) and world == 1:
            tb = torch_cuda_baseline(args, dev, H, W)
            if "value" in tb:
                tb["speedup_of_this_repo"] = (1e3 * args.pairs / line["ms_per_step"]) / tb["value"]   # per GPU, frames/s over frames/s
            line["torch_cuda_baseline"] = tb
        if not args.no_cpu and world == 1:
            cores = os.cpu_count() or 1
            torch.set_num_threads(cores)
            oracle, sd, x0, x1, tt = cpu_dat_setup()
            cpu_dat_step(oracle, sd, x0, x1, tt)
            t0 = time.perf_counter()
            n = 2
            for _ in range(n):
                cpu_dat_step(oracle, sd, x0, x1, tt)
            dt = (time.perf_counter() - t0) / n
            line["cpu_baseline"] = {"value": cpu_scale(H, W) / dt, "unit": UNIT, "cores": cores, "kind": "port",
                                    "sample": cpu_sample_text(cores), "sec_per_sample_forward": dt,
                                    "measured_full_config": CPU_FULL_CONFIG_MEASURED}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def run_tiled(args):
    """BASELINE configs[4] / SURVEY.md 8(e): one 3840x2160 pair, spatial tiles with a halo over the ranks.  Every rank
    holds the whole input pair, computes the global rgb mean itself (identical on all ranks: no all-reduce), runs the
    forward on its tiles + halo and writes their interiors; the stitched frame is the union of the ranks' interiors
    (gathering it is not on the timed path).  value = pairs/s of the WHOLE pair = 1 / max-over-ranks step time."""
    import torch.distributed as dist
    import videoframeinterpolation_b200 as v
    from videoframeinterpolation_b200 import _lib
    from videoframeinterpolation_b200.shard import tile_plan, tiles_of_rank
    from videoframeinterpolation_b200.workloads import init_dat_weights, seed_for
    rank, world, local = dist_env(args.gpus)
    assert torch.cuda.is_available(), "bench.py needs a GPU"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.allow_tf32 = bool(args.tf32)
    torch.backends.cudnn.benchmark = True
    rows, cols = (int(x) for x in args.tile_grid.lower().split("x"))
    torch.manual_seed(0)
    net = init_dat_weights(v.DATwConstantnCv1(v.default_args())).to(dev).eval().to(memory_format=torch.channels_last)
    H, W = args.height, args.width
    g = torch.Generator().manual_seed(seed_for(5, 0))           # the same pair on every rank
    h_x0, h_x1 = torch.rand(1, 3, H, W, generator=g).pin_memory(), torch.rand(1, 3, H, W, generator=g).pin_memory()
    padder = v.InputPadder((H, W), divisor=16)
    d_x0, d_x1 = padder.pad(h_x0.to(dev), h_x1.to(dev))
    t = torch.full((1, 1, 1, 1), 0.5, device=dev)
    ph, pw = d_x0.shape[-2:]
    mine = tiles_of_rank(tile_plan(ph, pw, (rows, cols), args.halo), rank, world)
    out_buf = torch.zeros((1, 3, ph, pw), device=dev)

    def fn(a, b, tt):
        return net.inference_tiled(a, b, tt, grid=(rows, cols), halo=args.halo, rank=rank, world=world, out=out_buf)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.no_grad():
        graphed = v.GraphedInference(net, d_x0, d_x1, t, fn=fn)
        _lib.reset_launch_count()
        fn(d_x0, d_x1, t)
        launches = _lib.launch_count()
        for _ in range(max(args.warmup, 3)):
            graphed.replay()
        clock = ClockSampler(local)
        clock.start()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.steps):
            graphed.replay()
        e1.record()
        barrier()
        clocks = clock.stop()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        ms_per_step = float(ms.item()) / args.steps
    if rank == 0:
        crop_px = sum((tl.crop[0].stop - tl.crop[0].start) * (tl.crop[1].stop - tl.crop[1].start) for tl in tile_plan(ph, pw, (rows, cols), args.halo))
        line = {"metric": METRIC.replace("1080p", f"{W}x{H} tiled"), "value": 1e3 / ms_per_step, "unit": UNIT, "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"DAT inference on ONE synthetic {W}x{H} frame pair (padded to {pw}x{ph}), {rows}x{cols} tiles with a "
                                       f"{args.halo}-px halo dealt round-robin to {world} rank(s) (BASELINE configs[4])",
                           "tiles_per_rank": len(mine), "redundant_pixels": crop_px / float(ph * pw) - 1.0,
                           "parallelism": f"spatial tiles x{world}, no collective", "launch": "CUDA graph replay",
                           "l2": "inputs + activations per step exceed the 126 MB L2"},
                "gpu_launches": int(launches) * args.steps, "clocks": clocks}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def _quiet_native_stdout():
    """Send file descriptor 1 to stderr for the rest of the run (NCCL prints its version banner to stdout) and return
    a text stream on the original stdout for the one JSON line."""
    sys.stdout.flush()
    real = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real, "w", buffering=1)


def main():
    _quiet_native_stdout()
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.tile_grid:
        run_tiled(args)
    else:
        run_vfi(args)


if __name__ == "__main__":
    main()
