#!/usr/bin/env python
"""Time the reference's CPU path (oracle port with torch CPU grid_sample / torchvision deform_conv2d) at a given frame size
and report peak RSS -- the feasibility probe behind bench.py's --impl reference arm.
    python tools/cpu_ref_probe.py [H W] [--mem-gb N]   (address-space cap so that a too-large forward raises instead of
                                                          driving the box out of memory)"""
import os, resource, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
args = [a for a in sys.argv[1:] if not a.startswith("--")]
H, W = (int(args[0]), int(args[1])) if len(args) >= 2 else (1088, 1920)
cap = [a for a in sys.argv[1:] if a.startswith("--mem-gb")]
total_gb = os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES") / 2**30
cap_gb = float(sys.argv[sys.argv.index("--mem-gb") + 1]) if "--mem-gb" in sys.argv else 0.8 * total_gb
resource.setrlimit(resource.RLIMIT_AS, (int(cap_gb * 2**30), int(cap_gb * 2**30)))
import torch
import bench
cores = os.cpu_count() or 1
torch.set_num_threads(cores)
print(f"host: {cores} cores, {total_gb:.0f} GiB RAM, cap {cap_gb:.0f} GiB; frame {W}x{H}", flush=True)
bench.CPU_SAMPLE_HW = (H, W)
oracle, sd, x0, x1, t = bench.cpu_dat_setup()
try:
    for i in range(2):
        t0 = time.perf_counter()
        y = bench.cpu_dat_step(oracle, sd, x0, x1, t)
        dt = time.perf_counter() - t0
        print(f"forward {i}: {dt:.2f} s, peak RSS {resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 2**20:.1f} GiB", flush=True)
except (MemoryError, RuntimeError) as e:
    print(f"FAILED at {W}x{H}: {type(e).__name__}: {str(e)[:300]}", flush=True)
