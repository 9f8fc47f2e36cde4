"""In-tree build of libvfi_b200.so (hand-written CUDA for sm_100a behind the C ABI of include/vfi_b200.h).

    python -m videoframeinterpolation_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU; the built .so sits next to this file (git-ignored, shipped to the GPU
box by gpurun).  One object per translation unit, compiled in parallel, then linked.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

PKG = Path(__file__).resolve().parent
CSRC = PKG / "csrc"
OBJ = PKG / "build"
LIB = PKG / "libvfi_b200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "--expt-relaxed-constexpr"]


def sources():
    return sorted(CSRC.glob("*.cu"))


def _newest_header() -> float:
    hs = list(CSRC.glob("*.cuh")) + [PKG.parent / "include" / "vfi_b200.h"]
    return max(h.stat().st_mtime for h in hs)


def _compile(src: Path, verbose: bool) -> Path:
    obj = OBJ / (src.stem + ".o")
    cmd = [NVCC, *FLAGS, "-c", str(src), "-o", str(obj)]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{r.stdout}\n{r.stderr}")
    if verbose:
        print(r.stderr)
    return obj


def build(force: bool = False, verbose: bool = False) -> Path:
    OBJ.mkdir(exist_ok=True)
    srcs = sources()
    hdr = _newest_header()
    stale = [s for s in srcs if force or not (OBJ / (s.stem + ".o")).exists()
             or (OBJ / (s.stem + ".o")).stat().st_mtime < max(s.stat().st_mtime, hdr)]
    if stale:
        with ThreadPoolExecutor(max_workers=min(8, len(stale))) as ex:
            list(ex.map(lambda s: _compile(s, verbose), stale))
    objs = [OBJ / (s.stem + ".o") for s in srcs]
    if stale or not LIB.exists() or LIB.stat().st_mtime < max(o.stat().st_mtime for o in objs):
        cmd = [NVCC, "-shared", "-o", str(LIB), *map(str, objs), "-gencode", "arch=compute_100a,code=sm_100a",
               "-Xcompiler", "-fPIC", "-lcudart_static", "-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(a.force, a.verbose))
