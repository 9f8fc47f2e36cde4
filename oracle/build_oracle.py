"""Build recipe for the oracle's native pieces.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

1. ``oracle/softsplat_ref.c`` (our plain-C restatement of the reference splat kernel)
       -> ``oracle/_build/libsplat_oracle.so``                      (always; gcc)
2. when the reference checkout is present (build container only): the reference's OWN kernel text
   (``modules/softsplat.py:289-341``, a CUDA string whose sizes/strides are baked in per shape by
   ``cuda_kernel``, lines 32-221) is expanded for each requested shape and compiled as C++ behind a
   ten-line host shim (``__global__`` etc. defined away, ``atomicAdd`` -> ``+=``)
       -> ``oracle/_ref/splat_ref_N{n}_C{c}_H{h}_W{w}.so``
   Nothing from the reference is copied into the repository: the text is read from where it lies and
   only the compiled object lands in ``oracle/_ref/`` (git-ignored, travels to the GPU box).
"""
from __future__ import annotations

import ctypes
import inspect
import os
import subprocess
import sys
import tempfile
import types
from pathlib import Path

HERE = Path(__file__).resolve().parent
BUILD = HERE / "_build"
REFDIR = HERE / "_ref"
REFERENCE = Path(os.environ.get("VFI_REFERENCE_DIR", "/root/reference"))

_SHIM = r"""
#include <cmath>
#include <cassert>
struct D3 { int x, y, z; };
static D3 blockIdx = {0, 0, 0}, blockDim = {1, 1, 1}, threadIdx = {0, 0, 0}, gridDim = {1, 1, 1};
#define __global__
#define __restrict__
#define __launch_bounds__(x)
static inline void atomicAdd(float* p, float v) { *p += v; }
using std::isfinite; using std::floor;
/* one serial "thread" walks the whole grid-stride loop, so a CUDA thread's early exit is a skip */
#define return continue
"""


def build_c_oracle(force: bool = False) -> Path:
    BUILD.mkdir(exist_ok=True)
    src, dst = HERE / "softsplat_ref.c", BUILD / "libsplat_oracle.so"
    if force or not dst.exists() or dst.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(dst), str(src), "-lm"])
    return dst


def load_c_oracle():
    lib = ctypes.CDLL(str(build_c_oracle()))
    lib.oracle_splat_sum_f32.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int64] * 4
    lib.oracle_splat_sum_f32.restype = None
    return lib


def c_splat_sum(ten_in, ten_flow):
    """Run the C restatement on CPU torch tensors."""
    import torch
    ten_in = ten_in.float().contiguous()
    ten_flow = ten_flow.float().contiguous()
    out = torch.zeros_like(ten_in)
    N, C, H, W = ten_in.shape
    load_c_oracle().oracle_splat_sum_f32(ten_in.data_ptr(), ten_flow.data_ptr(), out.data_ptr(), N, C, H, W)
    return out


# ---------------------------------------------------------------------------------------------
# reference import shim (build container only)
# ---------------------------------------------------------------------------------------------

def reference_available() -> bool:
    return (REFERENCE / "modules" / "softsplat.py").exists()


def import_reference():
    """Make ``import modules.warp`` etc. work against the read-only checkout."""
    import torch
    if "cupy" not in sys.modules:
        cupy = types.ModuleType("cupy")
        cupy.int32, cupy.float32 = int, float
        cupy.ndarray = type("ndarray", (), {})
        cupy.memoize = lambda for_each_device=False: (lambda f: f)
        sys.modules["cupy"] = cupy
    for name in ("timm", "timm.models", "timm.models.layers"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["timm.models.layers"].trunc_normal_ = torch.nn.init.trunc_normal_
    if str(REFERENCE) not in sys.path:
        sys.path.insert(0, str(REFERENCE))


def _reference_kernel_source(N: int, C: int, H: int, W: int) -> str:
    import torch
    import_reference()
    import modules.softsplat as ss
    ss.objCudacache["device"] = "cpu-oracle"
    src = inspect.getsource(ss.softsplat_func.forward)
    start = src.index("cuda_kernel('softsplat_out', '''") + len("cuda_kernel('softsplat_out', '''")
    template = src[start:src.index("''',", start)]
    t_in = torch.zeros(N, C, H, W)
    t_flow = torch.zeros(N, 2, H, W)
    key = ss.cuda_kernel("softsplat_out", template, {"tenIn": t_in, "tenFlow": t_flow, "tenOut": torch.zeros_like(t_in)})
    return ss.objCudacache[key]["strKernel"]


def ref_so_path(N: int, C: int, H: int, W: int) -> Path:
    return REFDIR / f"splat_ref_N{N}_C{C}_H{H}_W{W}.so"


def build_reference_splat(N: int, C: int, H: int, W: int, force: bool = False) -> Path:
    """Compile the reference kernel text for one (contiguous) shape.  Needs the reference checkout."""
    dst = ref_so_path(N, C, H, W)
    if dst.exists() and not force:
        return dst
    if not reference_available():
        raise FileNotFoundError(f"reference checkout not found at {REFERENCE}; {dst.name} must be prebuilt")
    REFDIR.mkdir(exist_ok=True)
    text = _SHIM + _reference_kernel_source(N, C, H, W)
    with tempfile.TemporaryDirectory() as tmp:
        cpp = Path(tmp) / "k.cpp"
        cpp.write_text(text)
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-w", "-shared", "-fPIC", "-o", str(dst), str(cpp)])
    return dst


def reference_splat_sum(ten_in, ten_flow):
    """Run the reference's own kernel text (compiled for this shape) on CPU tensors."""
    import torch
    ten_in = ten_in.float().contiguous()
    ten_flow = ten_flow.float().contiguous()
    N, C, H, W = ten_in.shape
    lib = ctypes.CDLL(str(build_reference_splat(N, C, H, W)))
    lib.softsplat_out.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.softsplat_out.restype = None
    out = torch.zeros_like(ten_in)
    lib.softsplat_out(out.numel(), ten_in.data_ptr(), ten_flow.data_ptr(), out.data_ptr())
    return out


# ---------------------------------------------------------------------------------------------
# the reference's kernel text compiled for sm_100a: the "reference kernel on the same box" bar of SURVEY.md 8(d)
# ---------------------------------------------------------------------------------------------
_CUDA_LAUNCHER = r"""
#include <cuda_runtime.h>
/* launch configuration of modules/softsplat.py:342-348: ceil(n / 512) blocks of 512 threads on the caller's stream */
extern "C" int ref_softsplat_launch(int n, const float* in, const float* flow, float* out, void* stream) {
    softsplat_out<<<(n + 512 - 1) / 512, 512, 0, (cudaStream_t)stream>>>(n, in, flow, out);
    return (int)cudaGetLastError();
}
"""
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")


def ref_cuda_so_path(N: int, C: int, H: int, W: int) -> Path:
    return REFDIR / f"splat_ref_cuda_N{N}_C{C}_H{H}_W{W}.so"


def build_reference_splat_cuda(N: int, C: int, H: int, W: int, force: bool = False) -> Path:
    """The reference's own kernel text (sizes / strides baked in for this contiguous shape, as its NVRTC path does)
    compiled UNMODIFIED with nvcc for sm_100a behind a 4-line launcher that replays its launch configuration."""
    dst = ref_cuda_so_path(N, C, H, W)
    if dst.exists() and not force:
        return dst
    if not reference_available():
        raise FileNotFoundError(f"reference checkout not found at {REFERENCE}; {dst.name} must be prebuilt")
    REFDIR.mkdir(exist_ok=True)
    text = "#include <cassert>\n" + _reference_kernel_source(N, C, H, W) + _CUDA_LAUNCHER
    with tempfile.TemporaryDirectory() as tmp:
        cu = Path(tmp) / "k.cu"
        cu.write_text(text)
        subprocess.check_call([NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-w", "-shared", "-Xcompiler", "-fPIC",
                               "-o", str(dst), str(cu), "-lcudart_static"])
    return dst


def reference_splat_sum_cuda(ten_in, ten_flow, out=None):
    """Launch the reference's kernel (prebuilt for this shape) on CUDA tensors, on the current stream."""
    import torch
    N, C, H, W = ten_in.shape
    lib = ctypes.CDLL(str(build_reference_splat_cuda(N, C, H, W)))
    lib.ref_softsplat_launch.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.ref_softsplat_launch.restype = ctypes.c_int
    out = torch.zeros_like(ten_in) if out is None else out.zero_()
    rc = lib.ref_softsplat_launch(out.numel(), ten_in.data_ptr(), ten_flow.data_ptr(), out.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == 0, f"reference splat kernel launch failed: cudaError {rc}"
    return out


# shapes of the C+1-channel tensors the bench's CPU-baseline leg may time with the reference text
BENCH_REF_SHAPES = [(1, 4, 1080, 1920), (1, 4, 270, 480)]
# C + 1 channels (the reference concatenates the weight channel, modules/softsplat.py:260-274) at the microbench's 1080p rows
BENCH_REF_CUDA_SHAPES = [(1, 4, 1080, 1920), (1, 65, 1080, 1920)]


def build_all(verbose: bool = True) -> None:
    p = build_c_oracle()
    if verbose:
        print(f"[oracle] built {p}")
    if reference_available():
        for shp in BENCH_REF_SHAPES:
            q = build_reference_splat(*shp)
            if verbose:
                print(f"[oracle] built {q}")
        for shp in BENCH_REF_CUDA_SHAPES:
            q = build_reference_splat_cuda(*shp)
            if verbose:
                print(f"[oracle] built {q}")
    elif verbose:
        print("[oracle] reference checkout absent: oracle/_ref/ uses prebuilt files only")


if __name__ == "__main__":
    build_all()
