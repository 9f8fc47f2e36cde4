"""Deterministic, version-independent test data (numpy RandomState keyed by a name).

Used by ``tests/golden/make_golden.py`` (run once in the build container against the real reference)
and by the tests themselves, so fixtures only need to store the reference's OUTPUTS.
"""
from __future__ import annotations

import math
import zlib

import numpy as np
import torch


def _rng(name: str, seed: int) -> np.random.RandomState:
    return np.random.RandomState((zlib.crc32(name.encode()) ^ (seed * 2654435761)) & 0x7FFFFFFF)


def normal(name: str, shape, seed: int = 0, std: float = 1.0, mean: float = 0.0) -> torch.Tensor:
    a = _rng(name, seed).standard_normal(size=tuple(shape)).astype(np.float32) * np.float32(std) + np.float32(mean)
    return torch.from_numpy(a)


def uniform(name: str, shape, seed: int = 0, lo: float = 0.0, hi: float = 1.0) -> torch.Tensor:
    a = _rng(name, seed).uniform(lo, hi, size=tuple(shape)).astype(np.float32)
    return torch.from_numpy(a)


def flow_field(name: str, B: int, H: int, W: int, seed: int = 0, sigma: float = 3.0, outlier_frac: float = 0.02,
               nonfinite_frac: float = 0.0) -> torch.Tensor:
    """(B,2,H,W) flow: N(0, sigma^2) per pixel, ``outlier_frac`` of pixels leaving the frame
    (uniform in [-W, W]), optionally a few NaN / +-Inf entries (splat skip path)."""
    r = _rng(name, seed)
    f = r.standard_normal(size=(B, 2, H, W)).astype(np.float32) * np.float32(sigma)
    if outlier_frac > 0:
        m = r.uniform(size=(B, 1, H, W)) < outlier_frac
        big = r.uniform(-W, W, size=(B, 2, H, W)).astype(np.float32)
        f = np.where(m, big, f)
    if nonfinite_frac > 0:
        m = r.uniform(size=(B, 2, H, W)) < nonfinite_frac
        bad = r.choice(np.array([np.nan, np.inf, -np.inf], dtype=np.float32), size=(B, 2, H, W))
        f = np.where(m, bad, f)
    return torch.from_numpy(np.ascontiguousarray(f.astype(np.float32)))


def fill_state_dict(sd, seed: int = 0):
    """Overwrite every entry of a state dict in place with name-keyed deterministic values.

    Convs / linears: N(0, 1/fan_in); biases: N(0, 0.05^2); PReLU slopes 0.25 +- 0.05; LayerNorm
    weights 1 +- 0.1.  Heads the reference zero-initialises (modules/dcnv2.py:35,
    modules/deformable_attn.py:64) get non-zero values too, otherwise every residual offset is 0.
    """
    for name, ten in sd.items():
        shape = tuple(ten.shape)
        if ten.ndim >= 2:
            fan_in = int(np.prod(shape[1:]))
            val = normal(name, shape, seed, std=1.0 / math.sqrt(fan_in))
        elif name.endswith("bias"):
            val = normal(name, shape, seed, std=0.05)
        elif "norm" in name:
            val = normal(name, shape, seed, std=0.1, mean=1.0)
        else:                                     # PReLU slopes
            val = normal(name, shape, seed, std=0.05, mean=0.25)
        with torch.no_grad():
            ten.copy_(val)
    return sd
