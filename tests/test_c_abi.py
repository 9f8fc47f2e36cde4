"""The C-ABI library loads on a CPU-only box and exports every entry point include/vfi_b200.h declares, and the ctypes
signature table of the host side covers exactly those (no compute calls: there is no GPU here)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "vfi_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return set(re.findall(r"\b(vfi_[a-z0-9_]+)\s*\(", text))


def test_library_exports_every_declared_symbol():
    from videoframeinterpolation_b200 import _lib, build
    path = build.build(force=False)
    handle = ctypes.CDLL(str(path))
    declared = _declared()
    assert len(declared) >= 25, sorted(declared)
    missing = [n for n in sorted(declared) if not hasattr(handle, n)]
    assert not missing, f"declared in include/vfi_b200.h but not exported: {missing}"
    assert _lib.lib().vfi_version() >= 100
    assert isinstance(_lib.lib().vfi_last_error_string(), (bytes, type(None)))


def test_ctypes_table_matches_header():
    from videoframeinterpolation_b200 import _lib
    declared, bound = _declared(), set(_lib.SIGNATURES)
    assert bound <= declared | {"vfi_launch_count", "vfi_launch_count_reset"}, sorted(bound - declared)
    unbound = declared - bound
    assert not unbound, f"declared but not bound by the host side: {sorted(unbound)}"


def test_cpu_tensors_are_refused_not_computed():
    import pytest
    import torch
    import videoframeinterpolation_b200 as v
    with pytest.raises(Exception):
        v.bwarp(torch.zeros(1, 3, 8, 8), torch.zeros(1, 2, 8, 8))
    with pytest.raises(Exception):
        v.deform_conv2d(torch.zeros(1, 8, 8, 8), torch.zeros(1, 18, 8, 8), torch.zeros(8, 8, 3, 3))
