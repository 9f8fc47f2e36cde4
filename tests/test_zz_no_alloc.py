"""The allocation-free entry point vfi_frame_metrics_static_f32 (scratch in a file-scope device array; no counterpart in
the reference) against vfi_frame_metrics_f32 with a caller-supplied
workspace.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(a, b, val_range=1.0):
    from videoframeinterpolation_b200 import _lib
    dev = a.device
    B, C, H, W = a.shape
    st = _lib.stream_ptr(dev)
    ref = torch.empty((B, 3), dtype=torch.float64, device=dev)
    nbytes = int(_lib.lib().vfi_frame_metrics_workspace_bytes(B, C, H, W))
    ws = torch.empty((max(nbytes // 8, 1),), dtype=torch.float64, device=dev)
    _lib.call("vfi_frame_metrics_f32", dev, a.data_ptr(), b.data_ptr(), ref.data_ptr(), B, C, H, W, float(val_range),
              ws.data_ptr(), ws.numel() * 8, st)
    out = torch.full((B, 3), float("nan"), dtype=torch.float64, device=dev)
    _lib.call("vfi_frame_metrics_static_f32", dev, a.data_ptr(), b.data_ptr(), out.data_ptr(), B, C, H, W, float(val_range), st)
    torch.cuda.synchronize()
    return out.cpu(), ref.cpu()


def test_static_scratch_metrics_match_the_workspace_entry_point_twice_in_a_row():
    from videoframeinterpolation_b200 import _lib
    assert int(_lib.lib().vfi_frame_metrics_static_max_bytes()) == 1 << 20
    g = torch.Generator().manual_seed(5)
    # first call: large enough to fill a good part of the array; second call: a smaller, different shape, so stale
    # partials of the first call would show if the kernel read anything it had not written
    for shape in [(2, 3, 540, 960), (1, 3, 97, 131)]:
        a = torch.rand(shape, generator=g).cuda()
        b = (a + 0.05 * torch.randn(shape, generator=g).cuda()).clamp(0, 1)
        out, ref = _both(a, b)
        assert torch.isfinite(out).all()
        rel = float(((out - ref).abs() / ref.abs().clamp_min(1e-30)).max())
        assert rel <= 1e-5, f"{shape}: static-scratch metrics differ from the workspace entry point by {rel:.3e}"
        # the two entry points run the same kernels on the same partial layout: bitwise equal in practice
        assert torch.equal(out, ref), f"{shape}: not bitwise equal"


def test_static_scratch_refuses_shapes_beyond_its_array():
    from videoframeinterpolation_b200 import _lib
    B, C, H, W = 2, 32, 1024, 1024                     # 1.5 MiB of partials > the 1 MiB array
    assert int(_lib.lib().vfi_frame_metrics_workspace_bytes(B, C, H, W)) > int(_lib.lib().vfi_frame_metrics_static_max_bytes())
    a = torch.zeros((B, C, H, W), device="cuda")
    out = torch.zeros((B, 3), dtype=torch.float64, device="cuda")
    with pytest.raises(_lib.VfiError, match="built-in array"):
        _lib.call("vfi_frame_metrics_static_f32", a.device, a.data_ptr(), a.data_ptr(), out.data_ptr(), B, C, H, W, 1.0,
                  _lib.stream_ptr(a.device))
