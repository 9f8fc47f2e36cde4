"""vfi_nhwc_to_nchw_f32: the channels-last -> planar copy in front of the gather kernels is a pure permutation."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape", [(1, 72, 33, 47, 72), (2, 72, 16, 128, 76), (1, 6, 7, 5, 6), (3, 37, 9, 11, 40),
                                   (1, 200, 12, 13, 200), (1, 72, 544, 960, 72)])
def test_nhwc_to_nchw_is_exact(shape):
    from videoframeinterpolation_b200 import _lib
    B, C, H, W, ctot = shape
    torch.manual_seed(0)
    full = torch.randn(B, H, W, ctot, device="cuda").permute(0, 3, 1, 2)
    x = full[:, ctot - C:]
    assert not x.is_contiguous()
    with torch.no_grad():
        y = _lib.planar(x)
    assert y.is_contiguous() and y.shape == x.shape
    assert torch.equal(y, x.contiguous())


def test_planar_passthrough():
    from videoframeinterpolation_b200 import _lib
    x = torch.randn(1, 8, 5, 6, device="cuda")
    assert _lib.planar(x) is x
