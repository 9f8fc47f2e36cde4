"""ConvTranspose2d(4, 2, 1) == PixelShuffle(2) o Conv2d(3, 1, 1) with the re-indexed weights (host logic, CPU)."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F


@pytest.mark.parametrize("cin,cout,h,w", [(8, 8, 5, 7), (12, 2, 6, 4), (4, 20, 3, 3)])
def test_transpose_as_conv_matches_conv_transpose(cin, cout, h, w):
    from videoframeinterpolation_b200.conv import transpose_as_conv
    torch.manual_seed(0)
    ct = nn.ConvTranspose2d(cin, cout, 4, 2, 1).double()
    conv = transpose_as_conv(ct)
    assert conv.weight.shape == (4 * cout, cin, 3, 3)
    x = torch.randn(2, cin, h, w, dtype=torch.float64)
    with torch.no_grad():
        ref = ct(x)
        got = F.pixel_shuffle(conv(x), 2)
    assert got.shape == ref.shape
    assert torch.allclose(got, ref, atol=1e-12, rtol=1e-12)
    # the cache follows in-place parameter updates
    with torch.no_grad():
        ct.weight.mul_(2.0)
        assert transpose_as_conv(ct) is not conv
        assert torch.allclose(F.pixel_shuffle(transpose_as_conv(ct)(x), 2), ct(x), atol=1e-12, rtol=1e-12)
