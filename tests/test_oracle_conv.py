"""CPU checks of the oracle's TF32 convolution restatement (f2) -- known-answer values for the two roundings and
agreement with the fp32 convolution at TF32 precision."""
import torch
import torch.nn.functional as F

import oracle
from _util import rel_err


def test_tf32_roundings_known_answers():
    one_plus = torch.tensor([1.0 + 2.0 ** -11, 1.0 + 2.0 ** -10, 1.0 + 3 * 2.0 ** -11, -(1.0 + 2.0 ** -11), 3.0e38, 0.0])
    # TF32 keeps 10 mantissa bits: ulp(1) = 2^-10
    assert torch.equal(oracle.tf32_truncate(one_plus)[:4], torch.tensor([1.0, 1.0 + 2.0 ** -10, 1.0 + 2.0 ** -10, -1.0]))
    # ties away from zero
    assert torch.equal(oracle.tf32_round(one_plus)[:4],
                       torch.tensor([1.0 + 2.0 ** -10, 1.0 + 2.0 ** -10, 1.0 + 2.0 ** -9, -(1.0 + 2.0 ** -10)]))
    assert oracle.tf32_truncate(one_plus)[5] == 0.0 and oracle.tf32_round(one_plus)[5] == 0.0
    x = torch.randn(1000)
    assert float(((oracle.tf32_truncate(x) - x).abs() / x.abs()).max()) < 2.0 ** -10
    assert float(((oracle.tf32_round(x) - x).abs() / x.abs()).max()) <= 2.0 ** -11


def test_conv3x3_tf32_close_to_fp32_and_epilogue_order():
    g = torch.Generator().manual_seed(0)
    x = torch.randn(2, 16, 9, 11, generator=g)
    w = torch.randn(8, 16, 3, 3, generator=g) * 0.1
    b = torch.randn(8, generator=g)
    s = torch.rand(8, generator=g)
    r = torch.randn(2, 8, 9, 11, generator=g)
    y = oracle.conv3x3_tf32(x, w, b)
    assert rel_err(y, F.conv2d(x.double(), w.double(), b.double(), padding=1)) < 3e-3
    # representable inputs: exact
    xq, wq = oracle.tf32_truncate(x), oracle.tf32_round(w)
    assert torch.equal(oracle.conv3x3_tf32(xq, wq, b), F.conv2d(xq.double(), wq.double(), b.double(), padding=1))
    pre = oracle.conv3x3_tf32(x, w, b, s, r, residual_after_act=False)
    post = oracle.conv3x3_tf32(x, w, b, s, r, residual_after_act=True)
    assert torch.equal(pre, F.prelu(r.double() + y, s.double()))
    assert torch.equal(post, r.double() + F.prelu(y, s.double()))
