"""Reference checkpoint layouts load into the drop-in model (Trainer.py:26-29, 43-54)."""
import io

import torch

import golden_cases as gc  # noqa: F401  (sys.path set up by conftest)


def _model():
    import videoframeinterpolation_b200 as v
    return v.DATwConstantnCv1(v.default_args(nf=16, enc_res_blocks=1, dec_res_blocks=1))


def test_state_dict_keys_match_reference_layout():
    """Parameter names are the reference's (models/DAT.py, modules/*): spot-check the keys SURVEY.md 8(b) lists."""
    keys = set(_model().state_dict().keys())
    import oracle
    want = set(oracle.dat_param_shapes(oracle.dat_config(nf=16, enc_res_blocks=1, dec_res_blocks=1)).keys())
    assert want == keys, (sorted(want - keys)[:5], sorted(keys - want)[:5])
    assert any("conv_res_offset" in k for k in keys) and any("movement_extractor" in k for k in keys)
    assert any(k.endswith("q_proj.weight") for k in keys) and any("mlp.fc1" in k for k in keys)


def test_load_ddp_prefixed_and_optimizer_key_variants():
    import videoframeinterpolation_b200 as v
    src, dst = _model(), _model()
    with torch.no_grad():
        for p in src.parameters():
            p.normal_()
    opt = torch.optim.AdamW(src.parameters(), lr=1e-4)
    buf = io.BytesIO()
    v.save_reference_checkpoint(src, buf, epoch=3, step=1234, best_psnr=35.5, optimizer=opt, ddp_prefix=True)
    buf.seek(0)
    meta = v.load_reference_checkpoint(dst, buf)
    assert meta["epoch"] == 3 and meta["step"] == 1234 and meta["best_psnr"] == 35.5 and "optimizer" in meta
    for (ka, a), (kb, b) in zip(src.state_dict().items(), dst.state_dict().items()):
        assert ka == kb and torch.equal(a, b)
    # the spelling the reference's load_trained() reads ('optim'), and a bare state_dict
    meta = v.load_reference_checkpoint(_model(), {"model": src.state_dict(), "optim": opt.state_dict()})
    assert "optimizer" in meta
    assert v.load_reference_checkpoint(_model(), src.state_dict()) == {}
    assert set(v.strip_ddp_prefix({"module.a": 1, "b": 2}).keys()) == {"module.a", "b"}   # mixed keys: left alone
