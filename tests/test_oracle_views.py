"""Pin ``oracle.views`` to the reference: golden fixtures generated from the reference's own
classes (tests/golden/make_golden.py) and, when /root/reference is mounted, a live replay."""
import contextlib
import io
import os
import random

import numpy as np
import pytest
import torch

from oracle import views

CFG = {"pretrain": {"normalization": "mean_var", "augmentations": {
    "MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
    "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                         "time_crop_scale": [0.6, 1.5]}}}}


def test_two_views_match_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "views_reference.npz"))
    np.random.seed(int(g["seed"])); random.seed(int(g["seed"]))
    orc = views.TwoViewOracle(int(g["len_of_files"]))
    v1, v2, recs = orc.run_batch(torch.from_numpy(g["x"][:, 0]))
    assert np.array_equal(v1.numpy(), g["view1"])
    assert np.array_equal(v2.numpy(), g["view2"])
    arrs = views.records_to_arrays(recs)
    for key in ("crop", "src_clip", "bank_index"):
        assert np.array_equal(arrs[key], g[key]), key
    assert np.array_equal(arrs["alpha"], g["alpha"])
    assert np.array_equal(arrs["mean_std"], g["mean_std"])


def test_explicit_parameters_reproduce_golden(golden_dir):
    """The pieces, driven by the recorded parameters only (no RNG), rebuild the reference views."""
    g = np.load(os.path.join(golden_dir, "views_reference.npz"))
    x = torch.from_numpy(g["x"][:, 0])                         # [B, F, T]
    xn = [(x[b:b + 1] - g["mean_std"][b, 0]) / g["mean_std"][b, 1] for b in range(x.shape[0])]
    for b in range(x.shape[0]):
        for v, key in enumerate(("view1", "view2")):
            src = int(g["src_clip"][b, v])
            if src < 0:
                mixed = xn[b]
            else:
                mixed = views.log_mixup_exp(xn[b], xn[src], 1.0 - float(g["alpha"][b, v]))
            out = views.resize_crop(mixed, tuple(int(c) for c in g["crop"][b, v]))
            # alpha is stored as fp32 in the fixture, the reference mixes with the python double
            np.testing.assert_allclose(out.numpy(), g[key][b], rtol=1e-5, atol=1e-6)


def test_running_norm_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "running_norm_reference.npz"))
    rn = views.RunningNormOracle(int(g["epoch_samples"]))
    ms = []
    for c in torch.from_numpy(g["x"]):
        out = rn(c)
        ms.append((float(rn.mean), float(rn.std)))
    assert np.array_equal(np.asarray(ms, np.float32), g["mean_std"])
    assert np.array_equal(out.numpy(), g["last_out"])
    assert rn.n == 250                                         # froze at max_update
    # off-by-one closed form: mu_k = mean(m_1..m_k) for k >= 1
    m = g["x"].reshape(300, -1).mean(1, dtype=np.float64)
    np.testing.assert_allclose(g["mean_std"][100, 0], m[1:101].mean(), rtol=1e-5)
    np.testing.assert_allclose(g["mean_std"][299, 0], m[1:250].mean(), rtol=1e-5)


def test_normalize_batch_matches_golden(golden_dir):
    g = np.load(os.path.join(golden_dir, "normalize_batch_reference.npz"))
    y = views.normalize_batch(torch.from_numpy(g["x"]))
    assert np.array_equal(y.numpy(), g["y"])


@pytest.mark.parametrize("shape", [(64, 96), (64, 101), (16, 33)])
def test_live_replay_against_reference(reference_aug, shape):
    f, t = shape
    gen = torch.Generator().manual_seed(f * 1000 + t)
    x = (torch.randn(30, 1, f, t, generator=gen) * 1.1 - 2.5).to(torch.float32)
    np.random.seed(31); random.seed(31)
    with contextlib.redirect_stdout(io.StringIO()):
        mod = reference_aug.AugmentationModule(CFG, 4)
    ref = [mod(c.clone()) for c in x]
    np.random.seed(31); random.seed(31)
    orc = views.TwoViewOracle(4)
    for c, (r1, r2) in zip(x, ref):
        o1, o2 = orc(c.clone())
        assert torch.equal(o1, r1) and torch.equal(o2, r2)
    assert orc.pre_norm.n == mod.pre_norm.ema_mean.n
    # bank bookkeeping: view 2 of a clip may draw the clip itself
    assert any(r.views[1].src_clip == r.clip for r in orc.records) or len(orc.records) < 5


def test_live_get_params_draw_order(reference_aug):
    np.random.seed(3); random.seed(3)
    a = [reference_aug.RandomResizeCrop.get_params((64, 144), (64, 96), (0.6, 1.5), (0.6, 1.5)) for _ in range(200)]
    np.random.seed(3); random.seed(3)
    b = [views.draw_crop((64, 144), (64, 96)) for _ in range(200)]
    assert [tuple(int(v) for v in p) for p in a] == b
    assert any(p[2] == 64 for p in b)          # the skipped-draw branch (h == canvas_h) is exercised


@pytest.mark.parametrize("mode", ["bicubic", "bilinear"])
@pytest.mark.parametrize("hw", [(38, 57), (64, 144), (1, 1), (5, 1), (64, 96)])
def test_scalar_resampler_matches_aten(mode, hw):
    gen = torch.Generator().manual_seed(hw[0] * 7 + hw[1])
    p = torch.randn(*hw, generator=gen)
    want = torch.nn.functional.interpolate(p[None, None], size=(64, 96), mode=mode, align_corners=True)[0, 0]
    got = views.resample_numpy(p.numpy(), (64, 96), mode)
    np.testing.assert_allclose(got, want.numpy(), rtol=1e-5, atol=5e-6)


def test_identity_crop_is_exact():
    gen = torch.Generator().manual_seed(9)
    x = torch.randn(1, 64, 96, generator=gen)
    out = views.resize_crop(x, (0, (144 - 96) // 2, 64, 96))
    assert torch.equal(out, x)
