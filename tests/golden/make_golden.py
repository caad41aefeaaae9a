"""Generate the golden fixtures in this directory.  Run HERE (the build container), where the
reference tree is mounted read-only at /root/reference:

    python tests/golden/make_golden.py

* ``views_reference.npz``  -- the reference's OWN ``AugmentationModule``
  (/root/reference/src/augmentations/__init__.py:8-35) run on seeded synthetic log-mels with
  ``np.random.seed(31); random.seed(31)`` (the seed the reference scripts use,
  extras/delores-s/main.py:64).  Holds inputs, both views, and the per-clip / per-view
  parameters (mean, std, alpha, bank index, source clip, crop) recorded by replaying the same
  seeds through ``oracle.views.TwoViewOracle`` -- the replay is asserted bit-identical to the
  reference output before anything is written.
* ``running_norm_reference.npz`` -- the reference's ``RunningNorm`` (augmentations.py:259-282)
  over 300 tiny clips incl. the freeze after ``max_update``: per-clip (mean, std).
* ``normalize_batch_reference.npz`` -- the reference's ``NormalizeBatch`` (:305-324).
* ``frontend_crosscheck.npz`` -- librosa is absent, so these are NOT reference outputs: they are
  torchaudio's Slaney MelSpectrogram (fp32 FFT) + the oracle's own output on short seeded
  waveforms, kept as a cross-check / regression vector for ``oracle.frontend``.
"""
import importlib
import os
import random
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from oracle import frontend, views  # noqa: E402

CFG = {"pretrain": {"normalization": "mean_var", "augmentations": {
    "MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
    "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                         "time_crop_scale": [0.6, 1.5]}}}}


def main():
    ref = importlib.import_module("src.augmentations")

    # ---- two-view pipeline --------------------------------------------------------------
    g = torch.Generator().manual_seed(1234)
    n_clips, f, t = 6, 64, 96
    x = (torch.randn(n_clips, 1, f, t, generator=g) * 1.3 - 3.0).to(torch.float32)
    np.random.seed(31); random.seed(31)
    mod = ref.AugmentationModule(CFG, 3)
    out = [mod(c.clone()) for c in x]
    v1 = torch.stack([o[0] for o in out]).numpy()
    v2 = torch.stack([o[1] for o in out]).numpy()
    np.random.seed(31); random.seed(31)
    orc = views.TwoViewOracle(3)
    o1, o2, recs = orc.run_batch(x[:, 0])
    assert np.array_equal(o1.numpy(), v1) and np.array_equal(o2.numpy(), v2), "oracle != reference"
    arrs = views.records_to_arrays(recs)
    np.savez_compressed(os.path.join(HERE, "views_reference.npz"), x=x.numpy(), view1=v1, view2=v2,
                        len_of_files=3, seed=31, **arrs)

    # ---- running norm, long sequence with freeze ------------------------------------------
    g = torch.Generator().manual_seed(77)
    xs = (torch.randn(300, 1, 8, 12, generator=g) * torch.linspace(0.5, 2.0, 300).view(-1, 1, 1, 1)
          + torch.linspace(-4, -2, 300).view(-1, 1, 1, 1)).to(torch.float32)
    rn = ref.RunningNorm(epoch_samples=25)            # max_update = 250 -> clips 250.. use frozen stats
    ms = []
    last = None
    for c in xs:
        last = rn(c.clone())
        ms.append((float(rn.mean), float(rn.std)))
    np.savez_compressed(os.path.join(HERE, "running_norm_reference.npz"), x=xs.numpy(),
                        mean_std=np.asarray(ms, np.float32), epoch_samples=25, last_out=last.numpy())

    # ---- normalize batch ------------------------------------------------------------------
    g = torch.Generator().manual_seed(5)
    xb = (torch.randn(5, 1, 16, 20, generator=g) * 2.5 + 0.7).to(torch.float32)
    nb = ref.NormalizeBatch()(xb)
    np.savez_compressed(os.path.join(HERE, "normalize_batch_reference.npz"), x=xb.numpy(), y=nb.numpy())

    # ---- front-end cross-check (torchaudio, not the reference) ------------------------------
    import torchaudio
    g = torch.Generator().manual_seed(1234)
    tt = torch.arange(4000) / 16000.0
    waves = torch.stack([
        torch.randn(4000, generator=g) * 0.1,
        0.5 * torch.sin(2 * np.pi * (50 + (7900 - 50) * tt / (2 * tt[-1])) * tt) + 1e-3 * torch.randn(4000, generator=g),
        0.3 * torch.sin(2 * np.pi * 440 * tt) + 0.2 * torch.sin(2 * np.pi * 3000 * tt) + 1e-4 * torch.randn(4000, generator=g),
    ]).to(torch.float32)
    ta = torchaudio.transforms.MelSpectrogram(16000, n_fft=1024, hop_length=160, f_min=60, f_max=7800,
                                              n_mels=64, power=2, center=True, pad_mode="reflect",
                                              norm="slaney", mel_scale="slaney")
    lt = (ta(waves) + torch.finfo().eps).log().numpy()
    lo = frontend.log_mel_batch(waves.numpy())
    fb = torchaudio.functional.melscale_fbanks(513, 60, 7800, 64, 16000, "slaney", "slaney").T.numpy()
    np.savez_compressed(os.path.join(HERE, "frontend_crosscheck.npz"), waves=waves.numpy(),
                        logmel_torchaudio=lt, logmel_oracle=lo, fbank_torchaudio=fb.astype(np.float32))
    print("wrote fixtures:", sorted(p for p in os.listdir(HERE) if p.endswith(".npz")))


if __name__ == "__main__":
    main()
