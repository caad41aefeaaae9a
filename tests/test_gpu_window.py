"""GPU parity of the step before the front end: extract_window + L2 normalise
(/root/reference/src/utils/utils.py:166-182, src/dataset/upstream_dataset.py:60-63), and the raw-clip entry
of the fused pipeline with the reference's draw order (window offset, then the clip's view parameters)."""
import random

import numpy as np
import pytest
import torch

from oracle import frontend as ofe
from oracle import views as ov

pytestmark = pytest.mark.gpu

CFG = {"pretrain": {"normalization": "mean_var", "input": {"length_wave": 0.5},
                    "augmentations": {"MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
                                      "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                                                           "time_crop_scale": [0.6, 1.5]}}}}


def _raw_clips(lengths, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    return [(torch.randn(n, generator=g) * 0.1).to(torch.float32) for n in lengths]


def test_windows_match_oracle_bit_exact():
    import audiossl_b200 as A
    unit = 4000
    lengths = [4000, 1, 3999, 4001, 1234, 9000, 4003, 17, 40000, 2000]      # shorter, equal, longer, odd
    clips = _raw_clips(lengths, 5)
    starts = [0 if n <= unit else (n - unit) * (i + 1) // 11 for i, n in enumerate(lengths)]
    starts[5] = 9000 - unit                                                 # last admissible offset
    out, gain = A.extract_windows(clips, unit, starts=starts, with_gain=True)
    assert tuple(out.shape) == (len(clips), unit)
    got, gain = out.cpu().numpy(), gain.cpu().numpy()
    for b, (c, s) in enumerate(zip(clips, starts)):
        want, _ = ofe.extract_window(c.numpy(), unit, start=s)
        assert np.array_equal(got[b], want), f"clip {b} (len {lengths[b]}, start {s})"
        nrm = np.sqrt(np.sum(want.astype(np.float64) ** 2))
        assert abs(gain[b] * max(nrm, 1e-12) - 1.0) <= 1e-5
    # no offsets given: crop from the start, like random.randint never being called
    out0 = A.extract_windows(clips, unit).cpu().numpy()
    for b, c in enumerate(clips):
        assert np.array_equal(out0[b], ofe.extract_window(c.numpy(), unit, start=0)[0])


def test_single_clip_replays_python_random():
    import audiossl_b200 as A
    clip = _raw_clips([50000], 6)[0]
    random.seed(31)
    got = A.extract_window(clip, data_size=1.0).cpu().numpy()              # utils.py:168: unit = int(data_size * 16000)
    random.seed(31)
    want, start = ofe.extract_window(clip.numpy(), 16000, rng=random)
    assert start > 0 and np.array_equal(got, want)
    short = _raw_clips([700], 7)[0]
    state = random.getstate()
    got = A.extract_window(short, duration=1600).cpu().numpy()
    assert random.getstate() == state                                       # nothing drawn for a short clip
    assert np.array_equal(got, ofe.extract_window(short.numpy(), 1600)[0])


def test_l2_gain_feeds_the_front_end():
    import audiossl_b200 as A
    unit = 8000
    clips = _raw_clips([12000, 8000, 5000], 8)
    starts = [1000, 0, 0]
    win, gain = A.extract_windows(clips, unit, starts=starts, with_gain=True)
    mel = A.MelSpectrogramLibrosa()
    t = mel.plan.num_frames(unit)
    out = torch.empty((3, 64, t), dtype=torch.float32, device=win.device)
    mel.plan.logmel_into(win, out, gain=gain)
    for b, (c, s) in enumerate(zip(clips, starts)):
        w = ofe.l2_normalize(ofe.extract_window(c.numpy(), unit, start=s)[0])
        want = ofe.log_mel(w)
        got = out[b].cpu().numpy()
        assert (np.abs(got - want) <= 1e-5 + 1e-4 * np.abs(want)).all()


def test_raw_clip_pipeline_replays_reference_order():
    """BaseDataset.__getitem__ order per clip: extract_window's random.randint, then view 1, then view 2."""
    import audiossl_b200 as A
    unit = 8000
    lengths = [8000, 12000, 3000, 8001, 20000, 8000, 9500, 16000, 100, 11000, 8000, 30000]
    clips = _raw_clips(lengths, 9)
    np.random.seed(31); random.seed(31)
    oracle = ov.TwoViewOracle(8)
    want1, want2, want_start = [], [], []
    for c in clips:
        w, s = ofe.extract_window(c.numpy(), unit, rng=random)
        a, b = oracle(torch.from_numpy(ofe.log_mel(w)).unsqueeze(0))
        want1.append(a); want2.append(b); want_start.append(s)
    np.random.seed(31); random.seed(31)
    pipe = A.UpstreamInputPipeline(CFG, 8)
    assert pipe.unit_length == unit
    v1, v2 = pipe(clips)
    p = pipe.tfms.last_params
    arrs = ov.records_to_arrays(oracle.records)
    assert np.array_equal(p.start, np.array(want_start))
    assert np.array_equal(p.crop, arrs["crop"]) and np.array_equal(p.src_clip.astype(np.int32), arrs["src_clip"])
    for got, ref in ((v1, torch.stack(want1)), (v2, torch.stack(want2))):
        d = np.abs(got.cpu().numpy() - ref.numpy())
        assert (d <= 2e-5 + 1e-4 * np.abs(ref.numpy())).all(), d.max()


def test_empty_and_invalid():
    import audiossl_b200 as A
    out = A.extract_windows([], 100)
    assert tuple(out.shape) == (0, 100)
    with pytest.raises(ValueError):
        A.extract_windows([torch.zeros(2, 3)], 100)
    with pytest.raises(ValueError):
        A.UpstreamInputPipeline({"pretrain": {"normalization": "mean_var", "augmentations": {}}}, 4)([torch.zeros(10)])


def test_device_batch_loader_matches_reference_epoch():
    """One seeded epoch through DeviceBatchLoader == the reference's DataLoader(num_workers=0) order of clips and
    of random draws (window offset, then both views, clip by clip), views compared with the oracle."""
    import audiossl_b200 as A
    unit, bs = 8000, 4
    lengths = [8000, 12000, 3000, 8001, 20000, 8000, 9500, 16000, 100, 11000, 8000, 30000, 7000, 9000]
    clips = [c.numpy() for c in _raw_clips(lengths, 21)]
    torch.manual_seed(7); np.random.seed(31); random.seed(31)
    pipe = A.UpstreamInputPipeline(CFG, len(clips), max_batch=8)
    dm = A.BaselineDataModule(CFG, None, pipe, data_csv=clips, batch_size=bs, num_workers=2)
    dm.setup()
    assert dm.dataset_sizes["train"] == len(clips)
    got = [(a.cpu(), b.cpu()) for a, b in dm.train_dataloader()]
    assert len(got) == len(clips) // bs and tuple(got[0][0].shape) == (bs, 1, 64, 51)

    torch.manual_seed(7); np.random.seed(31); random.seed(31)
    order = A.DeviceBatchLoader(A.WaveformDataset(clips), None, batch_size=bs).batches()
    oracle = ov.TwoViewOracle(len(clips))
    for k, idx in enumerate(order):
        for r, i in enumerate(idx):
            w, _ = ofe.extract_window(clips[i], unit, rng=random)
            a, b = oracle(torch.from_numpy(ofe.log_mel(w)).unsqueeze(0))
            for have, ref in ((got[k][0][r], a), (got[k][1][r], b)):
                d = np.abs(have.numpy() - ref.numpy())
                assert (d <= 2e-5 + 1e-4 * np.abs(ref.numpy())).all(), (k, r, d.max())


def test_device_batch_loader_with_a_lookahead_pipeline_yields_the_same_epoch():
    """Raw clips through a lookahead pipeline (extract_window, log-mel, statistics of batch k; views of batch k-1,
    deferred launch; flush at the end of the epoch): the same batches, bit for bit, as the ordinary order."""
    import audiossl_b200 as A
    bs = 4
    lengths = [8000, 12000, 3000, 8001, 20000, 8000, 9500, 16000, 100, 11000, 8000, 30000, 7000, 9000, 8000, 8000, 12345]
    clips = [c.numpy() for c in _raw_clips(lengths, 33)]

    def epoch(lookahead):
        torch.manual_seed(7); np.random.seed(31); random.seed(31)
        pipe = A.UpstreamInputPipeline(CFG, len(clips), max_batch=8, lookahead=lookahead)
        dm = A.BaselineDataModule(CFG, None, pipe, data_csv=clips, batch_size=bs, num_workers=2)
        dm.setup()
        out = [(a.clone(), b.clone()) for a, b in dm.train_dataloader()]
        torch.cuda.synchronize()
        return out

    want, got = epoch(False), epoch(True)
    assert len(got) == len(want) == len(clips) // bs
    for (a1, a2), (b1, b2) in zip(got, want):
        assert torch.equal(a1, b1) and torch.equal(a2, b2)
