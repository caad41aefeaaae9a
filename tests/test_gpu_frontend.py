"""GPU parity: CUDA log-mel front end (through the C ABI) vs the CPU oracle.
Gate (BASELINE.json north_star): |a - b| <= 1e-5 + 1e-4 * |b| in fp32; frame counts bit-exact."""
import os

import numpy as np
import pytest
import torch

from oracle import frontend as ofe

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-4, 1e-5


def _close(got, want, rtol=RTOL, atol=ATOL):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    err = np.abs(got - want) - (atol + rtol * np.abs(want))
    assert err.max() <= 0, f"max violation {err.max():.3e} (max abs diff {np.abs(got - want).max():.3e})"


def families(n, length, seed):
    """SURVEY §8(d) synthetic waveform families, seeded on the CPU generator."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    t = torch.arange(length) / 16000.0
    out = []
    for k in range(n):
        kind = k % 6
        if kind == 0:
            w = torch.randn(length, generator=g) * 0.1
        elif kind == 1:      # pink-ish noise
            spec = torch.fft.rfft(torch.randn(length, generator=g))
            spec = spec / torch.sqrt(torch.arange(1, spec.numel() + 1, dtype=torch.float32))
            w = torch.fft.irfft(spec, n=length)
            w = 0.1 * w / w.std()
        elif kind == 2:      # chirp 50 Hz -> 7.9 kHz
            f = 50 + (7900 - 50) * t / (2 * t[-1])
            w = 0.5 * torch.sin(2 * np.pi * f * t) + 1e-3 * torch.randn(length, generator=g)
        elif kind == 3:      # two tones
            w = 0.3 * torch.sin(2 * np.pi * 440 * t) + 0.2 * torch.sin(2 * np.pi * 3000 * t) + 1e-4 * torch.randn(length, generator=g)
        elif kind == 4:
            w = torch.zeros(length)
            if k % 12 == 4:
                w[length // 2] = 1.0
        else:
            w = torch.randn(length, generator=g) * 0.1
            w = w / w.norm()
        out.append(w)
    return torch.stack(out).to(torch.float32)


@pytest.fixture(scope="module")
def mel():
    import audiossl_b200 as A
    return A.MelSpectrogramLibrosa()


def test_filterbank_matches_oracle(mel):
    fb = mel.mfb
    want = ofe.mel_filterbank()
    assert fb.shape == want.shape
    assert np.abs(fb - want).max() <= 1e-9 + 1e-6 * np.abs(want).max()
    assert np.array_equal(fb != 0, want != 0)


@pytest.mark.parametrize("length", [16000, 15200, 4000, 513, 16001, 1600])
def test_logmel_parity_families(mel, length):
    import audiossl_b200 as A
    waves = families(12, length, 1234 + length)
    got = A.extract_log_mel_spectrogram(waves.cuda(), mel).cpu().numpy()
    want = ofe.log_mel_batch(waves.numpy())
    assert got.shape == want.shape == (12, 64, 1 + length // 160)
    _close(got, want)


def test_single_clip_shape_and_power_output(mel):
    import audiossl_b200 as A
    w = families(1, 8000, 7)[0]
    lm = A.extract_log_mel_spectrogram(w, mel)            # CPU tensor in, like the reference call site
    assert lm.is_cuda and tuple(lm.shape) == (64, 51)
    _close(lm.cpu().numpy(), ofe.log_mel(w.numpy()))
    p = mel(w).cpu().numpy()                              # MelSpectrogramLibrosa.__call__ -> mel power
    want = ofe.mel_power(w.numpy())
    assert np.abs(p - want).max() <= 1e-5 * np.abs(want).max() + 1e-12


def test_long_clip_and_batch_1024(mel):
    import audiossl_b200 as A
    w = families(2, 160000, 99)
    _close(A.extract_log_mel_spectrogram(w.cuda(), mel).cpu().numpy(), ofe.log_mel_batch(w.numpy()))
    # full BASELINE config 2 size: spot-check clips against the oracle + batch-position independence
    g = torch.Generator(device="cpu").manual_seed(1235)
    big = (torch.randn(1024, 16000, generator=g) * 0.1).to(torch.float32)
    out = A.extract_log_mel_spectrogram(big.cuda(), mel)
    assert tuple(out.shape) == (1024, 64, 101) and bool(torch.isfinite(out).all())
    for b in (0, 511, 1023):
        _close(out[b].cpu().numpy(), ofe.log_mel(big[b].numpy()))
    again = A.extract_log_mel_spectrogram(big[500:524].cuda(), mel)
    assert torch.equal(again, out[500:524])


def test_l2_normalised_input(mel):
    w = families(6, 15200, 5)
    got = mel.log_mel(w.cuda(), l2_normalize=True).cpu().numpy()
    want = np.stack([ofe.log_mel(ofe.l2_normalize(x)) for x in w.numpy()])
    _close(got, want)


def test_linearity_property_full_size(mel):
    """Scaling the waveform by 2 scales mel power by 4 exactly (power-of-two scaling is exact in fp32)."""
    g = torch.Generator(device="cpu").manual_seed(3)
    w = (torch.randn(64, 16000, generator=g) * 0.1).to(torch.float32).cuda()
    p1, p2 = mel(w), mel(2 * w)
    assert torch.allclose(p2, 4 * p1, rtol=1e-6, atol=1e-12)


def test_too_short_clip_is_rejected(mel):
    with pytest.raises(ValueError):
        mel(torch.zeros(512))


def test_other_plan_parameters():
    import audiossl_b200 as A
    m = A.MelSpectrogramLibrosa(n_mels=128, fmin=0, fmax=8000, shift=256)
    w = families(3, 12000, 21)
    want = np.stack([ofe.log_mel(x, mfb=ofe.mel_filterbank(n_mels=128, fmin=0, fmax=8000), hop=256) for x in w.numpy()])
    _close(A.extract_log_mel_spectrogram(w.cuda(), m).cpu().numpy(), want)


def test_mast_variant_128_mels_time_major():
    """extras/mast_new/mast/utils.py:251-259 (n_mels=128) and the [1, T, F] layout of moco_dataset.py:105-106."""
    import audiossl_b200 as A
    m = A.MelSpectrogramLibrosa(n_mels=128)
    w = families(6, 16000, 77)
    mfb = ofe.mel_filterbank(n_mels=128)
    want = np.stack([ofe.log_mel(x, mfb=mfb) for x in w.numpy()])
    got_t = m.log_mel(w.cuda(), time_major=True)
    assert tuple(got_t.shape) == (6, 101, 128)
    _close(got_t.permute(0, 2, 1).cpu().numpy(), want)
    assert torch.equal(got_t.permute(0, 2, 1), m.log_mel(w.cuda()))      # same values, only the layout differs


def test_unaligned_and_strided_waveforms(mel):
    """Bases that are not 16-byte aligned (no bulk copy: the per-warp patch loop stages the whole span) and rows with
    a stride (ld > L) give the same bits as the aligned, packed batch."""
    import audiossl_b200 as A
    w = families(9, 8000, 41)
    ref = A.extract_log_mel_spectrogram(w.cuda(), mel)
    flat = torch.zeros(9 * 8000 + 1, device="cuda")
    flat[1:] = w.cuda().reshape(-1)
    shifted = flat[1:].view(9, 8000)                          # base + 4 bytes
    assert shifted.data_ptr() % 16 != 0
    assert torch.equal(A.extract_log_mel_spectrogram(shifted, mel), ref)
    wide = torch.zeros(9, 8004, device="cuda")
    wide[:, :8000] = w.cuda()
    out = torch.empty_like(ref)
    mel.plan.logmel_into(wide[:, :8000], out)                  # ld = 8004
    assert torch.equal(out, ref)


def test_filterbank_parts_are_dealt_without_bank_conflicts():
    """The plan starts every interval's run where its 16-byte bank group is free and colours the interval totals: no
    quarter-warp of the log-mel kernel's filterbank pass reads a bank group twice (csrc/plan.cu; what the kernel's
    shared-memory pipe, its nearest limit, pays for otherwise).  Reported by the plan itself under ASSL_PLAN_DEBUG."""
    import re
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, os.path.join(root, "tools", "plan_debug.py")], capture_output=True, text=True,
                         timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [l for l in res.stderr.splitlines() if "assl plan:" in l]
    assert len(lines) >= 3, res.stderr[-2000:]
    first = lines[0]                                              # the reference's configuration: 64 mels, 60-7800 Hz
    assert first.startswith("n_mels 64 fmin 60.0 fmax 7800.0") and "96 parts in 3 slots" in first, first
    for l in lines:
        m = re.search(r"bank group twice: (\d+), with a total's colour twice: (\d+)", l)
        assert m and int(m.group(1)) == 0 and int(m.group(2)) == 0, l
