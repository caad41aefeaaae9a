"""``oracle.frontend`` (librosa 0.8.1 restatement): cross-checks and analytic known answers.
librosa itself is absent, so these stand in for the missing golden vectors."""
import os

import numpy as np
import pytest

from oracle import frontend

TOL = dict(rtol=1e-4, atol=1e-5)


def test_filterbank_structure():
    fb = frontend.mel_filterbank()
    assert fb.shape == (64, 513) and fb.dtype == np.float32
    assert int((fb != 0).sum()) == 966
    nz = np.nonzero(fb.any(axis=0))[0]
    assert nz.min() == 4 and nz.max() == 499
    taps = (fb != 0).sum(1)
    assert taps.min() == 5 and taps.max() == 45
    # every bin feeds at most two adjacent mel bands
    for k in range(513):
        m = np.nonzero(fb[:, k])[0]
        assert len(m) <= 2 and (len(m) < 2 or m[1] == m[0] + 1)


def test_filterbank_vs_torchaudio_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "frontend_crosscheck.npz"))
    assert np.abs(frontend.mel_filterbank() - g["fbank_torchaudio"]).max() < 1e-7


def test_logmel_vs_torchaudio_fixture(golden_dir):
    g = np.load(os.path.join(golden_dir, "frontend_crosscheck.npz"))
    lo = frontend.log_mel_batch(g["waves"])
    assert np.array_equal(lo, g["logmel_oracle"])                        # regression
    # torchaudio runs the FFT in fp32; the documented gate still holds
    np.testing.assert_allclose(lo, g["logmel_torchaudio"], rtol=1e-4, atol=1.5e-4)
    assert np.abs(lo[0] - g["logmel_torchaudio"][0]).max() < 2e-5          # noise clip: tight


def test_frame_count_and_reflect_padding():
    for n in (513, 1600, 15200, 16000, 16001):
        y = np.random.default_rng(n).standard_normal(n).astype(np.float32)
        assert frontend.stft(y).shape == (513, 1 + n // 160)
    with pytest.raises(ValueError):
        frontend.stft(np.zeros(512, np.float32))
    # frame 0, tap n reads y[|n-512|]
    y = np.arange(2000, dtype=np.float32)
    spec = frontend.stft(y)
    frame0 = np.abs(np.arange(1024) - 512).astype(np.float64)
    want = np.fft.rfft(frontend.hann_periodic(1024) * frame0).astype(np.complex64)
    assert np.array_equal(spec[:, 0], want)


def test_all_zero_waveform_known_answer():
    fb = frontend.mel_filterbank()
    lm = frontend.log_mel(np.zeros(4000, np.float32))
    want = np.log(fb.astype(np.float32) @ np.full(513, np.float32(2.220446e-16), np.float32) + np.float32(1.1920929e-07))
    np.testing.assert_allclose(lm, np.broadcast_to(want[:, None], lm.shape), **TOL)


def test_unit_impulse_is_flat_spectrum():
    y = np.zeros(4000, np.float32)
    y[160 * 10] = 1.0                        # centre of frame 10 -> window tap 512 -> w = 1
    spec = frontend.stft(y)
    np.testing.assert_allclose(np.abs(spec[:, 10]) ** 2, 1.0, rtol=1e-6)


def test_bin_centred_cosine_main_lobe():
    n = np.arange(8000)
    k0 = 100
    y = np.cos(2 * np.pi * k0 * n / 1024).astype(np.float32)
    p = np.abs(frontend.stft(y)[:, 20]) ** 2
    np.testing.assert_allclose(p[k0], (1024 / 4) ** 2, rtol=1e-4)
    np.testing.assert_allclose(p[[k0 - 1, k0 + 1]], (1024 / 8) ** 2, rtol=1e-4)
    assert p[k0 + 3] < 1e-3


def test_extract_window_and_l2():
    y = np.arange(100, dtype=np.float32)
    w, s = frontend.extract_window(y, 160)
    assert s == 0 and w.shape == (160,) and w[30] == 0 and w[29] == 0 and w[31] == 1
    w, s = frontend.extract_window(y, 40, start=7)
    assert np.array_equal(w, y[7:47])
    z = frontend.l2_normalize(y)
    np.testing.assert_allclose(np.linalg.norm(z), 1.0, rtol=1e-6)


def test_against_transformers_audio_utils():
    """Second independent cross-check (librosa itself is not installable here): transformers.audio_utils is a numpy
    implementation written to reproduce librosa's filterbank and STFT (Slaney scale + area norm, centred reflect
    padding, periodic Hann, fp64 FFT).  Filterbank to 1e-8, log-mel of a noise clip and of a chirp to 2e-6."""
    au = pytest.importorskip("transformers.audio_utils")
    fb = au.mel_filter_bank(num_frequency_bins=513, num_mel_filters=64, min_frequency=60, max_frequency=7800,
                            sampling_rate=16000, norm="slaney", mel_scale="slaney")
    assert np.abs(fb.T - frontend.mel_filterbank()).max() < 1e-8
    win = au.window_function(1024, "hann", periodic=True)
    assert np.abs(win - frontend.hann_periodic(1024)).max() < 1e-15
    rng = np.random.default_rng(7)
    t = np.arange(15200) / 16000.0
    clips = [(rng.standard_normal(16000) * 0.1).astype(np.float32),
             (0.5 * np.sin(2 * np.pi * (50.0 + 4000.0 * t) * t) + 1e-3 * rng.standard_normal(15200)).astype(np.float32)]
    for y in clips:
        power = au.spectrogram(y.astype(np.float64), win, frame_length=1024, hop_length=160, fft_length=1024, power=2.0,
                               center=True, pad_mode="reflect", onesided=True, mel_filters=fb, mel_floor=1e-30)
        theirs = np.log(power + 1.1920928955078125e-07)
        ours = frontend.log_mel(y)
        assert ours.shape == theirs.shape == (64, frontend.num_frames(len(y)))
        assert np.abs(ours - theirs).max() < 2e-6
