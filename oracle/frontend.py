"""Oracle: log-mel front end (numpy).  TEST INFRASTRUCTURE -- never imported by the product.

Restates what the reference computes in
  * ``MelSpectrogramLibrosa.__init__``   /root/reference/src/utils/utils.py:22-24
  * ``MelSpectrogramLibrosa.__call__``   /root/reference/src/utils/utils.py:26-28
  * ``extract_log_mel_spectrogram``      /root/reference/src/utils/utils.py:43-49

The arithmetic itself lives in librosa==0.8.1 (/root/reference/requirements.txt:71), an
un-vendored dependency that is absent from this image.  The functions below restate the
published librosa 0.8.1 algorithm for the two calls the reference makes:

  ``librosa.filters.mel(sr, n_fft, n_mels, fmin, fmax)``  -- Slaney mel scale, Slaney area
      normalisation, float32 weights;
  ``librosa.stft(y, n_fft, hop_length)``                  -- center=True, reflect padding,
      periodic Hann of length n_fft (scipy ``get_window('hann', fftbins=True)``), float64
      window * float32 frames, float64 rFFT, result stored as complex64.

Parity against librosa itself is UNPINNED (it cannot be imported here); the restatement is
cross-checked against torchaudio's Slaney filterbank / STFT and analytic known answers in
``tests/test_oracle_frontend.py``.
"""
from __future__ import annotations

import numpy as np

try:  # scipy is what librosa 0.8.1 calls for the window
    from scipy.signal import get_window as _get_window
except Exception:  # pragma: no cover
    _get_window = None

POWER_EPS = np.float32(np.finfo(float).eps)          # utils.py:28  (+ np.finfo(float).eps, kept fp32)
LOG_EPS = np.float32(np.finfo(np.float32).eps)       # utils.py:48  (+ torch.finfo().eps)


# --------------------------------------------------------------------------- mel scale
def hz_to_mel(freq):
    """librosa.convert.hz_to_mel(htk=False): linear below 1 kHz, log above."""
    freq = np.asanyarray(freq, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = freq / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    if freq.ndim:
        hi = freq >= min_log_hz
        mels = mels.copy()
        mels[hi] = min_log_mel + np.log(freq[hi] / min_log_hz) / logstep
    elif freq >= min_log_hz:
        mels = min_log_mel + np.log(freq / min_log_hz) / logstep
    return mels


def mel_to_hz(mels):
    """librosa.convert.mel_to_hz(htk=False)."""
    mels = np.asanyarray(mels, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * mels
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    if mels.ndim:
        hi = mels >= min_log_mel
        freqs = freqs.copy()
        freqs[hi] = min_log_hz * np.exp(logstep * (mels[hi] - min_log_mel))
    elif mels >= min_log_mel:
        freqs = min_log_hz * np.exp(logstep * (mels - min_log_mel))
    return freqs


def mel_filterbank(sr=16000, n_fft=1024, n_mels=64, fmin=60, fmax=7800):
    """``librosa.filters.mel`` as called at utils.py:24 -> float32 [n_mels, 1 + n_fft//2]."""
    if fmax is None:
        fmax = float(sr) / 2
    n_bins = 1 + n_fft // 2
    weights = np.zeros((n_mels, n_bins), dtype=np.float32)
    fftfreqs = np.linspace(0, float(sr) / 2, n_bins, endpoint=True)
    mel_pts = np.linspace(hz_to_mel(fmin), hz_to_mel(fmax), n_mels + 2)
    mel_f = mel_to_hz(mel_pts)
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))      # fp64 -> fp32 on assignment
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    # in-place fp32 *= fp64: computed in fp64, rounded back to fp32
    weights = (weights.astype(np.float64) * enorm[:, np.newaxis]).astype(np.float32)
    return weights


# --------------------------------------------------------------------------- STFT
def hann_periodic(n_fft):
    """scipy.signal.get_window('hann', n_fft, fftbins=True) (float64)."""
    if _get_window is not None:
        return _get_window("hann", n_fft, fftbins=True)
    n = np.arange(n_fft, dtype=np.float64)
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * n / n_fft)


def num_frames(n_samples, hop=160):
    """Frames librosa.stft(center=True) yields: 1 + L // hop."""
    return 1 + n_samples // hop


def stft(y, n_fft=1024, hop=160):
    """``librosa.stft(y, n_fft, hop_length)`` -> complex64 [1 + n_fft//2, T]."""
    y = np.asarray(y, dtype=np.float32)
    if y.ndim != 1:
        raise ValueError("oracle stft takes one clip [L]")
    if y.shape[0] <= n_fft // 2:
        raise ValueError("reflect padding needs L > n_fft/2")
    win = hann_periodic(n_fft).reshape(-1, 1)                       # float64 [n_fft, 1]
    y_pad = np.pad(y, n_fft // 2, mode="reflect")
    t = 1 + (y_pad.shape[0] - n_fft) // hop
    idx = np.arange(n_fft)[:, None] + hop * np.arange(t)[None, :]
    frames = y_pad[idx]                                             # float32 [n_fft, T]
    spec = np.fft.rfft(win * frames, axis=0)                        # float64 math
    return spec.astype(np.complex64)


def mel_power(y, mfb=None, n_fft=1024, hop=160):
    """``MelSpectrogramLibrosa.__call__`` (utils.py:26-28) -> float32 [n_mels, T]."""
    if mfb is None:
        mfb = mel_filterbank(n_fft=n_fft)
    spec = stft(y, n_fft=n_fft, hop=hop)
    # numpy 1.20 (requirements.txt:94) keeps abs(X)**2 + python-float eps in float32;
    # numpy >= 2 would promote, hence the explicit float32 epsilon.
    power = (np.abs(spec) ** 2).astype(np.float32) + POWER_EPS
    return np.matmul(mfb, power).astype(np.float32)


def log_mel(y, mfb=None, n_fft=1024, hop=160):
    """``extract_log_mel_spectrogram`` (utils.py:43-49) -> float32 [n_mels, T]."""
    m = mel_power(y, mfb=mfb, n_fft=n_fft, hop=hop)
    return np.log(m + LOG_EPS).astype(np.float32)


def log_mel_batch(waves, mfb=None, n_fft=1024, hop=160):
    """Apply :func:`log_mel` clip by clip (the reference is per-clip) -> [B, n_mels, T]."""
    waves = np.asarray(waves, dtype=np.float32)
    if mfb is None:
        mfb = mel_filterbank(n_fft=n_fft)
    return np.stack([log_mel(w, mfb=mfb, n_fft=n_fft, hop=hop) for w in waves])


# --------------------------------------------------------------------------- waveform prep
def extract_window(wav, unit_length, start=None, rng=None):
    """``extract_window`` (utils.py:166-182): centre zero-pad to ``unit_length`` then crop.

    ``start`` is the crop offset the reference draws with ``random.randint(0, len-unit)``
    (python ``random``); pass it explicitly or pass ``rng`` (a ``random.Random``/module).
    Returns (window, start).
    """
    wav = np.asarray(wav, dtype=np.float32)
    adj = unit_length - wav.shape[0]
    if adj > 0:
        half = adj // 2
        wav = np.pad(wav, (half, adj - half))
    adj = wav.shape[0] - unit_length
    if start is None:
        start = rng.randint(0, adj) if (adj > 0 and rng is not None) else 0
    return wav[start:start + unit_length], start


def l2_normalize(wav, eps=1e-12):
    """``F.normalize(waveform, dim=-1, p=2)`` (upstream_dataset.py:62-63), float32."""
    wav = np.asarray(wav, dtype=np.float32)
    nrm = np.sqrt(np.sum(wav.astype(np.float32) ** 2, dtype=np.float32))
    return (wav / np.maximum(nrm, np.float32(eps))).astype(np.float32)
