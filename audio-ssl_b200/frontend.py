"""Drop-in for the reference's log-mel front end, running on the GPU.

Mirrors ``MelSpectrogramLibrosa`` and ``extract_log_mel_spectrogram``
(/root/reference/src/utils/utils.py:20-28, 43-49): same names, same constructor arguments, same
result shapes (``[n_mels, T]`` per clip).  Additionally accepts a leading batch dimension
(``[B, L]`` -> ``[B, n_mels, T]``).  CPU tensors / numpy arrays are copied to the current CUDA
device; results are CUDA tensors.  There is no CPU implementation here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _native


def _stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def as_cuda_f32(x, device=None):
    if not torch.cuda.is_available():
        raise RuntimeError("audiossl_b200 needs a CUDA device: the CUDA kernels are the only implementation")
    if not isinstance(x, torch.Tensor):
        x = torch.as_tensor(np.asarray(x))
    if device is None:
        device = x.device if x.is_cuda else torch.device("cuda", torch.cuda.current_device())
    return x.to(device=device, dtype=torch.float32, non_blocking=True).contiguous()


class Plan:
    """Owns one ``assl_plan`` (window, twiddles, sparse filterbank on the current device)."""

    def __init__(self, fs=16000, n_fft=1024, shift=160, n_mels=64, fmin=60, fmax=7800):
        lib = _native.lib()
        self.cfg = _native.Cfg(int(fs), int(n_fft), int(shift), int(n_mels), float(fmin),
                               float(fmax) if fmax is not None else 0.0)
        self.handle = C.c_void_p()
        _native.check(lib.assl_plan_create(C.byref(self.handle), C.byref(self.cfg)), "assl_plan_create")
        self.device = torch.cuda.current_device()
        self.n_mels, self.hop, self.n_fft = int(n_mels), int(shift), int(n_fft)

    def num_frames(self, n_samples):
        t = _native.lib().assl_plan_num_frames(self.handle, int(n_samples))
        if t < 0:
            raise ValueError(f"clip of {n_samples} samples is too short: reflect padding needs more than n_fft/2")
        return int(t)

    def num_tiles(self, n_samples):
        return int(_native.lib().assl_plan_num_tiles(self.handle, int(n_samples)))

    def filterbank(self):
        out = np.empty((self.n_mels, self.n_fft // 2 + 1), np.float32)
        _native.check(_native.lib().assl_plan_filterbank(self.handle, out.ctypes.data_as(C.c_void_p)))
        return out

    def logmel_into(self, wave, out, slot0=0, cap=0, tile_sums=None, gain=None, power=False, time_major=False, pitch=0,
                    deferred=False):
        """wave [B, L] cuda fp32 -> out (ring or [B, n_mels, T]; [B, T, n_mels] with time_major); all on the
        current stream.  ``pitch``: floats per output row (0 = T; the engines' rings pad their rows to a multiple of 4).
        ``deferred``: assl_logmel_deferred_f32 -- the kernel reads ``wave`` / ``gain`` without waiting for the kernel
        launched before it (it waits before its first store): only when that kernel did not write them."""
        b, n = wave.shape
        name = "assl_logmel_deferred_f32" if deferred else "assl_logmel_pitched_f32"
        _native.check(getattr(_native.lib(), name)(
            self.handle, _ptr(wave), _ptr(gain), b, n, wave.stride(0), _ptr(out), int(slot0), int(cap), int(pitch),
            _ptr(tile_sums), (1 if power else 0) | (2 if time_major else 0), _stream_ptr()), name)

    def __del__(self):
        try:
            if self.handle:
                _native.lib().assl_plan_destroy(self.handle)
                self.handle = C.c_void_p()
        except Exception:
            pass


class HostPipeline:
    """``assl_pipeline``: the whole path with HOST buffers (pinned waveforms in, both views back in pinned host memory;
    uploads, kernels and downloads of consecutive calls overlap) -- what a DataLoader-side caller binds in place of
    BaseDataset.__getitem__'s per-clip CPU work (/root/reference/src/dataset/upstream_dataset.py:65,82,86).  Owns the
    copy streams and the call parity; the plan stays shareable.  ``exchange``: a ``StatsExchange`` over peer memory for a
    clip-sharded multi-GPU run (every call then scans all ranks' clips)."""

    def __init__(self, plan: Plan, exchange=None):
        self.plan, self.exchange = plan, exchange
        self.handle = C.c_void_p()
        ex = exchange.handle if exchange is not None else None
        _native.check(_native.lib().assl_pipeline_create(C.byref(self.handle), plan.handle, ex), "assl_pipeline_create")

    def workspace_bytes(self, batch, n_samples, cap):
        n = _native.lib().assl_pipeline_workspace_bytes(self.plan.handle, int(batch), int(n_samples), int(cap))
        if n < 0:
            raise ValueError("assl_pipeline_workspace_bytes: invalid batch / clip length / ring capacity")
        return int(n)

    def run(self, vcfg, ws, host_wave, batch, n_samples, slot0, cap, max_update, mix_slot, alpha, crop,
            host_view1=None, host_view2=None, stream=None):
        """One call.  ``ws`` device uint8 workspace; ``host_wave`` / ``host_view*`` pinned CPU tensors; ``mix_slot`` /
        ``alpha`` / ``crop`` contiguous numpy arrays that stay alive until ``sync()``."""
        hp = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)   # noqa: E731
        # int16 waveforms: 16-bit PCM as the audio file holds it (x / 32768 on the device, half the upload)
        name = "assl_pipeline_host_pcm16" if host_wave.dtype == torch.int16 else "assl_pipeline_host_f32"
        if name.endswith("f32") and host_wave.dtype != torch.float32:
            raise TypeError("host_wave must be float32 or int16 (PCM)")
        _native.check(getattr(_native.lib(), name)(
            self.handle, C.byref(vcfg), _ptr(ws), ws.numel(), hp(host_wave), int(batch), int(n_samples), int(slot0), int(cap),
            int(max_update), mix_slot.ctypes.data_as(C.c_void_p), alpha.ctypes.data_as(C.c_void_p),
            crop.ctypes.data_as(C.c_void_p), hp(host_view1), hp(host_view2), None, None,
            stream if stream is not None else _stream_ptr()), name)

    def sync(self):
        _native.check(_native.lib().assl_pipeline_host_sync(self.handle), "assl_pipeline_host_sync")

    def close(self):
        if self.handle:
            _native.lib().assl_pipeline_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class MelSpectrogramLibrosa:
    """Mel spectrogram with librosa 0.8.1 semantics (utils.py:20-28), computed by the fused kernel.

    ``__call__(audio)`` returns the mel POWER spectrogram like the reference; use
    :func:`extract_log_mel_spectrogram` (or :meth:`log_mel`) for the log-mel, which takes the log
    inside the same kernel.
    """

    def __init__(self, fs=16000, n_fft=1024, shift=160, n_mels=64, fmin=60, fmax=7800):
        self.fs, self.n_fft, self.shift, self.n_mels, self.fmin, self.fmax = fs, n_fft, shift, n_mels, fmin, fmax
        self.plan = Plan(fs, n_fft, shift, n_mels, fmin, fmax)

    @property
    def mfb(self):
        return self.plan.filterbank()

    def _run(self, audio, power, l2_normalize=False, time_major=False):
        wave = as_cuda_f32(audio)
        single = wave.dim() == 1
        if single:
            wave = wave.unsqueeze(0)
        if wave.dim() != 2:
            raise ValueError("expected a waveform [L] or a batch [B, L]")
        with torch.cuda.device(wave.device):
            t = self.plan.num_frames(wave.shape[1])
            shape = (wave.shape[0], t, self.n_mels) if time_major else (wave.shape[0], self.n_mels, t)
            out = torch.empty(shape, dtype=torch.float32, device=wave.device)
            gain = None
            if l2_normalize:
                gain = torch.empty(wave.shape[0], dtype=torch.float32, device=wave.device)
                _native.check(_native.lib().assl_wave_l2_gain_f32(
                    _ptr(wave), wave.shape[0], wave.shape[1], wave.stride(0), _ptr(gain), _stream_ptr()))
            self.plan.logmel_into(wave, out, gain=gain, power=power, time_major=time_major)
        return out[0] if single else out

    def __call__(self, audio):
        return self._run(audio, power=True)

    def log_mel(self, audio, l2_normalize=False, time_major=False):
        """Fused log-mel; ``l2_normalize`` folds ``F.normalize(waveform, p=2)`` (upstream_dataset.py:62-63) in;
        ``time_major`` writes ``[T, n_mels]`` per clip, the layout of the MAST variant
        (extras/mast_new/mast/moco_dataset.py:105-106), straight from the kernel."""
        return self._run(audio, power=False, l2_normalize=l2_normalize, time_major=time_major)


def extract_log_mel_spectrogram(waveform, to_mel_spec):
    """utils.py:43-49: ``(to_mel_spec(waveform) + eps).log()`` -- here one fused kernel."""
    if not isinstance(to_mel_spec, MelSpectrogramLibrosa):
        raise TypeError("to_mel_spec must be audiossl_b200.MelSpectrogramLibrosa")
    return to_mel_spec.log_mel(waveform)


def extract_windows(waves, unit_length, starts=None, with_gain=False, device=None):
    """Batched ``extract_window`` (utils.py:166-182) on the GPU.

    waves: a sequence of 1-D waveforms of any lengths (tensors or arrays).  Clips shorter than
    ``unit_length`` are centre zero-padded, longer ones cropped at ``starts[b]`` (the reference draws it
    with ``random.randint(0, len - unit_length)``; None = 0).  Returns ``[B, unit_length]`` fp32 on the
    GPU and, with ``with_gain``, also ``1 / max(||window||_2, 1e-12)`` per clip -- the reduction of
    ``F.normalize(p=2)`` (upstream_dataset.py:62-63) from the same pass, to be handed to the log-mel kernel.
    """
    if not torch.cuda.is_available():
        raise RuntimeError("audiossl_b200 needs a CUDA device: the CUDA kernels are the only implementation")
    unit_length = int(unit_length)
    ws = [w if isinstance(w, torch.Tensor) else torch.as_tensor(np.asarray(w)) for w in waves]
    if any(w.dim() != 1 for w in ws):
        raise ValueError("extract_windows takes 1-D waveforms")
    if device is None:
        device = next((w.device for w in ws if w.is_cuda), torch.device("cuda", torch.cuda.current_device()))
    lengths = np.array([w.shape[0] for w in ws], np.int64)
    with torch.cuda.device(device):
        raw = torch.cat([w.to(device=device, dtype=torch.float32, non_blocking=True) for w in ws]) if ws else \
            torch.empty(0, dtype=torch.float32, device=device)
    return extract_windows_packed(raw, lengths, unit_length, starts=starts, with_gain=with_gain)


def extract_windows_packed(raw, lengths, unit_length, starts=None, with_gain=False):
    """``extract_windows`` for clips already back to back in one CUDA buffer ``raw`` (clip b has ``lengths[b]``
    samples and starts where clip b-1 ends)."""
    unit_length = int(unit_length)
    lengths = np.asarray(lengths, np.int64)
    b = int(lengths.shape[0])
    offsets = np.concatenate(([0], np.cumsum(lengths)[:-1])).astype(np.int64) if b else np.zeros(0, np.int64)
    if b and int(lengths.sum()) > raw.numel():
        raise ValueError("clip lengths exceed the packed buffer")
    device = raw.device
    with torch.cuda.device(device):
        if raw.numel() == 0:
            raw = torch.zeros(1, dtype=torch.float32, device=device)
        meta = np.stack([offsets, lengths, np.zeros(b, np.int64) if starts is None else np.asarray(starts, np.int64)])
        meta_d = torch.from_numpy(meta).to(device)
        out = torch.empty((b, unit_length), dtype=torch.float32, device=device)
        gain = torch.empty(b, dtype=torch.float32, device=device) if with_gain else None
        _native.check(_native.lib().assl_extract_windows_f32(
            _ptr(raw), _ptr(meta_d[0]), _ptr(meta_d[1]), _ptr(meta_d[2]), b, unit_length, _ptr(out), unit_length,
            _ptr(gain), _stream_ptr()), "assl_extract_windows_f32")
    return (out, gain) if with_gain else out


def extract_window(wav, duration=16000, data_size=None, start=None):
    """utils.py:166-182 for one clip, same arguments; the crop offset comes from python's ``random`` like the
    reference's unless ``start`` is given."""
    import random
    unit_length = int(data_size * 16000) if data_size else int(duration)
    adj = int(wav.shape[0]) - unit_length
    if start is None:
        start = random.randint(0, adj) if adj > 0 else 0
    return extract_windows([wav], unit_length, starts=[start])[0]
