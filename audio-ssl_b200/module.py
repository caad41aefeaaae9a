"""``AugmentationModule`` drop-in (/root/reference/src/augmentations/__init__.py:5-35) and the fused
waveform -> two-views pipeline built on the same engine.

The reference composes ``RunningNorm`` (once per clip) with ``Sequential(MixupBYOLA, RandomResizeCrop)``
applied twice on ONE shared MixupBYOLA instance; here that is one statistics scan plus one fused kernel
launch per batch.  A ``Kmix`` entry with a ``centroid_path`` appends :class:`augmentations.Kmix` after the crop
like the reference's ``nn.Sequential`` (src/augmentations/__init__.py:24-25); it then sees the views in the
reference's order (view 1 of clip 0, view 2 of clip 0, view 1 of clip 1, ...) but draws after the batch's
MixupBYOLA / RandomResizeCrop draws instead of interleaved with them (its draws are data dependent, see
DESIGN.md 4.4).  ``Kmix`` with ``centroid_path: None`` (the shipped configs) and ``PatchDrop`` are ignored with a
warning: in the reference they cannot run (``torch.load(None)``, augmentations.py:136; ``PatchDrop.forward`` uses
an undefined name, :72-75).
"""
from __future__ import annotations

import warnings

import numpy as np
import torch

from .augmentations import _as_batch, _chunks
from .datamodule import PackedClips
from .engine import ViewEngine
from .frontend import Plan, as_cuda_f32
from .sampler import FastSampler, ReferenceSampler, ViewParams


class AugmentationModule:
    """The Augmentation Module: ``tfms(x) -> (view1, view2)``.

    config / len_of_files: exactly the reference's arguments (train_upstream.py:31).
    Extras (keyword only): ``interpolation`` ('bicubic' as the reference, or 'bilinear'), ``sampler``
    ('reference' replays np.random / random like the reference, 'fast' is a vectorised generator, or
    any object with ``draw(clips_done, batch, F, T) -> ViewParams``), ``max_batch`` (ring head room).
    """

    def __init__(self, config, len_of_files, *, interpolation="bicubic", sampler="reference",
                 max_batch=4096, seed=0, sync_stats=True, global_bank=False, lookahead=False):
        pre = config["pretrain"]
        aug = pre["augmentations"]
        self.kmix = None
        km = aug.get("Kmix")
        if km is not None and km.get("centroid_path") is not None:
            from .augmentations import Kmix
            self.kmix = Kmix(ratio=km["ratio"], log_mixup_exp=km["log_mixup_exp"], top_k=km["top_k"],
                             centroid_path=km["centroid_path"], max_batch=2 * max_batch)
        elif km is not None:
            warnings.warn("Kmix without a centroid_path is ignored (the reference fails in torch.load there)", stacklevel=2)
        if "PatchDrop" in aug:
            warnings.warn("PatchDrop is ignored (PatchDrop.forward raises NameError in the reference)", stacklevel=2)
        mix = aug.get("MixupBYOLA")
        rrc = aug.get("RandomResizeCrop")
        self.has_mix, self.has_crop = mix is not None, rrc is not None
        ratio = mix["ratio"] if mix else 0.0
        log_mixup = mix["log_mixup_exp"] if mix else True
        vscale = tuple(rrc["virtual_crop_scale"]) if rrc else (1.0, 1.0)
        fscale = tuple(rrc["freq_crop_scale"]) if rrc else (1.0, 1.0)
        tscale = tuple(rrc["time_crop_scale"]) if rrc else (1.0, 1.0)
        # __init__.py:13-14 -- RunningNorm(epoch_samples=2*len_of_files); any other setting means no pre-norm
        self.pre_norm = pre.get("normalization") == "mean_var"
        max_update = 2 * len_of_files * 10 if self.pre_norm else None
        self.engine = ViewEngine(n_memory=2048 if mix else 0, pushes_per_clip=2, ratio=ratio, log_mixup=log_mixup,
                                 virtual_crop_scale=vscale, interp=interpolation, normalize=self.pre_norm,
                                 max_update=max_update, max_batch=max_batch, sync_stats=sync_stats,
                                 global_bank=global_bank and mix is not None, lookahead=lookahead)
        kw = dict(ratio=ratio, n_memory=2048, freq_scale=fscale, time_scale=tscale, virtual_crop_scale=vscale,
                  pushes_per_clip=2, mixup=self.has_mix, crop=self.has_crop)
        if sampler == "reference":
            self.sampler = ReferenceSampler(**kw)
        elif sampler == "fast":
            self.sampler = FastSampler(seed=seed, **kw)
        else:
            self.sampler = sampler
        self.last_params = None

    # -- spectrogram in, two views out (what BaseDataset.__getitem__ calls, upstream_dataset.py:86)
    def __call__(self, x, params: ViewParams = None):
        xb, single = _as_batch(x)
        eng = self.engine
        v1s, v2s = [], []
        with torch.cuda.device(xb.device):
            eng.ensure(xb.shape[1], xb.shape[2], xb.device)
            for lo, hi in _chunks(xb.shape[0], eng.max_batch):
                part = xb[lo:hi].contiguous()
                eng.ring_put(part)
                if self.pre_norm:
                    eng.update_stats(eng.moments(part), 1)
                p = self._params(params, lo, hi)
                a, b = eng.launch_views(p, hi - lo, xb.device)
                v1s.append(a); v2s.append(b)
        v1 = v1s[0] if len(v1s) == 1 else torch.cat(v1s)
        v2 = v2s[0] if len(v2s) == 1 else torch.cat(v2s)
        v1, v2 = self._post(v1, v2)
        return (v1[:, 0], v2[:, 0]) if single else (v1, v2)

    def _post(self, v1, v2):
        """Augmentations the reference chains after the crop (Kmix): views in the reference's order v1[0], v2[0], v1[1], ..."""
        if self.kmix is None:
            return v1, v2
        both = self.kmix(torch.stack((v1, v2), 1).reshape(-1, *v1.shape[1:]))
        both = both.reshape(v1.shape[0], 2, *v1.shape[1:])
        return both[:, 0], both[:, 1]

    def _params(self, params, lo, hi, g0=None):
        eng = self.engine
        g0 = eng.clips_done if g0 is None else g0
        if params is None and eng.global_bank:
            # the FIFO runs over the global clip order: every rank draws the parameters of the WHOLE global batch
            # from the same generator state (same seed on every rank) and keeps its slice; src_clip = global ids
            n = hi - lo
            everyone = self.sampler.draw(eng.clips_done * eng.world, n * eng.world, eng.f, eng.t)
            a, b = eng.rank * n, (eng.rank + 1) * n
            p = ViewParams(everyone.alpha[a:b], everyone.src_clip[a:b], everyone.bank_index[a:b], everyone.crop[a:b])
        elif params is None:
            p = self.sampler.draw(g0, hi - lo, self.engine.f, self.engine.t)
        else:
            p = ViewParams(params.alpha[lo:hi], params.src_clip[lo:hi], params.bank_index[lo:hi], params.crop[lo:hi])
        self.last_params = p
        return p

    def state_dict(self):
        sd = self.engine.state_dict()
        if self.kmix is not None:
            sd["kmix"] = self.kmix.state_dict()
        return sd

    def load_state_dict(self, sd):
        self.engine.load_state_dict(sd)
        if self.kmix is not None and "kmix" in sd:
            self.kmix.load_state_dict(sd["kmix"])


class UpstreamInputPipeline:
    """Waveforms in, two augmented views out, everything on the GPU:

        [extract_window + L2 gain, one pass, for raw clips of any length]
        -> log-mel (fused kernel, written straight into the bank ring, per-tile sums on the side)
        -> running-statistics scan -> two-view kernel.

    This is the batched equivalent of ``BaseDataset.__getitem__`` lines :60-65, :82, :86
    (/root/reference/src/dataset/upstream_dataset.py) for a whole batch.  Call it with a ``[B, L]`` batch
    of equal-length windows, or with a list of raw 1-D clips of different lengths: those go through
    ``extract_window`` first (``config["pretrain"]["input"]["length_wave"]`` seconds, or ``unit_length``
    samples), their crop offsets drawn by the sampler in the reference's order.

    ``inputs_ready=True`` (with ``lookahead=True``, ``[B, L]`` batches): the caller promises that the waveforms of a call
    were complete on the device BEFORE the kernel launched last on the stream started (a resident dataset; an upload
    guarded by an event the stream has waited on).  The log-mel kernel is then launched with assl_logmel_deferred_f32:
    it reads and transforms its first frame pairs under the tail of the views launched before it and waits for them only
    before its first store.  Same results.
    """

    def __init__(self, config, len_of_files, *, to_mel_spec=None, unit_length=None, inputs_ready=False, **kw):
        self.inputs_ready = bool(inputs_ready)
        self.tfms = AugmentationModule(config, len_of_files, **kw)
        self.plan = to_mel_spec.plan if to_mel_spec is not None else Plan()
        self.l2 = config["pretrain"].get("normalization") == "l2"
        length_wave = config["pretrain"].get("input", {}).get("length_wave")
        self.unit_length = int(unit_length) if unit_length is not None else \
            (int(length_wave * 16000) if length_wave else None)          # utils.py:168-169
        self.last_lms = None
        self._queue = []                     # lookahead: batches whose front end has run, waiting for their views

    def __call__(self, waveforms, params: ViewParams = None):
        if isinstance(waveforms, PackedClips):
            return self._ragged(waveforms, params)
        if isinstance(waveforms, (list, tuple)):
            return self._ragged(list(waveforms), params)
        if self.tfms.engine.lookahead:
            return self._call_lookahead(waveforms, params)
        wave = as_cuda_f32(waveforms)
        if wave.dim() == 1:
            wave = wave.unsqueeze(0)
        eng, tf = self.tfms.engine, self.tfms
        from . import _native
        from .frontend import _ptr, _stream_ptr
        v1s, v2s = [], []
        with torch.cuda.device(wave.device):
            eng.ensure(self.plan.n_mels, self.plan.num_frames(wave.shape[1]), wave.device)
            for lo, hi in _chunks(wave.shape[0], eng.max_batch):
                part = wave[lo:hi]
                gain = None
                if self.l2:
                    gain = torch.empty(hi - lo, dtype=torch.float32, device=wave.device)
                    _native.check(_native.lib().assl_wave_l2_gain_f32(_ptr(part), hi - lo, part.shape[1], part.stride(0),
                                                                      _ptr(gain), _stream_ptr()))
                a, c = self._windows_to_views(part, gain, lambda lo=lo, hi=hi: tf._params(params, lo, hi))
                v1s.append(a); v2s.append(c)
        v1 = v1s[0] if len(v1s) == 1 else torch.cat(v1s)
        v2 = v2s[0] if len(v2s) == 1 else torch.cat(v2s)
        return self.tfms._post(v1, v2)

    # -- lookahead=True: software-pipelined step order  log-mel(k) . statistics(k) . views(k-1)
    #    (across GPUs: log-mel(k) . scan(k-1)+send(k) in one launch . views(k-2), see ViewEngine.split_stats)
    def _call_lookahead(self, waveforms, params):
        """Run the front end and the statistics of THIS batch and return the two views of the PREVIOUS one (``None`` on
        the first call; ``flush()`` returns the last batch's).  The views of batch k-1 are launched right behind the
        statistics kernel of batch k with assl_two_views_deferred_f32: they fill the SMs the log-mel kernel's tail
        leaves idle and run beside the statistics kernel (and, across GPUs, beside its exchange).  Same views as the
        ordinary order, one call later; the draws happen in the same clip order."""
        wave = as_cuda_f32(waveforms)
        if wave.dim() == 1:
            wave = wave.unsqueeze(0)
        eng, tf = self.tfms.engine, self.tfms
        from . import _native
        from .frontend import _ptr, _stream_ptr
        b = wave.shape[0]
        if b > eng.max_batch:
            raise ValueError(f"lookahead takes at most max_batch = {eng.max_batch} clips per call")
        with torch.cuda.device(wave.device):
            eng.ensure(self.plan.n_mels, self.plan.num_frames(wave.shape[1]), wave.device)
            gain = None
            if self.l2:
                gain = torch.empty(b, dtype=torch.float32, device=wave.device)
                _native.check(_native.lib().assl_wave_l2_gain_f32(_ptr(wave), b, wave.shape[1], wave.stride(0), _ptr(gain),
                                                                  _stream_ptr()))
            # deferred only when no kernel of this call wrote what the log-mel kernel reads (no gain pass, no conversion copy)
            ready = self.inputs_ready and gain is None and wave.data_ptr() == getattr(waveforms, "data_ptr", lambda: 0)()
            return self._front_then_previous_views(wave, gain, tf._params(params, 0, b, g0=eng.head), deferred=ready)

    def _front_then_previous_views(self, part, gain, p, deferred=False):
        """[b, L] windows of THIS batch -> log-mel into the ring -> statistics; then the views of the oldest pending batch
        (one batch back; two across GPUs, see ViewEngine.split_stats)."""
        eng, tf = self.tfms.engine, self.tfms
        b = part.shape[0]
        # parameters of this batch: uploaded before its kernels, used one or two calls later
        prep = eng.prepare_views(p, b, part.device, eng.head)
        tiles = self.plan.num_tiles(part.shape[1])
        sums = torch.empty((b, tiles, 2), dtype=torch.float64, device=part.device) if tf.pre_norm else None
        split = tf.pre_norm and eng.split_stats
        self.plan.logmel_into(part, eng.ring, slot0=eng.slot0, cap=eng.cap, tile_sums=sums, gain=gain, pitch=eng.pitch,
                              deferred=deferred)
        if split:
            # sharded run: the statistics of the batch before are scanned now -- every rank sent its sums a step ago, so
            # nobody waits for the slowest rank -- then this batch's sums go out; both run beside the deferred views below
            # (one launch: the scan may start under the log-mel kernel's tail -- it reads records and state only -- and the
            #  same cluster sends once that kernel has completed)
            eng.scan_then_send_stats(sums, tiles)
        elif tf.pre_norm:
            eng.update_stats(sums, tiles)
        eng.ahead += b
        out = None
        if len(self._queue) >= eng.depth:
            # (deferred: behind the statistics kernels -- or, without pre-norm, the log-mel kernel -- of THIS step, which
            #  the launch may overlap; everything it reads was complete before this step's log-mel kernel started)
            oldest = self._queue.pop(0)
            v1, v2 = eng.launch_prepared(oldest, deferred=True)
            eng.ahead -= oldest["batch"]
            out = tf._post(v1, v2)
        self._queue.append(prep)
        return out

    def flush(self):
        """Views of the oldest batch still waiting in a ``lookahead`` pipeline (``None`` when nothing is pending); call it
        until it returns None (one batch is pending on one GPU, up to two across GPUs) -- or use ``drain()``."""
        if not self._queue:
            return None
        eng = self.tfms.engine
        oldest = self._queue.pop(0)
        with torch.cuda.device(oldest["device"]):
            if not self._queue:
                eng.scan_pending_stats()                        # (sharded run: the last batch's statistics)
            v1, v2 = eng.launch_prepared(oldest, deferred=False)
        eng.ahead -= oldest["batch"]
        return self.tfms._post(v1, v2)

    def drain(self):
        """Every batch still pending, oldest first: a list of (view1, view2)."""
        out = []
        while self._queue:
            out.append(self.flush())
        return out

    def _windows_to_views(self, part, gain, draw):
        """[b, L] windows on the GPU -> log-mel into the ring -> statistics -> two views."""
        eng, tf = self.tfms.engine, self.tfms
        b = part.shape[0]
        tiles = self.plan.num_tiles(part.shape[1])
        sums = torch.empty((b, tiles, 2), dtype=torch.float64, device=part.device) if tf.pre_norm else None
        self.plan.logmel_into(part, eng.ring, slot0=eng.slot0, cap=eng.cap, tile_sums=sums, gain=gain, pitch=eng.pitch)
        if tf.pre_norm:
            eng.update_stats(sums, tiles)
        return eng.launch_views(draw(), b, part.device)

    def _ragged(self, clips, params):
        """Raw clips of any length: extract_window (+ L2 gain in the same pass) feeds the front end."""
        from .frontend import extract_windows, extract_windows_packed
        if self.tfms.engine.global_bank:
            raise ValueError("global_bank takes equal-length windows [B, L] (the window offsets of the other ranks' clips "
                             "would have to be drawn here too)")
        if self.unit_length is None:
            raise ValueError("raw clips need a window length: config['pretrain']['input']['length_wave'] or unit_length=")
        eng, tf = self.tfms.engine, self.tfms
        packed = clips if isinstance(clips, PackedClips) else None
        if packed is not None:
            device, lengths = packed.raw.device, packed.lengths
            offsets = np.concatenate((packed.offsets, [int(lengths.sum())]))
        else:
            device = next((c.device for c in clips if isinstance(c, torch.Tensor) and c.is_cuda),
                          torch.device("cuda", torch.cuda.current_device()))
            lengths = np.array([int(c.shape[0]) for c in clips], np.int64)
        v1s, v2s = [], []
        with torch.cuda.device(device):
            eng.ensure(self.plan.n_mels, self.plan.num_frames(self.unit_length), device)
            if eng.lookahead and len(lengths) > eng.max_batch:
                raise ValueError(f"lookahead takes at most max_batch = {eng.max_batch} clips per call")
            for lo, hi in _chunks(len(lengths), eng.max_batch):
                adj = lengths[lo:hi] - self.unit_length
                if params is None:       # the window offset of a clip is drawn before its view parameters
                    p = tf.sampler.draw(eng.head, hi - lo, eng.f, eng.t, window_adj=adj)
                else:
                    p = ViewParams(params.alpha[lo:hi], params.src_clip[lo:hi], params.bank_index[lo:hi],
                                   params.crop[lo:hi], None if params.start is None else params.start[lo:hi])
                tf.last_params = p
                if packed is not None:
                    win = extract_windows_packed(packed.raw[offsets[lo]:offsets[hi]], lengths[lo:hi], self.unit_length,
                                                 starts=p.start, with_gain=self.l2)
                else:
                    win = extract_windows(clips[lo:hi], self.unit_length, starts=p.start, with_gain=self.l2, device=device)
                part, gain = win if self.l2 else (win, None)
                if eng.lookahead:
                    return self._front_then_previous_views(part, gain, p)
                a, c = self._windows_to_views(part, gain, lambda p=p: p)
                v1s.append(a); v2s.append(c)
        v1 = v1s[0] if len(v1s) == 1 else torch.cat(v1s)
        v2 = v2s[0] if len(v2s) == 1 else torch.cat(v2s)
        return self.tfms._post(v1, v2)
