"""Drop-in classes for the reference's augmentation module, backed by the CUDA kernels.

Same names, constructor arguments and call shapes as /root/reference/src/augmentations/augmentations.py
and /root/reference/src/augmentations/__init__.py.  Each ``forward`` takes what the reference takes --
one clip ``[1, F, T]`` -- or a batch ``[B, 1, F, T]`` whose clips are processed in batch order with the
same sequential semantics (running statistics, FIFO bank) the reference has when it is called clip
after clip.  Tensors live on the GPU; CPU inputs are copied over.  Random draws come from the same
generators as the reference (``np.random`` global + python ``random``) in the same order unless a
different sampler is given; explicit parameters can be passed for parity runs.
"""
from __future__ import annotations

import ctypes as C
import warnings

import numpy as np
import torch
from torch import nn

from . import _native
from .engine import INTERP, ViewEngine
from .frontend import _ptr, _stream_ptr, as_cuda_f32
from .sampler import FastSampler, ReferenceSampler, ViewParams, canvas_size

F32_EPS = float(torch.finfo(torch.float32).eps)


def _as_batch(x):
    """[1,F,T] -> ([1,F,T] batch of one, True) ; [B,1,F,T] -> ([B,F,T], False)."""
    x = as_cuda_f32(x)
    if x.dim() == 3:
        if x.shape[0] != 1:
            raise ValueError("a single clip must be [1, F, T]; pass batches as [B, 1, F, T]")
        return x, True
    if x.dim() == 4:
        if x.shape[1] != 1:
            raise ValueError("batches must be [B, 1, F, T]")
        return x[:, 0], False
    raise ValueError("expected [1, F, T] or [B, 1, F, T]")


def _chunks(n, size):
    return [(s, min(s + size, n)) for s in range(0, n, size)]


class RunningNorm(nn.Module):
    """Online normalisation with running mean/std (augmentations.py:259-282), statistics updated once
    per clip in batch order, including the reference's divisor (the 2nd clip overwrites the 1st)."""

    def __init__(self, epoch_samples, max_update_epochs=10, axis=[1, 2]):
        super().__init__()
        if list(axis) != [1, 2]:
            raise NotImplementedError("only the reference's axis=[1, 2] (one scalar per clip) is built")
        self.max_update = epoch_samples * max_update_epochs
        self.axis = list(axis)
        self._state = None

    def __len__(self):
        return 0 if self._state is None else int(self._state[0].item())

    @property
    def mean(self):
        return None if self._state is None else float(self._state[3].item())

    @property
    def std(self):
        return None if self._state is None else float(self._state[4].item())

    def forward(self, image):
        x, single = _as_batch(image)
        lib = _native.lib()
        with torch.cuda.device(x.device):
            if self._state is None:
                self._state = torch.zeros(8, dtype=torch.float64, device=x.device)
            b, elems = x.shape[0], x[0].numel()
            sums = torch.empty((b, 2), dtype=torch.float64, device=x.device)
            ms = torch.empty((b, 2), dtype=torch.float32, device=x.device)
            y = torch.empty_like(x)
            _native.check(lib.assl_clip_moments_f32(_ptr(x), b, elems, _ptr(sums), _stream_ptr()))
            _native.check(lib.assl_running_stats(_ptr(self._state), _ptr(sums), b, 1, elems, int(self.max_update),
                                                 0, b, _ptr(ms), 0, 0, _stream_ptr()))
            _native.check(lib.assl_clip_normalize_f32(_ptr(x), _ptr(y), b, elems, _ptr(ms), 0, 0, _stream_ptr()))
        return y if single else y.unsqueeze(1)

    def __repr__(self):
        return self.__class__.__name__ + f'(max_update={self.max_update},axis={self.axis})'


class PrecomputedNorm(nn.Module):
    """Normalisation with fixed (mean, std) (augmentations.py:285-302)."""

    def __init__(self, stats, axis=[1, 2]):
        super().__init__()
        self.axis = axis
        self.mean, self.std = stats

    def forward(self, X):
        x = as_cuda_f32(X)
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            _native.check(_native.lib().assl_affine_norm_f32(_ptr(x), _ptr(y), x.numel(), float(self.mean),
                                                             float(self.std), _stream_ptr()))
        return y

    def __repr__(self):
        return self.__class__.__name__ + f'(mean={self.mean}, std={self.std}, axis={self.axis})'


class NormalizeBatch(nn.Module):
    """Batch normalisation of the input (augmentations.py:305-324): one mean / unbiased std over
    axes [0, 2, 3].  ``sync=True`` all-reduces the sums over the process group first."""

    def __init__(self, axis=[0, 2, 3], sync=False):
        super().__init__()
        if list(axis) != [0, 2, 3]:
            raise NotImplementedError("only the reference's axis=[0, 2, 3] is built")
        self.axis, self.sync = list(axis), sync

    def forward(self, X):
        x = as_cuda_f32(X)
        if x.dim() != 4 or x.shape[1] != 1:
            raise ValueError("NormalizeBatch takes [B, 1, F, T]")
        lib = _native.lib()
        y = torch.empty_like(x)
        with torch.cuda.device(x.device):
            sums = torch.zeros(3, dtype=torch.float64, device=x.device)
            _native.check(lib.assl_batch_sums(_ptr(x), x.numel(), _ptr(sums), _stream_ptr()))
            if self.sync:
                import torch.distributed as dist
                if dist.is_available() and dist.is_initialized():
                    dist.all_reduce(sums)
            _native.check(lib.assl_normalize_apply_f32(_ptr(x), _ptr(y), x.numel(), _ptr(sums), _stream_ptr()))
        return y

    def __repr__(self):
        return self.__class__.__name__ + f'(axis={self.axis})'


class RandomResizeCrop(nn.Module):
    """Random resize crop (augmentations.py:14-62): zero canvas, centred paste, random box, resample
    back to the input size.  ``interpolation`` is 'bicubic' like the reference; 'bilinear' is offered too."""

    def __init__(self, virtual_crop_scale=(1.0, 1.5), freq_scale=(0.6, 1.5), time_scale=(0.6, 1.5),
                 interpolation='bicubic'):
        super().__init__()
        self.virtual_crop_scale = virtual_crop_scale
        self.freq_scale = freq_scale
        self.time_scale = time_scale
        self.interpolation = interpolation
        assert time_scale[1] >= 1.0 and freq_scale[1] >= 1.0
        if interpolation not in INTERP:
            raise ValueError(f"interpolation must be one of {sorted(INTERP)}")

    @staticmethod
    def get_params(virtual_crop_size, in_size, time_scale, freq_scale):
        """Same draws in the same order as augmentations.py:30-38 (h from freq_scale, w from time_scale on
        numpy's global generator; i, j from python's ``random``, each skipped when the box spans the canvas)."""
        import random
        canvas_h, canvas_w = virtual_crop_size
        src_h, src_w = in_size
        h = np.clip(int(np.random.uniform(*freq_scale) * src_h), 1, canvas_h)
        w = np.clip(int(np.random.uniform(*time_scale) * src_w), 1, canvas_w)
        i = random.randint(0, canvas_h - h) if canvas_h > h else 0
        j = random.randint(0, canvas_w - w) if canvas_w > w else 0
        return i, j, h, w

    def forward(self, lms, crop=None):
        x, single = _as_batch(lms)
        b, f, t = x.shape
        ch, cw = canvas_size(f, t, self.virtual_crop_scale)
        if crop is None:
            crop = np.asarray([self.get_params((ch, cw), (f, t), self.time_scale, self.freq_scale) for _ in range(b)], np.int32)
        crop = np.asarray(crop, np.int32).reshape(b, 1, 4)
        params = ViewParams(np.zeros((b, 2), np.float32), np.full((b, 2), -1, np.int64), np.full((b, 2), -1, np.int32),
                            np.repeat(crop, 2, axis=1))
        eng = ViewEngine(n_memory=0, pushes_per_clip=1, virtual_crop_scale=self.virtual_crop_scale,
                         interp=self.interpolation, normalize=False, max_batch=max(b, 1))
        with torch.cuda.device(x.device):
            eng.f, eng.t, eng.pitch = f, t, t
            eng.cap = b
            eng.ring, eng.stats = x.contiguous(), None           # the input itself is the "ring" (dense rows)
            v1, _ = eng.launch_views(params, b, x.device)
        return v1[:, 0] if single else v1

    def __repr__(self):
        format_string = self.__class__.__name__ + f'(virtual_crop_size={self.virtual_crop_scale}'
        format_string += ', time_scale={0}'.format(tuple(round(s, 4) for s in self.time_scale))
        format_string += ', freq_scale={0})'.format(tuple(round(r, 4) for r in self.freq_scale))
        return format_string


class MixupBYOLA(nn.Module):
    """Mixup against a FIFO memory bank of past inputs (augmentations.py:82-116).  The bank is a
    device ring; a batch is mixed in order, so clip b can draw clips 0..b-1 of the same batch."""

    def __init__(self, ratio=0.4, n_memory=2048, log_mixup_exp=True, max_batch=4096):
        super().__init__()
        self.ratio = ratio
        self.n = n_memory
        self.log_mixup_exp = log_mixup_exp
        self._engine = ViewEngine(n_memory=n_memory, pushes_per_clip=1, ratio=ratio, log_mixup=log_mixup_exp,
                                  virtual_crop_scale=(1.0, 1.0), normalize=False, max_batch=max_batch)
        self._sampler = ReferenceSampler(ratio=ratio, n_memory=n_memory, virtual_crop_scale=(1.0, 1.0),
                                         pushes_per_clip=1, mixup=True, crop=False)

    @property
    def memory_bank_size(self):
        return min(self._engine.clips_done, self.n)

    def forward(self, x, params: ViewParams = None):
        xb, single = _as_batch(x)
        eng = self._engine
        outs = []
        with torch.cuda.device(xb.device):
            eng.ensure(xb.shape[1], xb.shape[2], xb.device)
            for lo, hi in _chunks(xb.shape[0], eng.max_batch):
                part = xb[lo:hi]
                eng.ring_put(part)
                p = params if params is not None else self._sampler.draw(eng.clips_done, hi - lo, eng.f, eng.t)
                v1, _ = eng.launch_views(p, hi - lo, xb.device)
                outs.append(v1)
        out = outs[0] if len(outs) == 1 else torch.cat(outs)
        return out[:, 0] if single else out

    def __repr__(self):
        format_string = self.__class__.__name__ + f'(ratio={self.ratio},n={self.n}'
        format_string += f',log_mixup_exp={self.log_mixup_exp})'
        return format_string


class Kmix(nn.Module):
    """K-mix (augmentations.py:119-189): MixupBYOLA whose bank entry is drawn from the cluster FARTHEST from the
    incoming clip's (clusters = nearest of the given centroids to the per-band mean over time).

    Same constructor as the reference; ``centroid_path`` is loaded with ``torch.load`` like there, or pass the
    ``[K, F]`` tensor as ``centroids=``.  The cluster features and ids are computed on the GPU for a whole batch
    (``assl_clip_clusters_f32``), come back to the host once per call (K small integers per clip), and the
    selection replays the reference's loop and its draws (``np.random.random()`` for alpha, then
    ``np.random.randint(len(candidates))`` -- ``len(bank)`` while the bank holds fewer than 128 entries) clip by
    clip in batch order; the mix itself is the MixupBYOLA kernel path.  ``top_k`` must be an integer <= 128: for
    anything else the reference raises on its first full bank (``len(bank) < None`` / ``bank[None]``, :143, :178).
    """

    def __init__(self, ratio=0.4, n_memory=2048, log_mixup_exp=True, top_k=None, centroid_path=None, centroids=None,
                 max_batch=4096):
        super().__init__()
        if not isinstance(top_k, int) or top_k > 128:
            raise ValueError("Kmix needs an integer top_k <= 128 (the reference cannot run otherwise)")
        if centroids is None:
            if centroid_path is None:
                raise ValueError("Kmix needs centroids: centroid_path (a torch-saved [K, F] tensor) or centroids=")
            centroids = torch.load(centroid_path)                 # Centroid path must be in torch tensor format
        c = torch.as_tensor(centroids, dtype=torch.float32).cpu()
        if c.dim() != 2:
            raise ValueError("centroids must be [K, F]")
        self.ratio, self.n, self.log_mixup_exp, self.top_k = ratio, n_memory, log_mixup_exp, top_k
        self.centroids = c
        self._unit = c / c.norm(dim=-1, keepdim=True)                                        # :146
        k = c.shape[0]
        self._order = torch.topk(torch.cdist(self._unit, self._unit, p=2), k=k, dim=1).indices.numpy()   # :152
        self._unit_dev = None
        self._engine = ViewEngine(n_memory=n_memory, pushes_per_clip=1, ratio=ratio, log_mixup=log_mixup_exp,
                                  virtual_crop_scale=(1.0, 1.0), normalize=False, max_batch=max_batch)
        self._bank_cluster = np.zeros(0, np.int32)            # cluster of every bank entry, FIFO order
        self.last_params = None
        self.last_candidates = None

    @property
    def memory_bank_size(self):
        return int(self._bank_cluster.shape[0])

    def state_dict(self):
        """Bank (the engine's ring) + the cluster of every bank entry; the np.random / random generator states are the
        caller's to save, as in the reference."""
        return {"engine": self._engine.state_dict(), "bank_cluster": self._bank_cluster.copy()}

    def load_state_dict(self, sd):
        self._engine.load_state_dict(sd["engine"])
        self._bank_cluster = np.asarray(sd["bank_cluster"], np.int32).copy()

    def _clusters(self, xb):
        lib = _native.lib()
        b, f, t = xb.shape
        if f != self.centroids.shape[1]:
            raise ValueError(f"centroids have {self.centroids.shape[1]} bands, the clips {f}")
        if self._unit_dev is None or self._unit_dev.device != xb.device:
            self._unit_dev = self._unit.to(xb.device).contiguous()
        ids = torch.empty((2, b), dtype=torch.int32, device=xb.device)
        _native.check(lib.assl_clip_clusters_f32(_ptr(xb), b, f, t, _ptr(self._unit_dev), self._unit_dev.shape[0],
                                                 _ptr(ids[0]), _ptr(ids[1]), _stream_ptr()), "assl_clip_clusters_f32")
        ids = ids.cpu().numpy()                                # the one device -> host read of the call
        return ids[0], ids[1]

    def _draw(self, g0, mem, point):
        """Replay Kmix.forward's draws for clips g0 .. g0+b-1 (:168-182); returns ViewParams + candidate counts."""
        b = mem.shape[0]
        alpha = np.zeros((b, 2), np.float32)
        src = np.full((b, 2), -1, np.int64)
        bidx = np.full((b, 2), -1, np.int32)
        ncand = np.zeros(b, np.int32)
        bank = self._bank_cluster
        for i in range(b):
            alpha[i, 0] = self.ratio * np.random.random()
            n = bank.shape[0]
            if n:
                if n >= 128:
                    cand = np.zeros(0, np.int64)
                    for tgt in self._order[point[i]]:
                        cand = np.nonzero(bank == tgt)[0]
                        if cand.size:
                            break
                    cand = cand[:128]
                    j = int(cand[np.random.randint(len(cand))])
                    ncand[i] = len(cand)
                else:
                    j = int(np.random.randint(n))
                    ncand[i] = n
                bidx[i, 0] = j
                src[i, 0] = g0 + i - n + j                    # bank position j holds clip (g0 + i) - n + j
            bank = np.concatenate((bank, mem[i:i + 1]))[-self.n:]
        self._bank_cluster = bank
        f, t = self._engine.f, self._engine.t
        crop = np.zeros((b, 2, 4), np.int32)
        crop[:, :, 2], crop[:, :, 3] = f, t                   # identity box: no resampling
        return ViewParams(alpha, src, bidx, crop), ncand

    def forward(self, x, params: ViewParams = None):
        xb, single = _as_batch(x)
        eng = self._engine
        outs = []
        with torch.cuda.device(xb.device):
            eng.ensure(xb.shape[1], xb.shape[2], xb.device)
            for lo, hi in _chunks(xb.shape[0], eng.max_batch):
                part = xb[lo:hi].contiguous()
                eng.ring_put(part)
                mem, point = self._clusters(part)
                if params is None:
                    p, self.last_candidates = self._draw(eng.clips_done, mem, point)
                else:
                    p = ViewParams(params.alpha[lo:hi], params.src_clip[lo:hi], params.bank_index[lo:hi], params.crop[lo:hi])
                    self._bank_cluster = np.concatenate((self._bank_cluster, mem))[-self.n:]
                self.last_params = p
                v1, _ = eng.launch_views(p, hi - lo, xb.device)
                outs.append(v1)
        out = outs[0] if len(outs) == 1 else torch.cat(outs)
        return out[:, 0] if single else out

    def __repr__(self):
        format_string = self.__class__.__name__ + f'(ratio={self.ratio},n={self.n}'
        format_string += f',log_mixup_exp={self.log_mixup_exp})'
        return format_string
