"""Oracle: two-view augmentation (torch CPU + numpy).  TEST INFRASTRUCTURE ONLY.

A functional restatement -- explicit state, explicit parameters, one clip at a time -- of
what the reference's augmentation objects compute:

  * running pre-normalisation   /root/reference/src/augmentations/augmentations.py:215-282
                                (RunningMean / RunningVariance / RunningNorm, incl. the
                                off-by-one divisor at :225-227)
  * log-mixup-exp vs a FIFO bank /root/reference/src/augmentations/augmentations.py:8-12, 82-111
  * random-resize-crop           /root/reference/src/augmentations/augmentations.py:14-55
  * two-view composition         /root/reference/src/augmentations/__init__.py:8-16, 32-35
  * batch post-normalisation     /root/reference/src/augmentations/augmentations.py:305-324
  * fixed-statistics norm        /root/reference/src/augmentations/augmentations.py:285-302

PINNED: ``tests/test_oracle_views.py`` replays the same seeds through the reference's own
classes (imported read-only from /root/reference when that tree is present) and through the
fixtures in ``tests/golden/`` that were generated from them; outputs and drawn integers must
be identical.  The resampling arithmetic is ATen's (``F.interpolate``), called directly as
the reference does; :func:`resample_numpy` is an independent scalar restatement of it.
"""
from __future__ import annotations

import random as _pyrandom
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F

F32_EPS = torch.finfo(torch.float32).eps
F32_MAX = torch.finfo(torch.float32).max


# --------------------------------------------------------------------------- records
@dataclass
class ViewRecord:
    """Everything random or stateful that went into one view of one clip."""
    alpha: float          # mixing weight on the bank entry (x gets 1 - alpha)
    bank_index: int       # index drawn into the FIFO at that moment, -1 when the bank was empty
    src_clip: int         # process-local id of the clip that bank entry came from, -1 if none
    crop: Tuple[int, int, int, int]   # (i, j, h, w) in canvas coordinates


@dataclass
class ClipRecord:
    clip: int
    mean: float
    std: float
    views: List[ViewRecord] = field(default_factory=list)


# --------------------------------------------------------------------------- pieces
def log_mixup_exp(xa, xb, alpha):
    """augmentations.py:8-12 -- alpha weights ``xa``."""
    mixed = alpha * xa.exp() + (1.0 - alpha) * xb.exp()
    return torch.log(mixed + torch.finfo(mixed.dtype).eps)


def canvas_size(f, t, virtual_crop_scale=(1.0, 1.5)):
    """augmentations.py:42 -- ``int(s * c)`` per axis."""
    return int(f * virtual_crop_scale[0]), int(t * virtual_crop_scale[1])


def draw_crop(canvas_hw, in_hw, freq_scale=(0.6, 1.5), time_scale=(0.6, 1.5),
              np_rng=np.random, py_rng=_pyrandom):
    """augmentations.py:30-38.  Two numpy draws, then up to two python-``random`` draws
    (each skipped when the crop spans the canvas on that axis)."""
    ch, cw = canvas_hw
    sh, sw = in_hw
    h = int(np.clip(int(np_rng.uniform(*freq_scale) * sh), 1, ch))
    w = int(np.clip(int(np_rng.uniform(*time_scale) * sw), 1, cw))
    i = py_rng.randint(0, ch - h) if ch > h else 0
    j = py_rng.randint(0, cw - w) if cw > w else 0
    return i, j, h, w


def resize_crop(x, crop, virtual_crop_scale=(1.0, 1.5), mode="bicubic"):
    """augmentations.py:40-55 with the crop passed in.  ``x``: float32 [C, F, T]."""
    c, f, t = x.shape
    ch, cw = canvas_size(f, t, virtual_crop_scale)
    canvas = torch.zeros((c, ch, cw), dtype=torch.float32)
    x0, y0 = (cw - t) // 2, (ch - f) // 2
    canvas[:, y0:y0 + f, x0:x0 + t] = x
    i, j, h, w = crop
    patch = canvas[:, i:i + h, j:j + w]
    out = F.interpolate(patch.unsqueeze(0), size=(f, t), mode=mode, align_corners=True)
    return out.squeeze(0).to(torch.float32)


def normalize_batch(x, axis=(0, 2, 3)):
    """augmentations.py:321-324 -- unbiased std, clamped to [eps, max]."""
    mean = x.mean(dim=list(axis), keepdim=True)
    std = torch.clamp(x.std(dim=list(axis), keepdim=True), F32_EPS, F32_MAX)
    return (x - mean) / std


def precomputed_norm(x, mean, std):
    """augmentations.py:299-300."""
    return (x - mean) / std


# --------------------------------------------------------------------------- running norm
class RunningNormOracle:
    """augmentations.py:215-282, state spelled out.

    ``n`` counts ``put`` calls.  The divisor used for the k-th update (k = n before the
    increment) is ``n`` itself, so after k >= 1 updates ``mu`` is the mean of samples
    1..k -- sample 0 is overwritten by sample 1 (:225-227).  The variance tracker applies
    the same rule to ``mean((x - mu)**2)`` with the *already updated* ``mu`` (:277-279).
    """

    def __init__(self, epoch_samples, max_update_epochs=10, axis=(1, 2)):
        self.max_update = epoch_samples * max_update_epochs
        self.axis = list(axis)
        self.n = 0
        self.mu = None
        self.s2 = None
        self.mean = None
        self.std = None

    def __call__(self, x):
        if self.n < self.max_update:
            m = x.mean(self.axis, keepdim=True)
            if self.n == 0:
                self.mu = m
            else:
                self.mu = self.mu + (m - self.mu) / self.n
            v = ((x - self.mu) ** 2).mean(self.axis, keepdim=True)
            if self.n == 0:
                self.s2 = v
            else:
                self.s2 = self.s2 + (v - self.s2) / self.n
            self.n += 1
            self.mean = self.mu
            # the reference takes the root with numpy on the torch tensor (:255-256); torch.sqrt can
            # differ from it in the last bit, so do the same
            self.std = torch.clamp(torch.as_tensor(np.sqrt(self.s2.numpy())), F32_EPS, F32_MAX)
        return (x - self.mean) / self.std


# --------------------------------------------------------------------------- the two-view pipeline
class TwoViewOracle:
    """``AugmentationModule`` (src/augmentations/__init__.py:8-35) for the working pipeline
    RunningNorm -> MixupBYOLA -> RandomResizeCrop, with ONE mixup bank shared by both views.

    Call it clip by clip with float32 ``[1, F, T]``; it returns ``(view1, view2)`` and appends a
    :class:`ClipRecord` to ``self.records``.  Randomness comes from the numpy *global* generator
    and python's ``random`` module in exactly the reference's order:
        per view: np.random.random() [alpha] -> np.random.randint(len(bank)) [only if bank]
                  -> np.random.uniform [h] -> np.random.uniform [w]
                  -> random.randint [i, if canvas_h > h] -> random.randint [j, if canvas_w > w]
    """

    def __init__(self, len_of_files, ratio=0.4, n_memory=2048, use_log_mixup=True,
                 virtual_crop_scale=(1.0, 1.5), freq_scale=(0.6, 1.5), time_scale=(0.6, 1.5),
                 normalization="mean_var", mode="bicubic", np_rng=np.random, py_rng=_pyrandom):
        self.ratio, self.n_memory, self.use_log_mixup = ratio, n_memory, use_log_mixup
        self.vscale, self.freq_scale, self.time_scale = virtual_crop_scale, freq_scale, time_scale
        self.mode = mode
        self.np_rng, self.py_rng = np_rng, py_rng
        self.pre_norm = RunningNormOracle(2 * len_of_files) if normalization == "mean_var" else None
        self.bank: List[torch.Tensor] = []
        self.bank_clip: List[int] = []       # which clip each bank entry came from
        self.clips_seen = 0
        self.records: List[ClipRecord] = []

    # -- MixupBYOLA.forward, augmentations.py:97-111
    def _mix(self, xn, clip_id):
        alpha = self.ratio * self.np_rng.random()
        if self.bank:
            r = int(self.np_rng.randint(len(self.bank)))
            z = self.bank[r]
            src = self.bank_clip[r]
            if self.use_log_mixup:
                mixed = log_mixup_exp(xn, z, 1.0 - alpha)
            else:
                mixed = alpha * z + (1.0 - alpha) * xn
        else:
            r, src, mixed = -1, -1, xn
        self.bank = (self.bank + [xn])[-self.n_memory:]
        self.bank_clip = (self.bank_clip + [clip_id])[-self.n_memory:]
        return mixed.to(torch.float32), alpha, r, src

    def __call__(self, x):
        x = torch.as_tensor(x, dtype=torch.float32)
        clip_id = self.clips_seen
        if self.pre_norm is not None:
            xn = self.pre_norm(x)
            rec = ClipRecord(clip_id, float(self.pre_norm.mean), float(self.pre_norm.std))
        else:
            xn = x
            rec = ClipRecord(clip_id, 0.0, 1.0)
        _, f, t = x.shape
        views = []
        for _ in range(2):
            mixed, alpha, r, src = self._mix(xn, clip_id)
            crop = draw_crop(canvas_size(f, t, self.vscale), (f, t), self.freq_scale,
                             self.time_scale, self.np_rng, self.py_rng)
            views.append(resize_crop(mixed, crop, self.vscale, self.mode))
            rec.views.append(ViewRecord(float(alpha), r, src, tuple(int(c) for c in crop)))
        self.clips_seen += 1
        self.records.append(rec)
        return views[0], views[1]

    # -- convenience: run a whole batch [B, F, T] in order, stack the outputs
    def run_batch(self, lms):
        lms = torch.as_tensor(lms, dtype=torch.float32)
        first = len(self.records)
        v1, v2 = [], []
        for b in range(lms.shape[0]):
            a, c = self(lms[b:b + 1])
            v1.append(a)
            v2.append(c)
        recs = self.records[first:]
        return torch.stack(v1), torch.stack(v2), recs


def records_to_arrays(recs):
    """Pack records into the arrays the C ABI takes: mean_std [B,2] f32, alpha [B,2] f32,
    src_clip [B,2] i32 (process-local clip ids, -1 = none), crop [B,2,4] i32 (i,j,h,w)."""
    b = len(recs)
    mean_std = np.zeros((b, 2), np.float32)
    alpha = np.zeros((b, 2), np.float32)
    src = np.full((b, 2), -1, np.int32)
    bank_index = np.full((b, 2), -1, np.int32)
    crop = np.zeros((b, 2, 4), np.int32)
    for k, r in enumerate(recs):
        mean_std[k] = (r.mean, r.std)
        for v, vr in enumerate(r.views):
            alpha[k, v] = vr.alpha
            src[k, v] = vr.src_clip
            bank_index[k, v] = vr.bank_index
            crop[k, v] = vr.crop
    return dict(mean_std=mean_std, alpha=alpha, src_clip=src, bank_index=bank_index, crop=crop)


# --------------------------------------------------------------------------- independent resampler
_A = -0.75


def _cc1(x):   # |x| <= 1
    return ((_A + 2.0) * x - (_A + 3.0)) * x * x + 1.0


def _cc2(x):   # 1 < |x| < 2
    return ((_A * x - 5.0 * _A) * x + 8.0 * _A) * x - 4.0 * _A


def resample_numpy(patch, out_hw, mode="bicubic"):
    """Scalar restatement (float32 arithmetic) of ATen's ``upsample_bicubic2d`` /
    ``upsample_bilinear2d`` with ``align_corners=True`` -- the kernel the reference reaches at
    augmentations.py:53-54.  ``patch``: float32 [h, w] -> float32 [out_h, out_w].
    Source coordinate ``dst * (in-1)/(out-1)``; bicubic taps are clamped to the patch."""
    patch = np.asarray(patch, dtype=np.float32)
    h, w = patch.shape
    oh, ow = out_hw
    f32 = np.float32

    def axis(n_in, n_out):
        scale = f32(n_in - 1) / f32(n_out - 1) if n_out > 1 else f32(0)
        src = (np.arange(n_out, dtype=np.float32) * scale).astype(np.float32)
        i0 = np.floor(src).astype(np.int64)
        t = (src - i0.astype(np.float32)).astype(np.float32)
        return i0, t

    y0, ty = axis(h, oh)
    x0, tx = axis(w, ow)
    if mode == "bilinear":
        y1 = np.minimum(y0 + 1, h - 1)
        x1 = np.minimum(x0 + 1, w - 1)
        wy1, wx1 = ty[:, None], tx[None, :]
        wy0, wx0 = f32(1) - wy1, f32(1) - wx1
        top = patch[y0][:, x0] * wx0 + patch[y0][:, x1] * wx1
        bot = patch[y1][:, x0] * wx0 + patch[y1][:, x1] * wx1
        return (top * wy0 + bot * wy1).astype(np.float32)
    if mode != "bicubic":
        raise ValueError(mode)

    def weights(t):
        return np.stack([_cc2(t + f32(1)), _cc1(t), _cc1(f32(1) - t), _cc2(f32(2) - t)]).astype(np.float32)

    wy, wx = weights(ty), weights(tx)                # [4, oh], [4, ow]
    out = np.zeros((oh, ow), np.float32)
    for a in range(4):
        ya = np.clip(y0 + a - 1, 0, h - 1)
        row = np.zeros((oh, ow), np.float32)
        for b in range(4):
            xb = np.clip(x0 + b - 1, 0, w - 1)
            row += patch[ya][:, xb] * wx[b][None, :]
        out += row * wy[a][:, None]
    return out
