"""Host-side draw of the per-view random parameters (alpha, bank entry, crop box).

``ReferenceSampler`` replays the reference's exact draw order on the SAME generators the reference
uses -- numpy's global ``np.random`` and python's ``random`` module -- so that a training script that
seeds them (``np.random.seed(31); random.seed(31)``, /root/reference/extras/delores-s/main.py:64) gets
bit-identical crop boxes and mixup selections (with a ``Kmix`` stage configured its data-dependent draws come after the
batch's draws instead of interleaved per view, so the replay is bit-exact only WITHOUT Kmix; see module.py):

    per view:  np.random.random()            -> alpha      (MixupBYOLA.forward, augmentations.py:99)
               np.random.randint(len(bank))  -> bank entry (:102, only when the bank is not empty)
               np.random.uniform(*freq)      -> h          (RandomResizeCrop.get_params, :34)
               np.random.uniform(*time)      -> w          (:35)
               random.randint(0, ch - h)     -> i          (:36, skipped when ch == h)
               random.randint(0, cw - w)     -> j          (:37, skipped when cw == w)

When the raw clips go through ``extract_window`` first (BaseDataset.__getitem__, upstream_dataset.py:60),
its ``random.randint(0, len - unit)`` (utils.py:179, skipped when the clip is not longer than the window)
comes BEFORE the clip's view draws on the same ``random`` stream; pass ``window_adj`` to ``draw``.

``FastSampler`` draws the same distributions from a vectorised ``numpy.random.Generator`` (not
sequence-compatible with the reference; for throughput runs).

Bank bookkeeping (both samplers): the two views of a clip share one FIFO (src/augmentations/__init__.py:10-12,
32-35) which receives the clip once per view, so with ``g`` clips already finished the push log is
[0, 2g): push q holds clip q // 2.  View v of clip g sees the last min(2g + v, n_memory) pushes.
"""
from __future__ import annotations

import random as _pyrandom
from dataclasses import dataclass

import numpy as np


@dataclass
class ViewParams:
    """Explicit parameters of a batch, host arrays laid out as the C ABI takes them."""
    alpha: np.ndarray       # [B, 2] float32
    src_clip: np.ndarray    # [B, 2] int64   process-local clip id of the bank entry, -1 = no mix
    bank_index: np.ndarray  # [B, 2] int32   index drawn into the FIFO (-1 when it was empty)
    crop: np.ndarray        # [B, 2, 4] int32 (i, j, h, w)
    start: np.ndarray = None  # [B] int64 extract_window crop offsets, when the sampler drew them


def canvas_size(f, t, virtual_crop_scale):
    """augmentations.py:42."""
    return int(f * virtual_crop_scale[0]), int(t * virtual_crop_scale[1])


def bank_window(pushes_before: int, n_memory: int):
    """(first push still in the FIFO, FIFO length) after ``pushes_before`` pushes."""
    length = min(pushes_before, n_memory)
    return pushes_before - length, length


class ReferenceSampler:
    def __init__(self, ratio=0.4, n_memory=2048, freq_scale=(0.6, 1.5), time_scale=(0.6, 1.5),
                 virtual_crop_scale=(1.0, 1.5), pushes_per_clip=2, mixup=True, crop=True,
                 np_rng=None, py_rng=None):
        self.ratio, self.n_memory = ratio, n_memory
        self.freq_scale, self.time_scale, self.vscale = tuple(freq_scale), tuple(time_scale), tuple(virtual_crop_scale)
        self.pushes_per_clip = pushes_per_clip
        self.mixup, self.crop = mixup, crop
        self.np_rng = np_rng if np_rng is not None else np.random
        self.py_rng = py_rng if py_rng is not None else _pyrandom

    def draw(self, clips_done: int, batch: int, f: int, t: int, window_adj=None) -> ViewParams:
        """``window_adj[b]`` = len(clip b) - unit_length: also draw the extract_window offsets, in place."""
        ch, cw = canvas_size(f, t, self.vscale)
        nv = self.pushes_per_clip
        alpha = np.zeros((batch, 2), np.float32)
        src = np.full((batch, 2), -1, np.int64)
        bidx = np.full((batch, 2), -1, np.int32)
        crop = np.zeros((batch, 2, 4), np.int32)
        crop[:, :, 2], crop[:, :, 3] = f, t
        crop[:, :, 0], crop[:, :, 1] = (ch - f) // 2, (cw - t) // 2     # identity box when cropping is off
        start = np.zeros(batch, np.int64) if window_adj is not None else None
        rnd, rint, uni, pint = self.np_rng.random, self.np_rng.randint, self.np_rng.uniform, self.py_rng.randint
        for b in range(batch):
            g = clips_done + b
            if window_adj is not None and window_adj[b] > 0:
                start[b] = pint(0, int(window_adj[b]))
            for v in range(nv):
                if self.mixup:
                    alpha[b, v] = self.ratio * rnd()
                    lo, length = bank_window(nv * g + v, self.n_memory)
                    if length:
                        r = int(rint(length))
                        bidx[b, v] = r
                        src[b, v] = (lo + r) // nv
                if self.crop:
                    h = int(np.clip(int(uni(*self.freq_scale) * f), 1, ch))
                    w = int(np.clip(int(uni(*self.time_scale) * t), 1, cw))
                    i = pint(0, ch - h) if ch > h else 0
                    j = pint(0, cw - w) if cw > w else 0
                    crop[b, v] = (i, j, h, w)
        return ViewParams(alpha, src, bidx, crop, start)


class FastSampler:
    """Same distributions, vectorised; its stream is NOT the reference's."""

    def __init__(self, ratio=0.4, n_memory=2048, freq_scale=(0.6, 1.5), time_scale=(0.6, 1.5),
                 virtual_crop_scale=(1.0, 1.5), pushes_per_clip=2, mixup=True, crop=True, seed=0):
        self.ratio, self.n_memory = ratio, n_memory
        self.freq_scale, self.time_scale, self.vscale = tuple(freq_scale), tuple(time_scale), tuple(virtual_crop_scale)
        self.pushes_per_clip = pushes_per_clip
        self.mixup, self.crop = mixup, crop
        self.rng = np.random.default_rng(seed)

    def draw(self, clips_done: int, batch: int, f: int, t: int, window_adj=None) -> ViewParams:
        ch, cw = canvas_size(f, t, self.vscale)
        nv = self.pushes_per_clip
        rng = self.rng
        start = None
        if window_adj is not None:
            adj = np.maximum(np.asarray(window_adj, np.int64), 0)
            start = np.minimum(np.floor(rng.random(batch) * (adj + 1)).astype(np.int64), adj)
        alpha = np.zeros((batch, 2), np.float32)
        src = np.full((batch, 2), -1, np.int64)
        bidx = np.full((batch, 2), -1, np.int32)
        crop = np.empty((batch, 2, 4), np.int32)
        crop[:, :, 0], crop[:, :, 1], crop[:, :, 2], crop[:, :, 3] = (ch - f) // 2, (cw - t) // 2, f, t
        if self.mixup:
            alpha[:, :nv] = (self.ratio * rng.random((batch, nv))).astype(np.float32)
            pushes = nv * (clips_done + np.arange(batch, dtype=np.int64))[:, None] + np.arange(nv, dtype=np.int64)[None, :]
            length = np.minimum(pushes, self.n_memory)
            r = np.floor(rng.random((batch, nv)) * length).astype(np.int64)
            r = np.minimum(r, np.maximum(length - 1, 0))
            has = length > 0
            bidx[:, :nv] = np.where(has, r, -1).astype(np.int32)
            src[:, :nv] = np.where(has, (pushes - length + r) // nv, -1)
        if self.crop:
            h = np.clip((rng.uniform(*self.freq_scale, (batch, 2)) * f).astype(np.int64), 1, ch)
            w = np.clip((rng.uniform(*self.time_scale, (batch, 2)) * t).astype(np.int64), 1, cw)
            i = np.floor(rng.random((batch, 2)) * (ch - h + 1)).astype(np.int64)
            j = np.floor(rng.random((batch, 2)) * (cw - w + 1)).astype(np.int64)
            crop[:, :, 0], crop[:, :, 1] = np.minimum(i, ch - h), np.minimum(j, cw - w)
            crop[:, :, 2], crop[:, :, 3] = h, w
        return ViewParams(alpha, src, bidx, crop, start)


def validate_crops(crop: np.ndarray, ch: int, cw: int) -> None:
    """What RandomResizeCrop.get_params guarantees; checked on the host before any launch."""
    c = np.asarray(crop).reshape(-1, 4)
    i, j, h, w = c[:, 0], c[:, 1], c[:, 2], c[:, 3]
    if not (np.all(h >= 1) and np.all(w >= 1) and np.all(i >= 0) and np.all(j >= 0)
            and np.all(i + h <= ch) and np.all(j + w <= cw)):
        raise ValueError(f"crop box outside the {ch}x{cw} canvas")
