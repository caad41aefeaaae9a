"""Device-side batch assembly: the step after the hot path.

The reference builds a batch on the CPU -- ``DataLoader(BaseDataset, shuffle=True, drop_last=True,
pin_memory=True)`` (/root/reference/src/dataset/upstream_dataset.py:118-125) runs ``__getitem__`` per clip
(decode, extract_window, log-mel, two views), collates the views and copies them to the GPU.  Here the
worker side only DECODES: raw clips are packed back to back into a pinned staging buffer, one
asynchronous copy moves them to the GPU on a side stream while the previous batch is still being
computed, and the fused pipeline (extract_window -> log-mel -> statistics -> two views) produces
``(img_1, img_2)``, each ``[B, 1, F, T]`` fp32 ON THE DEVICE -- the tuple the upstream experts' training_step
unpacks (src/upstream/delores_s/upstream_expert.py:198) -- with no CPU round trip.

Index order: ``torch.utils.data.RandomSampler`` + ``BatchSampler`` are used as they are, so a seeded run
visits the clips in the same order as the reference's DataLoader.  Augmentation draws follow the
reference's per-clip order for ``num_workers=0`` (one stream of ``random`` / ``np.random``); with worker
processes the reference's draws depend on per-worker seeds and are not reproducible by anyone.
"""
from __future__ import annotations

import threading
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass

import numpy as np
import torch
from torch.utils.data import BatchSampler, RandomSampler, SequentialSampler


@dataclass
class PackedClips:
    """Raw clips of one batch, back to back in one device buffer."""
    raw: torch.Tensor          # [sum(lengths)] fp32, CUDA
    lengths: np.ndarray        # [B] int64 (host)
    ready: torch.cuda.Event = None   # recorded after the upload; the consumer stream waits on it

    @property
    def offsets(self):
        return np.concatenate(([0], np.cumsum(self.lengths)[:-1])).astype(np.int64)

    def __len__(self):
        return int(self.lengths.shape[0])


class WaveformDataset:
    """``files[i]`` -> 1-D float32 waveform through ``load(path)`` (decoding is the caller's: the reference uses
    ``librosa.core.load(path, sr=16000)``, upstream_dataset.py:55); or wrap in-memory clips with ``load=None``."""

    def __init__(self, files, load=None):
        self.files, self.load = files, load

    def __len__(self):
        return len(self.files)

    def __getitem__(self, i):
        x = self.files[i] if self.load is None else self.load(self.files[i])
        if isinstance(x, torch.Tensor):
            x = x.detach().cpu().numpy()
        return np.ascontiguousarray(x, dtype=np.float32).reshape(-1)


class DeviceBatchLoader:
    """Iterate ``(img_1, img_2)`` device batches, like ``BaselineDataModule.train_dataloader()`` would after
    ``.to(device)``.

    dataset      indexable of raw 1-D waveforms (``WaveformDataset``)
    pipeline     ``UpstreamInputPipeline`` (owns the augmentation state, the bank and the window length)
    batch_size / shuffle / drop_last / generator   as ``torch.utils.data.DataLoader`` (defaults = the reference's)
    num_workers  decode threads (0 = decode in the prefetch thread)
    prefetch     staged batches in flight (pinned buffers in rotation)
    """

    def __init__(self, dataset, pipeline, batch_size=8, shuffle=True, drop_last=True, generator=None,
                 num_workers=0, prefetch=2, device=None):
        self.dataset, self.pipeline = dataset, pipeline
        self.batch_size, self.drop_last = int(batch_size), bool(drop_last)
        sampler = RandomSampler(dataset, generator=generator) if shuffle else SequentialSampler(dataset)
        self.batch_sampler = BatchSampler(sampler, self.batch_size, self.drop_last)
        self.num_workers, self.prefetch = int(num_workers), max(1, int(prefetch))
        self.device, self.generator = device, generator
        self._pinned = [None] * (self.prefetch + 1)

    def __len__(self):
        return len(self.batch_sampler)

    def batches(self):
        """The index lists of one epoch, in the order the reference's DataLoader would visit them under the same
        torch seed: its iterator draws one int64 (the workers' base seed) before the sampler draws the
        permutation seed, so the same value is drawn and dropped here."""
        torch.empty((), dtype=torch.int64).random_(generator=self.generator)
        return list(self.batch_sampler)

    # -- host side: decode + pack into pinned memory + asynchronous upload
    def _stage(self, indices, slot, pool, copy_stream, device):
        clips = list(pool.map(self.dataset.__getitem__, indices)) if pool else [self.dataset[i] for i in indices]
        lengths = np.array([c.shape[0] for c in clips], np.int64)
        total = int(lengths.sum())
        buf = self._pinned[slot]
        if buf is None or buf.numel() < max(total, 1):
            buf = torch.empty(max(total, 1) * 5 // 4, dtype=torch.float32).pin_memory()
            self._pinned[slot] = buf
        view = buf.numpy()
        o = 0
        for c in clips:
            view[o:o + c.shape[0]] = c
            o += c.shape[0]
        with torch.cuda.stream(copy_stream):
            raw = torch.empty(max(total, 1), dtype=torch.float32, device=device)
            raw[:max(total, 1)].copy_(buf[:max(total, 1)], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(copy_stream)
        return PackedClips(raw, lengths, ev)

    def __iter__(self):
        if not torch.cuda.is_available():
            raise RuntimeError("DeviceBatchLoader assembles batches on the GPU; no CUDA device is visible")
        device = self.device or torch.device("cuda", torch.cuda.current_device())
        copy_stream = torch.cuda.Stream(device=device)
        pool = ThreadPoolExecutor(self.num_workers) if self.num_workers > 0 else None
        index_lists = self.batches()
        staged, lock = {}, threading.Condition()
        n_slots = len(self._pinned)
        released = [threading.Event() for _ in range(n_slots)]      # pinned buffer free again (its upload finished)
        for e in released:
            e.set()
        failure = []

        def producer():
            try:
                for k, idx in enumerate(index_lists):
                    slot = k % n_slots
                    released[slot].wait()
                    released[slot].clear()
                    packed = self._stage(idx, slot, pool, copy_stream, device)
                    with lock:
                        staged[k] = packed
                        lock.notify_all()
                    with lock:                                       # at most ``prefetch`` batches ahead
                        while len(staged) > self.prefetch:
                            lock.wait(0.05)
            except BaseException as exc:                             # surfaced in the consumer
                failure.append(exc)
                with lock:
                    lock.notify_all()

        th = threading.Thread(target=producer, daemon=True)
        th.start()
        try:
            for k in range(len(index_lists)):
                with lock:
                    while k not in staged and not failure:
                        lock.wait(0.05)
                    if failure:
                        raise failure[0]
                    packed = staged.pop(k)
                    lock.notify_all()
                torch.cuda.current_stream(device).wait_event(packed.ready)
                packed.ready.synchronize()                           # upload done -> the pinned slot may be refilled
                released[k % n_slots].set()
                packed.raw.record_stream(torch.cuda.current_stream(device))
                out = self.pipeline(packed)
                if out is not None:                                  # (a lookahead pipeline returns an earlier batch at call k)
                    yield out
            while getattr(self.pipeline, "flush", None):             # whatever is still pending at the end of the epoch
                last = self.pipeline.flush()
                if last is None:
                    break
                yield last
        finally:
            for e in released:
                e.set()
            with lock:
                staged.clear()
                lock.notify_all()
            th.join(timeout=5)
            if pool:
                pool.shutdown(wait=False)


class BaselineDataModule:
    """``BaselineDataModule(config, args, tfms, data_csv, batch_size, num_workers)``
    (upstream_dataset.py:96-125) without Lightning: ``setup()`` then ``train_dataloader()`` returns a
    :class:`DeviceBatchLoader`.  ``tfms`` is an ``UpstreamInputPipeline`` (the reference passes the
    ``AugmentationModule``; here the transform and the front end are one fused object).  ``data_csv`` is the
    reference's csv with a ``files`` column, or any sequence of paths / in-memory clips; ``load`` decodes one path."""

    def __init__(self, config, args, tfms, data_csv="./", batch_size=8, num_workers=8, load=None):
        self.config, self.args, self.transformation = config, args, tfms
        self.data_dir_train, self.batch_size, self.num_workers, self.load = data_csv, batch_size, num_workers, load
        self.dataset_sizes = {}
        self.train_dataset = None

    def setup(self, stage=None):
        if stage == "fit" or stage is None:
            files = self.data_dir_train
            if isinstance(files, str):
                import pandas as pd
                files = list(pd.read_csv(files)["files"])
                if self.load is None:
                    raise ValueError("pass load=<callable(path) -> waveform>: audio decoding is outside this package")
            self.train_dataset = WaveformDataset(files, self.load)
            self.dataset_sizes["train"] = len(self.train_dataset)

    def train_dataloader(self):
        return DeviceBatchLoader(self.train_dataset, self.transformation, batch_size=self.batch_size, shuffle=True,
                                 drop_last=True, num_workers=self.num_workers)
