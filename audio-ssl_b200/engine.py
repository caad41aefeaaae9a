"""Device-side state and launch plumbing shared by the drop-in augmentation classes.

``ViewEngine`` owns (through the torch caching allocator -- the C library owns no data buffers):
  * the log-mel RING: ``cap`` slots of ``[F, T]`` fp32.  Clip ``g`` of this process lives in slot
    ``g % cap``; the spectrogram the front end writes there is also the clip's future mixup-bank
    entry, so the FIFO of the reference's ``MixupBYOLA`` (augmentations.py:95, 109) never copies;
  * the statistics ring ``[cap, 2]`` (mean, std current when each clip was normalised -- normalising
    a bank entry on the fly equals storing the normalised tensor as the reference does);
  * the running-norm state (8 doubles) of ``RunningNorm`` (augmentations.py:259-282).

One process per GPU; with ``torch.distributed`` initialised and ``sync_stats=True`` the per-clip
(sum, sum of squares) pairs are exchanged (rank-major = global clip order; over peer memory, see ``StatsExchange``) so
every rank runs the identical statistics scan; the bank stays rank-local, as each reference worker's does.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np
import torch

from . import _native
from .frontend import _ptr, _stream_ptr
from .sampler import ViewParams, canvas_size, validate_crops

INTERP = {"bicubic": 0, "bilinear": 1}


def _dist_world():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def heavy_first(crop, n_time=None, canvas_w=None):
    """Launch order of the views of a batch for ``assl_two_views_ordered_f32``: [2B] int32, view ids (2*clip + view)
    sorted by decreasing cost.  A view mixes h crop rows over the columns of its crop that lie inside the pasted clip
    (``n_time`` columns centred in a canvas of ``canvas_w``), in groups of 32, and resamples the h rows along time;
    without the canvas geometry the whole crop width counts."""
    c = np.asarray(crop, np.int64).reshape(-1, 4)
    width = c[:, 3]
    if n_time is not None and canvas_w is not None:
        x0 = (int(canvas_w) - int(n_time)) // 2
        width = np.clip(np.minimum(c[:, 1] + c[:, 3], x0 + int(n_time)) - np.maximum(c[:, 1], x0), 0, None)
    cost = c[:, 2] * (15 * ((width + 31) // 32) + 55)
    return np.argsort(-cost, kind="stable").astype(np.int32)


def shard_bounds(n_global: int, rank: int, world: int):
    """Contiguous block split of a global batch by clip index."""
    base, rem = divmod(n_global, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_clip_sums(local):
    """All-gather per-clip (sum, sum of squares) [b, 2] -> [world*b, 2] in rank-major order, which is
    the global clip order of the contiguous block split.  NCCL on GPUs, gloo in the CPU tests."""
    import torch.distributed as dist
    world = dist.get_world_size()
    gathered = torch.empty((world * local.shape[0], 2), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(gathered, local.contiguous())
    return gathered


class _RawBytes:
    """A device buffer of the C library as a uint8 ``__cuda_array_interface__`` object (torch.as_tensor wraps it)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


class StatsExchange:
    """Per-clip sums across the GPUs of one node (``assl_exchange_*``): every rank stores flagged records straight into
    every peer's buffer over NVLink and scans its own buffer -- one kernel (``run``: send + scan fused) or two
    (``send`` right behind the log-mel kernel, ``scan`` one step later, when the records have long landed: the
    lookahead order).  ``torch.distributed`` only carries the 64-byte IPC handles once.  When the devices cannot map
    each other's memory (no P2P, more than one node) or ``ASSL_STATS_EXCHANGE=nccl`` is set, the same kernels run with
    the STAGED transport: a rank writes its segment into its own buffer and an all-gather moves the segments.  Ranks
    may bring different numbers of clips to a call (the records carry a header)."""

    def __init__(self, handle, rank, world, max_clips, staged, device):
        self.handle, self.rank, self.world, self.max_clips = handle, rank, world, max_clips
        self.staged, self.device = staged, device

    @classmethod
    def open(cls, max_clips, device):
        import os
        import torch.distributed as dist
        rank, world = _dist_world()
        if world < 2 or world > 16:
            return None
        lib = _native.lib()
        want_peer = os.environ.get("ASSL_STATS_EXCHANGE", "peer") != "nccl"
        mine = (C.c_ubyte * 64)()
        handle = C.c_void_p()
        _native.check(lib.assl_exchange_create(C.byref(handle), rank, world, int(max_clips), mine), "assl_exchange_create")
        local = torch.tensor(list(mine), dtype=torch.uint8, device=device)
        everyone = torch.empty(world * 64, dtype=torch.uint8, device=device)
        dist.all_gather_into_tensor(everyone, local)
        raw = bytes(everyone.cpu().tolist())
        ok = want_peer and lib.assl_exchange_connect(handle, raw) == 0
        agree = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN)               # all ranks take the same transport
        staged = int(agree.item()) == 0
        if staged:
            if ok:                                                  # some other rank could not map: start over, staged
                lib.assl_exchange_destroy(handle)
                handle = C.c_void_p()
                _native.check(lib.assl_exchange_create(C.byref(handle), rank, world, int(max_clips), mine))
            _native.check(lib.assl_exchange_use_staging(handle), "assl_exchange_use_staging")
        secs = os.environ.get("ASSL_EXCHANGE_TIMEOUT_S")
        ex = cls(handle, rank, world, int(max_clips), staged, device)
        if secs:
            ex.set_timeout(float(secs))
        return ex

    def set_timeout(self, seconds):
        _native.check(_native.lib().assl_exchange_set_timeout(self.handle, float(seconds)), "assl_exchange_set_timeout")

    def check(self):
        """Raise once a scan gave up waiting for a peer (reads a host word the kernel sets: no synchronisation)."""
        if self.handle and _native.lib().assl_exchange_error(self.handle) != 0:
            raise RuntimeError("statistics exchange: a peer's records did not arrive within the timeout; the running "
                               "statistics of this rank were left untouched by that step and no longer agree with the "
                               "other ranks' (assl_exchange_error)")

    def _gather(self):
        """Staged transport: all-gather the segments of the call sent last into this rank's buffer."""
        import torch.distributed as dist
        buf, seg = C.c_void_p(), C.c_int64()
        _native.check(_native.lib().assl_exchange_staging(self.handle, 0, C.byref(buf), C.byref(seg)), "assl_exchange_staging")
        whole = torch.as_tensor(_RawBytes(buf.value, seg.value * self.world), device=self.device)
        mine = whole[self.rank * seg.value:(self.rank + 1) * seg.value].clone()
        dist.all_gather_into_tensor(whole, mine)
        self._keep = mine

    def send(self, sums, b, n_parts):
        self.check()
        _native.check(_native.lib().assl_exchange_send(self.handle, _ptr(sums), b, n_parts, _stream_ptr()), "assl_exchange_send")
        if self.staged:
            self._gather()

    def scan(self, state, elems, max_update, stats, slot0, cap, write_all=False, deferred=False):
        _native.check(_native.lib().assl_running_stats_exchange_scan(
            self.handle, _ptr(state), elems, int(max_update), _ptr(stats), slot0, cap, 1 if write_all else 0,
            1 if deferred else 0, _stream_ptr()), "assl_running_stats_exchange_scan")

    def scan_send(self, state, elems, max_update, stats, slot0, cap, sums, b, n_parts):
        """The deferred scan of the call sent a step ago and the send of this step's sums in ONE launch
        (assl_exchange_scan_send); two launches with the staged transport (the segments have to be gathered in between)."""
        self.check()
        if self.staged:
            self.scan(state, elems, max_update, stats, slot0, cap, deferred=True)
            self.send(sums, b, n_parts)
            return
        _native.check(_native.lib().assl_exchange_scan_send(
            self.handle, _ptr(state), elems, int(max_update), _ptr(stats), slot0, cap, 0, _ptr(sums), b, n_parts,
            _stream_ptr()), "assl_exchange_scan_send")

    def run(self, state, sums, b, n_parts, elems, max_update, stats, slot0, cap, write_all=False):
        """Send + scan of one call: ONE kernel over peer memory; send, all-gather, scan with the staged transport."""
        self.check()
        if self.staged:
            self.send(sums, b, n_parts)
            self.scan(state, elems, max_update, stats, slot0, cap, write_all)
            return
        _native.check(_native.lib().assl_running_stats_exchange(
            self.handle, _ptr(state), _ptr(sums), b, n_parts, elems, int(max_update), _ptr(stats), slot0, cap,
            1 if write_all else 0, _stream_ptr()), "assl_running_stats_exchange")

    def close(self):
        if self.handle:
            _native.lib().assl_exchange_destroy(self.handle)
            self.handle = None


class _RawCudaArray:
    """A cudaMalloc'd buffer of the C library seen through ``__cuda_array_interface__`` (torch.as_tensor wraps it)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f4", "data": (int(ptr), False),
                                         "version": 3, "strides": None}


class PeerRing:
    """The log-mel ring of the global-bank mode: allocated by the C library so that its CUDA IPC handle can be opened
    by the other ranks of the node; ``ptrs`` = every rank's ring as mapped into this process."""

    def __init__(self, shape, device):
        import torch.distributed as dist
        lib = _native.lib()
        rank, world = _dist_world()
        self.local = C.c_void_p()
        mine = (C.c_ubyte * 64)()
        nbytes = 4 * int(np.prod(shape))
        with torch.cuda.device(device):
            _native.check(lib.assl_peer_alloc(C.byref(self.local), nbytes, mine), "assl_peer_alloc")
            everyone = torch.empty(world * 64, dtype=torch.uint8, device=device)
            dist.all_gather_into_tensor(everyone, torch.tensor(list(mine), dtype=torch.uint8, device=device))
            raw = bytes(everyone.cpu().tolist())
            self.ptrs = (C.c_void_p * world)()
            for r in range(world):
                if r == rank:
                    self.ptrs[r] = self.local.value
                else:
                    p = C.c_void_p()
                    _native.check(lib.assl_peer_open(raw[64 * r:64 * (r + 1)], C.byref(p)), "assl_peer_open")
                    self.ptrs[r] = p.value
        self.rank, self.world = rank, world
        self.tensor = torch.as_tensor(_RawCudaArray(self.local.value, shape), device=device)


class ViewEngine:
    def __init__(self, n_memory=2048, pushes_per_clip=2, ratio=0.4, log_mixup=True,
                 virtual_crop_scale=(1.0, 1.5), interp="bicubic", normalize=True, max_update=None,
                 max_batch=4096, sync_stats=True, global_bank=False, lookahead=False):
        if interp not in INTERP:
            raise ValueError(f"interp must be one of {sorted(INTERP)}")
        self.n_memory, self.nv = int(n_memory), int(pushes_per_clip)
        self.log_mixup, self.vscale, self.interp = bool(log_mixup), tuple(virtual_crop_scale), interp
        self.normalize, self.max_update = bool(normalize), max_update
        if normalize and (max_update is None or max_update < 1):
            raise ValueError("RunningNorm needs max_update >= 1 (epoch_samples * max_update_epochs)")
        self.max_batch, self.sync_stats = int(max_batch), sync_stats
        self.history = -(-self.n_memory // self.nv)          # distinct clips the FIFO can still hold
        # lookahead (UpstreamInputPipeline(lookahead=True)): the front end of batch k+1 runs before the views of batch k,
        # so the ring holds one more batch and the write head runs ``ahead`` clips in front of ``clips_done``
        self.lookahead = bool(lookahead)
        self.ahead = 0
        # how many batches the views run behind the front end: 1 with lookahead; 2 when the statistics are exchanged
        # between GPUs as well (the scan of batch k-1 then runs BESIDE the views of batch k-2, off the critical path)
        self.depth = 0 if not self.lookahead else (2 if (self.normalize and sync_stats and _dist_world()[1] > 1) else 1)
        self.cap = self.history + (1 + self.depth) * self.max_batch
        # global bank (optional): the FIFO runs over the GLOBAL clip order; bank entries are read from the
        # peers' rings over NVLink.  Every step must bring exactly max_batch clips per rank.  The ring is
        # steps_back + 2 steps deep: a step's clips can be picked for steps_back more steps, and a rank is never
        # more than one statistics exchange ahead of the slowest reader.
        self.global_bank = bool(global_bank) and _dist_world()[1] > 1
        if self.global_bank and self.lookahead:
            raise ValueError("lookahead is not available with global_bank")
        if self.global_bank:
            if not (normalize and sync_stats):
                raise ValueError("global_bank needs the synchronised running statistics (normalize and sync_stats)")
            self.rank, self.world = _dist_world()
            gb = self.world * self.max_batch
            self.steps_back = -(-self.history // gb) if self.history else 0
            self.cap = (self.steps_back + 2) * self.max_batch
            self.gcap = (self.steps_back + 2) * gb
            self.peer_ring = None
        self.clips_done = 0
        self.ring = self.stats = self.state = None
        self.f = self.t = self.pitch = None
        self._exchange = False               # False: not tried yet; None: unavailable; else a StatsExchange

    # ------------------------------------------------------------------ buffers
    def ensure(self, f, t, device):
        if self.ring is None:
            self.f, self.t = int(f), int(t)
            # ring rows are padded to a multiple of 4 floats: 16-byte aligned rows, read with 128-bit loads by the two-view
            # kernel (assl_view_cfg.ring_pitch); the pad columns are never read as data
            self.pitch = (self.t + 3) & ~3
            if self.global_bank:
                self.peer_ring = PeerRing((self.cap, self.f, self.pitch), device)
                self.ring = self.peer_ring.tensor
                self.stats = torch.zeros((self.gcap, 2), dtype=torch.float32, device=device)   # over GLOBAL clips
            else:
                self.ring = torch.zeros((self.cap, self.f, self.pitch), dtype=torch.float32, device=device)
                self.stats = torch.zeros((self.cap, 2), dtype=torch.float32, device=device)
            self.stats[:, 1] = 1.0
            self.state = torch.zeros(8, dtype=torch.float64, device=device)
        elif (self.f, self.t) != (int(f), int(t)):
            raise ValueError(f"clip shape changed from {(self.f, self.t)} to {(f, t)}: the mixup bank holds one shape")

    @property
    def head(self):
        """Global index of the next clip the front end writes."""
        return self.clips_done + self.ahead

    @property
    def slot0(self):
        return self.head % self.cap

    def ring_put(self, x):
        """Copy a [B, F, T] batch into the ring at the current head (at most two slice copies)."""
        b, s0 = x.shape[0], self.slot0
        first = min(b, self.cap - s0)
        self.ring[s0:s0 + first, :, :self.t].copy_(x[:first])
        if first < b:
            self.ring[:b - first, :, :self.t].copy_(x[first:])

    # ------------------------------------------------------------------ statistics
    def exchange(self, device):
        """The statistics exchange of a clip-sharded run (opened on first use), or None for one process / sync_stats=False."""
        world = _dist_world()[1]
        if world > 1 and self.sync_stats and self._exchange is False:
            self._exchange = StatsExchange.open(self.max_batch, device)
        return self._exchange if (world > 1 and self.sync_stats) else None

    @property
    def split_stats(self):
        """Lookahead order across GPUs -- log-mel(k) . scan(k-1)+send(k) . views(k-2): the sums of batch k are SENT right
        behind its log-mel kernel and SCANNED one step later, beside the views of the batch before (launched deferred),
        so neither the wait for the slowest rank nor the scan itself sits between two log-mel kernels."""
        return self.depth == 2

    def update_stats(self, sums, n_parts):
        """sums: [B, n_parts, 2] fp64 on device for this rank's clips, in order: statistics of the batch at the ring head."""
        lib = _native.lib()
        b = sums.shape[0]
        elems = self.f * self.t
        ex = self.exchange(sums.device)
        if ex is not None and b > ex.max_clips:
            raise ValueError(f"{b} clips in one call, the statistics exchange was opened for max_batch = {ex.max_clips}")
        if self.global_bank:
            if ex is None or ex.staged or b != self.max_batch:
                raise RuntimeError("global_bank needs the peer-memory exchange and exactly max_batch clips per rank and step")
            ex.run(self.state, sums, b, n_parts, elems, self.max_update, self.stats,
                   (self.clips_done * ex.world) % self.gcap, self.gcap, write_all=True)
        elif ex is not None:
            ex.run(self.state, sums, b, n_parts, elems, self.max_update, self.stats, self.slot0, self.cap)
        else:
            _native.check(lib.assl_running_stats(_ptr(self.state), _ptr(sums), b, n_parts, elems,
                                                 int(self.max_update), 0, b, _ptr(self.stats),
                                                 self.slot0, self.cap, _stream_ptr()), "assl_running_stats")
        self._keep = sums        # (a deferred launch may still run beside the kernel that reads them)

    def send_stats(self, sums, n_parts):
        """split_stats: this rank's sums of the batch at the ring head go out; ``scan_pending_stats`` completes the step."""
        ex = self.exchange(sums.device)
        b = sums.shape[0]
        if b > ex.max_clips:
            raise ValueError(f"{b} clips in one call, the statistics exchange was opened for max_batch = {ex.max_clips}")
        ex.send(sums, b, n_parts)
        self._sent = (self.slot0, b)
        self._keep2, self._keep = getattr(self, "_keep", None), sums

    def scan_then_send_stats(self, sums, n_parts):
        """split_stats, one step: the statistics of the batch sent a step ago are scanned (deferred: nobody waits for the
        slowest rank, its records have long landed) and this batch's sums go out -- one launch when a batch is pending."""
        if getattr(self, "_sent", None) is None:
            return self.send_stats(sums, n_parts)
        ex = self.exchange(sums.device)
        b = sums.shape[0]
        if b > ex.max_clips:
            raise ValueError(f"{b} clips in one call, the statistics exchange was opened for max_batch = {ex.max_clips}")
        ex.scan_send(self.state, self.f * self.t, self.max_update, self.stats, self._sent[0], self.cap, sums, b, n_parts)
        self._sent = (self.slot0, b)
        self._keep2, self._keep = getattr(self, "_keep", None), sums

    def scan_pending_stats(self, deferred=False):
        """split_stats: the scan of the batch sent last (no-op when none is pending)."""
        if getattr(self, "_sent", None) is None:
            return False
        slot0, _ = self._sent
        self._exchange.scan(self.state, self.f * self.t, self.max_update, self.stats, slot0, self.cap, deferred=deferred)
        self._sent = None
        return True

    def check_exchange(self):
        if self._exchange not in (False, None):
            self._exchange.check()

    def moments(self, x):
        sums = torch.empty((x.shape[0], 1, 2), dtype=torch.float64, device=x.device)
        _native.check(_native.lib().assl_clip_moments_f32(_ptr(x), x.shape[0], x[0].numel(), _ptr(sums), _stream_ptr()))
        return sums

    # ------------------------------------------------------------------ views
    def launch_views(self, params: ViewParams, batch: int, device):
        """Two views of the ``batch`` clips at the ring head, then advance the head."""
        if self.global_bank:
            ch, cw = canvas_size(self.f, self.t, self.vscale)
            validate_crops(params.crop, ch, cw)
            return self._launch_views_global(params, batch, device, ch, cw)
        return self.launch_prepared(self.prepare_views(params, batch, device, self.clips_done))

    def prepare_views(self, params: ViewParams, batch: int, device, g0: int):
        """Validate and upload the parameters of the ``batch`` clips whose first global index is ``g0``."""
        ch, cw = canvas_size(self.f, self.t, self.vscale)
        validate_crops(params.crop, ch, cw)
        src = np.asarray(params.src_clip)
        clip = (g0 + np.arange(batch, dtype=np.int64)).reshape((-1,) + (1,) * (src.ndim - 1))
        # the FIFO, not the ring, bounds a pick: the last ``history`` clips (and, for view 2, the clip itself)
        if np.any(src > clip) or np.any((src >= 0) & (src < clip - self.history)):
            raise ValueError("mixup source outside the bank (the FIFO holds the last "
                             f"{self.history} clips)")
        mix_slot = np.where(src >= 0, src % self.cap, -1).astype(np.int32)
        clip_slot = ((g0 + np.arange(batch, dtype=np.int64)) % self.cap).astype(np.int32)
        to_dev = lambda a: torch.from_numpy(a).to(device, non_blocking=True)   # noqa: E731
        return dict(batch=batch, device=device, g0=g0, ch=ch, cw=cw, clip=to_dev(clip_slot), mix=to_dev(mix_slot),
                    alpha=to_dev(np.ascontiguousarray(params.alpha, np.float32)),
                    crop=to_dev(np.ascontiguousarray(params.crop, np.int32)),
                    order=to_dev(heavy_first(params.crop, self.t, cw)))

    def launch_prepared(self, prep, deferred=False):
        """Launch the two-view kernel for a prepared batch (the oldest one not yet turned into views).  ``deferred``:
        the launch may run beside the kernel launched before it (assl_two_views_deferred_f32, include/assl.h)."""
        if prep["g0"] != self.clips_done:
            raise RuntimeError("views must be launched in clip order")
        batch, device = prep["batch"], prep["device"]
        v1 = torch.empty((batch, 1, self.f, self.t), dtype=torch.float32, device=device)
        v2 = torch.empty_like(v1)
        vcfg = _native.ViewCfg(self.f, self.t, prep["ch"], prep["cw"], INTERP[self.interp], 1 if self.log_mixup else 0,
                               self.pitch)
        name = "assl_two_views_deferred_f32" if deferred else "assl_two_views_ordered_f32"
        _native.check(getattr(_native.lib(), name)(
            C.byref(vcfg), _ptr(self.ring), _ptr(self.stats if self.normalize else None), self.cap,
            _ptr(prep["clip"]), _ptr(prep["mix"]), _ptr(prep["alpha"]), _ptr(prep["crop"]), _ptr(prep["order"]), batch,
            _ptr(v1), _ptr(v2), _stream_ptr()), name)
        self.clips_done += batch
        return v1, v2

    def global_indices(self, src, step):
        """Global-bank bookkeeping of one step: ``src`` [B,2] GLOBAL clip ids of the bank picks (-1 = none) of this
        rank's clips -> int32 arrays (clip_slot [B], clip_stat [B], mix_slot [B,2], mix_rank [B,2], mix_stat [B,2]).
        Global clip G of step s = G // (world*B) sits on rank (G % (world*B)) // B, local clip s*B + G % B."""
        bsz, gb = self.max_batch, self.world * self.max_batch
        local0 = step * bsz
        mine = step * gb + self.rank * bsz + np.arange(bsz, dtype=np.int64)          # global ids of this rank's clips
        # (view 2 may pick the clip itself: the FIFO already holds view 1's push of it)
        if np.any(src > mine[:, None]) or np.any((src >= 0) & (src < mine[:, None] - self.history)):
            raise ValueError("mixup source outside the global bank")
        has = src >= 0
        s_step, s_in = np.divmod(np.where(has, src, 0), gb)
        s_rank, s_b = np.divmod(s_in, bsz)
        if np.any(has & (s_step < step - self.steps_back)):
            raise ValueError("mixup source older than the ring keeps")
        i32 = lambda a: np.ascontiguousarray(a, np.int32)   # noqa: E731
        return (i32((local0 + np.arange(bsz, dtype=np.int64)) % self.cap), i32(mine % self.gcap),
                i32(np.where(has, (s_step * bsz + s_b) % self.cap, -1)), i32(np.where(has, s_rank, 0)),
                i32(np.where(has, src % self.gcap, 0)))

    def _launch_views_global(self, params, batch, device, ch, cw):
        """``params.src_clip`` holds GLOBAL clip ids (the sampler drew over the global batch); this rank's slice."""
        if batch != self.max_batch:
            raise RuntimeError("global_bank needs exactly max_batch clips per rank and step")
        idx = self.global_indices(np.asarray(params.src_clip, np.int64), self.clips_done // self.max_batch)
        to_dev = lambda a: torch.from_numpy(a).to(device, non_blocking=True)   # noqa: E731
        d_clip, d_cstat, d_mix, d_rank, d_mstat = (to_dev(a) for a in idx)
        w = self.world
        d_alpha = to_dev(np.ascontiguousarray(params.alpha, np.float32))
        d_crop = to_dev(np.ascontiguousarray(params.crop, np.int32))
        d_order = to_dev(heavy_first(params.crop, self.t, cw))
        v1 = torch.empty((batch, 1, self.f, self.t), dtype=torch.float32, device=device)
        v2 = torch.empty_like(v1)
        vcfg = _native.ViewCfg(self.f, self.t, ch, cw, INTERP[self.interp], 1 if self.log_mixup else 0, self.pitch)
        _native.check(_native.lib().assl_two_views_global_f32(
            C.byref(vcfg), _ptr(self.ring), self.peer_ring.ptrs, w, _ptr(self.stats), _ptr(d_clip), _ptr(d_cstat),
            _ptr(d_mix), _ptr(d_rank), _ptr(d_mstat), _ptr(d_alpha), _ptr(d_crop), _ptr(d_order), batch, _ptr(v1), _ptr(v2),
            _stream_ptr()), "assl_two_views_global_f32")
        self.clips_done += batch
        return v1, v2

    # ------------------------------------------------------------------ checkpointing (the reference saves none of this)
    def state_dict(self):
        if self.ahead:
            raise RuntimeError("a batch is still waiting for its views: call flush() before state_dict()")
        self.check_exchange()
        return {"clips_done": self.clips_done, "cap": self.cap,
                "ring": None if self.ring is None else self.ring[:, :, :self.t].cpu(),
                "stats": None if self.stats is None else self.stats.cpu(),
                "state": None if self.state is None else self.state.cpu()}

    def load_state_dict(self, sd, device=None):
        if sd["ring"] is not None:
            if sd["cap"] != self.cap:
                raise ValueError("ring capacity differs (n_memory / max_batch changed)")
            device = device or torch.device("cuda", torch.cuda.current_device())
            self.ensure(sd["ring"].shape[1], sd["ring"].shape[2], device)
            self.ring[:, :, :self.t].copy_(sd["ring"]); self.stats.copy_(sd["stats"]); self.state.copy_(sd["state"])
        self.clips_done = int(sd["clips_done"])
