#!/usr/bin/env python3
"""Benchmark of the fused log-mel + two-view augmentation path (BASELINE.json metric: clips/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path over one batch of synthetic waveforms per GPU.  Headline workload (``value``,
``config.workload``): BASELINE config 2 -- n_fft 1024, hop 160, 64 mels, 60-7800 Hz, 1 s 16 kHz clips, batch 1024 per GPU
(weak scaling).  ``value`` is measured with inputs resident in HBM (CUDA events, max over ranks); ``e2e`` goes through
the C-ABI host entry point with pinned HOST buffers, H2D and D2H inside the timed region; ``roofline`` is for the
dominant kernel (log-mel) from its live CUDA-event time; ``cpu_baseline`` times the CPU oracle port on this box's host
cores on a bounded sample.  Behind the headline the same step loop is timed on BASELINE configs 3, 4 and three points of
config 5 (``configs``: GLOBAL batches sharded by clip over the ranks), and with more than one rank the run first checks,
through the product API, that the sharded pipeline equals one process on the unsharded batch
(``sharded_equals_unsharded``; exit code 3 if it does not).
"""
import argparse
import ctypes as C
import hashlib
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

CFG = {"pretrain": {"normalization": "mean_var", "augmentations": {
    "MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
    "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                         "time_crop_scale": [0.6, 1.5]}}}}
METRIC = "log-mel+2-view aug clips/sec"
UNIT = "clips/s"
N_MELS, HOP = 64, 160


def algorithmic_bytes(n_samples, frontend_only=False):
    """4L (waveform) + 4FT (log-mel, written once) [+ 8FT bank reads + 8FT two views]."""
    t = 1 + n_samples // HOP
    return 4 * n_samples + (4 if frontend_only else 20) * N_MELS * t


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
            return float(json.load(fh)["hbm_gbs"]), "MEASURED_PEAKS.json"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def kernel_source_hash():
    """sha256 over the log-mel kernel's sources: a DRAM-traffic figure captured under ncu belongs to one version of them."""
    h = hashlib.sha256()
    for f in ("logmel.cu", "common.cuh", "f32x2.cuh", "fft32p_gen.cuh"):
        with open(os.path.join(ROOT, "audio-ssl_b200", "csrc", f), "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def recorded_traffic():
    """profiles/traffic.json: dram bytes per launch of logmel_kernel from one `ncu --set full` capture of this very
    workload -- used only while the kernel's sources still hash to what the capture was taken from."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            t = json.load(fh)
        if t.get("source_sha") == kernel_source_hash():
            return t.get("logmel_kernel_bytes_per_launch"), t.get("capture")
        return None, "profiles/traffic.json is from another version of the kernel (source hash differs): not reported"
    except Exception:
        return None, "no capture on file"


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle port (numpy librosa restatement + restated augmentations), one clip at a time
# per worker process -- the reference's own execution model (DataLoader workers, upstream_dataset.py:118-125).
def _cpu_worker(args):
    """One DataLoader-style worker: its own augmentation state, one clip at a time, for ``budget_s`` seconds."""
    seed, budget_s, n_samples = args
    os.environ["OMP_NUM_THREADS"] = os.environ["OPENBLAS_NUM_THREADS"] = os.environ["MKL_NUM_THREADS"] = "1"
    import random

    import torch
    torch.set_num_threads(1)
    from oracle import frontend as ofe
    from oracle import views as ov
    rng = np.random.default_rng(seed)
    waves = (rng.standard_normal((32, n_samples)) * 0.1).astype(np.float32)
    np.random.seed(seed); random.seed(seed)
    mfb = ofe.mel_filterbank()
    tf = ov.TwoViewOracle(10 ** 6)
    done, t0 = 0, time.perf_counter()
    while time.perf_counter() - t0 < budget_s:
        for w in waves[:8]:
            tf(torch.from_numpy(ofe.log_mel(waves[done % 32], mfb=mfb)).unsqueeze(0))
            done += 1
    return done


class CpuArm:
    def __init__(self, n_samples, cores=None):
        import multiprocessing as mp
        self.n_samples = n_samples
        self.cores = cores or max(1, min(os.cpu_count() or 1, 64))
        os.environ.setdefault("OMP_NUM_THREADS", "1")
        os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
        self.pool = mp.get_context("spawn").Pool(self.cores)
        self.pool.map(_cpu_worker, [(0, 0.05, n_samples)] * self.cores)      # import + warm every worker

    def run(self, budget_s, seed=1):
        """All workers run for ``budget_s`` seconds; returns (clips/s, clips, wall seconds)."""
        t0 = time.perf_counter()
        done = self.pool.map(_cpu_worker, [(seed * 1000 + i, budget_s, self.n_samples) for i in range(self.cores)])
        dt = time.perf_counter() - t0
        return sum(done) / dt, sum(done), dt

    def close(self):
        self.pool.close(); self.pool.join()


def run_reference_arm(args):
    """--impl reference: the reference's CPU implementation of the path (oracle port: librosa is not
    installable here) on all host cores; each step is a time-bounded sample of the workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n_samples = int(args.seconds * 16000)
    arm = CpuArm(n_samples)
    step_s = min(args.cpu_seconds, 120.0 / max(1, args.steps + args.warmup))      # whole run stays within minutes
    for _ in range(args.warmup):
        arm.run(min(step_s, 1.0))
    clips, wall = 0, 0.0
    for s in range(args.steps):
        _, c, dt = arm.run(step_s, seed=s + 1)
        clips += c; wall += dt
    arm.close()
    val = clips / wall
    sample = f"{clips} clips in {args.steps} steps of {step_s:.1f} s, {args.seconds:g} s 16 kHz noise clips"
    _emit(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (fp64 FFT)", "data": "synthetic",
        "config": workload_config(args, per_gpu_batch=args.batch),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": arm.cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def workload_config(args, per_gpu_batch):
    """The same dict in both arms (the driver compares them)."""
    return {"workload": f"C2: log-mel (n_fft 1024, hop 160, 64 mels, 60-7800 Hz) + BYOL-A 2-view aug, "
                        f"{args.seconds:g} s 16 kHz clips, batch {per_gpu_batch} per GPU",
            "clip_seconds": args.seconds, "per_gpu_batch": per_gpu_batch, "global_batch": per_gpu_batch * args.gpus,
            "interp": "bicubic", "parallelism": f"clip-sharded x{args.gpus}",
            "l2": f"rotating {args.rotate} input batches + ring/views footprint > 126 MB L2"}


# ----------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled every 20 ms; ``mark()`` brackets the loaded periods and only
    samples taken inside one are reported (B200_PROFILING.md: clocks DURING the timed region)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc, self.loaded = index, [], None, []

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))
        except Exception:
            pass

    def mark(self, t0, t1):
        self.loaded.append((t0, t1))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        # a sample describes the ~20 ms before it arrived: keep those that arrived inside a loaded period
        rows = [r for t, r in self.rows if any(a + 0.03 <= t <= b + 0.005 for a, b in self.loaded)]
        num = lambda v: float(v) if v.replace(".", "", 1).isdigit() else None   # noqa: E731
        sm = [num(r[0]) for r in rows if r and num(r[0]) is not None]
        mx = [num(r[1]) for r in self.rows_all() if len(r) > 1 and num(r[1]) is not None]
        pw = [num(r[2]) for r in rows if len(r) > 2 and num(r[2]) is not None]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_mhz_min": min(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "power_w_max": max(pw) if pw else None,
                "reasons": reasons, "samples": len(sm), "samples_total": len(self.rows),
                "loaded_seconds": round(sum(b - a for a, b in self.loaded), 3),
                "how": "nvidia-smi -lms 20; only samples that arrived while the step loop was running on the GPU"}

    def rows_all(self):
        return [r for _, r in self.rows]


# ----------------------------------------------------------------------------------------------
class Workload:
    """Device-resident step loop for one (per-GPU batch, clip length): the C ABI driven through ctypes with raw pointers
    taken once (~5 us per call, so the host stays ahead of the GPU).  Step order with lookahead, one GPU:
    log-mel(k) . statistics(k) . views(k-1); across GPUs: log-mel(k) . scan(k-1)+send(k) (one launch) . views(k-2)."""

    def __init__(self, A, lib, dev, rank, world, batch, n_samples, exchange_factory, lookahead=True, gbank=False,
                 rotate_bytes=256 << 20, rotate_min=2, seed=1234):
        import torch
        from audiossl_b200 import _native
        from audiossl_b200.engine import heavy_first
        self.torch, self.lib, self.dev, self.rank, self.world = torch, lib, dev, rank, world
        self.B, self.L = batch, n_samples
        self.look, self.gbank = lookahead and not gbank, gbank
        self.plan = A.Plan()
        self.T, self.tiles = self.plan.num_frames(n_samples), self.plan.num_tiles(n_samples)
        B, L, T = self.B, self.L, self.T
        self.tf = A.AugmentationModule(CFG, len_of_files=10 ** 7, sampler="fast", max_batch=B,
                                       seed=seed + (0 if gbank else rank), global_bank=gbank, lookahead=self.look)
        eng = self.eng = self.tf.engine
        eng.ensure(N_MELS, T, dev)
        eng.ring.zero_()
        self.cap = eng.cap
        self.rotate = max(rotate_min, min(16, -(-rotate_bytes // (B * L * 4))))
        gen = torch.Generator(device=dev).manual_seed(seed + 2 + rank)          # seed 1234 + config id
        self.waves = [torch.randn(B, L, generator=gen, device=dev) * 0.1 for _ in range(self.rotate)]
        self.v1 = torch.empty((B, N_MELS, T), dtype=torch.float32, device=dev)
        self.v2 = torch.empty_like(self.v1)
        self.sums = [torch.empty((B, self.tiles, 2), dtype=torch.float64, device=dev) for _ in range(2)]
        ch, cw = int(N_MELS * 1.0), int(T * 1.5)
        self.vcfg = _native.ViewCfg(N_MELS, T, ch, cw, 0, 1, eng.pitch)     # ring rows padded to 16 bytes
        self.pitch = eng.pitch
        self.max_update = int(eng.max_update)
        self.exchange = exchange_factory(B) if world > 1 else None
        self.split = self.look and self.exchange is not None
        self.scan_deferred = int(os.environ.get("ASSL_BENCH_SCAN_DEFERRED", "1"))
        self.merged = self.split and not self.exchange.staged and os.environ.get("ASSL_BENCH_SCAN_SEND", "1") != "0"
        # the log-mel kernel of the next batch reads and transforms its first frame pairs under the tail of the views launched
        # before it and waits for them only before its first store (assl_logmel_deferred_f32, inputs_ready=True in the module
        #  -- the waveforms are resident and complete long before any step)
        self.front_overlap = self.look and os.environ.get("ASSL_FRONT_OVERLAP", "1") != "0"
        self.logmel = lib.assl_logmel_deferred_f32 if self.front_overlap else lib.assl_logmel_pitched_f32
        # parameter sets: the ring position of step s repeats with period cap / gcd(B, cap); the draws are those of steps far
        # into a run (bank full: every view mixes), so a set fits every step with the same ring position
        gsz = B * world if gbank else B
        period = self.cap // math.gcd(B, self.cap)
        if gbank:
            period = (eng.gcap // gsz) * period // math.gcd(eng.gcap // gsz, period)
        self.n_sets = period * max(1, -(-6 // period))
        s0 = self.n_sets * (-(-(4096 // B + 4) // self.n_sets))
        self.sets = []
        for j in range(self.n_sets):
            s = s0 + j
            if gbank:
                p = self.tf.sampler.draw(s * gsz, gsz, N_MELS, T)
                lo, hi = rank * B, (rank + 1) * B
                idx = eng.global_indices(np.asarray(p.src_clip[lo:hi], np.int64), s)
                arrs = list(idx) + [np.ascontiguousarray(p.alpha[lo:hi]), np.ascontiguousarray(p.crop[lo:hi]),
                                    heavy_first(p.crop[lo:hi], T, cw)]
            else:
                p = self.tf.sampler.draw(s * B, B, N_MELS, T)
                mix = np.where(p.src_clip >= 0, p.src_clip % self.cap, -1).astype(np.int32)
                clip = ((s * B + np.arange(B, dtype=np.int64)) % self.cap).astype(np.int32)
                arrs = [clip, mix, np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop), heavy_first(p.crop, T, cw)]
            self.sets.append([torch.from_numpy(a).to(dev) for a in arrs])
        ptr = lambda t: int(t.data_ptr())  # noqa: E731
        self.st = int(torch.cuda.current_stream().cuda_stream)
        self.p_waves = [ptr(w) for w in self.waves]
        self.p_ring, self.p_stats, self.p_state = ptr(eng.ring), ptr(eng.stats), ptr(eng.state)
        self.p_v1, self.p_v2 = ptr(self.v1), ptr(self.v2)
        self.p_sums = [ptr(t) for t in self.sums]
        self.p_sets = [[ptr(t) for t in sp] for sp in self.sets]
        self.vref = C.byref(self.vcfg)
        self.elems = N_MELS * T
        self.step_no = 0
        self.check = _native.check

    @property
    def launches_per_step(self):
        return 4 if (self.split and not self.merged) else 3

    def step(self, events=None):
        """One step; ``events`` = 4 CUDA events recorded around the three stages (serialises the dependent launches)."""
        lib, chk, st, B, L, cap = self.lib, self.check, self.st, self.B, self.L, self.cap
        s = self.step_no
        self.step_no += 1
        slot0 = (s * B) % cap
        sums = self.p_sums[s & 1]
        rec = (lambda i: events[i].record()) if events is not None else (lambda i: None)
        rec(0)
        chk(self.logmel(self.plan.handle, self.p_waves[s % self.rotate], None, B, L, L, self.p_ring, slot0, cap,
                        self.pitch, sums, 0, st))
        rec(1)
        if self.split:
            if s > 0 and self.merged:  # statistics of the PREVIOUS batch (every rank sent its sums a step ago) + this batch's send
                chk(lib.assl_exchange_scan_send(self.exchange.handle, self.p_state, self.elems, self.max_update, self.p_stats,
                                                ((s - 1) * B) % cap, cap, 0, sums, B, self.tiles, st))
            else:
                if s > 0:
                    chk(lib.assl_running_stats_exchange_scan(self.exchange.handle, self.p_state, self.elems, self.max_update,
                                                             self.p_stats, ((s - 1) * B) % cap, cap, 0, self.scan_deferred, st))
                chk(lib.assl_exchange_send(self.exchange.handle, sums, B, self.tiles, st))
                if self.exchange.staged:
                    self.exchange._gather()
        elif self.gbank:
            self.exchange.run(self.eng.state, self.sums[s & 1], B, self.tiles, self.elems, self.max_update, self.eng.stats,
                              (s * B * self.world) % self.eng.gcap, self.eng.gcap, write_all=True)
        elif self.exchange is not None:
            self.exchange.run(self.eng.state, self.sums[s & 1], B, self.tiles, self.elems, self.max_update, self.eng.stats,
                              slot0, cap)
        else:
            chk(lib.assl_running_stats(self.p_state, sums, B, self.tiles, self.elems, self.max_update, 0, B, self.p_stats,
                                       slot0, cap, st))
        rec(2)
        if self.look:                  # views of an EARLIER batch (one back; two across GPUs), beside this batch's statistics
            back = 2 if self.split else 1
            if s >= back:
                clip, mix, alpha, crop, order = self.p_sets[(s - back) % self.n_sets]
                chk(lib.assl_two_views_deferred_f32(self.vref, self.p_ring, self.p_stats, cap, clip, mix, alpha, crop, order,
                                                    B, self.p_v1, self.p_v2, st))
        elif self.gbank:
            clip, cstat, mix, mrank, mstat, alpha, crop, order = self.p_sets[s % self.n_sets]
            chk(lib.assl_two_views_global_f32(self.vref, self.p_ring, self.eng.peer_ring.ptrs, self.world, self.p_stats, clip,
                                              cstat, mix, mrank, mstat, alpha, crop, order, B, self.p_v1, self.p_v2, st))
        else:
            clip, mix, alpha, crop, order = self.p_sets[s % self.n_sets]
            chk(lib.assl_two_views_ordered_f32(self.vref, self.p_ring, self.p_stats, cap, clip, mix, alpha, crop, order, B,
                                               self.p_v1, self.p_v2, st))
        rec(3)

    def warm(self, min_steps):
        """At least ``min_steps`` untimed steps, and enough to fill the ring (every bank pick then reads a real log-mel)."""
        for _ in range(max(min_steps, -(-self.cap // self.B) + 1)):
            self.step()

    def timed(self, steps, fence):
        """``steps`` steps between two events, no event inside; returns (ms total, host enqueue ms per step)."""
        torch = self.torch
        fence()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        h0 = time.perf_counter()
        for _ in range(steps):
            self.step()
        host_ms = 1e3 * (time.perf_counter() - h0) / steps
        t1.record()
        fence()
        return t0.elapsed_time(t1), host_ms

    def close(self):
        if self.exchange is not None:
            self.exchange.check()
            self.exchange.close()


def sharded_check(A, dev, rank, world):
    """Through the product API, on small batches: (1) the sharded lookahead pipeline (rank-local banks switched off, i.e.
    mixup off; running statistics exchanged over peer memory, send/scan split) == ONE process on the unsharded batch in
    the ordinary order; (2) with the global mixup bank (bank entries read from the peers' rings over NVLink) and mixup
    ON the sharded run == one process on the unsharded batch.  Bit for bit (torch.equal); returns (ok, max |diff|, note)."""
    import torch
    import torch.distributed as dist
    b, L, steps = 64, 8000, 6
    g = torch.Generator(device="cpu").manual_seed(77)
    crop_only = {"pretrain": {"normalization": "mean_var", "augmentations": {
        "RandomResizeCrop": CFG["pretrain"]["augmentations"]["RandomResizeCrop"]}}}
    T = A.Plan().num_frames(L)
    take = lambda p, lo, hi: A.ViewParams(p.alpha[lo:hi], p.src_clip[lo:hi], p.bank_index[lo:hi], p.crop[lo:hi])   # noqa: E731
    worst, ok = 0.0, True

    def compare(got, want):
        nonlocal worst, ok
        for a, w in zip(got, want):
            w = w[rank * b:(rank + 1) * b]
            if not torch.equal(a, w):
                ok = False
                worst = max(worst, float((a - w).abs().max()))

    # (1) statistics exchange, lookahead order
    sharded = A.UpstreamInputPipeline(crop_only, 10 ** 6, max_batch=b, lookahead=True)
    single = A.UpstreamInputPipeline(crop_only, 10 ** 6, max_batch=world * b, sync_stats=False)
    draws = A.FastSampler(mixup=False, seed=5)
    got, want = [], []
    for s in range(steps):
        wave = (torch.randn(world * b, L, generator=g) * 0.1).to(dev)
        p = draws.draw(s * world * b, world * b, N_MELS, T)
        got.append(sharded(wave[rank * b:(rank + 1) * b], params=take(p, rank * b, (rank + 1) * b)))
        want.append(single(wave, params=p))
    got = [o for o in got if o is not None] + sharded.drain()
    torch.cuda.synchronize()
    for a, w in zip(got, want):
        compare(a, w)
    staged = sharded.tfms.engine._exchange.staged
    transport = "staged (all-gather)" if staged else "peer memory over NVLink"
    sharded.tfms.engine.check_exchange()
    dist.barrier()
    sharded.tfms.engine._exchange.close()
    # (2) global bank, mixup on
    note = f"statistics exchange ({transport}, send/scan split, lookahead order), {steps} steps of {b} clips per rank"
    if not staged:
        sharded = A.UpstreamInputPipeline(CFG, 10 ** 6, max_batch=b, sampler="fast", seed=11, global_bank=True)
        single = A.UpstreamInputPipeline(CFG, 10 ** 6, max_batch=world * b, sampler="fast", seed=11, sync_stats=False)
        remote = 0
        for s in range(steps):
            wave = (torch.randn(world * b, L, generator=g) * 0.1).to(dev)
            a = sharded(wave[rank * b:(rank + 1) * b])
            w = single(wave)
            torch.cuda.synchronize()
            compare(a, w)
            src = sharded.tfms.last_params.src_clip
            remote += int(np.sum((src >= 0) & ((src % (world * b)) // b != rank)))
        sharded.tfms.engine.check_exchange()
        dist.barrier()
        sharded.tfms.engine._exchange.close()
        note += f"; global mixup bank with mixup on, {steps} steps, {remote} bank entries read from other GPUs by rank 0"
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    w = torch.tensor([worst], dtype=torch.float64, device=dev)
    dist.all_reduce(w, op=dist.ReduceOp.MAX)
    return bool(int(flag.item())), float(w.item()), note


def run_ours(args):
    import torch
    import torch.distributed as dist

    import audiossl_b200 as A
    from audiossl_b200 import _native
    from audiossl_b200.engine import StatsExchange

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py times the CUDA path; no GPU is visible (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _native.lib()

    def fence():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    B, L, W, K = args.batch, int(args.seconds * 16000), max(args.warmup, 3), args.steps
    gbank = bool(args.global_bank) and world > 1
    peak, peak_src = measured_peak_gbs()

    # ---- multi-GPU: the sharded run has to equal the single-stream run before anything is timed
    shard_ok, shard_diff, shard_note = None, None, None
    if world > 1 and not args.no_check:
        shard_ok, shard_diff, shard_note = sharded_check(A, dev, rank, world)

    open_exchange = lambda b: StatsExchange.open(b, dev)   # noqa: E731
    wl = Workload(A, lib, dev, rank, world, B, L, open_exchange, lookahead=bool(args.lookahead), gbank=gbank,
                  rotate_min=args.rotate)
    if gbank and (wl.exchange is None or wl.exchange.staged):
        raise RuntimeError("--global-bank needs the peer-memory exchange")
    T = wl.T
    a_fe, a_all = algorithmic_bytes(L, True), algorithmic_bytes(L)

    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    wl.warm(W)
    fence()
    # clocks and power at steady state: the same step loop for >= 1 s before the timed region
    est_ms, _ = wl.timed(K, fence)
    est_ms = max(max_over_ranks(est_ms) / K, 1e-3)
    t_a = time.perf_counter()
    wl.timed(max(K, int(args.presteady_s * 1e3 / est_ms)), fence)
    # ---- the timed region: exactly K steps, no event between the kernels
    ms_total, host_enqueue_ms = wl.timed(K, fence)
    ms_total = max_over_ranks(ms_total)
    # ---- sustained: the same loop for >= 2 s
    n_sus = max(K, int(args.sustained_s * 1e3 / max(ms_total / K, 1e-3)))
    sus_ms, _ = wl.timed(n_sus, fence)
    sus_ms = max_over_ranks(sus_ms)
    t_b = time.perf_counter()
    if sampler:
        sampler.mark(t_a, t_b)
    # ---- per-kernel durations: K further steps with an event between the kernels
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(K)]
    for e in evs:
        wl.step(events=e)
    fence()
    k_ms = [statistics.fmean(e[i].elapsed_time(e[i + 1]) for e in evs) for i in range(3)]
    exchange_kind = ("none (one GPU)" if world == 1 else
                     ("staged: all-gather of the records" if wl.exchange.staged else "peer memory over NVLink") +
                     (", send(k) behind log-mel(k); scan(k-1) one step later, beside the views of batch k-2" if wl.split
                      else ", fused send + scan"))
    launches_per_step = wl.launches_per_step
    lookahead_on = wl.look
    front_overlap = wl.front_overlap
    wl.close()
    del wl
    torch.cuda.empty_cache()

    # ---- BASELINE configs 3, 4 and three points of config 5: GLOBAL batch sharded by clip over the ranks
    configs = {}
    if not args.no_configs:
        table = [("C3", 4096, 0.95, "DeLoRes-S upstream input pipeline (length_wave 0.95 s), global batch 4096"),
                 ("C4", 2048, 10.0, "DeCAR / DeLoRes-M 10 s clips, global batch 2048, running-norm exchange"),
                 ("C5_256x30s", 256, 30.0, "sweep point: 30 s clips, global batch 256"),
                 ("C5_4096x1s", 4096, 1.0, "sweep point: 1 s clips, global batch 4096"),
                 ("C5_16384x1s", 16384, 1.0, "sweep point: 1 s clips, global batch 16384")]
        for name, gb, secs, what in table:
            if gb % world:
                continue
            b, n = gb // world, int(secs * 16000)
            w2 = Workload(A, lib, dev, rank, world, b, n, open_exchange, lookahead=True, rotate_min=2, seed=1234 + len(configs))
            w2.warm(3)
            ms, _ = w2.timed(K, fence)
            ms = max_over_ranks(ms)
            cps = gb * K / (ms * 1e-3)
            bpc = algorithmic_bytes(n)
            configs[name] = {"what": what, "global_batch": gb, "per_gpu_batch": b, "clip_seconds": secs,
                             "value": cps, "unit": UNIT, "ms_per_step": ms / K, "steps": K,
                             "pipeline_roofline": {"achieved": cps / world * bpc / 1e9, "peak": peak, "unit": "GB/s",
                                                   "frac": cps / world * bpc / 1e9 / peak, "bytes_per_clip": bpc}}
            w2.close()
            del w2
            torch.cuda.empty_cache()

    # ---- end to end through the C ABI with HOST buffers (H2D + kernels + D2H inside the timed region); across GPUs
    #      with the statistics exchange, i.e. the sharded path
    plan = A.Plan()
    cap = 1024 + B
    max_update = 2 * 10 ** 7 * 10
    ex = StatsExchange.open(B, dev) if world > 1 else None
    if ex is not None and ex.staged:
        ex.close(); ex = None
    pipe = A.HostPipeline(plan, ex)
    ws_bytes = pipe.workspace_bytes(B, L, cap)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    ws[:64].zero_()
    gen = torch.Generator(device="cpu").manual_seed(1234 + 2 + rank)
    host_waves = [(torch.randn(B, L, generator=gen) * 0.1).to(torch.float32).pin_memory() for _ in range(args.rotate)]
    hv1 = torch.empty((B, N_MELS, T), dtype=torch.float32).pin_memory()
    hv2 = torch.empty((B, N_MELS, T), dtype=torch.float32).pin_memory()
    fast = A.FastSampler(seed=99 + rank)
    vcfg = _native.ViewCfg(N_MELS, T, int(N_MELS * 1.0), int(T * 1.5), 0, 1)
    keep = []

    def host_step(s, waves=None):
        p = fast.draw(s * B, B, N_MELS, T)
        mix = np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32)
        alpha, crop = np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop)
        keep.append((mix, alpha, crop))                       # the async copies read these until the sync
        pipe.run(vcfg, ws, (waves or host_waves)[s % args.rotate], B, L, (s * B) % cap, cap, max_update, mix, alpha, crop,
                 hv1, hv2)

    for s in range(W):
        host_step(s)
    pipe.sync()
    fence()
    n_e2e = K
    t0 = time.perf_counter()
    for s in range(W, W + n_e2e):
        host_step(s)
    pipe.sync()                                                # every download has landed in host memory
    torch.cuda.synchronize()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    checksum = float(hv1[0, 0, :4].sum())
    # the same with the waveforms as 16-bit PCM in host memory (what a 16-bit audio file holds; assl_pipeline_host_pcm16
    # widens x / 32768 on the device): half the upload.  Reported beside ``e2e``, which stays the fp32-buffer number.
    host_pcm = [(w * 32768.0).round().clamp_(-32768, 32767).to(torch.int16).pin_memory() for w in host_waves]
    for s in range(W + n_e2e, 2 * W + n_e2e):
        host_step(s, host_pcm)
    pipe.sync()
    fence()
    t0 = time.perf_counter()
    for s in range(2 * W + n_e2e, 2 * W + 2 * n_e2e):
        host_step(s, host_pcm)
    pipe.sync()
    torch.cuda.synchronize()
    pcm_s = max_over_ranks(time.perf_counter() - t0)
    clocks = sampler.stop() if sampler else None
    if ex is not None:
        ex.check()
    pipe.close()

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        if shard_ok is False:
            sys.exit(3)
        return
    clips_per_s = world * B * K / (ms_total * 1e-3)
    ach = B * a_fe / (k_ms[0] * 1e-3) / 1e9
    traffic, traffic_note = recorded_traffic()
    out = {
        "metric": METRIC, "value": clips_per_s, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "host_enqueue_ms_per_step": host_enqueue_ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, B),
        "path": {"stats_exchange": exchange_kind,
                 "step_order": ("log-mel(k), scan(k-1)+send(k) in one launch (assl_exchange_scan_send), views(k-2) (deferred launch)" if (lookahead_on and world > 1) else
                                "log-mel(k), statistics(k), views(k-1): the views of the previous batch are launched deferred "
                                "(assl_two_views_deferred_f32, UpstreamInputPipeline(lookahead=True)); every step runs all the "
                                "kernels" if lookahead_on else "log-mel(k), statistics(k), views(k)"),
                 "logmel_launch": ("assl_logmel_deferred_f32 (resident inputs: the kernel transforms its first frame pairs under "
                                   "the tail of the views before it and waits for them before its first store; "
                                   "UpstreamInputPipeline(inputs_ready=True))" if front_overlap else "assl_logmel_pitched_f32"),
                 "mixup_bank": ("global FIFO, bank entries read from the peers' rings over NVLink (optional mode)" if gbank
                                else "rank-local FIFO (as each reference worker's)")},
        "sustained": {"value": world * B * n_sus / (sus_ms * 1e-3), "unit": UNIT, "steps": n_sus, "seconds": sus_ms * 1e-3,
                      "note": "the same step loop run for >= 2 s right after the timed region (steady-state clocks and power)"},
        "roofline": {"bound": "hbm", "kernel": "logmel_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                     "frac": ach / peak, "traffic": traffic, "traffic_source": traffic_note, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": B * a_fe, "kernel_ms": k_ms[0],
                     "kernel_timing": f"CUDA events around each launch over {K} further steps of the same loop, right after the "
                                      f"timed region (events between the kernels would serialise their dependent launches)"},
        "pipeline_roofline": {"achieved": clips_per_s / world * a_all / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": clips_per_s / world * a_all / 1e9 / peak, "bytes_per_clip": a_all},
        "kernel_ms": {"logmel": k_ms[0], "running_stats (across GPUs: scan of the previous batch + send)": k_ms[1],
                      "two_views": k_ms[2]},
        "e2e": {"value": world * B * n_e2e / e2e_s, "unit": UNIT, "h2d_bytes_per_step": B * L * 4 + B * 40,
                "d2h_bytes_per_step": 2 * B * N_MELS * T * 4, "ms_per_step": 1e3 * e2e_s / n_e2e, "checksum": checksum,
                "api": "assl_pipeline_host_f32 + assl_pipeline_host_sync (pinned host waveforms in, both views back to pinned "
                       "host memory; uploads/kernels/downloads of consecutive calls overlap"
                       + ("; statistics exchanged between the ranks inside every call)" if ex is not None else ")")},
        "e2e_pcm16": {"value": world * B * n_e2e / pcm_s, "unit": UNIT, "h2d_bytes_per_step": B * L * 2 + B * 40,
                      "d2h_bytes_per_step": 2 * B * N_MELS * T * 4, "ms_per_step": 1e3 * pcm_s / n_e2e,
                      "api": "assl_pipeline_host_pcm16: the same call with the waveforms as 16-bit PCM in pinned host memory "
                             "(widened to x / 32768 on the device, the values librosa.core.load gives the reference)"},
        "gpu_launches": K * launches_per_step, "clocks": clocks,
    }
    if configs:
        out["configs"] = configs
    if shard_ok is not None:
        out["sharded_equals_unsharded"] = shard_ok
        out["sharded_check"] = {"max_abs_diff": shard_diff, "what": shard_note}
    if world == 1 and not args.no_cpu:
        arm = CpuArm(L)
        val, n_done, dt = arm.run(args.cpu_seconds)
        arm.close()
        out["cpu_baseline"] = {"value": val, "unit": UNIT, "cores": arm.cores, "kind": "port",
                               "sample": f"{n_done} clips of {args.seconds:g} s noise in {dt:.1f} s, "
                                         f"one single-threaded process per core (oracle port: numpy librosa restatement + "
                                         f"restated reference augmentations)"}
    _emit(json.dumps(out))
    if world > 1:
        dist.barrier(); dist.destroy_process_group()
    if shard_ok is False:
        sys.exit(3)


_REAL_STDOUT = None


def _emit(line):
    """The one JSON line goes to the process's original stdout."""
    out = _REAL_STDOUT or sys.stdout
    out.write(line + "\n")
    out.flush()


def main():
    # stdout carries exactly one line: anything a library prints there (NCCL's version banner on the first collective)
    # is sent to stderr instead
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=1024, help="clips per GPU per step")
    ap.add_argument("--seconds", type=float, default=1.0, help="clip length")
    ap.add_argument("--rotate", type=int, default=4, help="distinct input batches cycled through (L2 hygiene)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="wall seconds of one CPU-baseline sample")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip BASELINE configs 3-5 behind the headline")
    ap.add_argument("--no-check", action="store_true", help="skip the sharded == unsharded check (N > 1)")
    ap.add_argument("--presteady-s", type=float, default=1.0, help="seconds of the step loop before the timed region")
    ap.add_argument("--sustained-s", type=float, default=2.0, help="seconds of the sustained loop")
    ap.add_argument("--lookahead", type=int, default=1,
                    help="1: software-pipelined step order log-mel(k), statistics(k), views(k-1)")
    ap.add_argument("--global-bank", action="store_true",
                    help="N > 1 only, not the headline configuration: mixup FIFO over the global clip order, bank entries read "
                         "from the peers' rings over NVLink (SURVEY 8e, optional mode)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
