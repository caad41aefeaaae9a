#!/usr/bin/env python3
"""Benchmark of the fused log-mel + two-view augmentation path (BASELINE.json metric: clips/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one pass of the hot path over one batch of synthetic waveforms per GPU (weak scaling:
per-GPU batch fixed).  Workload at N=1: BASELINE config 2 -- n_fft 1024, hop 160, 64 mels, 60-7800 Hz,
1 s 16 kHz clips, batch 1024.  ``value`` is measured with inputs resident in HBM (CUDA events, max over
ranks); ``e2e`` goes through the C-ABI host entry point with pinned HOST buffers, H2D and D2H inside the
timed region.  ``roofline`` is for the dominant kernel (log-mel) from its live CUDA-event time;
``cpu_baseline`` times the CPU oracle port on this box's host cores on a bounded sample.
"""
import argparse
import ctypes as C
import json
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
    return {"workload": f"C2: log-mel (n_fft 1024, hop 160, 64 mels, 60-7800 Hz) + BYOL-A 2-view aug, "
                        f"{args.seconds:g} s 16 kHz clips, batch {per_gpu_batch} per GPU",
            "clip_seconds": args.seconds, "per_gpu_batch": per_gpu_batch, "global_batch": per_gpu_batch * args.gpus,
            "interp": "bicubic", "parallelism": f"clip-sharded x{args.gpus}",
            "l2": f"rotating {args.rotate} input batches + ring/views footprint > 126 MB L2"}


# ----------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def run_ours(args):
    import torch
    import torch.distributed as dist

    import audiossl_b200 as A
    from audiossl_b200 import _native
    from audiossl_b200.engine import StatsExchange, gather_clip_sums, heavy_first

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py times the CUDA path; no GPU is visible (there is no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":      # keeps NCCL's banner off stdout (one JSON line)
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    lib = _native.lib()
    P = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)  # noqa: E731
    stream = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)  # noqa: E731

    B, L, W, K = args.batch, int(args.seconds * 16000), args.warmup, args.steps
    plan = A.Plan()
    T, tiles = plan.num_frames(L), plan.num_tiles(L)
    gbank = bool(args.global_bank) and world > 1
    look = bool(args.lookahead) and not gbank
    # (lookahead: the ring holds one more batch -- the log-mel kernel of batch k+1 writes while the views of batch k read)
    tf = A.AugmentationModule(CFG, len_of_files=10 ** 7, sampler="fast", max_batch=2 * B if look else B,
                              seed=1234 + (0 if gbank else rank), global_bank=gbank)
    eng = tf.engine
    eng.ensure(N_MELS, T, dev)
    cap = eng.cap
    gen = torch.Generator(device="cpu").manual_seed(1234 + 2 + rank)          # seed 1234 + config id
    host_waves = [(torch.randn(B, L, generator=gen) * 0.1).to(torch.float32).pin_memory() for _ in range(args.rotate)]
    waves = [w.to(dev) for w in host_waves]
    v1 = torch.empty((B, N_MELS, T), dtype=torch.float32, device=dev)
    v2 = torch.empty_like(v1)
    sums = torch.empty((B, tiles, 2), dtype=torch.float64, device=dev)
    local_sums = torch.empty((B, 2), dtype=torch.float64, device=dev)
    ch, cw = int(N_MELS * 1.0), int(T * 1.5)
    vcfg = _native.ViewCfg(N_MELS, T, ch, cw, 0, 1)
    max_update = int(eng.max_update)

    def upload(p, clips_done):
        mix = np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32)
        clip = ((clips_done + np.arange(B, dtype=np.int64)) % cap).astype(np.int32)
        return [torch.from_numpy(a).to(dev) for a in (clip, mix, np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop),
                                                      heavy_first(p.crop, T, int(T * 1.5)))]

    def upload_global(step):
        """Global bank: every rank draws the global batch from the same generator and keeps its slice (global clip ids)."""
        p = tf.sampler.draw(step * B * world, B * world, N_MELS, T)
        lo, hi = rank * B, (rank + 1) * B
        idx = eng.global_indices(np.asarray(p.src_clip[lo:hi], np.int64), step)
        arrs = list(idx) + [np.ascontiguousarray(p.alpha[lo:hi]), np.ascontiguousarray(p.crop[lo:hi]), heavy_first(p.crop[lo:hi], T, int(T * 1.5))]
        return [torch.from_numpy(a).to(dev) for a in arrs]

    # steps [0, W) warm up, [W, W+K) are the timed region (no event between the kernels: an event record would
    # serialise the programmatic dependent launches), [W+K, W+2K) repeat the same work with an event between the
    # kernels for the per-kernel durations (roofline.kernel_ms)
    steps_params = [upload_global(s) if gbank else upload(tf.sampler.draw(s * B, B, N_MELS, T), s * B) for s in range(W + 2 * K)]
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(W + 2 * K)]

    exchange = StatsExchange.open(B, dev) if world > 1 else None        # peer-memory path; None -> NCCL all-gather
    if gbank and exchange is None:
        raise RuntimeError("--global-bank needs the peer-memory exchange")

    class _NoEvent:
        def record(self):
            pass

    # raw pointers taken once: the timed loop only makes ctypes calls (~5 us each), so the host stays ahead of the GPU
    ptr = lambda t: int(t.data_ptr())  # noqa: E731
    st = int(torch.cuda.current_stream().cuda_stream)
    p_waves = [ptr(w) for w in waves]
    p_ring, p_sums, p_stats, p_state, p_v1, p_v2 = ptr(eng.ring), ptr(sums), ptr(eng.stats), ptr(eng.state), ptr(v1), ptr(v2)
    p_params = [[ptr(t) for t in sp] for sp in steps_params]
    vref = C.byref(vcfg)
    elems = N_MELS * T

    def device_step(s, instrument=False):
        slot0 = (s * B) % cap
        e = evs[s] if instrument else [_NoEvent()] * 4
        e[0].record()
        _native.check(lib.assl_logmel_f32(plan.handle, p_waves[s % args.rotate], None, B, L, L, p_ring, slot0, cap, p_sums, 0, st))
        e[1].record()
        if gbank:
            exchange.run(eng.state, sums, B, tiles, elems, max_update, eng.stats, (s * B * world) % eng.gcap, eng.gcap,
                         write_all=True)
        elif exchange is not None:
            exchange.run(eng.state, sums, B, tiles, elems, max_update, eng.stats, slot0, cap)
        elif world > 1:
            _native.check(lib.assl_clip_sums(p_sums, B, tiles, P(local_sums), st))
            gathered = gather_clip_sums(local_sums)
            _native.check(lib.assl_running_stats(p_state, P(gathered), world * B, 1, elems, max_update,
                                                 rank * B, B, p_stats, slot0, cap, st))
        else:
            _native.check(lib.assl_running_stats(p_state, p_sums, B, tiles, elems, max_update, 0, B, p_stats, slot0, cap, st))
        e[2].record()
        if look:                                     # views of the PREVIOUS batch, beside this batch's statistics
            if s > 0:
                clip, mix, alpha, crop, order = p_params[s - 1]
                _native.check(lib.assl_two_views_deferred_f32(vref, p_ring, p_stats, cap, clip, mix, alpha, crop, order, B,
                                                              p_v1, p_v2, st))
        elif gbank:
            clip, cstat, mix, mrank, mstat, alpha, crop, order = p_params[s]
            _native.check(lib.assl_two_views_global_f32(vref, p_ring, eng.peer_ring.ptrs, world, p_stats, clip, cstat, mix,
                                                        mrank, mstat, alpha, crop, order, B, p_v1, p_v2, st))
        else:
            clip, mix, alpha, crop, order = p_params[s]
            _native.check(lib.assl_two_views_ordered_f32(vref, p_ring, p_stats, cap, clip, mix, alpha, crop, order, B,
                                                         p_v1, p_v2, st))
        e[3].record()

    def fence():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for s in range(W):
        device_step(s)
    fence()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    fence()
    t_begin, t_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_begin.record()
    h0 = time.perf_counter()
    for s in range(W, W + K):
        device_step(s)
    host_enqueue_ms = 1e3 * (time.perf_counter() - h0) / K
    t_end.record()
    fence()
    ms_total = t_begin.elapsed_time(t_end)
    for s in range(W + K, W + 2 * K):
        device_step(s, instrument=True)
    fence()
    k_ms = [statistics.fmean(evs[s][i].elapsed_time(evs[s][i + 1]) for s in range(W + K, W + 2 * K)) for i in range(3)]
    if world > 1:
        tmax = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        ms_total = float(tmax.item())

    # ---- end to end through the C ABI with HOST buffers (H2D + kernels + D2H inside the timed region)
    ws_bytes = lib.assl_pipeline_workspace_bytes(plan.handle, B, L, cap)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    ws[:64].zero_()
    hv1 = torch.empty((B, N_MELS, T), dtype=torch.float32).pin_memory()
    hv2 = torch.empty((B, N_MELS, T), dtype=torch.float32).pin_memory()
    fast = A.FastSampler(seed=99 + rank)
    keep = []

    def host_step(s):
        p = fast.draw(s * B, B, N_MELS, T)
        mix = np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32)
        alpha, crop = np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop)
        keep.append((mix, alpha, crop))                       # the async copies read these until the sync
        _native.check(lib.assl_pipeline_host_f32(
            plan.handle, C.byref(vcfg), P(ws), ws_bytes, C.c_void_p(host_waves[s % args.rotate].data_ptr()), B, L,
            (s * B) % cap, cap, max_update, mix.ctypes.data_as(C.c_void_p), alpha.ctypes.data_as(C.c_void_p),
            crop.ctypes.data_as(C.c_void_p), C.c_void_p(hv1.data_ptr()), C.c_void_p(hv2.data_ptr()), None, None, stream()),
            "assl_pipeline_host_f32")

    def host_fence():
        _native.check(lib.assl_pipeline_host_sync(plan.handle))
        fence()

    for s in range(W):
        host_step(s)
    host_fence()
    t0 = time.perf_counter()
    for s in range(W, W + K):
        host_step(s)
    _native.check(lib.assl_pipeline_host_sync(plan.handle))      # every download has landed in host memory
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        tmax = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        e2e_s = float(tmax.item())
    clocks = sampler.stop() if sampler else None
    checksum = float(hv1[0, 0, :4].sum())

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return
    peak, peak_src = measured_peak_gbs()
    clips_per_s = world * B * K / (ms_total * 1e-3)
    a_fe, a_all = algorithmic_bytes(L, True), algorithmic_bytes(L)
    ach = B * a_fe / (k_ms[0] * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as fh:
            traffic = json.load(fh).get("logmel_kernel_bytes_per_launch")
    except Exception:
        pass
    out = {
        "metric": METRIC, "value": clips_per_s, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_total / K, "host_enqueue_ms_per_step": host_enqueue_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": dict(workload_config(args, B), stats_exchange=("none (one GPU)" if world == 1 else
                                                                 "peer memory over NVLink, fused into the statistics kernel"
                                                                 if exchange is not None else "NCCL all-gather"),
                    step_order=("log-mel(k), statistics(k), views(k-1): the views of the previous batch are launched "
                                "deferred (assl_two_views_deferred_f32, UpstreamInputPipeline(lookahead=True)); every step "
                                "runs all three kernels" if look else "log-mel(k), statistics(k), views(k)"),
                    mixup_bank=("global FIFO, bank entries read from the peers' rings over NVLink (optional mode)" if gbank
                                else "rank-local FIFO (as each reference worker's)")),
        "roofline": {"bound": "hbm", "kernel": "logmel_kernel", "achieved": ach, "peak": peak, "unit": "GB/s",
                     "frac": ach / peak, "traffic": traffic, "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": B * a_fe, "kernel_ms": k_ms[0],
                     "kernel_timing": f"CUDA events around each launch over {K} further steps of the same loop, right after the "
                                      f"timed region (events between the kernels would serialise their dependent launches)"},
        "pipeline_roofline": {"achieved": clips_per_s / world * a_all / 1e9, "peak": peak, "unit": "GB/s",
                              "frac": clips_per_s / world * a_all / 1e9 / peak, "bytes_per_clip": a_all},
        "kernel_ms": {"logmel": k_ms[0], "running_stats(+gather)": k_ms[1], "two_views": k_ms[2]},
        "e2e": {"value": world * B * K / e2e_s, "unit": UNIT, "h2d_bytes_per_step": B * L * 4 + B * 40,
                "d2h_bytes_per_step": 2 * B * N_MELS * T * 4, "ms_per_step": 1e3 * e2e_s / K, "checksum": checksum,
                "api": "assl_pipeline_host_f32 + assl_pipeline_host_sync (pinned host waveforms in, both views back to pinned host memory; uploads/kernels/downloads of consecutive calls overlap)"},
        "gpu_launches": K * (3 if (world == 1 or exchange is not None) else 4), "clocks": clocks,
    }
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
