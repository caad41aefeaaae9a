"""World-size-2 ``gloo`` test (CPU) of the multi-GPU host logic: contiguous block split, rank-major
all-gather of per-clip sums, and the closed-form statistics scan every rank then runs.  The scan is
restated here in numpy exactly as ``running_stats_kernel`` evaluates it and checked against the oracle's
sequential recurrence over the UNSHARDED clip order."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def closed_form_stats(sums, elems, n0=0, sum_m=0.0, sum_v=0.0):
    """mu_k = mean(m_1..m_k), s2_k = mean(v_1..v_k), v_j = q_j - 2 mu_j m_j + mu_j^2 (k = 0: own stats)."""
    m, q = sums[:, 0] / elems, sums[:, 1] / elems
    k = n0 + np.arange(len(m))
    cm = np.where(k >= 1, m, 0.0)
    mu = np.where(k == 0, m, (sum_m + np.cumsum(cm)) / np.maximum(k, 1))
    v = np.maximum(q - 2 * mu * m + mu * mu, 0.0)
    cv = np.where(k >= 1, v, 0.0)
    s2 = np.where(k == 0, v, (sum_v + np.cumsum(cv)) / np.maximum(k, 1))
    return mu.astype(np.float32), np.sqrt(s2).astype(np.float32)


def _worker(rank, world, port, x_all, out_q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import audiossl_b200 as A
    from audiossl_b200.engine import gather_clip_sums
    lo, hi = A.shard_bounds(x_all.shape[0], rank, world)
    mine = x_all[lo:hi].double()
    local = torch.stack([mine.sum((1, 2)), (mine * mine).sum((1, 2))], dim=1)
    gathered = gather_clip_sums(local)
    mu, std = closed_form_stats(gathered.numpy(), x_all[0].numel())
    out_q.put((rank, lo, hi, gathered.numpy(), mu[lo:hi], std[lo:hi]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_stats_equal_single_stream_oracle():
    from oracle import views as ov
    gen = torch.Generator().manual_seed(2)
    x = (torch.randn(48, 16, 20, generator=gen) * torch.linspace(0.7, 1.6, 48).view(-1, 1, 1) - 3.0).float()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, x, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    orc = ov.RunningNormOracle(10 ** 6)
    want = []
    for c in x:
        orc(c[None])
        want.append((float(orc.mean), float(orc.std)))
    want = np.asarray(want, np.float32)
    full = torch.stack([x.double().sum((1, 2)), (x.double() ** 2).sum((1, 2))], 1).numpy()
    for rank, lo, hi, gathered, mu, std in res:
        assert (lo, hi) == (rank * 24, rank * 24 + 24)
        np.testing.assert_array_equal(gathered, full)             # rank-major gather == global clip order
        np.testing.assert_allclose(mu, want[lo:hi, 0], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(std, want[lo:hi, 1], rtol=1e-4, atol=1e-5)


def test_closed_form_continues_across_calls():
    from oracle import views as ov
    gen = torch.Generator().manual_seed(3)
    x = (torch.randn(30, 4, 6, generator=gen) * 2 + 1).float()
    orc = ov.RunningNormOracle(10 ** 6)
    want = []
    for c in x:
        orc(c[None]); want.append((float(orc.mean), float(orc.std)))
    want = np.asarray(want, np.float32)
    sums = torch.stack([x.double().sum((1, 2)), (x.double() ** 2).sum((1, 2))], 1).numpy()
    mu1, sd1 = closed_form_stats(sums[:11], 24)
    m, q = sums[:11, 0] / 24, sums[:11, 1] / 24
    v = q - 2 * mu1.astype(np.float64) * m + mu1.astype(np.float64) ** 2
    mu2, sd2 = closed_form_stats(sums[11:], 24, n0=11, sum_m=m[1:].sum(), sum_v=v[1:].sum())
    np.testing.assert_allclose(np.concatenate([mu1, mu2]), want[:, 0], rtol=1e-4, atol=1e-5)
    np.testing.assert_allclose(np.concatenate([sd1, sd2]), want[:, 1], rtol=1e-4, atol=1e-5)


def _worker_global(rank, world, port, out_q):
    """Global-bank bookkeeping on the host (no GPU): every rank draws the global batch from the same generator,
    keeps its slice and maps global clip ids to (rank, ring slot, statistics slot)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import audiossl_b200 as A
    from audiossl_b200.engine import ViewEngine
    b = 96
    eng = ViewEngine(max_batch=b, max_update=10 ** 6, global_bank=True)
    smp = A.FastSampler(seed=4)
    rows = []
    for step in range(14):
        p = smp.draw(step * b * world, b * world, 64, 51)
        src = p.src_clip[rank * b:(rank + 1) * b].astype(np.int64)
        rows.append((src,) + eng.global_indices(src, step))
    out_q.put((rank, eng.cap, eng.gcap, eng.steps_back, rows))
    dist.barrier()
    dist.destroy_process_group()


def test_global_bank_index_mapping():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    world, b = 2, 96
    procs = [ctx.Process(target=_worker_global, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, cap, gcap, steps_back, rows in res:
        assert steps_back == -(-1024 // (world * b)) and cap == (steps_back + 2) * b and gcap == cap * world
        # replay: where every global clip was written (rank, slot) and which statistics slot it got
        where = {}
        for step, (src, clip_slot, clip_stat, mix_slot, mix_rank, mix_stat) in enumerate(rows):
            for r in range(world):
                for i in range(b):
                    where[step * world * b + r * b + i] = (r, (step * b + i) % cap)
            mine = step * world * b + rank * b + np.arange(b)
            assert np.array_equal(clip_slot, [where[g][1] for g in mine]) and np.array_equal(clip_stat, mine % gcap)
            for i in range(b):
                for v in range(2):
                    g = src[i, v]
                    if g < 0:
                        assert mix_slot[i, v] == -1
                        continue
                    assert mine[i] - 1024 <= g <= mine[i]
                    assert (mix_rank[i, v], mix_slot[i, v]) == where[g] and mix_stat[i, v] == g % gcap
                    # nothing has overwritten that slot since: the next clip to land there is cap local clips later
                    assert (g // (world * b)) * b + g % b + cap > step * b + b - 1
    from audiossl_b200.engine import heavy_first
    crop = np.array([[[0, 0, 40, 60], [0, 0, 64, 151]], [[0, 0, 64, 100], [0, 0, 38, 61]]], np.int32)
    assert heavy_first(crop).tolist() == [1, 2, 0, 3]      # rows x 32-column groups: 64x5, 64x4, 40x2, 38x2
    # with the canvas geometry only the columns inside the pasted clip (canvas columns 25..125) count: 101, 75, 35, 36
    assert heavy_first(crop, 101, 151).tolist() == [1, 2, 0, 3]
    assert heavy_first(np.array([[[0, 0, 64, 60], [0, 40, 64, 60]]], np.int32), 101, 151).tolist() == [0, 1]
    assert heavy_first(np.array([[[0, 0, 64, 50], [0, 40, 64, 50]]], np.int32), 101, 151).tolist() == [1, 0]


def _worker_depth(rank, world, port, out_q):
    """Ring depth and mixup-source bounds of the lookahead orders on the host (no GPU)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from audiossl_b200.engine import ViewEngine
    b = 64
    rows = []
    for kw in (dict(), dict(lookahead=True), dict(lookahead=True, sync_stats=False), dict(lookahead=True, normalize=False)):
        e = ViewEngine(max_batch=b, max_update=10 ** 6 if kw.get("normalize", True) else None, **kw)
        rows.append((e.depth, e.cap, e.split_stats))
    out_q.put((rank, rows))
    dist.barrier()
    dist.destroy_process_group()


def test_lookahead_depth_across_ranks():
    """Across ranks with synchronised statistics the views run TWO batches behind the front end (the scan of batch k-1
    runs beside the views of batch k-2) and the ring is two batches deeper; otherwise one, or none."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 33500 + os.getpid() % 2000
    world, b = 2, 64
    procs = [ctx.Process(target=_worker_depth, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, rows in res:
        assert rows == [(0, 1024 + b, False), (2, 1024 + 3 * b, True), (1, 1024 + 2 * b, False), (1, 1024 + 2 * b, False)]
    # one process: lookahead is one batch deep
    from audiossl_b200.engine import ViewEngine
    e = ViewEngine(max_batch=b, max_update=10, lookahead=True)
    assert (e.depth, e.cap, e.split_stats) == (1, 1024 + 2 * b, False)
