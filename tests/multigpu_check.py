"""Multi-GPU check of the statistics exchange (not collected by pytest; needs >= 2 GPUs on one node):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 \
        tests/multigpu_check.py

1. the fused peer-memory exchange (assl_running_stats_exchange) == the split form (assl_exchange_send, then
   assl_running_stats_exchange_scan) == the staged transport (same kernels, records moved by an all-gather) == a
   hand-made NCCL path (assl_clip_sums + all-gather + assl_running_stats) == one process scanning the whole batch,
   bit for bit, over several steps (both parities, carries, a smaller last batch);
1b. ranks that bring DIFFERENT numbers of clips to a call (uneven block split) == one process on the concatenation;
2. the sharded front end + views (mixup off, explicit crops) == one process on the unsharded batch;
3. global-bank mode (mixup ON, bank entries read from the peers' rings over NVLink) == one process on the
   unsharded batch, bit for bit, over 9 steps;
4. lookahead order across GPUs (log-mel(k) . scan(k-1) . send(k) . views(k-2)) == the ordinary order;
5. NormalizeBatch(sync=True) over the ranks == one process on the concatenated batch;
6. a rank that shows up 2.5 s late is waited for (no error, same result); with the timeout set to 0.5 s the waiting
   ranks leave their statistics untouched and ``check()`` raises.
"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import audiossl_b200 as A
    from audiossl_b200 import _native
    from audiossl_b200.engine import StatsExchange, gather_clip_sums
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    lib = _native.lib()
    P = lambda t: C.c_void_p(t.data_ptr())                      # noqa: E731
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)   # noqa: E731

    ex = StatsExchange.open(300, dev)
    assert ex is not None and not ex.staged, "peer-memory exchange unavailable on this box"
    ex_split = StatsExchange.open(300, dev)
    os.environ["ASSL_STATS_EXCHANGE"] = "nccl"
    ex_staged = StatsExchange.open(300, dev)
    del os.environ["ASSL_STATS_EXCHANGE"]
    assert ex_staged.staged
    parts, elems, cap, max_update = 7, 640, 4096, 1500
    names = "abcde"
    state = {k: torch.zeros(8, dtype=torch.float64, device=dev) for k in names}
    stats = {k: torch.zeros((cap, 2), dtype=torch.float32, device=dev) for k in names}
    g = torch.Generator(device="cpu").manual_seed(7)
    done = 0
    for step, b in enumerate([300, 300, 300, 129, 300, 1]):
        allsums = torch.rand((world, b, parts, 2), generator=g, dtype=torch.float64) * 50 + 1   # same on every rank
        mine = allsums[rank].to(dev)
        slot0 = done % cap
        ex.run(state["a"], mine, b, parts, elems, max_update, stats["a"], slot0, cap)                       # fused
        local = torch.empty((b, 2), dtype=torch.float64, device=dev)                                        # hand-made NCCL
        _native.check(lib.assl_clip_sums(P(mine), b, parts, P(local), st()))
        gathered = gather_clip_sums(local)
        _native.check(lib.assl_running_stats(P(state["b"]), P(gathered), world * b, 1, elems, max_update, rank * b, b,
                                             P(stats["b"]), slot0, cap, st()))
        whole = allsums.reshape(world * b, parts, 2).to(dev)                                                # one process
        _native.check(lib.assl_running_stats(P(state["c"]), P(whole), world * b, parts, elems, max_update, rank * b, b,
                                             P(stats["c"]), slot0, cap, st()))
        ex_split.send(mine, b, parts)                                                                      # split
        ex_split.scan(state["d"], elems, max_update, stats["d"], slot0, cap, deferred=bool(step & 1))
        ex_staged.run(state["e"], mine, b, parts, elems, max_update, stats["e"], slot0, cap)               # staged transport
        torch.cuda.synchronize()
        for k in names[1:]:
            assert torch.equal(state["a"], state[k]), (step, k, state["a"], state[k])
            assert torch.equal(stats["a"], stats[k]), (step, k)
        assert float(state["a"][7]) == 0.0
        done += b
    assert float(state["a"][0]) == min(done * world, max_update)
    for e in (ex, ex_split, ex_staged):
        e.check()

    # ---- 1b. uneven block split: rank r brings 40 + 3 r clips (and nothing at all in one step)
    sa, sb = (torch.zeros(8, dtype=torch.float64, device=dev) for _ in range(2))
    ta, tb = (torch.zeros((cap, 2), dtype=torch.float32, device=dev) for _ in range(2))
    for step in range(4):
        counts = [40 + 3 * r if (step != 2 or r != world - 1) else 0 for r in range(world)]
        every = [torch.rand((c, parts, 2), generator=g, dtype=torch.float64) * 50 + 1 for c in counts]
        mine = every[rank].to(dev) if counts[rank] else torch.zeros((1, parts, 2), dtype=torch.float64, device=dev)
        ex_split.send(mine, counts[rank], parts)
        ex_split.scan(sa, elems, max_update, ta, 100 * step, cap)
        whole = torch.cat(every).to(dev)
        first = sum(counts[:rank])
        _native.check(lib.assl_running_stats(P(sb), P(whole), sum(counts), parts, elems, max_update, first, counts[rank],
                                             P(tb), 100 * step, cap, st()))
        torch.cuda.synchronize()
        assert torch.equal(sa, sb) and torch.equal(ta, tb), ("uneven", step)
    ex_split.check()

    # ---- 1c. scan of the call sent a step ago + send of this step's sums in ONE launch (assl_exchange_scan_send), uneven
    #      and empty batches included: after every step the statistics are those of one process one step earlier
    sa, sb = (torch.zeros(8, dtype=torch.float64, device=dev) for _ in range(2))
    ta, tb = (torch.zeros((cap, 2), dtype=torch.float32, device=dev) for _ in range(2))
    snaps, pending = [], None
    for step in range(7):
        counts = [(300 if step % 3 == 0 else 25 + 11 * r) if (step != 4 or r != 0) else 0 for r in range(world)]
        every = [torch.rand((c, parts, 2), generator=g, dtype=torch.float64) * 50 + 1 for c in counts]
        mine = every[rank].to(dev) if counts[rank] else torch.zeros((1, parts, 2), dtype=torch.float64, device=dev)
        slot0 = (300 * step) % cap
        if pending is None:
            ex_split.send(mine, counts[rank], parts)
        else:
            ex_split.scan_send(sa, elems, max_update, ta, pending, cap, mine, counts[rank], parts)
            torch.cuda.synchronize()
            assert torch.equal(sa, snaps[-1][0]) and torch.equal(ta, snaps[-1][1]), ("scan_send", step)
        pending = slot0
        _native.check(lib.assl_running_stats(P(sb), P(torch.cat(every).to(dev)), sum(counts), parts, elems, max_update,
                                             sum(counts[:rank]), counts[rank], P(tb), slot0, cap, st()))
        torch.cuda.synchronize()
        snaps.append((sb.clone(), tb.clone()))
    ex_split.scan(sa, elems, max_update, ta, pending, cap)
    torch.cuda.synchronize()
    assert torch.equal(sa, sb) and torch.equal(ta, tb), "scan_send: last scan"
    ex_split.check()

    # ---- 2. sharded pipeline vs one process
    cfg = {"pretrain": {"normalization": "mean_var", "input": {"length_wave": 0.5},
                        "augmentations": {"RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                                                               "time_crop_scale": [0.6, 1.5]}}}}
    b = 96
    wave = (torch.randn(world * b, 8000, generator=g) * 0.1).to(dev)
    T = A.Plan().num_frames(8000)
    params = A.FastSampler(mixup=False, seed=3).draw(0, world * b, 64, T)
    take = lambda p, lo, hi: A.ViewParams(p.alpha[lo:hi], p.src_clip[lo:hi], p.bank_index[lo:hi], p.crop[lo:hi])   # noqa: E731
    sharded = A.UpstreamInputPipeline(cfg, 10 ** 6, max_batch=b)
    v1, v2 = sharded(wave[rank * b:(rank + 1) * b], params=take(params, rank * b, (rank + 1) * b))
    assert sharded.tfms.engine._exchange is not None
    single = A.UpstreamInputPipeline(cfg, 10 ** 6, max_batch=world * b, sync_stats=False)
    w1, w2 = single(wave, params=params)
    torch.cuda.synchronize()
    assert torch.equal(v1, w1[rank * b:(rank + 1) * b]) and torch.equal(v2, w2[rank * b:(rank + 1) * b])
    # ---- 3. global bank: mixup picks over the GLOBAL clip order, bank entries read from the peers' rings over NVLink.
    #      9 steps: the ring (steps_back + 2 steps deep) wraps, picks reach back several steps and across ranks.
    cfg_mix = {"pretrain": {"normalization": "mean_var", "input": {"length_wave": 0.5},
                            "augmentations": {"MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
                                              "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                                                                   "time_crop_scale": [0.6, 1.5]}}}}
    sharded = A.UpstreamInputPipeline(cfg_mix, 10 ** 6, max_batch=b, sampler="fast", seed=11, global_bank=True)
    single = A.UpstreamInputPipeline(cfg_mix, 10 ** 6, max_batch=world * b, sampler="fast", seed=11, sync_stats=False)
    remote = 0
    for step in range(9):
        wave = (torch.randn(world * b, 8000, generator=g) * 0.1).to(dev)
        v1, v2 = sharded(wave[rank * b:(rank + 1) * b])
        w1, w2 = single(wave)
        torch.cuda.synchronize()
        assert torch.equal(v1, w1[rank * b:(rank + 1) * b]) and torch.equal(v2, w2[rank * b:(rank + 1) * b]), step
        src = sharded.tfms.last_params.src_clip
        remote += int(np.sum((src >= 0) & ((src % (world * b)) // b != rank)))
    assert remote > 0                                             # bank entries really came from other GPUs
    # ---- 4. lookahead step order across GPUs: log-mel(k) . scan(k-1) . send(k) . views(k-2), the deferred views of batch
    #      k-2 beside the log-mel tail of batch k and the scan of batch k-1.  Sharded lookahead pipeline (rank-local banks, global statistics) == the
    #      same pipeline in the ordinary order (fused exchange).
    import random
    rng_state = (np.random.get_state(), random.getstate())
    batches = [(torch.randn(b, 8000, generator=g) * 0.1).to(dev) for _ in range(14)]          # > 1024 clips: the bank wraps
    np.random.seed(31 + rank); random.seed(31 + rank)
    plain = A.UpstreamInputPipeline(cfg_mix, 10 ** 6, max_batch=b)
    want = [plain(w) for w in batches]
    np.random.seed(31 + rank); random.seed(31 + rank)
    look = A.UpstreamInputPipeline(cfg_mix, 10 ** 6, max_batch=b, lookahead=True)
    assert look.tfms.engine.split_stats
    got = [o for o in (look(w) for w in batches) if o is not None] + look.drain()
    assert len(got) == len(want)
    torch.cuda.synchronize()
    assert look.tfms.engine._exchange is not None
    for (a1, a2), (b1, b2) in zip(got, want):
        assert torch.equal(a1, b1) and torch.equal(a2, b2)
    assert torch.equal(look.tfms.engine.state, plain.tfms.engine.state)
    np.random.set_state(rng_state[0]); random.setstate(rng_state[1])
    assert float(sharded.tfms.engine.state[7]) == 0.0
    look.tfms.engine.check_exchange()

    # ---- 5. NormalizeBatch(sync=True): global batch statistics (optional mode)
    x = torch.randn(world * 24, 1, 64, 51, generator=g).to(dev) * 3 + 1
    y = A.NormalizeBatch(sync=True)(x[rank * 24:(rank + 1) * 24])
    want_y = (x - x.mean()) / x.std().clamp(torch.finfo().eps, torch.finfo().max)
    torch.cuda.synchronize()
    assert torch.allclose(y, want_y[rank * 24:(rank + 1) * 24], rtol=1e-5, atol=1e-5)

    # ---- 6. a late rank.  (a) 2.5 s late with the default timeout: waited for, same statistics as everyone.
    dist.barrier()
    sl, tl = torch.zeros(8, dtype=torch.float64, device=dev), torch.zeros((cap, 2), dtype=torch.float32, device=dev)
    allsums = torch.rand((world, 50, parts, 2), generator=g, dtype=torch.float64) * 50 + 1
    if rank == world - 1:
        time.sleep(2.5)
    ex.run(sl, allsums[rank].to(dev), 50, parts, elems, max_update, tl, 0, cap)
    torch.cuda.synchronize()
    ex.check()
    ref_s, ref_t = torch.zeros_like(sl), torch.zeros_like(tl)
    _native.check(lib.assl_running_stats(P(ref_s), P(allsums.reshape(world * 50, parts, 2).to(dev)), world * 50, parts, elems,
                                         max_update, rank * 50, 50, P(ref_t), 0, cap, st()))
    torch.cuda.synchronize()
    assert torch.equal(sl, ref_s) and torch.equal(tl, ref_t)
    #      (b) timeout 0.5 s, last rank 2 s late: the waiting ranks give up, leave their statistics untouched, raise.
    dist.barrier()
    ex.set_timeout(0.5)
    before = sl.clone()
    if rank == world - 1:
        time.sleep(2.0)
    ex.run(sl, allsums[rank].to(dev), 50, parts, elems, max_update, tl, 0, cap)
    torch.cuda.synchronize()
    raised = False
    try:
        ex.check()
    except RuntimeError:
        raised = True
    if rank == world - 1:
        assert not raised and float(sl[7]) == 0.0            # everyone else's records were there long before
    else:
        assert raised and float(sl[7]) == 1.0
        assert torch.equal(sl[:5], before[:5])               # n_put, carries, mean, std: untouched
    dist.barrier()
    #      (c) the same lateness inside scan + send in one launch: the last rank sends call 1 two seconds late, the others'
    #      scan of call 1 gives up (statistics untouched) -- and STILL sends call 2 from the same kernel, or the late rank,
    #      whose own scan of call 1 finds every record, would wait for call 2 in vain.
    ex_split.set_timeout(0.5)
    sc, tc = torch.zeros(8, dtype=torch.float64, device=dev), torch.zeros((cap, 2), dtype=torch.float32, device=dev)
    mine = allsums[rank].to(dev)
    if rank == world - 1:
        time.sleep(2.0)
    ex_split.send(mine, 50, parts)                                            # call 1
    _native.check(lib.assl_exchange_scan_send(ex_split.handle, P(sc), elems, max_update, P(tc), 0, cap, 0, P(mine), 50, parts,
                                              st()), "assl_exchange_scan_send")                 # scan of call 1 + call 2 out
    torch.cuda.synchronize()
    late = lib.assl_exchange_error(ex_split.handle) != 0
    if rank == world - 1:
        assert not late and float(sc[7]) == 0.0 and float(sc[0]) == min(50 * world, max_update)
    else:
        assert late and float(sc[7]) == 1.0 and float(sc[:5].abs().sum()) == 0.0
    dist.barrier()                                                            # everyone has sent call 2 by now
    ex_split.set_timeout(30.0)
    n_before = float(sc[0])
    ex_split.scan(sc, elems, max_update, tc, 50, cap)                        # call 2: every record is there, on every rank
    torch.cuda.synchronize()
    assert float(sc[0]) == min(n_before + 50 * world, max_update)
    dist.barrier()
    for e in (ex, ex_split, ex_staged):
        e.close()
    if rank == 0:
        print(f"multigpu_check: OK on {world} GPUs (fused == split == staged == NCCL path == single stream, bit for bit; "
              f"uneven shards; global bank == unsharded run, {remote} remote bank entries on rank 0; lookahead order == "
              f"ordinary order; NormalizeBatch(sync); late rank waited for / timeout raised / scan+send still sends)")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
