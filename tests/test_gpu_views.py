"""GPU parity: running norm, mixup, random-resize-crop, two-view module and the fused pipeline
(all through the C ABI) vs the CPU oracle and the golden fixtures generated from the reference.
Gates: drawn integers (crop boxes, bank picks) bit-exact; values |a-b| <= 1e-5 + 1e-4*|b|."""
import os
import random

import numpy as np
import pytest
import torch

from oracle import frontend as ofe
from oracle import views as ov

pytestmark = pytest.mark.gpu

RTOL, ATOL = 1e-4, 1e-5
CFG = {"pretrain": {"normalization": "mean_var", "augmentations": {
    "MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
    "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                         "time_crop_scale": [0.6, 1.5]}}}}


def _close(got, want, rtol=RTOL, atol=ATOL, what=""):
    got = got.detach().cpu().numpy() if isinstance(got, torch.Tensor) else np.asarray(got)
    want = want.detach().cpu().numpy() if isinstance(want, torch.Tensor) else np.asarray(want)
    assert got.shape == want.shape, (got.shape, want.shape)
    err = np.abs(got.astype(np.float64) - want) - (atol + rtol * np.abs(want))
    assert err.max() <= 0, f"{what} max violation {err.max():.3e}, max abs diff {np.abs(got - want).max():.3e}"


def _seed(s=31):
    np.random.seed(s); random.seed(s)


def test_two_views_golden_fixture(golden_dir):
    """Same seeds as the reference run that produced the fixture -> same integers, same views."""
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "views_reference.npz"))
    _seed(int(g["seed"]))
    tf = A.AugmentationModule(CFG, int(g["len_of_files"]))
    v1, v2 = tf(torch.from_numpy(g["x"]))
    p = tf.last_params
    assert np.array_equal(p.crop, g["crop"])
    assert np.array_equal(p.src_clip.astype(np.int32), g["src_clip"])
    assert np.array_equal(p.bank_index, g["bank_index"])
    assert np.array_equal(p.alpha, g["alpha"])
    ms = tf.engine.stats[:g["x"].shape[0]].cpu().numpy()
    _close(ms, g["mean_std"], what="running mean/std")
    _close(v1, g["view1"], what="view1")
    _close(v2, g["view2"], what="view2")


@pytest.mark.parametrize("shape,batches", [((64, 96), [5, 1, 30]), ((64, 101), [64]), ((16, 33), [7, 9]), ((64, 1001), [3])])
@pytest.mark.parametrize("interp", ["bicubic", "bilinear"])
def test_module_matches_oracle_clip_by_clip_and_batched(shape, batches, interp):
    import audiossl_b200 as A
    f, t = shape
    gen = torch.Generator().manual_seed(f * 1000 + t)
    n = sum(batches)
    x = (torch.randn(n, 1, f, t, generator=gen) * 1.2 - 2.8).to(torch.float32)
    _seed()
    orc = ov.TwoViewOracle(4, mode=interp)
    w1, w2, recs = orc.run_batch(x[:, 0])
    arrs = ov.records_to_arrays(recs)
    _seed()
    tf = A.AugmentationModule(CFG, 4, interpolation=interp)
    lo = 0
    for b in batches:
        v1, v2 = tf(x[lo:lo + b].cuda())
        p = tf.last_params
        assert np.array_equal(p.crop, arrs["crop"][lo:lo + b])
        assert np.array_equal(p.src_clip.astype(np.int32), arrs["src_clip"][lo:lo + b])
        _close(v1, w1[lo:lo + b], what=f"view1[{lo}:{lo + b}]")
        _close(v2, w2[lo:lo + b], what=f"view2[{lo}:{lo + b}]")
        lo += b


def test_single_clip_call_shapes():
    import audiossl_b200 as A
    _seed()
    tf = A.AugmentationModule(CFG, 10)
    a, b = tf(torch.randn(1, 64, 96))
    assert tuple(a.shape) == tuple(b.shape) == (1, 64, 96) and a.is_cuda


def test_bank_wraps_past_n_memory():
    """More than 1024 clips so the FIFO drops entries and the ring wraps."""
    import audiossl_b200 as A
    f, t = 8, 12
    gen = torch.Generator().manual_seed(4)
    x = (torch.randn(2600, 1, f, t, generator=gen) - 3).to(torch.float32)
    _seed(5)
    orc = ov.TwoViewOracle(50)
    w1, w2, recs = orc.run_batch(x[:, 0])
    _seed(5)
    tf = A.AugmentationModule(CFG, 50, max_batch=700)
    v1, v2 = tf(x.cuda())
    _close(v1[:, 0], w1[:, 0], what="view1")
    _close(v2[:, 0], w2[:, 0], what="view2")
    assert min(r.views[0].src_clip for r in recs[2000:]) > 0          # old entries really left the FIFO


def test_running_norm_golden_with_freeze(golden_dir):
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "running_norm_reference.npz"))
    rn = A.RunningNorm(int(g["epoch_samples"]))
    x = torch.from_numpy(g["x"])
    outs = [rn(x[:100].cuda()), rn(x[100:101, 0].cuda()).unsqueeze(0), rn(x[101:].cuda())]   # spans the freeze at 250
    assert len(rn) == 250
    _close(np.float32([rn.mean, rn.std]), g["mean_std"][-1])
    _close(outs[2][-1], g["last_out"])
    # every clip's output against the oracle replay
    orc = ov.RunningNormOracle(int(g["epoch_samples"]))
    want = torch.stack([orc(c) for c in x])
    _close(torch.cat(outs), want)


def test_normalize_batch_and_precomputed(golden_dir):
    import audiossl_b200 as A
    g = np.load(os.path.join(golden_dir, "normalize_batch_reference.npz"))
    _close(A.NormalizeBatch()(torch.from_numpy(g["x"])), g["y"])
    x = torch.randn(3, 1, 8, 9)
    _close(A.PrecomputedNorm((-2.5, 1.7))(x), ov.precomputed_norm(x, -2.5, 1.7))


def test_standalone_mixup_and_crop():
    import audiossl_b200 as A
    gen = torch.Generator().manual_seed(11)
    x = (torch.randn(40, 1, 64, 96, generator=gen) * 0.8).to(torch.float32)
    # MixupBYOLA alone (one push per call), oracle = the same recurrence spelled out
    _seed(8)
    bank, want = [], []
    for c in x:
        alpha = 0.4 * np.random.random()
        if bank:
            z = bank[np.random.randint(len(bank))]
            want.append(ov.log_mixup_exp(c, z, 1.0 - alpha))
        else:
            want.append(c)
        bank = (bank + [c])[-16:]
    _seed(8)
    mx = A.MixupBYOLA(ratio=0.4, n_memory=16)
    got = torch.cat([mx(x[:13].cuda()), mx(x[13:].cuda())])
    _close(got, torch.stack(want))
    # RandomResizeCrop alone
    _seed(9)
    boxes = [ov.draw_crop((64, 144), (64, 96)) for _ in range(40)]
    want = torch.stack([ov.resize_crop(c, bx) for c, bx in zip(x, boxes)])
    _seed(9)
    got = A.RandomResizeCrop()(x.cuda())
    _close(got, want)
    ident = A.RandomResizeCrop()(x[:2].cuda(), crop=[(0, 24, 64, 96)] * 2)
    assert torch.equal(ident.cpu(), x[:2])


def test_fused_pipeline_waveform_to_views():
    """Waveforms -> log-mel -> running norm -> two views on the GPU vs oracle front end + oracle views."""
    import audiossl_b200 as A
    from test_gpu_frontend import families
    waves = families(24, 15200, 77)
    lms = torch.from_numpy(ofe.log_mel_batch(waves.numpy()))
    _seed()
    orc = ov.TwoViewOracle(6)
    w1, w2, recs = orc.run_batch(lms)
    _seed()
    pipe = A.UpstreamInputPipeline(CFG, 6)
    v1a, v2a = pipe(waves[:10].cuda())
    v1b, v2b = pipe(waves[10:].cuda())
    # per-family gates of the fused path (noise-like 2e-5, tonal 1e-4): measured in test_gpu_tolerance.py
    from test_gpu_tolerance import assert_fused_close
    assert_fused_close(torch.cat([v1a, v1b]), w1, what="view1")
    assert_fused_close(torch.cat([v2a, v2b]), w2, what="view2")


def test_c_abi_host_pipeline_matches_device_pipeline():
    """assl_pipeline_host_f32 (host buffers, overlapped copy queues, double-buffered workspace) over several
    calls == the device-resident pipeline with the same explicit parameters."""
    import ctypes as C
    import audiossl_b200 as A
    from audiossl_b200 import _native
    from test_gpu_frontend import families
    lib = _native.lib()
    B, L, calls = 40, 15200, 4
    waves = [families(B, L, 300 + k).pin_memory() for k in range(calls)]
    pipe = A.UpstreamInputPipeline(CFG, 5, sampler="fast", seed=3, max_batch=64)
    plan = pipe.plan
    T = plan.num_frames(L)
    cap = pipe.tfms.engine.cap
    host = A.HostPipeline(plan)
    ws_bytes = host.workspace_bytes(B, L, cap)
    assert ws_bytes > 0
    ws = torch.zeros(ws_bytes, dtype=torch.uint8, device="cuda")
    vcfg = _native.ViewCfg(64, T, 64, int(T * 1.5), 0, 1)
    hv = [(torch.empty(B, 64, T).pin_memory(), torch.empty(B, 64, T).pin_memory()) for _ in range(calls)]
    keep, want = [], []
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for k in range(calls):
        v1, v2 = pipe(waves[k].cuda())
        want.append((v1[:, 0].cpu(), v2[:, 0].cpu()))
        p = pipe.tfms.last_params
        mix = np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32)
        alpha, crop = np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop)
        keep.append((mix, alpha, crop))
        host.run(vcfg, ws, waves[k], B, L, (k * B) % cap, cap, 2 * 5 * 10, mix, alpha, crop, hv[k][0], hv[k][1], stream=st)
    host.sync()
    torch.cuda.synchronize()
    for k in range(calls):
        # chunked statistics scans may round the running mean/std differently in the last bit
        _close(hv[k][0], want[k][0], rtol=1e-6, atol=2e-6, what=f"call {k} view1")
        _close(hv[k][1], want[k][1], rtol=1e-6, atol=2e-6, what=f"call {k} view2")


def test_pcm16_host_pipeline_is_the_f32_pipeline_on_x_over_32768():
    """assl_pipeline_host_pcm16 (16-bit PCM in host memory, widened on the device) == assl_pipeline_host_f32 fed
    x / 32768 -- what librosa.core.load hands the reference for a 16-bit file (upstream_dataset.py:55) -- bit for bit
    over several overlapped calls; and the widening kernel on its own, aligned and unaligned."""
    import ctypes as C
    import audiossl_b200 as A
    from audiossl_b200 import _native
    lib = _native.lib()
    B, L, calls = 48, 15200, 4
    g = torch.Generator(device="cpu").manual_seed(77)
    pcm = [torch.randint(-32768, 32768, (B, L), generator=g, dtype=torch.int32).to(torch.int16).pin_memory() for _ in range(calls)]
    pcm[0][0, :4] = torch.tensor([-32768, 32767, 0, -1], dtype=torch.int16)
    f32 = [(p.to(torch.float32) / 32768.0).pin_memory() for p in pcm]
    plan = A.Plan()
    T = plan.num_frames(L)
    cap = 1024 + B
    vcfg = _native.ViewCfg(64, T, 64, int(T * 1.5), 0, 1)
    fast = A.FastSampler(seed=5)
    params = []
    for k in range(calls):
        p = fast.draw(k * B, B, 64, T)
        params.append((np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32), np.ascontiguousarray(p.alpha),
                       np.ascontiguousarray(p.crop)))
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def run(waves):
        host = A.HostPipeline(plan)
        ws = torch.zeros(host.workspace_bytes(B, L, cap), dtype=torch.uint8, device="cuda")
        hv = [(torch.empty(B, 64, T).pin_memory(), torch.empty(B, 64, T).pin_memory()) for _ in range(calls)]
        for k in range(calls):
            mix, alpha, crop = params[k]
            host.run(vcfg, ws, waves[k], B, L, (k * B) % cap, cap, 10 ** 6, mix, alpha, crop, hv[k][0], hv[k][1], stream=st)
        host.sync()
        torch.cuda.synchronize()
        host.close()
        return hv

    want, got = run(f32), run(pcm)
    for k in range(calls):
        assert torch.equal(got[k][0], want[k][0]) and torch.equal(got[k][1], want[k][1]), f"call {k}"
    with pytest.raises(TypeError):
        A.HostPipeline(plan).run(vcfg, torch.zeros(8, dtype=torch.uint8, device="cuda"), f32[0].double(), B, L, 0, cap, 10,
                                 *params[0])
    # the widening step alone: 16-byte aligned (eight samples per thread, ragged tail) and unaligned (one per thread)
    P = lambda t: C.c_void_p(t.data_ptr())                            # noqa: E731
    src = pcm[0].reshape(-1).cuda()
    for off, n in ((0, 100003), (1, 4099), (0, 7), (8, 64)):
        out = torch.full((n + 1,), float("nan"), device="cuda")
        dst = out[1:] if off == 1 else out[:n]
        _native.check(lib.assl_pcm16_to_f32(P(src[off:]), n, P(dst), None), "assl_pcm16_to_f32")
        torch.cuda.synchronize()
        assert torch.equal(dst[:n], src[off:off + n].to(torch.float32) / 32768.0)
    assert lib.assl_pcm16_to_f32(None, 4, P(out), None) == 1


@pytest.mark.parametrize("n_parts", [1, 5, 51])
def test_running_stats_many_chunks_and_freeze(n_parts):
    """20000 clips in one call (three passes of the 8-CTA cluster, a partial last chunk), statistics freezing in the
    middle of a chunk, then a second call that starts frozen: against the reference recurrence in float64."""
    import ctypes as C
    from audiossl_b200 import _native
    lib = _native.lib()
    n, elems, max_update = 20000, 640, 13333
    rng = np.random.default_rng(n_parts)
    parts = rng.random((n, n_parts, 2)) * 40 + 1
    parts[:, :, 1] += 2000                                            # keep q - mu^2 positive
    d_parts = torch.from_numpy(parts).cuda()
    state = torch.zeros(8, dtype=torch.float64, device="cuda")
    out = torch.zeros((n, 2), dtype=torch.float32, device="cuda")
    P = lambda t: C.c_void_p(t.data_ptr())                            # noqa: E731
    for lo, hi in ((0, 15000), (15000, n)):
        _native.check(lib.assl_running_stats(P(state), P(d_parts[lo:hi]), hi - lo, n_parts, elems, max_update, 0, hi - lo,
                                             P(out[lo:hi]), 0, 0, None))
    got = out.cpu().numpy()
    tot = parts.sum(axis=1) / elems
    mu = sv = 0.0
    frozen = None
    for k in range(n):
        m, q = tot[k]
        if k < max_update:
            mu = m if k == 0 else mu + (m - mu) / k                  # divisor = samples seen BEFORE (augmentations.py:226)
            v = max(q - 2 * mu * m + mu * mu, 0.0)
            sv = v if k == 0 else sv + (v - sv) / k
            frozen = (mu, np.sqrt(sv))
        want = frozen
        assert abs(got[k, 0] - want[0]) <= 1e-6 * abs(want[0]) + 1e-7, k
        assert abs(got[k, 1] - want[1]) <= 1e-5 * abs(want[1]) + 1e-7, k
    st = state.cpu().numpy()
    assert st[0] == max_update and abs(st[3] - frozen[0]) <= 1e-6 * abs(frozen[0])


def test_launch_order_does_not_change_the_views():
    """assl_two_views_ordered_f32: any permutation of the views (heaviest first in the product) gives the same bits."""
    import ctypes as C
    import audiossl_b200 as A
    from audiossl_b200 import _native
    from audiossl_b200.engine import heavy_first
    lib = _native.lib()
    b, f, t = 37, 64, 101
    g = torch.Generator(device="cpu").manual_seed(3)
    ring = torch.randn(b, f, t, generator=g).cuda()
    stats = torch.stack((torch.randn(b, generator=g) * 0.1, torch.rand(b, generator=g) + 0.5), 1).contiguous().cuda()
    p = A.FastSampler(seed=5).draw(b, b, f, t)                  # bank picks among the b clips before this batch
    mix = torch.from_numpy(np.where(p.src_clip >= 0, p.src_clip % b, -1).astype(np.int32)).cuda()
    clip = torch.arange(b, dtype=torch.int32).cuda()
    alpha, crop = torch.from_numpy(p.alpha).cuda(), torch.from_numpy(p.crop).cuda()
    vcfg = _native.ViewCfg(f, t, f, int(t * 1.5), 0, 1)
    P = lambda x: C.c_void_p(x.data_ptr()) if x is not None else None   # noqa: E731
    outs = []
    order = heavy_first(p.crop)
    assert sorted(order.tolist()) == list(range(2 * b))
    cost = p.crop.reshape(-1, 4)[order]
    assert cost[0, 2] * cost[0, 3] >= cost[-1, 2] * cost[-1, 3]
    shuffled = torch.from_numpy(np.random.default_rng(0).permutation(2 * b).astype(np.int32)).cuda()
    for o in (None, torch.from_numpy(order).cuda(), shuffled):
        v1, v2 = torch.empty(b, f, t, device="cuda"), torch.empty(b, f, t, device="cuda")
        _native.check(lib.assl_two_views_ordered_f32(C.byref(vcfg), P(ring), P(stats), b, P(clip), P(mix), P(alpha), P(crop),
                                                     P(o), b, P(v1), P(v2), None))
        outs.append((v1, v2))
    for v1, v2 in outs[1:]:
        assert torch.equal(v1, outs[0][0]) and torch.equal(v2, outs[0][1])


@pytest.mark.parametrize("t", [33, 96, 101, 330, 1001])
def test_padded_ring_rows_give_the_same_bits_as_dense_rows(t):
    """The two-view kernel reads 16-byte aligned ring rows (assl_view_cfg.ring_pitch, a multiple of 4) with 128-bit loads
    and dense rows with scalar loads: same arithmetic, same bits, for every tile size and the T that is not a multiple
    of 4.  Garbage in the pad columns must not leak into a view."""
    import ctypes as C
    import audiossl_b200 as A
    from audiossl_b200 import _native
    lib = _native.lib()
    b, f = 24, 64
    g = torch.Generator(device="cpu").manual_seed(t)
    dense = torch.randn(b, f, t, generator=g).cuda()
    pitch = (t + 3) & ~3
    padded = torch.full((b, f, pitch), float("nan"), device="cuda")      # NaN pads: any use of them would show
    padded[:, :, :t] = dense
    stats = torch.stack((torch.randn(b, generator=g) * 0.1, torch.rand(b, generator=g) + 0.5), 1).contiguous().cuda()
    p = A.FastSampler(seed=5 + t).draw(b, b, f, t)
    mix = torch.from_numpy(np.where(p.src_clip >= 0, p.src_clip % b, -1).astype(np.int32)).cuda()
    clip = torch.arange(b, dtype=torch.int32).cuda()
    alpha, crop = torch.from_numpy(p.alpha).cuda(), torch.from_numpy(p.crop).cuda()
    P = lambda x: C.c_void_p(x.data_ptr()) if x is not None else None   # noqa: E731
    outs = []
    for ring, pt in ((dense, 0), (padded, pitch)):
        vcfg = _native.ViewCfg(f, t, f, int(t * 1.5), 0, 1, pt)
        v1, v2 = torch.empty(b, f, t, device="cuda"), torch.empty(b, f, t, device="cuda")
        _native.check(lib.assl_two_views_f32(C.byref(vcfg), P(ring), P(stats), b, P(clip), P(mix), P(alpha), P(crop), b,
                                             P(v1), P(v2), None))
        outs.append((v1, v2))
    torch.cuda.synchronize()
    assert torch.isfinite(outs[1][0]).all() and torch.isfinite(outs[1][1]).all()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


def test_delores_s_shard_512_clips_of_15200_samples():
    """BASELINE config 3's per-GPU shard as one full batch (512 x 0.95 s, /root/reference/src/upstream/delores_s/config.yaml:18)
    through UpstreamInputPipeline: crop boxes and bank picks bit-exact, views inside the fused-path gates."""
    import audiossl_b200 as A
    from test_gpu_frontend import families
    from test_gpu_tolerance import assert_fused_close
    waves = families(512, 15200, 512)
    lms = torch.from_numpy(ofe.log_mel_batch(waves.numpy()))
    _seed()
    w1, w2, recs = ov.TwoViewOracle(2000).run_batch(lms)
    _seed()
    pipe = A.UpstreamInputPipeline(CFG, 2000, max_batch=512)
    v1, v2 = pipe(waves.cuda())
    arrs = ov.records_to_arrays(recs)
    p = pipe.tfms.last_params
    assert np.array_equal(p.crop, arrs["crop"]) and np.array_equal(p.src_clip.astype(np.int32), arrs["src_clip"])
    assert_fused_close(v1, w1, what="view1")
    assert_fused_close(v2, w2, what="view2")


def test_two_views_with_128_mel_bands():
    """The MAST variant's 128 bands (extras/mast_new/mast/utils.py:251-259) through the two-view kernel: a canvas of 128
    rows (twice the shared memory per CTA)."""
    import audiossl_b200 as A
    g = torch.Generator(device="cpu").manual_seed(128)
    x = torch.randn(20, 128, 101, generator=g) * 2.0 - 4.0
    _seed(5)
    w1, w2, recs = ov.TwoViewOracle(50).run_batch(x)
    _seed(5)
    tf = A.AugmentationModule(CFG, 50, max_batch=32)
    v1, v2 = tf(x.unsqueeze(1).cuda())
    arrs = ov.records_to_arrays(recs)
    assert np.array_equal(tf.last_params.crop, arrs["crop"])
    _close(v1, w1, what="view1")
    _close(v2, w2, what="view2")


@pytest.mark.parametrize("t", [96, 330])
def test_canvas_taller_than_the_clip(t):
    """virtual_crop_scale[0] > 1: canvas rows above and below the pasted clip (zero-filled apart from the clip rows in the
    two-view kernel) and a frequency scale above 1 (tap rows that move by two per output row)."""
    import audiossl_b200 as A
    vs = (1.4, 1.5)
    g = torch.Generator(device="cpu").manual_seed(14 + t)
    x = torch.randn(24, 1, 64, t, generator=g) * 2.0 - 3.0
    ch, cw = int(64 * vs[0]), int(t * vs[1])
    _seed(3)
    boxes = [ov.draw_crop((ch, cw), (64, t)) for _ in range(24)]
    boxes[0] = (0, 0, ch, cw)                                   # the whole canvas
    boxes[1] = (0, 5, 10, 40)                                   # canvas rows only at the top ...
    boxes[2] = (ch - 12, 0, 12, cw)                             # ... and at the bottom
    assert max(b[2] for b in boxes) > 64
    for mode in ("bicubic", "bilinear"):
        want = torch.stack([ov.resize_crop(c, bx, vs, mode) for c, bx in zip(x, boxes)])
        got = A.RandomResizeCrop(virtual_crop_scale=vs, interpolation=mode)(x.cuda(), crop=boxes)
        _close(got, want, what=mode)
