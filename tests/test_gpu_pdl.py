"""Programmatic dependent launch (csrc/common.cuh): the kernels of the path may be scheduled before the previous
kernel on the stream has finished; they must still see everything it wrote.  Producers here are FOREIGN kernels
(torch element-wise ops) on a side stream with no synchronisation before the call, and the whole pipeline is compared
bit for bit with ASSL_PDL=0 (ordinary launches) in a second process."""
import os
import subprocess
import sys
import zlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CFG = {"pretrain": {"normalization": "mean_var", "augmentations": {
    "MixupBYOLA": {"ratio": 0.4, "log_mixup_exp": True},
    "RandomResizeCrop": {"virtual_crop_scale": [1.0, 1.5], "freq_crop_scale": [0.6, 1.5],
                         "time_crop_scale": [0.6, 1.5]}}}}


def test_inputs_written_by_foreign_kernels_right_before_the_call():
    import audiossl_b200 as A
    mel = A.MelSpectrogramLibrosa()
    g = torch.Generator(device="cpu").manual_seed(5)
    base = (torch.randn(256, 16000, generator=g) * 0.1).cuda()
    scales = [0.25 + 0.05 * i for i in range(12)]
    want = []
    for s in scales:                                   # reference results: one step at a time, synchronised
        w = base * s
        torch.cuda.synchronize()
        want.append(A.extract_log_mel_spectrogram(w, mel).clone())
        torch.cuda.synchronize()
    side = torch.cuda.Stream()
    got = []
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        buf = torch.empty_like(base)
        for s in scales:                               # producer kernel -> log-mel kernel, back to back, one buffer
            torch.mul(base, s, out=buf)
            got.append(A.extract_log_mel_spectrogram(buf, mel).clone())
    side.synchronize()
    for a, b in zip(got, want):
        assert torch.equal(a, b)


_SCRIPT = r"""
import random, sys, zlib
import numpy as np, torch
sys.path.insert(0, {root!r})
import audiossl_b200 as A
cfg = {cfg!r}
g = torch.Generator(device="cpu").manual_seed(11)
waves = (torch.randn(6, 96, 15200, generator=g) * 0.1).cuda()
np.random.seed(31); random.seed(31)
pipe = A.UpstreamInputPipeline(cfg, 10 ** 6)
crc = 0
for k in range(6):                                     # six steps streamed without a sync: every kernel follows a kernel
    v1, v2 = pipe(waves[k])
    keep = (v1, v2) if k == 0 else keep + (v1, v2)
torch.cuda.synchronize()
for v in keep:
    crc = zlib.crc32(v.cpu().numpy().tobytes(), crc)
print("CRC", crc)
"""


def _run(pdl):
    env = dict(os.environ, ASSL_PDL=pdl)
    res = subprocess.run([sys.executable, "-c", _SCRIPT.format(root=ROOT, cfg=CFG)], env=env, capture_output=True,
                         text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    return [l for l in res.stdout.splitlines() if l.startswith("CRC")][-1]


def test_pipeline_is_bit_identical_with_and_without_pdl():
    assert _run("1") == _run("0")


def test_lookahead_pipeline_returns_the_same_views_one_call_later():
    """Step order log-mel(k), statistics(k), views(k-1) (assl_two_views_deferred_f32) vs the ordinary order: same
    draws, same running statistics, same bank -> bit-identical views, delivered one call later; flush() gives the last."""
    import random

    import audiossl_b200 as A
    g = torch.Generator(device="cpu").manual_seed(21)
    sizes = [64, 64, 17, 64, 40, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 64, 33]   # > 1024 clips: bank wraps
    batches = [(torch.randn(n, 15200, generator=g) * 0.1).cuda() for n in sizes]

    np.random.seed(31); random.seed(31)
    plain = A.UpstreamInputPipeline(CFG, 10 ** 6, max_batch=64)
    want = [plain(w) for w in batches]
    torch.cuda.synchronize()

    np.random.seed(31); random.seed(31)
    look = A.UpstreamInputPipeline(CFG, 10 ** 6, max_batch=64, lookahead=True)
    assert look.tfms.engine.cap == plain.tfms.engine.cap + 64
    got = []
    for k, w in enumerate(batches):                    # streamed, no synchronisation between the calls
        out = look(w)
        assert (out is None) == (k == 0)
        if out is not None:
            got.append(out)
    with pytest.raises(RuntimeError):
        look.tfms.state_dict()                          # a batch is still pending
    got.append(look.flush())
    assert look.flush() is None
    torch.cuda.synchronize()
    assert len(got) == len(want)
    for (a1, a2), (b1, b2) in zip(got, want):
        assert a1.shape == b1.shape and torch.equal(a1, b1) and torch.equal(a2, b2)
    sd = look.tfms.state_dict()
    assert sd["clips_done"] == sum(sizes)


def test_deferred_logmel_gives_the_same_views():
    """inputs_ready=True: the log-mel kernel is launched with assl_logmel_deferred_f32 (reads and transforms its first
    frame pairs under the tail of the views launched before it, waits only before its first store).  Resident
    waveforms, many steps streamed without a synchronisation, batches big enough for every SM to hold a CTA of the
    views when the next log-mel kernel arrives: bit-identical to the ordinary lookahead order."""
    import random

    import audiossl_b200 as A
    g = torch.Generator(device="cpu").manual_seed(23)
    batches = [(torch.randn(1024, 16000, generator=g) * 0.1).cuda() for _ in range(4)]
    order = [k % 4 for k in range(14)]
    torch.cuda.synchronize()

    def run(ready):
        np.random.seed(31); random.seed(31)
        pipe = A.UpstreamInputPipeline(CFG, 10 ** 7, max_batch=1024, lookahead=True, inputs_ready=ready)
        outs = []
        for k in order:
            out = pipe(batches[k])
            if out is not None:
                outs.append((out[0].clone(), out[1].clone()))
        outs.append(pipe.flush())
        torch.cuda.synchronize()
        return outs

    want, got = run(False), run(True)
    assert len(got) == len(want) == len(order)
    for (a1, a2), (b1, b2) in zip(got, want):
        assert torch.equal(a1, b1) and torch.equal(a2, b2)


def test_deferred_logmel_alone_matches_the_ordinary_launch():
    """assl_logmel_deferred_f32 through the C ABI right behind a long foreign kernel that writes NONE of its inputs."""
    import audiossl_b200 as A
    from audiossl_b200 import _native
    from audiossl_b200.frontend import _ptr, _stream_ptr
    plan = A.Plan()
    g = torch.Generator(device="cpu").manual_seed(29)
    wave = (torch.randn(300, 15200, generator=g) * 0.1).cuda()
    t = plan.num_frames(15200)
    want = torch.empty((300, 64, t), device="cuda")
    plan.logmel_into(wave, want)
    torch.cuda.synchronize()
    got = torch.full((300, 64, t), float("nan"), device="cuda")
    filler = torch.empty(64 << 20, device="cuda")
    for _ in range(3):
        filler.normal_()                                # (busy SMs when the deferred launch arrives)
        plan.logmel_into(wave, got, deferred=True)
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    assert _native.lib().assl_logmel_deferred_f32(None, _ptr(wave), None, 1, 15200, 15200, _ptr(got), 0, 0, 0, None, 0,
                                                  _stream_ptr()) == 1
