#!/usr/bin/env python3
"""Where does the host-buffer path (assl_pipeline_host_f32) lose time against the PCIe ceiling?  Variants of the bench
e2e loop: full, no download, parameters drawn ahead of the timed loop."""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import audiossl_b200 as A  # noqa: E402
from audiossl_b200 import _native  # noqa: E402

B, L, K, W = 1024, 16000, 60, 5
lib = _native.lib()
plan = A.Plan()
T = plan.num_frames(L)
cap = 2048
vcfg = _native.ViewCfg(64, T, 64, int(T * 1.5), 0, 1)
pipe = A.HostPipeline(plan)
ws_bytes = pipe.workspace_bytes(B, L, cap)
ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
ws[:64].zero_()
g = torch.Generator().manual_seed(1)
waves = [(torch.randn(B, L, generator=g) * 0.1).pin_memory() for _ in range(4)]
hv1 = torch.empty((B, 64, T)).pin_memory()
hv2 = torch.empty((B, 64, T)).pin_memory()
fast = A.FastSampler(seed=9)
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)  # noqa: E731


def params(s):
    p = fast.draw(s * B, B, 64, T)
    mix = np.where(p.src_clip >= 0, p.src_clip % cap, -1).astype(np.int32)
    return mix, np.ascontiguousarray(p.alpha), np.ascontiguousarray(p.crop)


def loop(download=True, ahead=False):
    pre = [params(s) for s in range(W + K)] if ahead else None
    keep = []
    host_ms = 0.0

    def step(s):
        nonlocal host_ms
        t = time.perf_counter()
        mix, alpha, crop = pre[s] if ahead else params(s)
        keep.append((mix, alpha, crop))
        pipe.run(vcfg, ws, waves[s % 4], B, L, (s * B) % cap, cap, 10 ** 9, mix, alpha, crop,
                 hv1 if download else None, hv2 if download else None, stream=st())
        host_ms += time.perf_counter() - t
    for s in range(W):
        step(s)
    pipe.sync(); torch.cuda.synchronize()
    host_ms = 0.0
    t0 = time.perf_counter()
    for s in range(W, W + K):
        step(s)
    pipe.sync(); torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(f"download={download!s:5} ahead={ahead!s:5}  {1e3 * dt / K:6.3f} ms/step  {B * K / dt / 1e3:7.1f} k clips/s  "
          f"H2D {B * L * 4 * K / dt / 1e9:5.1f} GB/s  host enqueue {1e3 * host_ms / K:5.3f} ms/step")


import os
loop(True, False)
loop(True, True)
loop(False, True)
for c in (1, 2, 4, 8):
    os.environ['ASSL_PIPE_CHUNKS'] = str(c)
    print('chunks', c, end='  ')
    loop(True, True)
