#!/usr/bin/env python3
"""Pinned host <-> device copy bandwidth of the box (one direction, then both at once): the ceiling of bench.py's e2e."""
import time

import torch

n = 64 << 20                       # 64 Mi floats = 256 MiB
h_in, h_out = torch.empty(n).pin_memory(), torch.empty(n).pin_memory()
d_in, d_out = torch.empty(n, device="cuda"), torch.empty(n, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down, reps=8):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        if up:
            with torch.cuda.stream(s1):
                d_in.copy_(h_in, non_blocking=True)
        if down:
            with torch.cuda.stream(s2):
                h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    return reps * n * 4 / dt / 1e9


for _ in range(2):
    run(True, True, 2)
print(f"H2D alone   {run(True, False):6.1f} GB/s")
print(f"D2H alone   {run(False, True):6.1f} GB/s")
print(f"both at once {run(True, True):6.1f} GB/s each direction")
