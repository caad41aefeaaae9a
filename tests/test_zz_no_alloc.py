"""assl_running_norm_stats_static_f32 -- the entry point that allocates nothing (scratch in a file-scope device array;
no counterpart in the reference) -- against the existing entry points
assl_clip_moments_f32 + assl_running_stats, through the C ABI, on two calls in a row.

Collected last (the file name sorts behind every other test file).
Passes alone in a fresh process and after the other tests: it owns every buffer it touches and the static scratch
needs no initial state.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
RTOL, ATOL = 1e-4, 1e-5          # the gate of the other GPU tests


def _close(got, want, what):
    got, want = got.detach().cpu().numpy().astype(np.float64), want.detach().cpu().numpy().astype(np.float64)
    err = np.abs(got - want) - (ATOL + RTOL * np.abs(want))
    assert err.max() <= 0, f"{what}: max violation {err.max():.3e}"


def test_static_scratch_entry_matches_the_allocating_path_on_two_calls():
    import audiossl_b200 as A
    from audiossl_b200 import _native
    lib = _native.lib()
    P = lambda t: C.c_void_p(t.data_ptr())                      # noqa: E731
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)   # noqa: E731
    assert A is not None
    f, t, cap, max_update = 64, 101, 1024, 700
    elems = f * t
    g = torch.Generator(device="cpu").manual_seed(2024)
    # second call SMALLER than the first: stale entries of the shared scratch must not reach the result; together the
    # calls cross max_update (700 < 600 + 250), so the freeze branch runs too
    batches = [(torch.randn(600, f, t, generator=g) * 3.0 - 5.0).cuda(), (torch.randn(250, f, t, generator=g) * 2.0 + 1.0).cuda()]
    state_a = torch.zeros(8, dtype=torch.float64, device="cuda")
    state_b = torch.zeros(8, dtype=torch.float64, device="cuda")
    stats_a = torch.zeros((cap, 2), dtype=torch.float32, device="cuda")
    stats_b = torch.zeros((cap, 2), dtype=torch.float32, device="cuda")
    done = 0
    for x in batches:
        b = x.shape[0]
        slot0 = done % cap
        _native.check(lib.assl_running_norm_stats_static_f32(P(state_a), P(x), b, elems, max_update, P(stats_a), slot0, cap, st()),
                      "assl_running_norm_stats_static_f32")
        sums = torch.empty((b, 2), dtype=torch.float64, device="cuda")          # the caller-owned scratch of the existing path
        _native.check(lib.assl_clip_moments_f32(P(x), b, elems, P(sums), st()), "assl_clip_moments_f32")
        _native.check(lib.assl_running_stats(P(state_b), P(sums), b, 1, elems, max_update, 0, b, P(stats_b), slot0, cap, st()),
                      "assl_running_stats")
        torch.cuda.synchronize()
        _close(stats_a, stats_b, f"mean/std after {done + b} clips")
        _close(state_a, state_b, f"state after {done + b} clips")
        done += b
    assert float(state_a[0]) == max_update                      # frozen at max_update puts
    assert stats_a[:done, 1].min() > 0                           # every clip got a (mean, std)
    # same kernels on the same numbers: in fact identical
    assert torch.equal(stats_a, stats_b) and torch.equal(state_a, state_b)


def test_static_scratch_entry_rejects_what_it_cannot_hold():
    from audiossl_b200 import _native
    lib = _native.lib()
    P = lambda t: C.c_void_p(t.data_ptr())                      # noqa: E731
    state = torch.zeros(8, dtype=torch.float64, device="cuda")
    stats = torch.zeros((8, 2), dtype=torch.float32, device="cuda")
    x = torch.zeros(4, device="cuda")
    too_many = 4096 + 1                                          # ASSL_STATIC_MAX_CLIPS + 1
    rc = lib.assl_running_norm_stats_static_f32(P(state), P(x), too_many, 1, 10, P(stats), 0, 0, None)
    assert rc == 6 and b"static scratch" in lib.assl_status_string(rc)          # ASSL_ERR_TOO_LARGE, nothing launched
    assert lib.assl_running_norm_stats_static_f32(None, P(x), 1, 1, 10, P(stats), 0, 0, None) == 1
    assert lib.assl_running_norm_stats_static_f32(P(state), P(x), 0, 1, 10, P(stats), 0, 0, None) == 0      # empty batch: no-op
    torch.cuda.synchronize()
    assert float(state.abs().sum()) == 0.0
