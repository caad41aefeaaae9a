"""The generated register FFT (audio-ssl_b200/csrc/fft32p_gen.cuh) is emitted from an op list that can
also be executed in numpy (fp32, FMA = one rounding).  Check the op list against numpy's FFT, check the
32x32 two-frame packing / unpacking index algebra the kernel uses, and check the header is up to date."""
import io
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_fft32  # noqa: E402


def test_two_frame_packing_through_generated_fft():
    """The kernel's whole data flow in numpy: window/2, 32x32 split, twiddles, transpose, unpack."""
    n = 1024
    rng = np.random.default_rng(2)
    x1, x2 = rng.standard_normal(n).astype(np.float32), rng.standard_normal(n).astype(np.float32)
    w = 0.5 - 0.5 * np.cos(2 * np.pi * np.arange(n) / n)
    wh = (0.5 * w).astype(np.float32)
    # lane n1 holds samples n1 + 32*n2  -> arrays [n2, n1]
    re, im = (wh * x1).reshape(32, 32), (wh * x2).reshape(32, 32)
    yr, yi = gen_fft32.run_numpy_packed(re, im)                                   # [k2, n1]
    k2, n1 = np.arange(32)[:, None], np.arange(32)[None, :]
    tw = np.exp(-2j * np.pi * k2 * n1 / n).astype(np.complex64)
    y = ((yr + 1j * yi) * tw).astype(np.complex64)
    zr, zi = gen_fft32.run_numpy_packed(y.real.T.copy(), y.imag.T.copy())          # stage 2 over n1: [k1, k2]
    z = zr + 1j * zi
    p1, p2 = np.zeros(513), np.zeros(513)
    for lane in range(32):
        for k1 in range(17):
            k = lane + 32 * k1
            if k > 512 or (lane != 0 and k1 == 16):
                continue
            a = z[k1, lane]
            b = z[(32 - k1) % 32, 0] if lane == 0 else z[31 - k1, 32 - lane]
            p1[k] = (a.real + b.real) ** 2 + (a.imag - b.imag) ** 2
            p2[k] = (a.imag + b.imag) ** 2 + (a.real - b.real) ** 2
    r1 = np.abs(np.fft.rfft(w * x1)) ** 2
    r2 = np.abs(np.fft.rfft(w * x2)) ** 2
    assert np.abs(p1 - r1).max() < 2e-6 * r1.max() and np.abs(p2 - r2).max() < 2e-6 * r2.max()


def test_packed_header_is_current():
    buf = io.StringIO()
    gen_fft32.emit_cuda_packed(buf)
    with open(os.path.join(ROOT, "audio-ssl_b200", "csrc", "fft32p_gen.cuh")) as fh:
        assert fh.read() == buf.getvalue(), "regenerate with: python tools/gen_fft32.py > audio-ssl_b200/csrc/fft32p_gen.cuh"


def test_packed_op_list_matches_numpy_fft():
    """The {re, im}-pair network the kernel runs (FFMA2 / FADD2): 194 two-wide ops, fp32-grade accuracy."""
    rng = np.random.default_rng(3)
    x = rng.standard_normal((32, 4000)) + 1j * rng.standard_normal((32, 4000))
    re, im = gen_fft32.run_numpy_packed(x.real.astype(np.float32), x.imag.astype(np.float32))
    ref = np.fft.fft(x.real.astype(np.float32).astype(np.complex128) + 1j * x.imag.astype(np.float32), axis=0)
    err = np.abs((re + 1j * im) - ref)
    scale = np.sqrt((np.abs(ref) ** 2).mean())
    assert err.max() < 1e-6 * scale * 8 and np.sqrt((err ** 2).mean()) < 2e-7 * scale
    ops, _ = gen_fft32.build_fft32_packed()
    assert len(ops) == 194


def test_factorised_twiddles_and_analytic_window():
    """The kernel keeps w^8, w^16, w^24, w^1..w^7 per lane and rotates twice for the other exponents, and
    evaluates the Hann window from one sin/cos pair per lane; both stay at fp32 rounding level."""
    n = 1024
    lane = np.arange(32)
    w = np.exp(-2j * np.pi * lane / n)
    base = {k: (w ** k).astype(np.complex64) for k in (8, 16, 24, 1, 2, 3, 4, 5, 6, 7)}
    worst = 0.0
    for k2 in range(1, 32):
        a, b = divmod(k2, 8)
        t = np.ones(32, np.complex64)
        if b:
            t = (t * base[b]).astype(np.complex64)
        if a:
            t = (t * base[8 * a]).astype(np.complex64)
        worst = max(worst, np.abs(t - w ** k2).max())
    assert worst < 3e-7
    n2 = np.arange(32)[:, None]
    theta = (lane[None, :] / 512.0)
    wa = (-0.25 * np.cos(np.pi * theta)).astype(np.float32)
    wb = (0.25 * np.sin(np.pi * theta)).astype(np.float32)
    c, s = np.cos(2 * np.pi * n2 / 32).astype(np.float32), np.sin(2 * np.pi * n2 / 32).astype(np.float32)
    win = (np.float32(0.25) + wa * c + wb * s).astype(np.float32)
    ref = 0.5 * (0.5 - 0.5 * np.cos(2 * np.pi * (lane[None, :] + 32 * n2) / n))
    assert np.abs(win - ref).max() < 1.5e-7
