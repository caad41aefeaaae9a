#!/usr/bin/env python3
"""Generator for ``audio-ssl_b200/csrc/fft32p_gen.cuh``: a fully unrolled, register-resident 32-point complex FFT
(forward, e^{-2 pi i jk/32}) on {re, im} register pairs, written as straight-line two-wide FP32 code.

The op list is built once and has two back ends:
  * ``emit_cuda_packed`` -- prints the CUDA device function;
  * ``run_numpy_packed`` -- executes the very same ops on float32 numpy arrays (FMA modelled as float64
    multiply-add rounded once), so the emitted arithmetic can be checked against numpy's FFT in this GPU-less
    container (tests/test_fft32_codegen.py).

Algorithm: radix-2 decimation in time, natural-order in, natural-order out (the bit-reversal is a renaming of
registers); the trivial twiddles (1, -i, (+-1-i)/sqrt2) are specialised.

Usage:  python tools/gen_fft32.py > audio-ssl_b200/csrc/fft32p_gen.cuh
"""
import math
import sys

import numpy as np

N = 32


def bitrev(i, bits):
    r = 0
    for _ in range(bits):
        r = (r << 1) | (i & 1)
        i >>= 1
    return r


def lit(c):
    return f"{np.float32(c)!r}".replace("np.float32(", "").rstrip(")") + "f"


# ---------------------------------------------------------------------------------------------------
# Packed variant: one 64-bit register pair holds {re, im} of a point and every butterfly is written
# with the sm_100 two-wide FP32 instructions (fma/add/sub.rn.f32x2 -> SASS FFMA2 / FADD2, whose
# operand modifiers swap and negate halves for free):
#     W = 1      o1 = a + b                       o2 = a - b                        (2 instr)
#     W = -i     o1 = a + {bi, -br}               o2 = a - {bi, -br}                (2)
#     W = c(1-i) s  = b + {bi, -br}               o1 = a + c s      o2 = a - c s    (3)
#     W = c(-1-i) s = {bi, -br} - b               o1 = a + c s      o2 = a - c s    (3)
#     general    t  = a + wr b      o1 = t + wi {-bi, br}           o2 = 2a - o1    (3)
# 46 trivial x 2 + 34 x 3 = 194 instructions per 32-point transform.

def build_fft32_packed():
    """Returns (ops, out) where ops = [(dst, kind, args)], values are names of complex registers."""
    ops, n = [], [0]

    def new(kind, *args):
        n[0] += 1
        d = f"t{n[0]}"
        ops.append((d, kind, args))
        return d

    bits = 5
    z = [f"z[{bitrev(i, bits)}]" for i in range(N)]
    m = 2
    while m <= N:
        half = m // 2
        for s in range(0, N, m):
            for j in range(half):
                ia, ib = s + j, s + j + half
                a, b = z[ia], z[ib]
                if j == 0:
                    o1, o2 = new("add", a, b), new("sub", a, b)
                elif 4 * j == m:
                    o1, o2 = new("add_mj", a, b), new("sub_mj", a, b)
                elif 8 * j == m:
                    c = math.sqrt(0.5)
                    sv = new("add_mj", b, b)                  # {br + bi, bi - br}
                    o1, o2 = new("fmac", c, sv, a), new("fmac", -c, sv, a)
                elif 8 * j == 3 * m:
                    c = math.sqrt(0.5)
                    sv = new("mj_sub", b)                     # {bi - br, -br - bi}
                    o1, o2 = new("fmac", c, sv, a), new("fmac", -c, sv, a)
                else:
                    ang = -2.0 * math.pi * j / m
                    wr, wi = math.cos(ang), math.sin(ang)
                    t = new("fmac", wr, b, a)
                    o1 = new("fma_j", wi, b, t)               # t + wi * {-bi, br}
                    o2 = new("two_minus", a, o1)              # 2a - o1
                z[ia], z[ib] = o1, o2
        m *= 2
    return ops, z


def run_numpy_packed(re_in, im_in):
    """Execute the packed op list per component in float32 (FMA = one rounding)."""
    ops, out = build_fft32_packed()
    env = {f"z[{i}]": (np.asarray(re_in[i], np.float32), np.asarray(im_in[i], np.float32)) for i in range(N)}
    f32, f64 = np.float32, np.float64

    def fma(a, b, c):
        return (np.asarray(a, f64) * np.asarray(b, f64) + np.asarray(c, f64)).astype(f32)

    for d, kind, args in ops:
        if kind == "add":
            a, b = env[args[0]], env[args[1]]; env[d] = ((a[0] + b[0]).astype(f32), (a[1] + b[1]).astype(f32))
        elif kind == "sub":
            a, b = env[args[0]], env[args[1]]; env[d] = ((a[0] - b[0]).astype(f32), (a[1] - b[1]).astype(f32))
        elif kind == "add_mj":      # a + (-i) b = {ar + bi, ai - br}
            a, b = env[args[0]], env[args[1]]; env[d] = ((a[0] + b[1]).astype(f32), (a[1] - b[0]).astype(f32))
        elif kind == "sub_mj":      # a - (-i) b = {ar - bi, ai + br}
            a, b = env[args[0]], env[args[1]]; env[d] = ((a[0] - b[1]).astype(f32), (a[1] + b[0]).astype(f32))
        elif kind == "mj_sub":      # (-i) b - b = {bi - br, -br - bi}
            b = env[args[0]]; env[d] = ((b[1] - b[0]).astype(f32), (-b[0] - b[1]).astype(f32))
        elif kind == "fmac":        # a + c * s
            c, sv, a = args[0], env[args[1]], env[args[2]]; env[d] = (fma(f32(c), sv[0], a[0]), fma(f32(c), sv[1], a[1]))
        elif kind == "fma_j":       # t + w * {-bi, br}
            w, b, t = args[0], env[args[1]], env[args[2]]; env[d] = (fma(f32(w), -b[1], t[0]), fma(f32(w), b[0], t[1]))
        elif kind == "two_minus":   # 2a - o
            a, o = env[args[0]], env[args[1]]; env[d] = (fma(f32(2.0), a[0], -o[0]), fma(f32(2.0), a[1], -o[1]))
        else:
            raise ValueError(kind)
    return np.stack([env[n][0] for n in out]), np.stack([env[n][1] for n in out])


def emit_cuda_packed(out=sys.stdout):
    ops, res = build_fft32_packed()
    w = out.write
    w("// GENERATED by tools/gen_fft32.py -- do not edit; regenerate instead.\n")
    w("// 32-point forward complex FFT on {re, im} register pairs, natural order in and out.\n")
    w(f"// {len(ops)} two-wide FP32 instructions (FFMA2 / FADD2); helpers in f32x2.cuh.\n")
    w("#pragma once\n\n#include \"f32x2.cuh\"\n\nnamespace assl {\n\n")
    w("__device__ __forceinline__ void fft32p(c32 (&z)[32]) {\n")
    for d, kind, args in ops:
        if kind in ("add", "sub", "add_mj", "sub_mj"):
            w(f"  const c32 {d} = c_{kind}({args[0]}, {args[1]});\n")
        elif kind == "mj_sub":
            w(f"  const c32 {d} = c_mj_sub({args[0]});\n")
        elif kind in ("fmac", "fma_j"):
            w(f"  const c32 {d} = c_{kind}({lit(args[0])}, {args[1]}, {args[2]});\n")
        elif kind == "two_minus":
            w(f"  const c32 {d} = c_two_minus({args[0]}, {args[1]});\n")
    for i in range(N):
        w(f"  z[{i}] = {res[i]};\n")
    w("}\n\n}  // namespace assl\n")


if __name__ == "__main__":
    emit_cuda_packed()
