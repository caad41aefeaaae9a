// Two-wide FP32 arithmetic on sm_100a: fma/add/sub/mul.rn.f32x2 operate on a 64-bit register pair and
// compile to FFMA2 / FADD2 / FMUL2, which take one issue slot for two IEEE fp32 results.  ptxas folds
// the half swaps / negations / scalar broadcasts written below into operand modifiers
// (.F32x2.LO_HI, .NP / .PN, .F32), so the helpers cost exactly one instruction each.
//
// c32 holds a complex number {re (low half), im (high half)}.
#pragma once

namespace assl {

typedef unsigned long long c32;

__device__ __forceinline__ c32 pk(float lo, float hi) {
  c32 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk(c32 v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ float lo_of(c32 v) { float a, b; upk(v, a, b); return a; }
__device__ __forceinline__ float hi_of(c32 v) { float a, b; upk(v, a, b); return b; }

__device__ __forceinline__ c32 fma2(c32 a, c32 b, c32 c) {
  c32 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ c32 add2(c32 a, c32 b) {
  c32 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ c32 sub2(c32 a, c32 b) {
  c32 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ c32 mul2(c32 a, c32 b) {
  c32 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}

// {b.im, -b.re} = (-i) b   and   {-b.im, b.re} = i b, as operands (folded into modifiers)
__device__ __forceinline__ c32 mj(c32 b) { float r, i; upk(b, r, i); return pk(i, -r); }
__device__ __forceinline__ c32 pj(c32 b) { float r, i; upk(b, r, i); return pk(-i, r); }
__device__ __forceinline__ c32 neg2(c32 b) { float r, i; upk(b, r, i); return pk(-r, -i); }
__device__ __forceinline__ c32 swp(c32 b) { float r, i; upk(b, r, i); return pk(i, r); }
__device__ __forceinline__ c32 bc(float s) { return pk(s, s); }

// butterfly pieces used by the generated FFT (tools/gen_fft32.py)
__device__ __forceinline__ c32 c_add(c32 a, c32 b) { return add2(a, b); }
__device__ __forceinline__ c32 c_sub(c32 a, c32 b) { return sub2(a, b); }
__device__ __forceinline__ c32 c_add_mj(c32 a, c32 b) { return add2(a, mj(b)); }          // a + (-i) b
__device__ __forceinline__ c32 c_sub_mj(c32 a, c32 b) { return add2(a, pj(b)); }          // a - (-i) b
__device__ __forceinline__ c32 c_mj_sub(c32 b) { return sub2(mj(b), b); }                 // (-i) b - b
__device__ __forceinline__ c32 c_fmac(float c, c32 s, c32 a) { return fma2(bc(c), s, a); }      // a + c s
__device__ __forceinline__ c32 c_fma_j(float w, c32 b, c32 t) { return fma2(bc(w), pj(b), t); } // t + w (i b)
__device__ __forceinline__ c32 c_two_minus(c32 a, c32 o) { return fma2(bc(2.0f), a, neg2(o)); } // 2a - o

}  // namespace assl
