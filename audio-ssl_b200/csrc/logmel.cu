// Fused log-mel front end: frame + periodic Hann + rFFT + |X|^2 + sparse mel + log, one kernel.
//
// Replaces MelSpectrogramLibrosa.__call__ (/root/reference/src/utils/utils.py:26-28) and
// extract_log_mel_spectrogram (/root/reference/src/utils/utils.py:43-49).
//
// Work decomposition
//   One persistent CTA per SM; every WARP is its own pipeline over frame pairs (pair p of the flattened
//   (clip, pair) list goes to warp p mod #warps (numbered across the SMs), so at any time the resident warps work on neighbouring frames and the
//   6.4x overlap between hop-160 frames is served by L2).  There is no CTA-wide barrier after the prologue:
//   the warps drift apart, so while some sit in the FFTs (FP32 pipe) others are in the transposition or
//   the filterbank (shared-memory pipe) instead of all of them queueing for the same unit.
//     * staging: the pair's waveform span (n_fft + hop samples) is brought into the warp's own buffer by a
//       bulk async copy (TMA, cp.async.bulk + the warp's mbarrier) issued by lane 0 as soon as the previous
//       span is in registers, so it lands behind a whole pair of compute; librosa's reflect padding only
//       touches the first/last pairs of a clip and is patched in by the same warp;
//     * the two real frames are the real and imaginary part of one 1024-point complex FFT laid out
//       32 x 32: lane n1 holds samples n1 + 32*n2 (stride-32, conflict-free LDS; when hop is a multiple
//       of 32 the second frame is the first shifted by hop/32 registers, so 37 loads feed both);
//     * all arithmetic is two-wide (f32x2.cuh: FFMA2/FADD2/FMUL2 on {re, im} register pairs): a
//       register-resident 32-point FFT (fft32p_gen.cuh, 194 instructions), the inter-stage twiddles,
//       a transpose through the warp's padded shared buffer (128-bit row stores, 64-bit column loads),
//       the second 32-point FFT, and the two-for-one unpack (bin k of lane l pairs with bin 1024-k of
//       lane 32-l: one shuffle per half) straight into the two power spectra;
//     * the power spectra go back to the warp's buffer and the same warp applies the filterbank in
//       "interval" form: the bins between two neighbouring band centres feed the falling edge of one
//       band and the rising edge of the next, so every power value is read once; a lane owns the same
//       few intervals for the whole launch, a second short pass adds the two halves of each band,
//       takes the log and writes [n_mels, T] row-major.
#include "common.cuh"
#include "fft32p_gen.cuh"

#include <cstdlib>

namespace assl {

constexpr int kMaxWarps = 16;                       // per CTA (128 registers x 512 threads = the register file)
constexpr int kExStride = 34;                       // float2 units: lane-major rows of 32 + 2 pad (272 B: 128-bit
                                                    // row stores of a quarter warp hit 8 distinct 16-byte bank groups)
constexpr int kExWarp = 32 * kExStride + 2;         // 1090 float2 = 8720 B per warp
constexpr int kScratch = 640;                       // float2 offset of the interval totals inside a warp's buffer
                                                    // (padded power reads stay below 640; kScratchEntries x 16 B fit behind)
static_assert(kScratch + 2 * kScratchEntries <= 32 * kExStride, "interval totals overflow the warp buffer");
constexpr int kTwStride = 10;                       // float2 units: per lane w^8, w^16, w^24, w^1..w^7 (w = W_1024^lane);
                                                    // 80-byte rows: an odd number of 16-byte groups, conflict-free

// cos / sin of 2 pi n2 / 32, n2 = 0..16 (window phase of register n2, see the kernel prologue)
__device__ constexpr double kCosPhi[17] = {
    1.0, 0.98078528040323043, 0.92387953251128674, 0.83146961230254524, 0.70710678118654757,
    0.55557023301960218, 0.38268343236508978, 0.19509032201612825, 0.0, -0.19509032201612825,
    -0.38268343236508978, -0.55557023301960218, -0.70710678118654757, -0.83146961230254524,
    -0.92387953251128674, -0.98078528040323043, -1.0};
__device__ constexpr double kSinPhi[17] = {
    0.0, 0.19509032201612825, 0.38268343236508978, 0.55557023301960218, 0.70710678118654757,
    0.83146961230254524, 0.92387953251128674, 0.98078528040323043, 1.0, 0.98078528040323043,
    0.92387953251128674, 0.83146961230254524, 0.70710678118654757, 0.55557023301960218,
    0.38268343236508978, 0.19509032201612825, 0.0};

struct LogmelArgs {
  const float* wave;
  const float* gain;
  float* lms;
  double* pair_sums;  // [B, ppc, 2] (sum, sum^2) of what each frame pair wrote, or null
  long long L, ld, slot0, cap;
  int B, T, hop, n_mels, out_kind, bulk_ok;
  int pitch;          // floats per output row ([n_mels, pitch] slots; T for dense rows)
  int ppc;            // frame pairs per clip = ceil(T / 2)
  unsigned ppc_mul;   // n / ppc == (t + ((n - t) >> ppc_sh1)) >> ppc_sh2 with t = umulhi(ppc_mul, n)
  int ppc_sh1, ppc_sh2;
  int total_pairs;
  int wave_floats;    // per-warp staging buffer (n_fft + hop, rounded up to 4)
  int n_slots;        // rows of lane_iv in use
  const float2* tw;
  const int4* lane_iv;   // [kIvSlots][32] interval part owned by (slot, lane), see assl_plan::d_lane_iv
  const int4* band;      // [n_mels] {byte offset of U(m) in the warp's scratch, of D(m+1), enorm, 2.22e-16 * sum of the
                         //           band's weights (the "+ eps" of utils.py:28, folded)} (fp32 bits)
  unsigned* ticket;      // {next ticket, warps done}: pairs beyond the first two rounds are claimed, not dealt; the last
                         // warp to finish zeroes both (assl_plan::d_tickets, one pair per launch in flight); null = dealt
  int q_edge_lo, q_edge_hi;   // pairs q < q_edge_lo / q >= q_edge_hi of a clip touch the reflect padding (or the tail
                              // remainder of an L that is not a multiple of 4): their span is patched by the warp
  int span4;             // (n_fft + hop) & ~3: what one bulk copy covers for every other pair
  int deferred;          // assl_logmel_deferred_f32: does not wait for the grid before it at its start, but at its end
};

__host__ __device__ inline size_t logmel_smem_bytes(int warps, int wave_floats) {
  size_t b = 0;
  b += sizeof(float2) * warps * kExWarp;    // exchange / power / interval-sum buffers
  b += sizeof(float2) * 32 * kTwStride;     // twiddles
  b += sizeof(float) * warps * wave_floats; // staging
  b += sizeof(int4) * kMaxMels;             // per-band {scratch offsets of its two edges, enorm, folded eps}
  b += sizeof(int4) * kIvSlots * 32;        // interval ownership
  b += sizeof(unsigned long long) * (kMaxWarps + 32);  // one mbarrier per warp + the ticket counter's address, per lane
  return b;
}

// ---- mbarrier / bulk-copy helpers (sm_90+ PTX; SASS: SYNCS / UBLKCP)
__device__ __forceinline__ unsigned smem_u32(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(void* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(void* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(void* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ float ln_ftz(float x) {           // MUFU.LG2 * ln 2, like __logf minus its denormal pre-scaling
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y * 0.6931471805599453f;
}
// One ticket, drawn by the calling lane alone.  (atomicAdd() would be warp-aggregated by the compiler: vote + popc + one
// ATOMG + a SHFL that hands the base to the active lanes -- and that SHFL waits for the atomic's round trip on the spot,
// while the whole point here is to consume the value a pair of compute later.  ptxas does the same to a PTX atom on an
// address it can prove warp-uniform, so each lane reads the counter's address from its own shared-memory word.)
__device__ __forceinline__ unsigned draw_ticket(unsigned* counter) {
  unsigned t;
  asm volatile("atom.global.add.u32 %0, [%1], 1;" : "=r"(t) : "l"(counter) : "memory");
  return t;
}
__device__ __forceinline__ ulonglong2 lds128(unsigned addr) {
  ulonglong2 v;
  asm volatile("ld.shared.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "r"(addr) : "memory");
  return v;
}
// The waveform is read once (its 6.4x frame overlap is served while the neighbouring warps work on it): evict-first
// in L2, so that it does not push the log-mel ring out (the two-view kernel reads its bank entries from there).
#ifndef ASSL_WAVE_EVICT_FIRST
#define ASSL_WAVE_EVICT_FIRST 1
#endif
__device__ __forceinline__ unsigned long long l2_evict_first_policy() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar, unsigned long long pol) {
#if ASSL_WAVE_EVICT_FIRST
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol) : "memory");
#else
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
#endif
}

// Frame pair p of the flattened list -> clip and pair inside the clip (one multiply-high, see assl_logmel_f32).
struct PairPos {
  int clip, q;
};
__device__ __forceinline__ PairPos pair_pos(const LogmelArgs& a, int p) {
  PairPos s;
  const unsigned t = __umulhi(a.ppc_mul, static_cast<unsigned>(p));
  s.clip = static_cast<int>((t + ((static_cast<unsigned>(p) - t) >> a.ppc_sh1)) >> a.ppc_sh2);
  s.q = p - s.clip * a.ppc;
  return s;
}
// The part [lo, hi) (span-relative samples, multiples of 4) of a pair's span [base, base + span), base = 2 q hop -
// n_fft/2 clip-relative, that a bulk copy can cover.  Everything fits 32 bits: L < 2^31 - 4096 (assl_logmel_f32).
__device__ __forceinline__ void pair_cover(const LogmelArgs& a, int q, int span, int* lo, int* hi) {
  *lo = 0; *hi = a.span4;
  if (q < a.q_edge_lo || q >= a.q_edge_hi) {         // warp-uniform; ~5 of the 51 pairs of a 1-s clip
    const int base = 2 * q * a.hop - kNfft / 2;
    const int l = base < 0 ? -base : 0;
    int h = static_cast<int>(a.L) - base;
    h = (h > span ? span : h) & ~3;
    *lo = l; *hi = h > l ? h : l;
  }
  if (!a.bulk_ok) { *lo = 0; *hi = 0; }              // no bulk copy: the whole span goes through the patch loop
}

// H32 = hop / 32 when hop is a multiple of 32 (both frames of a pair share registers), else 0
template <int H32>
__global__ void __launch_bounds__(kMaxWarps * 32, 1) logmel_kernel(const LogmelArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int n_warps = blockDim.x >> 5;
  float2* s_ex = reinterpret_cast<float2*>(smem_raw);
  float2* s_tw = s_ex + n_warps * kExWarp;
  float* s_wave = reinterpret_cast<float*>(s_tw + 32 * kTwStride);
  int4* s_band = reinterpret_cast<int4*>(s_wave + n_warps * a.wave_floats);   // [n_mels]
  int4* s_lane = s_band + kMaxMels;                                          // [kIvSlots][32]
  unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_lane + kIvSlots * 32);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n_threads = blockDim.x;
  void* bar = s_bar + warp;                  // this warp's span has landed (tx bytes)
  float* buf = s_wave + warp * a.wave_floats;

  if (lane == 0) mbar_init(bar, 1);
  if (tid < 32) s_bar[kMaxWarps + tid] = reinterpret_cast<unsigned long long>(a.ticket);   // (see draw_ticket)
  const int span = kNfft + a.hop;                    // samples the two frames of a pair touch
  const int stride = static_cast<int>(gridDim.x) * n_warps;
  // producer side of the warp's pipeline (lane 0): bulk-copy the in-range part of a pair's span
  const unsigned long long wave_policy = l2_evict_first_policy();
  auto issue = [&](PairPos s) {
    int lo, hi;
    pair_cover(a, s.q, span, &lo, &hi);
    const unsigned bytes = static_cast<unsigned>(hi - lo) * 4u;
    mbar_expect_tx(bar, bytes);
    if (bytes) bulk_g2s(buf + lo, a.wave + (s.clip * a.ld + (2 * s.q * a.hop - kNfft / 2 + lo)), bytes, bar, wave_policy);
  };
  // Pair -> warp.  The first two rounds are dealt (warp-major numbering: consecutive pairs go to different SMs, so
  // that the resident warps work on neighbouring frames and the 6.4x overlap of hop-160 frames is served by L2);
  // every later pair is CLAIMED from a ticket counter, one atomic per pair fired a whole pair of compute before its
  // value is needed.  52224 pairs over 2368 warps are 22.05 rounds: dealt statically, the 23rd round (one warp per SM)
  // and every SM that got its CTA late set the kernel's duration; claimed, the warps all run out of work within one
  // pair of each other, and an SM that starts early (under the tail of the kernel before) simply takes more pairs.
  const int first = warp * static_cast<int>(gridDim.x) + static_cast<int>(blockIdx.x);
  // Launched behind another kernel of the path (PDL) this CTA gets its SM while that grid's last wave drains and
  // loads the plan's tables (immutable, never written by a kernel) meanwhile; the waveform, the ring and the sums
  // are touched only after pdl_wait() below.

  // Inter-stage twiddles w^k2 (w = W_1024^lane): only w^8, w^16, w^24 and w^1..w^7 are kept per lane; the
  // other 21 are applied as two rotations, (z * w^b) * w^(8a), trading 42 FFMA2/FMUL2 per transform for two
  // thirds of the table's shared-memory traffic
  for (int i = tid; i < 32 * kTwStride; i += n_threads) {
    const int n1 = i / kTwStride, e = i - n1 * kTwStride;
    const int k2 = e < 3 ? 8 * (e + 1) : e - 2;
    s_tw[i] = a.tw[k2 * 32 + n1];
  }
  // Window (periodic Hann, times the 1/2 of the two-frame unpack) without a table: for n = lane + 32 n2
  //   w = 1/4 - 1/4 cos(theta + phi),  theta = 2 pi lane / 1024 (per lane),  phi = 2 pi n2 / 32 (compile time)
  //     = 1/4 + wa cos(phi) + wb sin(phi),  wa = -1/4 cos(theta),  wb = 1/4 sin(theta)
  float wa, wb;
  sincospif(static_cast<float>(lane) * (1.0f / 512.0f), &wb, &wa);
  wa *= -0.25f; wb *= 0.25f;
  for (int i = tid; i < a.n_mels; i += n_threads) s_band[i] = a.band[i];
  for (int i = tid; i < kIvSlots * 32; i += n_threads) s_lane[i] = a.lane_iv[i];
  // (s_ex needs no clearing: every power bin a part reads is either written by this pair or selected away,
  //  never multiplied, and every interval total is started before it is read)
  // Deferred (assl_logmel_deferred_f32): the grid launched before this one writes nothing this grid READS (the
  // caller's promise), so a warp loads and transforms its first pair the moment its CTA has an SM -- under that grid's
  // tail -- and waits only before its first STORE (below): the slots and sums it overwrites may still be in use there.
  // (the trigger comes after the wait -- here, or with a warp's first store -- so that a kernel launched behind this one
  //  knows every older grid has completed: that is what lets a deferred two-view launch, assl_two_views_deferred_f32,
  //  start without waiting for this grid)
  if (!a.deferred) { pdl_wait(); pdl_launch_dependents(); }
  int cur = first, nxt = first + stride;             // the pair in work; the pair whose span is being copied
  PairPos pc = pair_pos(a, cur);
  unsigned ticket = 0;                               // lane 0: the claim in flight (becomes the pair after nxt)
  // the first copy flies while the other warps finish the tables (only this warp's lane 0 touches its barrier and buffer)
  if (lane == 0 && cur < a.total_pairs) {
    if (a.bulk_ok) {
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      issue(pc);
    }
  }
  __syncthreads();                                   // the only CTA-wide barrier
  unsigned* const ticket_ctr = reinterpret_cast<unsigned*>(s_bar[kMaxWarps + lane]);
  if (lane == 0 && a.ticket && nxt < a.total_pairs) ticket = draw_ticket(ticket_ctr);

  c32* ex = reinterpret_cast<c32*>(s_ex + warp * kExWarp);
  const unsigned scratch_addr = smem_u32(ex + kScratch);
  unsigned it = 0;
  while (cur < a.total_pairs) {
    const int clip = pc.clip, wpair = pc.q;
    // ---- waveform span: wait for the bulk copy, patch what it could not cover (reflect padding of
    //      np.pad(mode='reflect'): p<0 -> -p, p>L-1 -> 2(L-1)-p; tail remainder when L % 4 != 0)
    if (a.bulk_ok) mbar_wait(bar, it & 1);
    if (wpair < a.q_edge_lo || wpair >= a.q_edge_hi || !a.bulk_ok) {   // warp-uniform
      int lo, hi;
      pair_cover(a, wpair, span, &lo, &hi);
      const int base = 2 * wpair * a.hop - kNfft / 2;
      const int last = static_cast<int>(a.L - 1);
      const int n_fix = lo + (span - hi);
      const float* src = a.wave + clip * a.ld;
      for (int q = lane; q < n_fix; q += 32) {
        const int i = q < lo ? q : hi + (q - lo);
        long long p = static_cast<long long>(base) + i;
        if (p < 0) p = -p;
        if (p > last) p = 2ll * last - p;
        p = p < 0 ? 0 : (p > last ? last : p);       // only frames >= T can land here; never read back
        buf[i] = __ldcg(src + p);                    // (L2, like the bulk copies: a deferred launch has made no acquire yet)
      }
      __syncwarp();
    }

    c32 z[32];                                       // {frame 2q, frame 2q+1} = {re, im} of one complex transform
    {
      const float* w0 = buf + lane;
      float s[32 + (H32 > 0 ? H32 : 32)];              // frame 2q: s[n2]; frame 2q+1: s[n2 + H32] (or the second half)
      constexpr int kB = H32 > 0 ? H32 : 32;
#pragma unroll
      for (int j = 0; j < 32 + (H32 > 0 ? H32 : 0); ++j) s[j] = w0[32 * j];
      if (H32 == 0) {
        const float* w1 = w0 + a.hop;
#pragma unroll
        for (int j = 0; j < 32; ++j) s[32 + j] = w1[32 * j];
      }
      // (the asm keeps the compiler from hoisting 32 window values out of the pair loop and spilling them)
      asm volatile("" : "+f"(wa), "+f"(wb));
      {
        const float wl = 0.25f + wa, wh = 0.25f - wa;
        z[0] = pk(wl * s[0], wl * s[kB]);
        z[16] = pk(wh * s[16], wh * s[16 + kB]);
      }
#pragma unroll
      for (int n2 = 1; n2 < 16; ++n2) {              // phi and 2 pi - phi share the cosine
        const float c = static_cast<float>(kCosPhi[n2]), sn = static_cast<float>(kSinPhi[n2]);
        const float t = fmaf(wa, c, 0.25f);
        const float wl = fmaf(wb, sn, t), wh = fmaf(wb, -sn, t);
        z[n2] = pk(wl * s[n2], wl * s[n2 + kB]);
        z[32 - n2] = pk(wh * s[32 - n2], wh * s[32 - n2 + kB]);
      }
    }
    // the span is in registers: the next pair's copy may overwrite the buffer and lands behind this pair's compute
    __syncwarp();
    const PairPos pn = pair_pos(a, nxt);
    if (a.bulk_ok && lane == 0 && nxt < a.total_pairs) issue(pn);
    // the pair after that: the ticket drawn one pair ago has long arrived; draw the next one
    int nn = nxt + stride;
    if (a.ticket) {
      nn = 2 * stride + static_cast<int>(__shfl_sync(0xffffffffu, ticket, 0));
      if (lane == 0 && nn < a.total_pairs) ticket = draw_ticket(ticket_ctr);
    }

    fft32p(z);                                       // over n2 -> k2, this lane is n1
    {   // inter-stage twiddles W^(n1 k2), k2 = 8a + b: z * t = z * t.re + (i z) * t.im
      const ulonglong2* tl = reinterpret_cast<const ulonglong2*>(s_tw + lane * kTwStride);
      c32 t[10];                                     // w^8, w^16, w^24, w^1 .. w^7
#pragma unroll
      for (int q = 0; q < 5; ++q) { const ulonglong2 v = tl[q]; t[2 * q] = v.x; t[2 * q + 1] = v.y; }
      auto rot = [](c32 v, c32 w) { return fma2(pj(v), bc(hi_of(w)), mul2(v, bc(lo_of(w)))); };
#pragma unroll
      for (int b = 1; b < 8; ++b) {
#pragma unroll
        for (int g = 0; g < 4; ++g) z[8 * g + b] = rot(z[8 * g + b], t[2 + b]);
      }
#pragma unroll
      for (int g = 1; g < 4; ++g) {
#pragma unroll
        for (int b = 0; b < 8; ++b) z[8 * g + b] = rot(z[8 * g + b], t[g - 1]);
      }
    }
    __syncwarp();                                    // the previous pair's interval sums have been read
    {
      ulonglong2* row = reinterpret_cast<ulonglong2*>(ex + lane * kExStride);
#pragma unroll
      for (int q = 0; q < 16; ++q) row[q] = make_ulonglong2(z[2 * q], z[2 * q + 1]);
    }
    __syncwarp();
#pragma unroll
    for (int n1 = 0; n1 < 32; ++n1) z[n1] = ex[n1 * kExStride + lane];
    fft32p(z);                                       // over n1 -> k1, this lane is k2: Z[lane + 32*k1]

    // unpack: X1[k] = (Z[k] + conj Z[N-k]) / 2, X2[k] = (Z[k] - conj Z[N-k]) / 2i  (the 1/2 sits in the window)
    // with b = Z[N-k]:  |X1|^2 = (re+br)^2 + (im-bi)^2,  |X2|^2 = (im+bi)^2 + (re-br)^2
    {
      const int partner = (32 - lane) & 31;
      const bool self = lane == 0;                   // bins 32 k1 pair with 32 (32 - k1): same lane, other register
      c32 pw2[16];
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) {
        const c32 src = self ? z[(32 - k1) & 31] : z[31 - k1];
        const c32 b = pk(__shfl_sync(0xffffffffu, lo_of(src), partner), __shfl_sync(0xffffffffu, hi_of(src), partner));
        const c32 u = add2(z[k1], b), v = swp(sub2(z[k1], b));
        pw2[k1] = fma2(u, u, mul2(v, v));
      }
      __syncwarp();                                  // everyone is done reading the exchange buffer
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) ex[lane + 32 * k1] = pw2[k1];
      if (lane == 0) {                               // Nyquist bin: its own partner; bin 513 is tap padding
        const c32 n2 = add2(z[16], z[16]);
        ex[512] = mul2(n2, n2);
        ex[513] = pk(0.f, 0.f);
      }
    }
    __syncwarp();                                    // both power spectra are in the warp's buffer

    // ---- filterbank, interval form.  Interval j = bins between the centres of bands j-1 and j; bin k there has
    //      weight enorm(j-1) (1 - t) in band j-1 and enorm(j) t in band j, t linear in k (plan.cu).  A lane owns
    //      one part (kPartBins bins) of an interval per slot: S0 = sum P, S1 = sum r P over the part's own bins,
    //      U = sum t P = alpha S1 + beta S0, D = S0 - U.  Interval totals {D(j), U(j)} live behind the power
    //      spectra, one 16-byte entry each at a plan-chosen position; band m = enorm(m) (U(m) + D(m+1)).
    //      (Bins of a part that belong to the neighbouring interval are zeroed by selects, R2P + 2 SEL per bin.  Guarding
    //      the adds instead -- ``if``, or @p in PTX -- buys nothing: ptxas turns either into FADD2 + 2 MOV/SEL.)
    for (int s = 0; s < a.n_slots; ++s) {
      const int4 iv = s_lane[s * 32 + lane];         // the same interval parts for the whole launch
      const unsigned pv = smem_u32(ex + (iv.x & 0x3ff));
      const int vm = iv.x >> 10;                     // bits 0..7: which of the part's bins belong to the interval
      c32 s0a = pk(0.f, 0.f), s0b = s0a, s1a = s0a, s1b = s0a;   // {frame 2q, frame 2q+1}; even / odd bins apart
#pragma unroll
      for (int j = 0; j < kPartBins / 2; ++j) {
        const ulonglong2 p = lds128(pv + 16 * j);     // (kept whole: the plan's dealing of parts to quarter-warps makes
                                                      //  the 128-bit form conflict-free, not the 64-bit halves)
        const c32 x0 = (vm & (1 << (2 * j))) ? p.x : 0ull, x1 = (vm & (2 << (2 * j))) ? p.y : 0ull;   // foreign bins -> 0
        s0a = add2(s0a, x0); if (j) s1a = fma2(bc(static_cast<float>(2 * j)), x0, s1a);
        s0b = add2(s0b, x1); s1b = fma2(bc(static_cast<float>(2 * j + 1)), x1, s1b);
      }
      if (iv.x & (1 << 27)) {                        // parts of one interval sit in different slots: fixed order
        const c32 S0 = add2(s0a, s0b), S1 = add2(s1a, s1b);
        c32 su = fma2(bc(__int_as_float(iv.y)), S1, mul2(bc(__int_as_float(iv.z)), S0));
        c32 sd = sub2(S0, su);
        const unsigned tot = scratch_addr + (static_cast<unsigned>(iv.x >> 14) & 0xff0u);   // 16-byte entry (iv.x >> 18) & 0xff
        if (!(iv.x & (1 << 26))) { const ulonglong2 o = lds128(tot); sd = add2(o.x, sd); su = add2(o.y, su); }
        asm volatile("st.shared.v2.u64 [%0], {%1, %2};" ::"r"(tot), "l"(sd), "l"(su) : "memory");
      }
      __syncwarp();
    }

    if (a.deferred && it == 0) {                     // (warp-uniform) nothing global is written before the grid before
      pdl_wait();                                    //  this one has completed
      pdl_launch_dependents();                       // (the CTA's first call counts)
    }
    // ---- bands: add the two edges, fold eps and gain, log, store, statistics
    const int ta = 2 * wpair;
    const bool two = ta + 1 < a.T;                   // the last pair of an odd-length clip has one frame
    long long slot = a.slot0 + clip;
    if (a.cap > 0 && slot >= a.cap) slot -= a.cap;
    const bool power = a.out_kind & 1, tmajor = a.out_kind & 2;      // [n_mels, T] rows, or time-major [T, n_mels]
    const int band_stride = tmajor ? 1 : a.pitch, frame_stride = tmajor ? a.n_mels : 1;
    float* dst = a.lms + slot * (static_cast<long long>(a.n_mels) * a.pitch) + (tmajor ? ta * a.n_mels : ta) + lane * band_stride;
    float g2 = 1.f;
    if (a.gain) { const float g = a.gain[clip]; g2 = g * g; }   // |g x|^2 = g^2 |x|^2: per-clip gain applied here
    float s1 = 0.f, s2 = 0.f;
    for (int m = lane; m < a.n_mels; m += 32, dst += 32 * band_stride) {
      const int4 bd = s_band[m];                     // {offset of U(m), offset of D(m+1), enorm, folded eps}
      c32 su, sd;                                    // rising edge + falling edge
      asm volatile("ld.shared.u64 %0, [%1];" : "=l"(su) : "r"(scratch_addr + bd.x) : "memory");
      asm volatile("ld.shared.u64 %0, [%1];" : "=l"(sd) : "r"(scratch_addr + bd.y) : "memory");
      float acc0, acc1;
      upk(add2(su, sd), acc0, acc1);
      const float e = __int_as_float(bd.z) * g2, eps = __int_as_float(bd.w);
      acc0 = fmaf(acc0, e, eps);
      acc1 = fmaf(acc1, e, eps);
      const float l0 = power ? acc0 : ln_ftz(acc0 + kLogEps);      // the argument is >= 1.19e-7: no denormal path needed
      const float l1 = power ? acc1 : ln_ftz(acc1 + kLogEps);
      dst[0] = l0;
      s1 += l0; s2 = fmaf(l0, l0, s2);
      if (two) { dst[frame_stride] = l1; s1 += l1; s2 = fmaf(l1, l1, s2); }
    }
    if (a.pair_sums) {                               // per frame pair (sum, sum^2), fixed butterfly order
      // The first step splits the work: the lower half-warp goes on with the sums, the upper one with the sums of
      // squares (each lane hands over the value it does not keep), so the remaining four steps move ONE value per lane --
      // five shuffles instead of ten, and the same additions in the same order as a butterfly over both values.
      const bool upper = lane >= 16;
      float keep = upper ? s2 : s1;
      keep += __shfl_xor_sync(0xffffffffu, upper ? s1 : s2, 16);
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) keep += __shfl_xor_sync(0xffffffffu, keep, o);
      if ((lane & 15) == 0) a.pair_sums[2 * cur + (lane >> 4)] = static_cast<double>(keep);   // lane 0: sum, lane 16: sum^2
    }
    cur = nxt; nxt = nn; pc = pn; ++it;
  }
  // the ticket counter is handed back zeroed: every warp counts itself out after its last claim has returned, the
  // last one resets (the next launch on this counter is many launches away, see assl_plan::d_tickets)
  if (a.ticket && lane == 0) {
    const unsigned total_warps = static_cast<unsigned>(stride);
    if (atomicAdd(a.ticket + 1, 1u) == total_warps - 1) { a.ticket[0] = 0u; a.ticket[1] = 0u; }
  }  // deferred launch: whoever waits for THIS grid to complete must also get the completion of the grid before it; every
  // warp that had a pair has waited for it already, this covers a launch too small to give CTA 0 one
  if (a.deferred && blockIdx.x == 0 && warp == 0) pdl_wait();
}

// gain[b] = 1 / max(||wave_b||_2, 1e-12): F.normalize(p=2), upstream_dataset.py:62-63
__global__ void __launch_bounds__(256) wave_l2_gain_kernel(const float* wave, long long L, long long ld, float* gain) {
  __shared__ float s_red[8];
  const float* src = wave + blockIdx.x * ld;
  float acc = 0.f;
  for (long long i = threadIdx.x; i < L; i += 256) { const float v = src[i]; acc = fmaf(v, v, acc); }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int w = 0; w < 8; ++w) t += s_red[w];
    gain[blockIdx.x] = 1.0f / fmaxf(sqrtf(t), 1e-12f);
  }
}

template <int H32>
int launch_logmel(const LogmelArgs& a, size_t smem, unsigned grid, int warps, cudaStream_t st) {
  if (cudaFuncSetAttribute(logmel_kernel<H32>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  if (launch_pdl(logmel_kernel<H32>, dim3(grid), dim3(warps * 32), smem, st, a) != cudaSuccess) return ASSL_ERR_CUDA;
  return launch_status();
}

}  // namespace assl

extern "C" int64_t assl_plan_num_tiles(const assl_plan* p, int64_t n_samples) {
  const int64_t t = assl_plan_num_frames(p, n_samples);
  if (t < 0) return -1;
  return (t + 1) / 2;                 // one (sum, sum^2) entry per frame pair
}

extern "C" int assl_logmel_f32(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B,
                               int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap,
                               double* tile_sums, int32_t out_kind, void* stream) {
  return assl_logmel_pitched_f32(p, wave, wave_gain, B, L, ld, lms, slot0, cap, 0, tile_sums, out_kind, stream);
}

namespace assl {
static int logmel_launch(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B, int64_t L, int64_t ld,
                         float* lms, int64_t slot0, int64_t cap, int64_t pitch, double* tile_sums, int32_t out_kind,
                         int deferred, void* stream);
}
extern "C" int assl_logmel_pitched_f32(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B,
                                       int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap, int64_t pitch,
                                       double* tile_sums, int32_t out_kind, void* stream) {
  return assl::logmel_launch(p, wave, wave_gain, B, L, ld, lms, slot0, cap, pitch, tile_sums, out_kind, 0, stream);
}
extern "C" int assl_logmel_deferred_f32(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B,
                                        int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap, int64_t pitch,
                                        double* tile_sums, int32_t out_kind, void* stream) {
  return assl::logmel_launch(p, wave, wave_gain, B, L, ld, lms, slot0, cap, pitch, tile_sums, out_kind, 1, stream);
}

static int assl::logmel_launch(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B, int64_t L, int64_t ld,
                               float* lms, int64_t slot0, int64_t cap, int64_t pitch, double* tile_sums, int32_t out_kind,
                               int deferred, void* stream) {
  using namespace assl;
  if (out_kind < 0 || out_kind > 3) return ASSL_ERR_INVALID_ARG;
  if (!p || !wave || !lms || B < 0 || ld < L || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (L <= p->cfg.n_fft / 2) return ASSL_ERR_INVALID_ARG;       // np.pad(reflect) needs L > n_fft/2
  if (cap > 0 && (B > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (tile_sums && reinterpret_cast<uintptr_t>(tile_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  const int64_t T = 1 + L / p->cfg.hop;
  if (pitch == 0) pitch = T;
  if (pitch < T || pitch > (1ll << 30) || ((out_kind & 2) && pitch != T)) return ASSL_ERR_INVALID_ARG;   // time-major slots are dense
  LogmelArgs a;
  a.pitch = static_cast<int>(pitch);
  a.ppc = static_cast<int>((T + 1) / 2);
  const int64_t total_pairs = B * a.ppc;
  if (T > (1ll << 24) || total_pairs > (1ll << 30) || L > (1ll << 31) - 4096) return ASSL_ERR_UNSUPPORTED;
  a.wave = wave; a.gain = wave_gain; a.lms = lms; a.pair_sums = tile_sums;
  a.B = static_cast<int>(B); a.L = L; a.ld = ld; a.slot0 = slot0; a.cap = cap;
  a.T = static_cast<int>(T); a.hop = p->cfg.hop; a.n_mels = p->cfg.n_mels;
  a.total_pairs = static_cast<int>(total_pairs);
  a.wave_floats = (kNfft + a.hop + 3) & ~3;
  a.n_slots = p->n_slots; a.out_kind = out_kind; a.deferred = deferred;
  {   // n / ppc by multiply-high (Granlund-Montgomery, exact for every 32-bit n)
    const unsigned d = static_cast<unsigned>(a.ppc);
    int l = 0;
    while ((1ull << l) < d) ++l;
    a.ppc_mul = static_cast<unsigned>(((1ull << 32) * ((1ull << l) - d)) / d + 1);
    a.ppc_sh1 = l < 1 ? l : 1;
    a.ppc_sh2 = l > 1 ? l - 1 : 0;
  }
  // the bulk (TMA) copy needs 16-byte aligned global addresses: clip starts and span starts (2*q*hop - 512)
  a.bulk_ok = (reinterpret_cast<uintptr_t>(wave) % 16 == 0) && (ld % 4 == 0) && (p->cfg.hop % 2 == 0);
  a.tw = p->d_twiddle;
  a.lane_iv = p->d_lane_iv; a.band = p->d_band;
  {   // which pairs of a clip touch the reflect padding / the unaligned tail (pair_cover)
    const int64_t span = kNfft + a.hop;
    a.span4 = static_cast<int>(span & ~3ll);
    a.q_edge_lo = (kNfft / 2 + 2 * a.hop - 1) / (2 * a.hop);                 // 2 q hop < n_fft / 2
    const int64_t room = L - span + kNfft / 2;                                // 2 q hop <= room: the span ends inside the clip
    a.q_edge_hi = (span % 4 != 0 || room < 0) ? 0 : static_cast<int>(room / (2 * a.hop) + 1);
  }
  // as many warps per CTA (one CTA per SM) as shared memory allows
  int warps = kMaxWarps;
  if (const char* e = getenv("ASSL_LOGMEL_WARPS")) { const int w = atoi(e); if (w >= 1 && w <= kMaxWarps) warps = w; }
  while (warps > 1 && logmel_smem_bytes(warps, a.wave_floats) > 227 * 1024) --warps;
  const size_t smem = logmel_smem_bytes(warps, a.wave_floats);
  if (smem > 227 * 1024) return ASSL_ERR_UNSUPPORTED;
  const int64_t ctas = (total_pairs + warps - 1) / warps;
  const unsigned grid = static_cast<unsigned>(ctas < p->sm_count ? ctas : p->sm_count);
  // pairs beyond the first two rounds are claimed from a ticket counter (ASSL_LOGMEL_TICKETS=0: dealt, for A/B runs).
  // Counters are handed out round-robin; a launch leaves its counter zeroed, and kTicketSlots launches would have to be
  // in flight at once for two of them to share one.
  a.ticket = nullptr;
  static const bool tickets_on = [] { const char* e = getenv("ASSL_LOGMEL_TICKETS"); return !(e && e[0] == '0'); }();
  if (tickets_on && total_pairs > 2ll * grid * warps)
    a.ticket = p->d_tickets + 2 * (p->ticket_seq.fetch_add(1u, std::memory_order_relaxed) % assl_plan::kTicketSlots);
  if (a.hop == 160) return launch_logmel<5>(a, smem, grid, warps, as_stream(stream));
  return launch_logmel<0>(a, smem, grid, warps, as_stream(stream));
}

extern "C" int assl_wave_l2_gain_f32(const float* wave, int64_t B, int64_t L, int64_t ld, float* gain, void* stream) {
  if (!wave || !gain || B < 0 || L < 1 || ld < L) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::wave_l2_gain_kernel<<<static_cast<unsigned>(B), 256, 0, assl::as_stream(stream)>>>(wave, L, ld, gain);
  return assl::launch_status();
}
