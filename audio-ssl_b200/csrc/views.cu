// Fused two-view generation: pre-normalise -> log-mixup-exp against a bank entry -> zero-canvas
// crop -> bicubic / bilinear resample, one CTA per (clip, view, tile of output frames).
//
// Replaces the body of AugmentationModule.__call__ (/root/reference/src/augmentations/__init__.py:32-35):
//   RunningNorm.forward's (x - mean) / std            augmentations.py:282
//   MixupBYOLA.forward / log_mixup_exp                augmentations.py:99-106, 8-12
//   RandomResizeCrop.forward                          augmentations.py:40-55
// with every random draw (alpha, bank entry, crop box) passed in.
//
// The resampler follows ATen's upsample_bicubic2d / upsample_bilinear2d with align_corners=True
// (what F.interpolate reaches at augmentations.py:53-54): source coordinate dst * (in-1)/(out-1)
// in fp32, A = -0.75 cubic convolution weights, taps clamped to the CROP, rows nested over columns.
// It is evaluated separably: stage A mixes the needed crop span once into shared memory (3 MUFU
// per source pixel instead of per tap; 128-bit loads from the ring), stage B resamples along time, stage C along frequency.
#include "common.cuh"

namespace assl {

constexpr int kViewThreads = 256;
constexpr int kViewWarps = kViewThreads / 32;

struct ViewArgs {
  assl_view_cfg c;
  const float* lms;
  const float* mean_std;
  long long cap, B;
  const int* clip_slot;
  const int* mix_slot;
  const float* alpha;
  const int* crop;
  const int* order;   // optional [2B]: launch position -> view id (2 * clip + view), heaviest views first
  // global-bank mode (assl_two_views_global_f32): bank entries may sit in a peer GPU's ring, statistics in a ring
  // indexed by global clip; all null / unused otherwise
  const int* mix_rank;    // [B,2] whose ring holds the bank entry
  const int* clip_stat;   // [B]   slot of the clip's (mean, std)
  const int* mix_stat;    // [B,2] slot of the bank entry's (mean, std)
  const float* peer_lms[ASSL_EXCHANGE_MAX_WORLD];   // every rank's ring as mapped into this process
  float* view1;
  float* view2;
  int tiles_t;     // output-time tiles per view
  int src_w;       // floats per staged row (a multiple of 4)
  int pitch;       // floats per ring row (>= n_time; a multiple of 4 on the vector path)
  int y0, x0;      // where the clip sits in the canvas (augmentations.py:47)
  int deferred;    // assl_two_views_deferred_f32: may overlap the grid launched before it (waits for it at the END)
};

__device__ __forceinline__ float ex2_approx(float x) {    // MUFU.EX2, flush-to-zero (|rel err| <= 2 ulp)
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ float lg2_approx(float x) {    // MUFU.LG2 without the denormal pre-scaling (callers pass >= 1.19e-7)
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// The views are written once and read by another kernel much later: streaming stores (evict-first in L2) keep them
// from pushing the log-mel ring -- the bank entries the next steps read -- out of the 126 MB L2.
#ifndef ASSL_VIEWS_STREAM_STORES
#define ASSL_VIEWS_STREAM_STORES 1
#endif
__device__ __forceinline__ void store_view(float* p, float v) {
#if ASSL_VIEWS_STREAM_STORES
  __stcs(p, v);
#else
  *p = v;
#endif
}

// ---- mbarrier + bulk async copy (TMA; SASS: SYNCS / UBLKCP): a warp brings the bank entry's rows into its staged rows
__device__ __forceinline__ unsigned vsmem(const void* p) { return static_cast<unsigned>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void vbar_init(void* bar) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(vsmem(bar)) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void vbar_expect(void* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(vsmem(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void vbar_wait(void* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "VWAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra VDONE_%=;\n\t"
      "bra VWAIT_%=;\n\t"
      "VDONE_%=:\n\t}" ::"r"(vsmem(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_row(void* dst, const void* src, unsigned bytes, void* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(vsmem(dst)), "l"(src), "r"(bytes), "r"(vsmem(bar)) : "memory");
}

__device__ __forceinline__ float cubic1(float x) {   // |x| <= 1
  const float A = -0.75f;
  return ((A + 2.f) * x - (A + 3.f)) * x * x + 1.f;
}
__device__ __forceinline__ float cubic2(float x) {   // 1 < |x| < 2
  const float A = -0.75f;
  return ((A * x - 5.f * A) * x + 8.f * A) * x - 4.f * A;
}

// ATen area_pixel_compute_scale (align_corners=True) + guard_index_and_lambda
__device__ __forceinline__ void source_index(int dst, int n_in, int n_out, int* idx, float* lam) {
  const float scale = n_out > 1 ? static_cast<float>(n_in - 1) / static_cast<float>(n_out - 1) : 0.f;
  const float real = scale * static_cast<float>(dst);
  int i = static_cast<int>(floorf(real));
  i = i < n_in - 1 ? i : n_in - 1;
  *idx = i;
  *lam = fminf(fmaxf(real - static_cast<float>(i), 0.f), 1.f);
}

__device__ __forceinline__ float4 tap_weights(float t, int bilinear) {
  if (bilinear) return make_float4(0.f, 1.f - t, t, 0.f);
  return make_float4(cubic2(t + 1.f), cubic1(t), cubic1(1.f - t), cubic2(2.f - t));
}

// shared memory: frequency taps [F] (weights, row byte offsets), time taps [TW] (weights, first-tap byte offset), rows
__host__ __device__ inline size_t views_smem_bytes(int canvas_h, int src_w, int n_freq, int tw) {
  return sizeof(float) * static_cast<size_t>(canvas_h) * src_w + sizeof(float4) * static_cast<size_t>(n_freq) +
         sizeof(int) * static_cast<size_t>((n_freq + 3) & ~3) + (sizeof(float4) + sizeof(int)) * static_cast<size_t>(tw) + sizeof(unsigned long long) * kViewWarps;   // + one mbarrier per warp
}

// What a CTA knows about its view, tile and staged rows.  A staged row holds the crop columns xs_lo .. xs_hi of one crop
// row, xs_lo = (first tap of the tile's first output column) = ix_lo - 1 and xs_hi = ix_hi + 2 UNCLAMPED, so that the four
// taps of an output column are four consecutive floats; the (at most one + two) slots that fall outside the crop hold
// copies of its first / last column (ATen clamps the taps to the crop).  Slot u of a row = crop column xs_lo + u - xo =
// clip column (d4 + u), where the shift xo in 0..3 makes d4 a multiple of 4: 16-byte groups of slots are 16-byte groups
// of the ring row, and one LDG.128 -> math -> STS.128 moves four columns.
struct Geo {
  int h, w, ci, cj;      // the crop box, clamped into the canvas
  int ox0, ncol;         // this tile's output columns
  int xs_lo, xo, nslots; // staged span
  int d4;                // clip column of slot 0 (multiple of 4, may be negative)
  int j_in0, j_in1;      // float4 groups of a row that overlap the clip's columns [0, T)
  int n4;                // float4 groups per staged row in use
};

// The per-view parameters, all loaded in ONE hop right after the view id is known (they were four dependent hops:
// order -> slots -> statistics -> rows); the (mean, std) pairs follow while the geometry is worked out.
struct Sel {
  const float* xp;       // the clip's log-mel in the ring
  const float* zp;       // the bank entry's (this ring or a peer's)
  float mu, sd, zmu, zsd, al;
  bool mix;              // there is a bank entry
};

// ---- stage A for one warp's rows: normalise + mix into shared memory.
//      exp(xn) = ex2(x*a + b) with a = log2e/std, b = -mean*log2e/std: one FFMA + one MUFU per operand.
//      LOGMIX (log-mixup-exp against a bank entry, the path's normal case) is a separate instantiation; the other
//      modes (no bank entry yet, plain mixup) share the generic one.  G = 32-lane groups of float4 the in-clip part of a
//      row needs (1 for clips up to 125 frames per tile span, else 2).  VEC: ring rows are 16-byte aligned (pitch % 4 == 0).
//      BULK (aligned rows of this GPU's ring): the bank entry's in-clip row segments are brought into the staged rows -- the
//      very slots the results go to -- by bulk async copies, one per row, each issued by a different lane, all of a warp's
//      rows in flight at once on the warp's mbarrier; the clip's own rows come through registers, four rows deep; the mix
//      happens in place.
template <int G, bool LOGMIX, bool VEC, bool BULK, typename Mid>
__device__ __forceinline__ void stage_rows(const ViewArgs& a, const Geo& g, const Sel& sel, float* s_src, void* bar, int lane, int warp,
                                           Mid mid) {
  const int F = a.c.n_freq, T = a.c.n_time;
  const float* xp = sel.xp;
  const float* zp = sel.zp;
  const bool mix = LOGMIX || sel.mix;
  const float inv = 1.f / sel.sd, zinv = 1.f / sel.zsd;
  const float wx = 1.f - sel.al, wz = sel.al;
  const float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  const float xa = LOGMIX ? inv * kLog2e : inv, xb = -sel.mu * xa;
  const float za = LOGMIX ? zinv * kLog2e : zinv, zb = -sel.zmu * za;

  // this lane's float4 groups: j = j_in0 + lane + 32 q; their clip columns col .. col + 3.  Everything that does not
  // depend on the row is worked out here, once: the row loop below only advances pointers.
  int col[G];
  bool on[G];
#pragma unroll
  for (int q = 0; q < G; ++q) {
    const int j = g.j_in0 + lane + 32 * q;
    on[q] = j < g.j_in1;
    col[q] = min(max(g.d4 + 4 * (on[q] ? j : g.j_in0), 0), VEC ? a.pitch - 4 : T - 1);   // (idle lanes: any address inside the row)
  }
  // zero canvas: the groups left and right of the clip's columns (all groups, in a canvas row above / below the clip)
  const bool row_a = lane < g.n4, row_b = lane + 32 < g.n4;                   // groups lane and lane + 32 exist
  const bool zero_a = row_a && (lane < g.j_in0 || lane >= g.j_in1);
  const bool zero_b = row_b && (lane + 32 < g.j_in0 || lane + 32 >= g.j_in1);
  // the crop's edge columns are copied into the slots just outside it (taps clamp to the crop) in one pass over the
  // warp's rows at the end: which slots, if any
  const int u_first = g.xo + (0 - g.xs_lo);                    // slot of crop column 0
  const int u_last = g.xo + (g.w - 1 - g.xs_lo);               // slot of crop column w - 1
  const int u_end = g.xo + g.nslots;                           // one past the last slot in use

  auto load4 = [&](const float* p, int c) -> float4 {          // (dense rows: c = the group's first column)
    if (VEC) return __ldcg(reinterpret_cast<const float4*>(p));
    float4 r;
    r.x = __ldcg(p);
    r.y = c + 1 < T ? __ldcg(p + 1) : 0.f;
    r.z = c + 2 < T ? __ldcg(p + 2) : 0.f;
    r.w = c + 3 < T ? __ldcg(p + 3) : 0.f;
    return r;
  };
  auto mix1 = [&](float x, float z) -> float {
    const float xn = fmaf(x, xa, xb);
    if (LOGMIX) {
      const float zn = fmaf(z, za, zb);
      return kLn2 * lg2_approx(fmaf(wx, ex2_approx(xn), fmaf(wz, ex2_approx(zn), kLogEps)));
    }
    if (mix) return fmaf(wz, fmaf(z, za, zb), wx * xn);
    return xn;
  };

  // Rows y = warp + 8 k, k < K: the same warp resamples them along time afterwards, so stage A -> stage B needs no
  // CTA-wide barrier.  Clip rows 8 k - r0 grow with k, so the warp's rows INSIDE the clip are one range [k0, k1) -- all of
  // them when the canvas is as high as the clip -- and the loop over it carries no per-row tests; canvas rows above /
  // below are zero-filled apart.  The loads run kDepth - 1 rows ahead of the arithmetic (kDepth register sets): a warp
  // waits for global memory once, not once per row.
  constexpr int kDepth = BULK ? (G == 1 ? 4 : 2) : (G == 1 ? 3 : 1);   // (without BULK two groups per lane leave no room
                                                                       //  for a second register set)
  const int K = warp < g.h ? (g.h - warp + kViewWarps - 1) / kViewWarps : 0;
  const int r0 = a.y0 - g.ci - warp;
  const int k0 = min(r0 > 0 ? (r0 + kViewWarps - 1) / kViewWarps : 0, K);
  const int k1 = min(max(F + r0 > 0 ? (F + r0 + kViewWarps - 1) / kViewWarps : 0, k0), K);
  for (int k = 0; k < K; ++k) {                                 // canvas rows (none in the reference's configuration)
    if (k >= k0 && k < k1) { k = k1 - 1; continue; }
    float4* row4 = reinterpret_cast<float4*>(s_src + (warp + kViewWarps * k) * a.src_w);
    if (row_a) row4[lane] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row_b) row4[lane + 32] = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  const bool zbulk = BULK && mix && k1 > k0 && g.j_in1 > g.j_in0;
  if (zbulk) {
    const unsigned nb = static_cast<unsigned>(g.j_in1 - g.j_in0) * 16u;       // the in-clip groups of a row
    if (lane == 0) vbar_expect(bar, nb * static_cast<unsigned>(k1 - k0));
    __syncwarp();
    for (int r = k0 + lane; r < k1; r += 32)
      bulk_row(s_src + (warp + kViewWarps * r) * a.src_w + 4 * g.j_in0,
               zp + (static_cast<long long>(kViewWarps * r - r0) * a.pitch + (g.d4 + 4 * g.j_in0)), nb, bar);
  }
  const float* xq[G];
  const float* zq[G];
#pragma unroll
  for (int q = 0; q < G; ++q) {
    const long long off = static_cast<long long>(kViewWarps * k0 - r0) * a.pitch + col[q];
    xq[q] = xp + off;
    zq[q] = zp + off;
  }
  const long long row_step = static_cast<long long>(kViewWarps) * a.pitch;
  float* srow = s_src + (warp + kViewWarps * k0) * a.src_w;
  const int srow_step = kViewWarps * a.src_w;
  float4 xv[kDepth][G], zv[BULK ? 1 : kDepth][G];
  auto fetch = [&](int set, int k) {                            // (no-op past the last row: that set is never emitted)
#pragma unroll
    for (int q = 0; q < G; ++q) {
      if (k < k1) {                                             // warp-uniform
        xv[set][q] = load4(xq[q], col[q]);
        if (!BULK && mix) zv[set][q] = load4(zq[q], col[q]);    // CTA-uniform (compile-time true for LOGMIX)
      }
      xq[q] += row_step;
      if (!BULK) zq[q] += row_step;
    }
  };
  auto emit = [&](int set) {
    float4* row4 = reinterpret_cast<float4*>(srow);
#pragma unroll
    for (int q = 0; q < G; ++q) {
      const float4 x = xv[set][q];
      float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
      if (BULK) { if (zbulk && on[q]) z = row4[g.j_in0 + lane + 32 * q]; }
      else if (mix) z = zv[set][q];
      float4 o;
      o.x = mix1(x.x, z.x);
      o.y = col[q] + 1 < T ? mix1(x.y, z.y) : 0.f;              // (the last group may reach past the clip's last column:
      o.z = col[q] + 2 < T ? mix1(x.z, z.z) : 0.f;              //  canvas)
      o.w = col[q] + 3 < T ? mix1(x.w, z.w) : 0.f;
      if (on[q]) row4[g.j_in0 + lane + 32 * q] = o;
    }
    if (zero_a) row4[lane] = make_float4(0.f, 0.f, 0.f, 0.f);   // canvas left and right of the clip's columns
    if (zero_b) row4[lane + 32] = make_float4(0.f, 0.f, 0.f, 0.f);
    srow += srow_step;
  };
#pragma unroll
  for (int d = 0; d < kDepth - 1; ++d) fetch(d, k0 + d);
  mid();                                                        // (the CTA's tap tables, while the first rows are in flight)
  if (zbulk) vbar_wait(bar, 0);                                 // the bank entry's rows have landed in the staged rows
  for (int k = k0; k < k1; k += kDepth) {                       // (unrolled by kDepth so that the register sets are static)
#pragma unroll
    for (int d = 0; d < kDepth; ++d) {
      if (k + d < k1) {                                         // warp-uniform
        fetch((d + kDepth - 1) % kDepth, k + d + kDepth - 1);
        emit(d);
      }
    }
  }
  __syncwarp();
  // edge slots: lane <-> row.  A copy of the crop's first / last column; where that column is canvas (or the whole row
  // is) it already holds the zero that belongs there.
  for (int r = lane; r < K; r += 32) {
    float* row = s_src + (warp + kViewWarps * r) * a.src_w;
    if (g.xs_lo < 0) row[u_first - 1] = row[u_first];
    if (u_last + 1 < u_end) {                                   // (a tile that ends inside the crop has no such slots)
      const float e = row[u_last];
      row[u_last + 1] = e;
      if (u_last + 2 < u_end) row[u_last + 2] = e;
    }
  }
  __syncwarp();
}

template <int G, bool VEC, bool BULK, typename Mid>
__device__ __forceinline__ void stage_rows_any(const ViewArgs& a, const Geo& g, const Sel& sel, float* s_src, void* bar, int lane,
                                               int warp, Mid mid) {
  const bool logmix = sel.mix && a.c.log_mixup;                 // CTA-uniform
  if (logmix) stage_rows<G, true, VEC, BULK>(a, g, sel, s_src, bar, lane, warp, mid);
  else stage_rows<G, false, VEC, BULK>(a, g, sel, s_src, bar, lane, warp, mid);
}

// TW = output frames per tile (32 / 64 / 96 / 128)
template <int TW, bool GB, bool VEC>
__global__ void __launch_bounds__(kViewThreads, 4) two_views_kernel(const ViewArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int F = a.c.n_freq, T = a.c.n_time;
  float4* s_wy = reinterpret_cast<float4*>(smem_raw);                // [F]   frequency tap weights
  int* s_top = reinterpret_cast<int*>(s_wy + F);                     // [F]   row of an output row's fourth frequency tap (padded to 4)
  float4* s_wx = reinterpret_cast<float4*>(s_top + ((F + 3) & ~3));  // [TW]  time tap weights of the tile's output columns
  int* s_xu = reinterpret_cast<int*>(s_wx + TW);                     // [TW]  byte offset of a column's first tap in a row
  unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_xu + TW);   // [8] one mbarrier per warp (bulk copies)
  float* s_src = reinterpret_cast<float*>(s_bar + kViewWarps);       // [h][src_w]  mixed crop span, then (in place,
                                                                     //             row by row) the time-resampled rows
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (VEC && !GB) {                                            // this warp's barrier: only this warp ever touches it
    if (lane == 0) vbar_init(s_bar + warp);
    __syncwarp();
  }
  unsigned blk = blockIdx.x;
  int tile = 0;
  if (a.tiles_t > 1) { tile = static_cast<int>(blk % static_cast<unsigned>(a.tiles_t)); blk /= static_cast<unsigned>(a.tiles_t); }
  pdl_launch_dependents();                     // the next kernel may take SMs as this grid's last wave drains
  if (!a.deferred) pdl_wait();                 // nothing global is read before the previous grid has completed
  if (a.order) blk = static_cast<unsigned>(__ldcg(a.order + blk));   // CTAs start in launch order: the host puts the big crops first
  const int v = static_cast<int>(blk & 1u);
  const long long b = blk >> 1;

  // ---- every per-view parameter in one hop ...
  // (every global read of this kernel goes to L2 -- ld.global.cg and bulk copies, never the SM's L1: a deferred launch
  //  starts without griddepcontrol.wait, i.e. without the acquire that would drop lines an earlier kernel left there)
  const int4 cr = __ldcg(reinterpret_cast<const int4*>(a.crop) + (b * 2 + v));
  const int xslot = __ldcg(a.clip_slot + b);
  const int zslot = __ldcg(a.mix_slot + (b * 2 + v));
  Sel sel;
  sel.al = __ldcg(a.alpha + (b * 2 + v));
  sel.mix = zslot >= 0;
  // GB (global bank, a separate instantiation so the default kernel keeps its register budget): the bank entry may
  // sit in a peer's ring (loads over NVLink) and the (mean, std) pairs live in a ring over global clips
  int xstat = xslot, zstat = zslot;
  const float* zbase = a.lms;
  if (GB) {
    if (a.clip_stat) { xstat = a.clip_stat[b]; zstat = a.mix_stat[b * 2 + v]; }
    if (a.mix_rank && sel.mix) zbase = a.peer_lms[a.mix_rank[b * 2 + v]];
  }
  {
    const long long plane = static_cast<long long>(F) * a.pitch;
    sel.xp = a.lms + xslot * plane;
    sel.zp = zbase + (sel.mix ? zslot : 0) * plane;
  }
  // ... then the two (mean, std) pairs, in flight while the geometry and the first row addresses are worked out
  sel.mu = 0.f; sel.sd = 1.f; sel.zmu = 0.f; sel.zsd = 1.f;
  if (a.mean_std) {
    const float2 xs = __ldcg(reinterpret_cast<const float2*>(a.mean_std) + xstat);
    sel.mu = xs.x; sel.sd = xs.y;
    if (sel.mix) { const float2 zs = __ldcg(reinterpret_cast<const float2*>(a.mean_std) + zstat); sel.zmu = zs.x; sel.zsd = zs.y; }
  }

  Geo g;
  {
    // the host mirror validates crops; the clamps only keep a bad box from leaving shared memory
    g.h = min(max(cr.z, 1), a.c.canvas_h); g.w = min(max(cr.w, 1), a.c.canvas_w);
    g.ci = min(max(cr.x, 0), a.c.canvas_h - g.h); g.cj = min(max(cr.y, 0), a.c.canvas_w - g.w);
  }
  const int bil = a.c.interp;
  g.ox0 = tile * TW;
  g.ncol = min(TW, T - g.ox0);
  {
    int ix_lo, ix_hi; float lam;
    source_index(g.ox0, g.w, T, &ix_lo, &lam);
    source_index(g.ox0 + g.ncol - 1, g.w, T, &ix_hi, &lam);
    g.xs_lo = ix_lo - 1;
    g.nslots = min(ix_hi + 2 - g.xs_lo + 1, a.src_w - 3);
    const int d0 = g.cj + g.xs_lo - a.x0;                            // clip column of the first slot
    g.xo = d0 & 3;
    g.d4 = d0 - g.xo;
    g.n4 = (g.xo + g.nslots + 3) >> 2;
    const int lo = g.d4 < 0 ? (-g.d4) >> 2 : 0;                      // first group at or right of clip column 0
    const int hi = g.d4 < T ? (T - g.d4 + 3) >> 2 : 0;               // one past the last group left of clip column T
    g.j_in0 = min(lo, g.n4);
    g.j_in1 = max(min(hi, g.n4), g.j_in0);
  }

  // ---- taps, once per CTA: frequency (shared by every column of the tile), time (shared by every row).  Built by every
  //      warp right after it has issued its first row loads (stage A calls this back), i.e. under their latency.
  auto tap_tables = [&]() {
    for (int oy = tid; oy < F; oy += kViewThreads) {
      int iy; float ty;
      source_index(oy, g.h, F, &iy, &ty);
      s_wy[oy] = tap_weights(ty, bil);
      s_top[oy] = iy + 2;                                  // the fourth tap's row, UNCLAMPED: stage C's position in the crop
    }
    for (int c = tid; c < TW; c += kViewThreads) {
      int ix; float tx;
      source_index(g.ox0 + min(c, g.ncol - 1), g.w, T, &ix, &tx);
      s_wx[c] = tap_weights(tx, bil);
      s_xu[c] = 4 * min(max(ix - 1 - g.xs_lo + g.xo, 0), a.src_w - 4);
    }
  };

  // ---- stage A (this warp's rows), with as many 32-lane groups of float4 as the in-clip part of THIS span needs
  constexpr bool kBulk = VEC && !GB;                            // (peer rings and dense rows: loads through registers)
  if (g.j_in1 - g.j_in0 <= 32) stage_rows_any<1, VEC, kBulk>(a, g, sel, s_src, s_bar + warp, lane, warp, tap_tables);
  else stage_rows_any<2, VEC, kBulk>(a, g, sel, s_src, s_bar + warp, lane, warp, tap_tables);
  __syncthreads();                             // (the tap tables; the rows themselves stay with their warp until stage C)

  // ---- stage B, along time: lanes <-> output columns, the four taps of a column are four consecutive floats of the row
  //      (1 address add + 4 LDS + 4 FMA per column), each row overwritten in place by the warp that owns it
  constexpr int NC = TW / 32;
  {
    float4 wt[NC];
    int xu[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) { wt[c] = s_wx[lane + 32 * c]; xu[c] = s_xu[lane + 32 * c]; }
    const unsigned src0 = static_cast<unsigned>(__cvta_generic_to_shared(s_src));
    for (int y = warp; y < g.h; y += kViewWarps) {
      const unsigned rowb = src0 + static_cast<unsigned>(y * a.src_w) * 4u;
      float acc[NC];
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const unsigned p = rowb + static_cast<unsigned>(xu[c]);
        float t0, t1, t2, t3;
        asm volatile("ld.shared.f32 %0, [%4];\n\tld.shared.f32 %1, [%4+4];\n\tld.shared.f32 %2, [%4+8];\n\tld.shared.f32 %3, [%4+12];"
                     : "=f"(t0), "=f"(t1), "=f"(t2), "=f"(t3) : "r"(p) : "memory");
        acc[c] = t0 * wt[c].x;
        acc[c] = fmaf(t1, wt[c].y, acc[c]);
        acc[c] = fmaf(t2, wt[c].z, acc[c]);
        acc[c] = fmaf(t3, wt[c].w, acc[c]);
      }
      __syncwarp();
      float* row = s_src + y * a.src_w;
#pragma unroll
      for (int c = 0; c < NC; ++c)
        if (lane + 32 * c < g.ncol) row[lane + 32 * c] = acc[c];
    }
  }
  __syncthreads();

  // ---- stage C, along frequency: lanes <-> columns, a warp takes a block of consecutive output rows, coalesced streaming
  //      stores.  The frequency scale (h - 1) / (F - 1) is at most 1 (the crop is not taller than the canvas of F rows
  //      unless virtual_crop_scale[0] > 1), so the four tap rows of consecutive output rows are the same or move down by
  //      one: a lane keeps them in registers and loads ONE new value per column when they move (a quarter of the
  //      shared-memory reads of four loads per output; the pipe that carries them is the busiest one in this kernel).
  {
    float* outp = (v == 0 ? a.view1 : a.view2) + b * (static_cast<long long>(F) * T) + g.ox0 + lane;
    const char* col0 = reinterpret_cast<const char*>(s_src + lane);
    bool live[NC];
#pragma unroll
    for (int c = 0; c < NC; ++c) live[c] = lane + 32 * c < g.ncol;
    const int rows_per_warp = (F + kViewWarps - 1) / kViewWarps;
    const int oy0 = warp * rows_per_warp, oy1 = min(F, oy0 + rows_per_warp);
    const int rb = a.src_w * 4;
    float* orow = outp + oy0 * T;
    // The walk is over the crop's rows: position r = row of the fourth tap (unclamped; taps = rows r-3 .. r clamped to the
    // crop, which is exactly ATen's clamping of iy-1 .. iy+2).  Every output row whose position is r is written from the
    // four rows in registers, then the oldest register takes row r+1.  Unrolled by four so that the registers keep
    // their names while their roles rotate: nothing is moved.
    float ra[NC], rb4[NC], rc[NC], rd[NC];
    const int hm1 = g.h - 1;
    auto load_row = [&](float (&dst)[NC], int row) {
      const float* src = reinterpret_cast<const float*>(col0 + min(max(row, 0), hm1) * rb);
#pragma unroll
      for (int c = 0; c < NC; ++c) dst[c] = src[32 * c];
    };
    int oy = oy0;
    int r = oy0 < oy1 ? s_top[oy0] : 0, pos = r;
    load_row(ra, r - 3); load_row(rb4, r - 2); load_row(rc, r - 1); load_row(rd, r);
    auto emit = [&](const float (&t0)[NC], const float (&t1)[NC], const float (&t2)[NC], const float (&t3)[NC]) {
      while (oy < oy1 && pos == r) {                            // (warp-uniform)
        const float4 wy = s_wy[oy];
#pragma unroll
        for (int c = 0; c < NC; ++c) {                          // (every load stayed inside the staged rows: no guard)
          float acc = t0[c] * wy.x;
          acc = fmaf(t1[c], wy.y, acc);
          acc = fmaf(t2[c], wy.z, acc);
          acc = fmaf(t3[c], wy.w, acc);
          if (live[c]) store_view(orow + 32 * c, acc);
        }
        ++oy; orow += T;
        pos = s_top[min(oy, F - 1)];
      }
    };
    while (oy < oy1) {
      emit(ra, rb4, rc, rd); if (oy >= oy1) break; load_row(ra, ++r);
      emit(rb4, rc, rd, ra); if (oy >= oy1) break; load_row(rb4, ++r);
      emit(rc, rd, ra, rb4); if (oy >= oy1) break; load_row(rc, ++r);
      emit(rd, ra, rb4, rc); if (oy >= oy1) break; load_row(rd, ++r);
    }
  }
  // deferred launch: this grid ran beside the one launched before it; whoever waits for THIS grid to complete
  // (the next kernel's pdl_wait) must also get that one's completion.  ONE CTA holds the grid open for that: if every
  // CTA waited here, the first wave would sit on its SM slots until the kernel before has finished (the statistics
  // scan + send take longer than a view) and nothing else of this grid could start.
  if (a.deferred && blockIdx.x == 0) pdl_wait();
}

template <int TW, bool GB, bool VEC>
int launch_views_kernel(ViewArgs& a, size_t smem, unsigned grid, cudaStream_t st) {
  if (cudaFuncSetAttribute(two_views_kernel<TW, GB, VEC>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  if (launch_pdl(two_views_kernel<TW, GB, VEC>, dim3(grid), dim3(kViewThreads), smem, st, a) != cudaSuccess) return ASSL_ERR_CUDA;
  return launch_status();
}

template <int TW>
int launch_views(ViewArgs& a, bool vec, cudaStream_t st) {
  a.tiles_t = (a.c.n_time + TW - 1) / TW;
  const size_t smem = views_smem_bytes(a.c.canvas_h, a.src_w, a.c.n_freq, TW);
  if (smem > 227 * 1024 / 2) return ASSL_ERR_UNSUPPORTED;        // (at least two CTAs per SM; a smaller tile is tried next)
  if ((a.src_w + 3) / 4 > 64 + 32) return ASSL_ERR_UNSUPPORTED;  // stage A covers the in-clip part with at most 2 x 32 groups
  const long long grid = a.B * 2 * a.tiles_t;
  if (grid > 0x7fffffffll) return ASSL_ERR_UNSUPPORTED;
  const bool global_bank = a.mix_rank || a.clip_stat;
  const unsigned gr = static_cast<unsigned>(grid);
  if (global_bank) return vec ? launch_views_kernel<TW, true, true>(a, smem, gr, st) : launch_views_kernel<TW, true, false>(a, smem, gr, st);
  return vec ? launch_views_kernel<TW, false, true>(a, smem, gr, st) : launch_views_kernel<TW, false, false>(a, smem, gr, st);
}

}  // namespace assl

extern "C" int assl_two_views_f32(const assl_view_cfg* vc, const float* lms, const float* mean_std, int64_t cap,
                                  const int32_t* clip_slot, const int32_t* mix_slot, const float* alpha,
                                  const int32_t* crop, int64_t B, float* view1, float* view2, void* stream) {
  return assl_two_views_ordered_f32(vc, lms, mean_std, cap, clip_slot, mix_slot, alpha, crop, nullptr, B, view1, view2, stream);
}

extern "C" int assl_two_views_ordered_f32(const assl_view_cfg* vc, const float* lms, const float* mean_std, int64_t cap,
                                          const int32_t* clip_slot, const int32_t* mix_slot, const float* alpha,
                                          const int32_t* crop, const int32_t* order, int64_t B, float* view1,
                                          float* view2, void* stream) {
  return assl_two_views_global_f32(vc, lms, nullptr, 0, mean_std, clip_slot, nullptr, mix_slot, nullptr, nullptr, alpha, crop,
                                   order, B, view1, view2, stream);
}

static int two_views_impl(const assl_view_cfg* vc, const float* lms, const float* const* peer_lms,
                          int32_t world, const float* mean_std, const int32_t* clip_slot,
                          const int32_t* clip_stat, const int32_t* mix_slot, const int32_t* mix_rank,
                          const int32_t* mix_stat, const float* alpha, const int32_t* crop,
                          const int32_t* order, int64_t B, float* view1, float* view2, int deferred, void* stream) {
  using namespace assl;
  const int64_t cap = 0;
  if (!vc || !lms || !clip_slot || !mix_slot || !alpha || !crop || !view1 || !view2 || B < 0) return ASSL_ERR_INVALID_ARG;
  if (world < 0 || world > ASSL_EXCHANGE_MAX_WORLD || (world > 0 && (!peer_lms || !mix_rank))) return ASSL_ERR_INVALID_ARG;
  if ((clip_stat == nullptr) != (mix_stat == nullptr)) return ASSL_ERR_INVALID_ARG;
  if (vc->n_freq < 1 || vc->n_time < 1 || vc->canvas_h < vc->n_freq || vc->canvas_w < vc->n_time) return ASSL_ERR_INVALID_ARG;
  if (vc->interp != 0 && vc->interp != 1) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  ViewArgs a;
  a.c = *vc; a.lms = lms; a.mean_std = mean_std; a.cap = cap; a.B = B;
  a.clip_slot = clip_slot; a.mix_slot = mix_slot; a.alpha = alpha; a.crop = crop; a.order = order;
  a.mix_rank = world > 0 ? mix_rank : nullptr; a.clip_stat = clip_stat; a.mix_stat = mix_stat;
  for (int r = 0; r < ASSL_EXCHANGE_MAX_WORLD; ++r) a.peer_lms[r] = r < world ? peer_lms[r] : nullptr;
  a.view1 = view1; a.view2 = view2;
  a.deferred = deferred;
  a.y0 = (vc->canvas_h - vc->n_freq) / 2;
  a.x0 = (vc->canvas_w - vc->n_time) / 2;
  const int T = vc->n_time;
  a.pitch = vc->ring_pitch > 0 ? vc->ring_pitch : T;
  if (a.pitch < T) return ASSL_ERR_INVALID_ARG;
  // 16-byte groups of a ring row are loaded as one word when every row starts on a 16-byte boundary, in every ring read
  bool vec = a.pitch % 4 == 0 && reinterpret_cast<uintptr_t>(lms) % 16 == 0;
  for (int r = 0; r < world; ++r) vec = vec && reinterpret_cast<uintptr_t>(peer_lms[r]) % 16 == 0;
  // floats per staged row: the widest span a TW-column tile can touch -- (TW-1) * max scale + 4 taps, never more than the
  // crop can be wide + the two taps outside it on either side -- + 3 (alignment shift), rounded up to whole 16-byte groups
  const double smax = T > 1 ? static_cast<double>(vc->canvas_w - 1) / (T - 1) : 0.0;
  auto span = [&](int tw) {
    const int cols = tw < T ? tw : T;
    long long s = static_cast<long long>((cols - 1) * smax) + 5;
    if (s > vc->canvas_w + 3) s = vc->canvas_w + 3;
    if (s + 3 < tw) s = tw - 3;                                  // (stages B and C read column lane + 32 c of a row unguarded)
    return static_cast<int>((s + 3 + 3) / 4 * 4);
  };
  int rc = ASSL_ERR_UNSUPPORTED;
  // tile width = lanes x column groups of stages B and C: the narrowest that holds the clip (T = 96, DeLoRes-S' 0.95 s,
  // takes three groups, not four with one idle), 128 for anything wider (tiled over time)
  if (T > 96) { a.src_w = span(128); rc = launch_views<128>(a, vec, as_stream(stream)); }
  if (rc == ASSL_ERR_UNSUPPORTED && T > 64) { a.src_w = span(96); rc = launch_views<96>(a, vec, as_stream(stream)); }
  if (rc == ASSL_ERR_UNSUPPORTED && T > 32) { a.src_w = span(64); rc = launch_views<64>(a, vec, as_stream(stream)); }
  if (rc == ASSL_ERR_UNSUPPORTED) { a.src_w = span(32); rc = launch_views<32>(a, vec, as_stream(stream)); }
  return rc;
}

extern "C" int assl_two_views_global_f32(const assl_view_cfg* vc, const float* lms, const float* const* peer_lms,
                                         int32_t world, const float* mean_std, const int32_t* clip_slot,
                                         const int32_t* clip_stat, const int32_t* mix_slot, const int32_t* mix_rank,
                                         const int32_t* mix_stat, const float* alpha, const int32_t* crop,
                                         const int32_t* order, int64_t B, float* view1, float* view2, void* stream) {
  return two_views_impl(vc, lms, peer_lms, world, mean_std, clip_slot, clip_stat, mix_slot, mix_rank, mix_stat, alpha, crop,
                        order, B, view1, view2, 0, stream);
}

extern "C" int assl_two_views_deferred_f32(const assl_view_cfg* vc, const float* lms, const float* mean_std, int64_t cap,
                                           const int32_t* clip_slot, const int32_t* mix_slot, const float* alpha,
                                           const int32_t* crop, const int32_t* order, int64_t B, float* view1,
                                           float* view2, void* stream) {
  (void)cap;
  return two_views_impl(vc, lms, nullptr, 0, mean_std, clip_slot, nullptr, mix_slot, nullptr, nullptr, alpha, crop, order, B,
                        view1, view2, 1, stream);
}
