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
// per source pixel instead of per tap), stage B resamples along time, stage C along frequency.
#include "common.cuh"

namespace assl {

constexpr int kViewThreads = 256;
constexpr int kSrcPad = 8;

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
  int src_w;       // staged source columns per tile (row stride of the source tile)
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

__device__ __forceinline__ void tap_weights(float t, int bilinear, float w[4]) {
  if (bilinear) {
    w[0] = 0.f; w[1] = 1.f - t; w[2] = t; w[3] = 0.f;
  } else {
    w[0] = cubic2(t + 1.f); w[1] = cubic1(t); w[2] = cubic1(1.f - t); w[3] = cubic2(2.f - t);
  }
}

__host__ __device__ inline size_t views_smem_bytes(int canvas_h, int src_w, int n_freq) {
  return sizeof(float) * static_cast<size_t>(canvas_h) * src_w +
         (sizeof(float) + sizeof(int)) * 4 * static_cast<size_t>(n_freq);
}

// ---- stage A: normalise + mix the crop span [0,h) x [xs_lo, xs_lo+nsx) into shared memory.
//      Only the part of the span that lies inside the pasted clip is computed: lanes <-> its columns
//      xs = xs_in0 + lane + 32 c, c < G = ceil(n_in / 32) (G chosen per CTA), i.e. clip columns lx_in0 + lane + 32 c --
//      one 64-bit row pointer per lane and row, every load at an immediate offset, no per-element predicate or
//      select (only the last group is clamped / guarded).  The zero canvas around the clip (augmentations.py:45-47)
//      is two guarded stores per row.  exp(xn) = ex2(x*a + b) with a = log2e/std, b = -mean*log2e/std: one FFMA +
//      one MUFU per operand.  LOGMIX (log-mixup-exp against a bank entry, the path's normal case) is a separate
//      instantiation; the other modes (no bank entry yet, plain mixup) share the generic one.
template <int G, bool GB, bool LOGMIX>
__device__ __forceinline__ void mix_crop_span(const ViewArgs& a, float* s_src, long long b, int v, int h, int ci,
                                              int lx_in0, int xs_in0, int n_in, int nsx, int lane, int warp) {
  const int F = a.c.n_freq, T = a.c.n_time;
  const int plane = F * T;
  // GB (global bank, a separate instantiation so the default kernel keeps its register budget): the bank entry may
  // sit in a peer's ring (loads over NVLink) and the (mean, std) pairs live in a ring over global clips
  const int xslot = a.clip_slot[b];
  const int xstat = (GB && a.clip_stat) ? a.clip_stat[b] : xslot;
  const float* xp = a.lms + static_cast<long long>(xslot) * plane;
  float mu = 0.f, inv = 1.f;
  if (a.mean_std) { mu = a.mean_std[2 * xstat]; inv = 1.f / a.mean_std[2 * xstat + 1]; }
  const int zslot = a.mix_slot[b * 2 + v];
  const bool mix = LOGMIX || zslot >= 0;
  const int zstat = (GB && a.mix_stat) ? a.mix_stat[b * 2 + v] : zslot;
  const float* zbase = a.lms;
  if (GB) { if (a.mix_rank && mix) zbase = a.peer_lms[a.mix_rank[b * 2 + v]]; }
  const float* zp = zbase + static_cast<long long>(mix ? zslot : 0) * plane;
  float zmu = 0.f, zinv = 1.f;
  if (mix && a.mean_std) { zmu = a.mean_std[2 * zstat]; zinv = 1.f / a.mean_std[2 * zstat + 1]; }
  const float al = a.alpha[b * 2 + v];
  const float wx = 1.f - al, wz = al;
  const float kLog2e = 1.4426950408889634f, kLn2 = 0.6931471805599453f;
  const float xa = LOGMIX ? inv * kLog2e : inv, xb = -mu * xa;
  const float za = LOGMIX ? zinv * kLog2e : zinv, zb = -zmu * za;
  constexpr int kColGroups = G;                               // 32 * G >= n_in, 32 * (G - 1) < n_in
  const int col0 = lx_in0 + lane;                             // this lane's clip column of group 0
  const int col_last = lx_in0 + min(lane + 32 * (kColGroups - 1), n_in - 1);   // last group: clamped into the clip
  const bool last_ok = lane + 32 * (kColGroups - 1) < n_in;
  const int ml = xs_in0, mr = nsx - (xs_in0 + n_in);          // canvas columns of the span left / right of the clip
  // (pointers opaque to the compiler, which otherwise rebuilds every address as ring base + 64-bit offset)
  asm volatile("" : "+l"(xp));
  asm volatile("" : "+l"(zp));
  // Rows are taken kRowBatch at a time per warp and ALL their global loads are issued before any
  // arithmetic, so a warp keeps 2 * kRowBatch * G loads in flight instead of 2.
  constexpr int kRowBatch = 2;                               // (3 or 4 rows for the narrow instances: no gain measured)
  for (int yb = warp * kRowBatch; yb < h; yb += (kViewThreads / 32) * kRowBatch) {
    float xv[kRowBatch][kColGroups], zv[kRowBatch][kColGroups];
#pragma unroll
    for (int r = 0; r < kRowBatch; ++r) {
      const int lyc = min(max(ci + yb + r - a.y0, 0), F - 1);  // warp-uniform; rows outside the clip are zero-filled below
      const float* xr = xp + (lyc * T + col0);
      const float* zr = zp + (lyc * T + col0);
#pragma unroll
      for (int c = 0; c < kColGroups - 1; ++c) {
        xv[r][c] = __ldg(xr + 32 * c);
        zv[r][c] = 0.f;
        if (mix) zv[r][c] = __ldg(zr + 32 * c);                // CTA-uniform (compile-time true for LOGMIX)
      }
      xv[r][kColGroups - 1] = __ldg(xp + (lyc * T + col_last));
      zv[r][kColGroups - 1] = 0.f;
      if (mix) zv[r][kColGroups - 1] = __ldg(zp + (lyc * T + col_last));
    }
#pragma unroll
    for (int r = 0; r < kRowBatch; ++r) {
      const int y = yb + r;
      const int ly = ci + y - a.y0;
      float* row = s_src + y * a.src_w;
      if (y < h) {
        if (ly >= 0 && ly < F) {                               // warp-uniform
          float* drow = row + xs_in0 + lane;
#pragma unroll
          for (int c = 0; c < kColGroups; ++c) {
            const float xn = fmaf(xv[r][c], xa, xb);
            float val;
            if (LOGMIX) {
              const float zn = fmaf(zv[r][c], za, zb);
              val = kLn2 * lg2_approx(fmaf(wx, ex2_approx(xn), fmaf(wz, ex2_approx(zn), kLogEps)));
            } else if (mix) {
              val = fmaf(wz, fmaf(zv[r][c], za, zb), wx * xn);
            } else {
              val = xn;
            }
            if (c < kColGroups - 1 || last_ok) drow[32 * c] = val;
          }
          if (lane < ml) row[lane] = 0.f;                      // zero canvas left and right of the pasted clip
          if (lane < mr) row[xs_in0 + n_in + lane] = 0.f;
          if (ml > 32 || mr > 32) {                            // (long clips: a tile that mostly misses the clip)
#pragma unroll 1
            for (int x = 32 + lane; x < ml; x += 32) row[x] = 0.f;
#pragma unroll 1
            for (int x = 32 + lane; x < mr; x += 32) row[xs_in0 + n_in + x] = 0.f;
          }
        } else {
#pragma unroll 1
          for (int x = lane; x < nsx; x += 32) row[x] = 0.f;   // canvas row above / below the clip
        }
      }
    }
  }
}

template <int G, bool GB>
__device__ __forceinline__ void mix_crop_span_any(const ViewArgs& a, float* s_src, long long b, int v, int h, int ci,
                                                  int lx_in0, int xs_in0, int n_in, int nsx, int lane, int warp) {
  const bool logmix = a.mix_slot[b * 2 + v] >= 0 && a.c.log_mixup;          // CTA-uniform
  if (logmix) mix_crop_span<G, GB, true>(a, s_src, b, v, h, ci, lx_in0, xs_in0, n_in, nsx, lane, warp);
  else mix_crop_span<G, GB, false>(a, s_src, b, v, h, ci, lx_in0, xs_in0, n_in, nsx, lane, warp);
}

// ---- stage B (along time: lanes <-> output columns, taps and weights in registers, warps <-> crop rows, each row
//      overwritten in place by the warp that owns it) and stage C (along frequency, coalesced stores: lanes <->
//      columns, warps <-> output rows).  NCM groups of 32 columns go through that mapping; with LEFT the tile has
//      32 * NCM + n_left columns, 1 <= n_left <= kLeftMax, and the left-over columns are handled 32 / n_left rows
//      at a time (lane -> (row, column)), so no lane idles on a mostly empty column group.
constexpr int kLeftMax = 8;

template <int NCM, bool LEFT>
__device__ __forceinline__ void resample_tile(const ViewArgs& a, float* s_src, const float4* s_wy, const int4* s_ya,
                                              float* outp, int h, int w, int ncol, int ox0, int xs_lo, int nsx, int bil,
                                              int lane, int warp) {
  const int F = a.c.n_freq, T = a.c.n_time;
  constexpr int kWarps = kViewThreads / 32;
  // left-over columns: lane -> (row rr of the iteration, column cc)
  const int n_left = LEFT ? ncol - 32 * NCM : 1;
  const int rpi = 32 / n_left;                                 // rows per iteration
  const int rr = lane / n_left, cc = lane - rr * n_left;
  const bool lv = LEFT && rr < rpi;
  {
    int xi[NCM][4];
    float wt[NCM][4];
#pragma unroll
    for (int c = 0; c < NCM; ++c) {
      const int col = LEFT ? lane + 32 * c : min(lane + 32 * c, ncol - 1);
      int ix; float tx;
      source_index(ox0 + col, w, T, &ix, &tx);
      tap_weights(tx, bil, wt[c]);
#pragma unroll
      for (int q = 0; q < 4; ++q) xi[c][q] = min(max(min(max(ix - 1 + q, 0), w - 1) - xs_lo, 0), nsx - 1);
    }
    float accl[2] = {0.f, 0.f};
    if (LEFT) {                                                // every left-over output of this warp's rows, before
      int ix; float tx;                                        // any of those rows is overwritten
      float wl[4];
      int xl[4];
      source_index(ox0 + 32 * NCM + cc, w, T, &ix, &tx);
      tap_weights(tx, bil, wl);
#pragma unroll
      for (int q = 0; q < 4; ++q) xl[q] = min(max(min(max(ix - 1 + q, 0), w - 1) - xs_lo, 0), nsx - 1);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int y = warp + kWarps * (j * rpi + rr);
        const float* row = s_src + ((lv && y < h) ? y : 0) * a.src_w;
        float acc = row[xl[0]] * wl[0];
        acc = fmaf(row[xl[1]], wl[1], acc);
        acc = fmaf(row[xl[2]], wl[2], acc);
        accl[j] = fmaf(row[xl[3]], wl[3], acc);
      }
      __syncwarp();
    }
    for (int y = warp; y < h; y += kWarps) {                   // a warp owns the row: read every tap, then overwrite
      float* row = s_src + y * a.src_w;
      float acc[NCM];
#pragma unroll
      for (int c = 0; c < NCM; ++c) {
        acc[c] = row[xi[c][0]] * wt[c][0];
        acc[c] = fmaf(row[xi[c][1]], wt[c][1], acc[c]);
        acc[c] = fmaf(row[xi[c][2]], wt[c][2], acc[c]);
        acc[c] = fmaf(row[xi[c][3]], wt[c][3], acc[c]);
      }
      __syncwarp();
#pragma unroll
      for (int c = 0; c < NCM; ++c)
        if (LEFT || lane + 32 * c < ncol) row[lane + 32 * c] = acc[c];
    }
    if (LEFT) {
      __syncwarp();
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int y = warp + kWarps * (j * rpi + rr);
        if (lv && y < h) s_src[y * a.src_w + 32 * NCM + cc] = accl[j];
      }
    }
  }
  __syncthreads();

  {
    for (int oy = warp; oy < F; oy += kWarps) {
      const float4 wy = s_wy[oy];
      const int4 ya = s_ya[oy];
      const float* r0 = s_src + ya.x * a.src_w + lane;
      const float* r1 = s_src + ya.y * a.src_w + lane;
      const float* r2 = s_src + ya.z * a.src_w + lane;
      const float* r3 = s_src + ya.w * a.src_w + lane;
      float* orow = outp + static_cast<long long>(oy) * T + lane;
#pragma unroll
      for (int c = 0; c < NCM; ++c) {
        if (LEFT || lane + 32 * c < ncol) {
          float acc = r0[32 * c] * wy.x;
          acc = fmaf(r1[32 * c], wy.y, acc);
          acc = fmaf(r2[32 * c], wy.z, acc);
          acc = fmaf(r3[32 * c], wy.w, acc);
          store_view(orow + 32 * c, acc);
        }
      }
    }
    if (LEFT) {
      const int col = 32 * NCM + cc;
      for (int k0 = 0; warp + kWarps * k0 < F; k0 += rpi) {
        const int oy = warp + kWarps * (k0 + rr);
        if (lv && oy < F) {
          const float4 wy = s_wy[oy];
          const int4 ya = s_ya[oy];
          float acc = s_src[ya.x * a.src_w + col] * wy.x;
          acc = fmaf(s_src[ya.y * a.src_w + col], wy.y, acc);
          acc = fmaf(s_src[ya.z * a.src_w + col], wy.z, acc);
          acc = fmaf(s_src[ya.w * a.src_w + col], wy.w, acc);
          store_view(outp + static_cast<long long>(oy) * T + col, acc);
        }
      }
    }
  }
}

// TW = output frames per tile; CG = 32-column groups of the staged crop span a warp loads per row
template <int TW, int CG, bool GB>
__global__ void __launch_bounds__(kViewThreads, 4) two_views_kernel(const ViewArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int F = a.c.n_freq, T = a.c.n_time;
  float4* s_wy = reinterpret_cast<float4*>(smem_raw);                // [F]         frequency tap weights
  int4* s_ya = reinterpret_cast<int4*>(s_wy + F);                    // [F]         frequency tap rows
  float* s_src = reinterpret_cast<float*>(s_ya + F);                 // [h][src_w]  mixed crop span, then (in place,
                                                                     //             row by row) the time-resampled rows

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  long long blk = blockIdx.x;
  const int tile = static_cast<int>(blk % a.tiles_t); blk /= a.tiles_t;
  pdl_launch_dependents();                     // the next kernel may take SMs as this grid's last wave drains
  if (!a.deferred) pdl_wait();                 // nothing global is read before the previous grid has completed
  if (a.order) blk = a.order[blk];             // CTAs start in launch order: the host puts the big crops first
  const int v = static_cast<int>(blk & 1);
  const long long b = blk >> 1;

  const int* cr = a.crop + (b * 2 + v) * 4;
  // the host mirror validates crops; the clamps only keep a bad box from leaving shared memory
  const int h = min(max(cr[2], 1), a.c.canvas_h), w = min(max(cr[3], 1), a.c.canvas_w);
  const int ci = min(max(cr[0], 0), a.c.canvas_h - h), cj = min(max(cr[1], 0), a.c.canvas_w - w);
  const int bil = a.c.interp;

  const int ox0 = tile * TW;
  const int ncol = min(TW, T - ox0);
  // crop columns this tile's taps can touch
  int ix_lo, ix_hi; float lam;
  source_index(ox0, w, T, &ix_lo, &lam);
  source_index(ox0 + ncol - 1, w, T, &ix_hi, &lam);
  const int xs_lo = max(ix_lo - 1, 0);
  const int xs_hi = min(ix_hi + 2, w - 1);
  const int nsx = min(xs_hi - xs_lo + 1, a.src_w);

  // ---- frequency-axis taps, shared by every column of the tile
  for (int oy = tid; oy < F; oy += kViewThreads) {
    int iy; float ty;
    source_index(oy, h, F, &iy, &ty);
    float wt[4];
    tap_weights(ty, bil, wt);
    s_wy[oy] = make_float4(wt[0], wt[1], wt[2], wt[3]);
    s_ya[oy] = make_int4(min(max(iy - 1, 0), h - 1), min(max(iy, 0), h - 1),
                         min(max(iy + 1, 0), h - 1), min(max(iy + 2, 0), h - 1));
  }

  // ---- stage A, with as many 32-column groups as the part of THIS crop span inside the clip needs
  {
    // columns of the span that fall inside the pasted clip: lx = cj + xs_lo + xs - x0 in [0, T)
    const int lx0 = cj + xs_lo - a.x0;
    const int xs_in0 = min(max(0, -lx0), nsx), xs_in1 = max(min(nsx, T - lx0), xs_in0);
    const int n_in = xs_in1 - xs_in0, lx_in0 = lx0 + xs_in0;
#define ASSL_STAGE_A(G) mix_crop_span_any<G, GB>(a, s_src, b, v, h, ci, lx_in0, xs_in0, n_in, nsx, lane, warp)
    const int g = min((n_in + 31) >> 5, CG);                     // (the launcher guarantees CG >= ceil(n_in / 32))
    if (g == 0) {                                                // the tile's span misses the clip: all canvas
      for (int y = warp; y < h; y += kViewThreads / 32) {
#pragma unroll 1
        for (int x = lane; x < nsx; x += 32) s_src[y * a.src_w + x] = 0.f;
      }
    } else if (g == 1) { ASSL_STAGE_A(1); }
    else if (CG >= 2 && g == 2) { ASSL_STAGE_A((CG >= 2 ? 2 : 1)); }
    else if (CG >= 3 && g == 3) { ASSL_STAGE_A((CG >= 3 ? 3 : 1)); }
    else if (CG >= 4 && g == 4) { ASSL_STAGE_A((CG >= 4 ? 4 : 1)); }
    else if (CG >= 5 && g == 5) { ASSL_STAGE_A((CG >= 5 ? 5 : 1)); }
    else if (CG >= 6 && g == 6) { ASSL_STAGE_A((CG >= 6 ? 6 : 1)); }
    else if (CG >= 7) { ASSL_STAGE_A((CG >= 7 ? 7 : 1)); }
#undef ASSL_STAGE_A
  }
  __syncthreads();

  // ---- stages B and C.  T = 101 is three full groups of 32 output columns and 5 left over: those are resampled
  //      with lanes <-> (row, column) instead of wasting 27 lanes of a fourth group on every row
  constexpr int NC = TW / 32;                                  // output columns per lane: lane + 32*c
  float* outp = (v == 0 ? a.view1 : a.view2) + b * (static_cast<long long>(F) * T) + ox0;
  const int n_left = ncol - 32 * (NC - 1);
  if (NC > 1 && n_left >= 1 && n_left <= kLeftMax && (a.c.canvas_h + 7) / 8 <= 2 * (32 / n_left))
    resample_tile<(NC > 1 ? NC - 1 : 1), true>(a, s_src, s_wy, s_ya, outp, h, w, ncol, ox0, xs_lo, nsx, bil, lane, warp);
  else
    resample_tile<NC, false>(a, s_src, s_wy, s_ya, outp, h, w, ncol, ox0, xs_lo, nsx, bil, lane, warp);
  // deferred launch: this grid ran beside the one launched before it; whoever waits for THIS grid to complete
  // (the next kernel's pdl_wait) must also get that one's completion
  if (a.deferred) pdl_wait();
}

template <int TW, int CG, bool GB>
int launch_views_kernel(ViewArgs& a, size_t smem, unsigned grid, cudaStream_t st) {
  if (cudaFuncSetAttribute(two_views_kernel<TW, CG, GB>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  if (launch_pdl(two_views_kernel<TW, CG, GB>, dim3(grid), dim3(kViewThreads), smem, st, a) != cudaSuccess) return ASSL_ERR_CUDA;
  return launch_status();
}

template <int TW, int CG>
int launch_views_cg(ViewArgs& a, cudaStream_t st) {
  const size_t smem = views_smem_bytes(a.c.canvas_h, a.src_w, a.c.n_freq);
  if (smem > 227 * 1024) return ASSL_ERR_UNSUPPORTED;
  const long long grid = a.B * 2 * a.tiles_t;
  if (grid > 0x7fffffffll) return ASSL_ERR_UNSUPPORTED;
  const bool global_bank = a.mix_rank || a.clip_stat;
  return global_bank ? launch_views_kernel<TW, CG, true>(a, smem, static_cast<unsigned>(grid), st)
                     : launch_views_kernel<TW, CG, false>(a, smem, static_cast<unsigned>(grid), st);
}

template <int TW>
int launch_views(ViewArgs& a, cudaStream_t st) {
  a.tiles_t = (a.c.n_time + TW - 1) / TW;
  if (a.src_w < (a.c.n_time < TW ? a.c.n_time : TW)) return ASSL_ERR_UNSUPPORTED;   // rows are reused for the outputs
  constexpr int kMaxCG = (TW * 3 / 2 + 16 + 31) / 32;            // covers span() for virtual_crop_scale <= 1.5
  const int need = (a.src_w - 1 + 31) / 32;                      // stage A loads 32 columns per group
  if (need > kMaxCG) return ASSL_ERR_UNSUPPORTED;                 // virtual_crop_scale[1] well above 1.5
  if (need <= kMaxCG - 2) return launch_views_cg<TW, (kMaxCG - 2 > 0 ? kMaxCG - 2 : 1)>(a, st);
  if (need <= kMaxCG - 1) return launch_views_cg<TW, kMaxCG - 1>(a, st);
  return launch_views_cg<TW, kMaxCG>(a, st);
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
  // widest crop span a TW-column tile can touch: (TW-1) * max scale + 4 taps, never more than the canvas
  const int T = vc->n_time;
  const double smax = T > 1 ? static_cast<double>(vc->canvas_w - 1) / (T - 1) : 0.0;
  auto span = [&](int tw) {
    const int cols = tw < T ? tw : T;
    long long s = static_cast<long long>((cols - 1) * smax) + 5;
    if (s > vc->canvas_w) s = vc->canvas_w;
    return static_cast<int>((s + kSrcPad - 1) / kSrcPad * kSrcPad) + 1;     // odd-ish stride spreads banks
  };
  int rc = ASSL_ERR_UNSUPPORTED;
  if (T > 64) { a.src_w = span(128); rc = launch_views<128>(a, as_stream(stream)); }
  if (rc == ASSL_ERR_UNSUPPORTED && T > 32) { a.src_w = span(64); rc = launch_views<64>(a, as_stream(stream)); }
  if (rc == ASSL_ERR_UNSUPPORTED) { a.src_w = span(32); rc = launch_views<32>(a, as_stream(stream)); }
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
