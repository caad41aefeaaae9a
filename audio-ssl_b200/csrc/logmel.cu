// Fused log-mel front end: frame + periodic Hann + rFFT + |X|^2 + sparse mel + log, one kernel.
//
// Replaces MelSpectrogramLibrosa.__call__ (/root/reference/src/utils/utils.py:26-28) and
// extract_log_mel_spectrogram (/root/reference/src/utils/utils.py:43-49).
//
// Work decomposition
//   persistent CTAs (2 per SM) loop over tiles of kFramesPerTile consecutive frames of one clip;
//   the tile's waveform span is brought into shared memory by ONE bulk async copy (TMA, cp.async.bulk
//   + mbarrier) issued a tile ahead, so the 6.4x overlap between hop-160 frames never goes back to
//   L2/HBM and the global latency hides behind the previous tile's FFTs; librosa's reflect padding
//   only touches the first/last tiles of a clip and is patched in by a short scalar fix-up;
//   each warp transforms TWO real frames at once as one 1024-point complex FFT laid out 32 x 32:
//   lane n1 holds samples n1 + 32*n2 (stride-32, conflict-free LDS; when hop is a multiple of 32 the
//   second frame's samples are the first frame's shifted by hop/32 registers, so 37 loads feed both),
//   runs a register-resident 32-point FFT (fft32_gen.cuh), multiplies by the inter-stage twiddles,
//   transposes through a padded per-warp shared buffer, runs the second 32-point FFT, and unpacks the
//   two spectra with one shuffle per partner value (bin k of lane l pairs with bin 1024-k of lane 32-l);
//   the two power spectra go back to the warp's own shared buffer, then the whole CTA applies the
//   sparse (triangular-support) filterbank, takes the log and writes [n_mels, T] row-major.
#include "common.cuh"
#include "fft32_gen.cuh"

namespace assl {

constexpr int kWarps = 8;
constexpr int kThreads = kWarps * 32;
constexpr int kExStride = 33;                       // float2 units: lane-major rows padded by one
constexpr int kExWarp = 32 * kExStride + 2;         // 1058 float2 = 2116 words == 4 (mod 32)

struct LogmelArgs {
  const float* wave;
  const float* gain;
  float* lms;
  double* tile_sums;
  long long L, ld, slot0, cap;
  int B, T, hop, n_mels, nnz, out_kind, bulk_ok;
  int ppc;            // frame pairs per clip = ceil(T / 2)
  int straddle;       // 1: a tile is 8 consecutive pairs of the flattened (clip, pair) list and may span two clips
  int total_tiles, n_parts, wave_floats;
  const float* win;
  const float2* tw;
  const int* mk0;
  const int* mcnt;
  const int* moff;
  const float* mw;
  const float* mconst;   // [n_mels] 2.22e-16 * sum of the band's weights (the "+ eps" of utils.py:28, folded)
};

__host__ __device__ inline size_t logmel_smem_bytes(int wave_floats, int nnz) {
  size_t b = 0;
  b += sizeof(float2) * kWarps * kExWarp;   // exchange / power buffers
  b += sizeof(float2) * 1024;               // twiddles
  b += sizeof(float) * 1024;                // window
  b += sizeof(float) * ((wave_floats + 3) & ~3);
  b += sizeof(float) * ((nnz + 3) & ~3);
  b += sizeof(int) * 3 * kMaxMels + sizeof(float) * kMaxMels;
  b += sizeof(float2) * 2 * kWarps * kWarps;   // per (warp, frame pair) (sum, sum^2), double-buffered by tile parity
  b += 2 * sizeof(unsigned long long);      // mbarriers
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
__device__ __forceinline__ void mbar_arrive(void* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
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
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, void* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// A tile = up to 8 frame pairs; its pairs belong to one clip, or (straddle mode) to two consecutive clips.
struct Segment {
  int clip;      // clip index in the batch
  int q0;        // first pair of the clip in this tile
  int n;         // pairs of this clip in the tile (0 = segment absent)
  int off;       // where the segment's waveform span starts in s_wave (floats, multiple of 4)
  int span;      // samples the segment's frames touch: (2n - 1) * hop + 1024
  long long base;   // clip-relative sample index of the span start: 2*q0*hop - 512 (may be negative)
};

__device__ __forceinline__ void tile_segments(const LogmelArgs& a, int tile, Segment seg[2]) {
  int p0, np;
  if (a.straddle) {
    p0 = tile * kWarps;
    const int total = a.B * a.ppc;
    np = total - p0 < kWarps ? total - p0 : kWarps;
  } else {                                 // clips shorter than 8 pairs: one clip per tile
    p0 = tile * a.ppc;
    np = a.ppc;
  }
  const int b0 = p0 / a.ppc;
  const int q0 = p0 - b0 * a.ppc;
  const int n0 = np < a.ppc - q0 ? np : a.ppc - q0;
  seg[0].clip = b0; seg[0].q0 = q0; seg[0].n = n0; seg[0].off = 0;
  seg[0].span = (2 * n0 - 1) * a.hop + kNfft;
  seg[0].base = 2ll * q0 * a.hop - kNfft / 2;
  seg[1].clip = b0 + 1; seg[1].q0 = 0; seg[1].n = np - n0; seg[1].off = (seg[0].span + 3) & ~3;
  seg[1].span = (2 * seg[1].n - 1) * a.hop + kNfft;
  seg[1].base = -kNfft / 2;
}

// part of a segment's span the bulk copy covers (in range of the clip), span-relative [lo, hi), multiples of 4
__device__ __forceinline__ void bulk_range(const Segment& g, long long L, int* lo, int* hi) {
  const long long l = g.base < 0 ? -g.base : 0;
  long long h = L - g.base;
  h = h > g.span ? g.span : h;
  h &= ~3ll;
  *lo = static_cast<int>(l);
  *hi = static_cast<int>(h > l ? h : l);
}

// H32 = hop / 32 when hop is a multiple of 32 (both frames of a pair share registers), else 0
template <int H32>
__global__ void __launch_bounds__(kThreads, 2) logmel_kernel(const LogmelArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* s_ex = reinterpret_cast<float2*>(smem_raw);
  float2* s_tw = s_ex + kWarps * kExWarp;
  float* s_win = reinterpret_cast<float*>(s_tw + 1024);
  float* s_wave = s_win + 1024;
  float* s_mw = s_wave + ((a.wave_floats + 3) & ~3);
  int* s_mk0 = reinterpret_cast<int*>(s_mw + ((a.nnz + 3) & ~3));
  int* s_mcnt = s_mk0 + kMaxMels;
  int* s_moff = s_mcnt + kMaxMels;
  float* s_mconst = reinterpret_cast<float*>(s_moff + kMaxMels);
  float2* s_red = reinterpret_cast<float2*>(s_mconst + kMaxMels);          // [2][kWarps][kWarps]
  unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_red + 2 * kWarps * kWarps);
  void* bar_wave = s_bar;        // tile's waveform spans have landed (tx bytes)
  void* bar_mel = s_bar + 1;     // every thread is done reading the previous tile's power spectra

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  if (tid == 0) { mbar_init(bar_wave, 1); mbar_init(bar_mel, kThreads); }
  for (int i = tid; i < 1024; i += kThreads) { s_tw[i] = a.tw[i]; s_win[i] = a.win[i]; }
  for (int i = tid; i < a.nnz; i += kThreads) s_mw[i] = a.mw[i];
  for (int i = tid; i < a.n_mels; i += kThreads) {
    s_mk0[i] = a.mk0[i]; s_mcnt[i] = a.mcnt[i]; s_moff[i] = a.moff[i]; s_mconst[i] = a.mconst[i];
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  // producer: bulk-copy the in-range spans of a tile into s_wave (one elected thread)
  auto issue_tile = [&](int tile) {
    Segment seg[2];
    tile_segments(a, tile, seg);
    unsigned bytes[2] = {0u, 0u};
    int lo[2] = {0, 0};
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (seg[s].n > 0) {
        int hi;
        bulk_range(seg[s], a.L, &lo[s], &hi);
        bytes[s] = static_cast<unsigned>(hi - lo[s]) * 4u;
      }
    }
    mbar_expect_tx(bar_wave, bytes[0] + bytes[1]);
#pragma unroll
    for (int s = 0; s < 2; ++s)
      if (bytes[s]) bulk_g2s(s_wave + seg[s].off + lo[s], a.wave + seg[s].clip * a.ld + seg[s].base + lo[s], bytes[s], bar_wave);
  };
  if (a.bulk_ok && tid == 0 && static_cast<int>(blockIdx.x) < a.total_tiles) issue_tile(blockIdx.x);

  // flush the statistics a tile left in s_red: one (sum, sum^2) pair per clip segment, fixed summation
  // order.  Called by warp 0: lane f adds up frame pair f over the 8 producing warps, then a masked
  // butterfly combines the pairs of each segment.
  auto flush_stats = [&](int tile, unsigned parity) {
    Segment seg[2];
    tile_segments(a, tile, seg);
    const float2* r = s_red + parity * kWarps * kWarps;
    double r1 = 0.0, r2 = 0.0;
    if (lane < kWarps) {
#pragma unroll
      for (int w = 0; w < kWarps; ++w) { const float2 v = r[w * kWarps + lane]; r1 += v.x; r2 += v.y; }
    }
    const int n0 = seg[0].n, np = n0 + seg[1].n;
    double a1 = lane < n0 ? r1 : 0.0, a2 = lane < n0 ? r2 : 0.0;            // segment 0
    double b1 = (lane >= n0 && lane < np) ? r1 : 0.0, b2 = (lane >= n0 && lane < np) ? r2 : 0.0;   // segment 1
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      a1 += __shfl_xor_sync(0xffffffffu, a1, o); a2 += __shfl_xor_sync(0xffffffffu, a2, o);
      b1 += __shfl_xor_sync(0xffffffffu, b1, o); b2 += __shfl_xor_sync(0xffffffffu, b2, o);
    }
    if (lane < 2 && (lane == 0 ? n0 : seg[1].n) > 0) {
      const int clip = seg[0].clip + lane;
      const int first_tile = a.straddle ? (clip * a.ppc) / kWarps : tile;
      double* dst = a.tile_sums + (static_cast<long long>(clip) * a.n_parts + (tile - first_tile)) * 2;
      dst[0] = lane == 0 ? a1 : b1;
      dst[1] = lane == 0 ? a2 : b2;
    }
  };

  unsigned it = 0;
  int prev_tile = -1;
  for (int tile = blockIdx.x; tile < a.total_tiles; tile += gridDim.x, ++it) {
    Segment seg[2];
    tile_segments(a, tile, seg);
    const int n0 = seg[0].n, np = n0 + seg[1].n;

    // ---- waveform spans: wait for the bulk copies, patch what they could not cover (reflect padding of
    //      np.pad(mode='reflect'): p<0 -> -p, p>L-1 -> 2(L-1)-p; tail remainder when L % 4 != 0)
    if (a.bulk_ok) mbar_wait(bar_wave, it & 1);
    else __syncthreads();                            // previous tile's input reads are done
    bool fixed = false;
#pragma unroll
    for (int s = 0; s < 2; ++s) {
      if (seg[s].n > 0) {
        int lo = 0, hi = 0;
        if (a.bulk_ok) bulk_range(seg[s], a.L, &lo, &hi);
        if (lo > 0 || hi < seg[s].span) {             // CTA-uniform
          const int last = static_cast<int>(a.L - 1);
          const int n_fix = lo + (seg[s].span - hi);
          const float* src = a.wave + seg[s].clip * a.ld;
          float* dst = s_wave + seg[s].off;
          for (int q = tid; q < n_fix; q += kThreads) {
            const int i = q < lo ? q : hi + (q - lo);
            long long p = seg[s].base + i;
            if (p < 0) p = -p;
            if (p > last) p = 2ll * last - p;
            p = p < 0 ? 0 : (p > last ? last : p);    // only frames >= T can land here; never read back
            dst[i] = __ldg(src + p);
          }
          fixed = true;
        }
      }
    }
    if (fixed) __syncthreads();

    // this warp's pair: which clip, which frames, where its samples start
    const bool have = warp < np;                      // warp-uniform
    const int wseg = warp >= n0 ? 1 : 0;
    const int wpair = wseg ? warp - n0 : seg[0].q0 + warp;        // pair index inside its clip
    float2* ex = s_ex + warp * kExWarp;
    float re[32], im[32];
    if (have) {
      const float* w0 = s_wave + (wseg ? seg[1].off + wpair * 2 * a.hop : warp * 2 * a.hop) + lane;
      if (H32 > 0) {
        float s[32 + H32];
#pragma unroll
        for (int j = 0; j < 32 + H32; ++j) s[j] = w0[32 * j];
#pragma unroll
        for (int n2 = 0; n2 < 32; ++n2) {
          const float w = s_win[lane + 32 * n2];
          re[n2] = w * s[n2];
          im[n2] = w * s[n2 + H32];
        }
      } else {
        const float* w1 = w0 + a.hop;
#pragma unroll
        for (int n2 = 0; n2 < 32; ++n2) {
          const float w = s_win[lane + 32 * n2];
          re[n2] = w * w0[32 * n2];
          im[n2] = w * w1[32 * n2];
        }
      }
    }
    // every warp has its samples in registers -> the next tile's copy may overwrite s_wave
    if (a.bulk_ok) {
      if (warp == 0) {
        asm volatile("bar.sync 1, %0;" ::"n"(kThreads) : "memory");
        if (lane == 0 && tile + static_cast<int>(gridDim.x) < a.total_tiles) issue_tile(tile + gridDim.x);
      } else {
        asm volatile("bar.arrive 1, %0;" ::"n"(kThreads) : "memory");
      }
    }
    if (have) {
      fft32(re, im);                               // over n2 -> k2, this lane is n1
#pragma unroll
      for (int k2 = 1; k2 < 32; ++k2) {
        const float2 t = s_tw[k2 * 32 + lane];
        const float r = re[k2] * t.x - im[k2] * t.y;
        const float i = re[k2] * t.y + im[k2] * t.x;
        re[k2] = r; im[k2] = i;
      }
    }
    if (it > 0) mbar_wait(bar_mel, (it - 1) & 1);   // previous tile's power spectra are consumed
    if (have) {
#pragma unroll
      for (int k2 = 0; k2 < 32; ++k2) ex[lane * kExStride + k2] = make_float2(re[k2], im[k2]);
      __syncwarp();
#pragma unroll
      for (int n1 = 0; n1 < 32; ++n1) { const float2 v = ex[n1 * kExStride + lane]; re[n1] = v.x; im[n1] = v.y; }
      fft32(re, im);                               // over n1 -> k1, this lane is k2: Z[lane + 32*k1]

      // unpack: X1[k] = (Z[k] + conj Z[N-k]) / 2, X2[k] = (Z[k] - conj Z[N-k]) / 2i  (the 1/2 sits in the window)
      const int partner = (32 - lane) & 31;
      float p1[16], p2[16];
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) {
        float br = __shfl_sync(0xffffffffu, re[31 - k1], partner);
        float bi = __shfl_sync(0xffffffffu, im[31 - k1], partner);
        if (lane == 0) { br = re[(32 - k1) & 31]; bi = im[(32 - k1) & 31]; }
        const float sr = re[k1] + br, di = im[k1] - bi;
        const float si = im[k1] + bi, dr = re[k1] - br;
        p1[k1] = sr * sr + di * di;
        p2[k1] = si * si + dr * dr;
      }
      __syncwarp();                                // everyone is done reading the exchange buffer
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) ex[lane + 32 * k1] = make_float2(p1[k1], p2[k1]);
      if (lane == 0) {                             // Nyquist bin: its own partner
        const float nr = 2.f * re[16], ni = 2.f * im[16];
        ex[512] = make_float2(nr * nr, ni * ni);
      }
    }
    __syncthreads();                               // power spectra of the whole tile are in shared memory

    // flush the previous tile's statistics (all warps wrote them before reaching the barrier above)
    if (a.tile_sums && warp == 0 && prev_tile >= 0) flush_stats(prev_tile, (it - 1) & 1);

    // ---- sparse filterbank + log: thread = (band, frame pair); 32 bands per pass
    const int fp = tid & 7;                         // frame pair = the warp that produced it
    const int fseg = fp >= n0 ? 1 : 0;
    const int ta = 2 * (fseg ? fp - n0 : seg[0].q0 + fp);
    const int fclip = seg[0].clip + fseg;
    const bool fhave = fp < np && ta < a.T;
    const bool fhave2 = fhave && ta + 1 < a.T;
    const float2* pw = s_ex + fp * kExWarp;
    float* out = a.lms + ring_slot(a.slot0, fclip, a.cap) * (static_cast<long long>(a.n_mels) * a.T) + ta;
    float g2 = 1.f;
    if (a.gain && fhave) { const float g = a.gain[fclip]; g2 = g * g; }   // |g x|^2 = g^2 |x|^2: per-clip gain applied here
    float s1 = 0.f, s2 = 0.f;
    for (int mb = 0, pass = 0; mb < a.n_mels; mb += 32, ++pass) {
      // odd passes walk their 32 bands backwards so wide and narrow bands share a warp's time
      const int m = (pass & 1) ? mb + 31 - (tid >> 3) : mb + (tid >> 3);
      if (m < a.n_mels && fhave) {
        const int cnt = s_mcnt[m];
        const float* wgt = s_mw + s_moff[m];
        const float2* pk = pw + s_mk0[m];
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 4
        for (int j = 0; j < cnt; ++j) {
          const float w = wgt[j];
          const float2 p = pk[j];
          acc0 = fmaf(w, p.x, acc0);
          acc1 = fmaf(w, p.y, acc1);
        }
        const float c = s_mconst[m];
        acc0 = fmaf(acc0, g2, c);
        acc1 = fmaf(acc1, g2, c);
        const float l0 = a.out_kind ? acc0 : __logf(acc0 + kLogEps);
        const float l1 = a.out_kind ? acc1 : __logf(acc1 + kLogEps);
        float* o = out + m * a.T;
        o[0] = l0;
        s1 += l0; s2 = fmaf(l0, l0, s2);
        if (fhave2) { o[1] = l1; s1 += l1; s2 = fmaf(l1, l1, s2); }
      }
    }
    mbar_arrive(bar_mel);                          // this thread no longer reads the power buffers
    if (a.tile_sums) {                             // per (warp, frame pair) partial sums; lanes with equal fp combine
      s1 += __shfl_xor_sync(0xffffffffu, s1, 8);  s2 += __shfl_xor_sync(0xffffffffu, s2, 8);
      s1 += __shfl_xor_sync(0xffffffffu, s1, 16); s2 += __shfl_xor_sync(0xffffffffu, s2, 16);
      if (lane < kWarps) s_red[((it & 1) * kWarps + warp) * kWarps + lane] = make_float2(s1, s2);
    }
    prev_tile = tile;
  }
  __syncthreads();
  if (a.tile_sums && warp == 0 && prev_tile >= 0) flush_stats(prev_tile, (it - 1) & 1);
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
int launch_logmel(const LogmelArgs& a, size_t smem, unsigned grid, cudaStream_t st) {
  if (cudaFuncSetAttribute(logmel_kernel<H32>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  logmel_kernel<H32><<<grid, kThreads, smem, st>>>(a);
  return launch_status();
}

}  // namespace assl

namespace {
// frame pairs per clip, and how many tiles (parts of tile_sums) can touch one clip
inline void tiling(int64_t T, int* ppc, int* straddle, int* n_parts) {
  *ppc = static_cast<int>((T + 1) / 2);
  *straddle = *ppc >= assl::kWarps ? 1 : 0;
  *n_parts = *straddle ? (*ppc - 1) / assl::kWarps + 2 : 1;
}
}  // namespace

extern "C" int64_t assl_plan_num_tiles(const assl_plan* p, int64_t n_samples) {
  const int64_t t = assl_plan_num_frames(p, n_samples);
  if (t < 0) return -1;
  int ppc, straddle, n_parts;
  tiling(t, &ppc, &straddle, &n_parts);
  return n_parts;
}

extern "C" int assl_logmel_f32(const assl_plan* p, const float* wave, const float* wave_gain, int64_t B,
                               int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap,
                               double* tile_sums, int32_t out_kind, void* stream) {
  using namespace assl;
  if (out_kind != 0 && out_kind != 1) return ASSL_ERR_INVALID_ARG;
  if (!p || !wave || !lms || B < 0 || ld < L || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (L <= p->cfg.n_fft / 2) return ASSL_ERR_INVALID_ARG;       // np.pad(reflect) needs L > n_fft/2
  if (cap > 0 && (B > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  const int64_t T = 1 + L / p->cfg.hop;
  LogmelArgs a;
  tiling(T, &a.ppc, &a.straddle, &a.n_parts);
  const int64_t total_pairs = B * a.ppc;
  if (T > (1ll << 24) || total_pairs > (1ll << 30) || L > (1ll << 31) - 4096) return ASSL_ERR_UNSUPPORTED;
  a.wave = wave; a.gain = wave_gain; a.lms = lms; a.tile_sums = tile_sums;
  a.B = static_cast<int>(B); a.L = L; a.ld = ld; a.slot0 = slot0; a.cap = cap;
  a.T = static_cast<int>(T); a.hop = p->cfg.hop; a.n_mels = p->cfg.n_mels;
  a.total_tiles = a.straddle ? static_cast<int>((total_pairs + kWarps - 1) / kWarps) : static_cast<int>(B);
  // two spans per tile at most: (16 - 2) hops + 2 windows, each span start rounded to 4 floats
  a.wave_floats = 2 * kWarps * p->cfg.hop + 2 * kNfft + 8;
  a.nnz = p->nnz; a.out_kind = out_kind;
  // the bulk (TMA) copy needs 16-byte aligned global addresses: clip starts and span starts (2*q*hop - 512)
  a.bulk_ok = (reinterpret_cast<uintptr_t>(wave) % 16 == 0) && (ld % 4 == 0) && (p->cfg.hop % 2 == 0);
  a.win = p->d_win_half; a.tw = p->d_twiddle;
  a.mk0 = p->d_mel_k0; a.mcnt = p->d_mel_cnt; a.moff = p->d_mel_off; a.mw = p->d_mel_w; a.mconst = p->d_mel_const;
  const size_t smem = logmel_smem_bytes(a.wave_floats, a.nnz);
  if (smem > 227 * 1024) return ASSL_ERR_UNSUPPORTED;
  // parts of tile_sums no tile writes (a clip touches at most n_parts tiles, usually fewer) must read as zero
  if (tile_sums && cudaMemsetAsync(tile_sums, 0, sizeof(double) * 2 * B * a.n_parts, as_stream(stream)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  const unsigned grid = static_cast<unsigned>(a.total_tiles < 2 * p->sm_count ? a.total_tiles : 2 * p->sm_count);
  if (a.hop == 160) return launch_logmel<5>(a, smem, grid, as_stream(stream));
  return launch_logmel<0>(a, smem, grid, as_stream(stream));
}

extern "C" int assl_wave_l2_gain_f32(const float* wave, int64_t B, int64_t L, int64_t ld, float* gain, void* stream) {
  if (!wave || !gain || B < 0 || L < 1 || ld < L) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::wave_l2_gain_kernel<<<static_cast<unsigned>(B), 256, 0, assl::as_stream(stream)>>>(wave, L, ld, gain);
  return assl::launch_status();
}
