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
constexpr int kFramesPerTile = 2 * kWarps;
constexpr int kExStride = 33;                       // float2 units: lane-major rows padded by one
constexpr int kExWarp = 32 * kExStride + 2;         // 1058 float2 = 2116 words == 4 (mod 32)

struct LogmelArgs {
  const float* wave;
  const float* gain;
  float* lms;
  double* tile_sums;
  long long B, L, ld, slot0, cap;
  int T, hop, n_mels, tiles_per_clip, tile_samples, nnz, out_kind, bulk_ok;
  const float* win;
  const float2* tw;
  const int* mk0;
  const int* mcnt;
  const int* moff;
  const float* mw;
  const float* mconst;   // [n_mels] 2.22e-16 * sum of the band's weights (the "+ eps" of utils.py:28, folded)
};

__host__ __device__ inline size_t logmel_smem_bytes(int tile_samples, int nnz) {
  size_t b = 0;
  b += sizeof(float2) * kWarps * kExWarp;   // exchange / power buffers
  b += sizeof(float2) * 1024;               // twiddles
  b += sizeof(float) * 1024;                // window
  b += sizeof(float) * ((tile_samples + 3) & ~3);
  b += sizeof(float) * ((nnz + 3) & ~3);
  b += sizeof(int) * 3 * kMaxMels + sizeof(float) * kMaxMels;
  b += sizeof(double) * 2 * 2 * kWarps;     // per-warp (sum, sum^2), double-buffered by tile parity
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

// in-range part of a tile that the bulk copy covers, tile-relative [lo, hi), both multiples of 4
__device__ __forceinline__ void bulk_range(long long base, long long L, int tile_samples, int* lo, int* hi) {
  const long long l = base < 0 ? -base : 0;
  long long h = L - base;
  h = h > tile_samples ? tile_samples : h;
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
  float* s_mw = s_wave + ((a.tile_samples + 3) & ~3);
  int* s_mk0 = reinterpret_cast<int*>(s_mw + ((a.nnz + 3) & ~3));
  int* s_mcnt = s_mk0 + kMaxMels;
  int* s_moff = s_mcnt + kMaxMels;
  float* s_mconst = reinterpret_cast<float*>(s_moff + kMaxMels);
  double* s_red = reinterpret_cast<double*>(s_mconst + kMaxMels);          // [2][kWarps][2]
  unsigned long long* s_bar = reinterpret_cast<unsigned long long*>(s_red + 4 * kWarps);
  void* bar_wave = s_bar;        // tile's waveform span has landed (tx bytes)
  void* bar_mel = s_bar + 1;     // every thread is done reading the previous tile's power spectra

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long total_tiles = a.B * a.tiles_per_clip;

  if (tid == 0) { mbar_init(bar_wave, 1); mbar_init(bar_mel, kThreads); }
  for (int i = tid; i < 1024; i += kThreads) { s_tw[i] = a.tw[i]; s_win[i] = a.win[i]; }
  for (int i = tid; i < a.nnz; i += kThreads) s_mw[i] = a.mw[i];
  for (int i = tid; i < a.n_mels; i += kThreads) {
    s_mk0[i] = a.mk0[i]; s_mcnt[i] = a.mcnt[i]; s_moff[i] = a.moff[i]; s_mconst[i] = a.mconst[i];
  }
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncthreads();

  // producer: bulk-copy the in-range span of a tile into s_wave (one elected thread)
  auto issue_tile = [&](long long tile) {
    const long long b = tile / a.tiles_per_clip;
    const int t0 = static_cast<int>(tile - b * a.tiles_per_clip) * kFramesPerTile;
    const long long base = static_cast<long long>(t0) * a.hop - kNfft / 2;
    int lo, hi;
    bulk_range(base, a.L, a.tile_samples, &lo, &hi);
    const unsigned bytes = static_cast<unsigned>(hi - lo) * 4u;
    mbar_expect_tx(bar_wave, bytes);
    if (bytes) bulk_g2s(s_wave + lo, a.wave + b * a.ld + base + lo, bytes, bar_wave);
  };
  if (a.bulk_ok && tid == 0 && blockIdx.x < total_tiles) issue_tile(blockIdx.x);

  unsigned it = 0;
  long long prev_tile = -1;
  for (long long tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
    const long long b = tile / a.tiles_per_clip;
    const int tile_in_clip = static_cast<int>(tile - b * a.tiles_per_clip);
    const int t0 = tile_in_clip * kFramesPerTile;
    const long long base = static_cast<long long>(t0) * a.hop - kNfft / 2;
    const float* src = a.wave + b * a.ld;

    // ---- waveform span: wait for the bulk copy, patch what it could not cover (reflect padding of
    //      np.pad(mode='reflect'): p<0 -> -p, p>L-1 -> 2(L-1)-p; tail remainder when L % 4 != 0)
    int lo = 0, hi = 0;
    if (a.bulk_ok) {
      bulk_range(base, a.L, a.tile_samples, &lo, &hi);
      mbar_wait(bar_wave, it & 1);
    } else {
      __syncthreads();                              // previous tile's input reads are done
    }
    if (lo > 0 || hi < a.tile_samples) {            // CTA-uniform
      const int last = static_cast<int>(a.L - 1);
      const int n_fix = lo + (a.tile_samples - hi);
      for (int q = tid; q < n_fix; q += kThreads) {
        const int i = q < lo ? q : hi + (q - lo);
        long long p = base + i;
        if (p < 0) p = -p;
        if (p > last) p = 2ll * last - p;
        p = p < 0 ? 0 : (p > last ? last : p);      // only frames >= T can land here; never read back
        s_wave[i] = __ldg(src + p);
      }
      __syncthreads();
    }

    float2* ex = s_ex + warp * kExWarp;
    const bool have = t0 + 2 * warp < a.T;          // warp-uniform
    float re[32], im[32];
    if (have) {
      const float* w0 = s_wave + (2 * warp) * a.hop + lane;
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
        if (lane == 0 && tile + gridDim.x < total_tiles) issue_tile(tile + gridDim.x);
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
    if (a.tile_sums && tid == 0 && prev_tile >= 0) {
      const double* r = s_red + ((it - 1) & 1) * 2 * kWarps;
      double r1 = 0.0, r2 = 0.0;
      for (int w = 0; w < kWarps; ++w) { r1 += r[2 * w]; r2 += r[2 * w + 1]; }
      a.tile_sums[2 * prev_tile] = r1;
      a.tile_sums[2 * prev_tile + 1] = r2;
    }

    // ---- sparse filterbank + log: thread = (band, frame pair); 32 bands per pass
    const int fp = tid & 7;
    const float2* pw = s_ex + fp * kExWarp;
    const int ta = t0 + 2 * fp;
    float* out = a.lms + ring_slot(a.slot0, b, a.cap) * (static_cast<long long>(a.n_mels) * a.T);
    float g2 = 1.f;
    if (a.gain) { const float g = a.gain[b]; g2 = g * g; }   // |g x|^2 = g^2 |x|^2: the per-clip gain is applied here
    double s1 = 0.0, s2 = 0.0;
    for (int mb = 0, pass = 0; mb < a.n_mels; mb += 32, ++pass) {
      // odd passes walk their 32 bands backwards so wide and narrow bands share a warp's time
      const int m = (pass & 1) ? mb + 31 - (tid >> 3) : mb + (tid >> 3);
      if (m < a.n_mels && ta < a.T) {
        const int k0 = s_mk0[m], cnt = s_mcnt[m];
        const float* wgt = s_mw + s_moff[m];
        float acc0 = 0.f, acc1 = 0.f;
#pragma unroll 4
        for (int j = 0; j < cnt; ++j) {
          const float w = wgt[j];
          const float2 p = pw[k0 + j];
          acc0 = fmaf(w, p.x, acc0);
          acc1 = fmaf(w, p.y, acc1);
        }
        const float c = s_mconst[m];
        acc0 = fmaf(acc0, g2, c);
        acc1 = fmaf(acc1, g2, c);
        const float l0 = a.out_kind ? acc0 : logf(acc0 + kLogEps);
        out[static_cast<long long>(m) * a.T + ta] = l0;
        s1 += l0; s2 += static_cast<double>(l0) * l0;
        if (ta + 1 < a.T) {
          const float l1 = a.out_kind ? acc1 : logf(acc1 + kLogEps);
          out[static_cast<long long>(m) * a.T + ta + 1] = l1;
          s1 += l1; s2 += static_cast<double>(l1) * l1;
        }
      }
    }
    mbar_arrive(bar_mel);                          // this thread no longer reads the power buffers
    if (a.tile_sums) {                             // deterministic per-tile (sum, sum of squares)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        s1 += __shfl_down_sync(0xffffffffu, s1, o);
        s2 += __shfl_down_sync(0xffffffffu, s2, o);
      }
      if (lane == 0) {
        double* r = s_red + (it & 1) * 2 * kWarps;
        r[2 * warp] = s1; r[2 * warp + 1] = s2;
      }
    }
    prev_tile = tile;
  }
  __syncthreads();
  if (a.tile_sums && tid == 0 && prev_tile >= 0) {
    const double* r = s_red + ((it - 1) & 1) * 2 * kWarps;
    double r1 = 0.0, r2 = 0.0;
    for (int w = 0; w < kWarps; ++w) { r1 += r[2 * w]; r2 += r[2 * w + 1]; }
    a.tile_sums[2 * prev_tile] = r1;
    a.tile_sums[2 * prev_tile + 1] = r2;
  }
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

extern "C" int64_t assl_plan_num_tiles(const assl_plan* p, int64_t n_samples) {
  const int64_t t = assl_plan_num_frames(p, n_samples);
  return t < 0 ? -1 : (t + assl::kFramesPerTile - 1) / assl::kFramesPerTile;
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
  if (T > (1ll << 30)) return ASSL_ERR_UNSUPPORTED;
  LogmelArgs a;
  a.wave = wave; a.gain = wave_gain; a.lms = lms; a.tile_sums = tile_sums;
  a.B = B; a.L = L; a.ld = ld; a.slot0 = slot0; a.cap = cap;
  a.T = static_cast<int>(T); a.hop = p->cfg.hop; a.n_mels = p->cfg.n_mels;
  a.tiles_per_clip = static_cast<int>((T + kFramesPerTile - 1) / kFramesPerTile);
  a.tile_samples = (kFramesPerTile - 1) * p->cfg.hop + kNfft;
  a.nnz = p->nnz; a.out_kind = out_kind;
  // the bulk (TMA) copy needs 16-byte aligned global addresses: clip starts and tile starts (t0*hop - 512, t0 % 16 == 0)
  a.bulk_ok = (reinterpret_cast<uintptr_t>(wave) % 16 == 0) && (ld % 4 == 0);
  a.win = p->d_win_half; a.tw = p->d_twiddle;
  a.mk0 = p->d_mel_k0; a.mcnt = p->d_mel_cnt; a.moff = p->d_mel_off; a.mw = p->d_mel_w; a.mconst = p->d_mel_const;
  const size_t smem = logmel_smem_bytes(a.tile_samples, a.nnz);
  if (smem > 227 * 1024) return ASSL_ERR_UNSUPPORTED;
  const long long total = B * a.tiles_per_clip;
  const unsigned grid = static_cast<unsigned>(total < 2ll * p->sm_count ? total : 2ll * p->sm_count);
  if (a.hop == 160) return launch_logmel<5>(a, smem, grid, as_stream(stream));
  return launch_logmel<0>(a, smem, grid, as_stream(stream));
}

extern "C" int assl_wave_l2_gain_f32(const float* wave, int64_t B, int64_t L, int64_t ld, float* gain, void* stream) {
  if (!wave || !gain || B < 0 || L < 1 || ld < L) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::wave_l2_gain_kernel<<<static_cast<unsigned>(B), 256, 0, assl::as_stream(stream)>>>(wave, L, ld, gain);
  return assl::launch_status();
}
