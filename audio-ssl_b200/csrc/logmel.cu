// Fused log-mel front end: frame + periodic Hann + rFFT + |X|^2 + sparse mel + log, one kernel.
//
// Replaces MelSpectrogramLibrosa.__call__ (/root/reference/src/utils/utils.py:26-28) and
// extract_log_mel_spectrogram (/root/reference/src/utils/utils.py:43-49).
//
// Work decomposition
//   persistent CTAs (2 per SM) loop over tiles of kFramesPerTile consecutive frames of one clip;
//   the tile's waveform span (with librosa's reflect padding resolved on load) is staged once in
//   shared memory, so the 6.4x overlap between hop-160 frames never goes back to L2/HBM;
//   each warp transforms TWO real frames at once as one 1024-point complex FFT laid out 32 x 32:
//   lane n1 holds samples n1 + 32*n2 (stride-32, conflict-free LDS), runs a register-resident
//   32-point FFT (fft32_gen.cuh), multiplies by the inter-stage twiddles, transposes through a
//   padded per-warp shared buffer, runs the second 32-point FFT, and unpacks the two spectra
//   with one shuffle per partner value (bin k of lane l pairs with bin 1024-k of lane 32-l);
//   the two power spectra go back to the warp's own shared buffer, then the whole CTA applies
//   the sparse (triangular-support) filterbank, takes the log and writes [n_mels, T] row-major.
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
  int T, hop, n_mels, tiles_per_clip, tile_samples, nnz, out_kind;
  const float* win;
  const float2* tw;
  const int* mk0;
  const int* mcnt;
  const int* moff;
  const float* mw;
};

__host__ __device__ inline size_t logmel_smem_bytes(int tile_samples, int nnz, int n_mels) {
  size_t b = 0;
  b += sizeof(float2) * kWarps * kExWarp;   // exchange / power buffers
  b += sizeof(float2) * 1024;               // twiddles
  b += sizeof(float) * 1024;                // window
  b += sizeof(float) * ((tile_samples + 3) & ~3);
  b += sizeof(float) * ((nnz + 3) & ~3);
  b += sizeof(int) * 3 * kMaxMels;
  b += sizeof(double) * 2 * kWarps;
  (void)n_mels;
  return b;
}

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
  double* s_red = reinterpret_cast<double*>(s_moff + kMaxMels);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

  // plan constants -> shared, once per CTA
  for (int i = tid; i < 1024; i += kThreads) { s_tw[i] = a.tw[i]; s_win[i] = a.win[i]; }
  for (int i = tid; i < a.nnz; i += kThreads) s_mw[i] = a.mw[i];
  for (int i = tid; i < a.n_mels; i += kThreads) { s_mk0[i] = a.mk0[i]; s_mcnt[i] = a.mcnt[i]; s_moff[i] = a.moff[i]; }

  const long long total_tiles = a.B * a.tiles_per_clip;
  for (long long tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    const long long b = tile / a.tiles_per_clip;
    const int tile_in_clip = static_cast<int>(tile - b * a.tiles_per_clip);
    const int t0 = tile_in_clip * kFramesPerTile;

    __syncthreads();   // previous tile fully consumed (and the constants above are visible)

    // ---- stage the waveform span; p = hop*t + n - 512 reflected without edge repeat (np.pad 'reflect')
    {
      const float* src = a.wave + b * a.ld;
      const float g = a.gain ? a.gain[b] : 1.0f;
      const long long base = static_cast<long long>(t0) * a.hop - kNfft / 2;
      const long long last = a.L - 1;
      for (int i = tid; i < a.tile_samples; i += kThreads) {
        long long p = base + i;
        if (p < 0) p = -p;
        if (p > last) p = 2 * last - p;
        p = p < 0 ? 0 : (p > last ? last : p);     // only frames >= T can land here; never read back
        s_wave[i] = __ldg(src + p) * g;
      }
    }
    __syncthreads();

    float2* ex = s_ex + warp * kExWarp;
    if (t0 + 2 * warp < a.T) {                    // warp-uniform
      float re[32], im[32];
      const float* w0 = s_wave + (2 * warp) * a.hop + lane;
      const float* w1 = w0 + a.hop;
#pragma unroll
      for (int n2 = 0; n2 < 32; ++n2) {
        const float w = s_win[lane + 32 * n2];
        re[n2] = w * w0[32 * n2];
        im[n2] = w * w1[32 * n2];
      }
      fft32(re, im);                               // over n2 -> k2, this lane is n1
#pragma unroll
      for (int k2 = 1; k2 < 32; ++k2) {
        const float2 t = s_tw[k2 * 32 + lane];
        const float r = re[k2] * t.x - im[k2] * t.y;
        const float i = re[k2] * t.y + im[k2] * t.x;
        re[k2] = r; im[k2] = i;
      }
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
        p1[k1] = sr * sr + di * di + kPowerEps;
        p2[k1] = si * si + dr * dr + kPowerEps;
      }
      __syncwarp();                                // everyone is done reading the exchange buffer
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) ex[lane + 32 * k1] = make_float2(p1[k1], p2[k1]);
      if (lane == 0) {                             // Nyquist bin: its own partner
        const float nr = 2.f * re[16], ni = 2.f * im[16];
        ex[512] = make_float2(nr * nr + kPowerEps, ni * ni + kPowerEps);
      }
    }
    __syncthreads();

    // ---- sparse filterbank + log: thread = (band, frame pair); 32 bands per pass
    const int fp = tid & 7;
    const float2* pw = s_ex + fp * kExWarp;
    const int ta = t0 + 2 * fp;
    float* out = a.lms + ring_slot(a.slot0, b, a.cap) * (static_cast<long long>(a.n_mels) * a.T);
    double s1 = 0.0, s2 = 0.0;
    for (int mb = 0, pass = 0; mb < a.n_mels; mb += 32, ++pass) {
      // odd passes walk their 32 bands backwards so wide and narrow bands share a warp's time
      const int m = (pass & 1) ? mb + 31 - (tid >> 3) : mb + (tid >> 3);
      if (m < a.n_mels && ta < a.T) {
        const int k0 = s_mk0[m], cnt = s_mcnt[m];
        const float* wgt = s_mw + s_moff[m];
        float acc0 = 0.f, acc1 = 0.f;
        for (int j = 0; j < cnt; ++j) {
          const float w = wgt[j];
          const float2 p = pw[k0 + j];
          acc0 = fmaf(w, p.x, acc0);
          acc1 = fmaf(w, p.y, acc1);
        }
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
    if (a.tile_sums) {                             // deterministic per-tile (sum, sum of squares)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        s1 += __shfl_down_sync(0xffffffffu, s1, o);
        s2 += __shfl_down_sync(0xffffffffu, s2, o);
      }
      if (lane == 0) { s_red[2 * warp] = s1; s_red[2 * warp + 1] = s2; }
      __syncthreads();
      if (tid == 0) {
        double r1 = 0.0, r2 = 0.0;
        for (int w = 0; w < kWarps; ++w) { r1 += s_red[2 * w]; r2 += s_red[2 * w + 1]; }
        a.tile_sums[2 * tile] = r1;
        a.tile_sums[2 * tile + 1] = r2;
      }
    }
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
  a.win = p->d_win_half; a.tw = p->d_twiddle;
  a.mk0 = p->d_mel_k0; a.mcnt = p->d_mel_cnt; a.moff = p->d_mel_off; a.mw = p->d_mel_w;
  const size_t smem = logmel_smem_bytes(a.tile_samples, a.nnz, a.n_mels);
  if (smem > 227 * 1024) return ASSL_ERR_UNSUPPORTED;
  if (cudaFuncSetAttribute(logmel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  const long long total = B * a.tiles_per_clip;
  const long long grid = total < 2ll * p->sm_count ? total : 2ll * p->sm_count;
  logmel_kernel<<<static_cast<unsigned>(grid), kThreads, smem, as_stream(stream)>>>(a);
  return launch_status();
}

extern "C" int assl_wave_l2_gain_f32(const float* wave, int64_t B, int64_t L, int64_t ld, float* gain, void* stream) {
  if (!wave || !gain || B < 0 || L < 1 || ld < L) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::wave_l2_gain_kernel<<<static_cast<unsigned>(B), 256, 0, assl::as_stream(stream)>>>(wave, L, ld, gain);
  return assl::launch_status();
}
