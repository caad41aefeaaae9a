// Plan construction: the host-side restatement of librosa.filters.mel as the reference calls it
// (/root/reference/src/utils/utils.py:24) plus the window / twiddle tables of the FFT kernel.
#include <cmath>
#include <cstring>
#include <new>

#include "common.cuh"

namespace {

const double kPi = 3.14159265358979323846;

// librosa.convert.hz_to_mel / mel_to_hz, htk=False (Slaney): linear below 1 kHz, log above.
double hz_to_mel(double f) {
  const double f_sp = 200.0 / 3.0, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = std::log(6.4) / 27.0;
  return f >= min_log_hz ? min_log_mel + std::log(f / min_log_hz) / logstep : f / f_sp;
}
double mel_to_hz(double m) {
  const double f_sp = 200.0 / 3.0, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = std::log(6.4) / 27.0;
  return m >= min_log_mel ? min_log_hz * std::exp(logstep * (m - min_log_mel)) : f_sp * m;
}

// np.linspace(a, b, n)[i] the way numpy evaluates it (start + i*step, last point pinned).
double linspace_at(double a, double b, int n, int i) {
  if (n == 1) return a;
  if (i == n - 1) return b;
  const double step = (b - a) / (n - 1);
  return a + i * step;
}

template <typename T>
int upload(T** dst, const std::vector<T>& src) {
  if (cudaMalloc(reinterpret_cast<void**>(dst), src.size() * sizeof(T)) != cudaSuccess) return ASSL_ERR_CUDA;
  if (cudaMemcpy(*dst, src.data(), src.size() * sizeof(T), cudaMemcpyHostToDevice) != cudaSuccess)
    return ASSL_ERR_CUDA;
  return ASSL_OK;
}

}  // namespace

extern "C" int assl_version(void) { return ASSL_VERSION; }

extern "C" const char* assl_status_string(int status) {
  switch (status) {
    case ASSL_OK: return "ok";
    case ASSL_ERR_INVALID_ARG: return "invalid argument";
    case ASSL_ERR_UNSUPPORTED: return "unsupported shape or configuration";
    case ASSL_ERR_CUDA: return "CUDA runtime error";
    case ASSL_ERR_NO_DEVICE: return "no CUDA device";
    default: return "unknown status";
  }
}

extern "C" int assl_plan_create(assl_plan** out, const assl_cfg* cfg) {
  if (!out || !cfg) return ASSL_ERR_INVALID_ARG;
  *out = nullptr;
  if (cfg->n_fft != assl::kNfft) return ASSL_ERR_UNSUPPORTED;
  if (cfg->hop < 1 || cfg->hop > cfg->n_fft) return ASSL_ERR_INVALID_ARG;
  if (cfg->n_mels < 1 || cfg->n_mels > assl::kMaxMels) return ASSL_ERR_UNSUPPORTED;
  if (cfg->sample_rate < 1) return ASSL_ERR_INVALID_ARG;
  const double sr = cfg->sample_rate;
  const double fmin = cfg->fmin, fmax = cfg->fmax > 0 ? cfg->fmax : sr / 2;
  if (fmin < 0 || fmax <= fmin || fmax > sr / 2 + 1e-9) return ASSL_ERR_INVALID_ARG;

  int device = 0;
  if (cudaGetDevice(&device) != cudaSuccess) return ASSL_ERR_NO_DEVICE;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return ASSL_ERR_NO_DEVICE;

  assl_plan* p = new (std::nothrow) assl_plan();
  if (!p) return ASSL_ERR_CUDA;
  p->cfg = *cfg;
  p->cfg.fmax = static_cast<float>(fmax);
  p->device = device;
  p->sm_count = prop.multiProcessorCount;

  const int n_mels = cfg->n_mels, n_bins = assl::kBins, n_fft = cfg->n_fft;

  // ---- filterbank, librosa 0.8.1 filters.mel: fp64 ramps -> fp32 weights -> *= enorm (fp64) -> fp32
  std::vector<double> mel_f(n_mels + 2);
  const double mlo = hz_to_mel(fmin), mhi = hz_to_mel(fmax);
  for (int i = 0; i < n_mels + 2; ++i) mel_f[i] = mel_to_hz(linspace_at(mlo, mhi, n_mels + 2, i));
  p->h_fb.assign(static_cast<size_t>(n_mels) * n_bins, 0.f);
  for (int m = 0; m < n_mels; ++m) {
    const double fd0 = mel_f[m + 1] - mel_f[m], fd1 = mel_f[m + 2] - mel_f[m + 1];
    const double enorm = 2.0 / (mel_f[m + 2] - mel_f[m]);
    for (int k = 0; k < n_bins; ++k) {
      const double fk = linspace_at(0.0, sr / 2, n_bins, k);
      const double lower = -(mel_f[m] - fk) / fd0;
      const double upper = (mel_f[m + 2] - fk) / fd1;
      const float w32 = static_cast<float>(std::fmax(0.0, std::fmin(lower, upper)));
      p->h_fb[static_cast<size_t>(m) * n_bins + k] = static_cast<float>(static_cast<double>(w32) * enorm);
    }
  }
  // sparse form: each band's support is one contiguous run of bins
  std::vector<int> k0(n_mels), cnt(n_mels), off(n_mels);
  std::vector<float> w;
  for (int m = 0; m < n_mels; ++m) {
    int lo = n_bins, hi = -1;
    for (int k = 0; k < n_bins; ++k)
      if (p->h_fb[static_cast<size_t>(m) * n_bins + k] != 0.f) { lo = lo < k ? lo : k; hi = k; }
    k0[m] = hi >= 0 ? lo : 0;
    cnt[m] = hi >= 0 ? hi - lo + 1 : 0;
    off[m] = static_cast<int>(w.size());
    for (int k = 0; k < cnt[m]; ++k) w.push_back(p->h_fb[static_cast<size_t>(m) * n_bins + k0[m] + k]);
  }
  std::vector<float> mconst(n_mels);
  for (int m = 0; m < n_mels; ++m) {
    float acc = 0.f;
    for (int k = 0; k < cnt[m]; ++k) acc += w[off[m] + k] * assl::kPowerEps;
    mconst[m] = acc;
  }
  if (w.empty()) w.push_back(0.f);
  p->nnz = static_cast<int>(w.size());

  // ---- window (scipy get_window('hann', fftbins=True), fp64 -> fp32) with the unpack's 1/2 folded in
  std::vector<float> win(n_fft);
  for (int n = 0; n < n_fft; ++n) win[n] = static_cast<float>(0.5 * (0.5 - 0.5 * std::cos(2.0 * kPi * n / n_fft)));
  // ---- inter-stage twiddles of the 32 x 32 decomposition
  std::vector<float2> tw(32 * 32);
  for (int k2 = 0; k2 < 32; ++k2)
    for (int n1 = 0; n1 < 32; ++n1) {
      const double ang = -2.0 * kPi * static_cast<double>(n1 * k2) / n_fft;
      tw[k2 * 32 + n1] = make_float2(static_cast<float>(std::cos(ang)), static_cast<float>(std::sin(ang)));
    }

  int rc = upload(&p->d_win_half, win);
  if (rc == ASSL_OK) rc = upload(&p->d_twiddle, tw);
  if (rc == ASSL_OK) rc = upload(&p->d_mel_k0, k0);
  if (rc == ASSL_OK) rc = upload(&p->d_mel_cnt, cnt);
  if (rc == ASSL_OK) rc = upload(&p->d_mel_off, off);
  if (rc == ASSL_OK) rc = upload(&p->d_mel_w, w);
  if (rc == ASSL_OK) rc = upload(&p->d_mel_const, mconst);
  if (rc != ASSL_OK) { assl_plan_destroy(p); return rc; }
  *out = p;
  return ASSL_OK;
}

extern "C" int assl_plan_destroy(assl_plan* p) {
  if (!p) return ASSL_OK;
  cudaFree(p->d_win_half);
  cudaFree(p->d_twiddle);
  cudaFree(p->d_mel_k0);
  cudaFree(p->d_mel_cnt);
  cudaFree(p->d_mel_off);
  cudaFree(p->d_mel_w);
  cudaFree(p->d_mel_const);
  if (p->aux_ready) {
    cudaStreamDestroy(p->s_in); cudaStreamDestroy(p->s_out);
    cudaEventDestroy(p->ev_fork);
    for (int i = 0; i < 2; ++i) { cudaEventDestroy(p->ev_kern[i]); cudaEventDestroy(p->ev_d2h[i]); }
    for (int i = 0; i < assl_plan::kMaxChunks; ++i) { cudaEventDestroy(p->ev_in[i]); cudaEventDestroy(p->ev_out[i]); }
  }
  delete p;
  return ASSL_OK;
}

extern "C" int64_t assl_plan_num_frames(const assl_plan* p, int64_t n_samples) {
  if (!p || n_samples <= p->cfg.n_fft / 2) return -1;
  return 1 + n_samples / p->cfg.hop;
}

extern "C" int assl_plan_filterbank(const assl_plan* p, float* host_out) {
  if (!p || !host_out) return ASSL_ERR_INVALID_ARG;
  std::memcpy(host_out, p->h_fb.data(), p->h_fb.size() * sizeof(float));
  return ASSL_OK;
}
