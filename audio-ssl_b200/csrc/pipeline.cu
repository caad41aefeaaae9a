// Whole path with HOST buffers: H2D -> log-mel -> running stats -> two views -> D2H.
// This is the entry a DataLoader-side caller of the reference would bind in place of
// BaseDataset.__getitem__'s per-clip CPU work (/root/reference/src/dataset/upstream_dataset.py:65,82,86).
//
// Three queues: uploads (s_in), kernels (the caller's stream), downloads (s_out).  The waveform / view
// buffers of the workspace are double-buffered by call parity so call k+1's upload and kernels overlap
// call k's download: PCIe is full duplex, so steady state costs max(upload, download) per batch instead
// of their sum.  An isolated call is additionally cut into chunks so chunk c+1 uploads while chunk c
// computes and chunk c-1 downloads (in a stream of calls whole-batch copies are faster, see below).  Kernels stay in order on the caller's stream (statistics scan and bank are sequential
// in clip order).  Host result buffers are complete after assl_pipeline_host_sync().
#include <cstdlib>
#include <new>

#include "common.cuh"

namespace {

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

inline int64_t ring_pitch(int64_t T) { return (T + 3) & ~static_cast<int64_t>(3); }   // 16-byte aligned ring rows

struct WsLayout {
  size_t state, stats_ring, lms_ring, tile_sums, clip_slot, mix_slot, alpha, crop, wave[2], pcm[2], view1[2], view2[2], total;
};

WsLayout layout(const assl_plan* p, int64_t B, int64_t L, int64_t cap) {
  const int64_t T = 1 + L / p->cfg.hop;
  const int64_t parts = assl_plan_num_tiles(p, L);
  const size_t plane = static_cast<size_t>(p->cfg.n_mels) * T * sizeof(float);
  const size_t ring_plane = static_cast<size_t>(p->cfg.n_mels) * ring_pitch(T) * sizeof(float);
  WsLayout w;
  size_t o = 0;
  w.state = o;      o = align_up(o + 8 * sizeof(double));
  w.stats_ring = o; o = align_up(o + static_cast<size_t>(cap) * 2 * sizeof(float));
  w.lms_ring = o;   o = align_up(o + static_cast<size_t>(cap) * ring_plane);
  w.tile_sums = o;  o = align_up(o + static_cast<size_t>(B) * parts * 2 * sizeof(double));
  w.clip_slot = o;  o = align_up(o + static_cast<size_t>(B) * sizeof(int32_t));
  w.mix_slot = o;   o = align_up(o + 2 * static_cast<size_t>(B) * 2 * sizeof(int32_t));   // x2: double-buffered
  w.alpha = o;      o = align_up(o + 2 * static_cast<size_t>(B) * 2 * sizeof(float));
  w.crop = o;       o = align_up(o + 2 * static_cast<size_t>(B) * 8 * sizeof(int32_t));
  for (int i = 0; i < 2; ++i) {
    w.wave[i] = o;  o = align_up(o + static_cast<size_t>(B) * L * sizeof(float));
    w.pcm[i] = o;   o = align_up(o + static_cast<size_t>(B) * L * sizeof(int16_t));     // 16-bit PCM uploads land here
    w.view1[i] = o; o = align_up(o + static_cast<size_t>(B) * plane);
    w.view2[i] = o; o = align_up(o + static_cast<size_t>(B) * plane);
  }
  w.total = o;
  return w;
}

// 16-bit PCM -> fp32 in [-1, 1): x / 32768 (exact), what soundfile / torchaudio.load hand the reference for a 16-bit
// file (librosa.core.load, /root/reference/src/dataset/upstream_dataset.py:55).  Eight samples per thread where the
// pointers allow 16-byte words.
__global__ void __launch_bounds__(256) pcm16_to_f32_kernel(const int16_t* __restrict__ in, float* __restrict__ out, long long n, int vec) {
  const long long i = blockIdx.x * 256ll + threadIdx.x;
  constexpr float k = 1.0f / 32768.0f;
  if (vec) {
    if (8 * i + 8 <= n) {
      const int4 v = __ldg(reinterpret_cast<const int4*>(in) + i);
      const int w[4] = {v.x, v.y, v.z, v.w};
      float f[8];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        f[2 * j] = static_cast<float>(static_cast<short>(w[j] & 0xffff)) * k;
        f[2 * j + 1] = static_cast<float>(w[j] >> 16) * k;
      }
      float4* o = reinterpret_cast<float4*>(out) + 2 * i;
      o[0] = make_float4(f[0], f[1], f[2], f[3]);
      o[1] = make_float4(f[4], f[5], f[6], f[7]);
    } else {
      for (long long j = 8 * i; j < n; ++j) out[j] = static_cast<float>(in[j]) * k;
    }
  } else if (i < n) {
    out[i] = static_cast<float>(in[i]) * k;
  }
}

__global__ void iota_slots_kernel(int32_t* slot, long long B, long long slot0, long long cap) {
  const long long b = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (b < B) slot[b] = static_cast<int32_t>(assl::ring_slot(slot0, b, cap));
}

}  // namespace

// The host-buffer pipeline's own state: copy streams, events, call parity.  One caller at a time per pipeline; the
// plan it was created on stays immutable and can be shared.
struct assl_pipeline {
  static constexpr int kMaxChunks = 8;
  const assl_plan* plan = nullptr;
  assl_exchange* exchange = nullptr;       // statistics exchange of a clip-sharded multi-GPU run, or null
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_fork = nullptr;
  cudaEvent_t ev_in[kMaxChunks] = {}, ev_out[kMaxChunks] = {};
  cudaEvent_t ev_kern[2] = {}, ev_d2h[2] = {};   // per buffer parity: kernels / downloads of that call done
  long long calls = 0;
};

extern "C" int assl_pipeline_create(assl_pipeline** out, const assl_plan* plan, assl_exchange* exchange) {
  if (!out || !plan) return ASSL_ERR_INVALID_ARG;
  *out = nullptr;
  assl_pipeline* pipe = new (std::nothrow) assl_pipeline();
  if (!pipe) return ASSL_ERR_CUDA;
  pipe->plan = plan; pipe->exchange = exchange;
  bool ok = cudaStreamCreateWithFlags(&pipe->s_in, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&pipe->s_out, cudaStreamNonBlocking) == cudaSuccess;
  auto mk = [&](cudaEvent_t* e) { ok = ok && cudaEventCreateWithFlags(e, cudaEventDisableTiming) == cudaSuccess; };
  mk(&pipe->ev_fork);
  for (int i = 0; i < 2; ++i) { mk(&pipe->ev_kern[i]); mk(&pipe->ev_d2h[i]); }
  for (int i = 0; i < assl_pipeline::kMaxChunks; ++i) { mk(&pipe->ev_in[i]); mk(&pipe->ev_out[i]); }
  if (!ok) { assl_pipeline_destroy(pipe); cudaGetLastError(); return ASSL_ERR_CUDA; }
  *out = pipe;
  return ASSL_OK;
}

extern "C" int assl_pipeline_destroy(assl_pipeline* pipe) {
  if (!pipe) return ASSL_OK;
  if (pipe->s_in) cudaStreamDestroy(pipe->s_in);
  if (pipe->s_out) cudaStreamDestroy(pipe->s_out);
  auto rm = [](cudaEvent_t e) { if (e) cudaEventDestroy(e); };
  rm(pipe->ev_fork);
  for (int i = 0; i < 2; ++i) { rm(pipe->ev_kern[i]); rm(pipe->ev_d2h[i]); }
  for (int i = 0; i < assl_pipeline::kMaxChunks; ++i) { rm(pipe->ev_in[i]); rm(pipe->ev_out[i]); }
  delete pipe;
  return ASSL_OK;
}

extern "C" int64_t assl_pipeline_workspace_bytes(const assl_plan* p, int64_t B, int64_t L, int64_t cap) {
  if (!p || B < 1 || cap < B || L <= p->cfg.n_fft / 2) return -1;
  return static_cast<int64_t>(layout(p, B, L, cap).total);
}

extern "C" int assl_pcm16_to_f32(const int16_t* pcm, int64_t n, float* out, void* stream) {
  if (!pcm || !out || n < 0) return ASSL_ERR_INVALID_ARG;
  if (n == 0) return ASSL_OK;
  const int vec = (reinterpret_cast<uintptr_t>(pcm) % 16 == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);
  const long long threads = vec ? (n + 7) / 8 : n;
  if ((threads + 255) / 256 > 0x7fffffffll) return ASSL_ERR_UNSUPPORTED;
  pcm16_to_f32_kernel<<<static_cast<unsigned>((threads + 255) / 256), 256, 0, assl::as_stream(stream)>>>(pcm, out, n, vec);
  return assl::launch_status();
}

extern "C" int assl_pipeline_host_sync(assl_pipeline* pipe) {
  if (!pipe) return ASSL_ERR_INVALID_ARG;
  bool ok = cudaStreamSynchronize(pipe->s_in) == cudaSuccess;
  ok = cudaStreamSynchronize(pipe->s_out) == cudaSuccess && ok;
  return ok ? ASSL_OK : ASSL_ERR_CUDA;
}

static int pipeline_host(assl_pipeline* pipe, const assl_view_cfg* vc, void* ws, int64_t ws_bytes, const void* host_wave,
                         bool pcm16, int64_t B, int64_t L, int64_t slot0, int64_t cap, int64_t max_update,
                         const int32_t* host_mix_slot, const float* host_alpha, const int32_t* host_crop, float* host_view1,
                         float* host_view2, float** dev_view1, float** dev_view2, void* stream) {
  using namespace assl;
  if (!pipe || !vc || !ws || !host_wave || !host_mix_slot || !host_alpha || !host_crop) return ASSL_ERR_INVALID_ARG;
  const assl_plan* p = pipe->plan;
  if (B < 1 || cap < B || slot0 < 0 || slot0 >= cap || L <= p->cfg.n_fft / 2) return ASSL_ERR_INVALID_ARG;
  const int64_t T = 1 + L / p->cfg.hop;
  if (vc->n_freq != p->cfg.n_mels || vc->n_time != T) return ASSL_ERR_INVALID_ARG;
  const WsLayout w = layout(p, B, L, cap);
  if (static_cast<size_t>(ws_bytes) < w.total) return ASSL_ERR_INVALID_ARG;
  assl_view_cfg vcr = *vc;
  vcr.ring_pitch = static_cast<int32_t>(ring_pitch(T));            // the ring lives in the workspace: its layout is ours
  int rc = ASSL_OK;
  cudaStream_t st = as_stream(stream);
  const int par = static_cast<int>(pipe->calls & 1);
  char* base = static_cast<char*>(ws);
  double* state = reinterpret_cast<double*>(base + w.state);
  float* stats = reinterpret_cast<float*>(base + w.stats_ring);
  float* lms = reinterpret_cast<float*>(base + w.lms_ring);
  double* tsums = reinterpret_cast<double*>(base + w.tile_sums);
  int32_t* clip_slot = reinterpret_cast<int32_t*>(base + w.clip_slot);
  int32_t* mix_slot = reinterpret_cast<int32_t*>(base + w.mix_slot) + par * B * 2;
  float* alpha = reinterpret_cast<float*>(base + w.alpha) + par * B * 2;
  int32_t* crop = reinterpret_cast<int32_t*>(base + w.crop) + par * B * 8;
  float* wave = reinterpret_cast<float*>(base + w.wave[par]);
  int16_t* pcm = reinterpret_cast<int16_t*>(base + w.pcm[par]);
  float* v1 = reinterpret_cast<float*>(base + w.view1[par]);
  float* v2 = reinterpret_cast<float*>(base + w.view2[par]);
  const int64_t elems = static_cast<int64_t>(p->cfg.n_mels) * T;
  const size_t plane = static_cast<size_t>(elems) * sizeof(float);
  const int64_t n_parts = assl_plan_num_tiles(p, L);
  // Chunks: an isolated call is cut in four so its upload, kernels and download overlap inside the call; when the
  // previous call is still in flight (a stream of calls) the double-buffered workspace already overlaps call k+1's
  // upload with call k's kernels and download, and whole-batch copies use the link better than quarter-batch ones
  // (measured on B200, 1024 x 1 s: 0.78 M clips/s with one chunk, 0.70 M with four).
  int n_chunks = B >= 256 ? 4 : (B >= 32 ? 2 : 1);
  if (pipe->calls > 0 && cudaEventQuery(pipe->ev_d2h[par ^ 1]) == cudaErrorNotReady) n_chunks = 1;
  if (pipe->exchange) n_chunks = 1;        // one exchange per call: every rank brings its whole batch to it
  if (const char* e = getenv("ASSL_PIPE_CHUNKS")) { const int v = atoi(e); if (v >= 1 && v <= assl_pipeline::kMaxChunks && v <= B) n_chunks = v; }

  bool ok = true;
  auto cp = [&](void* dst, const void* src, size_t bytes, cudaMemcpyKind kind, cudaStream_t s) {
    ok = ok && cudaMemcpyAsync(dst, src, bytes, kind, s) == cudaSuccess;
  };
  // order the uploads against whatever the caller queued on its stream before the first calls (its own
  // zero-fill of the workspace state ...), later against the kernels of the call that last used this parity
  if (pipe->calls < 2) {
    ok = ok && cudaEventRecord(pipe->ev_fork, st) == cudaSuccess;
    ok = ok && cudaStreamWaitEvent(pipe->s_in, pipe->ev_fork, 0) == cudaSuccess;
  } else {
    ok = ok && cudaStreamWaitEvent(pipe->s_in, pipe->ev_kern[par], 0) == cudaSuccess;
  }
  cp(mix_slot, host_mix_slot, static_cast<size_t>(B) * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, pipe->s_in);
  cp(alpha, host_alpha, static_cast<size_t>(B) * 2 * sizeof(float), cudaMemcpyHostToDevice, pipe->s_in);
  cp(crop, host_crop, static_cast<size_t>(B) * 8 * sizeof(int32_t), cudaMemcpyHostToDevice, pipe->s_in);
  iota_slots_kernel<<<static_cast<unsigned>((B + 255) / 256), 256, 0, st>>>(clip_slot, B, slot0, cap);
  // the view buffers of this parity were downloaded two calls ago; kernels that overwrite them wait for that
  if (pipe->calls >= 2) ok = ok && cudaStreamWaitEvent(st, pipe->ev_d2h[par], 0) == cudaSuccess;

  for (int c = 0; c < n_chunks && ok && rc == ASSL_OK; ++c) {
    const int64_t b0 = B * c / n_chunks, b1 = B * (c + 1) / n_chunks, nb = b1 - b0;
    if (nb == 0) continue;
    if (pcm16) {   // half the upload; widened on the device, on the upload stream (behind the link, not before the kernels)
      cp(pcm + b0 * L, static_cast<const int16_t*>(host_wave) + b0 * L, static_cast<size_t>(nb) * L * sizeof(int16_t),
         cudaMemcpyHostToDevice, pipe->s_in);
      if (ok && assl_pcm16_to_f32(pcm + b0 * L, nb * L, wave + b0 * L, pipe->s_in) != ASSL_OK) ok = false;
    } else {
      cp(wave + b0 * L, static_cast<const float*>(host_wave) + b0 * L, static_cast<size_t>(nb) * L * sizeof(float),
         cudaMemcpyHostToDevice, pipe->s_in);
    }
    ok = ok && cudaEventRecord(pipe->ev_in[c], pipe->s_in) == cudaSuccess;
    ok = ok && cudaStreamWaitEvent(st, pipe->ev_in[c], 0) == cudaSuccess;
    const int64_t s0 = (slot0 + b0) % cap;
    rc = assl_logmel_pitched_f32(p, wave + b0 * L, nullptr, nb, L, L, lms, s0, cap, ring_pitch(T), tsums + b0 * n_parts * 2, 0,
                                 stream);
    if (rc == ASSL_OK && pipe->exchange)   // sharded run: this rank's clips + every peer's, in global clip order
      rc = assl_running_stats_exchange(pipe->exchange, state, tsums, nb, n_parts, elems, max_update, stats, s0, cap, 0, stream);
    else if (rc == ASSL_OK)
      rc = assl_running_stats(state, tsums + b0 * n_parts * 2, nb, n_parts, elems, max_update, 0, nb, stats, s0, cap, stream);
    if (rc == ASSL_OK)
      rc = assl_two_views_f32(&vcr, lms, stats, cap, clip_slot + b0, mix_slot + b0 * 2, alpha + b0 * 2, crop + b0 * 8, nb,
                              v1 + b0 * elems, v2 + b0 * elems, stream);
    if (rc == ASSL_OK && (host_view1 || host_view2)) {
      ok = ok && cudaEventRecord(pipe->ev_out[c], st) == cudaSuccess;
      ok = ok && cudaStreamWaitEvent(pipe->s_out, pipe->ev_out[c], 0) == cudaSuccess;
      if (host_view1) cp(host_view1 + b0 * elems, v1 + b0 * elems, static_cast<size_t>(nb) * plane, cudaMemcpyDeviceToHost, pipe->s_out);
      if (host_view2) cp(host_view2 + b0 * elems, v2 + b0 * elems, static_cast<size_t>(nb) * plane, cudaMemcpyDeviceToHost, pipe->s_out);
    }
  }
  if (rc != ASSL_OK) return rc;
  ok = ok && cudaEventRecord(pipe->ev_kern[par], st) == cudaSuccess;
  ok = ok && cudaEventRecord(pipe->ev_d2h[par], pipe->s_out) == cudaSuccess;
  pipe->calls += 1;
  if (dev_view1) *dev_view1 = v1;
  if (dev_view2) *dev_view2 = v2;
  return ok ? launch_status() : ASSL_ERR_CUDA;
}

extern "C" int assl_pipeline_host_f32(assl_pipeline* pipe, const assl_view_cfg* vc, void* ws, int64_t ws_bytes,
                                      const float* host_wave, int64_t B, int64_t L, int64_t slot0, int64_t cap,
                                      int64_t max_update, const int32_t* host_mix_slot, const float* host_alpha,
                                      const int32_t* host_crop, float* host_view1, float* host_view2,
                                      float** dev_view1, float** dev_view2, void* stream) {
  return pipeline_host(pipe, vc, ws, ws_bytes, host_wave, false, B, L, slot0, cap, max_update, host_mix_slot, host_alpha,
                       host_crop, host_view1, host_view2, dev_view1, dev_view2, stream);
}

extern "C" int assl_pipeline_host_pcm16(assl_pipeline* pipe, const assl_view_cfg* vc, void* ws, int64_t ws_bytes,
                                        const int16_t* host_wave, int64_t B, int64_t L, int64_t slot0, int64_t cap,
                                        int64_t max_update, const int32_t* host_mix_slot, const float* host_alpha,
                                        const int32_t* host_crop, float* host_view1, float* host_view2,
                                        float** dev_view1, float** dev_view2, void* stream) {
  return pipeline_host(pipe, vc, ws, ws_bytes, host_wave, true, B, L, slot0, cap, max_update, host_mix_slot, host_alpha,
                       host_crop, host_view1, host_view2, dev_view1, dev_view2, stream);
}
