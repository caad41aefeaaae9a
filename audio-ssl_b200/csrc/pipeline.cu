// Whole path with HOST buffers: H2D -> log-mel -> running stats -> two views -> D2H, all enqueued
// on one stream.  This is the entry a DataLoader-side caller of the reference would bind in place
// of BaseDataset.__getitem__'s per-clip CPU work (/root/reference/src/dataset/upstream_dataset.py:65,82,86).
#include "common.cuh"

namespace {

inline size_t align_up(size_t v, size_t a = 256) { return (v + a - 1) / a * a; }

struct WsLayout {
  size_t state, stats_ring, lms_ring, wave, tile_sums, clip_slot, mix_slot, alpha, crop, view1, view2, total;
};

WsLayout layout(const assl_plan* p, int64_t B, int64_t L, int64_t cap) {
  const int64_t T = 1 + L / p->cfg.hop;
  const int64_t tiles = assl_plan_num_tiles(p, L);
  const size_t plane = static_cast<size_t>(p->cfg.n_mels) * T * sizeof(float);
  WsLayout w;
  size_t o = 0;
  w.state = o;      o = align_up(o + 8 * sizeof(double));
  w.stats_ring = o; o = align_up(o + static_cast<size_t>(cap) * 2 * sizeof(float));
  w.lms_ring = o;   o = align_up(o + static_cast<size_t>(cap) * plane);
  w.wave = o;       o = align_up(o + static_cast<size_t>(B) * L * sizeof(float));
  w.tile_sums = o;  o = align_up(o + static_cast<size_t>(B) * tiles * 2 * sizeof(double));
  w.clip_slot = o;  o = align_up(o + static_cast<size_t>(B) * sizeof(int32_t));
  w.mix_slot = o;   o = align_up(o + static_cast<size_t>(B) * 2 * sizeof(int32_t));
  w.alpha = o;      o = align_up(o + static_cast<size_t>(B) * 2 * sizeof(float));
  w.crop = o;       o = align_up(o + static_cast<size_t>(B) * 8 * sizeof(int32_t));
  w.view1 = o;      o = align_up(o + static_cast<size_t>(B) * plane);
  w.view2 = o;      o = align_up(o + static_cast<size_t>(B) * plane);
  w.total = o;
  return w;
}

__global__ void iota_slots_kernel(int32_t* slot, long long B, long long slot0, long long cap) {
  const long long b = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (b < B) slot[b] = static_cast<int32_t>(assl::ring_slot(slot0, b, cap));
}

}  // namespace

extern "C" int64_t assl_pipeline_workspace_bytes(const assl_plan* p, int64_t B, int64_t L, int64_t cap) {
  if (!p || B < 1 || cap < B || L <= p->cfg.n_fft / 2) return -1;
  return static_cast<int64_t>(layout(p, B, L, cap).total);
}

extern "C" int assl_pipeline_host_f32(const assl_plan* p, const assl_view_cfg* vc, void* ws, int64_t ws_bytes,
                                      const float* host_wave, int64_t B, int64_t L, int64_t slot0, int64_t cap,
                                      int64_t max_update, const int32_t* host_mix_slot, const float* host_alpha,
                                      const int32_t* host_crop, float* host_view1, float* host_view2,
                                      float** dev_view1, float** dev_view2, void* stream) {
  using namespace assl;
  if (!p || !vc || !ws || !host_wave || !host_mix_slot || !host_alpha || !host_crop) return ASSL_ERR_INVALID_ARG;
  if (B < 1 || cap < B || slot0 < 0 || slot0 >= cap || L <= p->cfg.n_fft / 2) return ASSL_ERR_INVALID_ARG;
  const int64_t T = 1 + L / p->cfg.hop;
  if (vc->n_freq != p->cfg.n_mels || vc->n_time != T) return ASSL_ERR_INVALID_ARG;
  const WsLayout w = layout(p, B, L, cap);
  if (static_cast<size_t>(ws_bytes) < w.total) return ASSL_ERR_INVALID_ARG;
  cudaStream_t st = as_stream(stream);
  char* base = static_cast<char*>(ws);
  double* state = reinterpret_cast<double*>(base + w.state);
  float* stats = reinterpret_cast<float*>(base + w.stats_ring);
  float* lms = reinterpret_cast<float*>(base + w.lms_ring);
  float* wave = reinterpret_cast<float*>(base + w.wave);
  double* tsums = reinterpret_cast<double*>(base + w.tile_sums);
  int32_t* clip_slot = reinterpret_cast<int32_t*>(base + w.clip_slot);
  int32_t* mix_slot = reinterpret_cast<int32_t*>(base + w.mix_slot);
  float* alpha = reinterpret_cast<float*>(base + w.alpha);
  int32_t* crop = reinterpret_cast<int32_t*>(base + w.crop);
  float* v1 = reinterpret_cast<float*>(base + w.view1);
  float* v2 = reinterpret_cast<float*>(base + w.view2);
  const size_t plane = static_cast<size_t>(p->cfg.n_mels) * T * sizeof(float);

  bool ok = true;
  ok &= cudaMemcpyAsync(wave, host_wave, static_cast<size_t>(B) * L * sizeof(float), cudaMemcpyHostToDevice, st) == cudaSuccess;
  ok &= cudaMemcpyAsync(mix_slot, host_mix_slot, static_cast<size_t>(B) * 2 * sizeof(int32_t), cudaMemcpyHostToDevice, st) == cudaSuccess;
  ok &= cudaMemcpyAsync(alpha, host_alpha, static_cast<size_t>(B) * 2 * sizeof(float), cudaMemcpyHostToDevice, st) == cudaSuccess;
  ok &= cudaMemcpyAsync(crop, host_crop, static_cast<size_t>(B) * 8 * sizeof(int32_t), cudaMemcpyHostToDevice, st) == cudaSuccess;
  if (!ok) return ASSL_ERR_CUDA;
  iota_slots_kernel<<<static_cast<unsigned>((B + 255) / 256), 256, 0, st>>>(clip_slot, B, slot0, cap);

  int rc = assl_logmel_f32(p, wave, nullptr, B, L, L, lms, slot0, cap, tsums, 0, stream);
  if (rc != ASSL_OK) return rc;
  rc = assl_running_stats(state, tsums, B, assl_plan_num_tiles(p, L), static_cast<int64_t>(p->cfg.n_mels) * T,
                          max_update, 0, B, stats, slot0, cap, stream);
  if (rc != ASSL_OK) return rc;
  rc = assl_two_views_f32(vc, lms, stats, cap, clip_slot, mix_slot, alpha, crop, B, v1, v2, stream);
  if (rc != ASSL_OK) return rc;
  if (host_view1) ok &= cudaMemcpyAsync(host_view1, v1, static_cast<size_t>(B) * plane, cudaMemcpyDeviceToHost, st) == cudaSuccess;
  if (host_view2) ok &= cudaMemcpyAsync(host_view2, v2, static_cast<size_t>(B) * plane, cudaMemcpyDeviceToHost, st) == cudaSuccess;
  if (dev_view1) *dev_view1 = v1;
  if (dev_view2) *dev_view2 = v2;
  return ok ? launch_status() : ASSL_ERR_CUDA;
}
