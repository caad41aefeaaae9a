// Cluster features for Kmix (/root/reference/src/augmentations/augmentations.py:119-189), SURVEY §8f-3.
//
// Kmix.get_index (:142-165) describes a spectrogram by its per-band mean over time and assigns it to the
// nearest of K centroids twice, with two different scalings:
//   * a bank entry uses the L2-NORMALISED band means against the L2-normalised centroids
//     (memory_avg / memory_avg.norm(), :148-151, 153) -> mem_cluster;
//   * the incoming clip uses its RAW band means against the same normalised centroids (:154) -> point_cluster.
// Every clip plays both roles (it is mixed first, then pushed into the bank), so one pass computes both ids.
// The choice among the bank entries of the farthest non-empty cluster and the random draw stay on the host
// (a few integers per clip); the mix itself is the MixupBYOLA arithmetic of the two-view kernel.
//
// One CTA per clip: band means with lanes along time (coalesced), then one thread per centroid.
#include "common.cuh"

namespace assl {

constexpr int kKmixThreads = 128;

__global__ void __launch_bounds__(kKmixThreads) clip_clusters_kernel(
    const float* __restrict__ x, int F, int T, const float* __restrict__ centroids_unit, int K,
    int* __restrict__ mem_cluster, int* __restrict__ point_cluster) {
  __shared__ float s_avg[kMaxMels];
  __shared__ float s_red[kKmixThreads / 32];
  __shared__ float s_best[2][kKmixThreads / 32];
  __shared__ int s_arg[2][kKmixThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* clip = x + static_cast<long long>(blockIdx.x) * F * T;

  for (int f = warp; f < F; f += kKmixThreads / 32) {            // x.squeeze(0).T.mean(dim=0): mean over time
    float acc = 0.f;
    for (int t = lane; t < T; t += 32) acc += clip[f * T + t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) s_avg[f] = acc / static_cast<float>(T);
  }
  __syncthreads();
  float sq = 0.f;
  for (int f = tid; f < F; f += kKmixThreads) sq = fmaf(s_avg[f], s_avg[f], sq);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  if (lane == 0) s_red[warp] = sq;
  __syncthreads();
  float nrm = 0.f;
  for (int w = 0; w < kKmixThreads / 32; ++w) nrm += s_red[w];
  const float inv = 1.0f / sqrtf(nrm);

  // squared distances to every centroid; ties go to the lowest index like torch.argmin
  float best_m = kF32Max, best_p = kF32Max;
  int arg_m = 0x7fffffff, arg_p = 0x7fffffff;
  for (int k = tid; k < K; k += kKmixThreads) {
    const float* c = centroids_unit + static_cast<long long>(k) * F;
    float dm = 0.f, dp = 0.f;
    for (int f = 0; f < F; ++f) {
      const float a = s_avg[f], cv = c[f];
      const float em = fmaf(a, inv, -cv), ep = a - cv;
      dm = fmaf(em, em, dm);
      dp = fmaf(ep, ep, dp);
    }
    if (dm < best_m) { best_m = dm; arg_m = k; }
    if (dp < best_p) { best_p = dp; arg_p = k; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, best_m, o), op = __shfl_xor_sync(0xffffffffu, best_p, o);
    const int im = __shfl_xor_sync(0xffffffffu, arg_m, o), ip = __shfl_xor_sync(0xffffffffu, arg_p, o);
    if (om < best_m || (om == best_m && im < arg_m)) { best_m = om; arg_m = im; }
    if (op < best_p || (op == best_p && ip < arg_p)) { best_p = op; arg_p = ip; }
  }
  if (lane == 0) { s_best[0][warp] = best_m; s_arg[0][warp] = arg_m; s_best[1][warp] = best_p; s_arg[1][warp] = arg_p; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < kKmixThreads / 32; ++w) {
      if (s_best[0][w] < best_m || (s_best[0][w] == best_m && s_arg[0][w] < arg_m)) { best_m = s_best[0][w]; arg_m = s_arg[0][w]; }
      if (s_best[1][w] < best_p || (s_best[1][w] == best_p && s_arg[1][w] < arg_p)) { best_p = s_best[1][w]; arg_p = s_arg[1][w]; }
    }
    mem_cluster[blockIdx.x] = arg_m;
    point_cluster[blockIdx.x] = arg_p;
  }
}

}  // namespace assl

extern "C" int assl_clip_clusters_f32(const float* x, int64_t B, int32_t n_freq, int32_t n_time,
                                      const float* centroids_unit, int32_t K, int32_t* mem_cluster,
                                      int32_t* point_cluster, void* stream) {
  if (B < 0 || n_freq < 1 || n_freq > assl::kMaxMels || n_time < 1 || K < 1) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  if (!x || !centroids_unit || !mem_cluster || !point_cluster) return ASSL_ERR_INVALID_ARG;
  if (B > 0x7fffffffll) return ASSL_ERR_UNSUPPORTED;
  assl::clip_clusters_kernel<<<static_cast<unsigned>(B), assl::kKmixThreads, 0, assl::as_stream(stream)>>>(
      x, n_freq, n_time, centroids_unit, K, mem_cluster, point_cluster);
  return assl::launch_status();
}
