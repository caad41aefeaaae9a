// extract_window for a batch of raw, variable-length waveforms, with the L2 gain of the window in the
// same pass.
//
// Replaces extract_window (/root/reference/src/utils/utils.py:166-182): a clip shorter than the unit
// length is zero-padded, left half = (unit - len) // 2; a longer one is cropped at ``start``, the
// offset the reference draws with random.randint(0, len - unit) (passed in, drawn on the host in the
// reference's order).  And the reduction of F.normalize(p=2, dim=-1)
// (/root/reference/src/dataset/upstream_dataset.py:62-63): gain = 1 / max(||window||_2, 1e-12), which
// assl_logmel_f32 applies on load (wave_gain).
//
// One CTA per clip: the raw clips sit back to back in one buffer at arbitrary (4-byte aligned) offsets,
// so loads are scalar and coalesced; the fixed-length rows written here are 16-byte aligned, which is
// what lets the log-mel kernel stage them with bulk (TMA) copies.  HBM bound: 4 bytes read + 4 written
// per window sample.
#include "common.cuh"

namespace assl {

constexpr int kWinThreads = 512;

__global__ void __launch_bounds__(kWinThreads) extract_windows_kernel(
    const float* __restrict__ raw, const long long* __restrict__ offsets, const long long* __restrict__ lengths,
    const long long* __restrict__ starts, long long unit, float* __restrict__ out, long long ld,
    float* __restrict__ gain) {
  __shared__ float s_red[kWinThreads / 32];
  const long long b = blockIdx.x;
  const long long len = lengths[b];
  const float* src = raw + offsets[b];
  // window sample p comes from clip sample p + shift (zero outside [0, len))
  long long shift;
  if (len < unit) {
    shift = -((unit - len) / 2);                     // F.pad(wav, (half, adj - half)); nothing to crop
  } else {
    long long st = starts ? starts[b] : 0;
    st = st < 0 ? 0 : (st > len - unit ? len - unit : st);
    shift = st;
  }
  float* dst = out + b * ld;
  float acc = 0.f;
  for (long long p = threadIdx.x; p < unit; p += kWinThreads) {
    const long long i = p + shift;
    const float v = (i >= 0 && i < len) ? __ldg(src + i) : 0.f;
    dst[p] = v;
    acc = fmaf(v, v, acc);
  }
  if (gain) {                                        // fixed-order block reduction: same bits every run
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = 0.f;
      for (int w = 0; w < kWinThreads / 32; ++w) t += s_red[w];
      gain[b] = 1.0f / fmaxf(sqrtf(t), 1e-12f);
    }
  }
}

}  // namespace assl

extern "C" int assl_extract_windows_f32(const float* raw, const int64_t* offsets, const int64_t* lengths,
                                        const int64_t* starts, int64_t B, int64_t unit_length, float* out,
                                        int64_t ld, float* gain, void* stream) {
  if (B < 0 || unit_length < 1 || ld < unit_length) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  if (!raw || !offsets || !lengths || !out) return ASSL_ERR_INVALID_ARG;
  if (B > 0x7fffffffll) return ASSL_ERR_UNSUPPORTED;
  static_assert(sizeof(long long) == sizeof(int64_t), "int64_t layout");
  assl::extract_windows_kernel<<<static_cast<unsigned>(B), assl::kWinThreads, 0, assl::as_stream(stream)>>>(
      raw, reinterpret_cast<const long long*>(offsets), reinterpret_cast<const long long*>(lengths),
      reinterpret_cast<const long long*>(starts), unit_length, out, ld, gain);
  return assl::launch_status();
}
