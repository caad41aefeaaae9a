// Shared declarations for the assl CUDA sources (sm_100a only, no host fallback).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <atomic>
#include <vector>

#include "../../include/assl.h"

namespace assl {

constexpr int kNfft = 1024;             // the FFT kernel is specialised for 32 x 32 points
constexpr int kBins = kNfft / 2 + 1;    // 513
constexpr int kMaxMels = 128;
constexpr int kIvSlots = 8;             // most filterbank-interval parts a lane of the log-mel kernel can own
constexpr int kPartBins = 8;            // bins per part (two per 128-bit load)
constexpr int kScratchEntries = 208;    // 16-byte interval totals that fit behind a warp's power spectra

constexpr float kPowerEps = 2.220446049250313e-16f;   // np.finfo(float).eps   (utils.py:28)
constexpr float kLogEps = 1.1920928955078125e-07f;    // torch.finfo().eps     (utils.py:48)
constexpr float kF32Max = 3.4028234663852886e+38f;

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

inline int launch_status() {
  return cudaPeekAtLastError() == cudaSuccess ? ASSL_OK : ASSL_ERR_CUDA;
}

// ---- programmatic dependent launch (PDL).  Every kernel of the path is launched with the programmatic-stream-
// serialization attribute, calls pdl_wait() before its first access to memory an earlier kernel may have written
// and pdl_launch_dependents() early, so the next kernel's CTAs are scheduled onto SMs as they drain and have their
// prologue behind them when the previous grid completes (pdl_wait returns only after the whole previous grid has
// finished and its writes are visible: the data dependencies are exactly those of plain stream order).  Launched
// without the attribute, both calls are no-ops.  ASSL_PDL=0 turns the attribute off (A/B measurements).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("ASSL_PDL"); return !(e && e[0] == '0'); }();
  return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__host__ __device__ inline long long ring_slot(long long slot0, long long b, long long cap) {
  return cap > 0 ? (slot0 + b) % cap : slot0 + b;
}

}  // namespace assl

// Plan: immutable after creation (the log-mel kernel's work-claiming counters are scratch that every launch hands back
// zeroed; the slot a launch uses is drawn from an atomic, so concurrent callers are safe).  Device tables are owned by
// the library.
struct assl_plan {
  assl_cfg cfg;
  int device;
  int sm_count;
  int n_slots;                  // interval parts a lane owns in the log-mel kernel (rows of d_lane_iv in use)
  float2* d_twiddle;            // [32][32]     tw[k2][n1] = exp(-2 pi i n1 k2 / 1024)
  int4* d_lane_iv;              // [slots][32]  interval part owned by (slot, lane) in the log-mel kernel:
                                //              {first bin (even) | valid-bin mask << 10 | scratch entry << 18 |
                                //               starts-the-total << 26 | owns-a-part << 27, alpha, beta, 0}
  int4* d_band;                 // [n_mels]     {byte offset of U(m) / of D(m+1) in a warp's interval totals, enorm,
                                //              2.22e-16 * sum of the band's weights (folded "+ eps" of utils.py:28)}
  std::vector<float> h_fb;      // dense [n_mels][513] copy for assl_plan_filterbank
  std::vector<int4> h_lane_iv;  // host copy of d_lane_iv
  std::vector<int> h_iv_pos;    // [n_mels+1] scratch entry of every interval's total
  // work-claiming counters of the log-mel kernel: {next ticket, warps done} per slot, zero between launches (the
  // kernel resets its own); the slot a launch uses is drawn here, so concurrent callers never share one
  static constexpr unsigned kTicketSlots = 64;
  unsigned* d_tickets = nullptr;
  mutable std::atomic<unsigned> ticket_seq{0};
};
