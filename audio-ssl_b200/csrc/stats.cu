// Running pre-normalisation statistics, batch post-normalisation, fixed-statistics normalisation.
//
// running_stats_kernel reproduces the reference's sequential RunningMean / RunningVariance updates
// (/root/reference/src/augmentations/augmentations.py:222-228, 249-256, 275-281) for a whole batch
// at once.  The reference divides by the number of samples seen BEFORE the current one, so after
// k >= 1 updates  mu_k = mean(m_1..m_k)  (sample 0 is overwritten) and the variance tracker does
// the same with v_j = mean((x_j - mu_j)^2) = q_j - 2 mu_j m_j + mu_j^2.  Both are prefix means, so the
// recurrence becomes two block-wide scans over per-clip (m_j, q_j) = (sum x, sum x^2) / elems in fp64.
#include <cooperative_groups.h>

#include <cstring>
#include <new>

#include "common.cuh"

namespace assl {

constexpr int kScanThreads = 1024;
constexpr int kStatCtas = 8;               // CTAs of the running-statistics cluster: 128 clips of a 1024-clip chunk each

// inclusive scan over the block; returns this thread's prefix, *total gets the block total
__device__ inline double block_scan_incl(double v, double* s_warp, double* total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const double n = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += n;
  }
  __syncthreads();                 // s_warp may still be read from the previous scan
  if (lane == 31) s_warp[warp] = v;
  __syncthreads();
  if (warp == 0) {
    double w = s_warp[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double n = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += n;
    }
    s_warp[lane] = w;
  }
  __syncthreads();
  if (warp > 0) v += s_warp[warp - 1];
  *total = s_warp[31];
  return v;
}

// (sum x, sum x^2) of clip ``idx`` over its ``n_parts`` entries, computed by 8 consecutive lanes: lane g adds parts
// g, g+8, ... in order, then a fixed xor tree.  Every path to a clip total goes through here (the single-GPU scan
// and the per-rank reduction before the all-gather), so the statistics do not depend on the number of GPUs.
// All 32 lanes of the warp must call it; clips past n_clips give zeros.
__device__ inline void clip_total8(const double* sums, long long idx, long long n_clips, long long n_parts, int g,
                                   double* sx, double* sxx) {
  double a = 0.0, b = 0.0;
  if (idx < n_clips) {
    const double2* sp = reinterpret_cast<const double2*>(sums) + idx * n_parts;
    const double2 zero = make_double2(0.0, 0.0);
    for (long long p = g; p < n_parts; p += 64) {  // eight loads in flight (a 1-s clip's 51 parts: one round trip);
      double2 v[8];                                //  added in the order g, g+8, g+16, ...
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = p + 8 * u < n_parts ? sp[p + 8 * u] : zero;
#pragma unroll
      for (int u = 0; u < 8; ++u) { a += v[u].x; b += v[u].y; }
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  *sx = a; *sxx = b;
}

// Peer-memory exchange of the per-clip sums (assl_running_stats_exchange): where every rank's buffer of the
// current parity is, as mapped into this process.  A record is 32 bytes per clip, written and read as 8-byte
// words {32 bits of data, 32-bit call number} -- each word carries its own flag, so the receiver needs no fence
// and no separate flag round trip: it polls the record until all four words show this call's number (the way
// NCCL's LL protocol moves small messages).
struct PeerArgs {
  int world = 0;                 // 0 = single process, no exchange
  int rank = 0;
  long long b_local = 0;         // clips of this rank (the same on every rank)
  long long n_parts_local = 0;
  const double* pair_sums = nullptr;                          // this rank's [b_local, n_parts_local, 2]
  unsigned seq = 0;                                           // this call's number (never 0)
  uint4* peer_rec[ASSL_EXCHANGE_MAX_WORLD] = {};              // rank j's records [world * b_local][2]
};

__device__ __forceinline__ void st_volatile_v4(uint4* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

// state: [0] n_put  [1] sum of m_j (j>=1)  [2] sum of v_j (j>=1)  [3] mean  [4] std  [5..6] reserved  [7] exchange error
//
// Launched as ONE thread-block cluster of kStatCtas CTAs.  The clips are taken kStatCtas * 1024 at a time, one
// chunk of 1024 per CTA:
//   1. per-clip (m, q) = (sum x, sum x^2) / elems of every chunk land in the owning CTA's shared memory -- for
//      per-pair sums all CTAs reduce 128 clips of each chunk and hand the result over through DSMEM (that read is
//      the bulk of the bytes and a single SM cannot keep enough of it in flight); gathered per-clip sums
//      (n_parts = 1) are simply loaded by the owner;
//   2. every CTA scans its chunk, the chunk totals are exchanged through DSMEM and added in chunk order, so the
//      result is bit-identical to a sequential walk over the chunks while a global batch of 8 GPUs x 1024 clips
//      costs the time of one chunk.
__global__ void __cluster_dims__(kStatCtas, 1, 1) __launch_bounds__(kScanThreads) running_stats_kernel(
    double* state, const double* sums, long long n_clips, long long n_parts, double inv_elems,
    long long max_update, long long first, long long count, float* mean_std, long long slot0, long long cap,
    const PeerArgs px) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  __shared__ double s_warp[32];
  __shared__ double s_m[kScanThreads], s_q[kScanThreads];
  __shared__ double s_tot[2][kStatCtas];           // chunk totals of m / of v, a full copy in every CTA
  __shared__ float s_frozen[2];
  const int tid = threadIdx.x;
  pdl_launch_dependents();                           // (the two-view kernel's CTAs wait for this grid to complete)
  pdl_wait();                                        // the log-mel kernel's per-pair sums are complete and visible
  const long long n0 = static_cast<long long>(state[0]);
  double carry_m = state[1], carry_v = state[2];
  if (tid == 0) { s_frozen[0] = static_cast<float>(state[3]); s_frozen[1] = static_cast<float>(state[4]); }

  __shared__ int s_late;
  if (tid == 0) s_late = 0;
  if (px.world > 0) {
    // ---- exchange, sending side: totals of this rank's clips -> the records [rank * b_local, ...) of EVERY rank's
    //      buffer (stores over NVLink for the peers; the local copy goes the same way)
    for (long long base = 0; base < px.b_local; base += kScanThreads) {
      const long long idx = base + rank * (kScanThreads / 8) + (tid >> 3);
      double sx, sxx;
      clip_total8(px.pair_sums, idx, px.b_local, px.n_parts_local, tid & 7, &sx, &sxx);
      if ((tid & 7) == 0 && idx < px.b_local) {
        const uint4 ra = make_uint4(__double2loint(sx), px.seq, __double2hiint(sx), px.seq);
        const uint4 rb = make_uint4(__double2loint(sxx), px.seq, __double2hiint(sxx), px.seq);
        for (int j = 0; j < px.world; ++j) {
          uint4* rec = px.peer_rec[j] + 2 * (px.rank * px.b_local + idx);
          st_volatile_v4(rec, ra);
          st_volatile_v4(rec + 1, rb);
        }
      }
    }
  }

  for (long long sbase = 0; sbase < n_clips; sbase += static_cast<long long>(kStatCtas) * kScanThreads) {
    const long long left = n_clips - sbase;
    const int n_chunks = static_cast<int>(left >= static_cast<long long>(kStatCtas) * kScanThreads
                                              ? kStatCtas : (left + kScanThreads - 1) / kScanThreads);
    if (n_parts == 1) {                              // per-clip sums: the owner loads its chunk
      const long long idx = sbase + rank * kScanThreads + tid;
      if (idx < n_clips && px.world > 0) {           // receiving side: poll this clip's record until it is this call's
        const uint4* rec = px.peer_rec[px.rank] + 2 * idx;
        const long long t0 = clock64();
        uint4 ra, rb;
        while (true) {
          ra = ld_volatile_v4(rec); rb = ld_volatile_v4(rec + 1);
          if (ra.y == px.seq && ra.w == px.seq && rb.y == px.seq && rb.w == px.seq) break;
          if (clock64() - t0 > 4000000000ll) { s_late = 1; break; }   // ~2 s: a peer is gone; do not hang the GPU
          __nanosleep(64);
        }
        s_m[tid] = __hiloint2double(ra.z, ra.x) * inv_elems; s_q[tid] = __hiloint2double(rb.z, rb.x) * inv_elems;
      } else if (idx < n_clips) {
        const double2 v = reinterpret_cast<const double2*>(sums)[idx];
        s_m[tid] = v.x * inv_elems; s_q[tid] = v.y * inv_elems;
      }
    } else {                                         // per-pair sums: 8 lanes per clip, 128 clips of each chunk per CTA
      for (int c = 0; c < n_chunks; ++c) {
        double* dm = cluster.map_shared_rank(s_m, c);
        double* dq = cluster.map_shared_rank(s_q, c);
        const int cc = rank * (kScanThreads / 8) + (tid >> 3);
        double sx, sxx;
        clip_total8(sums, sbase + static_cast<long long>(c) * kScanThreads + cc, n_clips, n_parts, tid & 7, &sx, &sxx);
        if ((tid & 7) == 0) { dm[cc] = sx * inv_elems; dq[cc] = sxx * inv_elems; }
      }
    }
    cluster.sync();

    const long long idx = sbase + rank * kScanThreads + tid;
    const bool valid = idx < n_clips;
    const long long k = n0 + idx;                     // put index of this clip
    const double m = valid ? s_m[tid] : 0.0, q = valid ? s_q[tid] : 0.0;
    const bool active = valid && k < max_update;
    const bool counted = active && k >= 1;
    double tot;
    if (n_chunks == 1 && rank > 0) break;            // a single chunk: CTA 0 finishes alone (cluster-uniform condition)
    double pm = block_scan_incl(counted ? m : 0.0, s_warp, &tot);
    if (n_chunks == 1) {
      pm += carry_m; carry_m += tot;
    } else {
      if (tid < kStatCtas) cluster.map_shared_rank(&s_tot[0][0], tid)[rank] = tot;
      cluster.sync();
      double cin = carry_m;
      for (unsigned c = 0; c < rank; ++c) cin += s_tot[0][c];
      for (int c = 0; c < kStatCtas; ++c) carry_m += s_tot[0][c];
      pm += cin;
    }
    const double mu = (k == 0) ? m : pm / static_cast<double>(k);
    double v = q - 2.0 * mu * m + mu * mu;
    v = v > 0.0 ? v : 0.0;
    double pv = block_scan_incl(counted ? v : 0.0, s_warp, &tot);
    if (n_chunks == 1) {
      pv += carry_v; carry_v += tot;
    } else {
      if (tid < kStatCtas) cluster.map_shared_rank(&s_tot[1][0], tid)[rank] = tot;
      cluster.sync();
      double cin = carry_v;
      for (unsigned c = 0; c < rank; ++c) cin += s_tot[1][c];
      for (int c = 0; c < kStatCtas; ++c) carry_v += s_tot[1][c];
      pv += cin;
    }
    const double s2 = (k == 0) ? v : pv / static_cast<double>(k);
    float my_mean = 0.f, my_std = 1.f;
    if (active) {
      my_mean = static_cast<float>(mu);
      my_std = fminf(fmaxf(sqrtf(static_cast<float>(s2)), kLogEps), kF32Max);   // clamp(std, eps, max) :281
      if (k == max_update - 1 || idx == n_clips - 1) { state[3] = my_mean; state[4] = my_std; }
      if (k == max_update - 1) {                       // the statistics freeze here: every later clip uses these
        for (int c = 0; c < (n_chunks == 1 ? 1 : kStatCtas); ++c) {
          float* fz = cluster.map_shared_rank(s_frozen, c);
          fz[0] = my_mean; fz[1] = my_std;
        }
      }
    }
    if (n_chunks == 1) __syncthreads();              // (the last pass: nothing is overwritten after it)
    else cluster.sync();                             // frozen values are out; (m, q) and the totals may be overwritten
    if (valid && !active) { my_mean = s_frozen[0]; my_std = s_frozen[1]; }     // statistics frozen :276
    if (valid && idx >= first && idx < first + count) {
      const long long s = ring_slot(slot0, idx - first, cap);
      mean_std[2 * s] = my_mean;
      mean_std[2 * s + 1] = my_std;
    }
  }
  if (tid == 0 && s_late) state[7] = 1.0;           // (CTAs that left early saw their records arrive)
  if (rank == 0 && tid == 0) {
    long long n1 = n0;
    if (n0 < max_update) n1 = (n0 + n_clips < max_update) ? n0 + n_clips : max_update;
    state[0] = static_cast<double>(n1);
    state[1] = carry_m;
    state[2] = carry_v;
  }
}

__global__ void __launch_bounds__(256) clip_sums_kernel(const double* tile_sums, long long B, long long n_parts, double* clip_sums) {
  const long long b = blockIdx.x * 32ll + (threadIdx.x >> 3);          // 8 lanes per clip
  double sx, sxx;
  clip_total8(tile_sums, b, B, n_parts, threadIdx.x & 7, &sx, &sxx);
  if (b < B && (threadIdx.x & 7) == 0) { clip_sums[2 * b] = sx; clip_sums[2 * b + 1] = sxx; }
}

// per-clip (sum x, sum x^2) in fp64, one CTA per clip
__global__ void __launch_bounds__(256) clip_moments_kernel(const float* x, long long elems, double* sums) {
  __shared__ double s_red[16];
  const float* src = x + blockIdx.x * elems;
  double s1 = 0.0, s2 = 0.0;
  for (long long i = threadIdx.x; i < elems; i += 256) { const double v = src[i]; s1 += v; s2 += v * v; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_down_sync(0xffffffffu, s1, o); s2 += __shfl_down_sync(0xffffffffu, s2, o); }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_red[2 * warp] = s1; s_red[2 * warp + 1] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double r1 = 0.0, r2 = 0.0;
    for (int w = 0; w < 8; ++w) { r1 += s_red[2 * w]; r2 += s_red[2 * w + 1]; }
    sums[2 * blockIdx.x] = r1;
    sums[2 * blockIdx.x + 1] = r2;
  }
}

__global__ void __launch_bounds__(256) clip_normalize_kernel(const float* x, float* y, long long elems, const float* mean_std,
                                                              long long slot0, long long cap) {
  const long long s = ring_slot(slot0, blockIdx.y, cap);
  const float mean = mean_std[2 * s], std = mean_std[2 * s + 1];
  const float* src = x + blockIdx.y * elems;
  float* dst = y + blockIdx.y * elems;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < elems; i += gridDim.x * 256ll) dst[i] = (src[i] - mean) / std;
}

// ---- NormalizeBatch (augmentations.py:321-324): sums in fp64, one atomic triple per CTA
__global__ void __launch_bounds__(256) batch_sums_kernel(const float* x, long long n, double* sums) {
  __shared__ double s_red[16];
  double s1 = 0.0, s2 = 0.0;
  const long long stride = static_cast<long long>(gridDim.x) * 256;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += stride) {
    const double v = x[i];
    s1 += v; s2 += v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s1 += __shfl_down_sync(0xffffffffu, s1, o); s2 += __shfl_down_sync(0xffffffffu, s2, o); }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { s_red[2 * warp] = s1; s_red[2 * warp + 1] = s2; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double r1 = 0.0, r2 = 0.0;
    for (int w = 0; w < 8; ++w) { r1 += s_red[2 * w]; r2 += s_red[2 * w + 1]; }
    atomicAdd(&sums[1], r1);
    atomicAdd(&sums[2], r2);
    if (blockIdx.x == 0) atomicAdd(&sums[0], static_cast<double>(n));
  }
}

__global__ void __launch_bounds__(256) normalize_apply_kernel(const float* x, float* y, long long n, const double* sums) {
  const double cnt = sums[0];
  const double mean = sums[1] / cnt;
  double var = (sums[2] - cnt * mean * mean) / (cnt - 1.0);     // torch.std default: unbiased
  var = var > 0.0 ? var : 0.0;
  const float fmean = static_cast<float>(mean);
  const float fstd = fminf(fmaxf(static_cast<float>(sqrt(var)), kLogEps), kF32Max);
  const long long stride = static_cast<long long>(gridDim.x) * 256;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += stride) y[i] = (x[i] - fmean) / fstd;
}

__global__ void __launch_bounds__(256) affine_norm_kernel(const float* x, float* y, long long n, float mean, float std) {
  const long long stride = static_cast<long long>(gridDim.x) * 256;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < n; i += stride) y[i] = (x[i] - mean) / std;
}

inline unsigned ew_grid(long long n) {
  long long g = (n + 255) / 256;
  const long long cap = 148ll * 16;
  return static_cast<unsigned>(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace assl

extern "C" int assl_running_stats(double* state, const double* sums, int64_t n_clips, int64_t n_parts,
                                  int64_t elems, int64_t max_update, int64_t first, int64_t count,
                                  float* mean_std, int64_t slot0, int64_t cap, void* stream) {
  using namespace assl;
  if (!state || !sums || !mean_std || n_clips < 0 || n_parts < 1 || elems < 1 || max_update < 1) return ASSL_ERR_INVALID_ARG;
  if (first < 0 || count < 0 || first + count > n_clips || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (cap > 0 && (count > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;   // (sum, sum^2) pairs are read as one 16-byte word
  if (n_clips == 0) return ASSL_OK;
  if (launch_pdl(running_stats_kernel, dim3(kStatCtas), dim3(kScanThreads), 0, as_stream(stream), state, sums,
                 static_cast<long long>(n_clips), static_cast<long long>(n_parts), 1.0 / static_cast<double>(elems),
                 static_cast<long long>(max_update), static_cast<long long>(first), static_cast<long long>(count), mean_std,
                 static_cast<long long>(slot0), static_cast<long long>(cap), PeerArgs()) != cudaSuccess)
    return ASSL_ERR_CUDA;
  return launch_status();
}

extern "C" int assl_clip_moments_f32(const float* x, int64_t B, int64_t elems, double* sums, void* stream) {
  if (!x || !sums || B < 0 || elems < 1) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::clip_moments_kernel<<<static_cast<unsigned>(B), 256, 0, assl::as_stream(stream)>>>(x, elems, sums);
  return assl::launch_status();
}

extern "C" int assl_clip_normalize_f32(const float* x, float* y, int64_t B, int64_t elems, const float* mean_std,
                                       int64_t slot0, int64_t cap, void* stream) {
  if (!x || !y || !mean_std || B < 0 || B > 65535 || elems < 1 || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  long long gx = (elems + 255) / 256;
  gx = gx > 64 ? 64 : gx;
  assl::clip_normalize_kernel<<<dim3(static_cast<unsigned>(gx), static_cast<unsigned>(B)), 256, 0, assl::as_stream(stream)>>>(
      x, y, elems, mean_std, slot0, cap);
  return assl::launch_status();
}

extern "C" int assl_clip_sums(const double* tile_sums, int64_t B, int64_t n_parts, double* clip_sums, void* stream) {
  if (!tile_sums || !clip_sums || B < 0 || n_parts < 1) return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(tile_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  assl::clip_sums_kernel<<<static_cast<unsigned>((B + 31) / 32), 256, 0, assl::as_stream(stream)>>>(tile_sums, B, n_parts, clip_sums);
  return assl::launch_status();
}

extern "C" int assl_batch_sums(const float* x, int64_t n, double* sums, void* stream) {
  if (!x || !sums || n < 0) return ASSL_ERR_INVALID_ARG;
  if (n == 0) return ASSL_OK;
  assl::batch_sums_kernel<<<assl::ew_grid(n), 256, 0, assl::as_stream(stream)>>>(x, n, sums);
  return assl::launch_status();
}

extern "C" int assl_normalize_apply_f32(const float* x, float* y, int64_t n, const double* sums, void* stream) {
  if (!x || !y || !sums || n < 0) return ASSL_ERR_INVALID_ARG;
  if (n == 0) return ASSL_OK;
  assl::normalize_apply_kernel<<<assl::ew_grid(n), 256, 0, assl::as_stream(stream)>>>(x, y, n, sums);
  return assl::launch_status();
}

extern "C" int assl_affine_norm_f32(const float* x, float* y, int64_t n, float mean, float std, void* stream) {
  if (!x || !y || n < 0) return ASSL_ERR_INVALID_ARG;
  if (n == 0) return ASSL_OK;
  assl::affine_norm_kernel<<<assl::ew_grid(n), 256, 0, assl::as_stream(stream)>>>(x, y, n, mean, std);
  return assl::launch_status();
}

// ---- peer-memory exchange object -------------------------------------------------------------------------
struct assl_exchange {
  int rank = 0, world = 0;
  long long max_clips = 0;
  unsigned char* local = nullptr;                         // this rank's buffer (cudaMalloc: IPC-exportable)
  unsigned char* peer[ASSL_EXCHANGE_MAX_WORLD] = {};       // every rank's buffer as mapped here (peer[rank] = local)
  size_t rec_bytes = 0, bytes = 0;                        // per parity: [world * max_clips] records of 32 bytes
  unsigned long long seq = 0;
  bool connected = false;
};

extern "C" int assl_exchange_create(assl_exchange** out, int32_t rank, int32_t world, int64_t max_clips, void* handle_out) {
  if (!out || !handle_out || world < 2 || world > ASSL_EXCHANGE_MAX_WORLD || rank < 0 || rank >= world || max_clips < 1)
    return ASSL_ERR_INVALID_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == ASSL_EXCHANGE_HANDLE_BYTES, "IPC handle size");
  *out = nullptr;
  assl_exchange* x = new (std::nothrow) assl_exchange();
  if (!x) return ASSL_ERR_CUDA;
  x->rank = rank; x->world = world; x->max_clips = max_clips;
  x->rec_bytes = (32 * static_cast<size_t>(world) * max_clips + 255) & ~static_cast<size_t>(255);
  x->bytes = 2 * x->rec_bytes;
  cudaIpcMemHandle_t h;
  if (cudaMalloc(reinterpret_cast<void**>(&x->local), x->bytes) != cudaSuccess || cudaMemset(x->local, 0, x->bytes) != cudaSuccess ||
      cudaIpcGetMemHandle(&h, x->local) != cudaSuccess) {
    cudaFree(x->local); delete x; cudaGetLastError();
    return ASSL_ERR_CUDA;
  }
  std::memcpy(handle_out, &h, sizeof(h));
  x->peer[rank] = x->local;
  *out = x;
  return ASSL_OK;
}

extern "C" int assl_exchange_connect(assl_exchange* x, const void* all_handles) {
  if (!x || !all_handles) return ASSL_ERR_INVALID_ARG;
  if (x->connected) return ASSL_OK;
  for (int r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const unsigned char*>(all_handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return ASSL_ERR_CUDA; }
    x->peer[r] = static_cast<unsigned char*>(p);
  }
  x->connected = true;
  return ASSL_OK;
}

extern "C" int assl_exchange_destroy(assl_exchange* x) {
  if (!x) return ASSL_OK;
  for (int r = 0; r < x->world; ++r)
    if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
  cudaFree(x->local);
  delete x;
  return ASSL_OK;
}

extern "C" int assl_running_stats_exchange(assl_exchange* x, double* state, const double* pair_sums, int64_t B,
                                           int64_t n_parts, int64_t elems, int64_t max_update, float* mean_std,
                                           int64_t slot0, int64_t cap, int32_t write_all, void* stream) {
  using namespace assl;
  if (!x || !x->connected || !state || !pair_sums || !mean_std || B < 1 || B > x->max_clips || n_parts < 1 || elems < 1 ||
      max_update < 1 || slot0 < 0 || cap < 0)
    return ASSL_ERR_INVALID_ARG;
  const long long n_all = static_cast<long long>(x->world) * B;
  if (cap > 0 && ((write_all ? n_all : B) > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(pair_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  PeerArgs px;
  px.world = x->world; px.rank = x->rank; px.b_local = B; px.n_parts_local = n_parts; px.pair_sums = pair_sums;
  ++x->seq;
  px.seq = static_cast<unsigned>(x->seq & 0x7fffffffu) | 0x80000000u;       // never 0 (the buffer starts zeroed)
  const size_t parity = static_cast<size_t>(x->seq & 1);
  for (int r = 0; r < x->world; ++r) px.peer_rec[r] = reinterpret_cast<uint4*>(x->peer[r] + parity * x->rec_bytes);
  if (launch_pdl(running_stats_kernel, dim3(kStatCtas), dim3(kScanThreads), 0, as_stream(stream), state, pair_sums, n_all,
                 1ll, 1.0 / static_cast<double>(elems), static_cast<long long>(max_update),
                 write_all ? 0ll : static_cast<long long>(x->rank) * B, write_all ? n_all : static_cast<long long>(B), mean_std,
                 static_cast<long long>(slot0), static_cast<long long>(cap), px) != cudaSuccess)
    return ASSL_ERR_CUDA;
  return launch_status();
}

// ---- IPC-exportable buffers (the log-mel ring of the global-bank mode)
extern "C" int assl_peer_alloc(void** ptr, int64_t bytes, void* handle_out) {
  if (!ptr || !handle_out || bytes < 1) return ASSL_ERR_INVALID_ARG;
  *ptr = nullptr;
  void* p = nullptr;
  cudaIpcMemHandle_t h;
  if (cudaMalloc(&p, static_cast<size_t>(bytes)) != cudaSuccess || cudaIpcGetMemHandle(&h, p) != cudaSuccess) {
    cudaFree(p); cudaGetLastError();
    return ASSL_ERR_CUDA;
  }
  std::memcpy(handle_out, &h, sizeof(h));
  *ptr = p;
  return ASSL_OK;
}
extern "C" int assl_peer_free(void* ptr) { return cudaFree(ptr) == cudaSuccess ? ASSL_OK : ASSL_ERR_CUDA; }
extern "C" int assl_peer_open(const void* handle, void** ptr) {
  if (!handle || !ptr) return ASSL_ERR_INVALID_ARG;
  cudaIpcMemHandle_t h;
  std::memcpy(&h, handle, sizeof(h));
  if (cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); return ASSL_ERR_CUDA; }
  return ASSL_OK;
}
extern "C" int assl_peer_close(void* ptr) { return cudaIpcCloseMemHandle(ptr) == cudaSuccess ? ASSL_OK : ASSL_ERR_CUDA; }
