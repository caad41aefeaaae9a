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

// ---- exchange of the per-clip sums between the GPUs of a node (assl_exchange_send / assl_running_stats_exchange_scan /
// assl_running_stats_exchange).  Every rank owns a buffer of ``world`` SEGMENTS (one per sender) per parity; a segment
// is ``seg`` records of 32 bytes: record 0 = header {clips the sender brought this call}, records 1.. = per-clip
// (sum x, sum x^2).  A record is written and read as 8-byte words {32 bits of data, 32-bit call number} -- each word
// carries its own flag, so the receiver needs no fence and no separate flag round trip: it polls a record until all four
// words show this call's number (the way NCCL's LL protocol moves small messages).  Ranks may bring different numbers
// of clips (an uneven block split, a ragged last batch): the headers say how many.
struct ExArgs {
  int world = 0;                 // 0 = single process, no exchange
  int rank = 0;
  long long seg = 0;             // records per segment (max_clips + 1)
  unsigned seq = 0;              // this call's number (never 0); SCAN side when send_after is set
  unsigned send_seq = 0;         // SEND side: the number of the call whose records go out (= seq unless send_after)
  int send_after = 0;            // scan call ``seq`` (sent a step ago) FIRST, then wait for the grid before this one
                                 // and send call ``send_seq``: assl_exchange_scan_send
  int n_dst = 0;                 // SEND: the buffers this rank's segment is stored into -- every rank's (peer transport:
  uint4* dst[ASSL_EXCHANGE_MAX_WORLD] = {};   //  stores over NVLink) or only its own (staged: the caller all-gathers)
  const uint4* src = nullptr;    // SCAN: this rank's buffer of the call's parity
  int* host_err = nullptr;       // mapped host word, set when a peer's records did not arrive in time
  unsigned long long timeout_ns = 0;
  long long b_local = 0;         // SCAN: clips this rank brought to the call (known on the host since the send)
  int write_all = 0;             // SCAN: (mean, std) of every clip of the global batch, not only this rank's
  int deferred = 0;              // SCAN: does not wait for the grid before it at its start, but at its end
};

__device__ __forceinline__ void st_volatile_v4(uint4* p, uint4 v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// one record = (a, b) as two flagged 16-byte words
__device__ __forceinline__ void put_record(const ExArgs& x, long long index, double a, double b) {
  const uint4 ra = make_uint4(__double2loint(a), x.send_seq, __double2hiint(a), x.send_seq);
  const uint4 rb = make_uint4(__double2loint(b), x.send_seq, __double2hiint(b), x.send_seq);
  for (int j = 0; j < x.n_dst; ++j) {
    uint4* rec = x.dst[j] + 2 * (x.rank * x.seg + index);
    st_volatile_v4(rec, ra);
    st_volatile_v4(rec + 1, rb);
  }
}
// wait until record ``index`` of sender ``from`` shows this call's number; false after timeout_ns
__device__ __forceinline__ bool await_record(const ExArgs& x, int from, long long index, double* a, double* b) {
  const uint4* rec = x.src + 2 * (from * x.seg + index);
  unsigned long long t0 = 0;
  for (unsigned spins = 0;; ++spins) {
    const uint4 ra = ld_volatile_v4(rec), rb = ld_volatile_v4(rec + 1);
    if (ra.y == x.seq && ra.w == x.seq && rb.y == x.seq && rb.w == x.seq) {
      *a = __hiloint2double(ra.z, ra.x); *b = __hiloint2double(rb.z, rb.x);
      return true;
    }
    if (spins == 64) t0 = global_ns();               // (the common case never reads the timer)
    if (spins > 64) {
      if (global_ns() - t0 > x.timeout_ns) return false;
      __nanosleep(256);
    }
  }
}

// Sending side on its own (assl_exchange_send): totals of this rank's clips -> its segment in every destination buffer.
// Issued right behind the log-mel kernel; nobody waits here.
__global__ void __launch_bounds__(256) stats_send_kernel(const ExArgs x, const double* pair_sums, long long b, long long n_parts) {
  pdl_launch_dependents();
  pdl_wait();                                        // the log-mel kernel's per-pair sums are complete and visible
  const long long idx = blockIdx.x * 32ll + (threadIdx.x >> 3);          // 8 lanes per clip
  double sx, sxx;
  clip_total8(pair_sums, idx, b, n_parts, threadIdx.x & 7, &sx, &sxx);
  if ((threadIdx.x & 7) == 0 && idx < b) put_record(x, 1 + idx, sx, sxx);
  if (blockIdx.x == 0 && threadIdx.x == 0) put_record(x, 0, static_cast<double>(b), 0.0);
}

// Sending side inside the scan cluster: totals of this rank's clips -> its segment in EVERY destination buffer (stores over
// NVLink for the peers; the local copy goes the same way).  All kStatCtas CTAs take part, 128 clips each per pass.
__device__ __forceinline__ void send_own(const ExArgs& x, const double* send_sums, long long send_b, long long send_parts,
                                         unsigned rank, int tid) {
  for (long long base = 0; base < send_b; base += kScanThreads) {
    const long long idx = base + rank * (kScanThreads / 8) + (tid >> 3);
    double sx, sxx;
    clip_total8(send_sums, idx, send_b, send_parts, tid & 7, &sx, &sxx);
    if ((tid & 7) == 0 && idx < send_b) put_record(x, 1 + idx, sx, sxx);
  }
  if (rank == 0 && tid == 0) put_record(x, 0, static_cast<double>(send_b), 0.0);
}

// state: [0] n_put  [1] sum of m_j (j>=1)  [2] sum of v_j (j>=1)  [3] mean  [4] std  [5..6] reserved  [7] exchange error
//
// Launched as ONE thread-block cluster of kStatCtas CTAs.  The clips are taken kStatCtas * 1024 at a time, one
// chunk of 1024 per CTA:
//   1. per-clip (m, q) = (sum x, sum x^2) / elems of every chunk land in the owning CTA's shared memory -- for
//      per-pair sums all CTAs reduce 128 clips of each chunk and hand the result over through DSMEM (that read is
//      the bulk of the bytes and a single SM cannot keep enough of it in flight); per-clip sums (n_parts = 1, or the
//      records of an exchange) are simply loaded by the owner;
//   2. every CTA scans its chunk, the chunk totals are exchanged through DSMEM and added in chunk order, so the
//      result is bit-identical to a sequential walk over the chunks while a global batch of 8 GPUs x 1024 clips
//      costs the time of one chunk.
// With an exchange (x.world > 0) the clips are the records in this rank's buffer, sender after sender (= global clip
// order of a contiguous block split); ``send_sums`` != null: this rank's own records are produced and sent here first
// (the fused form, assl_running_stats_exchange).  n_clips / first / count then come from the headers.
__global__ void __cluster_dims__(kStatCtas, 1, 1) __launch_bounds__(kScanThreads) running_stats_kernel(
    double* state, const double* sums, long long n_clips, long long n_parts, double inv_elems,
    long long max_update, long long first, long long count, float* mean_std, long long slot0, long long cap,
    const ExArgs x, const double* send_sums, long long send_b, long long send_parts) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  __shared__ double s_warp[32];
  __shared__ double s_m[kScanThreads], s_q[kScanThreads];
  __shared__ double s_tot[2][kStatCtas];           // chunk totals of m / of v, a full copy in every CTA
  __shared__ float s_frozen[2];
  __shared__ long long s_off[ASSL_EXCHANGE_MAX_WORLD + 1];   // first global clip of every sender
  __shared__ int s_late, s_late_all[kStatCtas];
  const int tid = threadIdx.x;
  pdl_launch_dependents();                           // (the two-view kernel's CTAs wait for this grid to complete)
  if (!x.deferred) pdl_wait();                       // the log-mel kernel's per-pair sums are complete and visible
  // (read from L2, never the SM's L1: a deferred launch starts without griddepcontrol.wait, i.e. without the acquire that
  //  would drop a line of ``state`` an earlier scan on this SM left there)
  const long long n0 = static_cast<long long>(__ldcg(state));
  double carry_m = __ldcg(state + 1), carry_v = __ldcg(state + 2);
  if (tid == 0) { s_frozen[0] = static_cast<float>(__ldcg(state + 3)); s_frozen[1] = static_cast<float>(__ldcg(state + 4)); s_late = 0; }

  if (x.world > 0) {
    if (send_sums && !x.send_after) send_own(x, send_sums, send_b, send_parts, rank, tid);   // fused form: send, then scan
    // ---- receiving side.  Headers first: how many clips every sender brought
    __syncthreads();
    if (tid < x.world) {
      double c = 0.0, unused;
      const bool ok = await_record(x, tid, 0, &c, &unused);
      const long long cnt = static_cast<long long>(c);
      if (!ok || cnt < 0 || cnt >= x.seg) { s_late = 1; s_off[tid + 1] = 0; } else s_off[tid + 1] = cnt;
    }
    __syncthreads();
    if (tid == 0) {
      s_off[0] = 0;
      for (int j = 0; j < x.world; ++j) s_off[j + 1] += s_off[j];
    }
    __syncthreads();
    n_clips = s_off[x.world];
    first = x.write_all ? 0 : s_off[x.rank];
    count = x.write_all ? n_clips : s_off[x.rank + 1] - s_off[x.rank];
    // then every record of the call has to be here before any of them is consumed: a peer that does not show up must
    // leave the running statistics untouched, not half updated
    if (!s_late) {
      for (long long idx = rank * kScanThreads + tid; idx < n_clips; idx += static_cast<long long>(kStatCtas) * kScanThreads) {
        int j = 0;
        while (j + 1 < x.world && idx >= s_off[j + 1]) ++j;
        double u0, u1;
        if (!await_record(x, j, 1 + idx - s_off[j], &u0, &u1)) { s_late = 1; break; }
      }
    }
    __syncthreads();
    if (tid < kStatCtas) cluster.map_shared_rank(&s_late_all[0], tid)[rank] = s_late;
    cluster.sync();
    int late = 0;
    for (int c = 0; c < kStatCtas; ++c) late |= s_late_all[c];
    if (late) {
      // a peer's records did not arrive within the timeout (or a header is garbage): the statistics stay as they were,
      // this rank's clips get the last published (mean, std), the error flag goes up (assl_exchange_error)
      const long long mine = x.write_all ? 0 : x.b_local;
      const bool any = n0 > 0;
      for (long long i = rank * kScanThreads + tid; i < mine; i += static_cast<long long>(kStatCtas) * kScanThreads) {
        const long long s = ring_slot(slot0, i, cap);
        mean_std[2 * s] = any ? s_frozen[0] : 0.f;
        mean_std[2 * s + 1] = any ? s_frozen[1] : 1.f;
      }
      if (rank == 0 && tid == 0) {
        state[7] = 1.0;
        if (x.host_err) { *reinterpret_cast<volatile int*>(x.host_err) = 1; __threadfence_system(); }
      }
      if (x.send_after) {                             // the peers still get this rank's records of the next call
        pdl_wait();
        send_own(x, send_sums, send_b, send_parts, rank, tid);
      } else if (x.deferred && rank == 0) pdl_wait();
      return;
    }
    n_parts = 1;
  }

  for (long long sbase = 0; sbase < n_clips; sbase += static_cast<long long>(kStatCtas) * kScanThreads) {
    const long long left = n_clips - sbase;
    const int n_chunks = static_cast<int>(left >= static_cast<long long>(kStatCtas) * kScanThreads
                                              ? kStatCtas : (left + kScanThreads - 1) / kScanThreads);
    if (n_parts == 1) {                              // per-clip sums: the owner loads its chunk
      const long long idx = sbase + rank * kScanThreads + tid;
      if (idx < n_clips && x.world > 0) {            // this call's records (all of them are known to have arrived)
        int j = 0;
        while (j + 1 < x.world && idx >= s_off[j + 1]) ++j;
        double sx, sxx;
        await_record(x, j, 1 + idx - s_off[j], &sx, &sxx);
        s_m[tid] = sx * inv_elems; s_q[tid] = sxx * inv_elems;
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
  if (rank == 0 && tid == 0) {
    long long n1 = n0;
    if (n0 < max_update) n1 = (n0 + n_clips < max_update) ? n0 + n_clips : max_update;
    state[0] = static_cast<double>(n1);
    state[1] = carry_m;
    state[2] = carry_v;
  }
  // deferred launch: this grid ran beside the one launched before it; whoever waits for THIS grid to complete (the
  // next kernel's pdl_wait) must also get that one's completion (one CTA holds the grid open, the others free their SMs)
  if (x.send_after) {
    // scan-then-send (assl_exchange_scan_send): the records of the NEXT call go out from here, once the grid before this
    // one (the log-mel kernel whose per-pair sums they add up) has completed.  The cluster barrier first: a peer that
    // has all of this rank's new records (header included) may overwrite the parity this call's records sat in, so
    // every CTA must be done reading them before the first new record leaves.
    cluster.sync();
    pdl_wait();
    send_own(x, send_sums, send_b, send_parts, rank, tid);
  } else if (x.deferred && rank == 0) pdl_wait();
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
                 static_cast<long long>(slot0), static_cast<long long>(cap), ExArgs(), static_cast<const double*>(nullptr),
                 0ll, 0ll) != cudaSuccess)
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

// ---- exchange object -------------------------------------------------------------------------------------
struct assl_exchange {
  int rank = 0, world = 0;
  long long max_clips = 0, seg = 0;                       // seg = max_clips + 1 records per sender (record 0 = header)
  unsigned char* local = nullptr;                         // this rank's buffer (cudaMalloc: IPC-exportable)
  unsigned char* peer[ASSL_EXCHANGE_MAX_WORLD] = {};       // every rank's buffer as mapped here (peer[rank] = local)
  size_t rec_bytes = 0, bytes = 0;                        // per parity: world segments of seg records of 32 bytes
  unsigned long long sent = 0, scanned = 0;               // calls sent / scanned so far (sent - scanned is 0 or 1)
  long long sent_b[2] = {0, 0};                           // clips this rank brought to the call of either parity
  bool connected = false, staged = false;
  int* h_err = nullptr;                                   // mapped host word the scan kernel sets on a timeout
  int* d_err = nullptr;
  unsigned long long timeout_ns = 30000000000ull;         // 30 s: of the order of a collective watchdog, not of a step
};

namespace {
unsigned seq_word(unsigned long long call) { return static_cast<unsigned>(call & 0x7fffffffu) | 0x80000000u; }   // never 0

assl::ExArgs send_args(const assl_exchange* x, unsigned long long call) {
  assl::ExArgs a;
  a.world = x->world; a.rank = x->rank; a.seg = x->seg; a.seq = a.send_seq = seq_word(call);
  const size_t off = static_cast<size_t>(call & 1) * x->rec_bytes;
  if (x->staged) { a.n_dst = 1; a.dst[0] = reinterpret_cast<uint4*>(x->local + off); }
  else { a.n_dst = x->world; for (int r = 0; r < x->world; ++r) a.dst[r] = reinterpret_cast<uint4*>(x->peer[r] + off); }
  return a;
}
void scan_args(const assl_exchange* x, unsigned long long call, assl::ExArgs* a) {
  a->world = x->world; a->rank = x->rank; a->seg = x->seg; a->seq = seq_word(call);
  a->src = reinterpret_cast<const uint4*>(x->local + static_cast<size_t>(call & 1) * x->rec_bytes);
  a->host_err = x->d_err; a->timeout_ns = x->timeout_ns; a->b_local = x->sent_b[call & 1];
}
}  // namespace

extern "C" int assl_exchange_create(assl_exchange** out, int32_t rank, int32_t world, int64_t max_clips, void* handle_out) {
  if (!out || !handle_out || world < 2 || world > ASSL_EXCHANGE_MAX_WORLD || rank < 0 || rank >= world || max_clips < 1)
    return ASSL_ERR_INVALID_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == ASSL_EXCHANGE_HANDLE_BYTES, "IPC handle size");
  *out = nullptr;
  assl_exchange* x = new (std::nothrow) assl_exchange();
  if (!x) return ASSL_ERR_CUDA;
  x->rank = rank; x->world = world; x->max_clips = max_clips; x->seg = max_clips + 1;
  x->rec_bytes = (32 * static_cast<size_t>(world) * x->seg + 255) & ~static_cast<size_t>(255);
  x->bytes = 2 * x->rec_bytes;
  cudaIpcMemHandle_t h;
  bool ok = cudaMalloc(reinterpret_cast<void**>(&x->local), x->bytes) == cudaSuccess && cudaMemset(x->local, 0, x->bytes) == cudaSuccess &&
            cudaIpcGetMemHandle(&h, x->local) == cudaSuccess;
  ok = ok && cudaHostAlloc(reinterpret_cast<void**>(&x->h_err), sizeof(int), cudaHostAllocMapped) == cudaSuccess;
  if (ok) { *x->h_err = 0; ok = cudaHostGetDevicePointer(reinterpret_cast<void**>(&x->d_err), x->h_err, 0) == cudaSuccess; }
  if (!ok) {
    cudaFree(x->local); if (x->h_err) cudaFreeHost(x->h_err); delete x; cudaGetLastError();
    return ASSL_ERR_CUDA;
  }
  std::memcpy(handle_out, &h, sizeof(h));
  x->peer[rank] = x->local;
  *out = x;
  return ASSL_OK;
}

extern "C" int assl_exchange_connect(assl_exchange* x, const void* all_handles) {
  if (!x || !all_handles || x->staged) return ASSL_ERR_INVALID_ARG;
  if (x->connected) return ASSL_OK;
  for (int r = 0; r < x->world; ++r) {
    if (r == x->rank) continue;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, static_cast<const unsigned char*>(all_handles) + static_cast<size_t>(r) * sizeof(h), sizeof(h));
    void* p = nullptr;
    if (cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
      cudaGetLastError();
      for (int q = 0; q < r; ++q)
        if (q != x->rank && x->peer[q]) { cudaIpcCloseMemHandle(x->peer[q]); x->peer[q] = nullptr; }
      return ASSL_ERR_CUDA;
    }
    x->peer[r] = static_cast<unsigned char*>(p);
  }
  x->connected = true;
  return ASSL_OK;
}

extern "C" int assl_exchange_use_staging(assl_exchange* x) {
  if (!x || (x->connected && !x->staged)) return ASSL_ERR_INVALID_ARG;
  x->staged = true; x->connected = true;
  return ASSL_OK;
}

extern "C" int assl_exchange_staging(const assl_exchange* x, int64_t back, void** buffer, int64_t* segment_bytes) {
  if (!x || !buffer || !segment_bytes || back < 0 || back > 1 || x->sent == 0) return ASSL_ERR_INVALID_ARG;
  const unsigned long long call = x->sent - static_cast<unsigned long long>(back);
  if (call == 0) return ASSL_ERR_INVALID_ARG;
  *buffer = x->local + static_cast<size_t>(call & 1) * x->rec_bytes;
  *segment_bytes = static_cast<int64_t>(32 * x->seg);
  return ASSL_OK;
}

extern "C" int assl_exchange_set_timeout(assl_exchange* x, double seconds) {
  if (!x || !(seconds > 0.0)) return ASSL_ERR_INVALID_ARG;
  x->timeout_ns = static_cast<unsigned long long>(seconds * 1e9);
  return ASSL_OK;
}

extern "C" int assl_exchange_error(const assl_exchange* x) {
  if (!x) return ASSL_ERR_INVALID_ARG;
  return *reinterpret_cast<volatile int*>(x->h_err) ? ASSL_ERR_PEER_TIMEOUT : ASSL_OK;
}

extern "C" int assl_exchange_destroy(assl_exchange* x) {
  if (!x) return ASSL_OK;
  if (!x->staged)
    for (int r = 0; r < x->world; ++r)
      if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
  cudaFree(x->local);
  if (x->h_err) cudaFreeHost(x->h_err);
  delete x;
  return ASSL_OK;
}

extern "C" int assl_exchange_send(assl_exchange* x, const double* pair_sums, int64_t B, int64_t n_parts, void* stream) {
  using namespace assl;
  if (!x || !x->connected || (!pair_sums && B > 0) || B < 0 || B > x->max_clips || n_parts < 1) return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(pair_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  if (x->sent != x->scanned) return ASSL_ERR_INVALID_ARG;              // the previous call has to be scanned first (two parities)
  const unsigned long long call = x->sent + 1;
  const ExArgs a = send_args(x, call);
  const unsigned grid = static_cast<unsigned>(B > 0 ? (B + 31) / 32 : 1);
  if (launch_pdl(stats_send_kernel, dim3(grid), dim3(256), 0, as_stream(stream), a, pair_sums, static_cast<long long>(B),
                 static_cast<long long>(n_parts)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  x->sent = call; x->sent_b[call & 1] = B;
  return launch_status();
}

static int exchange_scan(assl_exchange* x, double* state, int64_t elems, int64_t max_update, float* mean_std, int64_t slot0,
                         int64_t cap, int32_t write_all, int32_t deferred, const double* send_sums, int64_t send_b,
                         int64_t send_parts, void* stream, bool send_after = false) {
  using namespace assl;
  const unsigned long long call = x->scanned + 1;
  // (scan_args overwrites seq with the scanned call's number; send_seq keeps the number of the call that goes out)
  ExArgs a = (send_sums || send_after) ? send_args(x, send_after ? x->sent + 1 : call) : ExArgs();
  scan_args(x, call, &a);
  a.write_all = write_all ? 1 : 0; a.deferred = deferred ? 1 : 0; a.send_after = send_after ? 1 : 0;
  if (launch_pdl(running_stats_kernel, dim3(kStatCtas), dim3(kScanThreads), 0, as_stream(stream), state,
                 static_cast<const double*>(nullptr), 0ll, 1ll, 1.0 / static_cast<double>(elems), static_cast<long long>(max_update),
                 0ll, 0ll, mean_std, static_cast<long long>(slot0), static_cast<long long>(cap), a, send_sums,
                 static_cast<long long>(send_b), static_cast<long long>(send_parts)) != cudaSuccess)
    return ASSL_ERR_CUDA;
  x->scanned = call;
  return launch_status();
}

extern "C" int assl_running_stats_exchange_scan(assl_exchange* x, double* state, int64_t elems, int64_t max_update,
                                                float* mean_std, int64_t slot0, int64_t cap, int32_t write_all,
                                                int32_t deferred, void* stream) {
  if (!x || !x->connected || !state || !mean_std || elems < 1 || max_update < 1 || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (x->scanned >= x->sent) return ASSL_ERR_INVALID_ARG;              // nothing sent that has not been scanned
  const long long b = x->sent_b[(x->scanned + 1) & 1];
  if (cap > 0 && ((write_all ? static_cast<long long>(x->world) * b : b) > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  return exchange_scan(x, state, elems, max_update, mean_std, slot0, cap, write_all, deferred, nullptr, 0, 0, stream);
}

extern "C" int assl_running_stats_exchange(assl_exchange* x, double* state, const double* pair_sums, int64_t B,
                                           int64_t n_parts, int64_t elems, int64_t max_update, float* mean_std,
                                           int64_t slot0, int64_t cap, int32_t write_all, void* stream) {
  if (!x || !x->connected || x->staged || !state || !pair_sums || !mean_std || B < 1 || B > x->max_clips || n_parts < 1 ||
      elems < 1 || max_update < 1 || slot0 < 0 || cap < 0)
    return ASSL_ERR_INVALID_ARG;
  if (x->sent != x->scanned) return ASSL_ERR_INVALID_ARG;              // a split send is still waiting for its scan
  if (cap > 0 && ((write_all ? static_cast<long long>(x->world) * B : B) > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(pair_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  x->sent_b[(x->sent + 1) & 1] = B;
  const int rc = exchange_scan(x, state, elems, max_update, mean_std, slot0, cap, write_all, 0, pair_sums, B, n_parts, stream);
  if (rc == ASSL_OK) x->sent += 1;
  return rc;
}

extern "C" int assl_exchange_scan_send(assl_exchange* x, double* state, int64_t elems, int64_t max_update, float* mean_std,
                                       int64_t slot0, int64_t cap, int32_t write_all, const double* pair_sums, int64_t B,
                                       int64_t n_parts, void* stream) {
  if (!x || !x->connected || !state || !mean_std || elems < 1 || max_update < 1 || slot0 < 0 || cap < 0 ||
      (!pair_sums && B > 0) || B < 0 || B > x->max_clips || n_parts < 1)
    return ASSL_ERR_INVALID_ARG;
  if (reinterpret_cast<uintptr_t>(pair_sums) % 16 != 0) return ASSL_ERR_INVALID_ARG;
  if (x->scanned + 1 != x->sent) return ASSL_ERR_INVALID_ARG;          // exactly one call sent and not yet scanned
  const long long b = x->sent_b[x->sent & 1];
  if (cap > 0 && ((write_all ? static_cast<long long>(x->world) * b : b) > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  const int rc = exchange_scan(x, state, elems, max_update, mean_std, slot0, cap, write_all, 1, pair_sums, B, n_parts, stream,
                               true);
  if (rc == ASSL_OK) { x->sent += 1; x->sent_b[x->sent & 1] = B; }
  return rc;
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

// ---- running-norm statistics with NO allocation anywhere (assl_running_norm_stats_static_f32) ----------------------
// No counterpart in the reference: added for callers that cannot allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  assl_clip_moments_f32 followed by
// assl_running_stats needs a caller-owned [B, 2] fp64 array between the two kernels; here that scratch is a file-scope
// device array of ASSL_STATIC_MAX_CLIPS entries (64 KB, held by every GPU that loads the library until exit).
// Every call on a device shares it and it keeps the last call's contents: the first kernel overwrites all B entries
// the second one reads, so nothing has to start at zero, but two calls that overlap on the GPU (different streams)
// race on it.
namespace assl {
static __device__ __align__(256) double g_static_clip_sums[2 * ASSL_STATIC_MAX_CLIPS];
}

namespace {
constexpr int kStaticMaxDevices = 64;
// In host code ``g_static_clip_sums`` is only a stand-in for the symbol, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, asked once per device with that device current.
int static_clip_sums(double** out) {
  static std::atomic<double*> addr[kStaticMaxDevices];
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return ASSL_ERR_NO_DEVICE; }
  if (dev < 0 || dev >= kStaticMaxDevices) return ASSL_ERR_UNSUPPORTED;
  double* p = addr[dev].load(std::memory_order_acquire);
  if (!p) {
    void* sym = nullptr;
    if (cudaGetSymbolAddress(&sym, assl::g_static_clip_sums) != cudaSuccess || !sym) { cudaGetLastError(); return ASSL_ERR_CUDA; }
    p = static_cast<double*>(sym);
    addr[dev].store(p, std::memory_order_release);       // (two threads racing here store the same address)
  }
  *out = p;
  return ASSL_OK;
}
}  // namespace

extern "C" int assl_running_norm_stats_static_f32(double* state, const float* x, int64_t B, int64_t elems,
                                                  int64_t max_update, float* mean_std, int64_t slot0, int64_t cap,
                                                  void* stream) {
  using namespace assl;
  if (!state || !x || !mean_std || B < 0 || elems < 1 || max_update < 1 || slot0 < 0 || cap < 0) return ASSL_ERR_INVALID_ARG;
  if (B > ASSL_STATIC_MAX_CLIPS) return ASSL_ERR_TOO_LARGE;       // the static scratch holds ASSL_STATIC_MAX_CLIPS clips
  if (cap > 0 && (B > cap || slot0 >= cap)) return ASSL_ERR_INVALID_ARG;
  if (B == 0) return ASSL_OK;
  double* sums = nullptr;
  const int rc = static_clip_sums(&sums);
  if (rc != ASSL_OK) return rc;
  cudaStream_t st = as_stream(stream);
  clip_moments_kernel<<<static_cast<unsigned>(B), 256, 0, st>>>(x, elems, sums);
  if (launch_status() != ASSL_OK) return ASSL_ERR_CUDA;
  if (launch_pdl(running_stats_kernel, dim3(kStatCtas), dim3(kScanThreads), 0, st, state, static_cast<const double*>(sums),
                 static_cast<long long>(B), 1ll, 1.0 / static_cast<double>(elems), static_cast<long long>(max_update), 0ll,
                 static_cast<long long>(B), mean_std, static_cast<long long>(slot0), static_cast<long long>(cap), ExArgs(),
                 static_cast<const double*>(nullptr), 0ll, 0ll) != cudaSuccess)
    return ASSL_ERR_CUDA;
  return launch_status();
}
