// Plan construction: the host-side restatement of librosa.filters.mel as the reference calls it
// (/root/reference/src/utils/utils.py:24) plus the window / twiddle tables of the FFT kernel.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
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
    case ASSL_ERR_PEER_TIMEOUT: return "a peer's statistics records did not arrive in time";
    case ASSL_ERR_TOO_LARGE: return "more clips than the static scratch holds (ASSL_STATIC_MAX_CLIPS)";
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
  // sparse "interval" form.  Interval j (0..n_mels) = bins with centre(j-1) < f_k <= centre(j), centre(j) =
  // mel_f[j+1]; a bin there has weight only in band j-1 (falling edge) and band j (rising edge), and both are
  // linear in the bin index:  with t = (f_k - mel_f[j]) / (mel_f[j+1] - mel_f[j])  they are  enorm(j-1) (1 - t)
  // and  enorm(j) t.  The kernel therefore needs no weight table: a lane adds up  S0 = sum P_k  and
  // S1 = sum r P_k  over the interval's bins it owns (r = bin index inside its part), forms  U = sum t P =
  // alpha S1 + beta S0  and  D = S0 - U, and band m = enorm(m) (U(m) + D(m+1)).
  const int n_iv = n_mels + 1;
  std::vector<int> bin_iv(n_bins);
  for (int k = 0; k < n_bins; ++k) {
    const double fk = linspace_at(0.0, sr / 2, n_bins, k);
    int j = 0;
    while (j < n_mels && fk > mel_f[j + 1]) ++j;
    bin_iv[k] = j;
    for (int m = 0; m < n_mels; ++m)     // the triangles must overlap only with their neighbours
      if (m != j && m != j - 1 && p->h_fb[static_cast<size_t>(m) * n_bins + k] != 0.f) { delete p; return ASSL_ERR_UNSUPPORTED; }
  }
  // an interval's run: trimmed to its non-zero bins, widened to an even first bin (128-bit power loads) and to
  // whole parts of kPartBins bins; the bins of a part that are not the interval's are masked out in the kernel
  std::vector<int> k0(n_iv, 0), cnt(n_iv, 0);
  auto bin_in = [&](int j, int k) {
    if (k >= n_bins || bin_iv[k] != j) return false;
    const float wd = j >= 1 ? p->h_fb[static_cast<size_t>(j - 1) * n_bins + k] : 0.f;
    const float wu = j < n_mels ? p->h_fb[static_cast<size_t>(j) * n_bins + k] : 0.f;
    return wd != 0.f || wu != 0.f;
  };
  for (int j = 0; j < n_iv; ++j) {
    int lo = n_bins, hi = -1;
    for (int k = 0; k < n_bins; ++k)
      if (bin_in(j, k)) { lo = lo < k ? lo : k; hi = k; }
    if (hi < 0) continue;
    const int first = lo & ~1;
    int n = hi - first + 1;
    n = (n + assl::kPartBins - 1) / assl::kPartBins * assl::kPartBins;
    k0[j] = first;
    cnt[j] = n;
  }
  // Where a run starts is free within its padding: any even first bin that still covers the interval's bins with the
  // same number of parts will do (the mask says which bins count).  The choice decides the 16-byte bank group
  // (first bin / 2) mod 8 of the part's 128-bit loads, and a quarter-warp's eight loads are conflict-free when its
  // eight parts sit in eight different groups -- so the starts are chosen to spread the parts evenly over the groups
  // (at most one part per group and quarter-warp): first the single moves that help most, then augmenting paths of
  // single-part intervals from an over-full group to one with room.
  {
    int total = 0;
    for (int j = 0; j < n_iv; ++j) total += cnt[j] > 0 ? cnt[j] / assl::kPartBins : 1;
    const int room = 4 * ((total + 31) / 32);                    // quarter-warps over all slots = most parts per group
    std::vector<int> lo_of(n_iv, 0), hi_of(n_iv, -1);
    for (int j = 0; j < n_iv; ++j)
      if (cnt[j] > 0) {
        lo_of[j] = n_bins; 
        for (int k = 0; k < n_bins; ++k)
          if (bin_in(j, k)) { lo_of[j] = lo_of[j] < k ? lo_of[j] : k; hi_of[j] = k; }
      }
    auto n_parts_of = [&](int j) { return cnt[j] > 0 ? cnt[j] / assl::kPartBins : 1; };
    auto group_of = [&](int start, int c) { return (start / 2 + 4 * c) & 7; };      // part c of a run: 8 bins = 4 groups on
    auto smallest_start = [&](int j) {                          // even, >= 0, the run still reaches the interval's last bin
      int s = hi_of[j] - cnt[j] + 1;
      s = s < 0 ? 0 : s;
      return (s + 1) & ~1;
    };
    int load[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int j = 0; j < n_iv; ++j)
      for (int c = 0; c < n_parts_of(j); ++c) ++load[group_of(k0[j], c)];
    auto excess = [&](const int* l) { int e = 0; for (int g = 0; g < 8; ++g) e += l[g] > room ? l[g] - room : 0; return e; };
    auto spread = [&](const int* l) { int e = 0; for (int g = 0; g < 8; ++g) e += l[g] * l[g]; return e; };
    auto moved = [&](int j, int s, int* l) {                    // loads with interval j's run started at s instead
      for (int c = 0; c < n_parts_of(j); ++c) { --l[group_of(k0[j], c)]; ++l[group_of(s, c)]; }
    };
    for (int pass = 0; pass < 256 && excess(load) > 0; ++pass) {              // best single move, while one helps
      int best_j = -1, best_s = 0, best_e = excess(load), best_q = spread(load);
      for (int j = 0; j < n_iv; ++j) {
        if (cnt[j] == 0) continue;
        for (int st = smallest_start(j); st <= (lo_of[j] & ~1); st += 2) {
          if (st == k0[j]) continue;
          int l[8];
          for (int g = 0; g < 8; ++g) l[g] = load[g];
          moved(j, st, l);
          const int e = excess(l), q = spread(l);
          if (e < best_e || (e == best_e && q < best_q)) { best_j = j; best_s = st; best_e = e; best_q = q; }
        }
      }
      if (best_j < 0) break;
      moved(best_j, best_s, load);
      k0[best_j] = best_s;
    }
    for (int pass = 0; pass < 256 && excess(load) > 0; ++pass) {              // augmenting paths, one part at a time
      int src = 0;
      while (load[src] <= room) ++src;
      int from[8], via_j[8], via_s[8], queue[8], qh = 0, qt = 0;
      for (int g = 0; g < 8; ++g) from[g] = -2;
      from[src] = -1; queue[qt++] = src;
      int dst = -1;
      while (qh < qt && dst < 0) {
        const int a = queue[qh++];
        for (int j = 0; j < n_iv && dst < 0; ++j) {
          if (cnt[j] != assl::kPartBins || group_of(k0[j], 0) != a) continue;  // single-part intervals sitting in group a
          bool used = false;
          for (int g = 0; g < 8; ++g) used = used || (from[g] >= 0 && via_j[g] == j);
          if (used) continue;
          for (int st = smallest_start(j); st <= (lo_of[j] & ~1); st += 2) {
            const int b = group_of(st, 0);
            if (from[b] != -2) continue;
            from[b] = a; via_j[b] = j; via_s[b] = st; queue[qt++] = b;
            if (load[b] < room) { dst = b; break; }
          }
        }
      }
      if (dst < 0) break;                                       // (stays correct, a quarter-warp then pays a second wavefront)
      for (int g = dst; from[g] >= 0; g = from[g]) { k0[via_j[g]] = via_s[g]; --load[from[g]]; ++load[g]; }
    }
  }
  // Lanes of the log-mel kernel own pieces of intervals for the whole launch.  Every interval is cut into
  // parts of exactly kPartBins bins (its weight run is zero-padded to whole parts), so all lanes run the
  // same straight-line code; a lane owns one part per slot.  Two parts of one interval never share a slot:
  // their sums are added to the interval's total slot after slot, in a fixed order.  A 128-bit shared load
  // is served a quarter-warp at a time, conflict-free when its eight 16-byte accesses fall in eight different
  // bank groups: parts are dealt to the quarter-warps round-robin in order of bank group, so a quarter sees
  // a group twice only when more parts share it than there are quarters.
  {
    struct Part { int k0, iv; };
    std::vector<Part> parts;
    for (int j = 0; j < n_iv; ++j)
      for (int c = 0; c < (cnt[j] > 0 ? cnt[j] : 1); c += assl::kPartBins)
        parts.push_back({cnt[j] > 0 ? k0[j] + c : 0, j});
    const int n_slots = (static_cast<int>(parts.size()) + 31) / 32;
    if (n_slots > assl::kIvSlots) { delete p; return ASSL_ERR_UNSUPPORTED; }
    auto group = [](const Part& x) { return (x.k0 / 2) & 7; };
    {   // padding parts own nothing (every bin masked): they read wherever the emptiest bank group is
      int load[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (const Part& x : parts) ++load[group(x)];
      while (static_cast<int>(parts.size()) < 32 * n_slots) {
        int g = 0;
        for (int c = 1; c < 8; ++c) if (load[c] < load[g]) g = c;
        parts.push_back({2 * g, -1});
        ++load[g];
      }
    }
    std::stable_sort(parts.begin(), parts.end(), [&](const Part& x, const Part& y) { return group(x) < group(y); });
    const int n_q = 4 * n_slots;                                 // quarter-warps over all slots
    std::vector<std::vector<Part>> quarter(n_q);
    for (size_t i = 0; i < parts.size(); ++i) quarter[i % n_q].push_back(parts[i]);   // 8 each
    auto slot_of = [&](int q) { return q % n_slots; };
    // repair: two parts of one interval in one slot -> swap one with a part of the same bank group elsewhere
    for (int pass = 0; pass < 64; ++pass) {
      bool clash = false;
      for (int q = 0; q < n_q; ++q)
        for (size_t i = 0; i < quarter[q].size(); ++i) {
          const Part x = quarter[q][i];
          if (x.iv < 0) continue;
          auto in_slot = [&](int sl, int iv, int skip_q, size_t skip_i) {
            for (int r = 0; r < n_q; ++r)
              if (slot_of(r) == sl)
                for (size_t t = 0; t < quarter[r].size(); ++t)
                  if (!(r == skip_q && t == skip_i) && quarter[r][t].iv == iv) return true;
            return false;
          };
          if (!in_slot(slot_of(q), x.iv, q, i)) continue;
          clash = true;
          bool done = false;
          for (int r = 0; r < n_q && !done; ++r) {
            if (slot_of(r) == slot_of(q)) continue;
            for (size_t t = 0; t < quarter[r].size() && !done; ++t) {
              const Part y = quarter[r][t];
              if (group(y) != group(x)) continue;
              if (in_slot(slot_of(r), x.iv, r, t)) continue;                    // x would clash there
              if (y.iv >= 0 && in_slot(slot_of(q), y.iv, q, i)) continue;      // y would clash here
              quarter[q][i] = y; quarter[r][t] = x; done = true;
            }
          }
        }
      if (!clash) break;
      if (pass == 63) { delete p; return ASSL_ERR_UNSUPPORTED; }
    }
    std::vector<int> first_slot(n_iv, 1 << 30);
    for (int q = 0; q < n_q; ++q)
      for (const Part& x : quarter[q])
        if (x.iv >= 0 && slot_of(q) < first_slot[x.iv]) first_slot[x.iv] = slot_of(q);
    // where an interval's running total lives in the warp's scratch (16-byte entries): a quarter-warp's eight
    // read-modify-writes are conflict-free when their entries differ mod 8, so intervals get one of 8 colours,
    // greedily, least clashes with the intervals they share a quarter with first
    std::vector<int> colour(n_iv, -1), pos(n_iv, 0);
    {
      std::vector<std::vector<int>> mates(n_iv);
      for (int q = 0; q < n_q; ++q)
        for (const Part& x : quarter[q])
          for (const Part& y : quarter[q])
            if (x.iv >= 0 && y.iv >= 0 && x.iv != y.iv) mates[x.iv].push_back(y.iv);
      std::vector<int> by_deg(n_iv);
      for (int j = 0; j < n_iv; ++j) by_deg[j] = j;
      std::stable_sort(by_deg.begin(), by_deg.end(), [&](int x, int y) { return mates[x].size() > mates[y].size(); });
      int used_per_colour[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int j : by_deg) {
        int clash[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int m : mates[j]) if (colour[m] >= 0) ++clash[colour[m]];
        int best = 0;
        for (int c = 1; c < 8; ++c)
          if (clash[c] < clash[best] || (clash[c] == clash[best] && used_per_colour[c] < used_per_colour[best])) best = c;
        colour[j] = best;
        ++used_per_colour[best];
      }
      // repair: while some interval shares its colour with a quarter-mate, give it the colour that clashes least
      // (sideways moves allowed a bounded number of times; 8 colours for quarters of 8 usually end at no clash at all)
      for (int pass = 0; pass < 64; ++pass) {
        bool any = false;
        for (int j : by_deg) {
          int clash[8] = {0, 0, 0, 0, 0, 0, 0, 0};
          for (int m : mates[j]) ++clash[colour[m]];
          if (clash[colour[j]] == 0) continue;
          any = true;
          int best = colour[j];
          for (int k = 1; k <= 8; ++k) {                         // (starting behind the current colour: ties rotate)
            const int c = (colour[j] + k + pass) & 7;
            if (clash[c] < clash[best] || (c != colour[j] && clash[c] == clash[best] && pass % 3 == 2)) best = c;
          }
          colour[j] = best;
        }
        if (!any) break;
      }
      int next_of[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int j = 0; j < n_iv; ++j) pos[j] = colour[j] + 8 * next_of[colour[j]]++;
    }
    if (getenv("ASSL_PLAN_DEBUG")) {                             // how conflict-free the dealing came out
      int group_clash = 0, colour_clash = 0;
      for (int q = 0; q < n_q; ++q) {
        int gs[8] = {0, 0, 0, 0, 0, 0, 0, 0}, cs[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (const Part& x : quarter[q]) { ++gs[group(x)]; if (x.iv >= 0) ++cs[colour[x.iv]]; }
        for (int c = 0; c < 8; ++c) { group_clash += gs[c] > 1 ? gs[c] - 1 : 0; colour_clash += cs[c] > 1 ? cs[c] - 1 : 0; }
      }
      fprintf(stderr, "assl plan: %d parts in %d slots; quarter-warps with a bank group twice: %d, with a total's colour twice: %d\n",
              static_cast<int>(parts.size()), n_slots, group_clash, colour_clash);
    }
    int n_pos = 0;
    for (int j = 0; j < n_iv; ++j) n_pos = n_pos > pos[j] + 1 ? n_pos : pos[j] + 1;
    if (n_pos > assl::kScratchEntries) { delete p; return ASSL_ERR_UNSUPPORTED; }
    // {first bin | valid-bin mask << 10 | scratch entry << 18 | starts-the-total << 26 | owns-an-interval << 27,
    //  alpha, beta (fp32 bits), 0}:  t of the part's bin r is alpha r + beta
    std::vector<int4> lane_iv(static_cast<size_t>(32) * assl::kIvSlots, make_int4(0, 0, 0, 0));
    const double df = sr / n_fft;
    for (int q = 0; q < n_q; ++q)
      for (size_t i = 0; i < quarter[q].size(); ++i) {
        const Part& x = quarter[q][i];
        const int sl = slot_of(q), lane = (q / n_slots) * 8 + static_cast<int>(i);
        if (x.iv < 0) { lane_iv[static_cast<size_t>(sl) * 32 + lane] = make_int4(x.k0, 0, 0, 0); continue; }   // (mask 0, owns nothing)
        int vm = 0;
        for (int r = 0; r < assl::kPartBins; ++r) vm |= (bin_in(x.iv, x.k0 + r) ? 1 : 0) << r;
        const double width = mel_f[x.iv + 1] - mel_f[x.iv];
        const float alpha = static_cast<float>(df / width), beta = static_cast<float>((x.k0 * df - mel_f[x.iv]) / width);
        int ia, ib;
        std::memcpy(&ia, &alpha, 4); std::memcpy(&ib, &beta, 4);
        const int word = x.k0 | (vm << 10) | (pos[x.iv] << 18) | ((first_slot[x.iv] == sl ? 1 : 0) << 26) | (1 << 27);
        lane_iv[static_cast<size_t>(sl) * 32 + lane] = make_int4(word, ia, ib, 0);
      }
    p->h_iv_pos = pos;
    p->n_slots = n_slots;
    p->h_lane_iv = lane_iv;
  }
  // per band {where its rising edge U(m) and its falling edge D(m+1) sit among a warp's interval totals (byte offsets:
  // entry pos[j] = 16 bytes {D(j), U(j)}, each a {frame 2q, frame 2q+1} pair), area normalisation enorm(m),
  // 2.22e-16 * sum of the band's weights (the "+ eps" of utils.py:28, folded)}
  std::vector<int4> band(n_mels);
  for (int m = 0; m < n_mels; ++m) {
    float acc = 0.f;
    for (int k = 0; k < n_bins; ++k) acc += p->h_fb[static_cast<size_t>(m) * n_bins + k] * assl::kPowerEps;
    const float enorm = static_cast<float>(2.0 / (mel_f[m + 2] - mel_f[m]));
    int ie, ia;
    std::memcpy(&ie, &enorm, 4); std::memcpy(&ia, &acc, 4);
    band[m] = make_int4(16 * p->h_iv_pos[m] + 8, 16 * p->h_iv_pos[m + 1], ie, ia);
  }

  // ---- window: scipy get_window('hann', fftbins=True) = 0.5 - 0.5 cos(2 pi n / n_fft) needs no table; the
  //      kernel evaluates it (times the 1/2 of the two-frame unpack) from one sin/cos pair per lane
  // ---- inter-stage twiddles of the 32 x 32 decomposition
  std::vector<float2> tw(32 * 32);
  for (int k2 = 0; k2 < 32; ++k2)
    for (int n1 = 0; n1 < 32; ++n1) {
      const double ang = -2.0 * kPi * static_cast<double>(n1 * k2) / n_fft;
      tw[k2 * 32 + n1] = make_float2(static_cast<float>(std::cos(ang)), static_cast<float>(std::sin(ang)));
    }

  int rc = upload(&p->d_twiddle, tw);
  if (rc == ASSL_OK) rc = upload(&p->d_lane_iv, p->h_lane_iv);
  if (rc == ASSL_OK) rc = upload(&p->d_band, band);
  if (rc == ASSL_OK) rc = upload(&p->d_tickets, std::vector<unsigned>(2 * assl_plan::kTicketSlots, 0u));
  if (rc != ASSL_OK) { assl_plan_destroy(p); return rc; }
  *out = p;
  return ASSL_OK;
}

extern "C" int assl_plan_destroy(assl_plan* p) {
  if (!p) return ASSL_OK;
  cudaFree(p->d_twiddle);
  cudaFree(p->d_lane_iv);
  cudaFree(p->d_band);
  cudaFree(p->d_tickets);
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
