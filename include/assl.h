/*
 * assl.h -- C ABI of the B200-native audio-ssl upstream input path.
 *
 * Drop-in boundary for ONE hot path of Sreyan88/audio-ssl: the per-clip log-mel front end followed
 * by the BYOL-A style two-view augmentation.  The reference is pure Python and has no FFI of its
 * own; each entry point below states the reference interface (file:line under /root/reference)
 * whose arithmetic it replaces.  The Python host side in ``audio-ssl_b200/`` binds these with
 * ctypes and mirrors the reference's class names.
 *
 * Conventions
 *   - every ``const float*`` / ``float*`` / ``double*`` / ``int32_t*`` argument is a raw DEVICE
 *     pointer unless its name starts with ``host_``; the caller owns every data buffer, the
 *     library owns only the plan-lifetime constants (window, twiddles, sparse filterbank);
 *   - every call is asynchronous on ``stream`` (a ``cudaStream_t`` passed as ``void*``), never
 *     synchronises, keeps no global mutable state; a plan is immutable after creation and may be shared
 *     between threads (per-call state lives in assl_pipeline / assl_exchange objects, one caller each);
 *   - return value: 0 = ASSL_OK, otherwise an ``assl_status``; nothing throws, aborts or prints;
 *   - fp32 data, CUDA sm_100a only; there is no CPU fallback.
 *
 * "Ring" addressing: log-mels and their running statistics live in caller-owned rings of ``cap``
 * clip slots so that the spectrogram written by the front end IS the future mixup-bank entry
 * (no copy).  Slot of clip b of a call = (slot0 + b) % cap.  cap == 0 means a plain [B, ...] array.
 */
#ifndef ASSL_H_
#define ASSL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ASSL_VERSION 100 /* 0.1.0 */

typedef enum assl_status {
  ASSL_OK = 0,
  ASSL_ERR_INVALID_ARG = 1,   /* null pointer, negative size, crop outside the canvas ...      */
  ASSL_ERR_UNSUPPORTED = 2,   /* shape/config outside what the kernels are built for           */
  ASSL_ERR_CUDA = 3,          /* allocation / copy / launch failure reported by the runtime    */
  ASSL_ERR_NO_DEVICE = 4,     /* no sm_100 device visible                                      */
  ASSL_ERR_PEER_TIMEOUT = 5,  /* a peer's statistics records did not arrive (assl_exchange_error) */
  ASSL_ERR_TOO_LARGE = 6      /* more clips than the static scratch of a no-allocation entry point holds */
} assl_status;

/* Front-end constants: MelSpectrogramLibrosa.__init__ defaults, src/utils/utils.py:22. */
typedef struct assl_cfg {
  int32_t sample_rate; /* fs     = 16000 */
  int32_t n_fft;       /* n_fft  = 1024 (the only size the kernel is specialised for)          */
  int32_t hop;         /* shift  = 160  (1 .. n_fft)                                           */
  int32_t n_mels;      /* n_mels = 64   (1 .. 128)                                             */
  float fmin;          /* fmin   = 60                                                          */
  float fmax;          /* fmax   = 7800 (<= 0 means sample_rate / 2)                           */
} assl_cfg;

typedef struct assl_plan assl_plan;

/* Parameters of one two-view call that do not vary per clip. */
typedef struct assl_view_cfg {
  int32_t n_freq;      /* F : mel bands per clip                                               */
  int32_t n_time;      /* T : frames per clip                                                  */
  int32_t canvas_h;    /* int(F * virtual_crop_scale[0])   augmentations.py:42                 */
  int32_t canvas_w;    /* int(T * virtual_crop_scale[1])                                       */
  int32_t interp;      /* 0 = bicubic (reference, augmentations.py:27), 1 = bilinear           */
  int32_t log_mixup;   /* 1 = log-mixup-exp (augmentations.py:8-12), 0 = linear mix (:104)     */
  int32_t ring_pitch;  /* floats per row of the log-mel ring / array: 0 = n_time (dense rows); a multiple of 4
                          (with a 16-byte aligned base) lets the kernel read the rows with 128-bit loads   */
} assl_view_cfg;

int assl_version(void);
const char* assl_status_string(int status);

/* -- plan ----------------------------------------------------------------------------------
 * Replaces MelSpectrogramLibrosa.__init__ (src/utils/utils.py:22-24): builds the Slaney mel
 * filterbank the way librosa.filters.mel does (fp64 -> fp32), stores it sparse, and uploads the
 * periodic Hann window and FFT twiddles.  Uses the current CUDA device. */
int assl_plan_create(assl_plan** out, const assl_cfg* cfg);
int assl_plan_destroy(assl_plan* plan);
/* Frames per clip of L samples: 1 + L / hop (librosa.stft, center=True).  <0 on error. */
int64_t assl_plan_num_frames(const assl_plan* plan, int64_t n_samples);
/* Parts per clip of ``tile_sums`` (second dim): one per frame pair, ceil(T / 2). */
int64_t assl_plan_num_tiles(const assl_plan* plan, int64_t n_samples);
/* Copy the dense filterbank [n_mels, 1 + n_fft/2] (fp32) to HOST memory -- for parity tests. */
int assl_plan_filterbank(const assl_plan* plan, float* host_out);

/* -- log-mel front end ---------------------------------------------------------------------
 * Replaces MelSpectrogramLibrosa.__call__ (src/utils/utils.py:26-28) + extract_log_mel_spectrogram
 * (src/utils/utils.py:43-49) for B clips at once:
 *   reflect-pad n_fft/2, frame n_fft / hop, periodic Hann, rFFT, |X|^2 + 2.22e-16, mel, log(. + 1.19e-7).
 * wave      [B, ld] fp32, clip b starts at wave + b*ld, L valid samples (L > n_fft/2, ld >= L)
 * wave_gain optional [B] fp32 per-clip scale applied on load (the reference's F.normalize(p=2),
 *           src/dataset/upstream_dataset.py:62-63, fused); NULL = 1
 * lms       ring/array of [n_mels, T] fp32 slots, see "Ring" above
 * tile_sums optional [B, num_tiles, 2] fp64, 16-byte aligned: per frame pair (sum x, sum x^2) of the values written,
 *           input of assl_running_stats; NULL to skip
 * out_kind  0 = log-mel (extract_log_mel_spectrogram), 1 = mel power (MelSpectrogramLibrosa.__call__
 *           alone, i.e. without the final log(. + eps)); add 2 for time-major slots [T, n_mels], the
 *           layout the MAST variant feeds its transforms (extras/mast_new/mast/moco_dataset.py:105-106
 *           permutes to [1, T, F]); the rings of the two-view path use the default [n_mels, T]    */
int assl_logmel_f32(const assl_plan* plan, const float* wave, const float* wave_gain, int64_t B,
                    int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap,
                    double* tile_sums, int32_t out_kind, void* stream);
/* The same into slots of [n_mels, pitch] floats, pitch >= T (0 = T): the rings of the two-view path pad their rows to a
 * multiple of 4 floats so that the two-view kernel reads them with 128-bit loads (assl_view_cfg.ring_pitch).  The
 * pad columns are never written.  Time-major slots (out_kind | 2) are dense: pitch must be 0 or T. */
int assl_logmel_pitched_f32(const assl_plan* plan, const float* wave, const float* wave_gain, int64_t B,
                            int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap, int64_t pitch,
                            double* tile_sums, int32_t out_kind, void* stream);
/* The same, DEFERRED: the launch does not wait for the kernel launched before it on the stream at its start; every warp
 * loads and transforms its first frame pair the moment its CTA has an SM -- under the tail of that kernel -- and waits
 * for it only before its first STORE (slots, tile_sums).  The caller guarantees that the kernel launched before it does
 * not WRITE what this launch reads (wave, wave_gain: do not use it right behind assl_wave_l2_gain_f32 or
 * assl_extract_window_f32).  Completion stays ordered: a kernel that waits for this launch also gets the completion of
 * everything launched before it.  (The lookahead order of UpstreamInputPipeline: log-mel(k+1) behind the views of
 * batch k-1.) */
int assl_logmel_deferred_f32(const assl_plan* plan, const float* wave, const float* wave_gain, int64_t B,
                             int64_t L, int64_t ld, float* lms, int64_t slot0, int64_t cap, int64_t pitch,
                             double* tile_sums, int32_t out_kind, void* stream);

/* Per-clip sum of squares of the waveform -> gain[b] = 1 / max(||wave_b||_2, 1e-12)
 * (F.normalize(p=2), src/dataset/upstream_dataset.py:62-63). */
int assl_wave_l2_gain_f32(const float* wave, int64_t B, int64_t L, int64_t ld, float* gain,
                          void* stream);

/* -- waveform window (the step before the front end) ---------------------------------------
 * Replaces extract_window (src/utils/utils.py:166-182) for B raw clips of different lengths stored
 * back to back in ``raw``: clip b = raw[offsets[b] .. offsets[b] + lengths[b]).  A clip shorter than
 * unit_length is zero-padded, (unit_length - len) // 2 samples on the left; a longer one is cropped
 * at starts[b], the value the reference draws with random.randint(0, len - unit_length)
 * (starts == NULL: 0).  out [B, ld] fp32 receives the windows.  gain, when not NULL, receives
 * 1 / max(||window_b||_2, 1e-12) -- the reduction of F.normalize(p=2)
 * (src/dataset/upstream_dataset.py:62-63) in the same pass; hand it to assl_logmel_f32. */
int assl_extract_windows_f32(const float* raw, const int64_t* offsets, const int64_t* lengths,
                             const int64_t* starts, int64_t B, int64_t unit_length, float* out,
                             int64_t ld, float* gain, void* stream);

/* -- running pre-normalisation -------------------------------------------------------------
 * Replaces RunningMean.put / RunningVariance.put / RunningNorm.forward's statistics update
 * (src/augmentations/augmentations.py:222-228, 249-256, 275-281), including the divisor that
 * makes the 2nd sample overwrite the 1st, for ``n_clips`` clips in sequence order.
 * state      [8] fp64, caller-owned, zero-filled = fresh: {n_put, sum_mean, sum_var, mean, std, ...}
 * sums       [n_clips, n_parts, 2] fp64 (sum x, sum x^2) per clip part (tile_sums, or the
 *            all-gathered per-clip sums with n_parts = 1); 16-byte aligned
 * elems      elements per clip (F*T)
 * max_update RunningNorm.max_update = epoch_samples * max_update_epochs (augmentations.py:271)
 * mean_std   ring/array of [2] fp32 (mean, std) slots; written for clips [first, first+count)
 *            of the sequence, clip first+b -> slot (slot0 + b) % cap                           */
int assl_running_stats(double* state, const double* sums, int64_t n_clips, int64_t n_parts,
                       int64_t elems, int64_t max_update, int64_t first, int64_t count,
                       float* mean_std, int64_t slot0, int64_t cap, void* stream);

/* Per-clip (sum x, sum x^2) of an arbitrary [B, elems] fp32 array -> sums [B, 2] fp64: the
 * x.mean(axis) / ((x-mu)**2).mean(axis) reductions of RunningMean.put / RunningVariance.put
 * (augmentations.py:223-226, 250) when the clips did not come from assl_logmel_f32. */
int assl_clip_moments_f32(const float* x, int64_t B, int64_t elems, double* sums, void* stream);
/* y[b] = (x[b] - mean_b) / std_b with (mean, std) = mean_std[(slot0 + b) % cap]: the last line of
 * RunningNorm.forward (augmentations.py:282) on its own. */
int assl_clip_normalize_f32(const float* x, float* y, int64_t B, int64_t elems, const float* mean_std,
                            int64_t slot0, int64_t cap, void* stream);

/* -- running pre-normalisation statistics, no allocation anywhere ---------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same result as assl_clip_moments_f32(x, B, elems, sums) followed by assl_running_stats(state, sums, B, 1, elems,
 * max_update, 0, B, mean_std, slot0, cap) -- the same two kernels -- but the [B, 2] fp64 array between them is a
 * file-scope device array inside the library instead of a caller buffer: the call (and everything under it) makes no
 * cudaMalloc / cudaFree and calls no allocating library routine.  The caller still owns x, state and mean_std.
 *   - B <= ASSL_STATIC_MAX_CLIPS (4096 clips = 64 KB, held by every GPU that loads the library until the process
 *     exits); a larger B returns ASSL_ERR_TOO_LARGE and launches nothing;
 *   - every call on a device shares that array and it keeps the last call's contents.  Nothing has to start at zero
 *     (the first kernel overwrites every entry the second one reads), but two calls that overlap on the GPU --
 *     different streams, no ordering between them -- race on it: keep calls on one device in stream order;
 *   - the array's device address is looked up once per device, on the first call made with that device current.   */
#define ASSL_STATIC_MAX_CLIPS 4096
int assl_running_norm_stats_static_f32(double* state, const float* x, int64_t B, int64_t elems,
                                       int64_t max_update, float* mean_std, int64_t slot0, int64_t cap,
                                       void* stream);

/* Reduce tile_sums [B, n_parts, 2] -> clip_sums [B, 2] (what ranks all-gather). */
int assl_clip_sums(const double* tile_sums, int64_t B, int64_t n_parts, double* clip_sums,
                   void* stream);

/* -- global mixup bank across the GPUs of one node (optional) --------------------------------
 * The reference's bank is per worker; with the batch sharded by clip that makes a multi-GPU run differ from
 * one process on the unsharded batch.  In global-bank mode the FIFO runs over the GLOBAL clip order: a view may
 * mix with a clip that lives in another rank's ring, and the kernel reads it there directly (P2P loads over
 * NVLink) -- no copy, no collective.  Together with the statistics exchange this makes the sharded run equal
 * to the single-stream run on the unsharded batch.
 *
 * assl_peer_alloc / assl_peer_open   a cudaMalloc'd buffer with its 64-byte CUDA IPC handle; the handle opened in
 *                        another process of the same node (assl_peer_close / assl_peer_free undo them)
 * assl_two_views_global_f32          as assl_two_views_ordered_f32, plus
 *     peer_lms  HOST array of ``world`` pointers: every rank's log-mel ring as mapped here (own entry = lms)
 *     mix_rank  [B,2] int32  whose ring holds the bank entry of view v (ignored where mix_slot < 0)
 *     clip_stat [B], mix_stat [B,2] int32  slots of the (mean, std) pairs in ``mean_std``, which in this mode is
 *               a ring over GLOBAL clips that every rank fills itself (assl_running_stats_exchange, write_all);
 *               both NULL = the log-mel slots are used, as in the other entry points
 * Ordering between ranks is the caller's: a peer's ring slot may be read once the statistics exchange of the
 * step that wrote it has returned on this rank's stream (engine.ViewEngine keeps the ring three steps deep so
 * that no rank overwrites a slot another rank may still read).                                            */
int assl_peer_alloc(void** ptr, int64_t bytes, void* handle_out);
int assl_peer_free(void* ptr);
int assl_peer_open(const void* handle, void** ptr);
int assl_peer_close(void* ptr);
int assl_two_views_global_f32(const assl_view_cfg* cfg, const float* lms, const float* const* peer_lms,
                              int32_t world, const float* mean_std, const int32_t* clip_slot,
                              const int32_t* clip_stat, const int32_t* mix_slot, const int32_t* mix_rank,
                              const int32_t* mix_stat, const float* alpha, const int32_t* crop,
                              const int32_t* order, int64_t B, float* view1, float* view2, void* stream);

/* -- running pre-normalisation across the GPUs of one node, over peer memory ----------------
 * With the batch sharded by clip (rank r owns a contiguous block of the global clip order), the running
 * statistics of the reference's single stream need every rank's per-clip sums.  Instead of reduce -> NCCL
 * all-gather -> scan (three launches and a collective's latency) every rank stores its per-clip sums straight
 * into every peer's buffer over NVLink (P2P stores of 8-byte words that each carry the call number as their own
 * flag, like NCCL's LL protocol) and scans the records in its own buffer.  Ranks may bring different numbers of
 * clips to a call (an uneven block split, a ragged last batch): a header record per sender says how many, and the
 * scan runs over their concatenation, sender after sender.  No host synchronisation.
 *
 * assl_exchange_create   allocates this rank's buffer (world segments of max_clips + 1 flagged 32-byte records, two
 *                        parities) and writes its 64-byte CUDA IPC handle to handle_out
 * assl_exchange_connect  all_handles = world * 64 bytes, rank-major (all-gather them with any host or device
 *                        collective); maps the peers' buffers.  ASSL_ERR_CUDA when the devices cannot reach each
 *                        other: call assl_exchange_use_staging instead
 * assl_exchange_use_staging  no peer mapping: assl_exchange_send writes this rank's segment into its OWN buffer only
 *                        and the caller moves the segments with any collective before the scan -- all-gather
 *                        ``segment_bytes`` from buffer + rank * segment_bytes into buffer (assl_exchange_staging
 *                        returns both for the call sent last, back = 0, or the one before, back = 1).  Same kernels,
 *                        same results; the fused assl_running_stats_exchange is not available
 * assl_exchange_send     this rank's clips of one call: pair_sums [B, n_parts, 2] (0 <= B <= max_clips) -> per-clip
 *                        totals -> its segment of every rank's buffer.  One small kernel behind the log-mel kernel;
 *                        nobody waits in it
 * assl_running_stats_exchange_scan   the scan of the call sent last: waits until every sender's records of that call
 *                        are in this rank's buffer, then as assl_running_stats over their concatenation; (mean, std) of
 *                        the local slice is written to mean_std slot (slot0 + b) % cap (write_all != 0: of ALL clips,
 *                        global clip g -> slot (slot0 + g) % cap -- the statistics ring of the global-bank mode, which
 *                        needs the same B on every rank).  deferred != 0: the launch does not wait for the kernel
 *                        launched before it on the stream at its start but at its end, like assl_two_views_deferred_f32
 *                        -- everything it depends on travels in the flagged records and in ``state`` (written by the
 *                        previous scan, which a log-mel launch in between has already waited for).  Issued one step
 *                        AFTER the send -- log-mel(k) . scan(k-1) . send(k) . views(k-2) -- the records have long
 *                        landed and no rank waits for the slowest one inside a step.  A call has to be scanned before
 *                        the next one is sent (two parities)
 * assl_exchange_scan_send  the two above in ONE launch for the lookahead order: the deferred scan of the call sent a step
 *                        ago (exactly one call must be pending), then -- in the same cluster, after the kernel launched
 *                        before it on the stream (the log-mel kernel of THIS step) has completed -- the send of this
 *                        step's pair_sums [B, n_parts, 2].  Step order log-mel(k) . scan(k-1)+send(k) . views(k-2): one
 *                        small kernel between the log-mel kernel and the views instead of two
 * assl_running_stats_exchange  send + scan of one call in ONE kernel (peer transport), for the ordinary step order:
 *                        arguments as assl_running_stats, except pair_sums [B, n_parts, 2] are THIS rank's clips
 * assl_exchange_error    ASSL_ERR_PEER_TIMEOUT once a scan gave up waiting for a peer (assl_exchange_set_timeout,
 *                        30 s by default): that scan left the running statistics UNTOUCHED (its clips got the last
 *                        published mean/std, state[7] = 1) and the ranks' statistics no longer agree -- treat it as
 *                        fatal.  Reads a mapped host word the kernel sets: no synchronisation, callable every step */
typedef struct assl_exchange assl_exchange;
#define ASSL_EXCHANGE_HANDLE_BYTES 64
#define ASSL_EXCHANGE_MAX_WORLD 16
int assl_exchange_create(assl_exchange** out, int32_t rank, int32_t world, int64_t max_clips,
                         void* handle_out);
int assl_exchange_connect(assl_exchange* x, const void* all_handles);
int assl_exchange_use_staging(assl_exchange* x);
int assl_exchange_staging(const assl_exchange* x, int64_t back, void** buffer, int64_t* segment_bytes);
int assl_exchange_set_timeout(assl_exchange* x, double seconds);
int assl_exchange_error(const assl_exchange* x);
int assl_exchange_destroy(assl_exchange* x);
int assl_exchange_send(assl_exchange* x, const double* pair_sums, int64_t B, int64_t n_parts, void* stream);
int assl_running_stats_exchange_scan(assl_exchange* x, double* state, int64_t elems, int64_t max_update,
                                     float* mean_std, int64_t slot0, int64_t cap, int32_t write_all,
                                     int32_t deferred, void* stream);
int assl_exchange_scan_send(assl_exchange* x, double* state, int64_t elems, int64_t max_update, float* mean_std,
                            int64_t slot0, int64_t cap, int32_t write_all, const double* pair_sums, int64_t B,
                            int64_t n_parts, void* stream);
int assl_running_stats_exchange(assl_exchange* x, double* state, const double* pair_sums, int64_t B,
                                int64_t n_parts, int64_t elems, int64_t max_update, float* mean_std,
                                int64_t slot0, int64_t cap, int32_t write_all, void* stream);

/* -- two views -----------------------------------------------------------------------------
 * Replaces, per clip and per view, RunningNorm.forward's (x-mean)/std (augmentations.py:282),
 * MixupBYOLA.forward's mix (:99-106, log_mixup_exp :8-12) and RandomResizeCrop.forward (:40-55:
 * zero canvas, centred paste, crop, F.interpolate(bicubic, align_corners=True)) -- i.e. the
 * body of AugmentationModule.__call__ (src/augmentations/__init__.py:32-35) -- with every random
 * draw passed in explicitly.
 * lms, mean_std  rings (cap slots) holding this call's clips AND the bank sources
 * mean_std may be NULL (no pre-normalisation: x is used as is)
 * clip_slot [B]      int32 ring slot of each clip's log-mel
 * mix_slot  [B,2]    int32 ring slot of the bank entry mixed into view v, -1 = no mix
 * alpha     [B,2]    fp32 weight of the bank entry (the clip gets 1 - alpha)
 * crop      [B,2,4]  int32 (i, j, h, w) in canvas coordinates, as RandomResizeCrop.get_params
 *                    (augmentations.py:30-38) returns them
 * view1/2   [B, F, T] fp32 outputs
 * The host-side mirror validates crops (1<=h<=canvas_h, i+h<=canvas_h, ...) before calling; the
 * kernel additionally clamps them into the canvas so a bad box cannot leave its shared memory.  */
int assl_two_views_f32(const assl_view_cfg* vcfg, const float* lms, const float* mean_std,
                       int64_t cap, const int32_t* clip_slot, const int32_t* mix_slot,
                       const float* alpha, const int32_t* crop, int64_t B, float* view1,
                       float* view2, void* stream);

/* The same, with the order in which the views are started: order [2B] int32 on the device, launch position ->
 * view id (2 * clip + view), a permutation.  A view's cost grows with its crop (rows x 32-column groups of the
 * staged span), CTAs start in launch order, so putting the big crops first shortens the tail of the launch
 * (1024 clips = 3.5 waves of CTAs).  NULL = clip order.  The results do not depend on it.                 */
int assl_two_views_ordered_f32(const assl_view_cfg* cfg, const float* lms, const float* mean_std,
                               int64_t cap, const int32_t* clip_slot, const int32_t* mix_slot,
                               const float* alpha, const int32_t* crop, const int32_t* order, int64_t B,
                               float* view1, float* view2, void* stream);

/* The same as assl_two_views_ordered_f32 for a software-pipelined step order: log-mel(k) . statistics(k) . views(k-1).
 * The launch does NOT wait for the kernel launched before it on the stream (it may run beside it, filling the SMs the
 * log-mel kernel's tail leaves idle while the statistics kernel of the next batch runs); it waits for that kernel at
 * its END, so whatever is launched next still sees stream order.  The caller guarantees (a) that the kernel launched
 * immediately before neither writes anything this call reads (ring slots of this batch and of its bank entries, their
 * mean/std pairs, the parameter arrays) nor touches view1/view2 -- with the ring that means cap >= bank clips + 2 B --
 * and (b) that everything this call reads was produced by kernels launched BEFORE the last assl_logmel_f32 on this
 * stream (that kernel releases its dependents only after everything older has completed).  Same results as
 * assl_two_views_ordered_f32; replaces nothing in the reference (its DataLoader overlaps batches across workers,
 * src/dataset/upstream_dataset.py:118-125).                                                                        */
int assl_two_views_deferred_f32(const assl_view_cfg* cfg, const float* lms, const float* mean_std,
                               int64_t cap, const int32_t* clip_slot, const int32_t* mix_slot,
                               const float* alpha, const int32_t* crop, const int32_t* order, int64_t B,
                               float* view1, float* view2, void* stream);

/* -- Kmix cluster features ------------------------------------------------------------------
 * The data-dependent half of Kmix.get_index (src/augmentations/augmentations.py:142-165): per clip, the
 * per-band mean over time is assigned to the nearest of K centroids, once L2-normalised
 * (a bank entry's role, :148-151, 153 -> mem_cluster) and once raw (the incoming clip's role, :154 ->
 * point_cluster).  centroids_unit [K, n_freq] holds the centroids already divided by their L2 norms (:146).
 * x [B, n_freq, n_time] fp32; outputs [B] int32, ties to the lowest index.  The choice among the bank entries
 * of the farthest non-empty cluster, the random draws and the bank bookkeeping stay on the host; the mix is
 * the MixupBYOLA arithmetic (assl_two_views_f32 with an identity crop). */
int assl_clip_clusters_f32(const float* x, int64_t B, int32_t n_freq, int32_t n_time,
                           const float* centroids_unit, int32_t K, int32_t* mem_cluster,
                           int32_t* point_cluster, void* stream);

/* -- batch post-normalisation --------------------------------------------------------------
 * Replaces NormalizeBatch.forward (src/augmentations/augmentations.py:321-324).
 * assl_batch_sums accumulates {count, sum x, sum x^2} of x[n] into sums[3] (fp64; the caller zeroes
 * it, may all-reduce it across ranks); assl_normalize_apply then writes
 * y = (x - mean) / clamp(std_unbiased, eps, max).  x == y is allowed.                          */
int assl_batch_sums(const float* x, int64_t n, double* sums, void* stream);
int assl_normalize_apply_f32(const float* x, float* y, int64_t n, const double* sums, void* stream);
/* PrecomputedNorm.forward (augmentations.py:299-300): y = (x - mean) / std. */
int assl_affine_norm_f32(const float* x, float* y, int64_t n, float mean, float std, void* stream);

/* -- whole path, HOST buffers (what a DataLoader-side caller binds) --------------------------
 * host_wave [B, L] (pinned, so the copies are asynchronous) -> H2D -> log-mel -> running stats -> two views
 * -> D2H into host_view1/2 [B, F, T].  Parameter arrays are HOST pointers with the layouts above (slots refer
 * to the ring inside ``ws``).  ``ws`` is a caller-owned DEVICE workspace of at least
 * assl_pipeline_workspace_bytes(), zero-filled before first use; the ring (cap slots) and the norm state
 * live at its start and persist between calls.  Kernels run in order on ``stream``; uploads and downloads
 * run on two copy streams of the pipeline object, chunked and double-buffered so that consecutive calls overlap
 * upload, kernels and download.  Every host buffer passed in must stay untouched, and host_view* is complete, only
 * after assl_pipeline_host_sync() returns.  host_view* may be NULL to leave the views on the device:
 * their device addresses are returned through dev_view1/2 and are valid in ``stream`` order until the call
 * after next.
 * assl_pipeline_create: the pipeline object holds the copy streams, events and the call parity (the plan stays
 * immutable and shareable); one caller at a time per PIPELINE.  ``exchange`` (may be NULL) is the connected
 * peer-transport statistics exchange of a clip-sharded multi-GPU run: every call then runs the fused send + scan
 * (assl_running_stats_exchange) over all ranks' clips instead of the local scan, and every rank must make the
 * same sequence of calls.                                                                                 */
typedef struct assl_pipeline assl_pipeline;
int assl_pipeline_create(assl_pipeline** out, const assl_plan* plan, assl_exchange* exchange);
int assl_pipeline_destroy(assl_pipeline* pipe);
int64_t assl_pipeline_workspace_bytes(const assl_plan* plan, int64_t B, int64_t L, int64_t cap);
int assl_pipeline_host_f32(assl_pipeline* pipe, const assl_view_cfg* vcfg, void* ws,
                           int64_t ws_bytes, const float* host_wave, int64_t B, int64_t L,
                           int64_t slot0, int64_t cap, int64_t max_update,
                           const int32_t* host_mix_slot, const float* host_alpha,
                           const int32_t* host_crop, float* host_view1, float* host_view2,
                           float** dev_view1, float** dev_view2, void* stream);
/* The same with 16-bit PCM waveforms in host memory -- the samples a 16-bit audio file holds; the reference gets them
 * as x / 32768 in fp32 from librosa.core.load (src/dataset/upstream_dataset.py:55) and so does the kernel chain here:
 * half the upload, widened on the device (exact), same views bit for bit as assl_pipeline_host_f32 on x / 32768. */
int assl_pipeline_host_pcm16(assl_pipeline* pipe, const assl_view_cfg* vcfg, void* ws,
                             int64_t ws_bytes, const int16_t* host_wave, int64_t B, int64_t L,
                             int64_t slot0, int64_t cap, int64_t max_update,
                             const int32_t* host_mix_slot, const float* host_alpha,
                             const int32_t* host_crop, float* host_view1, float* host_view2,
                             float** dev_view1, float** dev_view2, void* stream);
int assl_pipeline_host_sync(assl_pipeline* pipe);
/* Device buffers: out[i] = pcm[i] / 32768 (the widening step on its own). */
int assl_pcm16_to_f32(const int16_t* pcm, int64_t n, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ASSL_H_ */
