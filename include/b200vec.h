/* b200vec.h — C-ABI of the B200-native vector-build engine (libb200vec.so).
 *
 * Drop-in boundary for ONE hot path of ramiyako/vector: the composite I/Q vector build
 * (generate_vector in unified_gui.py / main.py) and the utils.py helpers it leans on.
 * The reference has no native interface; its boundary is a set of Python names.  Each
 * entry point below says which reference lines it replaces (paths relative to the
 * reference root).  INTEGRATION.md shows the ctypes binding a maintainer adds.
 *
 * Conventions
 *   - plain C types only; every pointer called *_dev is a CUDA device pointer on the
 *     context's device, every other pointer is host memory;
 *   - complex64 is interleaved float32 (re, im), 8 bytes per sample — numpy's layout;
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all
 *     calls are asynchronous on it unless stated otherwise;
 *   - every function returns 0 on success, non-zero on error and never throws;
 *     b200vec_last_error() gives a thread-local message for the last failure;
 *   - the caller owns every buffer it passes; the library owns only its context
 *     (packet bank, twiddle/window tables, select/FFT scratch) and allocates nothing
 *     inside build/stft/db/select/quantize calls once scratch is warm.
 *   - there is NO CPU fallback: without a CUDA device b200vec_create fails.
 */
#ifndef B200VEC_H
#define B200VEC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200VEC_ABI_VERSION 2
#define B200VEC_MAX_PACKETS 32      /* descriptors per build call (reference: MAX_PACKETS = 6, unified_gui.py:41) */
#define B200VEC_MAX_SLOTS   64      /* packet-bank slots per context */
#define B200VEC_MAX_RANKS   8       /* order statistics per select call */

typedef struct b200vec_ctx b200vec_ctx;

/* One packet's placement, all integers computed on the host exactly as the reference does
 * (unified_gui.py:1748-1757, main.py:247-253): instance k sits at start + k*period for
 * 0 <= k < count, and `count` is the number of instances that fit entirely. */
typedef struct b200vec_pkt_desc {
    int64_t start;     /* first instance, vector-absolute sample index, >= 0 */
    int64_t period;    /* > 0 */
    int64_t count;     /* K_p, may be 0 */
    int32_t slot;      /* bank slot holding the committed (resampled/shifted) packet */
    int32_t reserved;  /* bit 0 (optional hint for the max pass of the normalised builds): the caller vouches that
                        * every sample of this packet appears unsummed somewhere in the vector (for every packet-local
                        * index some instance has no other instance over that sample — e.g. one free-standing instance);
                        * max |y|^2 of the whole packet (kept by the bank) is then attained by the vector, and the max
                        * pass skips every sample lying under exactly one instance of it.  0 = no hint. */
} b200vec_pkt_desc;

enum b200vec_build_mode {
    B200VEC_BUILD_STORE   = 0,  /* place and store, no normalisation (normalize unchecked) */
    B200VEC_BUILD_MAXONLY = 1,  /* pass A: place in registers, reduce max |v|^2, store nothing */
    B200VEC_BUILD_SCALE   = 2   /* pass B: place, multiply by 1/sqrt(*maxsq_dev), store */
};

int         b200vec_abi_version(void);
const char *b200vec_last_error(void);

/* Lifetime.  `device` is a CUDA ordinal. */
int b200vec_create(int device, b200vec_ctx **out);
int b200vec_destroy(b200vec_ctx *ctx);

/* ---- packet bank -------------------------------------------------------------------
 * upload   : what load_packet_info returns (utils.py:89-105), complex64 [len].
 * resample : scipy.signal.resample(signal, len_out) as called at utils.py:117
 *            (Fourier method: forward DFT, copy the min(len,len_out) lowest-|f| bins,
 *            split/unite the unpaired Nyquist bin, inverse DFT); computed in fp64 on the
 *            device (Bluestein over power-of-two FFTs), rounded once to complex64.
 * shift    : apply_frequency_shift (utils.py:120-127): y[n] *= exp(j*(2*pi*f)*(n/fs)),
 *            phase in fp64 from the integer packet-local index, one rounding to complex64.
 *            freq_hz == 0 commits the samples unchanged (utils.py:122-123).
 *            shift also "commits" the slot: it writes the copies the placement kernels
 *            read.  A slot must be committed before b200vec_build.
 * attach   : make caller-owned device memory (complex64 [len], 8-byte aligned) a committed slot
 *            without copying it — e.g. one common period of an already built vector, which the
 *            next launch then places like a packet of period == len (dist.build_stft_slab's
 *            periodic tiling: the composite of unified_gui.py:1757-1769 repeats bit for bit with the
 *            least common multiple of the packet periods).  Works with b200vec_build* and the
 *            frame-major fused step; the memory must stay valid and unchanged while launches that
 *            name the slot run.
 * download : committed samples back to the host (synchronises the stream). */
int b200vec_bank_upload(b200vec_ctx *ctx, int slot, const void *src, int64_t len,
                        int src_is_device, void *stream);
int b200vec_bank_resample(b200vec_ctx *ctx, int slot, int64_t len_out, void *stream);
int b200vec_bank_shift(b200vec_ctx *ctx, int slot, double freq_hz, double fs_hz, void *stream);
int b200vec_bank_attach(b200vec_ctx *ctx, int slot, const void *samples_dev, int64_t len, void *stream);
int b200vec_bank_len(b200vec_ctx *ctx, int slot, int64_t *len_out);
int b200vec_bank_download(b200vec_ctx *ctx, int slot, void *dst_host, int64_t len, void *stream);

/* ---- vector build ------------------------------------------------------------------
 * Replaces np.zeros + the placement loop + normalisation:
 *   unified_gui.py:1712, 1757-1769, 1778-1782;  main.py:218, 253-257, 274-277.
 * Computes output samples [slab_lo, slab_hi) of the vector into out_dev[0 .. slab_hi-slab_lo)
 * (complex64, 16-byte aligned).  Every sample of the slab is written exactly once
 * (zeros included); contributions are summed in (packet, instance) order like the
 * reference's in-place adds.  maxsq_dev points at ONE device double holding max |v|^2:
 *   MAXONLY atomically raises it (caller zeroes it first; a multi-GPU caller all-reduces
 *   it with MAX between the passes), SCALE reads it and multiplies every sample by
 *   fl32(1 / fl32(sqrt(maxsq))) when maxsq > 0 — numpy's complex64 / float32-scalar. */
int b200vec_build(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                  int64_t slab_lo, int64_t slab_hi, void *out_dev, int mode,
                  double *maxsq_dev, void *stream);

/* The normalised build in ONE launch (unified_gui.py:1778-1782 folded in): max |v|^2 over the sample
 * ranges norm_ranges[2r] .. norm_ranges[2r+1] (n_ranges <= 4 [lo, hi) pairs, vector-absolute; ranges
 * whose max is the max of the whole vector — all of it, or one common period of the packet trains plus
 * head and tail), a grid barrier (cooperative launch), then the SCALE pass of b200vec_build.
 * *maxsq_dev receives max |v|^2.  8 B per output sample of HBM traffic, nothing read. */
int b200vec_build_norm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                       int64_t slab_lo, int64_t slab_hi, void *out_dev,
                       const int64_t *norm_ranges, int n_ranges, double *maxsq_dev, void *stream);

/* ---- scenario batch (BASELINE.json configs[4]: 256 vectors sweeping shift / period / offset) ----------
 * The reference builds a scenario sweep by running generate_vector once per scenario
 * (unified_gui.py:1722-1782 in a loop).  Here the whole sweep is two calls:
 *   prepare : every packet of every scenario, pkts[first[v] .. first[v+1]) for vector v.  Each distinct
 *             (src_slot, freq_hz) pair is shifted ONCE (apply_frequency_shift, utils.py:120-127, same
 *             arithmetic as b200vec_bank_shift) into a batch bank owned by the context, in one launch;
 *             src_slot names an uploaded (optionally resampled) packet, it need not be committed.
 *             Synchronises the stream.
 *   build   : ONE launch places, (per vector) normalises and stores all vectors:
 *             out_dev[v * out_ld + i] = sample i of vector v, 0 <= i < total.  normalize != 0: each vector
 *             is divided by its own max |v| (maxsq_dev: n_vectors doubles, receives max |v|^2 per vector);
 *             the kernel finds the maxima itself (cooperative launch, grid barrier between the passes).
 *             Independent vectors: no collective; shard a sweep over GPUs by vector. */
typedef struct b200vec_batch_pkt {
    int64_t start;     /* as b200vec_pkt_desc */
    int64_t period;
    int64_t count;
    int32_t src_slot;  /* bank slot holding the packet BEFORE the frequency shift */
    int32_t flags;     /* bit 0: the free-instance hint of b200vec_pkt_desc::reserved */
    double  freq_hz;   /* this scenario's shift; 0 = none */
} b200vec_batch_pkt;

int b200vec_batch_prepare(b200vec_ctx *ctx, const b200vec_batch_pkt *pkts, const int32_t *first, int n_vectors,
                          double fs_hz, void *stream);
int b200vec_batch_build(b200vec_ctx *ctx, int64_t total, void *out_dev, int64_t out_ld, int normalize,
                        double *maxsq_dev, void *stream);

/* ---- spectrogram -------------------------------------------------------------------
 * scipy.signal.spectrogram(x[::stride], window, nperseg, noverlap=nperseg-hop, nfft,
 * return_onesided=False, detrend=False, scaling) followed by fftshift(axes=0), as called at
 * utils.py:281-291/303-313/324-347/350-351 and vector_analyzer/spectrogram_analysis.py:40-53.
 * Frame f (global index) reads decimated samples f*hop .. f*hop+nperseg-1, i.e. vector
 * samples (f*hop+j)*stride; in_dev[0] is vector sample `in_origin`.  Frames
 * [frame_lo, frame_lo+n_frames) are written to out_dev[row*out_ld + (f-frame_lo)], row r
 * holding bin (r + nfft/2) mod nfft (the fftshift); float32 PSD = |X|^2 * scale.
 * *nonzero_flag_dev (int, optional) is OR-ed with 1 if any output > 0 (utils.py:316).
 * window_dev: nperseg float32 taps on the device (the host builds them the way
 * scipy.signal.get_window does and casts to float32, as _spectral_helper does).
 * nfft must be a power of two in [32, 16384]; nperseg <= nfft. */
int b200vec_stft(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                 int64_t stride, int nperseg, int hop, int nfft, const float *window_dev, float scale,
                 int64_t frame_lo, int64_t n_frames, float *out_dev, int64_t out_ld,
                 int *nonzero_flag_dev, void *stream);

/* The same call for a double-precision signal: scipy.signal.spectrogram on complex128 input (create_spectrogram
 * given np.complex128 / np.float64 data, utils.py:281-291) multiplies by the float64 window, transforms with the
 * complex128 FFT and returns a float64 PSD.  in_dev: complex128 (16-byte aligned), window_dev: nperseg float64 taps,
 * out_dev: float64 [nfft, out_ld].  nfft must be a power of two in [32, 8192] (16 B of shared memory per bin). */
int b200vec_stft_f64(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                     int64_t stride, int nperseg, int hop, int nfft, const double *window_dev, double scale,
                     int64_t frame_lo, int64_t n_frames, double *out_dev, int64_t out_ld,
                     int *nonzero_flag_dev, void *stream);

/* ---- fused build + full-rate spectrogram (the BASELINE cfg2 + cfg3 step in ONE kernel) -----
 * Pass B of b200vec_build (mode STORE or SCALE) and b200vec_stft for the fixed analysis shape of
 * vector_analyzer/spectrogram_analysis.py:35-49 (nperseg = nfft = 1024, hop = 512, stride 1,
 * two-sided, fftshifted rows) without re-reading the vector: a thread block places the samples of
 * 16 frames (unified_gui.py:1757-1769,1781), streams them to vec_out_dev and transforms them out
 * of shared memory.  HBM traffic 16 B per output sample (8 vector + 8 spectrogram), no reads.
 *   vec_out_dev[0 .. slab_hi-slab_lo) receives samples [slab_lo, slab_hi);
 *   frames [frame_lo, frame_lo+n_frames) go to spec_out_dev[row*out_ld + (f-frame_lo)];
 *   slab_lo must equal frame_lo*512 (slabs are cut on hop boundaries) and every frame must start
 *   inside the slab; samples past slab_hi that the last frames need (the <= 512-sample halo) are
 *   computed but not stored.  SCALE reads *maxsq_dev exactly like b200vec_build. */
int b200vec_build_stft1024(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                           int64_t slab_lo, int64_t slab_hi, void *vec_out_dev, int mode,
                           double *maxsq_dev, const float *window_dev, float scale,
                           int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                           int *nonzero_flag_dev, void *stream);

/* Same step, spectrogram stored FRAME-MAJOR: spec_out_dev[(f-frame_lo)*out_ld + row], out_ld >= 1024
 * and a multiple of 4, base 16-byte aligned.  The [1024, frames] array the reference hands to its
 * callers (vector_analyzer/spectrogram_analysis.py:49-53: Sxx after fftshift) is then the transpose
 * VIEW of this buffer — same shape, dtype and values, Fortran-ordered — and one frame's bins are
 * 4 KB of contiguous memory, which lets every warp transform and store its own frames with no
 * block-level synchronisation (wf1024.cuh).  Work items are handed out through a counter owned by
 * the context: do not run two of these calls concurrently on one context. */
int b200vec_build_stft1024_fm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                              int64_t slab_lo, int64_t slab_hi, void *vec_out_dev, int mode,
                              double *maxsq_dev, const float *window_dev, float scale,
                              int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                              int *nonzero_flag_dev, void *stream);

/* The whole normalised step in ONE launch (single-GPU builds; unified_gui.py:1778-1782 folded in):
 * the kernel first reduces max |v|^2 over the sample ranges norm_ranges[2r] .. norm_ranges[2r+1]
 * (host array of n_ranges <= 4 [lo, hi) pairs, vector-absolute; the caller passes ranges whose max
 * is the max of the whole vector — the entire vector, or one common period of the packet trains
 * plus head and tail), meets at a grid barrier (cooperative launch) and then runs the SCALE pass of
 * b200vec_build_stft1024_fm with that max.  *maxsq_dev (optional) receives max |v|^2.  A multi-GPU
 * build keeps the two-launch form: b200vec_build(MAXONLY), all-reduce MAX, ..._fm(SCALE). */
int b200vec_build_stft1024_fm_norm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                                   int64_t slab_lo, int64_t slab_hi, void *vec_out_dev,
                                   const int64_t *norm_ranges, int n_ranges, double *maxsq_dev,
                                   const float *window_dev, float scale,
                                   int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                                   int *nonzero_flag_dev, void *stream);

/* Periodic spectrogram tiling for dense periodic vectors (BASELINE.json configs[3]): where the vector repeats bit for
 * bit with period T, its full-rate spectrogram (hann 1024 / hop 512, vector_analyzer/spectrogram_analysis.py:35-49)
 * repeats with period lcm(T, 512) / 512 frames — frame f holds the same 1024 samples as frame f - period.  Copies,
 * in frame-major memory (see b200vec_build_stft1024_fm),
 *   spec[(frame_dst0 + i) * out_ld ..] = spec[(frame_src0 + (frame_dst0 + i - frame_src0) mod period_frames) * out_ld ..]
 * for 0 <= i < n_frames (1024 bins each); frame_dst0 >= frame_src0 + period_frames.  The caller (dist.build_stft_slab,
 * hostmath.tile_plan) guarantees that every destination frame lies inside the bit-exactly periodic part. */
int b200vec_replicate_frames(b200vec_ctx *ctx, float *spec_dev, int64_t out_ld, int64_t frame_src0, int64_t period_frames,
                             int64_t frame_dst0, int64_t n_frames, void *stream);

/* b200vec_stft for the full-rate analysis shape (hann-type window of 1024 taps, nperseg = nfft = 1024,
 * hop 512, stride 1; vector_analyzer/spectrogram_analysis.py:40-53, mat_analyzer.py:37-45) with the
 * spectrogram stored FRAME-MAJOR: out_dev[(f-frame_lo)*out_ld + row] (see b200vec_build_stft1024_fm).
 * in_dev[0] is vector sample in_origin; 16 B per output sample of HBM traffic (8 read + 8 written). */
int b200vec_stft1024_fm(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                        const float *window_dev, float scale, int64_t frame_lo, int64_t n_frames,
                        float *out_dev, int64_t out_ld, int *nonzero_flag_dev, void *stream);

/* The multi-GPU form of b200vec_build_stft1024_fm_norm: still ONE launch per GPU and step, and the
 * path's only collective — MAX of one double — runs INSIDE the kernel over NVLink peer memory
 * instead of as a separate NCCL all-reduce between two launches.  Every rank passes its share of the
 * norm ranges and the same `epoch` (1, 2, 3, ... one per step; 0 = count the steps on the device, in
 * the u32 at byte 24 of the home block — a step can then be replayed from a CUDA graph).
 * peer_blocks[r] is rank r's 64-byte,
 * zero-initialised, symmetric block mapped into this process (torch symmetric memory, cuMem or
 * cudaIpc): u64[0..1] max slots for even / odd epochs, u32 at byte 16 the arrival counter.  CTA 0
 * raises every GPU's slot with the local max (system-scope atomicMax), signals every GPU and waits
 * for `world` arrivals at home; all ranks then scale by the same exact max.  world <= 8. */
int b200vec_build_stft1024_fm_norm_peer(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                                        int64_t slab_lo, int64_t slab_hi, void *vec_out_dev,
                                        const int64_t *norm_ranges, int n_ranges,
                                        void *const *peer_blocks, int world, int rank, uint32_t epoch,
                                        double *maxsq_dev, const float *window_dev, float scale,
                                        int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                                        int *nonzero_flag_dev, void *stream);

/* *status_host = 1 when a b200vec_build_stft1024_fm_norm_peer launch on this context gave up waiting for the
 * other GPUs (the wait inside the kernel is bounded, ~2 s: a rank that never launched must not hang its
 * peers); the step's output is then scaled by the local max only and must be discarded.  Reading clears
 * the flag.  Synchronises the stream. */
int b200vec_peer_status(b200vec_ctx *ctx, int *status_host, void *stream);

/* ---- dB map and order statistics (normalize_spectrogram, utils.py:356-404) -----------
 * db      : out[i] = 10*log10(|s[i]| + floor)                      (utils.py:363-368)
 * select  : exact order statistics of a float32 array: ranks[] are 0-based positions in the
 *           ascending sort of the selected elements (all, or only those > 0 when
 *           positive_only — utils.py:365 takes the percentile of Sxx_abs[Sxx_abs > 0]).
 *           count_host receives how many elements were selected.  Two-call protocol so the
 *           host can derive ranks from the count the way np.percentile does:
 *           select_count() then select_ranks().  Both synchronise the stream. */
int b200vec_db(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val,
               float *out_dev, void *stream);
/* b200vec_db plus the same map applied to n_values (<= 8) host values in the same launch:
 * values_out_host[i] = 10*log10(|values_host[i]| + floor) computed by the kernel's own instruction sequence —
 * normalize_spectrogram's vmin / vmax (utils.py:372-373) are then dB-image values bit for bit, found as order
 * statistics of |S| (dB is monotone in |S|).  Synchronises the stream when n_values > 0. */
int b200vec_db_values(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val, float *out_dev,
                      const float *values_host, int n_values, float *values_out_host, void *stream);
int b200vec_select_count(b200vec_ctx *ctx, const float *data_dev, int64_t n, int positive_only,
                         int64_t *count_host, void *stream);
int b200vec_select_ranks(b200vec_ctx *ctx, const float *data_dev, int64_t n, int positive_only,
                         const int64_t *ranks_host, int n_ranks, float *values_host,
                         void *stream);

/* Order statistics of a large array in two full reads instead of three per call (what
 * normalize_spectrogram needs: the 5th percentile of the positive elements and the low / high percentiles,
 * minimum and maximum of the dB image — dB is monotone in |S|, so all eight ranks are ranks of |S|):
 *   begin  : 13-bit histogram of the order-preserving keys of data (|data| when use_abs) in one read;
 *            *count_pos_host = number of elements > 0.  Synchronises the stream.
 *   finish : exact values at ranks_host[] (0-based positions in the ascending sort of ALL n elements) —
 *            the host picks each rank's histogram bin, one more read compacts the elements of those bins,
 *            and the exact radix select of b200vec_select_ranks runs on the candidates only.
 *            Same array, unchanged, as in begin.  Synchronises the stream. */
int b200vec_select_begin(b200vec_ctx *ctx, const float *data_dev, int64_t n, int use_abs,
                         int64_t *count_pos_host, void *stream);
int b200vec_select_finish(b200vec_ctx *ctx, const float *data_dev, int64_t n, const int64_t *ranks_host,
                          int n_ranks, float *values_host, void *stream);

/* normalize_spectrogram's device work in ONE call (utils.py:356-373, called with 5 / 95 / 25 by plot_spectrogram,
 * utils.py:431): select_begin, numpy's float32 percentile index arithmetic on the host (npquant.h), select_finish,
 * the floor, the dB map.  db_out_dev[i] = 10*log10(|s[i]| + floor); values_host[0] = floor,
 * [1..6] = the dB-image values at the ranks (low percentile: lower / upper neighbour; high percentile: lower /
 * upper neighbour; minimum; maximum), [7], [8] = the float32 interpolation weights of the two percentiles;
 * *n_pos_host = number of elements > 0.  The caller finishes with numpy's lerp and utils.py:375-397.  n >= 1.
 * Synchronises the stream. */
int b200vec_normalize_db(b200vec_ctx *ctx, const float *s_dev, int64_t n, double low_percentile, double high_percentile,
                         float *db_out_dev, float *values_host, int64_t *n_pos_host, void *stream);

/* out_dev[i] = in_dev[i * stride], i < ceil(n_in / stride): the naive decimation sig[::factor] of
 * create_spectrogram (utils.py:192-195), made once so that the overlapping preview frames read a compact
 * copy. */
int b200vec_decimate(b200vec_ctx *ctx, const void *in_dev, int64_t n_in, int64_t stride, void *out_dev, void *stream);

/* ---- .wv payload (mat2wv, vector_analyzer/mat_to_wv_converter.py:28-50) ---------------
 * quantize    : out[2i] = (int16) trunc(fl(re * pre_scale) * 32767), out[2i+1] likewise for im.  pre_scale = 1
 *               is mat2wv(bNormalize=False), the form save_vector_wv / generate_vector use; the bNormalize
 *               branch passes fl32(1 / max |x|) (numpy divides complex64 by a float32 scalar that way).
 * power_stats : the level metrics of that branch for y = fl(x * pre_scale): stats_host[0] = max |y|^2,
 *               stats_host[1] = sum of fl32(fl32(|y|)^2) accumulated in fp64 (the reference's float32
 *               np.mean is a pairwise float32 sum: equal to ~1e-7 relative).  Synchronises the stream. */
int b200vec_wv_quantize(b200vec_ctx *ctx, const void *in_dev, int64_t n, float pre_scale, int16_t *out_dev,
                        void *stream);
int b200vec_power_stats(b200vec_ctx *ctx, const void *in_dev, int64_t n, float pre_scale, double *stats_host,
                        void *stream);

/* ---- MAT v5 payload (save_vector, utils.py:659-670 -> scipy.io.savemat) --------------------
 * MAT v5 stores a complex array as two planes: re_out[i] = Re in[i], im_out[i] = Im in[i]
 * (float32).  16 B per sample of HBM traffic; all three buffers 16-byte aligned. */
int b200vec_split_planes(b200vec_ctx *ctx, const void *in_dev, int64_t n, float *re_out_dev,
                         float *im_out_dev, void *stream);

/* ---- the steps either side of the path -------------------------
 * box_energy        : detect_packet_bounds' smoothing (utils.py:813-816): smoothed[i] =
 *                     np.convolve(np.abs(x)**2, np.ones(window)/window, 'same')[i] in float64, and
 *                     *max_out_dev = smoothed.max().  window <= 4096 (the reference uses fs // 1e6).
 * threshold_bounds  : first_last_host[0/1] = first / last i with smoothed[i] >= threshold
 *                     (utils.py:820-825); (n, -1) when none.  Synchronises the stream.
 * median2_rows      : scipy.ndimage.median_filter(S, size=(2, 1)) as plot_spectrogram applies it to
 *                     the dB map (utils.py:435-440): out[r][c] = max(S[r-1][c], S[r][c]), row 0 kept.
 *                     Element strides are explicit so both spectrogram layouts work; not in place. */
int b200vec_box_energy(b200vec_ctx *ctx, const void *in_dev, int64_t n, int window, double *smoothed_out_dev,
                       double *max_out_dev, void *stream);
/* Exact order statistics of a float64 device array: values_host[i] = element at 0-based position ranks_host[i]
 * of the ascending sort (n_ranks <= 4) — np.median of the smoothed energies' first tenth (utils.py:818) takes
 * the two middle ones.  One cooperative launch (radix select on 64-bit keys).  Synchronises the stream. */
int b200vec_select_f64(b200vec_ctx *ctx, const double *data_dev, int64_t n, const int64_t *ranks_host, int n_ranks,
                       double *values_host, void *stream);
int b200vec_threshold_bounds(b200vec_ctx *ctx, const double *smoothed_dev, int64_t n, double threshold,
                             int64_t *first_last_host, void *stream);
int b200vec_median2_rows(b200vec_ctx *ctx, const float *in_dev, int64_t rows, int64_t cols, int64_t in_row_stride,
                         int64_t in_col_stride, float *out_dev, int64_t out_row_stride, int64_t out_col_stride,
                         void *stream);

/* ---- allocation-free variant (no counterpart in the reference) ----
 * power_stats_static : the level metrics of b200vec_power_stats for callers that may not allocate (code under
 *                      CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It takes no
 *                      context and calls no cudaMalloc / cudaFree / allocating routine: the scratch of its two-stage
 *                      reduction is a file-scope `static __device__` array of 256 KiB + 256 B that every GPU
 *                      loading the library holds until exit (one partial per 65 536 samples, 16 384 partials).
 *                      LIMIT: n <= 2^30 samples per call; larger n is refused with an error, nothing is launched.
 *                      stats_out_dev (DEVICE, 2 doubles, caller-owned): [0] = max |y|^2, [1] = sum of
 *                      fl32(fl32(|y|)^2) in fp64, y = fl(x * pre_scale); the sum is taken in a fixed order, so
 *                      it is reproducible run to run and equals b200vec_power_stats' to fp64 rounding (~1e-15
 *                      relative), the max bit for bit.  Queued on `stream`, does not synchronise.  The first call
 *                      on a device resolves the array's device address (cudaGetSymbolAddress): make it before
 *                      any capture.  All calls on one GPU share the array: calls that overlap on the GPU (two
 *                      streams, two threads) race; serialise them. */
int b200vec_power_stats_static(int device, const void *in_dev, int64_t n, float pre_scale, double *stats_out_dev,
                               void *stream);

#ifdef __cplusplus
}
#endif
#endif /* B200VEC_H */
