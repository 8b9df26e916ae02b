// fused1024.cuh — the BASELINE step in one kernel: packet placement (+ scale) AND the
// nfft = 1024 spectrogram of the freshly built samples, without the vector ever being re-read.
//
// Replaces, for the full-rate analysis shape (hann 1024 / hop <= 512 / nfft 1024, two-sided,
// vector_analyzer/spectrogram_analysis.py:35-49):
//   unified_gui.py:1712,1757-1769,1781 (np.zeros, `+=` loop, divide by max)  -> P2
//   scipy.signal.spectrogram + fftshift                                     -> P3..P6
//
// One persistent 16-warp CTA per SM; a tile = 16 frames = 8192 new samples (+512 overlap).
//   P1  warp 0 derives, per packet, the instance range overlapping the NEXT tile (overlapped
//       with the current tile's arithmetic; two 64-bit floor divisions per packet).
//   P2  FUSED: every thread sums its 17 samples over the overlapping (packet, instance) pairs in
//       the reference's order (16-byte loads from the L2-resident bank), scales, streams the
//       first 8192 samples of the span to the vector in HBM and parks all 8704 in the smem stage.
//       PLAIN: the span arrives by TMA bulk copy (cp.async.bulk + mbarrier), issued one tile ahead.
//   P3  warp w = frame w: lane l loads x[l + 32 j] from the stage, first 32-point transform with
//       the window folded in (fft32.cuh), writes B_l[q] to the exchange buffer [frame][q][l].
//   P5  thread -> (frame f = lane & 15, q = 2 warp + (lane >> 4)): reads B_l[q] for all l
//       (conflict-free: frame pitch 1026), second transform with the W1024 twiddles folded in.
//   P6  |X|^2 (sqrt(scale) was folded into the window taps) goes STRAIGHT to HBM: for a fixed
//       s the 16 lanes of a half-warp hold the same output row and 16 consecutive frames, i.e.
//       one contiguous 64-byte run per half-warp — the reference's fftshifted [1024, frames]
//       layout needs no smem transpose and no write-out phase.
// Two __syncthreads per tile (stage full; exchange full).  HBM traffic: 8 B/sample vector store
// + 8 B/sample spectrogram store (FUSED), nothing read.
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "place.cuh"
#include "fft32.cuh"
#include "stft1024.cuh"      // TMA / mbarrier primitives, get_tw1024

namespace b200vec {

constexpr int kFMaxHop = 512;
constexpr int kExPitch = 1026;                                   // float2 per frame in the exchange buffer
constexpr int kTwPitch = 33;                                     // float2 per q row of the twiddle table: adjacent q rows hit different banks

// Shared-memory plan for W warps = W frames per tile.  W = 16: one 512-thread CTA per SM
// (209 664 B); W = 8: two 256-thread CTAs per SM (111 232 B each) whose phases interleave.
template <int W>
struct FL {
    static constexpr int kWarps = W;
    static constexpr int kThreads = 32 * W;
    static constexpr int kSpan = (W - 1) * kFMaxHop + 1024;          // samples per tile at hop 512
    static constexpr int kOwn = W * kFMaxHop;                        // samples stored per tile (FUSED)
    static constexpr int kVec = (kSpan / 2 + kThreads - 1) / kThreads;   // float4 per thread (9)
    static constexpr int kLastLive = kSpan / 2 - (kVec - 1) * kThreads;  // threads populated in the last round
    static constexpr int kQPerWarp = 32 / W;                         // P5: q values per warp
    static constexpr size_t kOffStage = 0;
    static constexpr size_t kOffExch = kOffStage + (size_t)kSpan * sizeof(float2);
    static constexpr size_t kOffTw = kOffExch + (size_t)W * kExPitch * sizeof(float2);
    static constexpr size_t kOffWin = kOffTw + 32 * kTwPitch * sizeof(float2);    // window taps * sqrt(scale): [8][32 lanes] float4, in order of use
    static constexpr size_t kOffMisc = kOffWin + 1024 * sizeof(float);
    static constexpr size_t kZeroBytes = W == 16 ? 8192 : 0;        // zero source for empty tiles (bulk stores); no room at W = 8
    static constexpr size_t kOffZero = kOffMisc + 512;
    static constexpr size_t kSmem = kOffZero + kZeroBytes;
};
constexpr int kFWarpsDefault = 16;

struct FMisc {                       // lives at kOffMisc
    long long rel[B200VEC_MAX_PACKETS];   // packet index of stage sample 0 for the first overlapping instance
    int cnt[B200VEC_MAX_PACKETS];         // overlapping instances (<= 0: none)
    int total;                            // overlapping (packet, instance) pairs over the span (saturating)
    int first;                            // first packet with cnt > 0
    unsigned long long mbar;
};
static_assert(sizeof(FMisc) <= 512, "FMisc must fit its smem slot");

struct FArgs {
    alignas(64) CUtensorMap tmap_z; // empty tiles: same tensor, box = W frames x 128 rows (8 KB of zeros at W = 16)
    int ztma;                       // empty tiles leave by bulk stores from a zeroed shared-memory block (W = 16, tensor map available)
    alignas(64) CUtensorMap tmap;   // TS kernels: float32 [1024 rows][n_frames], row pitch out_ld, box = W frames x 256 rows
    PlaceArgs place;             // FUSED: descriptors + the vector slab [lo, hi) this launch STORES
    long long soff[B200VEC_MAX_PACKETS];   // FUSED: sample offset from a packet's raw bank copies to its pre-scaled copies (0: unscaled build)
    const float2 *zeros;         // FUSED: >= kSpan zero samples in HBM/L2 (TMA source for the uncovered part of a span)
    float2 *vec;                 // FUSED: vector out, vec[0] = sample place.lo.  PLAIN: vector in, vec[0] = sample in_origin
    const unsigned long long *maxsq;
    int mode;                    // FUSED: B200VEC_BUILD_STORE or B200VEC_BUILD_SCALE
    int hop;                     // PLAIN: any even hop <= 512; FUSED: 512
    long long in_origin;
    const float *window;         // 1024 taps
    const float2 *tw;            // [q][l] = exp(-2*pi*i*q*l/1024)
    float scale;
    long long frame_lo, n_frames;
    float *out;
    long long out_ld;
    int *flag;
    long long n_tiles;
    int dbg;                     // B200VEC_FUSED_V bit mask, timing experiments only
};

// Pre-scaled bank copies for a normalised build: scaled[i] = fl(raw[i] * fl(1 / fl(sqrt(maxsq)))),
// pads included (0 * r = 0).  One launch for every distinct slot a build references.
struct BankScaleArgs {
    float2 *store[B200VEC_MAX_PACKETS];     // raw copies A|B; scaled copies follow at + count
    long long count[B200VEC_MAX_PACKETS];   // samples in the raw pair (2 * per)
    int n;
    const unsigned long long *maxsq;
};
__global__ void __launch_bounds__(256)
bank_scale_kernel(const __grid_constant__ BankScaleArgs sa) {
    const double m2 = __longlong_as_double((long long)*sa.maxsq);
    const float scl = m2 > 0.0 ? 1.0f / (float)sqrt(m2) : 1.0f;
    const float4 *src = reinterpret_cast<const float4 *>(sa.store[blockIdx.y]);
    const long long n4 = sa.count[blockIdx.y] / 2;        // per is even: whole float4s
    float4 *dst = reinterpret_cast<float4 *>(sa.store[blockIdx.y] + sa.count[blockIdx.y]);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
        float4 v = src[i];
        v.x *= scl; v.y *= scl; v.z *= scl; v.w *= scl;
        dst[i] = v;
    }
}

// P1: instance ranges of every packet over the span of tile t (warp 0, one lane per packet)
template <int W>
__device__ __forceinline__ void f_setup_tile(const FArgs &a, long long t, int lane, FMisc *m) {
    const long long s0 = (a.frame_lo + W * t) * kFMaxHop;
    int c32 = 0;
    if (lane < a.place.np) {
        const PlaceDesc &d = a.place.d[lane];
        long long klo = floor_div(s0 - d.len - d.start, d.period) + 1;
        long long khi = floor_div(s0 + FL<W>::kSpan - 1 - d.start, d.period);
        if (klo < 0) klo = 0;
        if (khi > d.count - 1) khi = d.count - 1;
        const long long c = khi - klo + 1;
        c32 = c > 0 ? (int)(c < 0x00ffffff ? c : 0x00ffffff) : 0;
        m->cnt[lane] = c32;
        m->rel[lane] = s0 - (d.start + klo * d.period);
    }
    const unsigned anyb = __ballot_sync(0xffffffffu, c32 > 0);
    const int total = __reduce_add_sync(0xffffffffu, c32);
    if (lane == 0) { m->total = total; m->first = anyb ? __ffs(anyb) - 1 : 0; }
}

__device__ __forceinline__ void tma_bulk_g2s_chunked(unsigned dst, const char *src, unsigned bytes, unsigned mbar) {
    for (unsigned off = 0; off < bytes; off += 16384) {
        const unsigned n = (bytes - off < 16384u) ? (bytes - off) : 16384u;
        tma_bulk_g2s(dst + off, src + off, n, mbar);
    }
}
__device__ __forceinline__ void tma_bulk_s2g(void *dst, unsigned src, unsigned bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// PLAIN: the span of tile t arrives from the vector in HBM.
template <int W>
__device__ __forceinline__ void f_issue_tma(const FArgs &a, long long t, unsigned stage_u32, unsigned mbar) {
    const long long f0 = t * W;
    const int nfr = (int)((a.n_frames - f0 < W) ? (a.n_frames - f0) : W);
    const unsigned bytes = (unsigned)(((nfr - 1) * a.hop + 1024) * sizeof(float2));
    const char *src = reinterpret_cast<const char *>(a.vec + ((a.frame_lo + f0) * a.hop - a.in_origin));
    mbar_expect_tx(mbar, bytes);
    tma_bulk_g2s_chunked(stage_u32, src, bytes, mbar);
}

// FUSED, span covered by exactly ONE (packet, instance): the span is that packet's pre-scaled bank
// copy (fl(y * r), the very value `vector / max` has where nothing else was added) plus zeros, so
// it is assembled entirely by bulk copies — no registers, no issue slots, one tile ahead.  The
// piece boundaries are rounded outwards to even samples (16 bytes); the extra sample is one of
// the bank's zero pads.
template <int W>
__device__ __forceinline__ void f_issue_single(const FArgs &a, const FMisc *m, unsigned stage_u32, unsigned mbar) {
    constexpr int span = FL<W>::kSpan;
    const int p = m->first;
    const PlaceDesc &d = a.place.d[p];
    const int rel = (int)m->rel[p];
    const int lo = rel < 0 ? ((-rel) & ~1) : 0;
    const int end = d.len - rel;                                   // > 0
    const int hi = end >= span ? span : ((end + 1) & ~1);
    const float2 *base = ((rel & 1) ? d.b : d.a) + a.soff[p] + rel;   // base + even s is 16-byte aligned
    mbar_expect_tx(mbar, (unsigned)(span * sizeof(float2)));
    if (lo > 0) tma_bulk_g2s_chunked(stage_u32, reinterpret_cast<const char *>(a.zeros), (unsigned)lo * 8u, mbar);
    tma_bulk_g2s_chunked(stage_u32 + (unsigned)lo * 8u, reinterpret_cast<const char *>(base + lo), (unsigned)(hi - lo) * 8u, mbar);
    if (hi < span) tma_bulk_g2s_chunked(stage_u32 + (unsigned)hi * 8u, reinterpret_cast<const char *>(a.zeros), (unsigned)(span - hi) * 8u, mbar);
}

#ifdef B2_PIPE_TIMING
__device__ long long g_pipe_prof[320][16][8];
__device__ long long g_pipe_cta[320][4];
#define PT_DECL long long pt_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}; long long pt_last = clock64(); const long long pt_t0 = pt_last;
#define PT(k) { const long long pt_now = clock64(); pt_acc[k] += pt_now - pt_last; pt_last = pt_now; }
#define PT_FLUSH if (lane == 0) { for (int k = 0; k < 8; ++k) g_pipe_prof[blockIdx.x][warp][k] = pt_acc[k]; if (warp == 0) { unsigned sm_; asm volatile("mov.u32 %0, %%smid;" : "=r"(sm_)); g_pipe_cta[blockIdx.x][0] = sm_; g_pipe_cta[blockIdx.x][1] = pt_t0; g_pipe_cta[blockIdx.x][2] = clock64(); } }
#else
#define PT_DECL
#define PT(k)
#define PT_FLUSH
#endif

template <bool FUSED, int W, bool TS>
__global__ void __launch_bounds__(32 * W, 16 / W)
tile1024_kernel(const __grid_constant__ FArgs a) {
    using L = FL<W>;
    extern __shared__ __align__(128) unsigned char f_smem[];
    float2 *stage = reinterpret_cast<float2 *>(f_smem + L::kOffStage);
    float2 *exch = reinterpret_cast<float2 *>(f_smem + L::kOffExch);
    float2 *tw = reinterpret_cast<float2 *>(f_smem + L::kOffTw);
    FMisc *misc = reinterpret_cast<FMisc *>(f_smem + L::kOffMisc);
    const unsigned stage_u32 = smem_u32(stage);
    const unsigned mbar = smem_u32(&misc->mbar);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 1024; i += L::kThreads) tw[(i >> 5) * kTwPitch + (i & 31)] = __ldg(a.tw + i);
    float4 *win4 = reinterpret_cast<float4 *>(f_smem + L::kOffWin);
    const bool ztma = FUSED && W == 16 && a.ztma;
    if (ztma) {
        float4 *z = reinterpret_cast<float4 *>(f_smem + L::kOffZero);
        for (int i = tid; i < (int)(L::kZeroBytes / 16); i += L::kThreads) z[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    {
        const float rs = sqrtf(a.scale);
        float *wf = reinterpret_cast<float *>(win4);
        for (int i = tid; i < 1024; i += L::kThreads) {     // i = (k * 32 + lane) * 4 + c  ->  tap j = brev5(4 k + c) of lane
            const int c = i & 3, ln = (i >> 2) & 31, k = i >> 7;
            wf[i] = __ldg(a.window + ln + 32 * brev5(4 * k + c)) * rs;
        }
    }
    float scl = 1.0f;
    if (FUSED && a.mode == B200VEC_BUILD_SCALE) {
        const double m2 = __longlong_as_double((long long)*a.maxsq);
        if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);  // numpy's complex64 / float32: multiply by the fp32 reciprocal
    }
    const int hop = FUSED ? kFMaxHop : a.hop;
    long long t = blockIdx.x;
    const bool use_single = FUSED && !(a.dbg & 16);              // B200VEC_FUSED_V=16: force the register path (A/B)
    if (tid == 0) mbar_init(mbar, 1);
    if (FUSED) {
        if (warp == 0 && t < a.n_tiles) {
            f_setup_tile<W>(a, t, lane, misc);
            __syncwarp();
            if (lane == 0 && use_single && misc->total == 1) f_issue_single<W>(a, misc, stage_u32, mbar);
        }
    } else if (tid == 0) {
        if (t < a.n_tiles) f_issue_tma<W>(a, t, stage_u32, mbar);
    }
    __syncthreads();

    unsigned anybits = 0, parity = 0;
    PT_DECL
    const int f5 = lane & (W - 1), q5 = L::kQPerWarp * warp + lane / W;      // P5/P6 role of this thread
    for (; t < a.n_tiles; t += gridDim.x) {
        const long long f0 = t * W;                        // first frame of the tile, relative to frame_lo
        const int nfr = (int)((a.n_frames - f0 < W) ? (a.n_frames - f0) : W);   // may be <= 0 (FUSED tail)
        bool live = true;                                        // block-uniform: the tile holds signal
        bool single = false;                                     // block-uniform: span assembled by bulk copies
        int rem = 0;
        float2 *dst = nullptr;
        // ------------------------------------------------------------------ P2
        if (FUSED) {
            const long long s0 = (a.frame_lo + f0) * kFMaxHop;
            const int total = misc->total;
            live = total != 0;
            single = use_single && total == 1;
            const long long rem64 = a.place.hi - s0;             // samples of this tile inside the slab
            rem = rem64 > L::kOwn ? L::kOwn : (rem64 < 0 ? 0 : (int)rem64);
            dst = a.vec + (s0 - a.place.lo);
            if (single) {                                        // the span was assembled by bulk copies (f_issue_single)
                mbar_wait(mbar, parity);
                parity ^= 1;
            } else if (ztma && !live) {
                // Empty span: vector samples and PSD columns are zeros.  One thread streams them out of
                // the zeroed block with bulk stores — no registers, no store instructions, no waiting.
                if (tid == 0) {
                    const unsigned zsrc = smem_u32(f_smem + L::kOffZero);
                    const unsigned bytes = (unsigned)(rem & ~1) * 8u;
                    for (unsigned off = 0; off < bytes; off += 8192u) {
                        const unsigned n = bytes - off < 8192u ? bytes - off : 8192u;
                        tma_bulk_s2g(reinterpret_cast<char *>(dst) + off, zsrc, n);
                    }
                    if (rem & 1) __stcs(dst + rem - 1, make_float2(0.f, 0.f));
                    if (nfr > 0) {
#pragma unroll
                        for (int r = 0; r < 8; ++r)
                            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                                         ::"l"(&a.tmap_z), "r"((int)f0), "r"(128 * r), "r"(zsrc) : "memory");
                    }
                    tma_store_commit();
                }
            } else {
            float4 acc[L::kVec];
#pragma unroll
            for (int j = 0; j < L::kVec; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) {
                for (int p = 0; p < a.place.np; ++p) {
                    const int cnt = misc->cnt[p];
                    if (cnt <= 0) continue;
                    const PlaceDesc &d = a.place.d[p];
                    const int len = d.len;
                    long long r64 = misc->rel[p];
                    for (int i = 0; i < cnt; ++i, r64 -= d.period) {
                        const int rel = (int)r64;
                        const float2 *base = ((rel & 1) ? d.b : d.a) + rel;
                        if (rel >= 0 && rel + L::kSpan <= len) {
#pragma unroll
                            for (int j = 0; j < L::kVec; ++j) {
                                if (j == L::kVec - 1 && tid >= (L::kLastLive)) break;
                                const float4 v = __ldg(reinterpret_cast<const float4 *>(base + 2 * (tid + L::kThreads * j)));
                                acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < L::kVec; ++j) {
                                if (j == L::kVec - 1 && tid >= (L::kLastLive)) break;
                                const int s = 2 * (tid + L::kThreads * j);
                                const int n = s + rel;
                                if (n >= -1 && n < len) {        // the bank's zero pads cover n = -1 and n + 1 = len
                                    const float4 v = __ldg(reinterpret_cast<const float4 *>(base + s));
                                    acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
                                }
                            }
                        }
                    }
                }
            }
            float4 *st4 = reinterpret_cast<float4 *>(stage);
#pragma unroll
            for (int j = 0; j < L::kVec; ++j) {
                if (j == L::kVec - 1 && tid >= (L::kLastLive)) break;
                float4 v = acc[j];
                v.x *= scl; v.y *= scl; v.z *= scl; v.w *= scl;
                const int s = 2 * (tid + L::kThreads * j);
                if (live) st4[tid + L::kThreads * j] = v;
                if (j < L::kVec - 1) {             // s < L::kOwn: this tile owns the sample
                    if (s + 1 < rem) __stcs(reinterpret_cast<float4 *>(dst + s), v);
                    else if (s < rem) __stcs(dst + s, make_float2(v.x, v.y));
                }
            }
            }
        } else {
            mbar_wait(mbar, parity);
            parity ^= 1;
        }
        PT(0)
        if (TS && tid == 0) tma_store_wait_read();                 // the last PSD tile has left the exchange memory
        __syncthreads();                                          // A: stage full (and last tile's exchange reads done)
        PT(1)
        if (FUSED && single && tid == 0 && rem > 0) {   // vector store of a bulk-assembled span: smem -> HBM bulk copy
            if (rem > 1) tma_bulk_s2g(dst, stage_u32, (unsigned)(rem & ~1) * 8u);
            tma_store_commit();
            if (rem & 1) __stcs(dst + rem - 1, stage[rem - 1]);
        }
        // ------------------------------------------------------------------ P3 + P4
        if (live && warp < nfr) {
            cf32 v[32];
            const float2 *sp = stage + warp * hop + lane;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float2 x = sp[32 * brev5(i)];
                v[i].x = x.x; v[i].y = x.y;
            }
            const float4 *wl = win4 + lane;
            fft32_windowed_q(v, [wl](int k) { const float4 t = wl[32 * k]; return f32x4{t.x, t.y, t.z, t.w}; });
            PT(2)
            float2 *ex = exch + warp * kExPitch + lane;
#pragma unroll
            for (int q = 0; q < 32; ++q) ex[32 * q] = make_float2(v[q].x, v[q].y);
        }
        if (FUSED && warp == 0) {
            if (t + gridDim.x < a.n_tiles) f_setup_tile<W>(a, t + gridDim.x, lane, misc);   // P1 (next tile)
            if (lane == 0 && single) tma_store_wait_read();      // the vector bulk store has finished reading the stage
        }
        PT(3)
        __syncthreads();                                          // B: exchange full, stage free
        PT(4)
        if (tid == 0 && t + gridDim.x < a.n_tiles) {
            if (!FUSED) f_issue_tma<W>(a, t + gridDim.x, stage_u32, mbar);
            else if (use_single && misc->total == 1) f_issue_single<W>(a, misc, stage_u32, mbar);
        }
        // ------------------------------------------------------------------ P5 + P6
        if (TS && live) {
            // The PSD tile [1024 rows][W frames] is staged in the exchange memory (all of it is in
            // registers after the loads) and leaves by 2-D bulk tensor stores: no store instructions,
            // no address arithmetic, and the drain overlaps the next tile's placement and transform.
            const bool act = f5 < nfr;
            cf32 v[32];
            if (act) {
                const float4 *ex4 = reinterpret_cast<const float4 *>(exch + f5 * kExPitch + q5 * 32);
#pragma unroll
                for (int l = 0; l < 32; l += 2) {
                    const float4 x = ex4[l / 2];
                    v[brev5(l)].x = x.x; v[brev5(l)].y = x.y;
                    v[brev5(l + 1)].x = x.z; v[brev5(l + 1)].y = x.w;
                }
            }
            __syncthreads();                                      // C: exchange consumed
            if (act) {
                const float2 *tq = tw + q5 * kTwPitch;
                fft32_twiddled(v, [tq](int l) { const float2 w = tq[l]; return cf32{w.x, w.y}; });
                PT(5)
                float *pt = reinterpret_cast<float *>(exch) + q5 * W + f5;
#pragma unroll
                for (int sp = 0; sp < 32; ++sp) {                 // output row q + 32 sp holds bin q + 32 s, s = (sp + 16) & 31
                    const int s = (sp + 16) & 31;
                    const float p = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
                    anybits |= __float_as_uint(p);
                    pt[32 * W * sp] = p;
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();                                      // D: PSD tile complete
            if (tid == 0) {
                const unsigned src = smem_u32(exch);
#pragma unroll
                for (int r = 0; r < 4; ++r)
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                                 ::"l"(&a.tmap), "r"((int)f0), "r"(256 * r), "r"(src + (unsigned)(256 * r * W * 4)) : "memory");
                tma_store_commit();
            }
        } else if (f5 < nfr) {
            // row q + 32 sp, sp = 8 hi + lo: four 64-bit row-group bases, 32-bit offsets inside a group
            float *o = a.out + (long long)q5 * a.out_ld + (f0 + f5);
            const unsigned step = 32u * (unsigned)a.out_ld;     // host guarantees 7 * 32 * out_ld < 2^32
            float *og[4];
#pragma unroll
            for (int h = 0; h < 4; ++h) og[h] = o + (long long)h * 256 * a.out_ld;
            if (live) {
                cf32 v[32];
                const float4 *ex4 = reinterpret_cast<const float4 *>(exch + f5 * kExPitch + q5 * 32);
#pragma unroll
                for (int l = 0; l < 32; l += 2) {
                    const float4 x = ex4[l / 2];
                    v[brev5(l)].x = x.x; v[brev5(l)].y = x.y;
                    v[brev5(l + 1)].x = x.z; v[brev5(l + 1)].y = x.w;
                }
                const float2 *tq = tw + q5 * kTwPitch;
                fft32_twiddled(v, [tq](int l) { const float2 w = tq[l]; return cf32{w.x, w.y}; });
                PT(5)
#pragma unroll
                for (int sp = 0; sp < 32; ++sp) {                 // output row q + 32 sp holds bin q + 32 s, s = (sp + 16) & 31
                    const int s = (sp + 16) & 31;
                    const float p = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
                    anybits |= __float_as_uint(p);                // p >= 0 or NaN: non-zero bits <=> p != 0
                    __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, p);
                }
            } else if (!ztma) {
#pragma unroll
                for (int sp = 0; sp < 32; ++sp) __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, 0.0f);
            }
        }
        PT(6)
    }
    PT_FLUSH
    if (tid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (a.flag && __syncthreads_or(anybits != 0) && tid == 0) atomicOr(a.flag, 1);
}

// W = frames per tile: 16 -> one 512-thread CTA per SM (the product path).  -DB2_AB builds also carry
// W = 8 (two 256-thread CTAs per SM) and the bulk-tensor-store (TS / ZTMA) variants, selected by
// B200VEC_FUSED_W / _TS / _ZTMA; all of them measured slower (profiles/r1_fused_ablation.txt).
inline int fused_w() {
    const char *e = ab_env("B200VEC_FUSED_W");
    if (e && e[0] == '8') return 8;
    return kFWarpsDefault;
}

typedef CUresult (*TensorMapEncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                      const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                      CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// 2-D tensor map of the spectrogram slab this launch writes: dim0 = frames (contiguous), dim1 = 1024
// rows at pitch out_ld; box = w frames x `rows` rows.  Columns past n_frames are clipped by the hardware,
// so a partial last tile needs no special case.  Returns false when the layout cannot be described
// (pitch or base not 16-byte aligned) — the caller then uses the direct-store kernel.
inline bool make_spec_tmap(CUtensorMap *tm, float *out, long long out_ld, long long n_frames, int w, int rows) {
    static TensorMapEncodeFn enc = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult qr;
        void *fn = nullptr;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qr) == cudaSuccess &&
            qr == cudaDriverEntryPointSuccess)
            enc = reinterpret_cast<TensorMapEncodeFn>(fn);
    }
    if (!enc || n_frames <= 0 || (out_ld & 3) || (reinterpret_cast<uintptr_t>(out) & 15)) return false;
    cuuint64_t dims[2] = {(cuuint64_t)n_frames, 1024};
    cuuint64_t strides[1] = {(cuuint64_t)out_ld * 4};
    cuuint32_t box[2] = {(cuuint32_t)w, (cuuint32_t)rows};
    cuuint32_t es[2] = {1, 1};
    return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <bool FUSED>
inline int launch_tile1024(b200vec_ctx *ctx, FArgs &a, long long n_frames, long long n_vec, cudaStream_t st) {
    const int w = fused_w();
    { const char *e = ab_env("B200VEC_FUSED_V"); a.dbg = e ? atoi(e) : 0; }
    const long long own = (long long)w * kFMaxHop;
    const long long t_frames = (n_frames + w - 1) / w, t_vec = (n_vec + own - 1) / own;
    a.n_tiles = t_frames > t_vec ? t_frames : t_vec;
    if (a.n_tiles == 0) return 0;
    const long long cap = (long long)ctx->sm_count * (16 / w);
    const long long grid = a.n_tiles < cap ? a.n_tiles : cap;
    a.ztma = 0;
#ifdef B2_AB
    const char *ts_env = ab_env("B200VEC_FUSED_TS");     // "1": PSD tiles leave by bulk tensor stores (measured slower: 360 vs 298 us)
    const bool ts = ts_env && ts_env[0] == '1' && make_spec_tmap(&a.tmap, a.out, a.out_ld, n_frames, w, 256);
    const char *z_env = ab_env("B200VEC_FUSED_ZTMA");    // "1": empty tiles by bulk stores from a zeroed smem block (measured 283 vs 274 us)
    a.ztma = FUSED && w == 16 && z_env && z_env[0] == '1' && make_spec_tmap(&a.tmap_z, a.out, a.out_ld, n_frames, w, 128);
    if (w == 8) {
        if (ts) tile1024_kernel<FUSED, 8, true><<<(unsigned)grid, FL<8>::kThreads, FL<8>::kSmem, st>>>(a);
        else tile1024_kernel<FUSED, 8, false><<<(unsigned)grid, FL<8>::kThreads, FL<8>::kSmem, st>>>(a);
        B2_CUDA(cudaGetLastError());
        return 0;
    }
    if (ts) {
        tile1024_kernel<FUSED, 16, true><<<(unsigned)grid, FL<16>::kThreads, FL<16>::kSmem, st>>>(a);
        B2_CUDA(cudaGetLastError());
        return 0;
    }
#endif
    tile1024_kernel<FUSED, 16, false><<<(unsigned)grid, FL<16>::kThreads, FL<16>::kSmem, st>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

#ifdef B2_AB
bool pipe1024_enabled();                                                                   // pipe1024.cuh
int launch_pipe1024_plain(b200vec_ctx *ctx, FArgs &a, long long n_frames, cudaStream_t st);   // pipe1024.cuh
#endif

inline int launch_tile1024_plain(b200vec_ctx *ctx, const S1Args &s, cudaStream_t st) {
    FArgs a;
    memset(&a, 0, sizeof(a));
    a.vec = const_cast<float2 *>(s.in);
    a.hop = s.hop; a.in_origin = s.in_origin;
    a.window = s.window; a.tw = s.tw; a.scale = s.scale;
    a.frame_lo = s.frame_lo; a.n_frames = s.n_frames; a.out = s.out; a.out_ld = s.out_ld; a.flag = s.flag;
#ifdef B2_AB
    if (pipe1024_enabled()) return launch_pipe1024_plain(ctx, a, s.n_frames, st);
#endif
    return launch_tile1024<false>(ctx, a, s.n_frames, 0, st);
}

template <bool FUSED, int W, bool TS>
inline cudaError_t fused1024_attr() {
    // Without the carve-out hint the driver sizes shared memory for ONE resident CTA, and the two
    // 256-thread CTAs per SM of the W = 8 layout silently run one after the other.
    cudaError_t e = cudaFuncSetAttribute(tile1024_kernel<FUSED, W, TS>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(tile1024_kernel<FUSED, W, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FL<W>::kSmem);
}
inline int fused1024_setup() {
    B2_CUDA((fused1024_attr<true, 16, false>())); B2_CUDA((fused1024_attr<false, 16, false>()));
#ifdef B2_AB
    B2_CUDA((fused1024_attr<true, 16, true>()));  B2_CUDA((fused1024_attr<false, 16, true>()));
    B2_CUDA((fused1024_attr<true, 8, true>()));   B2_CUDA((fused1024_attr<true, 8, false>()));
    B2_CUDA((fused1024_attr<false, 8, true>()));  B2_CUDA((fused1024_attr<false, 8, false>()));
#endif
    return 0;
}

}  // namespace b200vec
