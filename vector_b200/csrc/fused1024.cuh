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
#include "common.cuh"
#include "place.cuh"
#include "fft32.cuh"
#include "stft1024.cuh"      // TMA / mbarrier primitives, get_tw1024

namespace b200vec {

constexpr int kFWarps = 16;
constexpr int kFThreads = 32 * kFWarps;
constexpr int kFMaxHop = 512;
constexpr int kFSpan = (kFWarps - 1) * kFMaxHop + 1024;          // 8704 samples per tile at hop 512
constexpr int kFOwn = kFWarps * kFMaxHop;                        // 8192 samples stored per tile (FUSED)
constexpr int kFVec = (kFSpan / 2 + kFThreads - 1) / kFThreads;  // 9 float4 per thread (last one half-populated)
constexpr int kExPitch = 1026;                                   // float2 per frame in the exchange buffer
constexpr size_t kFOffStage = 0;
constexpr size_t kFOffExch = kFOffStage + (size_t)kFSpan * sizeof(float2);                 // 69 632
constexpr size_t kFOffTw = kFOffExch + (size_t)kFWarps * kExPitch * sizeof(float2);       // +131 328
constexpr size_t kFOffMisc = kFOffTw + 1024 * sizeof(float2);                             // +8 192
constexpr size_t kFSmem = kFOffMisc + 512;                                                 // 209 664 B

struct FMisc {                       // lives at kFOffMisc
    long long rel[B200VEC_MAX_PACKETS];   // packet index of stage sample 0 for the first overlapping instance
    int cnt[B200VEC_MAX_PACKETS];         // overlapping instances (<= 0: none)
    int any;
    int pad;
    unsigned long long mbar;
};
static_assert(sizeof(FMisc) <= 512, "FMisc must fit its smem slot");

struct FArgs {
    PlaceArgs place;             // FUSED: descriptors + the vector slab [lo, hi) this launch STORES
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
};

// P1: instance ranges of every packet over the span of tile t (warp 0, one lane per packet)
__device__ __forceinline__ void f_setup_tile(const FArgs &a, long long t, int lane, FMisc *m) {
    const long long s0 = (a.frame_lo + kFWarps * t) * kFMaxHop;
    bool has = false;
    if (lane < a.place.np) {
        const PlaceDesc &d = a.place.d[lane];
        long long klo = floor_div(s0 - d.len - d.start, d.period) + 1;
        long long khi = floor_div(s0 + kFSpan - 1 - d.start, d.period);
        if (klo < 0) klo = 0;
        if (khi > d.count - 1) khi = d.count - 1;
        const long long c = khi - klo + 1;
        m->cnt[lane] = c > 0 ? (int)(c < 0x7fffffff ? c : 0x7fffffff) : 0;
        m->rel[lane] = s0 - (d.start + klo * d.period);
        has = c > 0;
    }
    const unsigned anyb = __ballot_sync(0xffffffffu, has);
    if (lane == 0) m->any = anyb != 0;
}

template <bool FUSED>
__device__ __forceinline__ void f_issue_tma(const FArgs &a, long long t, unsigned stage_u32, unsigned mbar) {
    const long long f0 = t * kFWarps;
    const int nfr = (int)((a.n_frames - f0 < kFWarps) ? (a.n_frames - f0) : kFWarps);
    const unsigned bytes = (unsigned)(((nfr - 1) * a.hop + 1024) * sizeof(float2));
    const char *src = reinterpret_cast<const char *>(a.vec + ((a.frame_lo + f0) * a.hop - a.in_origin));
    mbar_expect_tx(mbar, bytes);
    for (unsigned off = 0; off < bytes; off += 16384) {
        const unsigned n = (bytes - off < 16384u) ? (bytes - off) : 16384u;
        tma_bulk_g2s(stage_u32 + off, src + off, n, mbar);
    }
}

template <bool FUSED>
__global__ void __launch_bounds__(kFThreads, 1)
tile1024_kernel(const __grid_constant__ FArgs a) {
    extern __shared__ __align__(128) unsigned char f_smem[];
    float2 *stage = reinterpret_cast<float2 *>(f_smem + kFOffStage);
    float2 *exch = reinterpret_cast<float2 *>(f_smem + kFOffExch);
    float2 *tw = reinterpret_cast<float2 *>(f_smem + kFOffTw);
    FMisc *misc = reinterpret_cast<FMisc *>(f_smem + kFOffMisc);
    const unsigned stage_u32 = smem_u32(stage);
    const unsigned mbar = smem_u32(&misc->mbar);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < 1024; i += kFThreads) tw[i] = __ldg(a.tw + i);
    float wreg[32];                                   // this lane's window taps * sqrt(scale), resident
    {
        const float rs = sqrtf(a.scale);
#pragma unroll
        for (int j = 0; j < 32; ++j) wreg[j] = __ldg(a.window + lane + 32 * j) * rs;
    }
    float scl = 1.0f;
    if (FUSED && a.mode == B200VEC_BUILD_SCALE) {
        const double m2 = __longlong_as_double((long long)*a.maxsq);
        if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);  // numpy's complex64 / float32: multiply by the fp32 reciprocal
    }
    const int hop = FUSED ? kFMaxHop : a.hop;
    long long t = blockIdx.x;
    if (FUSED) {
        if (warp == 0 && t < a.n_tiles) f_setup_tile(a, t, lane, misc);
    } else if (tid == 0) {
        mbar_init(mbar, 1);
        if (t < a.n_tiles) f_issue_tma<FUSED>(a, t, stage_u32, mbar);
    }
    __syncthreads();

    unsigned anybits = 0, parity = 0;
    const int f5 = lane & 15, q5 = 2 * warp + (lane >> 4);      // P5/P6 role of this thread
    for (; t < a.n_tiles; t += gridDim.x) {
        const long long f0 = t * kFWarps;                        // first frame of the tile, relative to frame_lo
        const int nfr = (int)((a.n_frames - f0 < kFWarps) ? (a.n_frames - f0) : kFWarps);   // may be <= 0 (FUSED tail)
        bool live = true;                                        // block-uniform: the tile holds signal
        // ------------------------------------------------------------------ P2
        if (FUSED) {
            const long long s0 = (a.frame_lo + f0) * kFMaxHop;
            live = misc->any != 0;
            float4 acc[kFVec];
#pragma unroll
            for (int j = 0; j < kFVec; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
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
                        if (rel >= 0 && rel + kFSpan <= len) {
#pragma unroll
                            for (int j = 0; j < kFVec; ++j) {
                                if (j == kFVec - 1 && tid >= (kFSpan / 2 - (kFVec - 1) * kFThreads)) break;
                                const float4 v = __ldg(reinterpret_cast<const float4 *>(base + 2 * (tid + kFThreads * j)));
                                acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
                            }
                        } else {
#pragma unroll
                            for (int j = 0; j < kFVec; ++j) {
                                if (j == kFVec - 1 && tid >= (kFSpan / 2 - (kFVec - 1) * kFThreads)) break;
                                const int s = 2 * (tid + kFThreads * j);
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
            float2 *dst = a.vec + (s0 - a.place.lo);
            float4 *st4 = reinterpret_cast<float4 *>(stage);
            const long long rem64 = a.place.hi - s0;             // samples of this tile inside the slab
            const int rem = rem64 > kFOwn ? kFOwn : (rem64 < 0 ? 0 : (int)rem64);
#pragma unroll
            for (int j = 0; j < kFVec; ++j) {
                if (j == kFVec - 1 && tid >= (kFSpan / 2 - (kFVec - 1) * kFThreads)) break;
                float4 v = acc[j];
                v.x *= scl; v.y *= scl; v.z *= scl; v.w *= scl;
                const int s = 2 * (tid + kFThreads * j);
                if (live) st4[tid + kFThreads * j] = v;
                if (j < kFVec - 1) {                             // s < kFOwn: this tile owns the sample
                    if (s + 1 < rem) __stcs(reinterpret_cast<float4 *>(dst + s), v);
                    else if (s < rem) __stcs(dst + s, make_float2(v.x, v.y));
                }
            }
        } else {
            mbar_wait(mbar, parity);
            parity ^= 1;
        }
        __syncthreads();                                          // A: stage full (and last tile's exchange reads done)
        // ------------------------------------------------------------------ P3 + P4
        if (live && warp < nfr) {
            cf32 v[32];
            const float2 *sp = stage + warp * hop + lane;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const float2 x = sp[32 * brev5(i)];
                v[i].x = x.x; v[i].y = x.y;
            }
            fft32_windowed(v, wreg);
            float2 *ex = exch + warp * kExPitch + lane;
#pragma unroll
            for (int q = 0; q < 32; ++q) ex[32 * q] = make_float2(v[q].x, v[q].y);
        }
        if (FUSED && warp == 0 && t + gridDim.x < a.n_tiles) f_setup_tile(a, t + gridDim.x, lane, misc);   // P1 (next tile)
        __syncthreads();                                          // B: exchange full, stage free
        if (!FUSED && tid == 0 && t + gridDim.x < a.n_tiles) f_issue_tma<FUSED>(a, t + gridDim.x, stage_u32, mbar);
        // ------------------------------------------------------------------ P5 + P6
        if (f5 < nfr) {
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
                const float2 *tq = tw + q5 * 32;
                fft32_twiddled(v, [tq](int l) { const float2 w = tq[l]; return cf32{w.x, w.y}; });
#pragma unroll
                for (int sp = 0; sp < 32; ++sp) {                 // output row q + 32 sp holds bin q + 32 s, s = (sp + 16) & 31
                    const int s = (sp + 16) & 31;
                    const float p = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
                    anybits |= __float_as_uint(p);                // p >= 0 or NaN: non-zero bits <=> p != 0
                    __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, p);
                }
            } else {
#pragma unroll
                for (int sp = 0; sp < 32; ++sp) __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, 0.0f);
            }
        }
    }
    if (a.flag && __syncthreads_or(anybits != 0) && tid == 0) atomicOr(a.flag, 1);
}

inline int launch_tile1024_plain(b200vec_ctx *ctx, const S1Args &s, cudaStream_t st) {
    FArgs a;
    memset(&a, 0, sizeof(a));
    a.vec = const_cast<float2 *>(s.in);
    a.hop = s.hop; a.in_origin = s.in_origin;
    a.window = s.window; a.tw = s.tw; a.scale = s.scale;
    a.frame_lo = s.frame_lo; a.n_frames = s.n_frames; a.out = s.out; a.out_ld = s.out_ld; a.flag = s.flag;
    a.n_tiles = (s.n_frames + kFWarps - 1) / kFWarps;
    const long long grid = a.n_tiles < ctx->sm_count ? a.n_tiles : ctx->sm_count;
    tile1024_kernel<false><<<(unsigned)grid, kFThreads, kFSmem, st>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

inline int fused1024_setup() {
    B2_CUDA(cudaFuncSetAttribute(tile1024_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFSmem));
    B2_CUDA(cudaFuncSetAttribute(tile1024_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFSmem));
    return 0;
}

}  // namespace b200vec
