// stft1024.cuh — K4 fast path: nfft = 1024, one warp per frame, FFT held in registers.
//
// This is the BASELINE shape (cfg3 full-rate hann 1024/512, and create_spectrogram's heavy
// preview, which caps nfft at 1024 — reference utils.py:268-269).
//
// 1024 = 32 x 32.  Lane l of a warp loads x[l + 32 j] (j = 0..31: coalesced 256-byte rows,
// window applied on the way in), runs a 32-point FFT over j entirely in registers (radix-2
// Stockham, unrolled at compile time, literal twiddles), multiplies by W1024^(l*q) from a
// conflict-free [q][l] shared table, swaps (l, q) across the warp through a padded shared
// buffer, and runs a second in-register 32-point FFT over l: lane q ends up holding bins
// q + 32 s.  |X|^2 * scale goes into a swizzled shared [1024 rows][16 frames] tile; when the
// block's 16 frames are done every output row is written as one contiguous 64-byte run
// (rows are the reference's fftshifted [nfft, frames] layout, frames are the fast axis).
//
// Two kernels share that FFT core:
//   stft1024_tma_kernel    unit stride, nperseg = 1024, hop <= 512 (the full-rate BASELINE shape).
//       One persistent 16-warp CTA per SM.  The (15*hop + 1024)-sample input span of the NEXT
//       16-frame tile is fetched by one TMA bulk copy (cp.async.bulk -> mbarrier) while the
//       current tile is transformed, so every input sample crosses HBM->SM exactly once and no
//       warp ever waits on a global load; window taps live in registers.
//   stft1024_gather_kernel any stride / nperseg <= 1024 / hop (the decimated preview of
//       create_spectrogram, utils.py:192-195: a strided gather, the decimated copy is never
//       materialised).  2 CTAs of 8 warps per SM, plain predicated loads.
#pragma once
#include "common.cuh"
#include <utility>

namespace b200vec {

constexpr int kS1Threads = 256;
constexpr int kS1Warps = kS1Threads / 32;
constexpr int kS1Frames = 16;                                  // frames per tile
constexpr int kS1ExchPerWarp = 16 * 33;                        // float2 units = one padded [32][33] float square
constexpr size_t kS1TileBytes = 1024 * kS1Frames * sizeof(float);
constexpr size_t kS1ExchBytes = (size_t)kS1Warps * kS1ExchPerWarp * sizeof(float2);
constexpr size_t kS1TwBytes = 1024 * sizeof(float2);
constexpr size_t kS1WinBytes = 1024 * sizeof(float);
constexpr size_t kS1Smem = kS1TileBytes + kS1ExchBytes + kS1TwBytes + kS1WinBytes;   // 111 616 B

struct S1Args {
    const float2 *in;
    long long in_origin, stride;
    int nperseg, hop;
    const float *window;
    const float2 *tw;            // [q][l] = exp(-2*pi*i*q*l/1024)
    float scale;
    long long frame_lo, n_frames;
    float *out;
    long long out_ld;
    int *flag;
};

// exp(-2*pi*i*k/32) = kC[k] - i*kS[k], k = 0..15 (float32-rounded literals)
static constexpr float kC32[16] = {1.0f, 0.980785251f, 0.923879504f, 0.831469595f, 0.707106769f, 0.555570245f,
                                   0.382683426f, 0.195090324f, 0.0f, -0.195090324f, -0.382683426f, -0.555570245f,
                                   -0.707106769f, -0.831469595f, -0.923879504f, -0.980785251f};
static constexpr float kS32[16] = {0.0f, 0.195090324f, 0.382683426f, 0.555570245f, 0.707106769f, 0.831469595f,
                                   0.923879504f, 0.980785251f, 1.0f, 0.980785251f, 0.923879504f, 0.831469595f,
                                   0.707106769f, 0.555570245f, 0.382683426f, 0.195090324f};

// v * exp(-2*pi*i*IDX/32)
template <int IDX>
__device__ __forceinline__ float2 mulw32(const float2 v) {
    constexpr float h = 0.707106769f;
    if constexpr (IDX == 0) {
        return v;
    } else if constexpr (IDX == 8) {                      // -i
        return make_float2(v.y, -v.x);
    } else if constexpr (IDX == 4) {                      // (1 - i)/sqrt2
        return make_float2((v.x + v.y) * h, (v.y - v.x) * h);
    } else if constexpr (IDX == 12) {                     // (-1 - i)/sqrt2
        return make_float2((v.y - v.x) * h, -(v.x + v.y) * h);
    } else {                                              // (x + iy)(c - is) = (xc + ys) + i(yc - xs)
        constexpr float c = kC32[IDX], s = kS32[IDX];
        return make_float2(fmaf(v.y, s, v.x * c), fmaf(-v.x, s, v.y * c));
    }
}

template <int NS, int J>
__device__ __forceinline__ void s1_bfly(const float2 (&x)[32], float2 (&y)[32]) {
    constexpr int k = J % NS;
    constexpr int j0 = (J / NS) * 2 * NS + k;
    const float2 a = x[J];
    const float2 t = mulw32<k * (16 / NS)>(x[J + 16]);   // exp(-2*pi*i*k/(2*NS))
    y[j0] = make_float2(a.x + t.x, a.y + t.y);
    y[j0 + NS] = make_float2(a.x - t.x, a.y - t.y);
}

template <int NS, int... J>
__device__ __forceinline__ void s1_stage(const float2 (&x)[32], float2 (&y)[32], std::integer_sequence<int, J...>) {
    (s1_bfly<NS, J>(x, y), ...);
}

// 32-point forward DFT, natural order in (a) and out (b); a is clobbered.
__device__ __forceinline__ void fft32(float2 (&a)[32], float2 (&b)[32]) {
    using S = std::make_integer_sequence<int, 16>;
    s1_stage<1>(a, b, S{});
    s1_stage<2>(b, a, S{});
    s1_stage<4>(a, b, S{});
    s1_stage<8>(b, a, S{});
    s1_stage<16>(a, b, S{});
}

// (l, q) swap across the warp: on entry lane l holds y[q] = B_l[q]; on exit lane q holds
// x[l] = B_l[q].  Real parts, then imaginary parts, through one padded [32][33] float square:
// every lane is active in every access, all accesses are conflict-free, and at most 64 data
// registers are live at any point (the re halves die as they are stored).
__device__ __forceinline__ void exchange_lq(float *e, int lane, const float2 (&y)[32], float2 (&x)[32]) {
#pragma unroll
    for (int q = 0; q < 32; ++q) e[q * 33 + lane] = y[q].x;
    __syncwarp();
#pragma unroll
    for (int l = 0; l < 32; ++l) x[l].x = e[lane * 33 + l];
    __syncwarp();
#pragma unroll
    for (int q = 0; q < 32; ++q) e[q * 33 + lane] = y[q].y;
    __syncwarp();
#pragma unroll
    for (int l = 0; l < 32; ++l) x[l].y = e[lane * 33 + l];
    __syncwarp();
}

__global__ void __launch_bounds__(kS1Threads, 2)
stft1024_gather_kernel(const S1Args a) {
    extern __shared__ __align__(16) unsigned char s1_smem[];
    float *tile = reinterpret_cast<float *>(s1_smem);
    float2 *exch_all = reinterpret_cast<float2 *>(s1_smem + kS1TileBytes);
    float2 *tw = reinterpret_cast<float2 *>(s1_smem + kS1TileBytes + kS1ExchBytes);
    float *win = reinterpret_cast<float *>(s1_smem + kS1TileBytes + kS1ExchBytes + kS1TwBytes);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float *exch = reinterpret_cast<float *>(exch_all + warp * kS1ExchPerWarp);
    for (int i = tid; i < 1024; i += kS1Threads) {
        tw[i] = __ldg(a.tw + i);
        win[i] = (i < a.nperseg) ? __ldg(a.window + i) : 0.f;
    }
    __syncthreads();

    int any = 0;
    const int swz = (lane >> 1) & 15;                       // tile column swizzle of this lane's rows
    const long long n_tiles = (a.n_frames + kS1Frames - 1) / kS1Frames;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const long long f0 = t * kS1Frames;
        const int nfr = (int)((a.n_frames - f0 < kS1Frames) ? (a.n_frames - f0) : kS1Frames);
#pragma unroll 1
        for (int h = 0; h < kS1Frames / kS1Warps; ++h) {
            const int c = warp + kS1Warps * h;              // tile column = frame within the tile (warp-uniform)
            if (c >= nfr) continue;
            const long long f = a.frame_lo + f0 + c;
            float2 x[32], y[32];
            {
                const long long base = f * a.hop;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int n = lane + 32 * j;
                    float2 v = make_float2(0.f, 0.f);
                    if (n < a.nperseg) {
                        v = __ldg(a.in + ((base + n) * a.stride - a.in_origin));
                        const float w = win[n];
                        v.x *= w; v.y *= w;
                    }
                    x[j] = v;
                }
            }
            fft32(x, y);                                    // y[q] = sum_j x[l + 32 j] W32^(j q)
#pragma unroll
            for (int q = 1; q < 32; ++q) {                  // times W1024^(l q)
                const float2 w = tw[q * 32 + lane];
                const float2 v = y[q];
                y[q] = make_float2(v.x * w.x - v.y * w.y, v.x * w.y + v.y * w.x);
            }
            exchange_lq(exch, lane, y, x);                  // lane q receives x[l] = y_l[q]
            fft32(x, y);                                    // y[s] = X[lane + 32 s]
            const int col = c ^ swz;
#pragma unroll
            for (int s = 0; s < 32; ++s) {
                const float p = (y[s].x * y[s].x + y[s].y * y[s].y) * a.scale;
                any |= (p != 0.f);
                const int row = lane + 32 * ((s + 16) & 31);                 // fftshift
                tile[row * kS1Frames + col] = p;
            }
        }
        __syncthreads();
        {
            const int c = tid & (kS1Frames - 1);
            const int r0 = tid / kS1Frames;                 // 0..15
            if (c < nfr) {
                float *o = a.out + f0 + c;
#pragma unroll 8
                for (int i = 0; i < 1024 / (kS1Threads / kS1Frames); ++i) {
                    const int r = r0 + (kS1Threads / kS1Frames) * i;
                    __stcs(o + (long long)r * a.out_ld, tile[r * kS1Frames + (c ^ ((r >> 1) & 15))]);
                }
            }
        }
        __syncthreads();
    }
    if (a.flag && __syncthreads_or(any) && tid == 0) atomicOr(a.flag, 1);
}

// ---- TMA / mbarrier primitives (sm_90+ PTX; SASS: UBLKCP / SYNCS) -------------------------
__device__ __forceinline__ unsigned smem_u32(const void *p) {
    return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned mbar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(mbar), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned mbar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity) {
    unsigned ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned mbar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(mbar) : "memory");
}

// Shared-memory plan of stft1024_tma_kernel<W>: W warps, W frames per tile (one per warp).
template <int W>
struct T1 {
    static constexpr int kThreads = 32 * W;
    static constexpr int kSpan = (W - 1) * 512 + 1024;                 // input samples per tile at hop <= 512
    static constexpr size_t kTile = (size_t)1024 * W * sizeof(float);  // PSD tile [1024 rows][W frames]
    static constexpr size_t kOffExch = kTile;
    static constexpr size_t kOffTw = kOffExch + (size_t)W * 32 * 33 * sizeof(float);
    static constexpr size_t kOffStage = kOffTw + kS1TwBytes;
    static constexpr size_t kOffBar = kOffStage + (size_t)kSpan * sizeof(float2);
    static constexpr size_t kSmem = kOffBar + 16;      // W=16: 211 216 B (1 CTA/SM); W=8: 111 632 B (2 CTAs/SM)
};

template <int W>
__device__ __forceinline__ void t1_issue_tile(const S1Args &a, long long tile, unsigned stage, unsigned mbar) {
    const long long f0 = tile * W;
    const int nfr = (int)((a.n_frames - f0 < W) ? (a.n_frames - f0) : W);
    const unsigned bytes = (unsigned)(((nfr - 1) * a.hop + 1024) * sizeof(float2));
    const char *src = reinterpret_cast<const char *>(a.in + ((a.frame_lo + f0) * a.hop - a.in_origin));
    mbar_expect_tx(mbar, bytes);
    for (unsigned off = 0; off < bytes; off += 16384) {           // a few <= 16 KB bulk copies, one barrier
        const unsigned n = (bytes - off < 16384u) ? (bytes - off) : 16384u;
        tma_bulk_g2s(stage + off, src + off, n, mbar);
    }
}

template <int W>
__global__ void __launch_bounds__(32 * W, 16 / W)
stft1024_tma_kernel(const S1Args a) {
    using L = T1<W>;
    constexpr int kRowsPerPass = L::kThreads / W;                 // = 32
    constexpr int kSwzShift = (W == 16) ? 1 : 2;                  // rows per bank group: 32 / W
    extern __shared__ __align__(128) unsigned char t1_smem[];
    float *tile = reinterpret_cast<float *>(t1_smem);
    float2 *tw = reinterpret_cast<float2 *>(t1_smem + L::kOffTw);
    const float2 *stage = reinterpret_cast<const float2 *>(t1_smem + L::kOffStage);
    const unsigned stage_u32 = smem_u32(t1_smem + L::kOffStage);
    const unsigned mbar = smem_u32(t1_smem + L::kOffBar);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float *exch = reinterpret_cast<float *>(t1_smem + L::kOffExch) + warp * (32 * 33);
    for (int i = tid; i < 1024; i += L::kThreads) tw[i] = __ldg(a.tw + i);
    float wreg[32];                                               // this lane's window taps, resident
#pragma unroll
    for (int j = 0; j < 32; ++j) wreg[j] = __ldg(a.window + lane + 32 * j);
    const long long n_tiles = (a.n_frames + W - 1) / W;
    if (tid == 0) {
        mbar_init(mbar, 1);
        if ((long long)blockIdx.x < n_tiles) t1_issue_tile<W>(a, blockIdx.x, stage_u32, mbar);
    }
    __syncthreads();

    unsigned anybits = 0;
    unsigned parity = 0;
    const int c = warp;                                           // tile column = frame within the tile
    float *tile_w = tile + lane * W + (c ^ ((lane >> kSwzShift) & (W - 1)));          // + 32*W*rowblock
    const int cc = tid & (W - 1);
    const int r0 = tid / W;                                       // 0..31
    const float *tile_r = tile + r0 * W + (cc ^ ((r0 >> kSwzShift) & (W - 1)));       // + 32*W*i
    const long long out_step = (long long)kRowsPerPass * a.out_ld;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const long long f0 = t * W;
        const int nfr = (int)((a.n_frames - f0 < W) ? (a.n_frames - f0) : W);
        float2 x[32], y[32];
        mbar_wait(mbar, parity);                                  // this tile's samples have landed
        parity ^= 1;
        if (c < nfr) {
            const float2 *p = stage + c * a.hop + lane;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const float2 v = p[32 * j];
                x[j] = make_float2(v.x * wreg[j], v.y * wreg[j]);
            }
        }
        __syncthreads();                                          // staging and PSD tile are both free now
        if (tid == 0 && t + gridDim.x < n_tiles) t1_issue_tile<W>(a, t + gridDim.x, stage_u32, mbar);
        if (c < nfr) {
            fft32(x, y);                                          // y[q] = sum_j x[l + 32 j] W32^(j q)
#pragma unroll
            for (int q = 1; q < 32; ++q) {                        // times W1024^(l q)
                const float2 w = tw[q * 32 + lane];
                const float2 v = y[q];
                y[q] = make_float2(v.x * w.x - v.y * w.y, v.x * w.y + v.y * w.x);
            }
            exchange_lq(exch, lane, y, x);
            fft32(x, y);                                          // y[s] = X[lane + 32 s]
#pragma unroll
            for (int s = 0; s < 32; s += 2) {
                const float p0 = (y[s].x * y[s].x + y[s].y * y[s].y) * a.scale;
                const float p1 = (y[s + 1].x * y[s + 1].x + y[s + 1].y * y[s + 1].y) * a.scale;
                anybits |= __float_as_uint(p0) | __float_as_uint(p1);     // p >= 0 or NaN: nonzero bits <=> p != 0
                tile_w[32 * W * ((s + 16) & 31)] = p0;            // row = lane + 32*((s+16)&31): the fftshift
                tile_w[32 * W * ((s + 17) & 31)] = p1;
            }
        }
        __syncthreads();                                          // tile complete
        if (cc < nfr) {
            float *o = a.out + f0 + cc + (long long)r0 * a.out_ld;
#pragma unroll 8
            for (int i = 0; i < 1024 / kRowsPerPass; ++i) {
                __stcs(o, tile_r[32 * W * i]);
                o += out_step;
            }
        }
    }
    if (a.flag && __syncthreads_or(anybits != 0) && tid == 0) atomicOr(a.flag, 1);
}

inline int stft1024_setup() {
    B2_CUDA(cudaFuncSetAttribute(stft1024_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kS1Smem));
    B2_CUDA(cudaFuncSetAttribute(stft1024_tma_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T1<16>::kSmem));
#ifdef B2_AB
    B2_CUDA(cudaFuncSetAttribute(stft1024_tma_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T1<8>::kSmem));
#endif
    return 0;
}

inline bool stft1024_enabled() {
    const char *e = ab_env("B200VEC_STFT1024");            // "0" forces the generic kernel (A/B builds)
    return !(e && e[0] == '0');
}

inline int get_tw1024(b200vec_ctx *ctx, const float2 **out) {
    const int key = -1024;                                  // [q][l] layout, kept apart from the radix-2 tables
    auto it = ctx->twiddles.find(key);
    if (it != ctx->twiddles.end()) { *out = it->second; return 0; }
    std::vector<float2> h(1024);
    for (int q = 0; q < 32; ++q)
        for (int l = 0; l < 32; ++l) {
            const double ang = -2.0 * M_PI * (double)((q * l) % 1024) / 1024.0;
            h[q * 32 + l] = make_float2((float)cos(ang), (float)sin(ang));
        }
    float2 *d = nullptr;
    B2_CUDA(cudaMalloc(&d, sizeof(float2) * 1024));
    B2_CUDA(cudaMemcpy(d, h.data(), sizeof(float2) * 1024, cudaMemcpyHostToDevice));
    ctx->twiddles[key] = d;
    *out = d;
    return 0;
}

int launch_tile1024_plain(b200vec_ctx *ctx, const S1Args &a, cudaStream_t st);   // fused1024.cuh

inline int launch_stft1024(b200vec_ctx *ctx, const float2 *in, long long in_origin, long long in_len,
                           long long stride, int nperseg, int hop, const float *window, float scale,
                           long long frame_lo, long long n_frames, float *out, long long out_ld,
                           int *flag, cudaStream_t st) {
    (void)in_len;
    S1Args a;
    a.in = in; a.in_origin = in_origin; a.stride = stride; a.nperseg = nperseg; a.hop = hop;
    a.window = window; a.scale = scale; a.frame_lo = frame_lo; a.n_frames = n_frames;
    a.out = out; a.out_ld = out_ld; a.flag = flag;
    if (get_tw1024(ctx, &a.tw)) return 1;
    const long long n_tiles = (n_frames + kS1Frames - 1) / kS1Frames;
    const char *no_tma = ab_env("B200VEC_STFT_TMA");          // "0" forces the gather kernel (A/B builds)
    const long long first = frame_lo * (long long)hop - in_origin;
    const bool tma_ok = stride == 1 && nperseg == 1024 && hop <= 512 && (hop & 1) == 0 &&
                        ((reinterpret_cast<uintptr_t>(in + first)) & 15) == 0 && !(no_tma && no_tma[0] == '0');
    if (tma_ok) {
        const char *wsel = ab_env("B200VEC_STFT_W");           // "8"/"16": the v2 tile-transposing kernels (A/B builds)
        if (!wsel && out_ld < (1LL << 24))                     // default: direct-store kernel (fused1024.cuh)
            return launch_tile1024_plain(ctx, a, st);
#ifdef B2_AB
        if (wsel && wsel[0] == '8') {
            const long long nt = (n_frames + 7) / 8;
            long long grid = nt < 2LL * ctx->sm_count ? nt : 2LL * ctx->sm_count;
            stft1024_tma_kernel<8><<<(unsigned)grid, 256, T1<8>::kSmem, st>>>(a);
            B2_CUDA(cudaGetLastError());
            return 0;
        }
#endif
        // rows longer than 2^24 frames (the direct-store kernel's 32-bit row offsets end there): tile-transposing kernel
        const long long nt = (n_frames + 15) / 16;
        long long grid = nt < ctx->sm_count ? nt : ctx->sm_count;
        stft1024_tma_kernel<16><<<(unsigned)grid, 512, T1<16>::kSmem, st>>>(a);
    } else {
        long long grid = n_tiles < 2LL * ctx->sm_count ? n_tiles : 2LL * ctx->sm_count;
        stft1024_gather_kernel<<<(unsigned)grid, kS1Threads, kS1Smem, st>>>(a);
    }
    B2_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200vec
