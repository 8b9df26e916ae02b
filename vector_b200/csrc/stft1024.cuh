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
// Per frame: one global read of the samples, one shared round trip, ~1.5 k warp instructions.
// 111 616 B of shared memory and <= 128 registers per thread -> 2 CTAs (16 warps) per SM;
// the grid is persistent: 2 x 148 CTAs stride over the frame tiles.
#pragma once
#include "common.cuh"
#include <utility>

namespace b200vec {

constexpr int kS1Threads = 256;
constexpr int kS1Warps = kS1Threads / 32;
constexpr int kS1Frames = 16;                                  // frames per tile
constexpr int kS1ExchPerWarp = 16 * 33;                        // float2: half of the (q, l) square, rows padded to 33
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

template <bool UNIT>       // UNIT: stride == 1 and nperseg == 1024 (no predicates, immediate offsets)
__global__ void __launch_bounds__(kS1Threads, 2)
stft1024_kernel(const S1Args a) {
    extern __shared__ __align__(16) unsigned char s1_smem[];
    float *tile = reinterpret_cast<float *>(s1_smem);
    float2 *exch_all = reinterpret_cast<float2 *>(s1_smem + kS1TileBytes);
    float2 *tw = reinterpret_cast<float2 *>(s1_smem + kS1TileBytes + kS1ExchBytes);
    float *win = reinterpret_cast<float *>(s1_smem + kS1TileBytes + kS1ExchBytes + kS1TwBytes);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2 *exch = exch_all + warp * kS1ExchPerWarp;
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
            if (UNIT) {
                const float2 *p = a.in + (f * a.hop - a.in_origin) + lane;
#pragma unroll
                for (int j = 0; j < 32; ++j) x[j] = __ldg(p + 32 * j);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float w = win[lane + 32 * j];
                    x[j].x *= w; x[j].y *= w;
                }
            } else {
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
            // (l, q) swap through shared memory, q-halves in turn: lane q receives x[l] = y_l[q]
#pragma unroll
            for (int q = 0; q < 16; ++q) exch[q * 33 + lane] = y[q];
            __syncwarp();
            if (lane < 16) {
#pragma unroll
                for (int l = 0; l < 32; ++l) x[l] = exch[lane * 33 + l];
            }
            __syncwarp();
#pragma unroll
            for (int q = 16; q < 32; ++q) exch[(q - 16) * 33 + lane] = y[q];
            __syncwarp();
            if (lane >= 16) {
#pragma unroll
                for (int l = 0; l < 32; ++l) x[l] = exch[(lane - 16) * 33 + l];
            }
            __syncwarp();
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

inline int stft1024_setup() {
    B2_CUDA(cudaFuncSetAttribute(stft1024_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kS1Smem));
    B2_CUDA(cudaFuncSetAttribute(stft1024_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kS1Smem));
    return 0;
}

inline bool stft1024_enabled() {
    const char *e = getenv("B200VEC_STFT1024");            // "0" forces the generic kernel (A/B tests)
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
    long long grid = n_tiles;
    const long long cap = 2LL * ctx->sm_count;              // persistent: 2 resident CTAs per SM
    if (grid > cap) grid = cap;
    if (stride == 1 && nperseg == 1024)
        stft1024_kernel<true><<<(unsigned)grid, kS1Threads, kS1Smem, st>>>(a);
    else
        stft1024_kernel<false><<<(unsigned)grid, kS1Threads, kS1Smem, st>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200vec
