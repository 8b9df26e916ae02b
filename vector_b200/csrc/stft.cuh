// stft.cuh — K4: framed, windowed, zero-padded two-sided FFT -> |X|^2*scale, fftshifted rows.
//
// Replaces scipy.signal.spectrogram(..., return_onesided=False, detrend=False) + fftshift as
// the reference calls it (utils.py:281-291,303-313,324-347,350-351;
// vector_analyzer/spectrogram_analysis.py:40-53).  scipy's legacy _spectral_helper does, for
// complex64 input: frames * window(complex64) -> fft(n=nfft) (complex64 pocketfft) ->
// conj(X)*X * scale -> .real.  The same steps in float32 here.
//
// Output layout is the reference's: C-contiguous [nfft, n_frames], row r = bin
// (r + nfft/2) mod nfft.  Frames are the fast axis, so each block transforms a group of
// T frames, parks the T power spectra in a shared [nfft][T+1] tile and writes every row as
// one contiguous run of T floats (128 B for T = 32).
//
// Generic kernel: any power-of-two nfft in [32, 16384], any nperseg <= nfft, any hop and
// any input stride (the decimated preview of create_spectrogram, utils.py:192-195, is a
// strided gather — the decimated copy is never materialised).
#pragma once
#include "common.cuh"
#include <math.h>

namespace b200vec {

constexpr int kStftThreads = 256;

struct StftArgs {
    const float2 *in;
    long long in_origin, in_len, stride;
    int nperseg, hop, nfft, log2n;
    const float *window;
    const float2 *tw;            // exp(-2*pi*i*k/nfft), k < nfft/2
    float scale;
    long long frame_lo, n_frames;
    float *out;
    long long out_ld;
    int *flag;
    int T;                       // frames per block (power of two, <= 32)
};

__device__ __forceinline__ float2 cmulf(float2 a, float2 b) {
    return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

__global__ void __launch_bounds__(kStftThreads)
stft_generic_kernel(const StftArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2 *buf = reinterpret_cast<float2 *>(smem_raw);
    float *tile = reinterpret_cast<float *>(smem_raw + (size_t)a.nfft * sizeof(float2));
    const int tid = threadIdx.x;
    const int n = a.nfft, T = a.T, TP = T + 1;
    const long long g0 = (long long)blockIdx.x * T;           // first local frame of this block
    const int nfr = (int)((a.n_frames - g0 < T) ? (a.n_frames - g0) : T);
    int any = 0;

    for (int t = 0; t < nfr; ++t) {
        const long long f = a.frame_lo + g0 + t;
        const long long base = f * a.hop;
        // gather + window + bit-reversed scatter
        for (int i = tid; i < n; i += kStftThreads) {
            float2 v = make_float2(0.f, 0.f);
            if (i < a.nperseg) {
                const long long gi = (base + i) * a.stride - a.in_origin;
                if (gi >= 0 && gi < a.in_len) {
                    v = __ldg(a.in + gi);
                    const float w = __ldg(a.window + i);
                    v.x *= w; v.y *= w;
                }
            }
            buf[__brev((unsigned)i) >> (32 - a.log2n)] = v;
        }
        __syncthreads();
        // radix-2 DIT, bit-reversed in -> natural out
        for (int h = 1, sh = a.log2n - 1; h < n; h <<= 1, --sh) {
            for (int b = tid; b < (n >> 1); b += kStftThreads) {
                const int j = b & (h - 1);
                const int i0 = ((b - j) << 1) + j;
                const int i1 = i0 + h;
                const float2 w = __ldg(a.tw + ((long long)j << sh));
                const float2 u = buf[i0];
                const float2 v = cmulf(buf[i1], w);
                buf[i0] = make_float2(u.x + v.x, u.y + v.y);
                buf[i1] = make_float2(u.x - v.x, u.y - v.y);
            }
            __syncthreads();
        }
        // power, fftshift
        for (int i = tid; i < n; i += kStftThreads) {
            const float2 X = buf[i];
            const float p = (X.x * X.x + X.y * X.y) * a.scale;
            any |= (p != 0.f);      // np.max(Sxx) == 0 is false for NaN too (utils.py:316)
            const int row = (i + (n >> 1)) & (n - 1);
            if (T > 1) tile[row * TP + t] = p;
            else a.out[(long long)row * a.out_ld + g0] = p;
        }
        __syncthreads();
    }
    if (T > 1) {
        // rows of nfr contiguous floats; a warp covers 32/T rows per store instruction
        const int lane_col = tid & (T - 1);
        const int rows_per_pass = kStftThreads / T;
        for (int r = tid / T; r < n; r += rows_per_pass) {
            if (lane_col < nfr)
                __stcs(a.out + (long long)r * a.out_ld + g0 + lane_col, tile[r * TP + lane_col]);
        }
    }
    if (a.flag && __syncthreads_or(any) && tid == 0) atomicOr(a.flag, 1);
}

// ---- float64 form: scipy's path for complex128 input (utils.py:281-291 with a double-precision signal: window,
// pocketfft and conj(X)*X*scale all in float64, float64 PSD out).  Same structure as the kernel above with every
// quantity in double; nfft <= 8192 (the frame buffer alone is 16 B per bin of shared memory).
struct Stft64Args {
    const double2 *in;
    long long in_origin, in_len, stride;
    int nperseg, hop, nfft, log2n;
    const double *window;
    const double2 *tw;           // exp(-2*pi*i*k/nfft), k < nfft/2, in double
    double scale;
    long long frame_lo, n_frames;
    double *out;
    long long out_ld;
    int *flag;
    int T;
};

__global__ void __launch_bounds__(kStftThreads)
stft_generic_f64_kernel(const Stft64Args a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double2 *buf = reinterpret_cast<double2 *>(smem_raw);
    double *tile = reinterpret_cast<double *>(smem_raw + (size_t)a.nfft * sizeof(double2));
    const int tid = threadIdx.x;
    const int n = a.nfft, T = a.T, TP = T + 1;
    const long long g0 = (long long)blockIdx.x * T;
    const int nfr = (int)((a.n_frames - g0 < T) ? (a.n_frames - g0) : T);
    int any = 0;

    for (int t = 0; t < nfr; ++t) {
        const long long base = (a.frame_lo + g0 + t) * a.hop;
        for (int i = tid; i < n; i += kStftThreads) {
            double2 v = make_double2(0.0, 0.0);
            if (i < a.nperseg) {
                const long long gi = (base + i) * a.stride - a.in_origin;
                if (gi >= 0 && gi < a.in_len) {
                    v = a.in[gi];
                    const double w = a.window[i];
                    v.x *= w; v.y *= w;
                }
            }
            buf[__brev((unsigned)i) >> (32 - a.log2n)] = v;
        }
        __syncthreads();
        for (int h = 1, sh = a.log2n - 1; h < n; h <<= 1, --sh) {
            for (int b = tid; b < (n >> 1); b += kStftThreads) {
                const int j = b & (h - 1);
                const int i0 = ((b - j) << 1) + j;
                const int i1 = i0 + h;
                const double2 w = a.tw[(long long)j << sh];
                const double2 u = buf[i0], x = buf[i1];
                const double2 v = make_double2(x.x * w.x - x.y * w.y, x.x * w.y + x.y * w.x);
                buf[i0] = make_double2(u.x + v.x, u.y + v.y);
                buf[i1] = make_double2(u.x - v.x, u.y - v.y);
            }
            __syncthreads();
        }
        for (int i = tid; i < n; i += kStftThreads) {
            const double2 X = buf[i];
            const double p = (X.x * X.x + X.y * X.y) * a.scale;
            any |= (p != 0.0);
            const int row = (i + (n >> 1)) & (n - 1);
            if (T > 1) tile[row * TP + t] = p;
            else a.out[(long long)row * a.out_ld + g0] = p;
        }
        __syncthreads();
    }
    if (T > 1) {
        const int lane_col = tid & (T - 1);
        const int rows_per_pass = kStftThreads / T;
        for (int r = tid / T; r < n; r += rows_per_pass) {
            if (lane_col < nfr)
                __stcs(a.out + (long long)r * a.out_ld + g0 + lane_col, tile[r * TP + lane_col]);
        }
    }
    if (a.flag && __syncthreads_or(any) && tid == 0) atomicOr(a.flag, 1);
}

inline int stft64_pick_T(int nfft) {
    const size_t budget = 200 * 1024;
    for (int T = 32; T > 1; T >>= 1)
        if ((size_t)nfft * 16 + (size_t)nfft * 8 * (T + 1) <= budget) return T;
    return 1;
}

inline int stft_pick_T(int nfft) {
    const size_t budget = 200 * 1024;
    for (int T = 32; T > 1; T >>= 1)
        if ((size_t)nfft * 8 + (size_t)nfft * 4 * (T + 1) <= budget) return T;
    return 1;
}

inline int get_twiddles(b200vec_ctx *ctx, int nfft, const float2 **out) {
    auto it = ctx->twiddles.find(nfft);
    if (it != ctx->twiddles.end()) { *out = it->second; return 0; }
    std::vector<float2> h(nfft / 2);
    for (int k = 0; k < nfft / 2; ++k) {
        const double ang = -2.0 * M_PI * (double)k / (double)nfft;
        h[k] = make_float2((float)cos(ang), (float)sin(ang));
    }
    float2 *d = nullptr;
    B2_CUDA(cudaMalloc(&d, sizeof(float2) * (nfft / 2)));
    B2_CUDA(cudaMemcpy(d, h.data(), sizeof(float2) * (nfft / 2), cudaMemcpyHostToDevice));
    ctx->twiddles[nfft] = d;
    *out = d;
    return 0;
}

inline int get_twiddles64(b200vec_ctx *ctx, int nfft, const double2 **out) {
    const int key = (1 << 24) + nfft;                        // float tables: +nfft (radix-2) and -1024 ([q][l] layout)
    auto it = ctx->twiddles.find(key);
    if (it != ctx->twiddles.end()) { *out = reinterpret_cast<const double2 *>(it->second); return 0; }
    std::vector<double2> h(nfft / 2);
    for (int k = 0; k < nfft / 2; ++k) {
        const double ang = -2.0 * M_PI * (double)k / (double)nfft;
        h[k] = make_double2(cos(ang), sin(ang));
    }
    double2 *d = nullptr;
    B2_CUDA(cudaMalloc(&d, sizeof(double2) * (nfft / 2)));
    B2_CUDA(cudaMemcpy(d, h.data(), sizeof(double2) * (nfft / 2), cudaMemcpyHostToDevice));
    ctx->twiddles[key] = reinterpret_cast<float2 *>(d);
    *out = d;
    return 0;
}

}  // namespace b200vec
