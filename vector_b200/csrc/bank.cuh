// bank.cuh — packet bank: K1 frequency shift / commit, K0 Fourier-method resample.
#pragma once
#include "common.cuh"

namespace b200vec {

// ---------------------------------------------------------------------------------------
// K1: apply_frequency_shift (reference utils.py:120-127) fused with the bank commit.
//   t = n / fs            (fp64 division, n the packet-local integer index, utils.py:125)
//   phase = (2*pi*f) * t  (w = 2*pi*f is formed on the host in the reference's own order)
//   y[n] = complex64( complex128(x[n]) * (cos phase + j sin phase) )   (utils.py:126-127)
// Products and sums are kept un-fused (__dmul_rn/__dadd_rn) like numpy's complex128 loop.
// ---------------------------------------------------------------------------------------
// sin / cos of a phase of any size without libdevice's Payne-Hanek path (taken above ~1.06e5 rad: a 20 MHz shift
// reaches 1.26e5 rad at the end of a 1 ms packet, and one slow lane stalls its warp): phases past 1e5 rad are
// first reduced by the nearest multiple of 2*pi with the product kept exact inside FMAs (2*pi = C1 + C2 + C3,
// |k| < 2^33: error below 1e-15 rad, far under what the single rounding to float32 of utils.py:127 can see);
// smaller phases go to sincos() unchanged, bit for bit as before.
__device__ __forceinline__ void sincos_phase(double ph, double *s, double *c) {
    if (fabs(ph) > 1.0e5 && fabs(ph) < 5.0e10) {
        const double k = rint(ph * 0.15915494309189535);              // 1 / (2*pi)
        double r = fma(-k, 6.2831853071795862, ph);                   // 2*pi rounded to double
        r = fma(-k, 2.4492935982947064e-16, r);                       // 2*pi - C1
        r = fma(-k, -5.9895396194366790e-33, r);                      // 2*pi - C1 - C2
        ph = r;
    }
    sincos(ph, s, c);
}

__global__ void __launch_bounds__(256)
shift_commit_kernel(const float2 *__restrict__ raw, long long len, double w, double fs,
                    int do_shift, float2 *__restrict__ a, float2 *__restrict__ b) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= len) return;
    float2 y = raw[n];
    if (do_shift) {
        const double t = (double)n / fs;
        const double ph = __dmul_rn(w, t);
        double s, c;
        sincos_phase(ph, &s, &c);
        const double xr = (double)y.x, xi = (double)y.y;
        const double re = __dsub_rn(__dmul_rn(xr, c), __dmul_rn(xi, s));
        const double im = __dadd_rn(__dmul_rn(xr, s), __dmul_rn(xi, c));
        y = make_float2((float)re, (float)im);
    }
    a[n] = y;
    b[n] = y;
}

// max |y|^2 (exact float32 products, in fp64) over a committed packet -> *out (bits, atomicMax; zeroed by the caller)
__global__ void __launch_bounds__(256)
bank_maxsq_kernel(const float2 *__restrict__ y, long long len, unsigned long long *__restrict__ out) {
    double m = 0.0;
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < len; n += (long long)gridDim.x * blockDim.x) {
        const float2 v = y[n];
        m = fmax(m, (double)v.x * (double)v.x + (double)v.y * (double)v.y);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.0) atomicMax(out, (unsigned long long)__double_as_longlong(m));
}

// ---------------------------------------------------------------------------------------
// K0: scipy.signal.resample(x, num) for complex input (reference call site utils.py:117),
// i.e. X = fft(x); Y = zeros(num); copy the m = min(num, n_x) lowest-|f| bins; if m is even
// unite (down-sampling) or split (up-sampling) the unpaired bin m/2; x_r = ifft(Y*num/n_x).
// Arbitrary lengths (56 100 = 2^2*3*5^2*11*17, 200 357, ...) go through Bluestein's chirp-z
// identity on power-of-two FFTs.  Everything is fp64 on the device and rounded once to
// complex64 — closer to the exact operator than scipy's own complex64 pocketfft path
// (oracle error budget: scipy-c64 vs fp64 = 2.3e-7 rel-L2).
//
// Power-of-two FFT: forward = radix-2 DIF (natural in, bit-reversed out), inverse = radix-2
// DIT (bit-reversed in, natural out).  A circular convolution needs no reordering at all:
// both spectra are multiplied in bit-reversed order.
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
    return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

// one DIF stage, half-span h: (u, v) -> (u + v, (u - v) * W_{2h}^{j}),  W = exp(-2*pi*i/(2h))
__global__ void __launch_bounds__(256)
fft64_dif_stage(double2 *__restrict__ x, long long half_m, long long h) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= half_m) return;
    const long long j = t & (h - 1);
    const long long i0 = ((t - j) << 1) + j;
    const long long i1 = i0 + h;
    const double2 u = x[i0], v = x[i1];
    double s, c;
    sincospi(-(double)j / (double)h, &s, &c);
    x[i0] = make_double2(u.x + v.x, u.y + v.y);
    x[i1] = cmul(make_double2(u.x - v.x, u.y - v.y), make_double2(c, s));
}

// one inverse DIT stage, half-span h: v' = v * conj(W_{2h}^{j}); (u, v) -> (u + v', u - v')
__global__ void __launch_bounds__(256)
fft64_dit_inv_stage(double2 *__restrict__ x, long long half_m, long long h) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= half_m) return;
    const long long j = t & (h - 1);
    const long long i0 = ((t - j) << 1) + j;
    const long long i1 = i0 + h;
    double s, c;
    sincospi((double)j / (double)h, &s, &c);
    const double2 u = x[i0];
    const double2 v = cmul(x[i1], make_double2(c, s));
    x[i0] = make_double2(u.x + v.x, u.y + v.y);
    x[i1] = make_double2(u.x - v.x, u.y - v.y);
}

inline void fft64_forward_bitrev(double2 *x, int log2m, cudaStream_t st) {
    const long long m = 1LL << log2m, half = m >> 1;
    const unsigned grid = (unsigned)((half + 255) / 256);
    for (long long h = half; h >= 1; h >>= 1) fft64_dif_stage<<<grid, 256, 0, st>>>(x, half, h);
}

inline void fft64_inverse_from_bitrev(double2 *x, int log2m, cudaStream_t st) {
    const long long m = 1LL << log2m, half = m >> 1;
    const unsigned grid = (unsigned)((half + 255) / 256);
    for (long long h = 1; h <= half; h <<= 1) fft64_dit_inv_stage<<<grid, 256, 0, st>>>(x, half, h);
}

// chirp c_sgn[n] = exp(-sgn * i*pi*n^2/N); n^2 is reduced mod 2N in integers first.
__device__ __forceinline__ double2 chirp(long long n, long long N, int sgn) {
    const unsigned long long r = ((unsigned long long)n * (unsigned long long)n) % (unsigned long long)(2 * N);
    double s, c;
    sincospi((double)r / (double)N, &s, &c);
    return make_double2(c, sgn > 0 ? -s : s);
}

// A[n] = x[n]*c[n] (n < N) else 0;  B[n] = conj(c[n]) for n < N, B[M-n] = conj(c[n]) for 1 <= n < N, else 0
template <typename TIn>
__global__ void __launch_bounds__(256)
bluestein_prepare(const TIn *__restrict__ x, long long N, long long M, int sgn,
                  double2 *__restrict__ A, double2 *__restrict__ B) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= M) return;
    double2 a = make_double2(0.0, 0.0), b = make_double2(0.0, 0.0);
    if (n < N) {
        const double2 c = chirp(n, N, sgn);
        a = cmul(make_double2((double)x[n].x, (double)x[n].y), c);
        b = make_double2(c.x, -c.y);
    } else if (M - n < N) {
        const double2 c = chirp(M - n, N, sgn);
        b = make_double2(c.x, -c.y);
    }
    A[n] = a;
    B[n] = b;
}

__global__ void __launch_bounds__(256)
pointwise_mul(double2 *__restrict__ A, const double2 *__restrict__ B, long long M) {
    const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n < M) A[n] = cmul(A[n], B[n]);
}

// X[k] = c[k] * conv[k] / M   (1/M: the inverse pow-2 FFT above is unnormalised)
__global__ void __launch_bounds__(256)
bluestein_finish(const double2 *__restrict__ conv, long long N, long long M, int sgn,
                 double out_scale, double2 *__restrict__ X) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= N) return;
    const double2 v = cmul(conv[k], chirp(k, N, sgn));
    const double s = out_scale / (double)M;
    X[k] = make_double2(v.x * s, v.y * s);
}

// Y (len n_out) from X (len n_in), scipy _signaltools.resample two-sided branch
__global__ void __launch_bounds__(256)
resample_bins(const double2 *__restrict__ X, long long n_in, long long n_out,
              double2 *__restrict__ Y) {
    const long long k = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n_out) return;
    const long long m = n_in < n_out ? n_in : n_out;
    const long long m2 = m / 2 + 1;
    const long long nneg = m - m2;                 // bins copied at the negative-frequency end
    double2 y = make_double2(0.0, 0.0);
    if (k < m2) y = X[k];
    else if (n_out - k <= nneg) y = X[n_in - (n_out - k)];
    if ((m & 1) == 0) {
        const long long h = m / 2;
        if (n_out < n_in) {                        // down-sampling: Y[-m/2] += X[-m/2]
            if (k == n_out - h) { const double2 e = X[n_in - h]; y.x += e.x; y.y += e.y; }
        } else if (n_in < n_out) {                 // up-sampling: Y[m/2] /= 2; Y[num-m/2] = Y[m/2]
            if (k == h || k == n_out - h) { const double2 e = X[h]; y = make_double2(e.x * 0.5, e.y * 0.5); }
        }
    }
    Y[k] = y;
}

__global__ void __launch_bounds__(256)
round_to_c64(const double2 *__restrict__ x, long long n, float2 *__restrict__ out) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = make_float2((float)x[i].x, (float)x[i].y);
}

inline int ensure_scratch(Scratch &s, size_t bytes) {
    if (s.bytes >= bytes) return 0;
    if (s.ptr) cudaFree(s.ptr);
    s.ptr = nullptr; s.bytes = 0;
    B2_CUDA(cudaMalloc(&s.ptr, bytes));
    s.bytes = bytes;
    return 0;
}

inline int ceil_log2(long long v) { int l = 1; while ((1LL << l) < v) ++l; return l; }   // >= 1 so that M/2 >= 1

// DFT of arbitrary length N with sign sgn (+1 forward, -1 backward/unnormalised), result * out_scale.
// work buffers A,B must hold M = 2^ceil_log2(2N-1) double2 each.
template <typename TIn>
inline int bluestein_dft(const TIn *x, long long N, int sgn, double out_scale, double2 *A,
                         double2 *B, double2 *X, cudaStream_t st) {
    const int lg = ceil_log2(2 * N - 1);
    const long long M = 1LL << lg;
    const unsigned gM = (unsigned)((M + 255) / 256), gN = (unsigned)((N + 255) / 256);
    bluestein_prepare<TIn><<<gM, 256, 0, st>>>(x, N, M, sgn, A, B);
    fft64_forward_bitrev(A, lg, st);
    fft64_forward_bitrev(B, lg, st);
    pointwise_mul<<<gM, 256, 0, st>>>(A, B, M);
    fft64_inverse_from_bitrev(A, lg, st);
    bluestein_finish<<<gN, 256, 0, st>>>(A, N, M, sgn, out_scale, X);
    B2_CUDA(cudaGetLastError());
    return 0;
}

inline int resample_slot(b200vec_ctx *ctx, BankSlot &sl, long long n_out, cudaStream_t st) {
    const long long n_in = sl.len_raw;
    const long long nmax = n_in > n_out ? n_in : n_out;
    const long long M = 1LL << ceil_log2(2 * nmax - 1);
    if (M > (1LL << 27)) return fail("b200vec_bank_resample", "packet too long for the resampler (2N-1 > 2^27)");
    for (int i = 0; i < 2; ++i)
        if (ensure_scratch(ctx->fft64[i], (size_t)M * sizeof(double2))) return 1;
    if (ensure_scratch(ctx->fft64[2], (size_t)nmax * sizeof(double2))) return 1;
    if (ensure_scratch(ctx->fft64[3], (size_t)n_out * sizeof(double2))) return 1;
    double2 *A = (double2 *)ctx->fft64[0].ptr, *B = (double2 *)ctx->fft64[1].ptr;
    double2 *X = (double2 *)ctx->fft64[2].ptr, *Y = (double2 *)ctx->fft64[3].ptr;
    if (bluestein_dft<float2>(sl.raw, n_in, +1, 1.0, A, B, X, st)) return 1;
    resample_bins<<<(unsigned)((n_out + 255) / 256), 256, 0, st>>>(X, n_in, n_out, Y);
    // ifft(Y * num/n_x) = (1/n_x) * sum_k Y[k] e^{+2*pi*i*k*n/num}; X is dead, reuse it
    double2 *Z = X;
    if (bluestein_dft<double2>(Y, n_out, -1, 1.0 / (double)n_in, A, B, Z, st)) return 1;
    // new raw
    if (sl.cap_raw < n_out) {
        B2_CUDA(cudaStreamSynchronize(st));
        float2 *nr = nullptr;
        B2_CUDA(cudaMalloc(&nr, (size_t)n_out * sizeof(float2)));
        if (sl.raw) cudaFree(sl.raw);
        sl.raw = nr; sl.cap_raw = n_out;
    }
    round_to_c64<<<(unsigned)((n_out + 255) / 256), 256, 0, st>>>(Z, n_out, sl.raw);
    B2_CUDA(cudaGetLastError());
    sl.len_raw = n_out;
    sl.len = 0;     // needs a new commit
    return 0;
}

}  // namespace b200vec
