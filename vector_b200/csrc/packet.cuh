// packet.cuh — the two helpers either side of the vector-build path that SURVEY.md §8(f) ranks next:
//   detect_packet_bounds (utils.py:811-825): |x|^2, 1 us box smoothing (np.convolve 'same', float64),
//     threshold between the noise floor and the peak, first / last sample above it;
//   plot_spectrogram's display conditioning (utils.py:435-440): scipy.ndimage.median_filter(Sxx_db,
//     size=(2, 1)) — for two taps the rank filter returns the larger one, with the first row reflected.
// Both are bandwidth-bound stencils; no tensor cores.
#pragma once
#include "common.cuh"

namespace b200vec {

constexpr int kBoxThreads = 256;
constexpr int kBoxMaxWindow = 4096;             // 1 us at up to 4 GSa/s

// smoothed[i] = sum_{m = i - window/2}^{i + (window-1)/2} e[m] * (1/window), e[m] = fl32(fl32(|x[m]|)^2), float64
// accumulation in ascending m (np.convolve(energy, np.ones(w)/w, 'same'); samples outside [0, n) count 0).
// maxbits is raised to max(smoothed) (non-negative doubles order like their bit patterns).
__global__ void __launch_bounds__(kBoxThreads)
box_energy_kernel(const float2 *__restrict__ x, long long n, int window, double *__restrict__ out,
                  unsigned long long *__restrict__ maxbits) {
    __shared__ double e[kBoxThreads + kBoxMaxWindow];
    __shared__ double red[kBoxThreads / 32];
    const int tid = threadIdx.x;
    const double c = 1.0 / (double)window;
    const int back = window / 2;
    double tmax = 0.0;
    const long long n_blocks = (n + kBoxThreads - 1) / kBoxThreads;
    for (long long b = blockIdx.x; b < n_blocks; b += gridDim.x) {
        const long long i0 = b * kBoxThreads;
        __syncthreads();
        for (int k = tid; k < kBoxThreads + window - 1; k += kBoxThreads) {
            const long long m = i0 - back + k;
            double v = 0.0;
            if (m >= 0 && m < n) {
                const float2 s = __ldg(x + m);
                const float a = (float)sqrt((double)s.x * (double)s.x + (double)s.y * (double)s.y);   // |x| rounded to float32
                v = (double)(a * a);
            }
            e[k] = v;
        }
        __syncthreads();
        const long long i = i0 + tid;
        if (i < n) {
            double acc = 0.0;
            for (int k = 0; k < window; ++k) acc += e[tid + k] * c;
            out[i] = acc;
            tmax = fmax(tmax, acc);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tmax = fmax(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
    if ((tid & 31) == 0) red[tid >> 5] = tmax;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < kBoxThreads / 32; ++w) tmax = fmax(tmax, red[w]);
        if (tmax > 0.0) atomicMax(maxbits, (unsigned long long)__double_as_longlong(tmax));
    }
}

// first_last[0] = min i, first_last[1] = max i with s[i] >= threshold (caller presets n and -1)
__global__ void __launch_bounds__(256)
threshold_bounds_kernel(const double *__restrict__ s, long long n, double threshold, long long *__restrict__ first_last) {
    long long lo = n, hi = -1;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        if (s[i] >= threshold) { if (i < lo) lo = i; if (i > hi) hi = i; }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const long long l2 = __shfl_xor_sync(0xffffffffu, lo, o), h2 = __shfl_xor_sync(0xffffffffu, hi, o);
        lo = l2 < lo ? l2 : lo; hi = h2 > hi ? h2 : hi;
    }
    if ((threadIdx.x & 31) == 0 && hi >= 0) {
        atomicMin(first_last, lo);
        atomicMax(first_last + 1, hi);
    }
}

// out[r][c] = max(in[r-1][c], in[r][c]), row -1 reflected onto row 0 (median of two taps = the larger one in
// scipy.ndimage's rank filter, mode='reflect'); arbitrary element strides so both spectrogram layouts work.
__global__ void __launch_bounds__(256)
median2_rows_kernel(const float *__restrict__ in, long long rows, long long cols, long long rs, long long cs,
                    float *__restrict__ out, long long ors, long long ocs, int cols_fast) {
    const long long total = rows * cols;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
        long long r, c;
        if (cols_fast) { r = t / cols; c = t - r * cols; } else { c = t / rows; r = t - c * rows; }
        const float cur = in[r * rs + c * cs];
        const float prev = r > 0 ? in[(r - 1) * rs + c * cs] : cur;
        out[r * ors + c * ocs] = fmaxf(prev, cur);         // dB values: no NaNs on this path
    }
}

}  // namespace b200vec
