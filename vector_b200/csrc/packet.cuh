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

// ---------------------------------------------------------------------------------------
// Exact order statistics of a float64 array (np.median of the smoothed energies' first tenth, utils.py:818) in
// ONE cooperative launch: MSD radix select on order-preserving 64-bit keys, six levels of 11 bits (the last
// of 9), the levels separated by grid barriers; block 0 picks the bins.  Inputs here are small (a tenth of
// a packet), so the six reads come out of L2.
constexpr int kSel64Threads = 512;
constexpr int kSel64MaxRanks = 4;
constexpr int kSel64Levels = 6;

struct Sel64State {
    unsigned long long prefix[kSel64MaxRanks];
    unsigned long long resid[kSel64MaxRanks];
    unsigned long long result[kSel64MaxRanks];      // key bits; the host undoes the order-preserving map
    int n_ranks;
    unsigned int barrier;
};

__device__ __forceinline__ unsigned long long d2key(double x) {
    const unsigned long long u = (unsigned long long)__double_as_longlong(x);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}

// scratch: Sel64State, then two histogram buffers [kSel64MaxRanks][2048] (levels alternate between them)
__global__ void __launch_bounds__(kSel64Threads, 1)
sel64_kernel(const double *__restrict__ x, long long n, unsigned char *__restrict__ scratch) {
    __shared__ unsigned int hs[kSel64MaxRanks][2048];
    __shared__ unsigned long long spref[kSel64MaxRanks];
    volatile Sel64State *st = reinterpret_cast<volatile Sel64State *>(scratch);
    unsigned int *gh = reinterpret_cast<unsigned int *>(scratch + 256);
    const int nr = st->n_ranks;
    unsigned int phase = 0;
    auto grid_sync = [&]() {
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            atomicAdd(const_cast<unsigned int *>(&st->barrier), 1u);
            ++phase;
            while (st->barrier < phase * gridDim.x) __nanosleep(20);
            __threadfence();
        }
        __syncthreads();
    };
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (int level = 0; level < kSel64Levels; ++level) {
        const int bits = level == kSel64Levels - 1 ? 64 - 11 * (kSel64Levels - 1) : 11;
        const int shift = level == kSel64Levels - 1 ? 0 : 64 - 11 * (level + 1);
        unsigned int *g = gh + (level & 1) * kSel64MaxRanks * 2048;
        for (int i = threadIdx.x; i < nr * 2048; i += kSel64Threads) (&hs[0][0])[i] = 0u;
        if (threadIdx.x < nr) spref[threadIdx.x] = st->prefix[threadIdx.x];
        __syncthreads();
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
            const unsigned long long key = d2key(x[i]);
            const unsigned int bin = (unsigned int)(key >> shift) & ((1u << bits) - 1u);
            const unsigned long long top = level == 0 ? 0ull : key >> (shift + bits);
            for (int r = 0; r < nr; ++r)
                if (level == 0 || top == spref[r]) atomicAdd(&hs[r][bin], 1u);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < nr * 2048; i += kSel64Threads)
            if ((&hs[0][0])[i]) atomicAdd(g + i, (&hs[0][0])[i]);
        grid_sync();
        if (blockIdx.x == 0) {
            const int r = threadIdx.x >> 5, lane = threadIdx.x & 31;
            if (r < nr) {                                    // warp r refines rank r: 64 bins per lane
                const unsigned int *h = g + r * 2048;
                unsigned long long s = 0;
                for (int k = 0; k < 64; ++k) s += __ldcg(h + lane * 64 + k);
                unsigned long long inc = s;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long up = __shfl_up_sync(0xffffffffu, inc, o);
                    if (lane >= o) inc += up;
                }
                const unsigned long long want = st->resid[r];
                const unsigned long long before = inc - s;
                const unsigned int owners = __ballot_sync(0xffffffffu, before <= want && want < inc);
                // the owning lane's 64 bins, re-read two per lane: one more L2 round trip instead of up to 64
                const int L = owners ? __ffs(owners) - 1 : 31;
                const unsigned long long before_l = __shfl_sync(0xffffffffu, before, L);
                const unsigned int c0 = __ldcg(h + L * 64 + 2 * lane), c1 = __ldcg(h + L * 64 + 2 * lane + 1);
                unsigned long long inc2 = (unsigned long long)c0 + c1;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long up = __shfl_up_sync(0xffffffffu, inc2, o);
                    if (lane >= o) inc2 += up;
                }
                unsigned long long cum = before_l + inc2 - c0 - c1;
                const bool mine = owners ? (cum <= want && want < cum + c0 + c1) : lane == 31;   // past the end: the last bin
                if (mine) {
                    int b = L * 64 + 2 * lane + 1;
                    if (owners && want < cum + c0) b = L * 64 + 2 * lane; else cum += c0;
                    st->resid[r] = want - cum;
                    const unsigned long long p = level == 0 ? (unsigned long long)b : ((spref[r] << bits) | (unsigned long long)b);
                    st->prefix[r] = p;
                    if (level == kSel64Levels - 1) st->result[r] = p;
                }
            }
            // the other buffer is the next level's: clear it (nobody reads it before the barrier below)
            unsigned int *g2 = gh + ((level + 1) & 1) * kSel64MaxRanks * 2048;
            for (int i = threadIdx.x; i < nr * 2048; i += kSel64Threads) g2[i] = 0u;
        }
        grid_sync();
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
