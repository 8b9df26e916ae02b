// select.cuh — K5: dB map and exact order statistics for normalize_spectrogram
// (reference utils.py:356-404: percentile(|S|[|S|>0], 5) -> floor; 10*log10(|S|+floor);
// percentile(Sxx_db, lo/hi)).
//
// np.percentile needs the exact values at one or two integer ranks of the sorted data.
// Instead of sorting 10^7..10^8 floats, a 3-level MSD radix select runs on order-preserving
// uint32 keys: 11 + 11 + 10 bits, one streaming pass per level, histograms in shared
// memory with warp-aggregated atomics (spectrograms are heavy with repeated values — 42 %
// exact zeros at BASELINE cfg2 — which would serialise naive atomics).
#pragma once
#include "common.cuh"

namespace b200vec {

constexpr int kSelThreads = 256;
constexpr int kBins1 = 2048, kBins2 = 2048, kBins3 = 1024;
// scratch layout in ctx->hist (unsigned int words)
constexpr int kHist1Off = 0;                                          // [2048]
constexpr int kHist2Off = kHist1Off + kBins1;                         // [MAX_RANKS][2048]
constexpr int kHist3Off = kHist2Off + B200VEC_MAX_RANKS * kBins2;     // [MAX_RANKS][1024]
constexpr int kStateOff = kHist3Off + B200VEC_MAX_RANKS * kBins3;     // SelState
constexpr int kHistWords = kStateOff + 64 + 4 * B200VEC_MAX_RANKS;

struct SelState {
    unsigned long long count;                         // selected elements
    unsigned long long resid[B200VEC_MAX_RANKS];      // rank within the current prefix bucket
    unsigned int prefix[B200VEC_MAX_RANKS];           // key prefix found so far
    unsigned int result[B200VEC_MAX_RANKS];           // final keys
    int n_ranks;
};

__device__ __forceinline__ unsigned int f2key(float x) {
    const unsigned int u = __float_as_uint(x);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float key2f(unsigned int k) {
    const unsigned int u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f; memcpy(&f, &u, 4); return f;
#endif
}

// add 1 to h[bin] for every active lane with pred, one shared atomic per distinct bin per warp
__device__ __forceinline__ void warp_hist_add(unsigned int *h, unsigned int bin, bool pred) {
    const unsigned int active = __ballot_sync(0xffffffffu, pred);
    if (!pred) return;
    const unsigned int peers = __match_any_sync(active, bin);
    if ((int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[bin], __popc(peers));
}

__global__ void __launch_bounds__(kSelThreads)
db_kernel(const float *__restrict__ s, long long n, float floor_val, float *__restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        out[i] = 10.0f * log10f(fabsf(__ldg(s + i)) + floor_val);
}

// level 1: histogram of the top 11 key bits over all selected elements
__global__ void __launch_bounds__(kSelThreads)
sel_hist1(const float *__restrict__ x, long long n, int positive_only, unsigned int *__restrict__ hist) {
    __shared__ unsigned int h[kBins1];
    for (int i = threadIdx.x; i < kBins1; i += kSelThreads) h[i] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_round = ((n + stride - 1) / stride) * stride;     // keep warps converged for the ballots
    for (long long i = i0; i < n_round; i += stride) {
        float v = 0.f; bool ok = false;
        if (i < n) { v = __ldg(x + i); ok = positive_only ? (v > 0.f) : true; }
        warp_hist_add(h, f2key(v) >> 21, ok);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kBins1; i += kSelThreads)
        if (h[i]) atomicAdd(&hist[kHist1Off + i], h[i]);
}

// levels 2 and 3: histogram of the next bits for the elements matching each rank's prefix
template <int LEVEL>
__global__ void __launch_bounds__(kSelThreads)
sel_histN(const float *__restrict__ x, long long n, int positive_only, unsigned int *__restrict__ hist) {
    constexpr int BINS = (LEVEL == 2) ? kBins2 : kBins3;
    constexpr int SHIFT_PREFIX = (LEVEL == 2) ? 21 : 10;
    extern __shared__ unsigned int hs[];          // [n_ranks][BINS]
    const SelState *st = reinterpret_cast<const SelState *>(hist + kStateOff);
    const int nr = st->n_ranks;
    unsigned int pref[B200VEC_MAX_RANKS];
#pragma unroll
    for (int r = 0; r < B200VEC_MAX_RANKS; ++r) pref[r] = (r < nr) ? st->prefix[r] : 0xffffffffu;
    for (int i = threadIdx.x; i < nr * BINS; i += kSelThreads) hs[i] = 0;
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n_round = ((n + stride - 1) / stride) * stride;
    for (long long i = i0; i < n_round; i += stride) {
        float v = 0.f; bool ok = false;
        if (i < n) { v = __ldg(x + i); ok = positive_only ? (v > 0.f) : true; }
        const unsigned int key = f2key(v);
        const unsigned int top = key >> SHIFT_PREFIX;
        const unsigned int bin = (LEVEL == 2) ? ((key >> 10) & 0x7ffu) : (key & 0x3ffu);
#pragma unroll
        for (int r = 0; r < B200VEC_MAX_RANKS; ++r) {
            if (r < nr) {
                // identical prefixes share one histogram (the first rank that has it)
                bool dup = false;
                for (int q = 0; q < r; ++q) dup |= (pref[q] == pref[r]);
                if (!dup) warp_hist_add(hs + r * BINS, bin, ok && top == pref[r]);
            }
        }
    }
    __syncthreads();
    const int off = (LEVEL == 2) ? kHist2Off : kHist3Off;
    for (int i = threadIdx.x; i < nr * BINS; i += kSelThreads)
        if (hs[i]) atomicAdd(&hist[off + i], hs[i]);
}

// single block: walk the level histogram of each rank, refine prefix/resid
template <int LEVEL>
__global__ void __launch_bounds__(1)
sel_pick(unsigned int *__restrict__ hist) {
    SelState *st = reinterpret_cast<SelState *>(hist + kStateOff);
    constexpr int BINS = (LEVEL == 1) ? kBins1 : (LEVEL == 2 ? kBins2 : kBins3);
    constexpr int BITS = (LEVEL == 3) ? 10 : 11;
    if (LEVEL == 1) {
        unsigned long long c = 0;
        for (int b = 0; b < BINS; ++b) c += hist[kHist1Off + b];
        st->count = c;
    }
    unsigned int old[B200VEC_MAX_RANKS];
    for (int r = 0; r < B200VEC_MAX_RANKS; ++r) old[r] = st->prefix[r];
    for (int r = 0; r < st->n_ranks; ++r) {
        int src = r;                       // ranks with an identical prefix share the first one's histogram
        if (LEVEL != 1)
            for (int q = r - 1; q >= 0; --q)
                if (old[q] == old[r]) src = q;
        const unsigned int *h = (LEVEL == 1) ? hist + kHist1Off
                              : (LEVEL == 2) ? hist + kHist2Off + src * kBins2
                                             : hist + kHist3Off + src * kBins3;
        unsigned long long want = st->resid[r], cum = 0;
        int b = 0;
        for (; b < BINS - 1; ++b) {
            const unsigned long long c = h[b];
            if (cum + c > want) break;
            cum += c;
        }
        st->resid[r] = want - cum;
        st->prefix[r] = (LEVEL == 1) ? (unsigned)b : ((st->prefix[r] << BITS) | (unsigned)b);
        if (LEVEL == 3) st->result[r] = st->prefix[r];
    }
}

}  // namespace b200vec
