// select.cuh — K5: dB map and exact order statistics for normalize_spectrogram
// (reference utils.py:356-404: percentile(|S|[|S|>0], 5) -> floor; 10*log10(|S|+floor);
// percentile(Sxx_db, lo/hi)).
//
// np.percentile needs the exact values at one or two integer ranks of the sorted data.
// Instead of sorting 10^7..10^8 floats, a 3-level MSD radix select runs on order-preserving
// uint32 keys: 11 + 11 + 10 bits, one streaming pass per level, histograms in shared
// memory, the lanes sharing the warp's first bin aggregated into one atomic (spectrograms are heavy
// with repeated values — 42 % exact zeros at BASELINE cfg2 — which would serialise naive atomics).
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
constexpr int kHistWords = kStateOff + 64 + 4 * B200VEC_MAX_RANKS + 4 + B200VEC_MAX_RANKS;   // + the grid-barrier counter of the one-launch selects + b200vec_db_values' results

struct SelState {
    unsigned long long count;                         // selected elements
    unsigned long long resid[B200VEC_MAX_RANKS];      // rank within the current prefix bucket
    unsigned int prefix[B200VEC_MAX_RANKS];           // key prefix found so far
    unsigned int result[B200VEC_MAX_RANKS];           // final keys
    int n_ranks;
};

// bins picked for compaction (b200vec_select_finish / k5_plan_kernel)
struct SelBins {
    unsigned int bin[B200VEC_MAX_RANKS];
    int slot[B200VEC_MAX_RANKS];   // level-2 histogram (hist2[slot]) this bin's elements are counted in
    int n;
};

struct K5Plan {
    SelBins sel;                          // bins to compact (ascending), their level-2 histogram slots
    unsigned long long pop;               // candidates the compaction will produce
    unsigned long long n_pos;             // elements > 0
    float gamma[3];                       // interpolation weights: low percentile, high percentile, floor percentile
    int n_ranks;                          // 6, or 8 when there are positive elements
    int sub_of[B200VEC_MAX_RANKS];        // rank r -> index in the candidate select's state, -1: the value is +0
    int overflow;                         // 1: more candidates than the buffer holds (the caller falls back)
};
struct K5Out {
    float floor_v;
    float six[6];                         // |S| values at the six dB ranks on entry of the dB map, their dB on exit
    float gamma[3];
    unsigned long long n_pos;
    int overflow;
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

// add 1 to h[bin] for every lane with pred (all 32 lanes call this together).  The lanes that share the
// first active lane's bin are counted by one atomic — whole warps of equal values (the exact zeros of empty
// frames, the floor value of the dB image) cost one shared atomic — the others add themselves.
__device__ __forceinline__ void warp_hist_add(unsigned int *h, unsigned int bin, bool pred) {
    const unsigned int active = __ballot_sync(0xffffffffu, pred);
    if (active == 0) return;
    const int leader = __ffs(active) - 1;
    const unsigned int b0 = __shfl_sync(0xffffffffu, bin, leader);
    const unsigned int same = __ballot_sync(0xffffffffu, pred && bin == b0);
    if ((int)(threadIdx.x & 31) == leader) atomicAdd(&h[b0], __popc(same));
    else if (pred && bin != b0) atomicAdd(&h[bin], 1u);
}

struct DbExtra {
    float v[B200VEC_MAX_RANKS];
    int n;
};
// ex: a few more values converted by the same instruction sequence (block 0) — the order statistics of |S| whose
// dB are the order statistics of the dB image
__global__ void __launch_bounds__(kSelThreads)
db_kernel(const float *__restrict__ s, long long n, float floor_val, float *__restrict__ out,
          const DbExtra ex, float *__restrict__ ex_out, K5Out *__restrict__ k5) {
    if (k5) {                                     // floor and the six values were formed on the device (k5_floor_kernel)
        if (k5->overflow) return;
        floor_val = k5->floor_v;
        if (blockIdx.x == 0 && threadIdx.x < 6) k5->six[threadIdx.x] = 10.0f * log10f(fabsf(k5->six[threadIdx.x]) + floor_val);
    } else if (blockIdx.x == 0 && (int)threadIdx.x < ex.n) {
        ex_out[threadIdx.x] = 10.0f * log10f(fabsf(ex.v[threadIdx.x]) + floor_val);
    }
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long i0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if ((((unsigned long long)s | (unsigned long long)out) & 15ull) == 0) {       // 16-byte path, streaming loads and stores
        const long long n4 = n / 4;
        for (long long i = i0; i < n4; i += stride) {
            const float4 a = __ldcs(reinterpret_cast<const float4 *>(s) + i);
            float4 r;
            r.x = 10.0f * log10f(fabsf(a.x) + floor_val); r.y = 10.0f * log10f(fabsf(a.y) + floor_val);
            r.z = 10.0f * log10f(fabsf(a.z) + floor_val); r.w = 10.0f * log10f(fabsf(a.w) + floor_val);
            __stcs(reinterpret_cast<float4 *>(out) + i, r);
        }
        for (long long i = 4 * n4 + i0; i < n; i += stride) out[i] = 10.0f * log10f(fabsf(__ldg(s + i)) + floor_val);
    } else {
        for (long long i = i0; i < n; i += stride) out[i] = 10.0f * log10f(fabsf(__ldg(s + i)) + floor_val);
    }
}

// four consecutive elements starting at 4 * g (zeros past n); one 16-byte load when x is 16-byte aligned
__device__ __forceinline__ void sel_load4(const float *__restrict__ x, long long n, long long g, bool aligned,
                                          float (&v)[4], bool (&in)[4]) {
    const long long i = 4 * g;
#pragma unroll
    for (int k = 0; k < 4; ++k) { v[k] = 0.f; in[k] = i + k < n; }
    if (aligned && i + 3 < n) {
        const float4 t = __ldg(reinterpret_cast<const float4 *>(x) + g);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) if (in[k]) v[k] = __ldg(x + i + k);
    }
}

constexpr int kSelUnroll = 4;      // 16-byte loads in flight per thread

// level 1: histogram of the top 11 key bits over all selected elements.  h: kBins1 words of shared memory.
__device__ __forceinline__ void sel_hist1_body(const float *__restrict__ x, long long n, int flags,
                                               unsigned int *__restrict__ hist, unsigned int *h) {
    const bool positive_only = flags & 1, use_abs = flags & 2;
    for (int i = threadIdx.x; i < kBins1; i += blockDim.x) h[i] = 0;
    __syncthreads();
    const bool aligned = (reinterpret_cast<unsigned long long>(x) & 15ull) == 0;
    const long long groups = (n + 3) / 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long g0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = stride * kSelUnroll;
    const long long g_round = ((groups + step - 1) / step) * step;           // keep warps converged for the ballots
    for (long long g = g0; g < g_round; g += step) {
        float v[kSelUnroll][4]; bool in[kSelUnroll][4];
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) sel_load4(x, n, g + u * stride, aligned, v[u], in[u]);
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u)
#pragma unroll
            for (int k = 0; k < 4; ++k)
            {
                const float val = use_abs ? fabsf(v[u][k]) : v[u][k];
                warp_hist_add(h, f2key(val) >> 21, in[u][k] && (positive_only ? val > 0.f : true));
            }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kBins1; i += blockDim.x)
        if (h[i]) atomicAdd(&hist[kHist1Off + i], h[i]);
}

__global__ void __launch_bounds__(kSelThreads)
sel_hist1(const float *__restrict__ x, long long n, int flags, unsigned int *__restrict__ hist) {
    __shared__ unsigned int h[kBins1];
    sel_hist1_body(x, n, flags, hist, h);
}

// levels 2 and 3: histogram of the next bits for the elements matching each rank's prefix.
// A 2048-entry table over the top 11 key bits says which ranks an element can belong to, so the
// common case (none) costs one shared load per element.
// hs: [n_ranks][BINS] words, lut: 2048 bytes, spref: B200VEC_MAX_RANKS words — all shared memory.  The state is
// read through volatile loads: inside the one-launch select (sel_small_kernel) it was written by another block
// earlier in the same kernel.
template <int LEVEL>
__device__ __forceinline__ void sel_histN_body(const float *__restrict__ x, long long n, int flags,
                                               unsigned int *__restrict__ hist, unsigned int *hs, unsigned char *lut,
                                               unsigned int *spref) {
    const bool positive_only = flags & 1, use_abs = flags & 2;
    constexpr int BINS = (LEVEL == 2) ? kBins2 : kBins3;
    constexpr int SHIFT_PREFIX = (LEVEL == 2) ? 21 : 10;
    const volatile SelState *st = reinterpret_cast<const volatile SelState *>(hist + kStateOff);
    const int nr = st->n_ranks;
    unsigned int pref[B200VEC_MAX_RANKS];
    unsigned int first = 0;                       // ranks whose prefix no earlier rank has: they own a histogram
#pragma unroll
    for (int r = 0; r < B200VEC_MAX_RANKS; ++r) {
        pref[r] = (r < nr) ? st->prefix[r] : 0xffffffffu;
        bool dup = r >= nr;
        for (int q = 0; q < r; ++q) dup |= (pref[q] == pref[r]);
        if (!dup) first |= 1u << r;
    }
    static_assert(B200VEC_MAX_RANKS <= 8, "one table byte holds one bit per rank");
    for (int i = threadIdx.x; i < nr * BINS; i += blockDim.x) hs[i] = 0;
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) lut[i] = 0;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int r = 0; r < B200VEC_MAX_RANKS; ++r) spref[r] = pref[r];
    if (threadIdx.x == 0)
        for (int r = 0; r < B200VEC_MAX_RANKS; ++r)
            if ((first >> r) & 1u) lut[(LEVEL == 2) ? pref[r] : (pref[r] >> 11)] |= (unsigned char)(1u << r);
    __syncthreads();
    const bool aligned = (reinterpret_cast<unsigned long long>(x) & 15ull) == 0;
    const long long groups = (n + 3) / 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long g0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long step = stride * kSelUnroll;
    const long long g_round = ((groups + step - 1) / step) * step;
    for (long long g = g0; g < g_round; g += step) {
        float v[kSelUnroll][4]; bool in[kSelUnroll][4];
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) sel_load4(x, n, g + u * stride, aligned, v[u], in[u]);
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float val = use_abs ? fabsf(v[u][k]) : v[u][k];
                const bool ok = in[u][k] && (positive_only ? val > 0.f : true);
                const unsigned int key = f2key(val);
                unsigned int mine = ok ? lut[key >> 21] : 0u;        // candidate histograms of this element
                if (!__any_sync(0xffffffffu, mine != 0)) continue;   // most elements match no prefix
                const unsigned int top = key >> SHIFT_PREFIX;
                const unsigned int bin = (LEVEL == 2) ? ((key >> 10) & 0x7ffu) : (key & 0x3ffu);
                for (unsigned int f = first; f; f &= f - 1) {        // warp-uniform: the ranks owning a histogram
                    const int r = __ffs(f) - 1;
                    warp_hist_add(hs + r * BINS, bin, ((mine >> r) & 1u) && top == spref[r]);
                }
            }
        }
    }
    __syncthreads();
    const int off = (LEVEL == 2) ? kHist2Off : kHist3Off;
    for (int i = threadIdx.x; i < nr * BINS; i += blockDim.x)
        if (hs[i]) atomicAdd(&hist[off + i], hs[i]);
}

template <int LEVEL>
__global__ void __launch_bounds__(kSelThreads)
sel_histN(const float *__restrict__ x, long long n, int flags, unsigned int *__restrict__ hist) {
    extern __shared__ unsigned int hs[];          // [n_ranks][BINS]
    __shared__ unsigned char lut[2048];
    __shared__ unsigned int spref[B200VEC_MAX_RANKS];
    sel_histN_body<LEVEL>(x, n, flags, hist, hs, lut, spref);
}

// one block of kPickThreads: block-wide prefix sum over the level histogram of each rank, then the
// bin holding the rank refines prefix/resid (first bin b with cum(b) + h[b] > want; the last bin if none)
constexpr int kPickThreads = 1024;
struct PickShared {
    unsigned int old[B200VEC_MAX_RANKS];
};
// Warp r refines rank r: lane l sums bins [l * PER, (l + 1) * PER) of the rank's level histogram, a warp scan
// finds the lane whose range holds the rank, that lane walks its bins (first bin b with cum(b) + h[b] > want;
// the last bin if none).  All ranks at once, two block barriers in total.
template <int LEVEL>
__device__ __forceinline__ void sel_pick_body(unsigned int *__restrict__ hist, PickShared &sh) {
    volatile SelState *st = reinterpret_cast<volatile SelState *>(hist + kStateOff);
    constexpr int BINS = (LEVEL == 1) ? kBins1 : (LEVEL == 2 ? kBins2 : kBins3);
    constexpr int BITS = (LEVEL == 3) ? 10 : 11;
    constexpr int PER = BINS / 32;
    const int t = threadIdx.x, lane = t & 31, r = t >> 5;
    const int nr = st->n_ranks;
    if (t < B200VEC_MAX_RANKS) sh.old[t] = st->prefix[t];
    __syncthreads();
    const int loops = (LEVEL == 1 && nr == 0) ? 1 : nr;     // level 1 also counts the selected elements
    if (r < loops) {
        int src = r;                       // ranks with an identical prefix share the first one's histogram
        if (LEVEL != 1)
            for (int q = r - 1; q >= 0; --q)
                if (sh.old[q] == sh.old[r]) src = q;
        const unsigned int *h = (LEVEL == 1) ? hist + kHist1Off
                              : (LEVEL == 2) ? hist + kHist2Off + src * kBins2
                                             : hist + kHist3Off + src * kBins3;
        const uint4 *h4 = reinterpret_cast<const uint4 *>(h + lane * PER);
        unsigned long long s = 0;
#pragma unroll 4
        for (int k = 0; k < PER / 4; ++k) {
            const uint4 c = __ldcg(h4 + k);
            s += (unsigned long long)c.x + c.y + c.z + c.w;
        }
        unsigned long long inc = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long up = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += up;
        }
        const unsigned long long total = __shfl_sync(0xffffffffu, inc, 31);
        const unsigned long long want = (r < nr) ? st->resid[r] : 0ull;
        const unsigned long long before = inc - s;
        const bool mine = before <= want && want < inc;
        const unsigned int owners = __ballot_sync(0xffffffffu, mine);
        if (LEVEL == 1 && r == 0 && lane == 0) st->count = total;
        if (r < nr) {
            if (owners == 0) {                               // rank past the end: the last bin
                if (lane == 31) {
                    const unsigned int last = __ldcg(h + BINS - 1);
                    st->resid[r] = want - (total - last);
                    st->prefix[r] = (LEVEL == 1) ? (unsigned)(BINS - 1) : ((sh.old[r] << BITS) | (unsigned)(BINS - 1));
                    if (LEVEL == 3) st->result[r] = st->prefix[r];
                }
            } else {
                // the owning lane's PER bins, re-read by the whole warp (PER / 32 bins per lane): one L2 round trip
                const int L = __ffs(owners) - 1;
                const unsigned long long before_l = __shfl_sync(0xffffffffu, before, L);
                constexpr int Q = PER / 32;
                unsigned int c[Q];
                unsigned long long s2 = 0;
#pragma unroll
                for (int k = 0; k < Q; ++k) { c[k] = __ldcg(h + L * PER + lane * Q + k); s2 += c[k]; }
                unsigned long long inc2 = s2;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const unsigned long long up = __shfl_up_sync(0xffffffffu, inc2, o);
                    if (lane >= o) inc2 += up;
                }
                unsigned long long cum = before_l + inc2 - s2;
                if (cum <= want && want < cum + s2) {
                    int b = L * PER + lane * Q + Q - 1;
#pragma unroll
                    for (int k = 0; k < Q; ++k) {
                        if (want < cum + c[k]) { b = L * PER + lane * Q + k; break; }
                        cum += c[k];
                    }
                    st->resid[r] = want - cum;
                    st->prefix[r] = (LEVEL == 1) ? (unsigned)b : ((sh.old[r] << BITS) | (unsigned)b);
                    if (LEVEL == 3) st->result[r] = st->prefix[r];
                }
            }
        }
    }
    __syncthreads();
}

template <int LEVEL>
__global__ void __launch_bounds__(kPickThreads)
sel_pick(unsigned int *__restrict__ hist) {
    __shared__ PickShared sh;
    sel_pick_body<LEVEL>(hist, sh);
}

// The whole 3-level select in ONE cooperative launch (candidate arrays, previews: inputs whose three reads cost
// less than the six dependent launches between them): blocks of kPickThreads threads, the histogram passes
// separated by grid barriers, block 0 picks the bins.  hist[kBarrierOff] is the barrier counter (zeroed by the
// caller with the rest of the scratch).
constexpr int kBarrierOff = kStateOff + 64 + 4 * B200VEC_MAX_RANKS;
__device__ __forceinline__ void sel_grid_sync(unsigned int *ctr, unsigned int &phase) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(ctr, 1u);
        ++phase;
        while (*reinterpret_cast<volatile unsigned int *>(ctr) < phase * gridDim.x) __nanosleep(20);
        __threadfence();
    }
    __syncthreads();
}
__global__ void __launch_bounds__(kPickThreads, 1)
sel_small_kernel(const float *__restrict__ x, long long n, int flags, unsigned int *__restrict__ hist) {
    extern __shared__ unsigned int hs[];          // [max(kBins1, n_ranks * kBins2)]
    __shared__ unsigned char lut[2048];
    __shared__ unsigned int spref[B200VEC_MAX_RANKS];
    __shared__ PickShared sh;
    unsigned int phase = 0;
    unsigned int *ctr = hist + kBarrierOff;
    const int nr = reinterpret_cast<const volatile SelState *>(hist + kStateOff)->n_ranks;
    sel_hist1_body(x, n, flags, hist, hs);
    sel_grid_sync(ctr, phase);
    if (blockIdx.x == 0) sel_pick_body<1>(hist, sh);
    if (nr == 0) return;
    sel_grid_sync(ctr, phase);
    sel_histN_body<2>(x, n, flags, hist, hs, lut, spref);
    sel_grid_sync(ctr, phase);
    if (blockIdx.x == 0) sel_pick_body<2>(hist, sh);
    sel_grid_sync(ctr, phase);
    sel_histN_body<3>(x, n, flags, hist, hs, lut, spref);
    sel_grid_sync(ctr, phase);
    if (blockIdx.x == 0) sel_pick_body<3>(hist, sh);
}

// ---------------------------------------------------------------------------------------
// Fast path of the order statistics: ONE full read for a 13-bit top histogram (+ the exact count of +0,
// the one value real spectrograms repeat by the tens of millions — it never touches an atomic), the host
// picks the bin of every rank, ONE more full read compacts the few elements of those bins into a candidate
// array, and the exact 3-level select above runs on the candidates only.  Falls back to the 3-level select
// on the whole array when the bins in question hold too many elements (a heavily repeated non-zero value).
constexpr int kTopBits = 13;
constexpr int kTopBins = 1 << kTopBits;                // sign + 8 exponent bits + 4 mantissa bits: 0.19 dB per bin
constexpr int kTopShift = 32 - kTopBits;
constexpr unsigned int kTopZeroBin = 0x80000000u >> kTopShift;      // first bin of the non-negative keys

// bin of the 13-bit top histogram; ABS: of |x| (sign bit cleared: always in the non-negative half)
template <bool ABS>
__device__ __forceinline__ unsigned int top_bin(unsigned int u) {
    if constexpr (ABS) return ((u & 0x7fffffffu) >> kTopShift) | kTopZeroBin;
    else return ((u & 0x80000000u) ? ~u : (u | 0x80000000u)) >> kTopShift;
}
template <bool ABS>
__device__ __forceinline__ bool is_poszero(unsigned int u) { return ABS ? (u & 0x7fffffffu) == 0u : u == 0u; }

// The hot loop handles whole 16-byte groups (aligned base, unrolled four deep, no per-element bounds logic: 2-3
// instructions per element next to the shared atomic); block 0 sweeps up the ragged ends element by element.
template <bool ABS>
__global__ void __launch_bounds__(kSelThreads)
sel_top_kernel(const float *__restrict__ x, long long n, unsigned int *__restrict__ hist,
               unsigned long long *__restrict__ n_poszero) {
    __shared__ unsigned int h[kTopBins];
    for (int i = threadIdx.x; i < kTopBins; i += kSelThreads) h[i] = 0;
    __syncthreads();
    // leading elements up to the first 16-byte boundary, whole groups, trailing elements
    long long head = (16 - (long long)(reinterpret_cast<unsigned long long>(x) & 15ull)) / 4 % 4;
    if (head > n) head = n;
    const long long groups = (n - head) / 4;
    const uint4 *x4 = reinterpret_cast<const uint4 *>(x + head);
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned int zc = 0;
    auto one = [&](unsigned int u) {
        if (is_poszero<ABS>(u)) ++zc;
        else atomicAdd(&h[top_bin<ABS>(u)], 1u);
    };
    long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; g + (kSelUnroll - 1) * stride < groups; g += stride * kSelUnroll) {
        uint4 v[kSelUnroll];
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) v[u] = __ldcs(x4 + g + u * stride);
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) {
            // whole groups of +0 (the empty frames of a spectrogram: 42 % of BASELINE cfg3) cost one OR and one compare
            if (((v[u].x | v[u].y | v[u].z | v[u].w) & (ABS ? 0x7fffffffu : 0xffffffffu)) == 0u) { zc += 4; continue; }
            one(v[u].x); one(v[u].y); one(v[u].z); one(v[u].w);
        }
    }
    for (; g < groups; g += stride) {
        const uint4 v = __ldcs(x4 + g);
        one(v.x); one(v.y); one(v.z); one(v.w);
    }
    if (blockIdx.x == 0) {
        const unsigned int *xu = reinterpret_cast<const unsigned int *>(x);
        for (long long i = threadIdx.x; i < head; i += kSelThreads) one(xu[i]);
        for (long long i = head + 4 * groups + threadIdx.x; i < n; i += kSelThreads) one(xu[i]);
    }
    zc = __reduce_add_sync(0xffffffffu, zc);
    if ((threadIdx.x & 31) == 0 && zc) atomicAdd(n_poszero, (unsigned long long)zc);
    __syncthreads();
    for (int i = threadIdx.x; i < kTopBins; i += kSelThreads)
        if (h[i]) atomicAdd(&hist[i], h[i]);
}

// cand[...] <- every element (after the optional abs) whose top bits name one of `sel`'s bins, +0 excluded.
// A byte table over the 13-bit bins answers "selected?" with one shared load per element; groups of four +0
// are skipped with one OR.  Candidates collect in a per-warp shared-memory buffer (positions from ballots, no
// atomics) that is flushed to `cand` in coalesced bursts, one global atomic per flush — a same-address atomic
// per group of four, as first written, was what bounded this pass (219 us on the 112 M-element map).
constexpr int kCompactBuf = 512;                      // floats per warp
constexpr int kCompactFlush = kCompactBuf - 128;      // flush when a further group (<= 128 hits) might not fit
// While it is at it the pass counts the candidates' next key bits: a 13-bit bin fixes key bits 31..19, the
// 9 bits below (18..10) index a per-bin shared histogram that is flushed into the level-2 histogram layout of
// the 3-level select (hist2[slot][key bits 20..10]) — the exact select on the candidates then starts at
// level 2 (sel_cand_kernel): one read of the candidate array instead of three.
template <bool ABS>
__global__ void __launch_bounds__(kSelThreads)
sel_compact_kernel(const float *__restrict__ x, long long n, const SelBins sel_arg,
                   float *__restrict__ cand, unsigned long long cap, unsigned long long *__restrict__ count,
                   unsigned int *__restrict__ hist, const K5Plan *__restrict__ plan) {
    // plan != NULL: the bins were picked on the device (k5_plan_kernel)
    if (plan && plan->overflow) return;
    __shared__ SelBins sel;
    if (threadIdx.x == 0) sel = plan ? plan->sel : sel_arg;
    __shared__ unsigned char lut[kTopBins];
    __shared__ float wbuf[kSelThreads / 32][kCompactBuf];
    __shared__ unsigned int h2[B200VEC_MAX_RANKS][512];
    for (int i = threadIdx.x; i < kTopBins / 4; i += kSelThreads) reinterpret_cast<unsigned int *>(lut)[i] = 0u;
    for (int i = threadIdx.x; i < B200VEC_MAX_RANKS * 512; i += kSelThreads) (&h2[0][0])[i] = 0u;
    __syncthreads();
    if (threadIdx.x < sel.n) lut[sel.bin[threadIdx.x]] = (unsigned char)(threadIdx.x + 1);
    __syncthreads();
    long long head = (16 - (long long)(reinterpret_cast<unsigned long long>(x) & 15ull)) / 4 % 4;
    if (head > n) head = n;
    const long long groups = (n - head) / 4;
    const uint4 *x4 = reinterpret_cast<const uint4 *>(x + head);
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int lane = threadIdx.x & 31;
    const unsigned int lt = (1u << lane) - 1u;
    float *buf = wbuf[threadIdx.x >> 5];
    int fill = 0;                                     // warp-uniform
    // 0: not a candidate, else 1 + index of the element's bin in sel.bin
    auto hit = [&](unsigned int u) -> unsigned int { return is_poszero<ABS>(u) ? 0u : lut[top_bin<ABS>(u)]; };
    auto val = [&](unsigned int u) { return __uint_as_float(ABS ? (u & 0x7fffffffu) : u); };
    auto key_of = [&](unsigned int u) {
        if (ABS) return (u & 0x7fffffffu) | 0x80000000u;
        return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    };
    auto flush = [&]() {                              // all lanes
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(count, (unsigned long long)fill);
        base = __shfl_sync(0xffffffffu, base, 0);
        __syncwarp();
        for (int i = lane; i < fill; i += 32)
            if (base + i < cap) cand[base + i] = buf[i];
        __syncwarp();
        fill = 0;
    };
    auto put = [&](unsigned int u, unsigned int h) {  // all lanes: one element each (h: 0 or 1 + bin index)
        const unsigned int m = __ballot_sync(0xffffffffu, h != 0u);
        if (h) {
            buf[fill + __popc(m & lt)] = val(u);
            atomicAdd(&h2[h - 1][(key_of(u) >> 10) & 0x1ffu], 1u);
        }
        fill += __popc(m);
    };
    auto group = [&](const uint4 &v) {                // all lanes: four elements each
        const bool live = ((v.x | v.y | v.z | v.w) & (ABS ? 0x7fffffffu : 0xffffffffu)) != 0u;
        if (!__any_sync(0xffffffffu, live)) return;
        put(v.x, live ? hit(v.x) : 0u); put(v.y, live ? hit(v.y) : 0u); put(v.z, live ? hit(v.z) : 0u); put(v.w, live ? hit(v.w) : 0u);
        if (fill > kCompactFlush) flush();
    };
    // warp-uniform trip counts: every lane of a warp runs the same number of iterations
    const long long w0 = ((long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31));
    long long g = w0 + lane;
    for (long long gw = w0; gw < groups; gw += stride * kSelUnroll, g += stride * kSelUnroll) {
        uint4 v[kSelUnroll];
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) {
            const long long gg = g + u * stride;
            v[u] = gg < groups ? __ldcs(x4 + gg) : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int u = 0; u < kSelUnroll; ++u) group(v[u]);
    }
    if (blockIdx.x == 0 && threadIdx.x < 32) {                // ragged ends: one warp, element by element
        const unsigned int *xu = reinterpret_cast<const unsigned int *>(x);
        const long long tail0 = head + 4 * groups;
        const long long n_rag = head + (n - tail0);
        for (long long i0 = 0; i0 < n_rag; i0 += 32) {
            const long long i = i0 + lane;
            unsigned int u = 0;
            if (i < n_rag) u = xu[i < head ? i : tail0 + (i - head)];
            put(u, i < n_rag ? hit(u) : 0u);
        }
    }
    if (fill > 0) flush();
    __syncthreads();
    for (int i = threadIdx.x; i < sel.n * 512; i += kSelThreads) {
        const int j = i >> 9;
        const unsigned int c = h2[j][i & 511];
        if (c) atomicAdd(&hist[kHist2Off + sel.slot[j] * kBins2 + ((sel.bin[j] & 3u) << 9) + (i & 511)], c);
    }
}

// The exact select on a candidate array whose level-2 histogram the compaction pass has already built and
// whose state holds the 11-bit prefixes and the ranks within them: pick, one read for the last 10 bits, pick.
__global__ void __launch_bounds__(kPickThreads, 1)
sel_cand_kernel(const float *__restrict__ x, long long n, unsigned int *__restrict__ hist, const K5Plan *__restrict__ plan) {
    if (plan) {                                   // candidate count known on the device only
        if (plan->overflow) return;
        n = (long long)plan->pop;
    }
    extern __shared__ unsigned int hs[];          // [n_ranks * kBins3]
    __shared__ unsigned char lut[2048];
    __shared__ unsigned int spref[B200VEC_MAX_RANKS];
    __shared__ PickShared sh;
    unsigned int phase = 0;
    unsigned int *ctr = hist + kBarrierOff;
    if (blockIdx.x == 0) sel_pick_body<2>(hist, sh);
    sel_grid_sync(ctr, phase);
    sel_histN_body<3>(x, n, 0, hist, hs, lut, spref);
    sel_grid_sync(ctr, phase);
    if (blockIdx.x == 0) sel_pick_body<3>(hist, sh);
}

// ---------------------------------------------------------------------------------------
// normalize_spectrogram without host round trips (b200vec_normalize_db): what the host did between the passes —
// count the positives, numpy's rank arithmetic (npquant.h), pick every rank's 13-bit bin, set up the candidate
// select, form the floor — runs in two tiny kernels, so the whole chain (histogram, plan, compaction, candidate
// select, floor, dB map) is queued at once and the host reads one small record at the end.
// one block: hist = the 13-bit histogram of |S| (+0 excluded), *n_poszero its +0 count
__global__ void __launch_bounds__(1024)
k5_plan_kernel(const unsigned int *__restrict__ hist13, const unsigned long long *__restrict__ n_poszero, long long n,
               double low_q, double high_q, unsigned long long cap, unsigned int *__restrict__ hist, K5Plan *__restrict__ plan) {
    __shared__ unsigned long long cum[kTopBins - kTopZeroBin + 1];       // exclusive prefix sums over the non-negative bins
    __shared__ unsigned long long wsum[32];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    constexpr int NB = kTopBins - kTopZeroBin;                            // 4096
    constexpr int PER = NB / 1024;                                        // 4
    unsigned int c[PER];
    unsigned long long s = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) { c[k] = hist13[kTopZeroBin + t * PER + k]; s += c[k]; }
    unsigned long long inc = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long up = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += up;
    }
    if (lane == 31) wsum[w] = inc;
    __syncthreads();
    if (w == 0) {
        unsigned long long x = wsum[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long up = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += up;
        }
        wsum[lane] = x;
    }
    __syncthreads();
    unsigned long long run = inc - s + (w ? wsum[w - 1] : 0ull);
#pragma unroll
    for (int k = 0; k < PER; ++k) { cum[t * PER + k] = run; run += c[k]; }
    if (t == 1023) cum[NB] = run;
    __syncthreads();
    if (t != 0) return;
    const unsigned long long n_pos = cum[NB], n_zero = *n_poszero;
    const PctRanks r1 = pct_ranks_f32(n, low_q), r2 = pct_ranks_f32(n, high_q);
    PctRanks r0 = {0, 0, 0.0f};
    long long ranks[B200VEC_MAX_RANKS] = {r1.lo, r1.hi, r2.lo, r2.hi, 0, n - 1, 0, 0};
    int nr = 6;
    if (n_pos > 0) {
        r0 = pct_ranks_f32((long long)n_pos, 5.0);
        ranks[6] = n - (long long)n_pos + r0.lo; ranks[7] = n - (long long)n_pos + r0.hi;   // the positives are the top n_pos of |S|
        nr = 8;
    }
    K5Plan p;
    p.n_pos = n_pos; p.n_ranks = nr; p.gamma[0] = r1.gamma; p.gamma[1] = r2.gamma; p.gamma[2] = r0.gamma;
    p.sel.n = 0; p.overflow = 0;
    int bin_of[B200VEC_MAX_RANKS];
    unsigned long long resid[B200VEC_MAX_RANKS];
    for (int r = 0; r < nr; ++r) {
        unsigned long long want = (unsigned long long)ranks[r];
        bin_of[r] = -1; resid[r] = 0;
        if (want < n_zero) continue;                                     // inside the +0 block
        want -= n_zero;
        int lo = 0, hi = NB;                                             // last b with cum[b] <= want
        while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (cum[mid] <= want) lo = mid; else hi = mid; }
        bin_of[r] = (int)kTopZeroBin + lo; resid[r] = want - cum[lo];
        bool seen = false;
        for (int q = 0; q < p.sel.n; ++q) seen = seen || p.sel.bin[q] == (unsigned)bin_of[r];
        if (!seen) p.sel.bin[p.sel.n++] = (unsigned)bin_of[r];
    }
    for (int i = 1; i < p.sel.n; ++i)                                    // ascending bins = ascending values
        for (int j = i; j > 0 && p.sel.bin[j - 1] > p.sel.bin[j]; --j) { const unsigned x = p.sel.bin[j]; p.sel.bin[j] = p.sel.bin[j - 1]; p.sel.bin[j - 1] = x; }
    p.pop = 0;
    auto cnt_of = [&](unsigned int bin) { return cum[bin - kTopZeroBin + 1] - cum[bin - kTopZeroBin]; };   // bin counts out of shared memory
    for (int i = 0; i < p.sel.n; ++i) p.pop += cnt_of(p.sel.bin[i]);
    if (p.pop > cap) p.overflow = 1;
    // the candidate select starts at level 2 (sel_compact_kernel builds those histograms): 11-bit prefix + rank within it
    SelState sloc;                                                      // built locally, written once
    SelState *st = &sloc;
    for (int r = 0; r < B200VEC_MAX_RANKS; ++r) { sloc.prefix[r] = 0u; sloc.resid[r] = 0ull; sloc.result[r] = 0u; }
    int n_sub = 0;
    for (int r = 0; r < nr; ++r) {
        p.sub_of[r] = -1;
        if (bin_of[r] < 0) continue;
        const unsigned int p11 = (unsigned)bin_of[r] >> 2;
        unsigned long long want = resid[r];
        for (int i = 0; i < p.sel.n; ++i)
            if ((p.sel.bin[i] >> 2) == p11 && p.sel.bin[i] < (unsigned)bin_of[r]) want += cnt_of(p.sel.bin[i]);
        st->prefix[n_sub] = p11; st->resid[n_sub] = want; st->result[n_sub] = 0u;
        p.sub_of[r] = n_sub++;
    }
    for (int r = nr; r < B200VEC_MAX_RANKS; ++r) p.sub_of[r] = -1;
    st->n_ranks = n_sub; st->count = p.pop;
    for (int i = 0; i < p.sel.n; ++i) {
        p.sel.slot[i] = 0;
        for (int q = 0; q < n_sub; ++q)
            if (st->prefix[q] == (p.sel.bin[i] >> 2)) { p.sel.slot[i] = q; break; }
    }
    *reinterpret_cast<SelState *>(hist + kStateOff) = sloc;
    *plan = p;
}

// one thread: the selected keys -> |S| values, the floor (utils.py:365-366) -> out (the dB map converts out->six)
__global__ void k5_floor_kernel(const unsigned int *__restrict__ hist, const K5Plan *__restrict__ plan, K5Out *__restrict__ out) {
    const SelState *st = reinterpret_cast<const SelState *>(hist + kStateOff);
    float v[B200VEC_MAX_RANKS];
    for (int r = 0; r < B200VEC_MAX_RANKS; ++r) {
        const int i = plan->sub_of[r];
        v[r] = i < 0 ? 0.0f : key2f(st->result[i]);
    }
    float floor_v = plan->n_pos > 0 ? lerp_f32(v[6], v[7], plan->gamma[2]) : 1e-12f;
    if (1e-12f > floor_v) floor_v = 1e-12f;                              // max(floor, 1e-12)
    out->floor_v = floor_v;
    for (int i = 0; i < 6; ++i) out->six[i] = v[i];
    for (int i = 0; i < 3; ++i) out->gamma[i] = plan->gamma[i];
    out->n_pos = plan->n_pos; out->overflow = plan->overflow;
}

// out[i] = in[i * stride]: the naive decimation of create_spectrogram (utils.py:192-195, sig[::factor])
// made once, so that the overlapping preview frames re-read a compact, L2-resident copy instead of one
// 32-byte sector per sample and frame.
__global__ void __launch_bounds__(256)
decimate_kernel(const float2 *__restrict__ in, long long n_out, long long stride, float2 *__restrict__ out) {
    const long long st = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_out; i += st)
        out[i] = __ldcs(in + i * stride);
}

}  // namespace b200vec
