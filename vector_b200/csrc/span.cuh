// span.cuh — K2/K3 (and the cfg5 batch form): output-stationary packet placement with fused max / scale,
// one WARP per 512-sample span — the structure of the fused step's builder (wf1024.cuh) without the
// transforms, for one vector or for a batch of independent vectors.
//
// Replaces the reference's np.zeros + `vector[pos:pos+L] += y` loop + normalisation
// (unified_gui.py:1712,1757-1769,1778-1782; main.py:218,253-257,274-277), per vector; a batch is the
// scenario sweep that runs that method once per scenario (BASELINE.json configs[4]).
//
// Work item = kSpChunk consecutive halves (512 samples each) of one vector, handed out by a global
// counter.  Per item a warp
//   S  finds, per packet (lane p), the instance range overlapping the item (two 64-bit floor divisions);
//   B  per half: lane l owns x[l + 32 j]; for every overlapping (packet, instance) in the reference's
//      order it adds the packet samples (8-byte loads, 256 B per warp instruction, from the L2-resident
//      bank), scales by fl32(1 / fl32(sqrt(max |v|^2))) and sends the 4 KB half to HBM as ONE bulk copy
//      (cp.async.bulk shared -> global) out of the warp's staging block; halves nothing overlaps leave
//      from a block of zeros.  Every output sample is written exactly once; no atomics on data.
// Normalisation: MAXONLY reduces max |v|^2 (exact fp64 products) per vector with no stores; SCALE reads
// it; NORM does both in one cooperative launch (max pass, grid barrier, scale pass) — per vector, so a
// batch of 256 scenario vectors is ONE launch.  HBM traffic: 8 B per output sample, nothing read.
#pragma once
#include "common.cuh"
#include "place.cuh"
#include "wf1024.cuh"        // bank-load / bulk-store helpers shared with the fused step

namespace b200vec {

constexpr int kSpWarps = 8;
constexpr int kSpThreads = 32 * kSpWarps;
constexpr int kSpChunk = 4;                                       // halves per work item (single vector)
constexpr int kSpChunkBatch = 16;                                 // batch: the per-item descriptor fetch (device memory) is amortised over 64 KB
constexpr int kSpNorm = 3;                                        // SpanArgs::mode: max pass + grid barrier + scale pass
constexpr size_t kSpSmem = (size_t)kSpWarps * 4096 + 4096;        // per-warp staging blocks + one block of zeros

struct SpanArgs {
    PlaceArgs place;             // single vector: descriptors by value (constant bank); place.lo/hi unused
    const PlaceDesc *gdesc;      // batch: [n_vec][pk_stride] descriptors in device memory (NULL: single vector)
    const int *gnp;              // batch: packets per vector
    int n_vec, pk_stride;
    long long lo, hi;            // every vector: samples [lo, hi) are built
    float2 *out;                 // vector v at out + v * out_ld; element 0 = sample lo
    long long out_ld;
    unsigned long long *maxsq;   // [n_vec] bits of max |v|^2 (double): MAXONLY / NORM raise it, SCALE / NORM read it
    int mode;                    // B200VEC_BUILD_STORE | _MAXONLY | _SCALE | kSpNorm
    const MaxSpan *spans;        // MAXONLY / NORM: the spans evaluated (maxspans.cuh), sorted by vector
    long long n_spans;
    unsigned int *ctr;           // [0] next work item, [2] grid barrier; zeroed on the stream before every launch
    int chunk;                   // halves per work item (set by launch_span)
};

template <bool BATCH>
__device__ __forceinline__ const PlaceDesc &sp_desc(const SpanArgs &a, const PlaceDesc *dv, int p) {
    if constexpr (BATCH) return dv[p];
    else return a.place.d[p];
}

// S: instances of packet `lane` overlapping [s0, s0 + span)
template <bool BATCH>
__device__ __forceinline__ void sp_overlap(const SpanArgs &a, const PlaceDesc *dv, int np, int lane, long long s0,
                                           int span, int &cnt_l, int &rel0_l) {
    cnt_l = 0; rel0_l = 0;
    if (lane < np) {
        long long start, period, count;
        int len;
        if constexpr (BATCH) {               // read-only path (L1): every warp of an SM reads the same few vectors' descriptors
            start = __ldg(&dv[lane].start); period = __ldg(&dv[lane].period);        // 56-byte records: 8-byte loads
            count = __ldg(&dv[lane].count); len = __ldg(&dv[lane].len);
        } else {
            const PlaceDesc &d = a.place.d[lane];
            start = d.start; period = d.period; count = d.count; len = d.len;
        }
        long long klo = floor_div(s0 - len - start, period) + 1;
        long long khi = floor_div(s0 + span - 1 - start, period);
        if (klo < 0) klo = 0;
        if (khi > count - 1) khi = count - 1;
        const long long cn = khi - klo + 1;
        cnt_l = cn > 0 ? (int)(cn < 0x00ffffff ? cn : 0x00ffffff) : 0;
        rel0_l = cnt_l ? (int)(s0 - (start + klo * period)) : 0;
    }
}

// B: the 512 samples starting `off` samples into the item; false when nothing overlaps them (x == 0).
template <bool BATCH>
__device__ __forceinline__ bool sp_build_half(const SpanArgs &a, const PlaceDesc *dv, unsigned active, int cnt_l,
                                              int rel0_l, int off, int lane, float2 (&x)[16], unsigned long long keep) {
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = make_float2(0.f, 0.f);
    bool live = false;
    for (unsigned m = active; m; m &= m - 1) {
        const int p = __ffs(m) - 1;
        const int cnt = __shfl_sync(0xffffffffu, cnt_l, p);
        const int rel0 = __shfl_sync(0xffffffffu, rel0_l, p);
        int len;
        long long period;
        const float2 *ya;
        if constexpr (BATCH) {
            len = __ldg(&dv[p].len); period = __ldg(&dv[p].period);
            ya = reinterpret_cast<const float2 *>(__ldg(reinterpret_cast<const unsigned long long *>(&dv[p].a)));
        } else {
            const PlaceDesc &d = a.place.d[p];
            len = d.len; period = d.period; ya = d.a;
        }
        long long r64 = (long long)rel0 + off;               // packet index of the half's sample 0 for instance klo + i
        for (int i = 0; i < cnt; ++i, r64 -= period) {
            if (r64 + 512 <= 0) break;                       // this and every later instance start after the half
            if (r64 >= len) continue;                        // instance ended before the half
            const int r = (int)r64;
            live = true;
            const float2 *base = ya + r + lane;
            if (r >= 0 && r + 512 <= len) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 v = wf_ldg_keep(base + 32 * j, keep);
                    x[j] = __fadd2_rn(x[j], v);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const int n = r + lane + 32 * j;
                    if (n >= 0 && n < len) {
                        const float2 v = wf_ldg_keep(base + 32 * j, keep);
                        x[j] = __fadd2_rn(x[j], v);
                    }
                }
            }
        }
    }
    return live;
}

// (Measured: capping registers for 4 or 5 CTAs per SM instead of 3, or two staging blocks per warp, does not help —
// 102 -> 111 / 118 us at cfg2 store-only — although the warps mostly wait on the L2 gather.)
// CHUNK = halves per work item, a compile-time constant: with a run-time item size the single-vector store pass
// ran at 96 us instead of 86 us (cfg2).
template <bool BATCH, int CHUNK = kSpChunk>
__global__ void __launch_bounds__(kSpThreads)
span_kernel(const __grid_constant__ SpanArgs a) {
    extern __shared__ __align__(16) unsigned char sp_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2 *vst = reinterpret_cast<float2 *>(sp_smem) + warp * 512;
    const unsigned zero_u32 = smem_u32(sp_smem + (size_t)kSpWarps * 4096);
    for (int i = tid; i < 256; i += kSpThreads)
        reinterpret_cast<float4 *>(sp_smem + (size_t)kSpWarps * 4096)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    wf_fence_async();
    __syncthreads();
    const unsigned long long keep = wf_policy_evict_last();
    const unsigned long long pol = wf_policy_evict_first();
    const long long gw = (long long)blockIdx.x * kSpWarps + warp, nw = (long long)gridDim.x * kSpWarps;
    // The first work item of pass B is claimed here, ahead of the max pass: thousands of warps leaving the grid
    // barrier together would otherwise queue up at the one counter (an atomic per clock at its L2 slice) before
    // any of them stores a sample.
    unsigned next = 0;
    if (lane == 0 && a.mode != B200VEC_BUILD_MAXONLY) next = atomicAdd(a.ctr, 1u);

    if (a.mode == B200VEC_BUILD_MAXONLY || a.mode == kSpNorm) {
        // ---------------------------------------------------------------- pass A: max |v|^2 per vector, no stores
        // The list is sorted by vector and the warps walk it side by side (warp w takes spans w, w + nw, ...): at
        // any moment the whole GPU works on a handful of vectors, whose packets (3 x 449 KB each) stay in L2 —
        // a batch's bank is far larger than the L2 (582 shifted packets = 261 MB at cfg5), and one contiguous
        // run per warp spread the warps over all vectors at once (2.1 GB of DRAM reads per launch).
        double tmax = 0.0;                                   // running max of |v|^2 for vector cur_v (exact products in fp64)
        float fthr = 0.0f;                                   // fp32 pre-filter threshold
        int cur_v = -1;
        for (long long u = gw; u < a.n_spans; u += nw) {
            const int4 sraw = __ldg(reinterpret_cast<const int4 *>(a.spans + u));      // one 16-byte read-only load
            MaxSpan sp;
            sp.s0 = (long long)(((unsigned long long)(unsigned)sraw.y << 32) | (unsigned)sraw.x); sp.n = sraw.z; sp.vec = sraw.w;
            const int v = BATCH ? sp.vec : 0;
            if (v != cur_v) {
                if (cur_v >= 0) {
                    tmax = warp_max(tmax);
                    if (lane == 0 && tmax > 0.0) atomicMax(a.maxsq + cur_v, (unsigned long long)__double_as_longlong(tmax));
                }
                cur_v = v; tmax = 0.0; fthr = 0.0f;
            }
            const PlaceDesc *dv = BATCH ? a.gdesc + (long long)v * a.pk_stride : nullptr;
            const int np = BATCH ? __ldg(a.gnp + v) : a.place.np;
            int cnt_l, rel0_l;
            sp_overlap<BATCH>(a, dv, np, lane, sp.s0, 512, cnt_l, rel0_l);
            const unsigned active = __ballot_sync(0xffffffffu, cnt_l > 0);
            if (!active) continue;
            float2 x[16];
            if (!sp_build_half<BATCH>(a, dv, active, cnt_l, rel0_l, 0, lane, x, keep)) continue;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float q = x[j].x * x[j].x + x[j].y * x[j].y;
                if (lane + 32 * j < sp.n && q >= fthr) {             // samples past the span's end never enter the max
                    const double e = (double)x[j].x * (double)x[j].x + (double)x[j].y * (double)x[j].y;
                    if (e > tmax) { tmax = e; fthr = q * 0.99999f; }
                }
            }
        }
        if (cur_v >= 0) {
            tmax = warp_max(tmax);
            if (lane == 0 && tmax > 0.0) atomicMax(a.maxsq + cur_v, (unsigned long long)__double_as_longlong(tmax));
        }
        for (long long v = gw; v < a.n_vec; v += nw) {       // whole-packet maxima the skipped spans stand for
            const PlaceDesc *dv = BATCH ? a.gdesc + v * a.pk_stride : nullptr;
            const int np = BATCH ? a.gnp[v] : a.place.np;
            if (lane < np) {
                const PlaceDesc &d = sp_desc<BATCH>(a, dv, lane);
                if (d.free_inst && d.count > 0) {
                    const unsigned long long pm = *d.pmax;
                    if (pm) atomicMax(a.maxsq + v, pm);
                }
            }
        }
        if (a.mode == B200VEC_BUILD_MAXONLY) return;
        __syncthreads();
        if (tid == 0) {                                      // grid barrier (cooperative launch: every CTA is resident)
            __threadfence();
            atomicAdd(a.ctr + 2, 1u);
            while (*reinterpret_cast<volatile unsigned int *>(a.ctr + 2) < gridDim.x) __nanosleep(40);
            __threadfence();
        }
        __syncthreads();
    }

    // -------------------------------------------------------------------- pass B: place, scale, store
    const long long n = a.hi - a.lo;
    const long long halves = (n + 511) / 512;
    constexpr int kChunk = CHUNK;
    const long long ipv = (halves + kChunk - 1) / kChunk;                     // items per vector
    const long long n_items = ipv * a.n_vec;
    const bool scaled = a.mode == B200VEC_BUILD_SCALE || a.mode == kSpNorm;
    float scl = 1.0f;
    int scl_v = -1;
    for (;;) {
        const long long c = __shfl_sync(0xffffffffu, next, 0);
        if (c >= n_items) break;
        if (lane == 0) next = atomicAdd(a.ctr, 1u);         // the next item's index arrives while this one is processed
        const int v = BATCH ? (int)(c / ipv) : 0;
        const long long h0 = (BATCH ? c - (long long)v * ipv : c) * kChunk;       // first half of the item
        const int nh = (int)(halves - h0 < kChunk ? halves - h0 : kChunk);
        if (scaled && v != scl_v) {                          // once per vector, not per item: an L2 round trip + fp64 sqrt
            const double m2 = __longlong_as_double((long long)__ldcg(a.maxsq + v));
            scl = m2 > 0.0 ? 1.0f / (float)sqrt(m2) : 1.0f;  // numpy's complex64 / float32: multiply by the fp32 reciprocal
            scl_v = v;
        }
        const PlaceDesc *dv = BATCH ? a.gdesc + (long long)v * a.pk_stride : nullptr;
        const int np = BATCH ? __ldg(a.gnp + v) : a.place.np;
        int cnt_l, rel0_l;
        sp_overlap<BATCH>(a, dv, np, lane, a.lo + h0 * 512, nh * 512, cnt_l, rel0_l);
        const unsigned active = __ballot_sync(0xffffffffu, cnt_l > 0);
        float2 *dst = a.out + (long long)v * a.out_ld;
#pragma unroll 1
        for (int h = 0; h < nh; ++h) {
            float2 x[16];
            const bool live = active && sp_build_half<BATCH>(a, dv, active, cnt_l, rel0_l, h * 512, lane, x, keep);
            if (!live) {
#pragma unroll
                for (int j = 0; j < 16; ++j) x[j] = make_float2(0.f, 0.f);
            }
            wf_store_half<true, true>(dst, (h0 + h) * 512, n, live, scl, lane, x, vst, zero_u32, pol);
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <bool BATCH, int CHUNK>
inline cudaError_t span_attr() {
    cudaError_t e = cudaFuncSetAttribute(span_kernel<BATCH, CHUNK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSpSmem);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(span_kernel<BATCH, CHUNK>, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared);
}
inline int span_setup() {
    B2_CUDA((span_attr<false, kSpChunk>()));
    B2_CUDA((span_attr<true, 16>())); B2_CUDA((span_attr<true, 8>())); B2_CUDA((span_attr<true, 4>()));
    return 0;
}

// a.place (or gdesc/gnp), n_vec, lo/hi, out, maxsq, mode and the max-pass span list are filled by the caller.
inline int launch_span(b200vec_ctx *ctx, SpanArgs &a, cudaStream_t st) {
    const long long n = a.hi - a.lo;
    if (a.n_vec <= 0) return 0;
    const bool batch = a.gdesc != nullptr;
    const bool pass_a = a.mode == B200VEC_BUILD_MAXONLY || a.mode == kSpNorm;
    const bool pass_b = a.mode != B200VEC_BUILD_MAXONLY;
    if (!pass_a) { a.spans = nullptr; a.n_spans = 0; }
    if ((!pass_b || n <= 0) && a.n_spans == 0 && !pass_a) return 0;
    B2_REQUIRE(!pass_b || n <= 0 || (((uintptr_t)a.out & 15) == 0 && (a.n_vec == 1 || (a.out_ld & 1) == 0)),
               "output must be 16-byte aligned (and the vector pitch even)");
    if (!ctx->sp_ctr) B2_CUDA(cudaMalloc(&ctx->sp_ctr, 4 * sizeof(unsigned int)));
    B2_CUDA(cudaMemsetAsync(ctx->sp_ctr, 0, 4 * sizeof(unsigned int), st));
    a.ctr = ctx->sp_ctr;
    static int occ[2] = {0, 0};            // resident CTAs per SM (register / shared-memory limited)
    if (!occ[batch]) {
        if (batch) B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], span_kernel<true, 16>, kSpThreads, kSpSmem));
        else B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], span_kernel<false, kSpChunk>, kSpThreads, kSpSmem));
        B2_REQUIRE(occ[batch] > 0, "span_kernel does not fit an SM");
    }
    const long long halves = n > 0 ? (n + 511) / 512 : 0;
    // batch: 16-half items amortise the per-item descriptor fetch, as long as every warp still gets >= ~24 of them
    // (32 vectors per GPU at cfg5 / 8 GPUs would leave 6 per warp: a 13 % tail); halve the item until it does
    const long long cap_warps = (long long)ctx->sm_count * occ[batch] * kSpWarps;
    int chunk = batch ? kSpChunkBatch : kSpChunk;
    while (batch && chunk > kSpChunk && (halves + chunk - 1) / chunk * a.n_vec < 24 * cap_warps) chunk >>= 1;
    a.chunk = chunk;
    const long long items = pass_b ? (halves + chunk - 1) / chunk * a.n_vec : 0;
    const long long units = a.n_spans;
    const long long work = items > units ? items : units;
    const long long cap = (long long)ctx->sm_count * occ[batch];       // one full wave: persistent warps, dynamic items
    long long grid = (work + kSpWarps - 1) / kSpWarps;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    B2_REQUIRE(items < 0x7fffffffLL - cap * kSpWarps, "too many work items for one launch (split the batch)");
    void *kargs[] = {&a};
    const void *fn = !batch ? (const void *)span_kernel<false, kSpChunk>
                   : chunk == 16 ? (const void *)span_kernel<true, 16>
                   : chunk == 8 ? (const void *)span_kernel<true, 8> : (const void *)span_kernel<true, 4>;
    if (a.mode == kSpNorm)                 // grid barrier between the passes: cooperative launch
        B2_CUDA(cudaLaunchCooperativeKernel(fn, dim3((unsigned)grid), dim3(kSpThreads), kargs, kSpSmem, st));
    else
        B2_CUDA(cudaLaunchKernel(fn, dim3((unsigned)grid), dim3(kSpThreads), kargs, kSpSmem, st));
    B2_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
// Batch bank: every distinct (source packet, frequency shift) of a scenario sweep shifted in ONE launch
// (apply_frequency_shift, utils.py:120-127 — the arithmetic of shift_commit_kernel), plus max |y|^2 of
// each shifted packet for the pass-A hint.
struct ShiftJob {
    const float2 *src;           // committed or uploaded source samples
    float2 *dst;                 // len samples in the batch bank (8-byte loads only: one copy)
    unsigned long long *pmax;    // bits of max |y|^2 (double), zeroed by the caller
    long long len;
    double w;                    // (2*pi)*f, formed on the host in the reference's order
    int do_shift;
};

__global__ void __launch_bounds__(256)
batch_shift_kernel(const ShiftJob *__restrict__ jobs, double fs) {
    const ShiftJob jb = jobs[blockIdx.y];
    double m = 0.0;
    for (long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x; n < jb.len; n += (long long)gridDim.x * blockDim.x) {
        float2 y = jb.src[n];
        if (jb.do_shift) {
            const double t = (double)n / fs;
            const double ph = __dmul_rn(jb.w, t);
            double s, c;
            sincos_phase(ph, &s, &c);
            const double xr = (double)y.x, xi = (double)y.y;
            const double re = __dsub_rn(__dmul_rn(xr, c), __dmul_rn(xi, s));
            const double im = __dadd_rn(__dmul_rn(xr, s), __dmul_rn(xi, c));
            y = make_float2((float)re, (float)im);
        }
        jb.dst[n] = y;
        m = fmax(m, (double)y.x * (double)y.x + (double)y.y * (double)y.y);
    }
    m = warp_max(m);
    if ((threadIdx.x & 31) == 0 && m > 0.0) atomicMax(jb.pmax, (unsigned long long)__double_as_longlong(m));
}

// ---------------------------------------------------------------------------------------
// Periodic spectrogram tiling (dense periodic vectors, BASELINE.json configs[3]).  Where the vector repeats bit for
// bit with period T (hostmath.tile_plan), its full-rate spectrogram repeats with period Pf = lcm(T, 512) / 512
// frames: frame f is the same 1024 samples, hence the same 4 KB, as frame f - Pf.  The steady part of a slab's
// spectrogram is therefore COPIED from the Pf frames before it (frame-major memory: dst[f] = src[f0 + (f - f0) mod Pf])
// instead of transformed — the tile (38 MB at cfg4) is read through L2 with evict_last, the copies stream out.
// One warp per frame and iteration, 16-byte accesses.
__global__ void __launch_bounds__(256)
replicate_frames_kernel(float *__restrict__ spec, long long ld, long long f_src0, long long period, long long f_dst0,
                        long long n_frames) {
    const unsigned long long keep = wf_policy_evict_last();
    const int lane = threadIdx.x & 31;
    const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long i = gw; i < n_frames; i += nw) {
        const long long f = f_dst0 + i;
        const float4 *src = reinterpret_cast<const float4 *>(spec + (f_src0 + (f - f_src0) % period) * ld) + lane;
        float4 *dst = reinterpret_cast<float4 *>(spec + f * ld) + lane;
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k)
            asm volatile("ld.global.nc.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;"
                         : "=f"(v[k].x), "=f"(v[k].y), "=f"(v[k].z), "=f"(v[k].w) : "l"(src + 32 * k), "l"(keep));
#pragma unroll
        for (int k = 0; k < 8; ++k) __stcs(dst + 32 * k, v[k]);
    }
}

}  // namespace b200vec
