// pipe1024.cuh — the BASELINE step as a warp-specialised pipeline: packet placement (+ scale),
// vector store AND the nfft = 1024 spectrogram in one persistent kernel (same contract and
// same arithmetic as fused1024.cuh; this is the scheduling that keeps the SM busy).
//
// Why: measured on B200 (profiles/README.md, scripts/fused_ab.py) the lock-step kernel spends its
// time in phases that cannot overlap — every warp loads, then every warp does FFT arithmetic,
// then every warp stores — so the shared-memory pipe (~4.5 k wavefront-cycles per tile), the FMA
// pipe (~4.9 k issue cycles per tile) and the SM->L2 store path take turns instead of running
// together.  Here the 16 warps of the one CTA per SM are split by ROLE:
//
//   warps 0-7  "column" warps (P3): wait for the span of tile t in the stage (bulk copies, issued
//              one tile ahead by warp 0), load frames w and w + 8 (lane l holds x[l + 32 j]),
//              first 32-point transform with the window folded in, write B_l[q] to the exchange
//              buffer.  Multi-instance spans: the extra (packet, instance) pairs are added here, in
//              the reference's order, straight from the L2-resident bank, and the summed samples
//              are scaled and stored to the vector from registers.
//   warps 8-15 "row" warps (P5/P6): wait for the exchange buffer of tile t, thread = (frame f,
//              residue q): second 32-point transform with the W1024 twiddles folded in, |X|^2,
//              fftshifted direct store (one 64-byte run per half-warp and output row).
//
// The two groups meet only through four mbarriers (stage full / stage free / exchange full /
// exchange free), so the column warps are already transforming tile t + 1 while the row warps
// finish and store tile t: loads, arithmetic and stores of different tiles overlap.
//
// Spans covered by zero or one (packet, instance) — 86 % of the tiles of the BASELINE workload —
// never touch a register on their way to the vector: bank (pre-scaled copy) -> bulk copy -> stage
// -> bulk copy -> HBM.
#pragma once
#include "fused1024.cuh"

namespace b200vec {

constexpr int kPGroup = 8;                       // warps per role
constexpr int kPThreads = 64 * kPGroup;          // 512
using PL = FL<16>;                               // same shared-memory plan as the lock-step kernel

struct PMisc {                                   // lives at PL::kOffMisc
    long long rel[B200VEC_MAX_PACKETS];          // tile being assembled: packet index of stage sample 0, first overlapping instance
    int cnt[B200VEC_MAX_PACKETS];                // overlapping instances per packet
    int total, first;
    int cls[2];                                  // by tile parity: 0 = empty span, 1 = bulk-assembled, 2 = bulk-assembled first instance + extras
    unsigned long long stage_full, stage_free, exch_full, exch_free;
};
static_assert(sizeof(PMisc) <= 512, "PMisc must fit its smem slot");

__device__ __forceinline__ void mbar_arrive(unsigned mbar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(mbar) : "memory");
}
__device__ __forceinline__ void tma_store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// instance ranges of every packet over the span of tile t (one warp, one lane per packet)
__device__ __forceinline__ void p_setup_tile(const FArgs &a, long long t, int lane, PMisc *m) {
    const long long s0 = (a.frame_lo + 16 * t) * kFMaxHop;
    int c32 = 0;
    if (lane < a.place.np) {
        const PlaceDesc &d = a.place.d[lane];
        long long klo = floor_div(s0 - d.len - d.start, d.period) + 1;
        long long khi = floor_div(s0 + PL::kSpan - 1 - d.start, d.period);
        if (klo < 0) klo = 0;
        if (khi > d.count - 1) khi = d.count - 1;
        const long long c = khi - klo + 1;
        c32 = c > 0 ? (int)(c < 0x00ffffff ? c : 0x00ffffff) : 0;
        m->cnt[lane] = c32;
        m->rel[lane] = s0 - (d.start + klo * d.period);
    }
    const unsigned anyb = __ballot_sync(0xffffffffu, c32 > 0);
    const int total = __reduce_add_sync(0xffffffffu, c32);
    if (lane == 0) { m->total = total; m->first = anyb ? __ffs(anyb) - 1 : 0; }
    __syncwarp();
}

// bulk-copy assembly of a span from the FIRST overlapping (packet, instance): zeros | samples | zeros.
// scaled = the pre-scaled bank copy (span holds final vector samples) or the raw one (extras follow).
__device__ __forceinline__ void p_issue_span(const FArgs &a, const PMisc *m, bool scaled, unsigned stage_u32, unsigned mbar) {
    constexpr int span = PL::kSpan;
    const int p = m->first;
    const PlaceDesc &d = a.place.d[p];
    const int rel = (int)m->rel[p];
    const int lo = rel < 0 ? ((-rel) & ~1) : 0;
    const int end = d.len - rel;                                   // > 0
    const int hi = end >= span ? span : ((end + 1) & ~1);
    const float2 *base = ((rel & 1) ? d.b : d.a) + (scaled ? a.soff[p] : 0) + rel;   // base + even s is 16-byte aligned
    mbar_expect_tx(mbar, (unsigned)(span * sizeof(float2)));
    if (lo > 0) tma_bulk_g2s_chunked(stage_u32, reinterpret_cast<const char *>(a.zeros), (unsigned)lo * 8u, mbar);
    tma_bulk_g2s_chunked(stage_u32 + (unsigned)lo * 8u, reinterpret_cast<const char *>(base + lo), (unsigned)(hi - lo) * 8u, mbar);
    if (hi < span) tma_bulk_g2s_chunked(stage_u32 + (unsigned)hi * 8u, reinterpret_cast<const char *>(a.zeros), (unsigned)(span - hi) * 8u, mbar);
}


template <bool FUSED>
__global__ void __launch_bounds__(kPThreads, 1)
pipe1024_kernel(const __grid_constant__ FArgs a) {
    extern __shared__ __align__(128) unsigned char p_smem[];
    float2 *stage = reinterpret_cast<float2 *>(p_smem + PL::kOffStage);
    float2 *exch = reinterpret_cast<float2 *>(p_smem + PL::kOffExch);
    float4 *tw2 = reinterpret_cast<float4 *>(p_smem + PL::kOffTw);       // [q][16] twiddle pairs
    PMisc *misc = reinterpret_cast<PMisc *>(p_smem + PL::kOffMisc);
    const unsigned stage_u32 = smem_u32(stage);
    const unsigned b_stage_full = smem_u32(&misc->stage_full), b_stage_free = smem_u32(&misc->stage_free);
    const unsigned b_exch_full = smem_u32(&misc->exch_full), b_exch_free = smem_u32(&misc->exch_free);

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, w = warp & (kPGroup - 1);
    for (int i = tid; i < 512; i += kPThreads) tw2[i] = __ldg(reinterpret_cast<const float4 *>(a.tw) + i);
    if (tid == 0) {
        mbar_init(b_stage_full, 1);
        mbar_init(b_stage_free, kPGroup);
        mbar_init(b_exch_full, kPGroup);
        mbar_init(b_exch_free, kPGroup);
    }
    __syncthreads();
    const int hop = FUSED ? kFMaxHop : a.hop;
    const long long t_first = blockIdx.x;
    if (t_first >= a.n_tiles) return;

    if (warp < kPGroup) {
        // ================================================================== column warps
        float wreg[32];                               // this lane's window taps * sqrt(scale), resident
        {
            const float rs = sqrtf(a.scale);
#pragma unroll
            for (int j = 0; j < 32; ++j) wreg[j] = __ldg(a.window + lane + 32 * j) * rs;
        }
        float scl = 1.0f;
        if (FUSED && a.mode == B200VEC_BUILD_SCALE) {
            const double m2 = __longlong_as_double((long long)*a.maxsq);
            if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);   // numpy's complex64 / float32: multiply by the fp32 reciprocal
        }
        if (warp == 0) {                              // span of the first tile
            if (FUSED) {
                p_setup_tile(a, t_first, lane, misc);
                if (lane == 0) {
                    const int c = misc->total == 0 ? 0 : (misc->total == 1 ? 1 : 2);
                    misc->cls[0] = c;
                    if (c == 0) mbar_arrive(b_stage_full);
                    else p_issue_span(a, misc, c == 1, stage_u32, b_stage_full);
                }
            } else if (lane == 0) {
                f_issue_tma<16>(a, t_first, stage_u32, b_stage_full);
            }
        }
        unsigned it = 0;
        PT_DECL
        for (long long t = t_first; t < a.n_tiles; t += gridDim.x, ++it) {
            const unsigned par = it & 1;
            const long long f0 = t * 16;
            const int nfr = (int)((a.n_frames - f0 < 16) ? (a.n_frames - f0) : 16);     // may be <= 0 (FUSED tail)
            PT(7)
            mbar_wait(b_stage_full, par);
            PT(0)
            const int cls = FUSED ? misc->cls[par] : 1;
            int rem = 0;
            float2 *dst = nullptr;
            if (FUSED) {
                const long long s0 = (a.frame_lo + f0) * kFMaxHop;
                const long long rem64 = a.place.hi - s0;             // samples of this tile inside the slab
                rem = rem64 > PL::kOwn ? PL::kOwn : (rem64 < 0 ? 0 : (int)rem64);
                dst = a.vec + (s0 - a.place.lo);
                if (cls == 1 && tid == 0 && rem > 0) {               // bulk-assembled span: smem -> HBM bulk copy
                    if (rem > 1) tma_bulk_s2g(dst, stage_u32, (unsigned)(rem & ~1) * 8u);
                    tma_store_commit();
                    if (rem & 1) __stcs(dst + rem - 1, stage[rem - 1]);
                }
                if (cls == 0) {                                      // empty span: the vector is zero here
                    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
                    for (int k = 0; k < PL::kOwn / (2 * 32 * kPGroup); ++k) {
                        const int s = 2 * (tid + 32 * kPGroup * k);
                        if (s + 1 < rem) __stcs(reinterpret_cast<float4 *>(dst + s), z);
                        else if (s < rem) __stcs(dst + s, make_float2(0.f, 0.f));
                    }
                }
            }
            if (FUSED && cls == 2) {
                // More than one (packet, instance) overlaps the span.  The stage holds the first one
                // (raw); the column threads add the others IN PLACE in the reference's order, scale,
                // store the tile's own samples to the vector and leave the final samples in the stage.
                constexpr int kCT = 32 * kPGroup;                        // 256 column threads
                constexpr int kChunk = 6;                                // float4 per thread and sweep
                float4 *st4 = reinterpret_cast<float4 *>(stage);
                const int first = misc->first;
#pragma unroll 1
                for (int c0 = 0; c0 < PL::kSpan / 2; c0 += kCT * kChunk) {
                    float4 acc[kChunk];
#pragma unroll
                    for (int j = 0; j < kChunk; ++j) {
                        const int i4 = c0 + tid + kCT * j;
                        acc[j] = i4 < PL::kSpan / 2 ? st4[i4] : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    for (int p = first; p < a.place.np; ++p) {
                        const int cnt = misc->cnt[p];
                        if (cnt <= 0) continue;
                        const PlaceDesc &d = a.place.d[p];
                        const int len = d.len;
                        long long r64 = misc->rel[p];
                        for (int i = 0; i < cnt; ++i, r64 -= d.period) {
                            if (p == first && i == 0) continue;
                            const int rel = (int)r64;
                            const float2 *base = ((rel & 1) ? d.b : d.a) + rel;
#pragma unroll
                            for (int j = 0; j < kChunk; ++j) {
                                const int s = 2 * (c0 + tid + kCT * j);
                                const int n = s + rel;
                                if (s < PL::kSpan && n >= -1 && n < len) {   // the bank's zero pads cover n = -1 and n + 1 = len
                                    const float4 y = __ldg(reinterpret_cast<const float4 *>(base + s));
                                    acc[j].x += y.x; acc[j].y += y.y; acc[j].z += y.z; acc[j].w += y.w;
                                }
                            }
                        }
                    }
#pragma unroll
                    for (int j = 0; j < kChunk; ++j) {
                        const int i4 = c0 + tid + kCT * j;
                        const int s = 2 * i4;
                        if (i4 < PL::kSpan / 2) {
                            float4 v4 = acc[j];
                            v4.x *= scl; v4.y *= scl; v4.z *= scl; v4.w *= scl;
                            st4[i4] = v4;
                            if (s + 1 < rem) __stcs(reinterpret_cast<float4 *>(dst + s), v4);
                            else if (s < rem) __stcs(dst + s, make_float2(v4.x, v4.y));
                        }
                    }
                }
                asm volatile("bar.sync 1, %0;" ::"n"(32 * kPGroup) : "memory");   // column warps only
            }
            PT(1)
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
                const int f = w + kPGroup * pass;
                const bool do_fft = cls != 0 && f < nfr;
                cf32 v[32];
                if (do_fft) {
                    const float2 *sp = stage + f * hop + lane;
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float2 x = sp[32 * brev5(i)];
                        v[i].x = x.x; v[i].y = x.y;
                    }
                }
                PT(2)
                if (pass == 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(b_stage_free);          // this warp has read everything it needs from the stage
                    if (warp == 0) {                                   // assemble the next span while this one is transformed
                        mbar_wait(b_stage_free, par);
                        const long long tn = t + gridDim.x;
                        if (tn < a.n_tiles) {
                            if (FUSED) {
                                if (lane == 0 && cls == 1) tma_store_wait_read();   // the vector bulk store has left the stage
                                __syncwarp();
                                p_setup_tile(a, tn, lane, misc);
                                if (lane == 0) {
                                    const int c = misc->total == 0 ? 0 : (misc->total == 1 ? 1 : 2);
                                    misc->cls[par ^ 1] = c;
                                    if (c == 0) mbar_arrive(b_stage_full);
                                    else p_issue_span(a, misc, c == 1, stage_u32, b_stage_full);
                                }
                            } else if (lane == 0) {
                                f_issue_tma<16>(a, tn, stage_u32, b_stage_full);
                            }
                        }
                        __syncwarp();
                    }
                }
                PT(5)
                if (do_fft) fft32_windowed(v, wreg);
                PT(2)
                if (pass == 0 && it > 0) mbar_wait(b_exch_free, (it - 1) & 1);   // the row warps hold tile t - 1 in registers
                PT(3)
                if (do_fft) {
                    float2 *ex = exch + f * kExPitch + lane;
#pragma unroll
                    for (int q = 0; q < 32; ++q) ex[32 * q] = make_float2(v[q].x, v[q].y);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(b_exch_full);
            PT(4)
        }
        PT_FLUSH
        if (tid == 0) tma_store_wait_all();
    } else {
        // ================================================================== row warps
        unsigned anybits = 0;
        const int f5 = lane & 15;
        unsigned it = 0;
        PT_DECL
        for (long long t = t_first; t < a.n_tiles; t += gridDim.x, ++it) {
            const unsigned par = it & 1;
            const long long f0 = t * 16;
            const int nfr = (int)((a.n_frames - f0 < 16) ? (a.n_frames - f0) : 16);
            PT(7)
            mbar_wait(b_exch_full, par);
            PT(0)
            const int cls = FUSED ? misc->cls[par] : 1;
            const bool active = f5 < nfr;
            const unsigned step = 32u * (unsigned)a.out_ld;         // host guarantees 7 * 32 * out_ld < 2^32
#pragma unroll 1
            for (int item = 0; item < 2; ++item) {
                const int q5 = 2 * w + (lane >> 4) + 16 * item;
                cf32 v[32];
                if (cls != 0 && active) {
                    const float4 *ex4 = reinterpret_cast<const float4 *>(exch + f5 * kExPitch + q5 * 32);
#pragma unroll
                    for (int l = 0; l < 32; l += 2) {
                        const float4 x = ex4[l / 2];
                        v[brev5(l)].x = x.x; v[brev5(l)].y = x.y;
                        v[brev5(l + 1)].x = x.z; v[brev5(l + 1)].y = x.w;
                    }
                }
                if (item == 1) {
                    __syncwarp();
                    if (lane == 0) mbar_arrive(b_exch_free);          // tile t now lives in registers
                }
                PT(1)
                if (!active) continue;
                // row q + 32 sp, sp = 8 hi + lo: four 64-bit row-group bases, 32-bit offsets inside a group
                float *o = a.out + (long long)q5 * a.out_ld + (f0 + f5);
                float *og[4];
#pragma unroll
                for (int h = 0; h < 4; ++h) og[h] = o + (long long)h * 256 * a.out_ld;
                if (cls != 0) {
                    const float4 *tq = tw2 + q5 * 16;
                    fft32_twiddled_pairs(v, [tq](int p) { const float4 t4 = tq[p]; return cf32x2{t4.x, t4.y, t4.z, t4.w}; });
#pragma unroll
                    for (int sp = 0; sp < 32; ++sp) {             // output row q + 32 sp holds bin q + 32 s, s = (sp + 16) & 31
                        const int s = (sp + 16) & 31;
                        const float p = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
                        anybits |= __float_as_uint(p);            // p >= 0 or NaN: non-zero bits <=> p != 0
                        __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, p);
                    }
                } else {
#pragma unroll
                    for (int sp = 0; sp < 32; ++sp) __stcs(og[sp >> 3] + (unsigned)(sp & 7) * step, 0.0f);
                }
                PT(2)
            }
        }
        PT_FLUSH
        if (a.flag && __any_sync(0xffffffffu, anybits != 0) && lane == 0) atomicOr(a.flag, 1);
    }
}

// pair-ordered twiddle table: tw2[q][P] = (W1024^(q la), W1024^(q lb)), la = brev5(2 P), lb = brev5(2 P + 1)
inline int get_tw1024_pairs(b200vec_ctx *ctx, const float2 **out) {
    const int key = -1025;
    auto it = ctx->twiddles.find(key);
    if (it != ctx->twiddles.end()) { *out = it->second; return 0; }
    std::vector<float2> h(1024);
    for (int q = 0; q < 32; ++q)
        for (int i = 0; i < 32; ++i) {
            const int l = brev5(i);
            const double ang = -2.0 * M_PI * (double)((q * l) % 1024) / 1024.0;
            h[q * 32 + i] = make_float2((float)cos(ang), (float)sin(ang));
        }
    float2 *d = nullptr;
    B2_CUDA(cudaMalloc(&d, sizeof(float2) * 1024));
    B2_CUDA(cudaMemcpy(d, h.data(), sizeof(float2) * 1024, cudaMemcpyHostToDevice));
    ctx->twiddles[key] = d;
    *out = d;
    return 0;
}

inline bool pipe1024_enabled() {
    // Measured on B200 (scripts/pipe_timing.py): 358 us vs 296 us for the lock-step kernel — with
    // one exchange buffer the two roles alternate on it and the 8 column warps carry the critical
    // path.  Kept as an A/B variant: B200VEC_PIPE=1 selects it.
    const char *e = ab_env("B200VEC_PIPE");
    return e && e[0] == '1';
}

template <bool FUSED>
inline int launch_pipe1024(b200vec_ctx *ctx, FArgs &a, long long n_frames, long long n_vec, cudaStream_t st) {
    const long long t_frames = (n_frames + 15) / 16, t_vec = (n_vec + PL::kOwn - 1) / PL::kOwn;
    a.n_tiles = t_frames > t_vec ? t_frames : t_vec;
    if (a.n_tiles == 0) return 0;
    if (get_tw1024_pairs(ctx, &a.tw)) return 1;
    const long long grid = a.n_tiles < ctx->sm_count ? a.n_tiles : ctx->sm_count;
    pipe1024_kernel<FUSED><<<(unsigned)grid, kPThreads, PL::kSmem, st>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

inline int launch_pipe1024_plain(b200vec_ctx *ctx, FArgs &a, long long n_frames, cudaStream_t st) {
    return launch_pipe1024<false>(ctx, a, n_frames, 0, st);
}

inline int pipe1024_setup() {
    B2_CUDA(cudaFuncSetAttribute(pipe1024_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PL::kSmem));
    B2_CUDA(cudaFuncSetAttribute(pipe1024_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PL::kSmem));
    return 0;
}

}  // namespace b200vec
