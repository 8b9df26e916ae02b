// wf1024.cuh — the BASELINE step (placement + scale + store + nfft = 1024 spectrogram) with every
// WARP working on its own frames and the spectrogram stored FRAME-MAJOR.
//
// Replaces, for the full-rate analysis shape (hann 1024 / hop 512 / nfft 1024, two-sided,
// vector_analyzer/spectrogram_analysis.py:35-49):
//   unified_gui.py:1712,1757-1769,1781 (np.zeros, `+=` loop, divide by max)
//   scipy.signal.spectrogram + fftshift(axes=0)
//
// Why a second fused kernel.  The tile kernel (fused1024.cuh) writes the reference's C-order
// [1024, frames] array directly; 64-byte row runs force 16 frames per tile, a 131 KB exchange
// buffer, one CTA per SM and two block barriers per tile, so its shared-memory phases, its FP
// phases and its store phases run one after the other (measured: 163 us skeleton + 74 us FFT
// + 60 us stores = the 297 us total, profiles/README.md).  Here the spectrogram buffer is
// [frames, 1024] (the array handed to Python is its transpose VIEW: same shape, dtype and values as
// the reference's, Fortran-ordered).  One frame's 1024 bins are then 4 KB of contiguous memory, a
// whole 1024-point transform fits inside ONE warp, and nothing is shared between warps:
//   * no block barrier in the main loop (a few __syncwarp per frame), no mbarrier;
//   * 16 independent warps per SM drift out of phase, so one warp's L2 gather, another's
//     shared-memory exchange, another's FMAs and another's stores overlap;
//   * a frame's 4 KB of PSD and a half's 4 KB of samples each leave as ONE bulk copy
//     (cp.async.bulk shared -> global, L2 evict_first) out of the warp's own staging blocks, so the
//     stores occupy neither registers nor the LSU queue the exchange needs (direct 128-byte-line
//     stores remain as the B200VEC_WF_BULK=0 A/B path); bank loads carry an L2 evict_last hint.
// Normalisation: the caller supplies the max (SCALE), or the kernel finds it first (kWfAutoNorm:
// every warp reduces its share of the max ranges, grid barrier under a cooperative launch), or —
// multi-GPU — additionally MAX-combines it with the other GPUs through NVLink peer memory
// (kWfAutoNormPeer): the path's only collective, inside the one launch.
// PLAIN instantiation: stand-alone STFT of an existing vector in the same layout.
//
// Work item = kWfChunk consecutive frames, handed out by a global atomic counter (dynamic: live
// and empty stretches of the vector cost very different amounts), single frames for the last
// two items' worth per warp so that the warps run dry together.  Per item a warp
//   S   finds, per packet (lane p), the instance range overlapping the item's span
//       (two 64-bit floor divisions), once;
//   B   builds the next 512 samples: lane l owns x[l + 32 j]; for every overlapping
//       (packet, instance) in the reference's order it adds the packet samples (8-byte loads,
//       256 B per warp instruction, from the L2-resident bank), scales, and streams them to the
//       vector (each sample is stored by exactly one item);
//   T1  first 32-point transform over j with the window folded in (fft32.cuh) on
//       [previous 512 | new 512], results to the warp's private exchange block [l][q] (pitch 33);
//   T2  lane q reads B_l[q] for all l, second transform with the W1024 twiddles folded in,
//       |X|^2 -> out[f][ (k + 512) mod 1024 ], k = q + 32 s.
// HBM traffic: 8 B/sample vector store + 8 B/sample spectrogram store, nothing read.
#pragma once
#include "common.cuh"
#include "place.cuh"
#include "fft32.cuh"
#include "stft1024.cuh"      // get_tw1024
#include "fused1024.cuh"     // tma_bulk_s2g, tma_store_commit
#include "maxspans.cuh"

namespace b200vec {

constexpr int kWfWarps = 16;
constexpr int kWfThreads = 32 * kWfWarps;
constexpr int kWfChunk = 4;                                       // frames per work item (default; B200VEC_WF_CHUNK=4|8 for A/B: 190 vs 203 us)
constexpr int kWfMaxNormRanges = 4;
constexpr int kWfAutoNorm = 3;                                    // WfArgs::mode: find the max in-kernel, then scale
constexpr int kWfAutoNormPeer = 4;                                // same, and MAX-combine it with the other GPUs through peer memory
constexpr int kWfMaxPeers = 8;
constexpr long long kWfPeerTimeoutClocks = 4000000000LL;          // ~2 s at 1.9 GHz: bound of the in-kernel wait for the other GPUs
constexpr int kWfExPitch = 33;                                    // float2 per l-row of a warp's exchange block
constexpr size_t kWfOffExch = 0;
constexpr size_t kWfOffTw = kWfOffExch + (size_t)kWfWarps * 32 * kWfExPitch * sizeof(float2);
constexpr size_t kWfOffWin = kWfOffTw + 32 * 32 * sizeof(float2);
constexpr size_t kWfSmemDirect = kWfOffWin + 1024 * sizeof(float);
// bulk-store variant: one 4 KB staging block per warp for the vector half + one shared block of zeros
constexpr size_t kWfOffVst = kWfSmemDirect;
constexpr size_t kWfOffZero = kWfOffVst + (size_t)kWfWarps * 4096;
constexpr size_t kWfSmemBulk = kWfOffZero + 4096;

struct WfArgs {
    PlaceArgs place;             // descriptors + the vector slab [lo, hi) this launch stores; lo = frame_lo * 512
    float2 *vec;                 // FUSED: vector out, vec[0] = sample place.lo.  PLAIN: vector in, vec[0] = first sample of frame 0 of this launch
    const unsigned long long *maxsq;
    int mode;                    // B200VEC_BUILD_STORE or B200VEC_BUILD_SCALE
    const float *window;         // 1024 taps
    const float2 *tw;            // [l][q] = exp(-2*pi*i*l*q/1024)  (symmetric)
    float scale;
    long long n_frames;          // frames [frame_lo, frame_lo + n_frames)
    float *out;                  // out[(f - frame_lo) * out_ld + row]
    long long out_ld;            // >= 1024, multiple of 4
    int *flag;
    long long n_big;             // work items [0, n_big) are CHUNK frames long, the rest one frame
    long long n_chunks;
    // in-kernel normalisation (mode == kWfAutoNorm[Peer]): max |v|^2 over these spans first — the pieces of the
    // caller's norm ranges that can hold the max (maxspans.cuh; device memory owned by the context)
    const MaxSpan *spans;
    long long n_spans;
    double *maxsq_out;           // receives max |v|^2 (may be NULL)
    int plain;                   // host side: launch the PLAIN instantiation
    // kWfAutoNormPeer: one symmetric 64-byte block per GPU, peer[r] = rank r's block mapped here:
    //   u64 [0], [1]  max |v|^2 bits of even / odd epochs      u32 at byte 16: arrivals (monotonic)
    unsigned long long *peer[kWfMaxPeers];
    int world, rank;
    unsigned int epoch;          // 1, 2, 3, ... identical on every rank for the same step; 0: count steps on the device
                                 // (u32 at byte 24 of the home block), which lets a step be replayed from a CUDA graph
    unsigned int *ctr;           // per-launch words, zeroed on the stream before every launch (and re-armed by the last CTA):
                                 //   [0] next work item, [1] finished CTAs, [2] grid barrier, [4..5] max |v|^2 bits
                                 // persistent words (peer mode): [6] epoch published, [8..9] combined max,
                                 //   [10] sticky error flag: 1 = the peer exchange timed out (b200vec_peer_status)
};

// Packet-bank loads ask L2 to keep their lines (evict_last) and the output leaves with evict_first, so the
// ~5 MB bank survives the 896 MB the kernel streams through the 126 MB L2 — between steps too: pass A of
// the next launch starts on a warm bank.
__device__ __forceinline__ unsigned long long wf_policy_evict_last() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ float2 wf_ldg_keep(const float2 *p, unsigned long long pol) {
    float2 v;
    asm volatile("ld.global.nc.L2::cache_hint.v2.f32 {%0, %1}, [%2], %3;" : "=f"(v.x), "=f"(v.y) : "l"(p), "l"(pol));
    return v;
}
__device__ __forceinline__ unsigned long long wf_policy_evict_first() {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void wf_bulk_s2g(void *dst, unsigned src, unsigned bytes, unsigned long long pol) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 ::"l"(dst), "r"(src), "r"(bytes), "l"(pol) : "memory");
}

// B: the 512 samples starting `off` samples into the item's span.  Returns false when no packet
// instance overlaps them (x is then exactly zero).
__device__ __forceinline__ bool wf_build_half(const WfArgs &a, unsigned active, int cnt_l, int rel0_l, int off,
                                              int lane, float2 (&x)[16], unsigned long long keep) {
#pragma unroll
    for (int j = 0; j < 16; ++j) x[j] = make_float2(0.f, 0.f);
    bool live = false;
    for (unsigned m = active; m; m &= m - 1) {
        const int p = __ffs(m) - 1;
        const int cnt = __shfl_sync(0xffffffffu, cnt_l, p);
        const int rel0 = __shfl_sync(0xffffffffu, rel0_l, p);
        const PlaceDesc &d = a.place.d[p];
        const int len = d.len;
        const long long period = d.period;
        long long r64 = (long long)rel0 + off;               // packet index of x's sample 0 for instance klo + i
        for (int i = 0; i < cnt; ++i, r64 -= period) {
            if (r64 + 512 <= 0) break;                       // this and every later instance start after the half
            if (r64 >= len) continue;                        // instance ended before the half
            const int r = (int)r64;
            live = true;
            const float2 *base = d.a + r + lane;
            if (r >= 0 && r + 512 <= len) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 v = wf_ldg_keep(base + 32 * j, keep);
                    x[j] = __fadd2_rn(x[j], v);              // FADD2: one issue slot per complex add
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

// mbarrier wait for a whole warp with a warp-uniform loop condition (lane 0 polls, the result is shuffled):
// a per-lane spin loop makes ptxas treat everything after it as possibly divergent and emit out-of-line
// slow paths — among them a copy of the second transform — behind the __syncwarps of the main loop.
__device__ __forceinline__ void wf_mbar_wait_warp(unsigned mbar, unsigned parity, int lane) {
    unsigned ok;
    do {
        ok = 0;
        if (lane == 0)
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
                         "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(mbar), "r"(parity) : "memory");
        ok = __shfl_sync(0xffffffffu, ok, 0);
    } while (!ok);
}

__device__ __forceinline__ void wf_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void wf_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// scale + store half h (slab-relative samples [s, s + 512)); n_vec = samples this launch stores.
// BULK: whole halves leave through the warp's 4 KB staging block as ONE bulk copy (cp.async.bulk
// shared -> global) issued by lane 0 — the store then occupies neither registers nor the LSU queue
// that the shared-memory exchange of the other warps needs; zero halves are copied from the block
// of zeros.  Before a staging block is rewritten lane 0 waits until its bulk copies have finished
// READING shared memory (wait_group.read 0): the copy issued last is at least one L2 gather or one
// 32-point transform old by then, so the wait is normally free.
template <bool STORE, bool BULK>
__device__ __forceinline__ void wf_store_half(float2 *vec, long long s, long long n_vec, bool live, float scl,
                                              int lane, float2 (&x)[16], float2 *vst, unsigned zero_u32,
                                              unsigned long long pol) {
    if (live) {
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = __fmul2_rn(x[j], make_float2(scl, scl));
    }
    if (!STORE || s >= n_vec) return;
    float2 *dst = vec + s;
    if (s + 512 <= n_vec) {
        if (BULK) {
            if (live) {
                if (lane == 0) wf_wait_read();
                __syncwarp();
#pragma unroll
                for (int j = 0; j < 16; ++j) vst[lane + 32 * j] = x[j];
                wf_fence_async();
                __syncwarp();
                if (lane == 0) { wf_bulk_s2g(dst, smem_u32(vst), 4096u, pol); tma_store_commit(); }
            } else if (lane == 0) {
                wf_bulk_s2g(dst, zero_u32, 4096u, pol); tma_store_commit();
            }
        } else if (live) {
#pragma unroll
            for (int j = 0; j < 16; ++j) __stcs(dst + lane + 32 * j, x[j]);
        } else {                                             // zeros: 16-byte stores, 512 B per warp instruction
            float4 *d4 = reinterpret_cast<float4 *>(dst) + lane;
#pragma unroll
            for (int k = 0; k < 8; ++k) __stcs(d4 + 32 * k, make_float4(0.f, 0.f, 0.f, 0.f));
        }
    } else {
        const int rem = (int)(n_vec - s);
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (lane + 32 * j < rem) __stcs(dst + lane + 32 * j, x[j]);
    }
}

// CHUNK = frames per work item.  DBG (timing experiments only, results are wrong when non-zero):
// bit 0 drops the spectrogram stores, bit 1 the vector stores, bit 2 both transforms.
// PLAIN: stand-alone STFT of an existing vector (b200vec_stft1024_fm): halves are read from a.vec
// (8-byte loads, 256 B per warp instruction, each sample read once per item + one shared half),
// nothing is placed, scaled or stored back.
template <int CHUNK, int DBG, bool BULK, bool PLAIN>
__global__ void __launch_bounds__(kWfThreads, 1)
wf1024_kernel(const __grid_constant__ WfArgs a) {
    extern __shared__ __align__(16) unsigned char w_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float2 *exch = reinterpret_cast<float2 *>(w_smem + kWfOffExch) + warp * (32 * kWfExPitch);
    float2 *tw = reinterpret_cast<float2 *>(w_smem + kWfOffTw);
    float4 *win4 = reinterpret_cast<float4 *>(w_smem + kWfOffWin);
    float2 *vst = reinterpret_cast<float2 *>(w_smem + kWfOffVst) + warp * 512;      // BULK only
    const unsigned zero_u32 = smem_u32(w_smem + kWfOffZero);                        // BULK only
    if (BULK) {
        for (int i = tid; i < 256; i += kWfThreads)
            reinterpret_cast<float4 *>(w_smem + kWfOffZero)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        wf_fence_async();
    }

    // twiddles in the order the first stage of T2 consumes them: tw[P][q] = W^(q la), la = brev5(2 P) — W^(q lb),
    // lb = la + 16, is formed from it in registers: one 8-byte load per butterfly
    auto build_tables = [&]() {
        for (int i = tid; i < 1024; i += kWfThreads) {
            const int q = i & 31, k = i >> 5;                // k = 2 P + c
            // half table (fft32.cuh): slot P holds W^(q la(P)), la = brev5(2 P); slot 0 (la = 0) holds c16 = W^(16 q) instead
            if ((k & 1) == 0) tw[(k >> 1) * 32 + q] = __ldg(a.tw + (k == 0 ? 16 : brev5(k)) * 32 + q);
        }
        const float rs = sqrtf(a.scale);                    // |X|^2 * scale = |FFT(x * w * sqrt(scale))|^2
        float *wf = reinterpret_cast<float *>(win4);
        for (int i = tid; i < 1024; i += kWfThreads) {      // i = (k * 32 + lane) * 4 + c  ->  tap j = brev5(4 k + c) of lane
            const int c = i & 3, ln = (i >> 2) & 31, k = i >> 7;
            wf[i] = __ldg(a.window + ln + 32 * brev5(4 * k + c)) * rs;
        }
    };
    // with an in-kernel max pass the tables are built while the grid barrier fills (nothing before it needs them)
    const bool norm_pass = !PLAIN && (a.mode == kWfAutoNorm || a.mode == kWfAutoNormPeer);
    if (!norm_pass) build_tables();
    const unsigned long long keep = wf_policy_evict_last();
    float scl = 1.0f;
    if (a.mode == B200VEC_BUILD_SCALE) {
        const double m2 = __longlong_as_double((long long)*a.maxsq);
        if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);        // numpy's complex64 / float32: multiply by the fp32 reciprocal
    }
    // The first work item is claimed ahead of the max pass: the warps leave the grid barrier together and would
    // otherwise queue up at the one counter before any of them places a sample.
    unsigned next = 0;
    if (lane == 0) next = atomicAdd(a.ctr, 1u);
    __syncthreads();
    if (norm_pass) {
        // Pass A inside the kernel: every warp evaluates its share of the spans that can hold the max of the
        // whole vector (maxspans.cuh: the pieces of hostmath.max_eval_ranges under >= 2 instances; what lies
        // under one instance is covered by the packet maxima below), the grid meets at a barrier (cooperative
        // launch: all CTAs are resident), and everybody reads the same exact max.
        unsigned long long *slot = reinterpret_cast<unsigned long long *>(a.ctr + 4);
        double tmax = 0.0;                                   // running max of |v|^2 (exact products in fp64)
        float fthr = 0.0f;                                   // fp32 pre-filter threshold
        const long long gw = (long long)blockIdx.x * kWfWarps + warp, nw = (long long)gridDim.x * kWfWarps;
        for (long long u = gw; u < a.n_spans; u += nw) {
            const MaxSpan sp = a.spans[u];
            const long long s0 = sp.s0;
            int cnt_l = 0, rel0_l = 0;
            if (lane < a.place.np) {
                const PlaceDesc &d = a.place.d[lane];
                long long klo = floor_div(s0 - d.len - d.start, d.period) + 1;
                long long khi = floor_div(s0 + 512 - 1 - d.start, d.period);
                if (klo < 0) klo = 0;
                if (khi > d.count - 1) khi = d.count - 1;
                const long long cn = khi - klo + 1;
                cnt_l = cn > 0 ? (int)(cn < 0x00ffffff ? cn : 0x00ffffff) : 0;
                rel0_l = cnt_l ? (int)(s0 - (d.start + klo * d.period)) : 0;
            }
            const unsigned active = __ballot_sync(0xffffffffu, cnt_l > 0);
            float2 x[16];
            if (!wf_build_half(a, active, cnt_l, rel0_l, 0, lane, x, keep)) continue;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float q = x[j].x * x[j].x + x[j].y * x[j].y;
                if (lane + 32 * j < sp.n && q >= fthr) {           // samples past the span's end never enter the max
                    const double e = (double)x[j].x * (double)x[j].x + (double)x[j].y * (double)x[j].y;
                    if (e > tmax) { tmax = e; fthr = q * 0.99999f; }
                }
            }
        }
        tmax = warp_max(tmax);
        if (lane == 0 && tmax > 0.0) atomicMax(slot, (unsigned long long)__double_as_longlong(tmax));
        if (blockIdx.x == 0 && warp == 0 && lane < a.place.np && a.place.d[lane].free_inst) {
            const unsigned long long pm = *a.place.d[lane].pmax;          // max |y|^2 over the whole packet
            if (pm) atomicMax(slot, pm);
        }
        __syncthreads();
        if (tid == 0) {                                      // grid barrier: arrive ...
            __threadfence();
            atomicAdd(a.ctr + 2, 1u);
        }
        build_tables();                                      // ... build the window / twiddle tables while the others arrive ...
        if (tid == 0) {                                      // ... wait
            while (*reinterpret_cast<volatile unsigned int *>(a.ctr + 2) < gridDim.x) __nanosleep(40);
            __threadfence();
        }
        __syncthreads();
        if (a.mode == kWfAutoNormPeer) {
            // The one collective of the path, inside the kernel: CTA 0 raises every GPU's max slot with this
            // GPU's max (system-scope atomics over NVLink peer memory), signals its arrival to every GPU and
            // waits for all `world` arrivals at home; the other CTAs wait for the combined value it publishes.
            // Slots alternate with the epoch's parity; a slot is cleared one epoch before it is used again,
            // ahead of the arrival signal that lets any peer start that epoch.
            unsigned long long *fin = reinterpret_cast<unsigned long long *>(a.ctr + 8);
            unsigned long long *home = a.peer[a.rank];
            const unsigned int epoch = a.epoch ? a.epoch : *reinterpret_cast<volatile unsigned int *>(home + 3) + 1u;
            if (blockIdx.x == 0 && warp == 0) {
                // lane r talks to GPU r: the `world` remote updates travel over NVLink side by side
                const unsigned long long mine = *reinterpret_cast<volatile unsigned long long *>(slot);
                const int cur = epoch & 1;
                if (lane == 0) home[cur ^ 1] = 0ull;
                __threadfence_system();
                if (lane < a.world) {
                    if (mine) atomicMax_system(a.peer[lane] + cur, mine);
                    __threadfence_system();                  // the max lands before the arrival that announces it
                    atomicAdd_system(reinterpret_cast<unsigned int *>(a.peer[lane] + 2), 1u);
                }
                __syncwarp();
                if (lane == 0) {
                    // home is local memory: poll it back to back.  The wait is bounded (a rank that never launched
                    // must not hang the other GPUs): on a timeout the step goes on with the local max and the
                    // sticky error flag tells the host (b200vec_peer_status).
                    volatile unsigned int *arr = reinterpret_cast<volatile unsigned int *>(home + 2);
                    const unsigned int want = epoch * (unsigned)a.world;
                    const long long t0 = clock64();
                    bool ok = true;
                    while ((int)(*arr - want) < 0) {
                        if (clock64() - t0 > kWfPeerTimeoutClocks) { ok = false; break; }
                    }
                    __threadfence_system();
                    unsigned long long all = *reinterpret_cast<volatile unsigned long long *>(home + cur);
                    if (!ok) { a.ctr[10] = 1u; if (mine > all) all = mine; }
                    *fin = all;
                    __threadfence();
                    atomicExch(a.ctr + 6, epoch);
                }
            }
            if (tid == 0) {
                const long long t0 = clock64();
                while (*reinterpret_cast<volatile unsigned int *>(a.ctr + 6) != epoch) {
                    __nanosleep(20);
                    if (clock64() - t0 > 2 * kWfPeerTimeoutClocks) break;      // CTA 0 has flagged the error by then
                }
                __threadfence();
            }
            __syncthreads();
            slot = fin;
        }
        const double m2 = __longlong_as_double((long long)*reinterpret_cast<volatile unsigned long long *>(slot));
        if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);
        if (blockIdx.x == 0 && tid == 0 && a.maxsq_out) *a.maxsq_out = m2;
    }

    const long long n_vec = a.place.hi - a.place.lo;
    const float4 *wl = win4 + lane;
    const float2 *tq2 = tw + lane;                           // half twiddle table, slot 0 = W^(16 q)
    float2 *ex_w = exch + lane * kWfExPitch;                 // T1 writes row l = lane
    const float2 *ex_r = exch + lane;                        // T2 reads column q = lane
    unsigned anybits = 0;
    const bool track = a.flag != nullptr;
    const unsigned long long pol = wf_policy_evict_first();

    // PLAIN: the input halves (4 KB contiguous each) arrive by TMA bulk loads into the warp's staging block
    // (unused otherwise in this mode), one half ahead: issued when the previous half has been taken, waited
    // for through the warp's mbarrier (it lives in the unused zero block).  Needs a 16-byte aligned input.
    const bool pl_tma = PLAIN && BULK && (reinterpret_cast<unsigned long long>(a.vec) & 15ull) == 0;
    const unsigned pl_bar = zero_u32 + 8u * warp;
    unsigned pl_phase = 0;
    bool pl_primed = false;                                  // the first half of the next item is already on its way
    if (pl_tma) {
        if (lane == 0) mbar_init(pl_bar, 1);
        __syncwarp();
    }
    auto pl_issue = [&](long long half) {                    // all lanes; the staging block has been read out
        __syncwarp();
        if (lane == 0) {
            wf_fence_async();
            mbar_expect_tx(pl_bar, 4096u);
            tma_bulk_g2s(smem_u32(vst), a.vec + half * 512, 4096u, pl_bar);
        }
    };
    auto pl_take = [&](float2 (&x)[16]) {
        wf_mbar_wait_warp(pl_bar, pl_phase, lane);
        pl_phase ^= 1u;
#pragma unroll
        for (int j = 0; j < 16; ++j) x[j] = vst[lane + 32 * j];
    };

    for (;;) {
        const long long c = __shfl_sync(0xffffffffu, next, 0);
        if (c >= a.n_chunks) break;
        if (lane == 0) next = atomicAdd(a.ctr, 1u);         // the next item's index arrives while this one is processed
        const bool big = c < a.n_big;
        const int cn = big ? CHUNK : 1;                      // frames in this item
        const long long h0 = big ? c * CHUNK : a.n_big * CHUNK + (c - a.n_big);   // its first half-frame / frame (slab-relative)
        // ---------------------------------------------------------------- S
        int cnt_l = 0, rel0_l = 0;
        if (!PLAIN && lane < a.place.np) {
            const PlaceDesc &d = a.place.d[lane];
            const long long s0 = a.place.lo + h0 * 512;
            long long klo = floor_div(s0 - d.len - d.start, d.period) + 1;
            long long khi = floor_div(s0 + (cn + 1) * 512 - 1 - d.start, d.period);
            if (klo < 0) klo = 0;
            if (khi > d.count - 1) khi = d.count - 1;
            const long long cn = khi - klo + 1;
            cnt_l = cn > 0 ? (int)(cn < 0x00ffffff ? cn : 0x00ffffff) : 0;
            rel0_l = cnt_l ? (int)(s0 - (d.start + klo * d.period)) : 0;
        }
        const unsigned active = __ballot_sync(0xffffffffu, cnt_l > 0);

        float2 xo[16], xn[16];
        bool live_o = true;
        if (PLAIN && pl_tma) {
            if (!__shfl_sync(0xffffffffu, (int)pl_primed, 0)) pl_issue(h0);   // shuffled: provably warp-uniform for ptxas
            pl_take(xo);
            pl_issue(h0 + 1);                                // frame h0 exists, so does half h0 + 1
        } else if (PLAIN) {
            const float2 *src = a.vec + h0 * 512 + lane;
#pragma unroll
            for (int j = 0; j < 16; ++j) xo[j] = __ldcs(src + 32 * j);
        }
        // Fused mode: the loop starts one iteration early (i = -1 only places and stores the item's first
        // half), so that the gather and the store are instantiated once — the kernel's code size sits at an
        // instruction-cache cliff (64 KB), and the hot loop is what every warp fetches out of step.
#pragma unroll 1
        for (int i = PLAIN ? 0 : -1; i < cn; ++i) {
            const long long h = h0 + i;
            const bool fv = i >= 0 && h < a.n_frames;                         // frame h exists
            const bool sv = !PLAIN && (i + 1 < cn) && (h + 1) * 512 < n_vec;      // this item stores half h + 1
            if (i >= 0 && !fv && !sv) break;
            bool live_n = true;
            if (PLAIN && pl_tma) {
                pl_take(xn);
                if (i + 1 < cn && h + 1 < a.n_frames) {
                    pl_issue(h + 2);                         // the next frame of this item
                } else {                                     // last frame of the item: start on the next item
                    const long long cx = __shfl_sync(0xffffffffu, next, 0);
                    pl_primed = cx < a.n_chunks;
                    if (pl_primed) pl_issue(cx < a.n_big ? cx * CHUNK : a.n_big * CHUNK + (cx - a.n_big));
                }
            } else if (PLAIN) {
                const float2 *src = a.vec + (h + 1) * 512 + lane;
#pragma unroll
                for (int j = 0; j < 16; ++j) xn[j] = __ldcs(src + 32 * j);
            } else {
                live_n = wf_build_half(a, active, cnt_l, rel0_l, (i + 1) * 512, lane, xn, keep);
                wf_store_half<!(DBG & 2), BULK>(a.vec, sv ? (h + 1) * 512 : n_vec, n_vec, live_n, scl, lane, xn, vst, zero_u32, pol);
            }
            if (fv) {
                float *of = a.out + h * a.out_ld;
                if (live_o || live_n) {
                    // ------------------------------------------------------ T1
                    cf32 v[32];
#pragma unroll
                    for (int k = 0; k < 32; ++k) {           // v[k] = x[brev5(k)]
                        const int j = brev5(k);
                        const float2 s = j < 16 ? xo[j] : xn[j - 16];
                        v[k].x = s.x; v[k].y = s.y;
                    }
                    if (!(DBG & 4)) fft32_windowed_q(v, [wl](int k) { const float4 t = wl[32 * k]; return f32x4{t.x, t.y, t.z, t.w}; });
                    if (BULK) {                              // the last PSD frame has left the exchange block
                        if (lane == 0) wf_wait_read();
                        __syncwarp();
                    }
#pragma unroll
                    for (int q = 0; q < 32; ++q) ex_w[q] = make_float2(v[q].x, v[q].y);
                    __syncwarp();
                    // ------------------------------------------------------ T2
#pragma unroll
                    for (int l = 0; l < 32; ++l) {           // v[brev5(l)] = B_l[q]
                        const float2 s = ex_r[l * kWfExPitch];
                        v[brev5(l)].x = s.x; v[brev5(l)].y = s.y;
                    }
                    __syncwarp();                            // the block may be overwritten by the next frame
                    if (!(DBG & 4)) {
                        const float2 c = tq2[0];
                        fft32_twiddled_half(v, [tq2](int P) { const float2 w = tq2[32 * P]; return cf32{w.x, w.y}; }, cf32{c.x, c.y});
                    }
                    // BULK: the frame's 4 KB are staged in the (now free) exchange block and leave as one bulk copy
                    float *o = BULK ? reinterpret_cast<float *>(exch) + lane : of + lane;
#pragma unroll
                    for (int sp = 0; sp < 32; ++sp) {        // row q + 32 sp holds bin q + 32 s, s = (sp + 16) & 31
                        const int s = (sp + 16) & 31;
                        const float p = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
                        if (track) anybits |= __float_as_uint(p);   // p >= 0 or NaN: non-zero bits <=> p != 0
                        if (BULK) o[32 * sp] = p;
                        else if (!(DBG & 1)) __stcs(o + 32 * sp, p);
                    }
                    if (BULK) {
                        wf_fence_async();
                        __syncwarp();
                        if (lane == 0 && !(DBG & 1)) { wf_bulk_s2g(of, smem_u32(exch), 4096u, pol); tma_store_commit(); }
                    }
                } else if (BULK) {
                    if (lane == 0 && !(DBG & 1)) { wf_bulk_s2g(of, zero_u32, 4096u, pol); tma_store_commit(); }
                } else {
                    float4 *o4 = reinterpret_cast<float4 *>(of) + lane;
#pragma unroll
                    for (int k = 0; k < 8; ++k) if (!(DBG & 1)) __stcs(o4 + 32 * k, make_float4(0.f, 0.f, 0.f, 0.f));
                }
            }
#pragma unroll
            for (int j = 0; j < 16; ++j) xo[j] = xn[j];
            live_o = live_n;
        }
    }
    if (BULK && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (a.flag && __any_sync(0xffffffffu, anybits != 0) && lane == 0) atomicOr(a.flag, 1);
    __syncthreads();
    if (tid == 0) {                                          // the last CTA out re-arms the counters for the next launch
        __threadfence();
        const unsigned done = atomicAdd(a.ctr + 1, 1u);
        if (done == gridDim.x - 1) {
            a.ctr[0] = 0; a.ctr[1] = 0; a.ctr[2] = 0; a.ctr[4] = 0; a.ctr[5] = 0;
            if (!PLAIN && a.mode == kWfAutoNormPeer && a.epoch == 0)            // every CTA has read the old count by now
                *reinterpret_cast<unsigned int *>(a.peer[a.rank] + 3) = a.ctr[6];
            __threadfence();
        }
    }
}

template <int CHUNK, int DBG, bool BULK, bool PLAIN = false>
inline cudaError_t wf1024_attr() {
    cudaError_t e = cudaFuncSetAttribute(wf1024_kernel<CHUNK, DBG, BULK, PLAIN>, cudaFuncAttributePreferredSharedMemoryCarveout,
                                         (int)cudaSharedmemCarveoutMaxShared);
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(wf1024_kernel<CHUNK, DBG, BULK, PLAIN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)(BULK ? kWfSmemBulk : kWfSmemDirect));
}
inline int wf1024_setup() {
    B2_CUDA((wf1024_attr<4, 0, true>()));
    B2_CUDA((wf1024_attr<4, 0, true, true>()));
#ifdef B2_AB
    B2_CUDA((wf1024_attr<8, 0, true>())); B2_CUDA((wf1024_attr<2, 0, true>()));
    B2_CUDA((wf1024_attr<4, 0, false>())); B2_CUDA((wf1024_attr<8, 0, false>())); B2_CUDA((wf1024_attr<2, 0, false>()));
#ifdef B2_WF_ABLATION
    B2_CUDA((wf1024_attr<4, 1, false>())); B2_CUDA((wf1024_attr<4, 2, false>())); B2_CUDA((wf1024_attr<4, 3, false>()));
    B2_CUDA((wf1024_attr<4, 4, false>())); B2_CUDA((wf1024_attr<4, 7, false>()));
#endif
#endif
    return 0;
}

// Product path: 4-frame work items, bulk stores, single-frame items at the tail.  -DB2_AB builds read
// B200VEC_WF_CHUNK (2|4|8), B200VEC_WF_BULK=0 (direct stores), B200VEC_WF_TAIL=0, B200VEC_WF_V (ablations).
inline int launch_wf1024(b200vec_ctx *ctx, WfArgs &a, cudaStream_t st) {
    int chunk = kWfChunk, dbg = 0;
    bool bulk = true, tail = true;
    if (const char *e = ab_env("B200VEC_WF_CHUNK")) chunk = atoi(e);
    if (const char *e = ab_env("B200VEC_WF_V")) dbg = atoi(e);
    if (const char *e = ab_env("B200VEC_WF_BULK")) bulk = e[0] != '0';
    if (const char *e = ab_env("B200VEC_WF_TAIL")) tail = e[0] != '0';
    B2_REQUIRE(chunk == 2 || chunk == 4 || chunk == 8, "B200VEC_WF_CHUNK must be 2, 4 or 8");
    const long long n_vec = a.plain ? 0 : a.place.hi - a.place.lo;
    const long long halves = (n_vec + 511) / 512;
    const long long units = a.n_frames > halves ? a.n_frames : halves;
    // Guided tail: the last ~2 items' worth of frames per warp are handed out one frame at a time, so
    // the warps run dry within one frame time of each other instead of one item time.
    const long long n_warps = (long long)ctx->sm_count * kWfWarps;
    long long singles = tail ? 2 * n_warps : 0;
    if (singles > units) singles = units;
    a.n_big = (units - singles) / chunk;
    a.n_chunks = a.n_big + (units - a.n_big * chunk);
    if (a.n_chunks == 0) return 0;
    B2_REQUIRE(a.n_chunks < 0x7fffffffLL, "slab too long for one launch (split it)");
    if (!ctx->wf_ctr) {
        B2_CUDA(cudaMalloc(&ctx->wf_ctr, 16 * sizeof(unsigned int)));
        B2_CUDA(cudaMemsetAsync(ctx->wf_ctr, 0, 16 * sizeof(unsigned int), st));
    }
    a.ctr = ctx->wf_ctr;
    // The per-launch words are re-armed by the last CTA of every launch; zeroing them on the stream as well
    // keeps a launch independent of whatever an earlier (failed, timed-out) launch left behind.  One fused
    // launch may be in flight per context (the counters are the context's).
    B2_CUDA(cudaMemsetAsync(ctx->wf_ctr, 0, 6 * sizeof(unsigned int), st));
    const long long cap = ctx->sm_count;
    const long long need = (a.n_chunks + kWfWarps - 1) / kWfWarps;
    const unsigned grid = (unsigned)(need < cap ? need : cap);
    // the in-kernel max needs a grid barrier: cooperative launch (all CTAs resident, or the launch fails)
    void *kargs[] = {&a};
    if (a.plain) {
        wf1024_kernel<4, 0, true, true><<<grid, kWfThreads, kWfSmemBulk, st>>>(a);
        B2_CUDA(cudaGetLastError());
        return 0;
    }
#define B2_WF_LAUNCH(C, D, B)                                                                              \
    do {                                                                                                   \
        if (a.mode == kWfAutoNorm || a.mode == kWfAutoNormPeer)                                            \
            B2_CUDA(cudaLaunchCooperativeKernel((const void *)wf1024_kernel<C, D, B, false>, dim3(grid), dim3(kWfThreads), \
                                                kargs, (B) ? kWfSmemBulk : kWfSmemDirect, st));            \
        else                                                                                               \
            wf1024_kernel<C, D, B, false><<<grid, kWfThreads, (B) ? kWfSmemBulk : kWfSmemDirect, st>>>(a);        \
    } while (0)
#ifndef B2_AB
    (void)dbg; (void)bulk;
    B2_WF_LAUNCH(4, 0, true);
#else
    if (dbg == 0) {
        const int sel = (chunk == 8 ? 4 : chunk == 2 ? 2 : 0) | (bulk ? 1 : 0);
        switch (sel) {
            case 0: B2_WF_LAUNCH(4, 0, false); break;
            case 1: B2_WF_LAUNCH(4, 0, true); break;
            case 2: B2_WF_LAUNCH(2, 0, false); break;
            case 3: B2_WF_LAUNCH(2, 0, true); break;
            case 4: B2_WF_LAUNCH(8, 0, false); break;
            default: B2_WF_LAUNCH(8, 0, true); break;
        }
    } else {
#ifdef B2_WF_ABLATION
        switch (dbg) {
            case 1: B2_WF_LAUNCH(4, 1, false); break;
            case 2: B2_WF_LAUNCH(4, 2, false); break;
            case 3: B2_WF_LAUNCH(4, 3, false); break;
            case 4: B2_WF_LAUNCH(4, 4, false); break;
            default: B2_WF_LAUNCH(4, 7, false); break;
        }
#else
        return fail("launch_wf1024", "B200VEC_WF_V needs a -DB2_AB -DB2_WF_ABLATION build");
#endif
    }
#endif
#undef B2_WF_LAUNCH
    B2_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace b200vec
