// b200vec.cu — C-ABI of libb200vec.so (see include/b200vec.h for the contract and the
// reference lines each entry point replaces).  sm_100a only; no CPU fallback.
#include <mutex>
#include "common.cuh"
#include "bank.cuh"
#include "place.cuh"
#include "stft.cuh"
#include "stft1024.cuh"
#include "fused1024.cuh"
#ifdef B2_AB
#include "pipe1024.cuh"       // warp-specialised variant, measured slower: A/B builds only
#endif
#include "wf1024.cuh"
#include "span.cuh"
#include "packet.cuh"
#include "npquant.h"
#include "select.cuh"

namespace b200vec {
thread_local std::string g_last_error;

// K6: mat2wv quantiser (vector_analyzer/mat_to_wv_converter.py:42-50): x*32767 in float32,
// truncation toward zero, int16 I/Q interleave.  Out-of-range values wrap modulo 2^16 like
// the x86 cvttss2si + narrowing the reference's astype(int16) compiles to.
// pre_scale: mat2wv's bNormalize branch divides by max |x| first (mat_to_wv_converter.py:31) — numpy's
// complex64 / float32 scalar, i.e. a multiply by the float32 reciprocal, rounded, THEN the x32767.
__global__ void __launch_bounds__(256)
wv_quantize_kernel(const float2 *__restrict__ in, long long n, float pre_scale, short2 *__restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 v = __ldg(in + i);
        const int re = __float2int_rz(__fmul_rn(__fmul_rn(v.x, pre_scale), 32767.0f));
        const int im = __float2int_rz(__fmul_rn(__fmul_rn(v.y, pre_scale), 32767.0f));
        out[i] = make_short2((short)re, (short)im);
    }
}

// mat2wv's level metrics (mat_to_wv_converter.py:31-35) of x * pre_scale: stats[0] = bits of max |y|^2 (exact
// float32 products summed in fp64), stats[1] = sum over e[i] = fl32(fl32(|y|)^2) in fp64 (np.abs(signal)**2).
__global__ void __launch_bounds__(256)
power_stats_kernel(const float2 *__restrict__ in, long long n, float pre_scale, unsigned long long *__restrict__ stats) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    double mx = 0.0, sum = 0.0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 v = __ldg(in + i);
        const float x = __fmul_rn(v.x, pre_scale), y = __fmul_rn(v.y, pre_scale);
        const double e = (double)x * (double)x + (double)y * (double)y;
        const float a = (float)sqrt(e);
        mx = fmax(mx, e);
        sum += (double)__fmul_rn(a, a);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (mx > 0.0) atomicMax(stats, (unsigned long long)__double_as_longlong(mx));
        atomicAdd(reinterpret_cast<double *>(stats + 1), sum);
    }
}

// K7: complex64 interleaved -> planar float32 real / imaginary parts, the layout MAT v5 stores a
// complex array in (utils.py:659-670 -> scipy.io.savemat writes the pr and pi sub-elements apart).
__global__ void __launch_bounds__(256)
split_planes_kernel(const float2 *__restrict__ in, long long n, float *__restrict__ re, float *__restrict__ im) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long n4 = n / 4;
    const float4 *in4 = reinterpret_cast<const float4 *>(in);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 a = __ldcs(in4 + 2 * i), b = __ldcs(in4 + 2 * i + 1);
        __stcs(reinterpret_cast<float4 *>(re) + i, make_float4(a.x, a.z, b.x, b.z));
        __stcs(reinterpret_cast<float4 *>(im) + i, make_float4(a.y, a.w, b.y, b.w));
    }
    for (long long i = 4 * n4 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 v = in[i];
        re[i] = v.x; im[i] = v.y;
    }
}

__global__ void sel_init(unsigned int *hist, SelState init) {
    *reinterpret_cast<SelState *>(hist + kStateOff) = init;
}

struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

constexpr long long kSelSmallMax = 16LL << 20;   // elements: at or below, the select runs as one cooperative launch
constexpr size_t kPinnedBytes = 64 * 1024;       // select state, the 13-bit histogram (32 KB + 16 B), small results

static int run_select(b200vec_ctx *ctx, const float *data, long long n, int flags,
                      const int64_t *ranks, int nr, cudaStream_t st, SelState *host_state) {
    B2_CUDA(cudaMemsetAsync(ctx->hist, 0, sizeof(unsigned int) * kHistWords, st));
    SelState init;
    memset(&init, 0, sizeof(init));
    init.n_ranks = nr;
    for (int r = 0; r < nr; ++r) init.resid[r] = (unsigned long long)ranks[r];
    sel_init<<<1, 1, 0, st>>>(ctx->hist, init);
    if (n > 0 && n <= kSelSmallMax) {
        // candidate arrays, previews: the three histogram reads cost less than six dependent launches — one
        // cooperative launch with grid barriers between the levels
        long long blocks = (n / 4 + kPickThreads * kSelUnroll - 1) / (kPickThreads * kSelUnroll);
        if (blocks > ctx->sm_count) blocks = ctx->sm_count;
        if (blocks < 1) blocks = 1;
        const size_t smem = (size_t)(nr * kBins2 > kBins1 ? nr * kBins2 : kBins1) * sizeof(unsigned int);
        void *kargs[] = {(void *)&data, (void *)&n, (void *)&flags, (void *)&ctx->hist};
        B2_CUDA(cudaLaunchCooperativeKernel((const void *)sel_small_kernel, dim3((unsigned)blocks), dim3(kPickThreads),
                                            kargs, smem, st));
        B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->hist + kStateOff, sizeof(SelState), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        memcpy(host_state, ctx->pinned, sizeof(SelState));
        return 0;
    }
    const unsigned grid = (unsigned)(ctx->sm_count * 8);
    if (n > 0) sel_hist1<<<grid, kSelThreads, 0, st>>>(data, n, flags, ctx->hist);
    sel_pick<1><<<1, kPickThreads, 0, st>>>(ctx->hist);
    if (nr > 0 && n > 0) {
        sel_histN<2><<<grid, kSelThreads, (size_t)nr * kBins2 * 4, st>>>(data, n, flags, ctx->hist);
        sel_pick<2><<<1, kPickThreads, 0, st>>>(ctx->hist);
        sel_histN<3><<<grid, kSelThreads, (size_t)nr * kBins3 * 4, st>>>(data, n, flags, ctx->hist);
        sel_pick<3><<<1, kPickThreads, 0, st>>>(ctx->hist);
    }
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->hist + kStateOff, sizeof(SelState),
                            cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_state, ctx->pinned, sizeof(SelState));
    return 0;
}

// ---- b200vec_power_stats_static: the level metrics with NO allocation anywhere on the call path ----------
// The scratch of the two-stage reduction lives in a file-scope device array that every GPU loading the library
// holds until exit: one (max, sum) partial per chunk of kStaticChunk samples + the ticket that elects the
// last block.  kStaticMaxPartials chunks bound the input: 2^30 samples (8 GiB of complex64) for 256 KiB + 256 B.
constexpr int kStaticChunk = 1 << 16;                 // samples per partial
constexpr int kStaticMaxPartials = 1 << 14;
constexpr long long kStaticMaxSamples = (long long)kStaticChunk * kStaticMaxPartials;
constexpr int kStaticMaxDevices = 64;

struct StaticScratch {
    double2 part[kStaticMaxPartials];                 // .x = max |y|^2, .y = sum fl32(fl32(|y|)^2) of chunk c
    unsigned int ticket;                              // blocks that have finished; must be 0 at launch
    unsigned int pad[63];
};
static __device__ __align__(256) unsigned char g_static_scratch[sizeof(StaticScratch)];

// Fixed-order block reduction (shuffles, then the 8 warp results in order): the result does not depend on
// the grid size or on which block got which chunk.
__device__ __forceinline__ void static_block_reduce(double &mx, double &sum, double *sh) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    }
    const int w = threadIdx.x >> 5;
    __syncthreads();                                  // sh may still be read from the previous round
    if ((threadIdx.x & 31) == 0) { sh[2 * w] = mx; sh[2 * w + 1] = sum; }
    __syncthreads();
    mx = sh[0]; sum = sh[1];
#pragma unroll
    for (int i = 1; i < 8; ++i) { mx = fmax(mx, sh[2 * i]); sum += sh[2 * i + 1]; }
}

__global__ void __launch_bounds__(256)
power_stats_static_kernel(const float2 *__restrict__ in, long long n, int n_parts, float pre_scale,
                          StaticScratch *__restrict__ sc, double *__restrict__ out) {
    __shared__ double sh[16];
    __shared__ bool last;
    for (int c = blockIdx.x; c < n_parts; c += gridDim.x) {
        const long long lo = (long long)c * kStaticChunk;
        const long long hi = lo + kStaticChunk < n ? lo + kStaticChunk : n;
        double mx = 0.0, sum = 0.0;
        for (long long i = lo + threadIdx.x; i < hi; i += 256) {
            const float2 v = __ldg(in + i);
            const float x = __fmul_rn(v.x, pre_scale), y = __fmul_rn(v.y, pre_scale);
            const double e = (double)x * (double)x + (double)y * (double)y;
            const float a = (float)sqrt(e);
            mx = fmax(mx, e);
            sum += (double)__fmul_rn(a, a);
        }
        static_block_reduce(mx, sum, sh);
        if (threadIdx.x == 0) sc->part[c] = make_double2(mx, sum);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&sc->ticket, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    double mx = 0.0, sum = 0.0;
    for (int c = threadIdx.x; c < n_parts; c += 256) {
        const double2 p = __ldcg(&sc->part[c]);
        mx = fmax(mx, p.x);
        sum += p.y;
    }
    static_block_reduce(mx, sum, sh);
    if (threadIdx.x == 0) {
        out[0] = mx; out[1] = sum;
        sc->ticket = 0;                               // left clean for the next call
    }
}

struct StaticDevice {
    StaticScratch *scratch = nullptr;                 // cudaGetSymbolAddress of g_static_scratch ON this device
    int sm_count = 0;
    bool ready = false;
};
static StaticDevice g_static_dev[kStaticMaxDevices];
static std::mutex g_static_mutex;

// Once per device, with that device current: in host code `g_static_scratch` is only a handle for the runtime,
// the address a kernel can use comes from cudaGetSymbolAddress.
static int static_device_setup(int device, StaticDevice **out) {
    std::lock_guard<std::mutex> lk(g_static_mutex);
    StaticDevice &d = g_static_dev[device];
    if (!d.ready) {
        int major = 0;
        B2_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, device));
        if (major != 10)
            return fail("b200vec_power_stats_static", "kernels are built for sm_100a only (need a Blackwell B200-class GPU)");
        B2_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
        void *p = nullptr;
        B2_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return fail("b200vec_power_stats_static", "cudaGetSymbolAddress returned a NULL scratch address");
        d.scratch = reinterpret_cast<StaticScratch *>(p);
        d.ready = true;
    }
    *out = &d;
    return 0;
}

}  // namespace b200vec

using namespace b200vec;

extern "C" {

int b200vec_abi_version(void) { return B200VEC_ABI_VERSION; }

const char *b200vec_last_error(void) { return g_last_error.c_str(); }

int b200vec_create(int device, b200vec_ctx **out) {
    B2_REQUIRE(out != nullptr, "out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail("b200vec_create", "no CUDA device available (this library has no CPU fallback)");
    B2_REQUIRE(device >= 0 && device < count, "device ordinal out of range");
    DeviceGuard g(device);
    cudaDeviceProp prop;
    B2_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail("b200vec_create", "kernels are built for sm_100a only (need a Blackwell B200-class GPU)");
    b200vec_ctx *ctx = new b200vec_ctx();
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    B2_CUDA(cudaMalloc(&ctx->hist, sizeof(unsigned int) * kHistWords));
    B2_CUDA(cudaMallocHost(&ctx->pinned, kPinnedBytes));
    B2_CUDA(cudaMalloc(&ctx->top_hist, kTopBins * sizeof(unsigned int) + 2 * sizeof(unsigned long long)));
    B2_CUDA(cudaFuncSetAttribute(sel_histN<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 B200VEC_MAX_RANKS * kBins2 * 4));
    B2_CUDA(cudaFuncSetAttribute(sel_histN<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 B200VEC_MAX_RANKS * kBins3 * 4));
    B2_CUDA(cudaFuncSetAttribute(sel_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 B200VEC_MAX_RANKS * kBins2 * 4));
    B2_CUDA(cudaFuncSetAttribute(stft_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024));
    B2_CUDA(cudaFuncSetAttribute(stft_generic_f64_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024));
    if (stft1024_setup() || fused1024_setup() || wf1024_setup() || span_setup()) { delete ctx; return 1; }
#ifdef B2_AB
    if (pipe1024_setup()) { delete ctx; return 1; }
#endif
    *out = ctx;
    return 0;
}

int b200vec_destroy(b200vec_ctx *ctx) {
    if (!ctx) return 0;
    DeviceGuard g(ctx->device);
    cudaDeviceSynchronize();
    for (auto &s : ctx->slots) {
        if (s.raw) cudaFree(s.raw);
        if (s.store) cudaFree(s.store);
    }
    for (auto &kv : ctx->twiddles) cudaFree(kv.second);
    for (auto &s : ctx->fft64) if (s.ptr) cudaFree(s.ptr);
    if (ctx->hist) cudaFree(ctx->hist);
    if (ctx->zeros) cudaFree(ctx->zeros);
    if (ctx->wf_ctr) cudaFree(ctx->wf_ctr);
    if (ctx->sp_ctr) cudaFree(ctx->sp_ctr);
    if (ctx->top_hist) cudaFree(ctx->top_hist);
    if (ctx->cand.ptr) cudaFree(ctx->cand.ptr);
    if (ctx->batch.bank.ptr) cudaFree(ctx->batch.bank.ptr);
    if (ctx->batch.tables.ptr) cudaFree(ctx->batch.tables.ptr);
    if (ctx->batch.pin) cudaFreeHost(ctx->batch.pin);
    if (ctx->sel64.ptr) cudaFree(ctx->sel64.ptr);
    if (ctx->k5.ptr) cudaFree(ctx->k5.ptr);
    for (SpanCache *c : {&ctx->batch.spans, &ctx->max_spans}) {
        if (c->dev.ptr) cudaFree(c->dev.ptr);
        if (c->pin) cudaFreeHost(c->pin);
        if (c->pin_free) cudaEventDestroy(c->pin_free);
    }
    if (ctx->slot_max) cudaFree(ctx->slot_max);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    delete ctx;
    return 0;
}

// ---------------------------------------------------------------------------------------
int b200vec_bank_upload(b200vec_ctx *ctx, int slot, const void *src, int64_t len,
                        int src_is_device, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    B2_REQUIRE(len > 0 && len < (1LL << 31) - 2 * kTile, "packet length must be in (0, 2^31)");
    B2_REQUIRE(src, "src is NULL");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    BankSlot &sl = ctx->slots[slot];
    if (sl.cap_raw < len) {
        B2_CUDA(cudaStreamSynchronize(st));
        if (sl.raw) cudaFree(sl.raw);
        sl.raw = nullptr; sl.cap_raw = 0;
        B2_CUDA(cudaMalloc(&sl.raw, (size_t)len * sizeof(float2)));
        sl.cap_raw = len;
    }
    B2_CUDA(cudaMemcpyAsync(sl.raw, src, (size_t)len * sizeof(float2),
                            src_is_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
    sl.len_raw = len;
    sl.len = 0;
    return 0;
}

int b200vec_bank_resample(b200vec_ctx *ctx, int slot, int64_t len_out, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    BankSlot &sl = ctx->slots[slot];
    B2_REQUIRE(sl.len_raw > 0, "slot is empty");
    B2_REQUIRE(len_out > 0, "scipy.signal.resample needs num > 0");
    DeviceGuard g(ctx->device);
    return resample_slot(ctx, sl, len_out, (cudaStream_t)stream);
}

int b200vec_bank_shift(b200vec_ctx *ctx, int slot, double freq_hz, double fs_hz, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    BankSlot &sl = ctx->slots[slot];
    B2_REQUIRE(sl.len_raw > 0, "slot is empty");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t len = sl.len_raw;
    if (sl.cap_store < len) {
        B2_CUDA(cudaStreamSynchronize(st));
        if (sl.store) cudaFree(sl.store);
        sl.store = nullptr; sl.cap_store = 0;
        // per copy: pad | samples (rounded up to even) | pad (+1 spare for the 8-byte phase of copy B)
        // then the same two copies again, pre-scaled by 1/max (filled per build by bank_scale_kernel)
        const int64_t per = ((len + 1) & ~1LL) + 2 * kBankPad + 2;
        B2_CUDA(cudaMalloc(&sl.store, (size_t)(4 * per) * sizeof(float2)));
        sl.cap_store = len;
    }
    const int64_t per = ((sl.cap_store + 1) & ~1LL) + 2 * kBankPad + 2;
    B2_CUDA(cudaMemsetAsync(sl.store, 0, (size_t)(4 * per) * sizeof(float2), st));
    sl.per = per;
    sl.external = false;
    sl.a = sl.store + kBankPad;                 // cudaMalloc is 256-byte aligned, kBankPad is even -> 16-byte aligned
    sl.b = sl.store + per + kBankPad + 1;       // per is even -> 8 mod 16
    // the reference forms the phase as ((2j*pi)*f)*t: w = fl(fl(2*pi)*f)   (utils.py:126)
    const double w = (2.0 * M_PI) * freq_hz;
    shift_commit_kernel<<<(unsigned)((len + 255) / 256), 256, 0, st>>>(
        sl.raw, len, w, fs_hz, freq_hz != 0.0 ? 1 : 0, sl.a, sl.b);
    B2_CUDA(cudaGetLastError());
    if (!ctx->slot_max) B2_CUDA(cudaMalloc(&ctx->slot_max, sizeof(unsigned long long) * B200VEC_MAX_SLOTS));
    B2_CUDA(cudaMemsetAsync(ctx->slot_max + slot, 0, sizeof(unsigned long long), st));
    bank_maxsq_kernel<<<(unsigned)((len + 1023) / 1024 < 64 ? (len + 1023) / 1024 : 64), 256, 0, st>>>(sl.a, len, ctx->slot_max + slot);
    B2_CUDA(cudaGetLastError());
    sl.len = len;
    return 0;
}

int b200vec_bank_len(b200vec_ctx *ctx, int slot, int64_t *len_out) {
    B2_REQUIRE(ctx && len_out, "NULL argument");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    *len_out = ctx->slots[slot].len ? ctx->slots[slot].len : ctx->slots[slot].len_raw;
    return 0;
}

int b200vec_bank_download(b200vec_ctx *ctx, int slot, void *dst_host, int64_t len, void *stream) {
    B2_REQUIRE(ctx && dst_host, "NULL argument");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    BankSlot &sl = ctx->slots[slot];
    B2_REQUIRE(sl.len > 0, "slot is not committed (call b200vec_bank_shift first)");
    B2_REQUIRE(len == sl.len, "length mismatch");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    B2_CUDA(cudaMemcpyAsync(dst_host, sl.a, (size_t)len * sizeof(float2), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------
static int fill_place_args(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                           int64_t slab_lo, int64_t slab_hi, PlaceArgs &args, int *slot_of = nullptr,
                           bool need_copy_b = false) {
    memset(&args, 0, sizeof(args));
    args.lo = slab_lo; args.hi = slab_hi;
    int np = 0;
    for (int i = 0; i < n_pkts; ++i) {
        const b200vec_pkt_desc &d = descs[i];
        B2_REQUIRE(d.slot >= 0 && d.slot < B200VEC_MAX_SLOTS, "descriptor slot out of range");
        const BankSlot &sl = ctx->slots[d.slot];
        B2_REQUIRE(sl.len > 0, "descriptor names an uncommitted bank slot");
        B2_REQUIRE(!need_copy_b || !sl.external, "attached (caller-owned) bank slots work with the frame-major / vector-only builds");
        B2_REQUIRE(d.period > 0, "period must be positive");
        B2_REQUIRE(d.start >= 0 && d.count >= 0, "negative start/count");
        if (d.count == 0) continue;
        if (slot_of) slot_of[np] = d.slot;
        PlaceDesc &p = args.d[np++];
        p.start = d.start; p.period = d.period; p.count = d.count;
        p.len = (int)sl.len; p.a = sl.a; p.b = sl.b;
        p.free_inst = d.reserved & 1; p.pmax = ctx->slot_max + d.slot;
    }
    args.np = np;
    return 0;
}

// Max-pass planner (maxspans.cuh): the must-evaluate spans of `ranges` for the packets of `pa`, uploaded into the
// context's cache (reused while descriptors and ranges do not change — every replay of a captured step, every
// repeated step of one scenario).
static int upload_spans(SpanCache &c, unsigned long long key, const std::vector<MaxSpan> &spans, cudaStream_t st) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    B2_CUDA(cudaStreamIsCapturing(st, &cs));
    B2_REQUIRE(cs == cudaStreamCaptureStatusNone,
               "the max-span list of this (descriptors, ranges) pair is not uploaded yet: run the step once before capturing it in a CUDA graph");
    if (ensure_scratch(c.dev, (spans.size() + 1) * sizeof(MaxSpan))) return 1;
    c.n = -1;
    if (!spans.empty()) {
        const size_t bytes = spans.size() * sizeof(MaxSpan);
        if (!c.pin_free) B2_CUDA(cudaEventCreateWithFlags(&c.pin_free, cudaEventDisableTiming));
        else B2_CUDA(cudaEventSynchronize(c.pin_free));               // the previous upload has left the staging block
        if (c.pin_bytes < bytes) {
            if (c.pin) cudaFreeHost(c.pin);
            c.pin = nullptr; c.pin_bytes = 0;
            B2_CUDA(cudaMallocHost(&c.pin, bytes + bytes / 4));
            c.pin_bytes = bytes + bytes / 4;
        }
        memcpy(c.pin, spans.data(), bytes);
        B2_CUDA(cudaMemcpyAsync(c.dev.ptr, c.pin, bytes, cudaMemcpyHostToDevice, st));
        B2_CUDA(cudaEventRecord(c.pin_free, st));
    }
    c.key = key; c.n = (long long)spans.size();
    return 0;
}

static int plan_max_spans(b200vec_ctx *ctx, const PlaceArgs &pa, const long long *ranges, int n_ranges,
                          cudaStream_t st, const MaxSpan **spans_dev, long long *n_spans) {
    MaxSpanPacket pk[B200VEC_MAX_PACKETS];
    for (int p = 0; p < pa.np; ++p)
        pk[p] = MaxSpanPacket{pa.d[p].start, pa.d[p].period, pa.d[p].count, (long long)pa.d[p].len, pa.d[p].free_inst};
    const unsigned long long key = max_spans_key(pk, pa.np, ranges, n_ranges);
    SpanCache &c = ctx->max_spans;
    if (c.n < 0 || c.key != key) {
        std::vector<MaxSpan> spans;
        max_spans(pk, pa.np, ranges, n_ranges, 0, spans);
        if (upload_spans(c, key, spans, st)) return 1;
    }
    *spans_dev = (const MaxSpan *)c.dev.ptr; *n_spans = c.n;
    return 0;
}

static int build_impl(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts, int64_t slab_lo, int64_t slab_hi,
                      void *out_dev, int mode, const int64_t *norm_ranges, int n_ranges, double *maxsq_dev,
                      void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(n_pkts >= 0 && n_pkts <= B200VEC_MAX_PACKETS, "too many packets");
    B2_REQUIRE(n_pkts == 0 || descs, "descs is NULL");
    B2_REQUIRE(slab_lo >= 0 && slab_hi >= slab_lo, "bad slab");
    B2_REQUIRE(mode == B200VEC_BUILD_MAXONLY || out_dev || slab_hi == slab_lo, "out_dev is NULL");
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || maxsq_dev, "maxsq_dev is NULL");
    B2_REQUIRE(((uintptr_t)out_dev & 15) == 0, "out_dev must be 16-byte aligned");
    SpanArgs a;
    memset(&a, 0, sizeof(a));
    if (fill_place_args(ctx, descs, n_pkts, slab_lo, slab_hi, a.place)) return 1;
    a.n_vec = 1; a.lo = slab_lo; a.hi = slab_hi;
    a.out = (float2 *)out_dev; a.out_ld = 0;
    a.maxsq = reinterpret_cast<unsigned long long *>(maxsq_dev);
    a.mode = mode;
    DeviceGuard g(ctx->device);
    if (mode == B200VEC_BUILD_MAXONLY) {
        if (slab_hi == slab_lo) return 0;
        const long long rg[2] = {slab_lo, slab_hi};
        if (plan_max_spans(ctx, a.place, rg, 1, (cudaStream_t)stream, &a.spans, &a.n_spans)) return 1;
    } else if (mode == kSpNorm) {
        B2_REQUIRE(n_ranges >= 0 && n_ranges <= kWfMaxNormRanges && (n_ranges == 0 || norm_ranges), "bad norm ranges");
        long long rg[2 * kWfMaxNormRanges];
        int nr = 0;
        for (int r = 0; r < n_ranges; ++r) {
            const int64_t lo = norm_ranges[2 * r], hi = norm_ranges[2 * r + 1];
            B2_REQUIRE(lo >= 0 && hi >= lo, "bad norm range");
            if (hi == lo) continue;
            rg[2 * nr] = lo; rg[2 * nr + 1] = hi; ++nr;
        }
        if (plan_max_spans(ctx, a.place, rg, nr, (cudaStream_t)stream, &a.spans, &a.n_spans)) return 1;
    }
#ifdef B2_AB
    if (mode != kSpNorm && ab_env("B200VEC_PLACE")) {      // round-1 kernel: needs both bank copies
        PlaceArgs pa;
        if (fill_place_args(ctx, descs, n_pkts, slab_lo, slab_hi, pa, nullptr, true)) return 1;
        return launch_place(ctx, pa, (float2 *)out_dev, maxsq_dev, mode, (cudaStream_t)stream);
    }
#endif
    if (mode == kSpNorm) B2_CUDA(cudaMemsetAsync(maxsq_dev, 0, sizeof(double), (cudaStream_t)stream));
    return launch_span(ctx, a, (cudaStream_t)stream);
}

int b200vec_build(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                  int64_t slab_lo, int64_t slab_hi, void *out_dev, int mode,
                  double *maxsq_dev, void *stream) {
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || mode == B200VEC_BUILD_MAXONLY || mode == B200VEC_BUILD_SCALE, "unknown mode");
    return build_impl(ctx, descs, n_pkts, slab_lo, slab_hi, out_dev, mode, nullptr, 0, maxsq_dev, stream);
}

int b200vec_build_norm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                       int64_t slab_lo, int64_t slab_hi, void *out_dev,
                       const int64_t *norm_ranges, int n_ranges, double *maxsq_dev, void *stream) {
    return build_impl(ctx, descs, n_pkts, slab_lo, slab_hi, out_dev, kSpNorm, norm_ranges, n_ranges, maxsq_dev, stream);
}

int b200vec_bank_attach(b200vec_ctx *ctx, int slot, const void *samples_dev, int64_t len, void *stream) {
    B2_REQUIRE(ctx && samples_dev, "NULL argument");
    B2_REQUIRE(slot >= 0 && slot < B200VEC_MAX_SLOTS, "slot out of range");
    B2_REQUIRE(len > 0 && len < (1LL << 31) - 2 * kTile, "packet length must be in (0, 2^31)");
    B2_REQUIRE(((uintptr_t)samples_dev & 7) == 0, "samples must be 8-byte aligned");
    DeviceGuard g(ctx->device);
    BankSlot &sl = ctx->slots[slot];
    if (!ctx->slot_max) B2_CUDA(cudaMalloc(&ctx->slot_max, sizeof(unsigned long long) * B200VEC_MAX_SLOTS));
    B2_CUDA(cudaMemsetAsync(ctx->slot_max + slot, 0, sizeof(unsigned long long), (cudaStream_t)stream));
    sl.a = const_cast<float2 *>((const float2 *)samples_dev);
    sl.b = nullptr;
    sl.external = true;
    sl.len = len;
    return 0;
}

// ---------------------------------------------------------------------------------------
// scenario batch (BASELINE.json configs[4]; the loop of unified_gui.py:1722-1782 once per scenario)
int b200vec_batch_prepare(b200vec_ctx *ctx, const b200vec_batch_pkt *pkts, const int32_t *first, int n_vectors,
                          double fs_hz, void *stream) {
    B2_REQUIRE(ctx && first && n_vectors >= 1, "NULL argument or empty batch");
    B2_REQUIRE(first[0] == 0, "first[0] must be 0");
    const int n_pk = first[n_vectors];
    B2_REQUIRE(n_pk == 0 || pkts, "pkts is NULL");
    int stride = 1;
    for (int v = 0; v < n_vectors; ++v) {
        const int c = first[v + 1] - first[v];
        B2_REQUIRE(c >= 0 && c <= B200VEC_MAX_PACKETS, "packets per vector out of range");
        if (c > stride) stride = c;
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    BatchState &bs = ctx->batch;
    // distinct (source slot, shift) pairs -> shift jobs
    std::map<std::pair<int, unsigned long long>, int> job_of;
    std::vector<int> pk_job(n_pk);
    std::vector<long long> job_off, job_len;
    std::vector<double> job_freq;
    std::vector<int> job_slot;
    long long bank_samples = 0;
    for (int i = 0; i < n_pk; ++i) {
        const b200vec_batch_pkt &q = pkts[i];
        B2_REQUIRE(q.src_slot >= 0 && q.src_slot < B200VEC_MAX_SLOTS, "source slot out of range");
        const BankSlot &sl = ctx->slots[q.src_slot];
        B2_REQUIRE(sl.len_raw > 0 && sl.raw, "source slot is empty (b200vec_bank_upload it first)");
        B2_REQUIRE(q.period > 0, "period must be positive");
        B2_REQUIRE(q.start >= 0 && q.count >= 0, "negative start/count");
        unsigned long long fb;
        const double f = q.freq_hz == 0.0 ? 0.0 : q.freq_hz;          // -0.0 and 0.0 are the same job
        memcpy(&fb, &f, 8);
        auto key = std::make_pair(q.src_slot, fb);
        auto it = job_of.find(key);
        if (it == job_of.end()) {
            const int j = (int)job_off.size();
            job_of[key] = j;
            job_off.push_back(bank_samples); job_len.push_back(sl.len_raw); job_freq.push_back(f); job_slot.push_back(q.src_slot);
            bank_samples += (sl.len_raw + 1) & ~1LL;                  // every packet starts on a 16-byte boundary
            pk_job[i] = j;
        } else {
            pk_job[i] = it->second;
        }
    }
    const int n_jobs = (int)job_off.size();
    B2_REQUIRE(n_jobs <= 65535, "too many distinct (packet, shift) pairs for one batch");
    const size_t b_desc = (size_t)n_vectors * stride * sizeof(PlaceDesc);
    const size_t b_np = ((size_t)n_vectors * sizeof(int) + 15) & ~(size_t)15;
    const size_t b_jobs = (size_t)(n_jobs > 0 ? n_jobs : 1) * sizeof(ShiftJob);
    const size_t b_pmax = (size_t)(n_jobs > 0 ? n_jobs : 1) * sizeof(unsigned long long);
    const size_t b_tables = b_desc + b_np + b_jobs + b_pmax;
    if (bs.bank.bytes < (size_t)bank_samples * sizeof(float2) || bs.tables.bytes < b_tables || bs.pin_bytes < b_tables)
        B2_CUDA(cudaStreamSynchronize(st));                            // nothing in flight reads the buffers being replaced
    if (ensure_scratch(bs.bank, (size_t)(bank_samples > 0 ? bank_samples : 1) * sizeof(float2))) return 1;
    if (ensure_scratch(bs.tables, b_tables)) return 1;
    if (bs.pin_bytes < b_tables) {
        if (bs.pin) cudaFreeHost(bs.pin);
        bs.pin = nullptr; bs.pin_bytes = 0;
        B2_CUDA(cudaMallocHost(&bs.pin, b_tables));
        bs.pin_bytes = b_tables;
    }
    char *hp = (char *)bs.pin, *dp = (char *)bs.tables.ptr;
    PlaceDesc *h_desc = (PlaceDesc *)hp;
    int *h_np = (int *)(hp + b_desc);
    ShiftJob *h_jobs = (ShiftJob *)(hp + b_desc + b_np);
    unsigned long long *h_pmax = (unsigned long long *)(hp + b_desc + b_np + b_jobs);
    unsigned long long *d_pmax = (unsigned long long *)(dp + b_desc + b_np + b_jobs);
    float2 *bank = (float2 *)bs.bank.ptr;
    memset(hp, 0, b_tables);
    long long max_len = 1;
    for (int j = 0; j < n_jobs; ++j) {
        const BankSlot &sl = ctx->slots[job_slot[j]];
        h_jobs[j].src = sl.raw; h_jobs[j].dst = bank + job_off[j]; h_jobs[j].pmax = d_pmax + j;
        h_jobs[j].len = job_len[j];
        h_jobs[j].w = (2.0 * M_PI) * job_freq[j];                      // ((2j*pi)*f) as utils.py:126 forms it
        h_jobs[j].do_shift = job_freq[j] != 0.0 ? 1 : 0;
        h_pmax[j] = 0;
        if (job_len[j] > max_len) max_len = job_len[j];
    }
    bs.h_pk.assign((size_t)n_vectors * stride, MaxSpanPacket{0, 1, 0, 0, 0});
    bs.h_np.assign(n_vectors, 0);
    for (int v = 0; v < n_vectors; ++v) {
        int np = 0;
        for (int i = first[v]; i < first[v + 1]; ++i) {
            const b200vec_batch_pkt &q = pkts[i];
            if (q.count == 0) continue;
            bs.h_pk[(size_t)v * stride + np] = MaxSpanPacket{q.start, q.period, q.count, job_len[pk_job[i]], q.flags & 1};
            bs.h_np[v] = np + 1;
            PlaceDesc &p = h_desc[(size_t)v * stride + np++];
            p.start = q.start; p.period = q.period; p.count = q.count;
            p.len = (int)job_len[pk_job[i]];
            p.free_inst = q.flags & 1;
            p.pmax = d_pmax + pk_job[i];
            p.a = bank + job_off[pk_job[i]];
            p.b = nullptr;
        }
        h_np[v] = np;
    }
    B2_CUDA(cudaMemcpyAsync(dp, hp, b_tables, cudaMemcpyHostToDevice, st));
    if (n_jobs > 0) {
        long long gx = (max_len + 255) / 256;
        if (gx > 64) gx = 64;
        batch_shift_kernel<<<dim3((unsigned)gx, (unsigned)n_jobs), 256, 0, st>>>((const ShiftJob *)(dp + b_desc + b_np), fs_hz);
        B2_CUDA(cudaGetLastError());
    }
    B2_CUDA(cudaStreamSynchronize(st));                                // the pinned staging block may be reused now
    bs.descs = dp; bs.np = (const int *)(dp + b_desc);
    bs.n_vec = n_vectors; bs.pk_stride = stride; bs.n_jobs = n_jobs;
    return 0;
}

int b200vec_batch_build(b200vec_ctx *ctx, int64_t total, void *out_dev, int64_t out_ld, int normalize,
                        double *maxsq_dev, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    BatchState &bs = ctx->batch;
    B2_REQUIRE(bs.n_vec > 0 && bs.descs, "no prepared batch (call b200vec_batch_prepare first)");
    B2_REQUIRE(total >= 0 && (total == 0 || out_dev), "bad arguments");
    B2_REQUIRE(out_ld >= total && (out_ld & 1) == 0 && ((uintptr_t)out_dev & 15) == 0,
               "out_ld >= total, even, and a 16-byte aligned base");
    B2_REQUIRE(!normalize || maxsq_dev, "maxsq_dev is NULL");
    if (total == 0) return 0;
    SpanArgs a;
    memset(&a, 0, sizeof(a));
    a.gdesc = (const PlaceDesc *)bs.descs; a.gnp = bs.np; a.n_vec = bs.n_vec; a.pk_stride = bs.pk_stride;
    a.lo = 0; a.hi = total; a.out = (float2 *)out_dev; a.out_ld = out_ld;
    a.maxsq = reinterpret_cast<unsigned long long *>(maxsq_dev);
    a.mode = normalize ? kSpNorm : B200VEC_BUILD_STORE;
    DeviceGuard g(ctx->device);
    if (normalize) {
        // per vector: the ranges that hold its max (one common period of its trains + head and tail, or the whole
        // vector), reduced to the spans under >= 2 instances (maxspans.cuh); planned once per (batch, total)
        long long tt[2] = {0, total};
        const unsigned long long key = max_spans_key(bs.h_pk.data(), (int)bs.h_pk.size(), tt, 1);
        if (bs.spans.n < 0 || bs.spans.key != key) {
            std::vector<MaxSpan> spans;
            for (int v = 0; v < bs.n_vec; ++v) {
                long long rg[4];
                const MaxSpanPacket *pk = bs.h_pk.data() + (size_t)v * bs.pk_stride;
                const int nr = max_eval_ranges(pk, bs.h_np[v], total, rg);
                max_spans(pk, bs.h_np[v], rg, nr, v, spans);
            }
            if (upload_spans(bs.spans, key, spans, (cudaStream_t)stream)) return 1;
        }
        a.spans = (const MaxSpan *)bs.spans.dev.ptr; a.n_spans = bs.spans.n;
    }
    if (normalize) B2_CUDA(cudaMemsetAsync(maxsq_dev, 0, sizeof(double) * bs.n_vec, (cudaStream_t)stream));
    return launch_span(ctx, a, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
int b200vec_replicate_frames(b200vec_ctx *ctx, float *spec_dev, int64_t out_ld, int64_t frame_src0, int64_t period_frames,
                             int64_t frame_dst0, int64_t n_frames, void *stream) {
    B2_REQUIRE(ctx && spec_dev, "NULL argument");
    B2_REQUIRE(out_ld >= 1024 && (out_ld & 3) == 0 && ((uintptr_t)spec_dev & 15) == 0,
               "frame-major spectrogram: out_ld >= 1024, a multiple of 4, 16-byte aligned base");
    B2_REQUIRE(frame_src0 >= 0 && period_frames >= 1 && n_frames >= 0 && frame_dst0 >= frame_src0 + period_frames,
               "the destination frames must lie past the source period");
    if (n_frames == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n_frames + 7) / 8;
    const long long cap = (long long)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    replicate_frames_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(spec_dev, out_ld, frame_src0, period_frames,
                                                                               frame_dst0, n_frames);
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_stft(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                 int64_t stride, int nperseg, int hop, int nfft, const float *window_dev,
                 float scale, int64_t frame_lo, int64_t n_frames, float *out_dev,
                 int64_t out_ld, int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(ctx && in_dev && out_dev && window_dev, "NULL argument");
    B2_REQUIRE(nfft >= 32 && nfft <= 16384 && (nfft & (nfft - 1)) == 0,
               "nfft must be a power of two in [32, 16384]");
    B2_REQUIRE(nperseg >= 1 && nperseg <= nfft, "nfft must be greater than or equal to nperseg.");
    B2_REQUIRE(hop >= 1, "noverlap must be less than nperseg.");
    B2_REQUIRE(stride >= 1 && in_len >= 0 && n_frames >= 0 && out_ld >= n_frames, "bad geometry");
    if (n_frames == 0) return 0;
    {   // every sample the requested frames read must lie inside in_dev[0, in_len)
        const long long first = frame_lo * (long long)hop * stride - in_origin;
        const long long last = ((frame_lo + n_frames - 1) * (long long)hop + nperseg - 1) * stride - in_origin;
        B2_REQUIRE(frame_lo >= 0 && first >= 0 && last < in_len, "frames read outside the input buffer");
    }
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    if (nfft == 1024 && stft1024_enabled()) {
        return launch_stft1024(ctx, (const float2 *)in_dev, in_origin, in_len, stride, nperseg, hop,
                               window_dev, scale, frame_lo, n_frames, out_dev, out_ld,
                               nonzero_flag_dev, st);
    }
    StftArgs a;
    a.in = (const float2 *)in_dev; a.in_origin = in_origin; a.in_len = in_len; a.stride = stride;
    a.nperseg = nperseg; a.hop = hop; a.nfft = nfft; a.log2n = ceil_log2(nfft);
    a.window = window_dev; a.scale = scale; a.frame_lo = frame_lo; a.n_frames = n_frames;
    a.out = out_dev; a.out_ld = out_ld; a.flag = nonzero_flag_dev;
    a.T = stft_pick_T(nfft);
    if (get_twiddles(ctx, nfft, &a.tw)) return 1;
    const size_t smem = (size_t)nfft * 8 + (a.T > 1 ? (size_t)nfft * 4 * (a.T + 1) : 0);
    const long long grid = (n_frames + a.T - 1) / a.T;
    B2_REQUIRE(grid < (1LL << 31), "too many frames for one launch");
    stft_generic_kernel<<<(unsigned)grid, kStftThreads, smem, st>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

// scipy's float64 path (complex128 input): window, FFT and PSD in double.
int b200vec_stft_f64(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                     int64_t stride, int nperseg, int hop, int nfft, const double *window_dev,
                     double scale, int64_t frame_lo, int64_t n_frames, double *out_dev,
                     int64_t out_ld, int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(ctx && in_dev && out_dev && window_dev, "NULL argument");
    B2_REQUIRE(nfft >= 32 && nfft <= 8192 && (nfft & (nfft - 1)) == 0,
               "nfft must be a power of two in [32, 8192] for the float64 STFT");
    B2_REQUIRE(nperseg >= 1 && nperseg <= nfft, "nfft must be greater than or equal to nperseg.");
    B2_REQUIRE(hop >= 1, "noverlap must be less than nperseg.");
    B2_REQUIRE(stride >= 1 && in_len >= 0 && n_frames >= 0 && out_ld >= n_frames, "bad geometry");
    B2_REQUIRE(((uintptr_t)in_dev & 15) == 0, "in_dev must be 16-byte aligned");
    if (n_frames == 0) return 0;
    {
        const long long first = frame_lo * (long long)hop * stride - in_origin;
        const long long last = ((frame_lo + n_frames - 1) * (long long)hop + nperseg - 1) * stride - in_origin;
        B2_REQUIRE(frame_lo >= 0 && first >= 0 && last < in_len, "frames read outside the input buffer");
    }
    DeviceGuard g(ctx->device);
    Stft64Args a;
    a.in = (const double2 *)in_dev; a.in_origin = in_origin; a.in_len = in_len; a.stride = stride;
    a.nperseg = nperseg; a.hop = hop; a.nfft = nfft; a.log2n = ceil_log2(nfft);
    a.window = window_dev; a.scale = scale; a.frame_lo = frame_lo; a.n_frames = n_frames;
    a.out = out_dev; a.out_ld = out_ld; a.flag = nonzero_flag_dev;
    a.T = stft64_pick_T(nfft);
    if (get_twiddles64(ctx, nfft, &a.tw)) return 1;
    const size_t smem = (size_t)nfft * 16 + (a.T > 1 ? (size_t)nfft * 8 * (a.T + 1) : 0);
    const long long grid = (n_frames + a.T - 1) / a.T;
    B2_REQUIRE(grid < (1LL << 31), "too many frames for one launch");
    stft_generic_f64_kernel<<<(unsigned)grid, kStftThreads, smem, (cudaStream_t)stream>>>(a);
    B2_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------------------------------
int b200vec_build_stft1024(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                           int64_t slab_lo, int64_t slab_hi, void *vec_out_dev, int mode,
                           double *maxsq_dev, const float *window_dev, float scale,
                           int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                           int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(n_pkts >= 0 && n_pkts <= B200VEC_MAX_PACKETS, "too many packets");
    B2_REQUIRE(n_pkts == 0 || descs, "descs is NULL");
    B2_REQUIRE(slab_lo >= 0 && slab_hi >= slab_lo, "bad slab");
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || mode == B200VEC_BUILD_SCALE, "mode must be STORE or SCALE");
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || maxsq_dev, "maxsq_dev is NULL");
    B2_REQUIRE(vec_out_dev || slab_hi == slab_lo, "vec_out_dev is NULL");
    B2_REQUIRE(((uintptr_t)vec_out_dev & 15) == 0, "vec_out_dev must be 16-byte aligned");
    B2_REQUIRE(window_dev && (spec_out_dev || n_frames == 0), "NULL argument");
    B2_REQUIRE(frame_lo >= 0 && n_frames >= 0 && out_ld >= n_frames, "bad geometry");
    B2_REQUIRE(out_ld < (1LL << 24), "out_ld must be below 2^24 frames per launch (split the slab)");
    B2_REQUIRE(slab_lo == frame_lo * kFMaxHop, "slab_lo must equal frame_lo * 512 (slabs are cut on hop boundaries)");
    B2_REQUIRE(n_frames * kFMaxHop <= slab_hi - slab_lo + kFMaxHop - 1, "frames must start inside the slab");
    FArgs a;
    int slot_of[B200VEC_MAX_PACKETS];
    if (fill_place_args(ctx, descs, n_pkts, slab_lo, slab_hi, a.place, slot_of, true)) return 1;
    DeviceGuard g(ctx->device);
    if (!ctx->zeros) {                                     // TMA source for the uncovered part of a span
        B2_CUDA(cudaMalloc(&ctx->zeros, FL<16>::kSpan * sizeof(float2)));
        B2_CUDA(cudaMemsetAsync(ctx->zeros, 0, FL<16>::kSpan * sizeof(float2), (cudaStream_t)stream));
    }
    a.zeros = ctx->zeros;
    for (int p = 0; p < B200VEC_MAX_PACKETS; ++p) a.soff[p] = 0;
    if (mode == B200VEC_BUILD_SCALE && a.place.np > 0) {   // pre-scaled bank copies: fl(y * fl(1 / max))
        BankScaleArgs sa;
        memset(&sa, 0, sizeof(sa));
        for (int p = 0; p < a.place.np; ++p) {
            BankSlot &sl = ctx->slots[slot_of[p]];
            a.soff[p] = 2 * sl.per;
            bool seen = false;
            for (int q = 0; q < p; ++q) seen = seen || slot_of[q] == slot_of[p];
            if (seen) continue;
            sa.store[sa.n] = sl.store; sa.count[sa.n] = 2 * sl.per; ++sa.n;
        }
        sa.maxsq = reinterpret_cast<const unsigned long long *>(maxsq_dev);
        bank_scale_kernel<<<dim3(64, sa.n), 256, 0, (cudaStream_t)stream>>>(sa);
        B2_CUDA(cudaGetLastError());
    }
    a.vec = (float2 *)vec_out_dev;
    a.maxsq = reinterpret_cast<const unsigned long long *>(maxsq_dev);
    a.mode = mode; a.hop = kFMaxHop; a.in_origin = slab_lo;
    a.window = window_dev; a.scale = scale;
    a.frame_lo = frame_lo; a.n_frames = n_frames; a.out = spec_out_dev; a.out_ld = out_ld;
    a.flag = nonzero_flag_dev;
    if (get_tw1024(ctx, &a.tw)) return 1;
#ifdef B2_AB
    if (pipe1024_enabled()) return launch_pipe1024<true>(ctx, a, n_frames, slab_hi - slab_lo, (cudaStream_t)stream);
#endif
    return launch_tile1024<true>(ctx, a, n_frames, slab_hi - slab_lo, (cudaStream_t)stream);
}

static int build_stft1024_fm_impl(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                                  int64_t slab_lo, int64_t slab_hi, void *vec_out_dev, int mode,
                                  const int64_t *norm_ranges, int n_ranges,
                                  void *const *peer_blocks, int world, int rank, unsigned int epoch,
                                  double *maxsq_dev, const float *window_dev, float scale,
                                  int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                                  int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(n_pkts >= 0 && n_pkts <= B200VEC_MAX_PACKETS, "too many packets");
    B2_REQUIRE(n_pkts == 0 || descs, "descs is NULL");
    B2_REQUIRE(slab_lo >= 0 && slab_hi >= slab_lo, "bad slab");
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || mode == B200VEC_BUILD_SCALE || mode == kWfAutoNorm ||
               mode == kWfAutoNormPeer, "bad mode");
    B2_REQUIRE(mode != B200VEC_BUILD_SCALE || maxsq_dev, "maxsq_dev is NULL");
    B2_REQUIRE(vec_out_dev || slab_hi == slab_lo, "vec_out_dev is NULL");
    B2_REQUIRE(((uintptr_t)vec_out_dev & 15) == 0, "vec_out_dev must be 16-byte aligned");
    B2_REQUIRE(window_dev && (spec_out_dev || n_frames == 0), "NULL argument");
    B2_REQUIRE(frame_lo >= 0 && n_frames >= 0, "bad geometry");
    B2_REQUIRE(out_ld >= 1024 && (out_ld & 3) == 0 && ((uintptr_t)spec_out_dev & 15) == 0,
               "frame-major spectrogram: out_ld >= 1024, a multiple of 4, 16-byte aligned base");
    B2_REQUIRE(slab_lo == frame_lo * 512, "slab_lo must equal frame_lo * 512 (slabs are cut on hop boundaries)");
    B2_REQUIRE(n_frames * 512 <= slab_hi - slab_lo + 511, "frames must start inside the slab");
    WfArgs a;
    if (fill_place_args(ctx, descs, n_pkts, slab_lo, slab_hi, a.place)) return 1;
    a.spans = nullptr; a.n_spans = 0; a.maxsq_out = nullptr; a.plain = 0;
    a.world = 1; a.rank = 0; a.epoch = 0;
    for (int r = 0; r < kWfMaxPeers; ++r) a.peer[r] = nullptr;
    if (mode == kWfAutoNormPeer) {
        B2_REQUIRE(peer_blocks && world >= 1 && world <= kWfMaxPeers && rank >= 0 && rank < world,
                   "bad peer arguments");
        for (int r = 0; r < world; ++r) {
            B2_REQUIRE(peer_blocks[r] && ((uintptr_t)peer_blocks[r] & 15) == 0, "peer block NULL or misaligned");
            a.peer[r] = (unsigned long long *)peer_blocks[r];
        }
        a.world = world; a.rank = rank; a.epoch = epoch;
    }
    DeviceGuard g(ctx->device);
    if (mode == kWfAutoNorm || mode == kWfAutoNormPeer) {
        B2_REQUIRE(n_ranges >= 0 && n_ranges <= kWfMaxNormRanges && (n_ranges == 0 || norm_ranges), "bad norm ranges");
        long long rg[2 * kWfMaxNormRanges];
        int nr = 0;
        for (int r = 0; r < n_ranges; ++r) {
            const int64_t lo = norm_ranges[2 * r], hi = norm_ranges[2 * r + 1];
            B2_REQUIRE(lo >= 0 && hi >= lo, "bad norm range");
            if (hi == lo) continue;
            rg[2 * nr] = lo; rg[2 * nr + 1] = hi; ++nr;
        }
        if (plan_max_spans(ctx, a.place, rg, nr, (cudaStream_t)stream, &a.spans, &a.n_spans)) return 1;
        a.maxsq_out = maxsq_dev;
    }
    a.vec = (float2 *)vec_out_dev;
    a.maxsq = reinterpret_cast<const unsigned long long *>(maxsq_dev);
    a.mode = mode;
    a.window = window_dev; a.scale = scale;
    a.n_frames = n_frames; a.out = spec_out_dev; a.out_ld = out_ld;
    a.flag = nonzero_flag_dev;
    if (get_tw1024(ctx, &a.tw)) return 1;
    return launch_wf1024(ctx, a, (cudaStream_t)stream);
}

int b200vec_build_stft1024_fm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                              int64_t slab_lo, int64_t slab_hi, void *vec_out_dev, int mode,
                              double *maxsq_dev, const float *window_dev, float scale,
                              int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                              int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || mode == B200VEC_BUILD_SCALE, "mode must be STORE or SCALE");
    return build_stft1024_fm_impl(ctx, descs, n_pkts, slab_lo, slab_hi, vec_out_dev, mode, nullptr, 0,
                                  nullptr, 1, 0, 0, maxsq_dev, window_dev, scale, frame_lo, n_frames, spec_out_dev, out_ld, nonzero_flag_dev, stream);
}

int b200vec_build_stft1024_fm_norm(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                                   int64_t slab_lo, int64_t slab_hi, void *vec_out_dev,
                                   const int64_t *norm_ranges, int n_ranges, double *maxsq_dev,
                                   const float *window_dev, float scale,
                                   int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                                   int *nonzero_flag_dev, void *stream) {
    return build_stft1024_fm_impl(ctx, descs, n_pkts, slab_lo, slab_hi, vec_out_dev, kWfAutoNorm, norm_ranges, n_ranges,
                                  nullptr, 1, 0, 0, maxsq_dev, window_dev, scale, frame_lo, n_frames, spec_out_dev,
                                  out_ld, nonzero_flag_dev, stream);
}

int b200vec_build_stft1024_fm_norm_peer(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                                        int64_t slab_lo, int64_t slab_hi, void *vec_out_dev,
                                        const int64_t *norm_ranges, int n_ranges,
                                        void *const *peer_blocks, int world, int rank, uint32_t epoch,
                                        double *maxsq_dev, const float *window_dev, float scale,
                                        int64_t frame_lo, int64_t n_frames, float *spec_out_dev, int64_t out_ld,
                                        int *nonzero_flag_dev, void *stream) {
    return build_stft1024_fm_impl(ctx, descs, n_pkts, slab_lo, slab_hi, vec_out_dev, kWfAutoNormPeer, norm_ranges,
                                  n_ranges, peer_blocks, world, rank, epoch, maxsq_dev, window_dev, scale, frame_lo,
                                  n_frames, spec_out_dev, out_ld, nonzero_flag_dev, stream);
}

int b200vec_stft1024_fm(b200vec_ctx *ctx, const void *in_dev, int64_t in_origin, int64_t in_len,
                        const float *window_dev, float scale, int64_t frame_lo, int64_t n_frames,
                        float *out_dev, int64_t out_ld, int *nonzero_flag_dev, void *stream) {
    B2_REQUIRE(ctx && in_dev && window_dev && (out_dev || n_frames == 0), "NULL argument");
    B2_REQUIRE(in_len >= 0 && frame_lo >= 0 && n_frames >= 0, "bad geometry");
    B2_REQUIRE(out_ld >= 1024 && (out_ld & 3) == 0 && ((uintptr_t)out_dev & 15) == 0,
               "frame-major spectrogram: out_ld >= 1024, a multiple of 4, 16-byte aligned base");
    if (n_frames == 0) return 0;
    const long long first = frame_lo * 512 - in_origin;
    const long long last = (frame_lo + n_frames - 1) * 512 + 1023 - in_origin;
    B2_REQUIRE(first >= 0 && last < in_len, "frames read outside the input buffer");
    WfArgs a;
    memset(&a, 0, sizeof(a));
    a.plain = 1;
    a.vec = const_cast<float2 *>((const float2 *)in_dev) + first;
    a.mode = B200VEC_BUILD_STORE;
    a.window = window_dev; a.scale = scale;
    a.n_frames = n_frames; a.out = out_dev; a.out_ld = out_ld;
    a.flag = nonzero_flag_dev;
    DeviceGuard g(ctx->device);
    if (get_tw1024(ctx, &a.tw)) return 1;
    return launch_wf1024(ctx, a, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
static int db_impl(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val, float *out_dev,
                   const float *values_host, int n_values, float *values_out_host, cudaStream_t st) {
    B2_REQUIRE(ctx && (n == 0 || (s_dev && out_dev)), "NULL argument");
    B2_REQUIRE(n_values >= 0 && n_values <= B200VEC_MAX_RANKS && (n_values == 0 || (values_host && values_out_host)),
               "bad value list");
    if (n == 0 && n_values == 0) return 0;
    DeviceGuard g(ctx->device);
    DbExtra ex;
    memset(&ex, 0, sizeof(ex));
    ex.n = n_values;
    for (int i = 0; i < n_values; ++i) ex.v[i] = values_host[i];
    long long grid = (n / 4 + kSelThreads - 1) / kSelThreads;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    if (grid < 1) grid = 1;
    float *ex_dev = reinterpret_cast<float *>(ctx->hist + kBarrierOff + 4);        // B200VEC_MAX_RANKS floats of scratch
    db_kernel<<<(unsigned)grid, kSelThreads, 0, st>>>(s_dev, n, floor_val, out_dev, ex, ex_dev, nullptr);
    B2_CUDA(cudaGetLastError());
    if (n_values > 0) {
        B2_CUDA(cudaMemcpyAsync(ctx->pinned, ex_dev, sizeof(float) * n_values, cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        memcpy(values_out_host, ctx->pinned, sizeof(float) * n_values);
    }
    return 0;
}

int b200vec_db(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val,
               float *out_dev, void *stream) {
    return db_impl(ctx, s_dev, n, floor_val, out_dev, nullptr, 0, nullptr, (cudaStream_t)stream);
}

int b200vec_db_values(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val, float *out_dev,
                      const float *values_host, int n_values, float *values_out_host, void *stream) {
    return db_impl(ctx, s_dev, n, floor_val, out_dev, values_host, n_values, values_out_host, (cudaStream_t)stream);
}

int b200vec_select_count(b200vec_ctx *ctx, const float *data_dev, int64_t n, int positive_only,
                         int64_t *count_host, void *stream) {
    B2_REQUIRE(ctx && count_host && (n == 0 || data_dev), "NULL argument");
    DeviceGuard g(ctx->device);
    SelState s;
    if (run_select(ctx, data_dev, n, positive_only ? 1 : 0, nullptr, 0, (cudaStream_t)stream, &s)) return 1;
    *count_host = (int64_t)s.count;
    return 0;
}

int b200vec_select_ranks(b200vec_ctx *ctx, const float *data_dev, int64_t n, int positive_only,
                         const int64_t *ranks_host, int n_ranks, float *values_host,
                         void *stream) {
    B2_REQUIRE(ctx && data_dev && ranks_host && values_host, "NULL argument");
    B2_REQUIRE(n_ranks >= 1 && n_ranks <= B200VEC_MAX_RANKS, "n_ranks out of range");
    B2_REQUIRE(n > 0, "empty input");
    DeviceGuard g(ctx->device);
    SelState s;
    if (run_select(ctx, data_dev, n, positive_only ? 1 : 0, ranks_host, n_ranks, (cudaStream_t)stream, &s))
        return 1;
    for (int r = 0; r < n_ranks; ++r) {
        B2_REQUIRE(ranks_host[r] >= 0 && (unsigned long long)ranks_host[r] < s.count,
                   "rank outside the selected element count");
        values_host[r] = key2f(s.result[r]);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// Fast order statistics (select.cuh, "Fast path"): begin = 13-bit top histogram (one read), finish = bin pick on
// the host, compaction of the bins in question (one read), exact select on the candidates.
int b200vec_select_begin(b200vec_ctx *ctx, const float *data_dev, int64_t n, int use_abs,
                         int64_t *count_pos_host, void *stream) {
    B2_REQUIRE(ctx && (n == 0 || data_dev), "NULL argument");
    B2_REQUIRE(n >= 0, "negative length");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    TopState &ts = ctx->top;
    ts.data = nullptr; ts.n = -1;
    const size_t bytes = kTopBins * sizeof(unsigned int) + 2 * sizeof(unsigned long long);
    B2_CUDA(cudaMemsetAsync(ctx->top_hist, 0, bytes, st));
    if (n > 0) {
        long long grid = (n / 4 + kSelThreads - 1) / kSelThreads + 1;
        const long long cap = (long long)ctx->sm_count * 6;         // 6 x 32 KB histograms per SM
        if (grid > cap) grid = cap;
        unsigned long long *pz = reinterpret_cast<unsigned long long *>(ctx->top_hist + kTopBins);
        if (use_abs) sel_top_kernel<true><<<(unsigned)grid, kSelThreads, 0, st>>>(data_dev, n, ctx->top_hist, pz);
        else sel_top_kernel<false><<<(unsigned)grid, kSelThreads, 0, st>>>(data_dev, n, ctx->top_hist, pz);
        B2_CUDA(cudaGetLastError());
    }
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->top_hist, bytes, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    const unsigned int *h = reinterpret_cast<const unsigned int *>(ctx->pinned);
    ts.hist.assign(h, h + kTopBins);
    memcpy(&ts.n_poszero, h + kTopBins, sizeof(unsigned long long));
    ts.data = data_dev; ts.n = n; ts.use_abs = use_abs ? 1 : 0;
    if (count_pos_host) {
        unsigned long long pos = 0;
        for (unsigned b = kTopZeroBin; b < (unsigned)kTopBins; ++b) pos += ts.hist[b];   // +0 is not in the histogram
        *count_pos_host = (int64_t)pos;
    }
    return 0;
}

int b200vec_select_finish(b200vec_ctx *ctx, const float *data_dev, int64_t n, const int64_t *ranks_host,
                          int n_ranks, float *values_host, void *stream) {
    B2_REQUIRE(ctx && data_dev && ranks_host && values_host, "NULL argument");
    B2_REQUIRE(n_ranks >= 1 && n_ranks <= B200VEC_MAX_RANKS, "n_ranks out of range");
    TopState &ts = ctx->top;
    B2_REQUIRE(ts.data == data_dev && ts.n == n && n > 0, "b200vec_select_begin was not called on this array");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    // ascending order of all elements: bins below the zero bin (negative keys), the +0 block, the bins from the zero bin up
    unsigned long long below_zero = 0;
    for (unsigned b = 0; b < kTopZeroBin; ++b) below_zero += ts.hist[b];
    int bin_of[B200VEC_MAX_RANKS];
    unsigned long long resid[B200VEC_MAX_RANKS];
    SelBins sel;
    memset(&sel, 0, sizeof(sel));
    for (int r = 0; r < n_ranks; ++r) {
        B2_REQUIRE(ranks_host[r] >= 0 && ranks_host[r] < n, "rank outside the array");
        unsigned long long want = (unsigned long long)ranks_host[r];
        bin_of[r] = -1;
        unsigned b0 = 0, b1 = kTopZeroBin;
        if (want >= below_zero) {
            if (want < below_zero + ts.n_poszero) { values_host[r] = 0.0f; continue; }     // inside the +0 block
            want -= below_zero + ts.n_poszero;
            b0 = kTopZeroBin; b1 = kTopBins;
        }
        unsigned long long cum = 0;
        for (unsigned b = b0; b < b1; ++b) {
            if (want < cum + ts.hist[b]) { bin_of[r] = (int)b; resid[r] = want - cum; break; }
            cum += ts.hist[b];
        }
        B2_REQUIRE(bin_of[r] >= 0, "histogram does not cover the rank (array changed since b200vec_select_begin?)");
        bool seen = false;
        for (int q = 0; q < sel.n; ++q) seen = seen || sel.bin[q] == (unsigned)bin_of[r];
        if (!seen) sel.bin[sel.n++] = (unsigned)bin_of[r];
    }
    if (sel.n == 0) return 0;
    for (int i = 1; i < sel.n; ++i)                                   // ascending bins = ascending values
        for (int j = i; j > 0 && sel.bin[j - 1] > sel.bin[j]; --j) { const unsigned t = sel.bin[j]; sel.bin[j] = sel.bin[j - 1]; sel.bin[j - 1] = t; }
    unsigned long long pop = 0;
    for (int i = 0; i < sel.n; ++i) pop += ts.hist[sel.bin[i]];
    int64_t sub_ranks[B200VEC_MAX_RANKS];
    int sub_of[B200VEC_MAX_RANKS], n_sub = 0;
    const unsigned long long cap_max = (unsigned long long)(n / 8 > (1 << 20) ? n / 8 : (1 << 20));
    SelState state;
    if (pop <= cap_max) {
        if (ctx->cand.bytes < pop * sizeof(float)) {
            B2_CUDA(cudaStreamSynchronize(st));
            if (ensure_scratch(ctx->cand, (size_t)(pop + pop / 4 + 1024) * sizeof(float))) return 1;
        }
        // The compaction pass also counts the candidates' key bits 18..10 into the level-2 histograms of the exact
        // select, so the select starts there: rank r enters with its 11-bit prefix (bin >> 2) and its position among
        // the candidates that share that prefix (the selected bins below its own + its position inside its bin).
        SelState init;
        memset(&init, 0, sizeof(init));
        for (int r = 0; r < n_ranks; ++r) {
            if (bin_of[r] < 0) continue;
            const unsigned int p11 = (unsigned)bin_of[r] >> 2;
            unsigned long long want = resid[r];
            for (int i = 0; i < sel.n; ++i)
                if ((sel.bin[i] >> 2) == p11 && sel.bin[i] < (unsigned)bin_of[r]) want += ts.hist[sel.bin[i]];
            init.prefix[n_sub] = p11; init.resid[n_sub] = want; sub_of[n_sub++] = r;
        }
        init.n_ranks = n_sub; init.count = pop;
        for (int i = 0; i < sel.n; ++i) {                             // hist2[slot]: the first rank with the bin's prefix owns it
            sel.slot[i] = 0;
            for (int q = 0; q < n_sub; ++q)
                if (init.prefix[q] == (sel.bin[i] >> 2)) { sel.slot[i] = q; break; }
        }
        B2_CUDA(cudaMemsetAsync(ctx->hist, 0, sizeof(unsigned int) * kHistWords, st));
        sel_init<<<1, 1, 0, st>>>(ctx->hist, init);
        unsigned long long *cnt = reinterpret_cast<unsigned long long *>(ctx->top_hist + kTopBins) + 1;
        B2_CUDA(cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), st));
        const unsigned grid = (unsigned)(ctx->sm_count * 4);
        if (ts.use_abs) sel_compact_kernel<true><<<grid, kSelThreads, 0, st>>>(data_dev, n, sel, (float *)ctx->cand.ptr, pop, cnt, ctx->hist, nullptr);
        else sel_compact_kernel<false><<<grid, kSelThreads, 0, st>>>(data_dev, n, sel, (float *)ctx->cand.ptr, pop, cnt, ctx->hist, nullptr);
        B2_CUDA(cudaGetLastError());
        long long blocks = ((long long)pop / 4 + kPickThreads * kSelUnroll - 1) / (kPickThreads * kSelUnroll);
        if (blocks > ctx->sm_count) blocks = ctx->sm_count;
        if (blocks < 1) blocks = 1;
        const float *cand = (const float *)ctx->cand.ptr;
        long long pop_ll = (long long)pop;
        const K5Plan *no_plan = nullptr;
        void *kargs[] = {(void *)&cand, (void *)&pop_ll, (void *)&ctx->hist, (void *)&no_plan};
        B2_CUDA(cudaLaunchCooperativeKernel((const void *)sel_cand_kernel, dim3((unsigned)blocks), dim3(kPickThreads),
                                            kargs, (size_t)n_sub * kBins3 * sizeof(unsigned int), st));
        B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->hist + kStateOff, sizeof(SelState), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync((char *)ctx->pinned + 1024, cnt, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        memcpy(&state, ctx->pinned, sizeof(SelState));
        unsigned long long got = 0;
        memcpy(&got, (char *)ctx->pinned + 1024, sizeof(got));
        B2_REQUIRE(got == pop, "candidate count differs from the histogram (array changed since b200vec_select_begin?)");
    } else {                                                          // a heavily repeated value: 3-level select on the whole array
        for (int r = 0; r < n_ranks; ++r) {
            if (bin_of[r] < 0) continue;
            sub_ranks[n_sub] = ranks_host[r]; sub_of[n_sub++] = r;
        }
        if (run_select(ctx, data_dev, n, ts.use_abs ? 2 : 0, sub_ranks, n_sub, st, &state)) return 1;
    }
    for (int i = 0; i < n_sub; ++i) values_host[sub_of[i]] = key2f(state.result[i]);
    return 0;
}

int b200vec_normalize_db(b200vec_ctx *ctx, const float *s_dev, int64_t n, double low_percentile, double high_percentile,
                         float *db_out_dev, float *values_host, int64_t *n_pos_host, void *stream) {
    B2_REQUIRE(ctx && s_dev && db_out_dev && values_host && n >= 1, "bad arguments");
    {
        // Fast path: the whole chain is queued at once — histogram, plan (positives, numpy's rank arithmetic, bin
        // picks: k5_plan_kernel), compaction, candidate select, floor, dB map — and ONE small record comes back.
        DeviceGuard g(ctx->device);
        cudaStream_t st = (cudaStream_t)stream;
        const unsigned long long cap = (unsigned long long)(n / 8 > (1 << 20) ? n / 8 : (1 << 20));
        if (ctx->cand.bytes < cap * sizeof(float)) {
            B2_CUDA(cudaStreamSynchronize(st));
            if (ensure_scratch(ctx->cand, (size_t)cap * sizeof(float))) return 1;
        }
        if (ensure_scratch(ctx->k5, sizeof(K5Plan) + sizeof(K5Out) + 64)) return 1;
        K5Plan *plan = reinterpret_cast<K5Plan *>(ctx->k5.ptr);
        K5Out *out = reinterpret_cast<K5Out *>((char *)ctx->k5.ptr + ((sizeof(K5Plan) + 15) & ~(size_t)15));
        const size_t top_bytes = kTopBins * sizeof(unsigned int) + 2 * sizeof(unsigned long long);
        unsigned long long *pz = reinterpret_cast<unsigned long long *>(ctx->top_hist + kTopBins);
        unsigned long long *cnt = pz + 1;
        ctx->top.data = nullptr; ctx->top.n = -1;                        // no begin/finish pair is open
        B2_CUDA(cudaMemsetAsync(ctx->top_hist, 0, top_bytes, st));
        B2_CUDA(cudaMemsetAsync(ctx->hist, 0, sizeof(unsigned int) * kHistWords, st));
        long long grid = (n / 4 + kSelThreads - 1) / kSelThreads + 1;
        if (grid > (long long)ctx->sm_count * 6) grid = (long long)ctx->sm_count * 6;
        sel_top_kernel<true><<<(unsigned)grid, kSelThreads, 0, st>>>(s_dev, n, ctx->top_hist, pz);
        k5_plan_kernel<<<1, 1024, 0, st>>>(ctx->top_hist, pz, n, low_percentile, high_percentile, cap, ctx->hist, plan);
        SelBins none;
        memset(&none, 0, sizeof(none));
        sel_compact_kernel<true><<<(unsigned)(ctx->sm_count * 4), kSelThreads, 0, st>>>(s_dev, n, none, (float *)ctx->cand.ptr,
                                                                                        cap, cnt, ctx->hist, plan);
        B2_CUDA(cudaGetLastError());
        const float *cand = (const float *)ctx->cand.ptr;
        long long n_unused = 0;
        const K5Plan *plan_c = plan;
        void *kargs[] = {(void *)&cand, (void *)&n_unused, (void *)&ctx->hist, (void *)&plan_c};
        B2_CUDA(cudaLaunchCooperativeKernel((const void *)sel_cand_kernel, dim3((unsigned)ctx->sm_count), dim3(kPickThreads),
                                            kargs, (size_t)B200VEC_MAX_RANKS * kBins3 * sizeof(unsigned int), st));
        k5_floor_kernel<<<1, 1, 0, st>>>(ctx->hist, plan, out);
        long long dgrid = (n / 4 + kSelThreads - 1) / kSelThreads;
        if (dgrid > (long long)ctx->sm_count * 16) dgrid = (long long)ctx->sm_count * 16;
        if (dgrid < 1) dgrid = 1;
        DbExtra ex;
        memset(&ex, 0, sizeof(ex));
        db_kernel<<<(unsigned)dgrid, kSelThreads, 0, st>>>(s_dev, n, 0.0f, db_out_dev, ex, nullptr, out);
        B2_CUDA(cudaGetLastError());
        B2_CUDA(cudaMemcpyAsync(ctx->pinned, out, sizeof(K5Out), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync((char *)ctx->pinned + 1024, cnt, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaMemcpyAsync((char *)ctx->pinned + 1040, &plan->pop, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
        B2_CUDA(cudaStreamSynchronize(st));
        K5Out h;
        memcpy(&h, ctx->pinned, sizeof(h));
        if (!h.overflow) {
            unsigned long long got = 0, pop = 0;
            memcpy(&got, (char *)ctx->pinned + 1024, sizeof(got));
            memcpy(&pop, (char *)ctx->pinned + 1040, sizeof(pop));
            B2_REQUIRE(got == pop, "candidate count differs from the histogram (array changed during the call?)");
            values_host[0] = h.floor_v;
            for (int i = 0; i < 6; ++i) values_host[1 + i] = h.six[i];
            values_host[7] = h.gamma[0]; values_host[8] = h.gamma[1];
            if (n_pos_host) *n_pos_host = (int64_t)h.n_pos;
            return 0;
        }
        // a heavily repeated non-zero value: more candidates than the buffer holds -> the host-driven path below
    }
    int64_t n_pos = 0;
    if (b200vec_select_begin(ctx, s_dev, n, 1, &n_pos, stream)) return 1;
    // numpy's float32 index arithmetic (npquant.h): dB is monotone in |S|, so the low / high percentiles, minimum
    // and maximum of the dB image and the 5th percentile of the positive elements are eight ranks of the sorted |S|
    const PctRanks r1 = pct_ranks_f32(n, low_percentile), r2 = pct_ranks_f32(n, high_percentile);
    PctRanks r0 = {0, 0, 0.0f};
    int64_t ranks[8] = {r1.lo, r1.hi, r2.lo, r2.hi, 0, n - 1, 0, 0};
    int nr = 6;
    if (n_pos > 0) {
        r0 = pct_ranks_f32(n_pos, 5.0);
        ranks[6] = n - n_pos + r0.lo; ranks[7] = n - n_pos + r0.hi;      // the positive elements are the top n_pos of |S|
        nr = 8;
    }
    float v[8] = {0};
    if (b200vec_select_finish(ctx, s_dev, n, ranks, nr, v, stream)) return 1;
    float floor_v = n_pos > 0 ? lerp_f32(v[6], v[7], r0.gamma) : (float)1e-12;
    if ((float)1e-12 > floor_v) floor_v = (float)1e-12;                   // max(floor, 1e-12), utils.py:366
    float six[6];
    if (b200vec_db_values(ctx, s_dev, n, floor_v, db_out_dev, v, 6, six, stream)) return 1;
    values_host[0] = floor_v;
    for (int i = 0; i < 6; ++i) values_host[1 + i] = six[i];
    values_host[7] = r1.gamma; values_host[8] = r2.gamma;
    if (n_pos_host) *n_pos_host = n_pos;
    return 0;
}

int b200vec_decimate(b200vec_ctx *ctx, const void *in_dev, int64_t n_in, int64_t stride, void *out_dev, void *stream) {
    B2_REQUIRE(ctx && n_in >= 0 && stride >= 1 && (n_in == 0 || (in_dev && out_dev)), "bad arguments");
    const long long n_out = (n_in + stride - 1) / stride;
    if (n_out == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n_out + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    decimate_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>((const float2 *)in_dev, n_out, stride, (float2 *)out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_peer_status(b200vec_ctx *ctx, int *status_host, void *stream) {
    B2_REQUIRE(ctx && status_host, "NULL argument");
    *status_host = 0;
    if (!ctx->wf_ctr) return 0;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->wf_ctr + 10, sizeof(unsigned int), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    *status_host = (int)*reinterpret_cast<unsigned int *>(ctx->pinned);
    if (*status_host) B2_CUDA(cudaMemsetAsync(ctx->wf_ctr + 10, 0, sizeof(unsigned int), st));   // reading clears it
    return 0;
}

// ---------------------------------------------------------------------------------------
int b200vec_wv_quantize(b200vec_ctx *ctx, const void *in_dev, int64_t n, float pre_scale, int16_t *out_dev,
                        void *stream) {
    B2_REQUIRE(ctx && (n == 0 || (in_dev && out_dev)), "NULL argument");
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    wv_quantize_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(
        (const float2 *)in_dev, n, pre_scale, (short2 *)out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_power_stats(b200vec_ctx *ctx, const void *in_dev, int64_t n, float pre_scale, double *stats_host,
                        void *stream) {
    B2_REQUIRE(ctx && stats_host && n >= 0 && (n == 0 || in_dev), "NULL argument");
    stats_host[0] = stats_host[1] = 0.0;
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long *dev = reinterpret_cast<unsigned long long *>(ctx->hist);     // select scratch doubles as the 2-slot result
    B2_CUDA(cudaMemsetAsync(dev, 0, 2 * sizeof(unsigned long long), st));
    long long grid = (n + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    power_stats_kernel<<<(unsigned)grid, 256, 0, st>>>((const float2 *)in_dev, n, pre_scale, dev);
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, dev, 2 * sizeof(double), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    memcpy(stats_host, ctx->pinned, 2 * sizeof(double));
    return 0;
}

#if defined(B2_AB) && defined(B2_PIPE_TIMING)
int b200vec_debug_pipe_prof(long long *out_host) {   // timing build only (scripts/pipe_timing.py)
    cudaDeviceSynchronize();
    return cudaMemcpyFromSymbol(out_host, b200vec::g_pipe_prof, sizeof(long long) * 320 * 16 * 8) != cudaSuccess;
}
int b200vec_debug_pipe_cta(long long *out_host) {
    cudaDeviceSynchronize();
    return cudaMemcpyFromSymbol(out_host, b200vec::g_pipe_cta, sizeof(long long) * 320 * 4) != cudaSuccess;
}
#endif

int b200vec_split_planes(b200vec_ctx *ctx, const void *in_dev, int64_t n, float *re_out_dev, float *im_out_dev,
                         void *stream) {
    B2_REQUIRE(ctx && n >= 0 && (n == 0 || (in_dev && re_out_dev && im_out_dev)), "NULL argument");
    B2_REQUIRE((((uintptr_t)in_dev | (uintptr_t)re_out_dev | (uintptr_t)im_out_dev) & 15) == 0,
               "buffers must be 16-byte aligned");
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n / 4 + 255) / 256 + 1;
    const long long cap = (long long)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    split_planes_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>((const float2 *)in_dev, n, re_out_dev, im_out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_box_energy(b200vec_ctx *ctx, const void *in_dev, int64_t n, int window, double *smoothed_out_dev,
                       double *max_out_dev, void *stream) {
    B2_REQUIRE(ctx && n >= 0 && (n == 0 || (in_dev && smoothed_out_dev)) && max_out_dev, "NULL argument");
    B2_REQUIRE(window >= 1 && window <= kBoxMaxWindow, "window must be in [1, 4096]");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    B2_CUDA(cudaMemsetAsync(max_out_dev, 0, sizeof(double), st));
    if (n == 0) return 0;
    long long grid = (n + kBoxThreads - 1) / kBoxThreads;
    const long long cap = (long long)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    box_energy_kernel<<<(unsigned)grid, kBoxThreads, 0, st>>>((const float2 *)in_dev, n, window, smoothed_out_dev,
                                                              reinterpret_cast<unsigned long long *>(max_out_dev));
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_select_f64(b200vec_ctx *ctx, const double *data_dev, int64_t n, const int64_t *ranks_host, int n_ranks,
                       double *values_host, void *stream) {
    B2_REQUIRE(ctx && data_dev && ranks_host && values_host, "NULL argument");
    B2_REQUIRE(n >= 1 && n_ranks >= 1 && n_ranks <= kSel64MaxRanks, "n >= 1 and 1 <= n_ranks <= 4");
    for (int r = 0; r < n_ranks; ++r) B2_REQUIRE(ranks_host[r] >= 0 && ranks_host[r] < n, "rank outside the array");
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    const size_t bytes = 256 + 2 * kSel64MaxRanks * 2048 * sizeof(unsigned int);
    if (ensure_scratch(ctx->sel64, bytes)) return 1;
    Sel64State init;
    memset(&init, 0, sizeof(init));
    init.n_ranks = n_ranks;
    for (int r = 0; r < n_ranks; ++r) init.resid[r] = (unsigned long long)ranks_host[r];
    static_assert(sizeof(Sel64State) <= 256, "state block");
    B2_CUDA(cudaMemsetAsync(ctx->sel64.ptr, 0, bytes, st));
    memcpy(ctx->pinned, &init, sizeof(init));
    B2_CUDA(cudaMemcpyAsync(ctx->sel64.ptr, ctx->pinned, sizeof(init), cudaMemcpyHostToDevice, st));
    long long blocks = (n + kSel64Threads * 8 - 1) / (kSel64Threads * 8);
    if (blocks > ctx->sm_count) blocks = ctx->sm_count;
    void *scratch = ctx->sel64.ptr;
    void *kargs[] = {(void *)&data_dev, (void *)&n, (void *)&scratch};
    B2_CUDA(cudaLaunchCooperativeKernel((const void *)sel64_kernel, dim3((unsigned)blocks), dim3(kSel64Threads), kargs, 0, st));
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->sel64.ptr, sizeof(init), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    memcpy(&init, ctx->pinned, sizeof(init));
    for (int r = 0; r < n_ranks; ++r) {
        const unsigned long long k = init.result[r];
        const unsigned long long u = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
        memcpy(&values_host[r], &u, sizeof(double));
    }
    return 0;
}

int b200vec_threshold_bounds(b200vec_ctx *ctx, const double *smoothed_dev, int64_t n, double threshold,
                             int64_t *first_last_host, void *stream) {
    B2_REQUIRE(ctx && first_last_host && n >= 0 && (n == 0 || smoothed_dev), "NULL argument");
    first_last_host[0] = n; first_last_host[1] = -1;
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    cudaStream_t st = (cudaStream_t)stream;
    long long *dev = reinterpret_cast<long long *>(ctx->hist);          // select scratch doubles as the 2-slot result
    long long *pin = reinterpret_cast<long long *>(ctx->pinned);
    pin[0] = n; pin[1] = -1;
    B2_CUDA(cudaMemcpyAsync(dev, pin, 2 * sizeof(long long), cudaMemcpyHostToDevice, st));
    long long grid = (n + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 8;
    if (grid > cap) grid = cap;
    threshold_bounds_kernel<<<(unsigned)grid, 256, 0, st>>>(smoothed_dev, n, threshold, dev);
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaMemcpyAsync(pin, dev, 2 * sizeof(long long), cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    first_last_host[0] = pin[0]; first_last_host[1] = pin[1];
    return 0;
}

int b200vec_median2_rows(b200vec_ctx *ctx, const float *in_dev, int64_t rows, int64_t cols, int64_t in_row_stride,
                         int64_t in_col_stride, float *out_dev, int64_t out_row_stride, int64_t out_col_stride,
                         void *stream) {
    B2_REQUIRE(ctx && rows >= 0 && cols >= 0, "bad geometry");
    if (rows == 0 || cols == 0) return 0;
    B2_REQUIRE(in_dev && out_dev && in_dev != out_dev, "NULL argument or in-place call");
    DeviceGuard g(ctx->device);
    long long grid = (rows * cols + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    const int cols_fast = llabs(in_col_stride) <= llabs(in_row_stride);   // walk the contiguous direction with threadIdx
    median2_rows_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(in_dev, rows, cols, in_row_stride, in_col_stride,
                                                                          out_dev, out_row_stride, out_col_stride, cols_fast);
    B2_CUDA(cudaGetLastError());
    return 0;
}

// The level metrics of b200vec_power_stats without a context and without any allocation: the scratch is the
// file-scope device array above (set-up per device on the first call: cudaGetSymbolAddress with that device
// current), the result goes to the caller's device buffer and nothing synchronises.  No counterpart in the
// reference.
int b200vec_power_stats_static(int device, const void *in_dev, int64_t n, float pre_scale, double *stats_out_dev,
                               void *stream) {
    B2_REQUIRE(stats_out_dev && n >= 0 && (n == 0 || in_dev), "NULL argument");
    B2_REQUIRE(device >= 0 && device < kStaticMaxDevices, "device ordinal out of range (0..63)");
    if (n > kStaticMaxSamples)
        return fail("b200vec_power_stats_static",
                    "n exceeds the static scratch: at most 2^30 (1073741824) samples per call; use b200vec_power_stats");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count <= 0)
        return fail("b200vec_power_stats_static", "no CUDA device available (this library has no CPU fallback)");
    B2_REQUIRE(device < count, "device ordinal out of range");
    DeviceGuard g(device);
    StaticDevice *d = nullptr;
    if (int rc = static_device_setup(device, &d)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) {
        B2_CUDA(cudaMemsetAsync(stats_out_dev, 0, 2 * sizeof(double), st));
        return 0;
    }
    B2_CUDA(cudaMemsetAsync(&d->scratch->ticket, 0, sizeof(unsigned int), st));     // the array keeps the last call's contents
    const int n_parts = (int)((n + kStaticChunk - 1) / kStaticChunk);
    const int cap = d->sm_count * 8;
    const int grid = n_parts < cap ? n_parts : cap;
    power_stats_static_kernel<<<(unsigned)grid, 256, 0, st>>>((const float2 *)in_dev, n, n_parts, pre_scale,
                                                               d->scratch, stats_out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"
