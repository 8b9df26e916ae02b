// b200vec.cu — C-ABI of libb200vec.so (see include/b200vec.h for the contract and the
// reference lines each entry point replaces).  sm_100a only; no CPU fallback.
#include "common.cuh"
#include "bank.cuh"
#include "place.cuh"
#include "stft.cuh"
#include "stft1024.cuh"
#include "fused1024.cuh"
#include "pipe1024.cuh"
#include "wf1024.cuh"
#include "packet.cuh"
#include "select.cuh"

namespace b200vec {
thread_local std::string g_last_error;

// K6: mat2wv quantiser (vector_analyzer/mat_to_wv_converter.py:42-50): x*32767 in float32,
// truncation toward zero, int16 I/Q interleave.  Out-of-range values wrap modulo 2^16 like
// the x86 cvttss2si + narrowing the reference's astype(int16) compiles to.
__global__ void __launch_bounds__(256)
wv_quantize_kernel(const float2 *__restrict__ in, long long n, short2 *__restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float2 v = __ldg(in + i);
        const int re = __float2int_rz(v.x * 32767.0f);
        const int im = __float2int_rz(v.y * 32767.0f);
        out[i] = make_short2((short)re, (short)im);
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

static int run_select(b200vec_ctx *ctx, const float *data, long long n, int positive_only,
                      const int64_t *ranks, int nr, cudaStream_t st, SelState *host_state) {
    B2_CUDA(cudaMemsetAsync(ctx->hist, 0, sizeof(unsigned int) * kHistWords, st));
    SelState init;
    memset(&init, 0, sizeof(init));
    init.n_ranks = nr;
    for (int r = 0; r < nr; ++r) init.resid[r] = (unsigned long long)ranks[r];
    sel_init<<<1, 1, 0, st>>>(ctx->hist, init);
    const unsigned grid = (unsigned)(ctx->sm_count * 8);
    if (n > 0) sel_hist1<<<grid, kSelThreads, 0, st>>>(data, n, positive_only, ctx->hist);
    sel_pick<1><<<1, kPickThreads, 0, st>>>(ctx->hist);
    if (nr > 0 && n > 0) {
        sel_histN<2><<<grid, kSelThreads, (size_t)nr * kBins2 * 4, st>>>(data, n, positive_only, ctx->hist);
        sel_pick<2><<<1, kPickThreads, 0, st>>>(ctx->hist);
        sel_histN<3><<<grid, kSelThreads, (size_t)nr * kBins3 * 4, st>>>(data, n, positive_only, ctx->hist);
        sel_pick<3><<<1, kPickThreads, 0, st>>>(ctx->hist);
    }
    B2_CUDA(cudaGetLastError());
    B2_CUDA(cudaMemcpyAsync(ctx->pinned, ctx->hist + kStateOff, sizeof(SelState),
                            cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaStreamSynchronize(st));
    memcpy(host_state, ctx->pinned, sizeof(SelState));
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
    B2_CUDA(cudaMallocHost(&ctx->pinned, 4096));
    B2_CUDA(cudaFuncSetAttribute(sel_histN<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 B200VEC_MAX_RANKS * kBins2 * 4));
    B2_CUDA(cudaFuncSetAttribute(sel_histN<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 B200VEC_MAX_RANKS * kBins3 * 4));
    B2_CUDA(cudaFuncSetAttribute(stft_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 200 * 1024));
    if (stft1024_setup() || fused1024_setup() || pipe1024_setup() || wf1024_setup()) { delete ctx; return 1; }
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
int b200vec_build(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                  int64_t slab_lo, int64_t slab_hi, void *out_dev, int mode,
                  double *maxsq_dev, void *stream) {
    B2_REQUIRE(ctx, "ctx is NULL");
    B2_REQUIRE(n_pkts >= 0 && n_pkts <= B200VEC_MAX_PACKETS, "too many packets");
    B2_REQUIRE(n_pkts == 0 || descs, "descs is NULL");
    B2_REQUIRE(slab_lo >= 0 && slab_hi >= slab_lo, "bad slab");
    B2_REQUIRE(mode == B200VEC_BUILD_MAXONLY || out_dev || slab_hi == slab_lo, "out_dev is NULL");
    B2_REQUIRE(mode == B200VEC_BUILD_STORE || maxsq_dev, "maxsq_dev is NULL");
    B2_REQUIRE(((uintptr_t)out_dev & 15) == 0, "out_dev must be 16-byte aligned");
    PlaceArgs args;
    memset(&args, 0, sizeof(args));
    args.lo = slab_lo; args.hi = slab_hi;
    int np = 0;
    for (int i = 0; i < n_pkts; ++i) {
        const b200vec_pkt_desc &d = descs[i];
        B2_REQUIRE(d.slot >= 0 && d.slot < B200VEC_MAX_SLOTS, "descriptor slot out of range");
        const BankSlot &sl = ctx->slots[d.slot];
        B2_REQUIRE(sl.len > 0, "descriptor names an uncommitted bank slot");
        B2_REQUIRE(d.period > 0, "period must be positive");
        B2_REQUIRE(d.start >= 0 && d.count >= 0, "negative start/count");
        if (d.count == 0) continue;
        PlaceDesc &p = args.d[np++];
        p.start = d.start; p.period = d.period; p.count = d.count;
        p.len = (int)sl.len; p.a = sl.a; p.b = sl.b;
        p.free_inst = d.reserved & 1; p.pmax = ctx->slot_max + d.slot;
    }
    args.np = np;
    DeviceGuard g(ctx->device);
    return launch_place(ctx, args, (float2 *)out_dev, maxsq_dev, mode, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------
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

// ---------------------------------------------------------------------------------------
static int fill_place_args(b200vec_ctx *ctx, const b200vec_pkt_desc *descs, int n_pkts,
                           int64_t slab_lo, int64_t slab_hi, PlaceArgs &args, int *slot_of = nullptr) {
    memset(&args, 0, sizeof(args));
    args.lo = slab_lo; args.hi = slab_hi;
    int np = 0;
    for (int i = 0; i < n_pkts; ++i) {
        const b200vec_pkt_desc &d = descs[i];
        B2_REQUIRE(d.slot >= 0 && d.slot < B200VEC_MAX_SLOTS, "descriptor slot out of range");
        const BankSlot &sl = ctx->slots[d.slot];
        B2_REQUIRE(sl.len > 0, "descriptor names an uncommitted bank slot");
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
    if (fill_place_args(ctx, descs, n_pkts, slab_lo, slab_hi, a.place, slot_of)) return 1;
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
    if (pipe1024_enabled()) return launch_pipe1024<true>(ctx, a, n_frames, slab_hi - slab_lo, (cudaStream_t)stream);
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
    a.n_norm = 0; a.norm_units = 0; a.maxsq_out = nullptr; a.plain = 0;
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
    if (mode == kWfAutoNorm || mode == kWfAutoNormPeer) {
        B2_REQUIRE(n_ranges >= 0 && n_ranges <= kWfMaxNormRanges && (n_ranges == 0 || norm_ranges), "bad norm ranges");
        for (int r = 0; r < n_ranges; ++r) {
            const int64_t lo = norm_ranges[2 * r], hi = norm_ranges[2 * r + 1];
            B2_REQUIRE(lo >= 0 && hi >= lo, "bad norm range");
            if (hi == lo) continue;
            a.norm_lo[a.n_norm] = lo; a.norm_hi[a.n_norm] = hi; ++a.n_norm;
            a.norm_units += (hi - lo + 1023) / 1024;
        }
        a.maxsq_out = maxsq_dev;
    }
    DeviceGuard g(ctx->device);
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
int b200vec_db(b200vec_ctx *ctx, const float *s_dev, int64_t n, float floor_val,
               float *out_dev, void *stream) {
    B2_REQUIRE(ctx && (n == 0 || (s_dev && out_dev)), "NULL argument");
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n + kSelThreads - 1) / kSelThreads;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    db_kernel<<<(unsigned)grid, kSelThreads, 0, (cudaStream_t)stream>>>(s_dev, n, floor_val, out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

int b200vec_select_count(b200vec_ctx *ctx, const float *data_dev, int64_t n, int positive_only,
                         int64_t *count_host, void *stream) {
    B2_REQUIRE(ctx && count_host && (n == 0 || data_dev), "NULL argument");
    DeviceGuard g(ctx->device);
    SelState s;
    if (run_select(ctx, data_dev, n, positive_only, nullptr, 0, (cudaStream_t)stream, &s)) return 1;
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
    if (run_select(ctx, data_dev, n, positive_only, ranks_host, n_ranks, (cudaStream_t)stream, &s))
        return 1;
    for (int r = 0; r < n_ranks; ++r) {
        B2_REQUIRE(ranks_host[r] >= 0 && (unsigned long long)ranks_host[r] < s.count,
                   "rank outside the selected element count");
        values_host[r] = key2f(s.result[r]);
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
int b200vec_wv_quantize(b200vec_ctx *ctx, const void *in_dev, int64_t n, int16_t *out_dev,
                        void *stream) {
    B2_REQUIRE(ctx && (n == 0 || (in_dev && out_dev)), "NULL argument");
    if (n == 0) return 0;
    DeviceGuard g(ctx->device);
    long long grid = (n + 255) / 256;
    const long long cap = (long long)ctx->sm_count * 16;
    if (grid > cap) grid = cap;
    wv_quantize_kernel<<<(unsigned)grid, 256, 0, (cudaStream_t)stream>>>(
        (const float2 *)in_dev, n, (short2 *)out_dev);
    B2_CUDA(cudaGetLastError());
    return 0;
}

#ifdef B2_PIPE_TIMING
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
    B2_REQUIRE(window >= 1 && window <= kBoxMaxWindow, "window must be in [1, 1024]");
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

}  // extern "C"
