// place.cuh — placement descriptors shared by every builder kernel (span.cuh, wf1024.cuh, fused1024.cuh),
// and (A/B builds only) the round-1 block-per-tile placement kernel.
//
// Replaces the reference's np.zeros + `vector[pos:pos+L] += y` loop + normalisation
// (unified_gui.py:1712,1757-1769,1778-1782; main.py:218,253-257,274-277).
//
// Each thread block owns tiles of kTile consecutive output samples.  For every packet it
// derives the instance range that overlaps the tile (two floor-divisions, done by one
// thread per packet), then every thread sums the overlapping (packet, instance) samples of
// its own output samples in registers — packet ascending, instance ascending, the same
// order as the reference's in-place float32 adds — and writes each sample exactly once
// with 16-byte streaming stores.  No atomics on the vector, no read-modify-write.
//
// HBM traffic: 8 B per output sample (store only).  Packet samples come from the bank
// (<= a few MB, L2-resident) with 16-byte loads (two phase-shifted copies, common.cuh).
#pragma once
#include "common.cuh"

namespace b200vec {

constexpr int kTile = 4096;           // output samples per tile (32 KB of complex64)
constexpr int kPlaceThreads = 256;
constexpr int kVec = kTile / (2 * kPlaceThreads);   // float4 (= 2 samples) per thread = 8

struct PlaceDesc {
    long long start, period, count;
    int len;
    int free_inst;                    // the caller vouches that every sample of this packet appears unsummed somewhere
                                      // in the vector (b200vec_pkt_desc::reserved bit 0): max |y|^2 over the whole
                                      // packet is then attained by the vector, and the max pass skips every sample
                                      // lying under exactly one instance of it
    const unsigned long long *pmax;   // bits of max |y|^2 (double) over the committed packet (bank, device memory)
    const float2 *a;                  // y[0], 16-byte aligned copy
    const float2 *b;                  // y[0], 8-mod-16 copy
};

struct PlaceArgs {
    PlaceDesc d[B200VEC_MAX_PACKETS];
    int np;
    long long lo, hi;                 // slab [lo, hi) in vector-absolute samples
};

__device__ __forceinline__ long long floor_div(long long a, long long b) {   // b > 0
    long long q = a / b;
    return (a % b != 0 && a < 0) ? q - 1 : q;
}

__device__ __forceinline__ double warp_max(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

#ifdef B2_AB
// The round-1 block-per-tile placement kernel (4096-sample tiles, 16-byte loads from the two bank copies).
// The product path is the warp-per-span kernel of span.cuh; this one stays in -DB2_AB builds as the
// bit-exact cross-check (B200VEC_PLACE=1).
// MODE 0: store; 1: max only; 2: scale by 1/sqrt(*maxsq) and store.
template <int MODE>
__global__ void __launch_bounds__(kPlaceThreads)
place_kernel(const __grid_constant__ PlaceArgs args, float2 *__restrict__ out,
             unsigned long long *__restrict__ maxsq, long long n_tiles) {
    __shared__ long long s_klo[B200VEC_MAX_PACKETS];
    __shared__ long long s_khi[B200VEC_MAX_PACKETS];
    __shared__ double s_red[kPlaceThreads / 32];

    const int tid = threadIdx.x;
    float scl = 1.0f;
    if (MODE == 2) {
        double m2 = __longlong_as_double((long long)*maxsq);
        if (m2 > 0.0) scl = 1.0f / (float)sqrt(m2);     // IEEE division (no fast-math)
    }
    double tmax = 0.0;       // running max of |v|^2 (exact products in fp64)
    float fthr = 0.0f;       // fp32 pre-filter threshold

    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long t0 = args.lo + tile * kTile;
        const long long t1 = (t0 + kTile < args.hi) ? t0 + kTile : args.hi;
        const int nvalid = (int)(t1 - t0);
        __syncthreads();
        if (tid < args.np) {
            const PlaceDesc &d = args.d[tid];
            // instance k overlaps [t0,t1) iff start+k*period < t1 and start+k*period+len > t0
            long long klo = floor_div(t0 - d.len - d.start, d.period) + 1;
            long long khi = floor_div(t1 - 1 - d.start, d.period);
            if (klo < 0) klo = 0;
            if (khi > d.count - 1) khi = d.count - 1;
            s_klo[tid] = klo;
            s_khi[tid] = khi;
        }
        __syncthreads();

        float4 acc[kVec];
#pragma unroll
        for (int j = 0; j < kVec; ++j) acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        bool any = false;

        for (int p = 0; p < args.np; ++p) {
            const long long klo = s_klo[p], khi = s_khi[p];
            if (klo > khi) continue;
            any = true;
            const PlaceDesc &d = args.d[p];
            const int len = d.len;
            for (long long k = klo; k <= khi; ++k) {
                const int rel = (int)(t0 - (d.start + k * d.period));   // packet index of tile sample 0
                const float2 *base = ((rel & 1) ? d.b : d.a) + rel;
                if (rel >= 0 && rel + kTile <= len) {                   // tile fully inside the instance
#pragma unroll
                    for (int j = 0; j < kVec; ++j) {
                        const int s = 2 * (tid + kPlaceThreads * j);
                        const float4 v = __ldg(reinterpret_cast<const float4 *>(base + s));
                        acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < kVec; ++j) {
                        const int s = 2 * (tid + kPlaceThreads * j);
                        const int n = s + rel;
                        if (n >= -1 && n < len) {                       // zero pads cover n=-1 and n+1=len
                            const float4 v = __ldg(reinterpret_cast<const float4 *>(base + s));
                            acc[j].x += v.x; acc[j].y += v.y; acc[j].z += v.z; acc[j].w += v.w;
                        }
                    }
                }
            }
        }

        if (MODE == 1) {
            if (!any) continue;                                         // block-uniform
#pragma unroll
            for (int j = 0; j < kVec; ++j) {
                const int s = 2 * (tid + kPlaceThreads * j);
                const float q0 = acc[j].x * acc[j].x + acc[j].y * acc[j].y;
                const float q1 = acc[j].z * acc[j].z + acc[j].w * acc[j].w;
                if (s < nvalid && q0 >= fthr) {                         // samples past the slab end never enter the max
                    const double e = (double)acc[j].x * (double)acc[j].x + (double)acc[j].y * (double)acc[j].y;
                    if (e > tmax) { tmax = e; fthr = q0 * 0.99999f; }
                }
                if (s + 1 < nvalid && q1 >= fthr) {
                    const double e = (double)acc[j].z * (double)acc[j].z + (double)acc[j].w * (double)acc[j].w;
                    if (e > tmax) { tmax = e; fthr = q1 * 0.99999f; }
                }
            }
        } else {
            float2 *dst = out + (t0 - args.lo);
#pragma unroll
            for (int j = 0; j < kVec; ++j) {
                const int s = 2 * (tid + kPlaceThreads * j);
                float4 v = acc[j];
                if (MODE == 2) { v.x *= scl; v.y *= scl; v.z *= scl; v.w *= scl; }
                if (s + 1 < nvalid) {
                    __stcs(reinterpret_cast<float4 *>(dst + s), v);
                } else if (s < nvalid) {
                    __stcs(dst + s, make_float2(v.x, v.y));
                }
            }
        }
    }

    if (MODE == 1) {
        tmax = warp_max(tmax);
        if ((tid & 31) == 0) s_red[tid >> 5] = tmax;
        __syncthreads();
        if (tid < 32) {
            double v = (tid < kPlaceThreads / 32) ? s_red[tid] : 0.0;
            v = warp_max(v);
            if (tid == 0 && v > 0.0)
                atomicMax(maxsq, (unsigned long long)__double_as_longlong(v));  // non-negative doubles order like their bits
        }
    }
}

inline int launch_place(b200vec_ctx *ctx, const PlaceArgs &args, float2 *out, double *maxsq,
                        int mode, cudaStream_t st) {
    const long long n = args.hi - args.lo;
    if (n <= 0) return 0;
    const long long n_tiles = (n + kTile - 1) / kTile;
    static int occ[3] = {0, 0, 0};         // resident CTAs per SM, per mode (register-limited)
    if (!occ[0]) {
        B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], place_kernel<0>, kPlaceThreads, 0));
        B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], place_kernel<1>, kPlaceThreads, 0));
        B2_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[2], place_kernel<2>, kPlaceThreads, 0));
    }
    long long grid = n_tiles;
    const long long cap = (long long)ctx->sm_count * occ[mode >= 0 && mode < 3 ? mode : 0];   // one full wave, grid-stride over tiles
    if (grid > cap) grid = cap;
    unsigned long long *m = reinterpret_cast<unsigned long long *>(maxsq);
    switch (mode) {
        case B200VEC_BUILD_STORE:
            place_kernel<0><<<(unsigned)grid, kPlaceThreads, 0, st>>>(args, out, m, n_tiles); break;
        case B200VEC_BUILD_MAXONLY:
            place_kernel<1><<<(unsigned)grid, kPlaceThreads, 0, st>>>(args, out, m, n_tiles); break;
        case B200VEC_BUILD_SCALE:
            place_kernel<2><<<(unsigned)grid, kPlaceThreads, 0, st>>>(args, out, m, n_tiles); break;
        default: return fail("b200vec_build", "unknown mode");
    }
    B2_CUDA(cudaGetLastError());
    return 0;
}
#endif  // B2_AB

}  // namespace b200vec
