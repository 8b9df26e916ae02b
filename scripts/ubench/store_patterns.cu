// store_patterns.cu — microbenchmark: how fast can B200 absorb the spectrogram's write pattern?
// Output matrix: float [1024 rows][pitch], one "tile" = all 1024 rows x F consecutive frames
// (what one FFT tile of F frames produces).  Build: nvcc -arch=sm_100a -O3 -o store_patterns store_patterns.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

constexpr int ROWS = 1024;

template <int MODE> __device__ __forceinline__ void st(float *p, float v) {
    if (MODE == 0) *p = v;
    else if (MODE == 1) __stcs(p, v);
    else if (MODE == 2) __stcg(p, v);
    else __stwt(p, v);
}

// F frames per tile; warp-instr covers (32 / F) rows x F frames (F <= 32) — the fused kernel's pattern at F = 16
template <int F, int MODE>
__global__ void __launch_bounds__(512) tile_rows_kernel(float *out, long long pitch, long long n_tiles) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int RPW = 32 / F;                       // rows per warp-instruction
    const int f = lane % F, r0 = warp * RPW + lane / F;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        float *o = out + (long long)r0 * pitch + t * F + f;
#pragma unroll 8
        for (int i = 0; i < ROWS / (16 * RPW); ++i) st<MODE>(o + (long long)i * 16 * RPW * pitch, (float)i);
    }
}

// same tile, but each thread writes float4 (4 consecutive frames): warp-instr covers 32*4/F rows... F >= 4
template <int F, int MODE>
__global__ void __launch_bounds__(512) tile_rows_v4_kernel(float *out, long long pitch, long long n_tiles) {
    const int tid = threadIdx.x;
    constexpr int TPR = F / 4;                        // threads per row
    const int f4 = tid % TPR, r0 = tid / TPR;         // 512 / TPR rows per pass
    constexpr int RPP = 512 / TPR;
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        float *o = out + (long long)r0 * pitch + t * F + 4 * f4;
#pragma unroll 4
        for (int i = 0; i < ROWS / RPP; ++i) {
            float4 v = make_float4(i, i, i, i);
            float4 *p = reinterpret_cast<float4 *>(o + (long long)i * RPP * pitch);
            if (MODE == 1) __stcs(p, v); else *p = v;
        }
    }
}

// linear writer: each CTA writes contiguous 64 KB chunks with STG.128
template <int MODE>
__global__ void __launch_bounds__(512) linear_kernel(float4 *out, long long n_chunks) {
    for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        float4 *o = out + c * 4096 + threadIdx.x;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float4 v = make_float4(i, i, i, i);
            if (MODE == 1) __stcs(o + 512 * i, v); else o[512 * i] = v;
        }
    }
}

// linear writer via TMA bulk store from smem (64 KB per bulk op, CH bytes each)
template <int CH>
__global__ void __launch_bounds__(128) linear_tma_kernel(char *out, long long n_chunks) {
    extern __shared__ __align__(128) unsigned char sm[];
    for (int i = threadIdx.x; i < CH / 4; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned s = (unsigned)__cvta_generic_to_shared(sm);
        for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + c * CH), "r"(s), "r"(CH) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

// tile writer via per-row TMA bulk stores: each row-run of F floats is one bulk copy (F*4 bytes >= 16)
template <int F>
__global__ void __launch_bounds__(512) tile_rows_bulk_kernel(float *out, long long pitch, long long n_tiles) {
    extern __shared__ __align__(128) unsigned char sm[];
    for (int i = threadIdx.x; i < ROWS * F; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const unsigned s = (unsigned)__cvta_generic_to_shared(sm);
    for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        for (int r = threadIdx.x; r < ROWS; r += blockDim.x) {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                         ::"l"(out + (long long)r * pitch + t * F), "r"(s + r * F * 4), "r"(F * 4) : "memory");
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncthreads();
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

template <class Fn> float timeit(const char *label, double bytes, Fn fn) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 3; ++i) fn();
    CK(cudaDeviceSynchronize());
    std::vector<float> ts;
    for (int i = 0; i < 10; ++i) {
        cudaEventRecord(a); fn(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); ts.push_back(ms);
    }
    CK(cudaGetLastError());
    std::sort(ts.begin(), ts.end());
    printf("%-52s med %8.1f us  min %8.1f us  -> %7.0f GB/s\n", label, ts[5] * 1e3, ts[0] * 1e3, bytes / ts[5] / 1e6);
    fflush(stdout);
    return ts[5];
}

int main() {
    const long long frames = 109374, pitch = 109376;
    const size_t bytes = (size_t)ROWS * pitch * 4;
    float *out; CK(cudaMalloc(&out, bytes + (1 << 20)));
    CK(cudaMemset(out, 0, bytes));
    const double nb = (double)ROWS * frames * 4;
    timeit("cudaMemsetAsync", (double)bytes, [&] { cudaMemsetAsync(out, 0, bytes); });
    for (int g : {148, 296, 592, 1184}) {
        char l[96];
        snprintf(l, 96, "linear STG.128 default, grid %d", g);
        timeit(l, (double)bytes, [&] { linear_kernel<0><<<g, 512>>>((float4 *)out, bytes / 65536); });
        snprintf(l, 96, "linear STG.128 .cs, grid %d", g);
        timeit(l, (double)bytes, [&] { linear_kernel<1><<<g, 512>>>((float4 *)out, bytes / 65536); });
    }
    CK(cudaFuncSetAttribute(linear_tma_kernel<65536>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
    CK(cudaFuncSetAttribute(linear_tma_kernel<16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384));
    for (int g : {148, 296}) {
        char l[96];
        snprintf(l, 96, "linear TMA bulk 64 KB, grid %d", g);
        timeit(l, (double)bytes, [&] { linear_tma_kernel<65536><<<g, 128, 65536>>>((char *)out, bytes / 65536); });
        snprintf(l, 96, "linear TMA bulk 16 KB, grid %d", g);
        timeit(l, (double)bytes, [&] { linear_tma_kernel<16384><<<g, 128, 16384>>>((char *)out, bytes / 16384); });
    }
#define TR(F, MODE, G, NAME) { char l[96]; snprintf(l, 96, "tile 1024 x %d STG.32 %s, grid %d", F, NAME, G); \
        timeit(l, nb, [&] { tile_rows_kernel<F, MODE><<<G, 512>>>(out, pitch, frames / F); }); }
    for (int g : {148, 296, 592}) {
        TR(8, 0, g, "default") TR(8, 1, g, ".cs")
        TR(16, 0, g, "default") TR(16, 1, g, ".cs") TR(16, 2, g, ".cg") TR(16, 3, g, ".wt")
        TR(32, 0, g, "default") TR(32, 1, g, ".cs")
    }
#define TV(F, MODE, G, NAME) { char l[96]; snprintf(l, 96, "tile 1024 x %d STG.128 %s, grid %d", F, NAME, G); \
        timeit(l, nb, [&] { tile_rows_v4_kernel<F, MODE><<<G, 512>>>(out, pitch, frames / F); }); }
    for (int g : {148, 296}) {
        TV(16, 0, g, "default") TV(16, 1, g, ".cs") TV(32, 0, g, "default") TV(32, 1, g, ".cs") TV(64, 0, g, "default") TV(64, 1, g, ".cs")
        TV(128, 1, g, ".cs")
    }
    CK(cudaFuncSetAttribute(tile_rows_bulk_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * 16 * 4));
    CK(cudaFuncSetAttribute(tile_rows_bulk_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * 32 * 4));
    CK(cudaFuncSetAttribute(tile_rows_bulk_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, ROWS * 8 * 4));
    timeit("tile 1024 x 8 per-row bulk 32 B, grid 296", nb, [&] { tile_rows_bulk_kernel<8><<<296, 512, ROWS * 8 * 4>>>(out, pitch, frames / 8); });
    timeit("tile 1024 x 16 per-row bulk 64 B, grid 148", nb, [&] { tile_rows_bulk_kernel<16><<<148, 512, ROWS * 16 * 4>>>(out, pitch, frames / 16); });
    timeit("tile 1024 x 16 per-row bulk 64 B, grid 296", nb, [&] { tile_rows_bulk_kernel<16><<<296, 512, ROWS * 16 * 4>>>(out, pitch, frames / 16); });
    timeit("tile 1024 x 32 per-row bulk 128 B, grid 148", nb, [&] { tile_rows_bulk_kernel<32><<<148, 512, ROWS * 32 * 4>>>(out, pitch, frames / 32); });
    return 0;
}
