// write_streams.cu — microbenchmark: what does B200 HBM absorb when every WARP streams two outputs
// (vector: 4 KB per half-frame, spectrogram: 4 KB per frame) the way wf1024_kernel does?
// Big buffers (default 2 x 2 GB) so that the dirty lines left in the 126 MB L2 at kernel end do not
// flatter the number.  Build: nvcc -arch=sm_100a -O3 -o write_streams write_streams.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

// MODE 0: vector STG.64 x16 (256 B / instr) + spectrogram STG.32 x32 (128 B / instr)  [the kernel today]
// MODE 1: both STG.128 x8 (512 B / instr)
// MODE 2: vector only (STG.64)      MODE 3: spectrogram only (STG.32)
// MODE 4: both STG.128, default cache policy (no .cs)
template <int MODE, int CHUNK>
__global__ void __launch_bounds__(512, 1) warp_streams(float2 *vec, float *spec, long long n_units, unsigned *ctr) {
    const int lane = threadIdx.x & 31;
    for (;;) {
        unsigned c = 0;
        if (lane == 0) c = atomicAdd(ctr, 1u);
        c = __shfl_sync(0xffffffffu, c, 0);
        if ((long long)c * CHUNK >= n_units) break;
        for (int i = 0; i < CHUNK; ++i) {
            const long long u = (long long)c * CHUNK + i;
            float2 *v = vec + u * 512;
            float *s = spec + u * 1024;
            if (MODE == 0 || MODE == 2) {
#pragma unroll
                for (int j = 0; j < 16; ++j) __stcs(v + lane + 32 * j, make_float2((float)j, 1.f));
            }
            if (MODE == 0 || MODE == 3) {
#pragma unroll
                for (int j = 0; j < 32; ++j) __stcs(s + lane + 32 * j, (float)j);
            }
            if (MODE == 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) __stcs(reinterpret_cast<float4 *>(v) + lane + 32 * j, make_float4(j, 1.f, 2.f, 3.f));
#pragma unroll
                for (int j = 0; j < 8; ++j) __stcs(reinterpret_cast<float4 *>(s) + lane + 32 * j, make_float4(j, 1.f, 2.f, 3.f));
            }
            if (MODE == 4) {
#pragma unroll
                for (int j = 0; j < 8; ++j) (reinterpret_cast<float4 *>(v) + lane + 32 * j)[0] = make_float4(j, 1.f, 2.f, 3.f);
#pragma unroll
                for (int j = 0; j < 8; ++j) (reinterpret_cast<float4 *>(s) + lane + 32 * j)[0] = make_float4(j, 1.f, 2.f, 3.f);
            }
        }
    }
}

// one linear stream, CTA-contiguous 64 KB chunks (STG.128)
__global__ void __launch_bounds__(512) linear_kernel(float4 *out, long long n_chunks) {
    for (long long c = blockIdx.x; c < n_chunks; c += gridDim.x) {
        float4 *o = out + c * 4096 + threadIdx.x;
#pragma unroll
        for (int i = 0; i < 8; ++i) __stcs(o + 512 * i, make_float4(i, i, i, i));
    }
}

template <class Fn> void timeit(const char *label, double bytes, Fn fn) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    for (int i = 0; i < 2; ++i) fn();
    CK(cudaDeviceSynchronize());
    std::vector<float> ts;
    for (int i = 0; i < 7; ++i) {
        cudaEventRecord(a); fn(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); ts.push_back(ms);
    }
    CK(cudaGetLastError());
    std::sort(ts.begin(), ts.end());
    printf("%-64s med %9.1f us  -> %7.0f GB/s\n", label, ts[3] * 1e3, bytes / ts[3] / 1e6);
    fflush(stdout);
}

int main(int argc, char **argv) {
    const long long units = argc > 1 ? atoll(argv[1]) : 524288;      // half-frames / frames; 4 KB of each stream per unit
    float2 *vec; float *spec; unsigned *ctr;
    CK(cudaMalloc(&vec, units * 4096)); CK(cudaMalloc(&spec, units * 4096)); CK(cudaMalloc(&ctr, 4));
    const double both = (double)units * 8192, one = (double)units * 4096;
    printf("units = %lld (%.0f MB per stream)\n", units, one / 1e6);
    timeit("cudaMemsetAsync, one stream", one, [&] { cudaMemsetAsync(vec, 0, units * 4096); });
    for (int g : {148, 592, 1184})
        { char l[96]; snprintf(l, 96, "linear STG.128 .cs, one stream, grid %d", g);
          timeit(l, one, [&] { linear_kernel<<<g, 512>>>((float4 *)vec, units * 4096 / 65536); }); }
#define WS(MODE, CHUNK, BYTES, NAME) timeit(NAME, BYTES, [&] { cudaMemsetAsync(ctr, 0, 4); warp_streams<MODE, CHUNK><<<148, 512>>>(vec, spec, units, ctr); });
    WS(0, 4, both, "per-warp 4 KB pieces, vec STG.64 + spec STG.32, chunk 4")
    WS(0, 8, both, "per-warp 4 KB pieces, vec STG.64 + spec STG.32, chunk 8")
    WS(0, 1, both, "per-warp 4 KB pieces, vec STG.64 + spec STG.32, chunk 1")
    WS(1, 4, both, "per-warp 4 KB pieces, both STG.128 .cs, chunk 4")
    WS(1, 1, both, "per-warp 4 KB pieces, both STG.128 .cs, chunk 1")
    WS(4, 4, both, "per-warp 4 KB pieces, both STG.128 default, chunk 4")
    WS(2, 4, one, "per-warp 4 KB pieces, vec only STG.64, chunk 4")
    WS(3, 4, one, "per-warp 4 KB pieces, spec only STG.32, chunk 4")
    return 0;
}
