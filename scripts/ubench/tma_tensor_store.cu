// tma_tensor_store.cu — can a 2-D TMA tensor store push the spectrogram's [1024 rows x F frames]
// tiles (64-byte runs at F = 16) as fast as 32 STG per thread?  Build: nvcc -arch=sm_100a -O3 ... -lcuda not needed
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)
typedef CUresult (*EncodeFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                             const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int F, int RB>   // F frames per tile, RB rows per TMA op
__global__ void __launch_bounds__(128) tstore_kernel(const __grid_constant__ CUtensorMap tm, long long n_tiles, int inflight) {
    extern __shared__ __align__(1024) unsigned char sm[];
    for (int i = threadIdx.x; i < 1024 * F; i += blockDim.x) reinterpret_cast<float *>(sm)[i] = 1.0f;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned s = (unsigned)__cvta_generic_to_shared(sm);
        for (long long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
            for (int r = 0; r < 1024; r += RB) {
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                             ::"l"(&tm), "r"((int)(t * F)), "r"(r), "r"(s + r * F * 4) : "memory");
            }
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (inflight == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            else asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

int main() {
    const long long frames = 109374, pitch = 109376;
    const size_t bytes = (size_t)1024 * pitch * 4;
    float *out; CK(cudaMalloc(&out, bytes));
    CK(cudaMemset(out, 0, bytes));
    EncodeFn enc = nullptr;
    cudaDriverEntryPointQueryResult qr;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&enc, cudaEnableDefault, &qr));
    if (!enc) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    auto make = [&](int F, int RB, CUtensorMap *tm) {
        cuuint64_t dims[2] = {(cuuint64_t)frames, 1024};
        cuuint64_t strides[1] = {(cuuint64_t)pitch * 4};
        cuuint32_t box[2] = {(cuuint32_t)F, (cuuint32_t)RB};
        cuuint32_t es[2] = {1, 1};
        CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, out, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d (F=%d RB=%d)\n", (int)r, F, RB); exit(1); }
    };
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    auto timeit = [&](const char *label, auto fn) {
        for (int i = 0; i < 3; ++i) fn();
        CK(cudaDeviceSynchronize());
        std::vector<float> ts;
        for (int i = 0; i < 10; ++i) { cudaEventRecord(a); fn(); cudaEventRecord(b); cudaEventSynchronize(b); float ms; cudaEventElapsedTime(&ms, a, b); ts.push_back(ms); }
        CK(cudaGetLastError());
        std::sort(ts.begin(), ts.end());
        printf("%-56s med %8.1f us -> %6.0f GB/s\n", label, ts[5] * 1e3, 1024.0 * frames * 4 / ts[5] / 1e6);
    };
    CUtensorMap tm;
#define RUN(F, RB, G, INF) { make(F, RB, &tm); CK(cudaFuncSetAttribute(tstore_kernel<F, RB>, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 * F * 4)); \
        char l[96]; snprintf(l, 96, "tensor store box %d x %d, grid %d, %s", F, RB, G, INF ? "3 groups in flight" : "wait each tile"); \
        timeit(l, [&] { tstore_kernel<F, RB><<<G, 128, 1024 * F * 4>>>(tm, frames / F, INF); }); }
    RUN(16, 256, 148, 0) RUN(16, 256, 148, 1) RUN(16, 256, 296, 1) RUN(16, 128, 148, 1) RUN(16, 64, 148, 1)
    RUN(32, 256, 148, 0) RUN(32, 256, 148, 1) RUN(8, 256, 296, 1) RUN(8, 256, 592, 1)
    return 0;
}
