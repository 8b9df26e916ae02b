// fma_rate.cu — FP32 issue rate per SM sub-partition on B200: 3-register FFMA, FFMA with an
// immediate, FADD, FMUL and the butterfly mix, at 1/2/4/8 warps per sub-partition.
#include <cuda_runtime.h>
#include <cstdio>
#define N_IT 4096
template <int MODE>
__global__ void k(float *out, float a, float b, long long *cyc) {
    float r[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) r[i] = threadIdx.x * 0.001f + i;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < N_IT; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            if (MODE == 0) r[i] = __fmaf_rn(r[i], a, b);                 // 3-reg, 2 uniform-ish operands
            if (MODE == 1) r[i] = __fmaf_rn(r[i], 0.98078525f, r[(i + 1) & 15]);   // imm multiplier
            if (MODE == 2) r[i] = r[i] + r[(i + 5) & 15];                // FADD
            if (MODE == 3) r[i] = __fmaf_rn(r[i], r[(i + 3) & 15], r[(i + 7) & 15]);   // 3 distinct regs
            if (MODE == 4) r[i] = r[i] * r[(i + 3) & 15];                // FMUL
        }
    }
    long long t1 = clock64();
    float s = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) s += r[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}
int main() {
    float *out; long long *cyc, h;
    cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
    const char *names[] = {"FFMA r,a,b (2 shared regs)", "FFMA r,imm,r'", "FADD r,r'", "FFMA r,r',r''", "FMUL r,r'"};
    for (int mode = 0; mode < 5; ++mode)
        for (int warps : {4, 8, 16, 32}) {
            for (int rep = 0; rep < 2; ++rep) {
                if (mode == 0) k<0><<<148, warps * 32>>>(out, 1.0001f, 0.5f, cyc);
                if (mode == 1) k<1><<<148, warps * 32>>>(out, 1.0001f, 0.5f, cyc);
                if (mode == 2) k<2><<<148, warps * 32>>>(out, 1.0001f, 0.5f, cyc);
                if (mode == 3) k<3><<<148, warps * 32>>>(out, 1.0001f, 0.5f, cyc);
                if (mode == 4) k<4><<<148, warps * 32>>>(out, 1.0001f, 0.5f, cyc);
                cudaDeviceSynchronize();
            }
            cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
            const double instr_per_smsp = (double)N_IT * 16 * warps / 4;
            printf("%-28s warps/SM %2d: %8lld cycles, %.3f warp-instr/clk/SMSP\n", names[mode], warps, h, instr_per_smsp / h);
        }
    return 0;
}
