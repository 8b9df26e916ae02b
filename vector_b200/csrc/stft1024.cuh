// stft1024.cuh — stub (fast nfft=1024 path not built yet)
#pragma once
#include "common.cuh"
namespace b200vec {
inline int stft1024_setup() { return 0; }
inline bool stft1024_enabled() { return false; }
inline int launch_stft1024(b200vec_ctx *, const float2 *, long long, long long, long long, int, int,
                           const float *, float, long long, long long, float *, long long, int *,
                           cudaStream_t) { return fail("stft1024", "not built"); }
}
