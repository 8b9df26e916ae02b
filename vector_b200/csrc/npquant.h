// npquant.h — numpy's percentile index arithmetic for float32 data, restated in C++ (host code of libb200vec).
//
// np.percentile(a, q) (method 'linear') on a float32 array evaluates the virtual index (n - 1) * (q / 100)
// IN FLOAT32 (numpy/lib/_function_base_impl.py: percentile -> _quantile -> _get_indexes / _get_gamma / _lerp):
// above 2^24 elements the neighbours themselves are decided by float32 rounding.  normalize_spectrogram
// (utils.py:365,372-373) calls it three times; vector_b200/hostmath.py::percentile_ranks / lerp hold the Python
// statement of the same arithmetic (pinned against numpy by tests/test_hostmath.py), tests/test_npquant.py pins
// this one against the Python one.  Plain C++ (no CUDA): the test harness builds it with g++.
#pragma once
#include <cmath>
#include <cstdint>

#ifdef __CUDACC__
#define B2_NPQ_HD __host__ __device__ inline
#else
#define B2_NPQ_HD inline
#endif

namespace b200vec {

struct PctRanks {
    int64_t lo, hi;      // 0-based positions in the ascending sort
    float gamma;         // interpolation weight, float32
};

B2_NPQ_HD PctRanks pct_ranks_f32(int64_t count, double q) {
    const float qq = (float)q / 100.0f;                       // np.true_divide(q, float32(100)): float32
    const float cm1 = (float)(count - 1);                     // python int -> float32 (weak scalar)
    const float vi = cm1 * qq;                                // virtual index, float32
    float prev = floorf(vi);
    float next = prev + 1.0f;                                 // may round back to prev above 2^24
    if (vi >= cm1) { prev = -1.0f; next = -1.0f; }            // _get_indexes: clamp to the last element
    if (vi < 0.0f) { prev = 0.0f; next = 0.0f; }
    const int64_t prev_i = (int64_t)prev, next_i = (int64_t)next;
    // _get_gamma: float32 array - intp array -> float64, then cast to the array's dtype
    const float gamma = (float)((double)vi - (double)prev_i);
    PctRanks r;
    r.lo = prev_i < 0 ? prev_i + count : prev_i;
    r.hi = next_i < 0 ? next_i + count : next_i;
    r.gamma = gamma;
    return r;
}

// numpy's _lerp in float32: a + (b - a) * t, switched to b - (b - a) * (1 - t) for t >= 0.5.  Every operation is
// rounded on its own (numpy evaluates ufunc by ufunc): on the device the intrinsics keep nvcc from contracting
// the multiply-add pairs into FMAs.
B2_NPQ_HD float lerp_f32(float a, float b, float t) {
#ifdef __CUDA_ARCH__
    const float d = __fsub_rn(b, a);
    float out = __fadd_rn(a, __fmul_rn(d, t));
    if (t >= 0.5f) out = __fsub_rn(b, __fmul_rn(d, __fsub_rn(1.0f, t)));
    return out;
#else
    const float d = b - a;
    float out = a + d * t;
    if (t >= 0.5f) out = b - d * (1.0f - t);
    return out;
#endif
}

}  // namespace b200vec
