// Host harness for vector_b200/csrc/npquant.h (tests/test_npquant.py).
#include "../../vector_b200/csrc/npquant.h"
using namespace b200vec;
extern "C" void b2host_pct_ranks(long long count, double q, long long *lo, long long *hi, float *gamma) {
    const PctRanks r = pct_ranks_f32(count, q);
    *lo = r.lo; *hi = r.hi; *gamma = r.gamma;
}
extern "C" float b2host_lerp(float a, float b, float t) { return lerp_f32(a, b, t); }
