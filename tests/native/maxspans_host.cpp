// Host harness for vector_b200/csrc/maxspans.cuh (the max-pass planner the library runs on the host):
// exposes max_spans / max_eval_ranges through a C interface for tests/test_maxspans.py.
#include "../../vector_b200/csrc/maxspans.cuh"
#include <cstring>
using namespace b200vec;

// pk: [np][5] = start, period, count, len, free_inst.  Returns the number of spans; writes up to cap of them as
// (s0, n) pairs.
extern "C" long long b2host_max_spans(const long long *pk, int np, const long long *ranges, int n_ranges,
                                      long long *out, long long cap) {
    std::vector<MaxSpanPacket> v(np);
    for (int p = 0; p < np; ++p) v[p] = MaxSpanPacket{pk[5 * p], pk[5 * p + 1], pk[5 * p + 2], pk[5 * p + 3], (int)pk[5 * p + 4]};
    std::vector<MaxSpan> spans;
    max_spans(v.data(), np, ranges, n_ranges, 0, spans);
    for (size_t i = 0; i < spans.size() && (long long)i < cap; ++i) { out[2 * i] = spans[i].s0; out[2 * i + 1] = spans[i].n; }
    return (long long)spans.size();
}

extern "C" int b2host_max_eval_ranges(const long long *pk, int np, long long total, long long *ranges) {
    std::vector<MaxSpanPacket> v(np);
    for (int p = 0; p < np; ++p) v[p] = MaxSpanPacket{pk[5 * p], pk[5 * p + 1], pk[5 * p + 2], pk[5 * p + 3], (int)pk[5 * p + 4]};
    return max_eval_ranges(v.data(), np, total, ranges);
}
