// maxspans.cuh — host-side planner of the max-abs pass (normalisation, unified_gui.py:1778-1782 / main.py:274-277).
//
// m = max |v| is needed before the first normalised sample can be stored, so the max pass is dead time for
// HBM: it has to be SHORT, not merely cheap.  The caller already restricts it to sample ranges whose max is
// the max of the whole vector (hostmath.max_eval_ranges: one common period of the packet trains + head and
// tail).  Inside those ranges most samples need no evaluation at all:
//   * where NO instance lies, v = 0;
//   * where exactly ONE instance lies, |v| = |y_p[n]| <= max |y_p|, and if the caller vouches
//     (b200vec_pkt_desc.reserved bit 0, hostmath.free_instance_flags) that every sample of packet p appears
//     unsummed somewhere in the vector, max |y_p| is attained by the vector itself: it enters the reduction
//     through the bank's per-slot packet max (PlaceDesc::pmax) and these samples add nothing.
// What is left — samples under >= 2 instances, or under an instance of an unflagged packet — is found
// here by one sweep over the sorted instance boundaries and handed to the kernels as a compact list of
// spans of <= 512 samples: one warp-sized piece of work each, ~1 per warp at the BASELINE config (1 238
// spans instead of 16 298 half-frames), so the pass is one L2 round trip long.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>

namespace b200vec {

struct MaxSpan {
    long long s0;                // first sample (vector-absolute)
    int n;                       // 1..512 samples
    int vec;                     // batch builds: vector index (0 otherwise)
};

struct MaxSpanPacket {
    long long start, period, count;
    long long len;
    int free_inst;
};

constexpr long long kMaxSpanInstances = 4000000;     // above this many instance boundaries: evaluate the ranges whole
constexpr long long kMaxSpanMergeGap = 128;          // must-evaluate pieces closer than this are merged (fewer, fuller spans)

inline long long ms_floor_div(long long a, long long b) {   // b > 0
    long long q = a / b;
    return (a % b != 0 && a < 0) ? q - 1 : q;
}

inline void ms_emit(std::vector<MaxSpan> &out, long long a, long long b, int vec) {
    for (long long s = a; s < b; s += 512) out.push_back(MaxSpan{s, (int)std::min<long long>(512, b - s), vec});
}

// Appends the must-evaluate spans of the ranges [lo_r, hi_r) to `out`.
inline void max_spans(const MaxSpanPacket *pk, int np, const long long *ranges, int n_ranges, int vec,
                      std::vector<MaxSpan> &out) {
    struct Ev { long long pos; int delta; int unflagged; };
    std::vector<Ev> ev;
    for (int r = 0; r < n_ranges; ++r) {
        const long long lo = ranges[2 * r], hi = ranges[2 * r + 1];
        if (hi <= lo) continue;
        long long n_inst = 0;
        for (int p = 0; p < np; ++p) {
            if (pk[p].count <= 0) continue;
            long long klo = std::max<long long>(0, ms_floor_div(lo - pk[p].len - pk[p].start, pk[p].period) + 1);
            long long khi = std::min<long long>(pk[p].count - 1, ms_floor_div(hi - 1 - pk[p].start, pk[p].period));
            if (khi >= klo) n_inst += khi - klo + 1;
        }
        if (n_inst > kMaxSpanInstances) { ms_emit(out, lo, hi, vec); continue; }
        ev.clear();
        ev.reserve((size_t)(2 * n_inst));
        for (int p = 0; p < np; ++p) {
            if (pk[p].count <= 0) continue;
            long long klo = std::max<long long>(0, ms_floor_div(lo - pk[p].len - pk[p].start, pk[p].period) + 1);
            long long khi = std::min<long long>(pk[p].count - 1, ms_floor_div(hi - 1 - pk[p].start, pk[p].period));
            const int unf = pk[p].free_inst ? 0 : 1;
            for (long long k = klo; k <= khi; ++k) {
                const long long s = pk[p].start + k * pk[p].period;
                ev.push_back(Ev{std::max(s, lo), +1, unf});
                ev.push_back(Ev{std::min(s + pk[p].len, hi), -1, unf});
            }
        }
        std::sort(ev.begin(), ev.end(), [](const Ev &x, const Ev &y) { return x.pos != y.pos ? x.pos < y.pos : x.delta < y.delta; });
        long long cov = 0, unf = 0, seg_a = -1, seg_b = -1;          // current merged must-evaluate piece [seg_a, seg_b)
        for (size_t i = 0; i < ev.size(); ++i) {
            cov += ev[i].delta; unf += ev[i].delta * ev[i].unflagged;
            if (i + 1 == ev.size() || ev[i + 1].pos == ev[i].pos) continue;     // state settles after the last event at a position
            if (cov >= 2 || unf >= 1) {
                const long long a = ev[i].pos, b = ev[i + 1].pos;
                if (seg_a >= 0 && a - seg_b <= kMaxSpanMergeGap) seg_b = b;
                else {
                    if (seg_a >= 0) ms_emit(out, seg_a, seg_b, vec);
                    seg_a = a; seg_b = b;
                }
            }
        }
        if (seg_a >= 0) ms_emit(out, seg_a, seg_b, vec);
    }
}

// hostmath.max_eval_ranges in C++ (batch builds plan their vectors here): sample ranges whose max is the max over
// [0, total).  With T = lcm of the periods, A = max_p max(0, s_p + L_p - P_p), B = min_p (s_p + K_p P_p - T) the
// composite repeats bit for bit, v(i + T) == v(i) for A <= i < B, so [0, A + T) U [B + T, total) suffices; the
// whole vector when the trains share no usable period.  Writes up to two [lo, hi) pairs, returns their number.
inline int max_eval_ranges(const MaxSpanPacket *pk, int np, long long total, long long *ranges) {
    if (total <= 0) return 0;
    ranges[0] = 0; ranges[1] = total;
    long long t = 1, a = 0, b = 0;
    bool any = false, ok = true;
    for (int p = 0; p < np && ok; ++p) {
        if (pk[p].count <= 0) continue;
        if (pk[p].count < 2) { ok = false; break; }
        long long x = t, y = pk[p].period;
        while (y) { const long long r = x % y; x = y; y = r; }
        t = t / x;
        if (t > total / pk[p].period) { ok = false; break; }          // lcm >= total (and no overflow)
        t *= pk[p].period;
        if (t >= total) { ok = false; break; }
        any = true;
    }
    if (!ok || !any) return 1;
    bool first = true;
    for (int p = 0; p < np; ++p) {
        if (pk[p].count <= 0) continue;
        const long long ap = std::max<long long>(0, pk[p].start + pk[p].len - pk[p].period);
        const long long bp = pk[p].start + pk[p].count * pk[p].period - t;
        a = first ? ap : std::max(a, ap);
        b = first ? bp : std::min(b, bp);
        first = false;
    }
    if (b <= a || a + t >= total) return 1;
    ranges[0] = 0; ranges[1] = a + t;
    if (b + t < total) { ranges[2] = b + t; ranges[3] = total; return 2; }
    return 1;
}

// FNV-1a over the planner's inputs: the uploaded list is reused while they do not change (graph replays,
// repeated steps of the same scenario).
inline unsigned long long max_spans_key(const MaxSpanPacket *pk, int np, const long long *ranges, int n_ranges) {
    unsigned long long h = 1469598103934665603ull;
    auto mix = [&h](unsigned long long v) {
        for (int i = 0; i < 8; ++i) { h ^= (v >> (8 * i)) & 0xffu; h *= 1099511628211ull; }
    };
    mix((unsigned long long)np); mix((unsigned long long)n_ranges);
    for (int p = 0; p < np; ++p) {
        mix((unsigned long long)pk[p].start); mix((unsigned long long)pk[p].period); mix((unsigned long long)pk[p].count);
        mix((unsigned long long)pk[p].len); mix((unsigned long long)pk[p].free_inst);
    }
    for (int r = 0; r < 2 * n_ranges; ++r) mix((unsigned long long)ranges[r]);
    return h;
}

}  // namespace b200vec
