// fft32.cuh — register-resident 32-point DFT building blocks for the nfft = 1024 STFT kernels.
//
// 1024 = 32 x 32 (n = l + 32 j, k = q + 32 s):
//     X[q + 32 s] = sum_l W32^(l s) * ( W1024^(l q) * sum_j x[l + 32 j] W32^(j q) )
// so a 1024-point transform is two 32-point transforms per thread with one data exchange in
// between.  Both 32-point transforms are in-place radix-2 decimation-in-time networks over a
// register array whose indices are all compile-time constants: the bit-reversed input order is
// a renaming of registers, not a permutation.
//
// Instruction economy (the STFT kernel is FP32-issue-bound, profiles/README.md):
//   * a DIT butterfly  (a, b) -> (a + w b, a - w b)  is 6 FMA-pipe instructions:
//         y0 = a + w b   as two chained FMAs per component,   y1 = 2 a - y0   as one FMA
//     (4 when w is 1 or -i);
//   * the window multiply is folded into the first stage of the first transform
//         y0 = wa*a + wb*b,  y1 = wa*a - wb*b        (2 FMUL + 4 FFMA instead of 4 FMUL + 4 FADD)
//   * the W1024^(l q) twiddles are folded into the first stage of the second transform
//         u = Ta*a, y0 = u + Tb*b, y1 = 2u - y0      (10 instead of 8 + 4)
//   32-point transform: 388 instructions bare, 420 with the window, 480 with the twiddles.
//
// The header compiles as plain C++ too (tests/test_fft_core.py builds a host harness with g++
// and checks the very same code against numpy's fp64 FFT), hence the HD macro and own types.
#pragma once
#include <math.h>

#ifdef __CUDACC__
#define B2_HD __host__ __device__ __forceinline__
#else
#define B2_HD inline
#endif

namespace b200vec {

struct cf32 {
    float x, y;
};

B2_HD float b2_fma(float a, float b, float c) {
#ifdef __CUDA_ARCH__
    return __fmaf_rn(a, b, c);
#else
    return fmaf(a, b, c);
#endif
}

B2_HD constexpr int brev5(int i) {
    return ((i & 1) << 4) | ((i & 2) << 2) | (i & 4) | ((i & 8) >> 2) | ((i & 16) >> 4);
}

// cos / sin of 2*pi*k/32, k = 0..15, float32-rounded literals
template <int K> struct W32c;
template <int K> struct W32s;
#define B2_W32(K, C, S)                                            \
    template <> struct W32c<K> { static constexpr float v = C; };  \
    template <> struct W32s<K> { static constexpr float v = S; };
B2_W32(0, 1.0f, 0.0f)
B2_W32(1, 0.980785251f, 0.195090324f)
B2_W32(2, 0.923879504f, 0.382683426f)
B2_W32(3, 0.831469595f, 0.555570245f)
B2_W32(4, 0.707106769f, 0.707106769f)
B2_W32(5, 0.555570245f, 0.831469595f)
B2_W32(6, 0.382683426f, 0.923879504f)
B2_W32(7, 0.195090324f, 0.980785251f)
B2_W32(8, 0.0f, 1.0f)
B2_W32(9, -0.195090324f, 0.980785251f)
B2_W32(10, -0.382683426f, 0.923879504f)
B2_W32(11, -0.555570245f, 0.831469595f)
B2_W32(12, -0.707106769f, 0.707106769f)
B2_W32(13, -0.831469595f, 0.555570245f)
B2_W32(14, -0.923879504f, 0.382683426f)
B2_W32(15, -0.980785251f, 0.195090324f)
#undef B2_W32

// (a, b) -> (a + w b, a - w b),  w = exp(-2*pi*i*K/32) = c - i s
template <int K>
B2_HD void dit_bfly(cf32 &a, cf32 &b) {
    if constexpr (K == 0) {
        const cf32 t = b;
        b.x = a.x - t.x; b.y = a.y - t.y;
        a.x = a.x + t.x; a.y = a.y + t.y;
    } else if constexpr (K == 8) {          // w = -i : w b = (b.y, -b.x)
        const cf32 t = b;
        b.x = a.x - t.y; b.y = a.y + t.x;
        a.x = a.x + t.y; a.y = a.y - t.x;
    } else {
        constexpr float c = W32c<K>::v, s = W32s<K>::v;
        // w b = (b.x c + b.y s,  b.y c - b.x s)
        const float y0x = b2_fma(b.x, c, b2_fma(b.y, s, a.x));
        const float y0y = b2_fma(b.y, c, b2_fma(-b.x, s, a.y));
        b.x = b2_fma(2.0f, a.x, -y0x);
        b.y = b2_fma(2.0f, a.y, -y0y);
        a.x = y0x; a.y = y0y;
    }
}

// stage with span M = 2^S (half = M/2) over v[0..31], twiddle W_M^k = W32^(k*32/M)
template <int M, int I>
B2_HD void dit_stage_one(cf32 (&v)[32]) {
    constexpr int half = M / 2;
    constexpr int blk = I / half, k = I % half;
    constexpr int i = blk * M + k;
    dit_bfly<k * (32 / M)>(v[i], v[i + half]);
}
#define B2_STAGE16(M, v)                                                                    \
    dit_stage_one<M, 0>(v); dit_stage_one<M, 1>(v); dit_stage_one<M, 2>(v); dit_stage_one<M, 3>(v);   \
    dit_stage_one<M, 4>(v); dit_stage_one<M, 5>(v); dit_stage_one<M, 6>(v); dit_stage_one<M, 7>(v);   \
    dit_stage_one<M, 8>(v); dit_stage_one<M, 9>(v); dit_stage_one<M, 10>(v); dit_stage_one<M, 11>(v); \
    dit_stage_one<M, 12>(v); dit_stage_one<M, 13>(v); dit_stage_one<M, 14>(v); dit_stage_one<M, 15>(v);

// stages 2..5 of the in-place DIT network (after a first stage of the caller's choice)
B2_HD void dit_tail(cf32 (&v)[32]) {
    B2_STAGE16(4, v)
    B2_STAGE16(8, v)
    B2_STAGE16(16, v)
    B2_STAGE16(32, v)
}

// ---- first transform: v[i] holds RAW sample x[brev5(i)] (j = brev5(i)), win[j] its window tap.
// On exit v[q] = sum_j win[j] x[j] W32^(j q).
template <int P>
B2_HD void win_bfly(cf32 (&v)[32], const float (&win)[32]) {
    constexpr int i = 2 * P;
    const float wa = win[brev5(i)], wb = win[brev5(i + 1)];
    const float ux = wa * v[i].x, uy = wa * v[i].y;
    const cf32 b = v[i + 1];
    v[i].x = b2_fma(wb, b.x, ux);     v[i].y = b2_fma(wb, b.y, uy);
    v[i + 1].x = b2_fma(-wb, b.x, ux); v[i + 1].y = b2_fma(-wb, b.y, uy);
}
B2_HD void fft32_windowed(cf32 (&v)[32], const float (&win)[32]) {
    win_bfly<0>(v, win); win_bfly<1>(v, win); win_bfly<2>(v, win); win_bfly<3>(v, win);
    win_bfly<4>(v, win); win_bfly<5>(v, win); win_bfly<6>(v, win); win_bfly<7>(v, win);
    win_bfly<8>(v, win); win_bfly<9>(v, win); win_bfly<10>(v, win); win_bfly<11>(v, win);
    win_bfly<12>(v, win); win_bfly<13>(v, win); win_bfly<14>(v, win); win_bfly<15>(v, win);
    dit_tail(v);
}

// Same transform with the taps fetched four at a time (shared memory instead of 32 resident
// registers): w4(k) = (win[brev5(4k)], win[brev5(4k+1)], win[brev5(4k+2)], win[brev5(4k+3)]).
struct f32x4 {
    float a, b, c, d;
};
template <int P>
B2_HD void win_bfly2(cf32 (&v)[32], const float wa, const float wb) {
    constexpr int i = 2 * P;
    const float ux = wa * v[i].x, uy = wa * v[i].y;
    const cf32 b = v[i + 1];
    v[i].x = b2_fma(wb, b.x, ux);     v[i].y = b2_fma(wb, b.y, uy);
    v[i + 1].x = b2_fma(-wb, b.x, ux); v[i + 1].y = b2_fma(-wb, b.y, uy);
}
template <int K, class W4>
B2_HD void win_quad(cf32 (&v)[32], const W4 &w4) {
    const f32x4 t = w4(K);
    win_bfly2<2 * K>(v, t.a, t.b);
    win_bfly2<2 * K + 1>(v, t.c, t.d);
}
template <class W4>
B2_HD void fft32_windowed_q(cf32 (&v)[32], const W4 &w4) {
    win_quad<0>(v, w4); win_quad<1>(v, w4); win_quad<2>(v, w4); win_quad<3>(v, w4);
    win_quad<4>(v, w4); win_quad<5>(v, w4); win_quad<6>(v, w4); win_quad<7>(v, w4);
    dit_tail(v);
}

// ---- second transform: v[i] holds RAW B_l[q] with l = brev5(i); tw(l) returns W1024^(l q).
// On exit v[s] = X[q + 32 s].
template <int P, class TW>
B2_HD void tw_bfly(cf32 (&v)[32], const TW &tw) {
    constexpr int i = 2 * P;
    constexpr int la = brev5(i), lb = brev5(i + 1);
    cf32 u;
    if constexpr (la == 0) {
        u = v[i];
    } else {
        const cf32 ta = tw(la);
        u.x = b2_fma(-v[i].y, ta.y, v[i].x * ta.x);
        u.y = b2_fma(v[i].y, ta.x, v[i].x * ta.y);
    }
    const cf32 tb = tw(lb);
    const cf32 b = v[i + 1];
    const float y0x = b2_fma(b.x, tb.x, b2_fma(-b.y, tb.y, u.x));
    const float y0y = b2_fma(b.x, tb.y, b2_fma(b.y, tb.x, u.y));
    v[i + 1].x = b2_fma(2.0f, u.x, -y0x);
    v[i + 1].y = b2_fma(2.0f, u.y, -y0y);
    v[i].x = y0x; v[i].y = y0y;
}
template <class TW>
B2_HD void fft32_twiddled(cf32 (&v)[32], const TW &tw) {
    tw_bfly<0>(v, tw); tw_bfly<1>(v, tw); tw_bfly<2>(v, tw); tw_bfly<3>(v, tw);
    tw_bfly<4>(v, tw); tw_bfly<5>(v, tw); tw_bfly<6>(v, tw); tw_bfly<7>(v, tw);
    tw_bfly<8>(v, tw); tw_bfly<9>(v, tw); tw_bfly<10>(v, tw); tw_bfly<11>(v, tw);
    tw_bfly<12>(v, tw); tw_bfly<13>(v, tw); tw_bfly<14>(v, tw); tw_bfly<15>(v, tw);
    dit_tail(v);
}

// Same transform with the twiddles fetched two at a time: tq[P] = (W1024^(q la), W1024^(q lb)) with
// la = brev5(2 P), lb = brev5(2 P + 1) — the order the first stage consumes them, one 16-byte
// load per butterfly (table built by get_tw1024_pairs).
struct cf32x2 {
    float ax, ay, bx, by;
};
template <int P, class TW2>
B2_HD void tw2_bfly(cf32 (&v)[32], const TW2 &tw2) {
    constexpr int i = 2 * P;
    const cf32x2 t = tw2(P);
    cf32 u;
    if constexpr (P == 0) {                     // la = 0: twiddle 1
        u = v[i];
    } else {
        u.x = b2_fma(-v[i].y, t.ay, v[i].x * t.ax);
        u.y = b2_fma(v[i].y, t.ax, v[i].x * t.ay);
    }
    const cf32 b = v[i + 1];
    const float y0x = b2_fma(b.x, t.bx, b2_fma(-b.y, t.by, u.x));
    const float y0y = b2_fma(b.x, t.by, b2_fma(b.y, t.bx, u.y));
    v[i + 1].x = b2_fma(2.0f, u.x, -y0x);
    v[i + 1].y = b2_fma(2.0f, u.y, -y0y);
    v[i].x = y0x; v[i].y = y0y;
}
template <class TW2>
B2_HD void fft32_twiddled_pairs(cf32 (&v)[32], const TW2 &tw2) {
    tw2_bfly<0>(v, tw2); tw2_bfly<1>(v, tw2); tw2_bfly<2>(v, tw2); tw2_bfly<3>(v, tw2);
    tw2_bfly<4>(v, tw2); tw2_bfly<5>(v, tw2); tw2_bfly<6>(v, tw2); tw2_bfly<7>(v, tw2);
    tw2_bfly<8>(v, tw2); tw2_bfly<9>(v, tw2); tw2_bfly<10>(v, tw2); tw2_bfly<11>(v, tw2);
    tw2_bfly<12>(v, tw2); tw2_bfly<13>(v, tw2); tw2_bfly<14>(v, tw2); tw2_bfly<15>(v, tw2);
    dit_tail(v);
}

}  // namespace b200vec
