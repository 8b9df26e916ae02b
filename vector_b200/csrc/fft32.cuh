// fft32.cuh — register-resident 32-point DFT building blocks for the nfft = 1024 STFT kernels.
//
// 1024 = 32 x 32 (n = l + 32 j, k = q + 32 s):
//     X[q + 32 s] = sum_l W32^(l s) * ( W1024^(l q) * sum_j x[l + 32 j] W32^(j q) )
// so a 1024-point transform is two 32-point transforms per thread with one data exchange in
// between.  Both 32-point transforms are in-place radix-2 decimation-in-time networks over a
// register array whose indices are all compile-time constants: the bit-reversed input order is
// a renaming of registers, not a permutation.
//
// Instruction economy (the STFT kernel is FP32-issue-bound, profiles/README.md).  sm_100a has packed
// FP32 instructions (FFMA2 / FADD2 / FMUL2: one issue slot, two IEEE fp32 results) whose operands take
// a half swap and per-half signs for free (R.F32x2.LO_HI.NP in SASS), so a complex number kept as one
// (re, im) register pair costs ONE instruction per complex add, scale or "+ s * (-i b)" term.  All
// arithmetic below goes through the c_* helpers: packed on the device, the same IEEE operations in the
// same order as scalar code on the host (tests/test_fft_core.py) — bit-identical results either way.
//   * a DIT butterfly  (a, b) -> (a + w b, a - w b)  is 3 packed instructions (6 scalar):
//         u = a + s (-i b),  y0 = u + c b   (two chained FMAs),   y1 = 2 a - y0   (one FMA)
//     (2 when w is 1 or -i);
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

struct alignas(8) cf32 {
    float x, y;
};

// ---- packed complex helpers.  Every function is one FFMA2 / FADD2 / FMUL2 on the device.
#ifdef __CUDA_ARCH__
#define B2_P2(v) make_float2((v).x, (v).y)
B2_HD cf32 b2_from2(float2 r) { return cf32{r.x, r.y}; }
// a + b, a - b
B2_HD cf32 c_add(cf32 a, cf32 b) { return b2_from2(__fadd2_rn(B2_P2(a), B2_P2(b))); }
B2_HD cf32 c_sub(cf32 a, cf32 b) { return b2_from2(__fadd2_rn(B2_P2(a), make_float2(-b.x, -b.y))); }
// a + (-i b) = a + (b.y, -b.x);   a - (-i b)
B2_HD cf32 c_add_mi(cf32 a, cf32 b) { return b2_from2(__fadd2_rn(B2_P2(a), make_float2(b.y, -b.x))); }
B2_HD cf32 c_sub_mi(cf32 a, cf32 b) { return b2_from2(__fadd2_rn(B2_P2(a), make_float2(-b.y, b.x))); }
// s * a (real scalar)
B2_HD cf32 c_scale(cf32 a, float s) { return b2_from2(__fmul2_rn(B2_P2(a), make_float2(s, s))); }
// s * b + c,  -s * b + c
B2_HD cf32 c_fma(cf32 b, float s, cf32 c) { return b2_from2(__ffma2_rn(B2_P2(b), make_float2(s, s), B2_P2(c))); }
B2_HD cf32 c_fnma(cf32 b, float s, cf32 c) { return b2_from2(__ffma2_rn(make_float2(-b.x, -b.y), make_float2(s, s), B2_P2(c))); }
// s * (-i b) + c = (s b.y + c.x, -s b.x + c.y);   s * (i b) + c = (-s b.y + c.x, s b.x + c.y)
B2_HD cf32 c_fma_mi(cf32 b, float s, cf32 c) { return b2_from2(__ffma2_rn(make_float2(b.y, -b.x), make_float2(s, s), B2_P2(c))); }
B2_HD cf32 c_fma_pi(cf32 b, float s, cf32 c) { return b2_from2(__ffma2_rn(make_float2(-b.y, b.x), make_float2(s, s), B2_P2(c))); }
// s * (i b)  (no addend)
B2_HD cf32 c_scale_pi(cf32 b, float s) { return b2_from2(__fmul2_rn(make_float2(-b.y, b.x), make_float2(s, s))); }
// 2 a - y
B2_HD cf32 c_2a_minus(cf32 a, cf32 y) { return b2_from2(__ffma2_rn(B2_P2(a), make_float2(2.0f, 2.0f), make_float2(-y.x, -y.y))); }
#undef B2_P2
#else
B2_HD cf32 c_add(cf32 a, cf32 b) { return cf32{a.x + b.x, a.y + b.y}; }
B2_HD cf32 c_sub(cf32 a, cf32 b) { return cf32{a.x - b.x, a.y - b.y}; }
B2_HD cf32 c_add_mi(cf32 a, cf32 b) { return cf32{a.x + b.y, a.y - b.x}; }
B2_HD cf32 c_sub_mi(cf32 a, cf32 b) { return cf32{a.x - b.y, a.y + b.x}; }
B2_HD cf32 c_scale(cf32 a, float s) { return cf32{a.x * s, a.y * s}; }
B2_HD cf32 c_fma(cf32 b, float s, cf32 c) { return cf32{fmaf(b.x, s, c.x), fmaf(b.y, s, c.y)}; }
B2_HD cf32 c_fnma(cf32 b, float s, cf32 c) { return cf32{fmaf(-b.x, s, c.x), fmaf(-b.y, s, c.y)}; }
B2_HD cf32 c_fma_mi(cf32 b, float s, cf32 c) { return cf32{fmaf(b.y, s, c.x), fmaf(-b.x, s, c.y)}; }
B2_HD cf32 c_fma_pi(cf32 b, float s, cf32 c) { return cf32{fmaf(-b.y, s, c.x), fmaf(b.x, s, c.y)}; }
B2_HD cf32 c_scale_pi(cf32 b, float s) { return cf32{-b.y * s, b.x * s}; }
B2_HD cf32 c_2a_minus(cf32 a, cf32 y) { return cf32{fmaf(2.0f, a.x, -y.x), fmaf(2.0f, a.y, -y.y)}; }
#endif
// t a  and  u + t b  for a twiddle t held as two scalars: t a = t.x a + t.y (i a), two instructions each
B2_HD cf32 c_twmul(cf32 a, cf32 t) { return c_fma_pi(a, t.y, c_scale(a, t.x)); }
B2_HD cf32 c_twmul_add(cf32 b, cf32 t, cf32 u) { return c_fma(b, t.x, c_fma_pi(b, t.y, u)); }

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
        b = c_sub(a, t);
        a = c_add(a, t);
    } else if constexpr (K == 8) {          // w = -i : w b = (b.y, -b.x)
        const cf32 t = b;
        b = c_sub_mi(a, t);
        a = c_add_mi(a, t);
    } else {
        constexpr float c = W32c<K>::v, s = W32s<K>::v;
        // w b = c b + s (-i b) = (b.x c + b.y s,  b.y c - b.x s)
        const cf32 y0 = c_fma(b, c, c_fma_mi(b, s, a));
        b = c_2a_minus(a, y0);
        a = y0;
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
    const cf32 u = c_scale(v[i], wa);
    const cf32 b = v[i + 1];
    v[i] = c_fma(b, wb, u);
    v[i + 1] = c_fnma(b, wb, u);
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
    const cf32 u = c_scale(v[i], wa);
    const cf32 b = v[i + 1];
    v[i] = c_fma(b, wb, u);
    v[i + 1] = c_fnma(b, wb, u);
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

// ---- second transform: v[i] holds RAW B_l[q] with l = brev5(i).  On exit v[s] = X[q + 32 s].
// The W1024^(l q) twiddles are folded into the first stage: butterfly P pairs l_a = brev5(2 P) with
// l_b = l_a + 16, so  u = T_a a,  y0 = u + T_b b,  y1 = 2 u - y0  with T_b = T_a * W1024^(16 q).
// Only HALF the twiddles are read: the table holds W1024^(q l_a) (one 8-byte load per butterfly) and
// c16 = W1024^(16 q) is a per-lane constant; T_b costs two more packed instructions instead of a second load —
// the shared-memory pipe, not the FMA pipe, is this kernel's co-bottleneck (profiles/README.md).
template <int P, class TW1>
B2_HD void tw1_bfly(cf32 (&v)[32], const TW1 &tw1, const cf32 c16) {
    constexpr int i = 2 * P;
    cf32 u, tb;
    if constexpr (P == 0) {                     // la = 0: twiddles 1 and c16
        u = v[i]; tb = c16;
    } else {
        const cf32 ta = tw1(P);
        tb = c_twmul(ta, c16);
        u = c_twmul(v[i], ta);
    }
    const cf32 y0 = c_twmul_add(v[i + 1], tb, u);
    v[i + 1] = c_2a_minus(u, y0);
    v[i] = y0;
}
template <class TW1>
B2_HD void fft32_twiddled_half(cf32 (&v)[32], const TW1 &tw1, const cf32 c16) {
    tw1_bfly<0>(v, tw1, c16); tw1_bfly<1>(v, tw1, c16); tw1_bfly<2>(v, tw1, c16); tw1_bfly<3>(v, tw1, c16);
    tw1_bfly<4>(v, tw1, c16); tw1_bfly<5>(v, tw1, c16); tw1_bfly<6>(v, tw1, c16); tw1_bfly<7>(v, tw1, c16);
    tw1_bfly<8>(v, tw1, c16); tw1_bfly<9>(v, tw1, c16); tw1_bfly<10>(v, tw1, c16); tw1_bfly<11>(v, tw1, c16);
    tw1_bfly<12>(v, tw1, c16); tw1_bfly<13>(v, tw1, c16); tw1_bfly<14>(v, tw1, c16); tw1_bfly<15>(v, tw1, c16);
    dit_tail(v);
}

// (A/B builds only, pipe1024.cuh) Same transform with the twiddles fetched two at a time: tq[P] = (W1024^(q la), W1024^(q lb)) with
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
        u = c_twmul(v[i], cf32{t.ax, t.ay});
    }
    const cf32 y0 = c_twmul_add(v[i + 1], cf32{t.bx, t.by}, u);
    v[i + 1] = c_2a_minus(u, y0);
    v[i] = y0;
}
template <class TW2>
B2_HD void fft32_twiddled_pairs(cf32 (&v)[32], const TW2 &tw2) {
    tw2_bfly<0>(v, tw2); tw2_bfly<1>(v, tw2); tw2_bfly<2>(v, tw2); tw2_bfly<3>(v, tw2);
    tw2_bfly<4>(v, tw2); tw2_bfly<5>(v, tw2); tw2_bfly<6>(v, tw2); tw2_bfly<7>(v, tw2);
    tw2_bfly<8>(v, tw2); tw2_bfly<9>(v, tw2); tw2_bfly<10>(v, tw2); tw2_bfly<11>(v, tw2);
    tw2_bfly<12>(v, tw2); tw2_bfly<13>(v, tw2); tw2_bfly<14>(v, tw2); tw2_bfly<15>(v, tw2);
    dit_tail(v);
}


// tw(l) returns W1024^(l q): the form the tile kernel and the host harness use (same arithmetic, bit for bit)
template <class TW>
B2_HD void fft32_twiddled(cf32 (&v)[32], const TW &tw) {
    const cf32 c16 = tw(16);
    fft32_twiddled_half(v, [&tw](int P) { return tw(brev5(2 * P)); }, c16);
}

}  // namespace b200vec
