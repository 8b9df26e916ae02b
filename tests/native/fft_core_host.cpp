// Host harness for vector_b200/csrc/fft32.cuh: runs the kernels' exact 1024-point transform
// (same header, same operation order, fmaf = FFMA) lane by lane on the CPU.  Built by
// tests/test_fft_core.py with g++; test infrastructure only.
#include "../../vector_b200/csrc/fft32.cuh"
#include <vector>
using namespace b200vec;

extern "C" int b2host_stft1024(const float *x /* interleaved re,im; n_frames*hop+1024-hop samples */,
                               int hop, int n_frames, const float *window /*1024*/, float sqrt_scale,
                               float *psd /* [n_frames][1024], natural bin order */) {
    std::vector<cf32> tw(1024);
    for (int q = 0; q < 32; ++q)
        for (int l = 0; l < 32; ++l) {
            const double ang = -2.0 * M_PI * (double)((q * l) % 1024) / 1024.0;
            tw[q * 32 + l] = cf32{(float)cos(ang), (float)sin(ang)};
        }
    std::vector<cf32> B(1024);
    for (int f = 0; f < n_frames; ++f) {
        const float *xf = x + 2 * (size_t)f * hop;
        for (int l = 0; l < 32; ++l) {
            cf32 v[32];
            float win[32];
            for (int j = 0; j < 32; ++j) win[j] = window[l + 32 * j] * sqrt_scale;
            for (int i = 0; i < 32; ++i) {
                const int n = l + 32 * brev5(i);
                v[i] = cf32{xf[2 * n], xf[2 * n + 1]};
            }
            fft32_windowed(v, win);
            for (int q = 0; q < 32; ++q) B[l * 32 + q] = v[q];
        }
        for (int q = 0; q < 32; ++q) {
            cf32 v[32];
            for (int i = 0; i < 32; ++i) v[i] = B[brev5(i) * 32 + q];
            const cf32 *t = &tw[q * 32];
            fft32_twiddled(v, [t](int l) { return t[l]; });
            for (int s = 0; s < 32; ++s)
                psd[(size_t)f * 1024 + q + 32 * s] = b2_fma(v[s].x, v[s].x, v[s].y * v[s].y);
        }
    }
    return 0;
}
