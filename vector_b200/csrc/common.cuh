// common.cuh — shared declarations for libb200vec (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>
#include <map>

#include "../../include/b200vec.h"
#include "maxspans.cuh"

namespace b200vec {

constexpr int kSmCount = 148;          // B200: 2 dies x 74 SMs; grids are sized in multiples of this

extern thread_local std::string g_last_error;

inline int fail(const char *where, const char *msg) {
    g_last_error = std::string(where) + ": " + msg;
    return 1;
}

#define B2_CUDA(call)                                                                   \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            return ::b200vec::fail(#call, cudaGetErrorString(e__));                     \
        }                                                                               \
    } while (0)

// A/B switches (environment variables selecting measured-slower kernel variants) exist only in
// -DB2_AB builds (make ab); the product library has one code path per op and reads no environment.
#ifdef B2_AB
inline const char *ab_env(const char *name) { return getenv(name); }
#else
inline const char *ab_env(const char *) { return nullptr; }
#endif

#define B2_REQUIRE(cond, msg)                                                           \
    do {                                                                                \
        if (!(cond)) return ::b200vec::fail(__func__, msg);                             \
    } while (0)

// A committed packet lives twice in HBM/L2 so that the placement kernel can always issue
// 16-byte loads: copy A has y[0] on a 16-byte boundary (used when the first sample of a
// float4 pair has an even packet index), copy B has y[0] 8 bytes past one (odd index).
// Both copies are surrounded by >= kBankPad zero samples, so a pair that straddles either
// end of the packet reads an exact 0 for the outside sample.
constexpr int kBankPad = 4;

struct BankSlot {
    float2 *raw = nullptr;      // uploaded / resampled samples, len_raw
    int64_t len_raw = 0, cap_raw = 0;
    float2 *store = nullptr;    // backing store of copies A and B
    int64_t cap_store = 0;      // capacity in samples PER COPY (excluding pads)
    float2 *a = nullptr;        // -> y[0] of copy A (16-byte aligned)
    float2 *b = nullptr;        // -> y[0] of copy B (8 mod 16)
    int64_t len = 0;            // committed length (0 = not committed)
    int64_t per = 0;            // samples per copy incl. pads; the pre-scaled copies live 2 * per after the raw ones
    bool external = false;      // `a` is caller-owned device memory (b200vec_bank_attach): no copy B, no pads
};

struct Scratch {
    void *ptr = nullptr;
    size_t bytes = 0;
};

// Uploaded must-evaluate span list of the max pass (maxspans.cuh), reused while the planner's inputs do not change
struct SpanCache {
    Scratch dev;
    void *pin = nullptr;        // pinned staging block of the upload
    size_t pin_bytes = 0;
    cudaEvent_t pin_free = nullptr;     // recorded after the upload that reads `pin`
    unsigned long long key = 0;
    long long n = -1;           // -1: nothing uploaded
};

// State of a prepared scenario batch (b200vec_batch_prepare -> b200vec_batch_build, span.cuh)
struct BatchState {
    Scratch bank;               // shifted packets, one copy each, 16-byte aligned starts
    Scratch tables;             // PlaceDesc[n_vec * pk_stride] | int np[n_vec] | ShiftJob[n_jobs] | u64 pmax[n_jobs]
    void *pin = nullptr;        // pinned staging of the tables
    size_t pin_bytes = 0;
    const void *descs = nullptr;
    const int *np = nullptr;
    int n_vec = 0, pk_stride = 0, n_jobs = 0;
    std::vector<MaxSpanPacket> h_pk;    // host copy of the placement descriptors, [n_vec][pk_stride] (max-span planner)
    std::vector<int> h_np;
    SpanCache spans;                    // per-vector must-evaluate spans for the `total` of the last normalised build
};

// Top-histogram state kept between b200vec_select_begin and b200vec_select_finish
struct TopState {
    std::vector<unsigned int> hist;     // host copy of the 13-bit histogram (+0 excluded)
    unsigned long long n_poszero = 0;
    const void *data = nullptr;
    long long n = -1;
    int use_abs = 0;
};

}  // namespace b200vec

struct b200vec_ctx {
    int device = 0;
    b200vec::BankSlot slots[B200VEC_MAX_SLOTS];
    std::map<int, float2 *> twiddles;     // nfft -> exp(-2*pi*i*k/nfft), k < nfft/2 (float2)
    b200vec::Scratch fft64[4];            // fp64 work buffers for the Bluestein resampler
    unsigned int *hist = nullptr;         // select scratch (see select.cuh)
    void *pinned = nullptr;               // small pinned staging block
    float2 *zeros = nullptr;              // one span of zero samples (bulk-copy source, fused1024.cuh)
    unsigned int *wf_ctr = nullptr;       // work-item counter of the warp-per-frame kernel (wf1024.cuh)
    unsigned int *sp_ctr = nullptr;       // work-item counter / grid barrier of the span kernel (span.cuh)
    b200vec::BatchState batch;            // prepared scenario batch
    b200vec::SpanCache max_spans;         // must-evaluate spans of the last normalised single-vector build
    b200vec::Scratch sel64;               // state + histograms of the float64 select (packet.cuh)
    b200vec::Scratch k5;                  // K5Plan + K5Out of b200vec_normalize_db (select.cuh)
    unsigned int *top_hist = nullptr;     // [kTopBins] u32 + one u64 (+0 count) + one u64 (candidate count)
    b200vec::Scratch cand;                // candidate array of the fast select
    b200vec::TopState top;
    unsigned long long *slot_max = nullptr;   // per bank slot: bits of max |y|^2 (double) of the committed packet
    int sm_count = b200vec::kSmCount;
};
