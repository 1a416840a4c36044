// mc_common.cuh -- shared device-side types of the simulator, the reductions and
// the radix select.  Host/device plain structs only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mc_b200.h"

namespace mc {

constexpr int SIM_BLOCK = 256;
constexpr int DIGIT0_BITS = 12;                 // first radix digit, histogrammed inside the sim kernel
constexpr int DIGIT0_BINS = 1 << DIGIT0_BITS;
constexpr int DIGIT_BITS = 10;                  // later digits
constexpr int DIGIT_BINS = 1 << DIGIT_BITS;
constexpr int MAX_TARGETS = 2 * MC_MAX_PCTS;    // two order statistics per percentile
constexpr int MAX_PASSES = 8;

// ---- order-preserving integer keys ------------------------------------------------
// key(a) < key(b)  <=>  a < b  for all non-NaN floats (negative portfolio values included)
__host__ __device__ __forceinline__ uint32_t key_of(float v) {
#ifdef __CUDA_ARCH__
    const uint32_t b = __float_as_uint(v);
#else
    union { float f; uint32_t u; } c; c.f = v; const uint32_t b = c.u;
#endif
    return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__host__ __device__ __forceinline__ uint64_t key_of(double v) {
#ifdef __CUDA_ARCH__
    const uint64_t b = (uint64_t)__double_as_longlong(v);
#else
    union { double f; uint64_t u; } c; c.f = v; const uint64_t b = c.u;
#endif
    return b ^ ((b >> 63) ? 0xFFFFFFFFFFFFFFFFull : 0x8000000000000000ull);
}
__host__ __device__ __forceinline__ float value_of(uint32_t k) {
    const uint32_t b = k ^ ((k >> 31) ? 0x80000000u : 0xFFFFFFFFu);
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    union { float f; uint32_t u; } c; c.u = b; return c.f;
#endif
}
__host__ __device__ __forceinline__ double value_of(uint64_t k) {
    const uint64_t b = k ^ ((k >> 63) ? 0x8000000000000000ull : 0xFFFFFFFFFFFFFFFFull);
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long)b);
#else
    union { double f; uint64_t u; } c; c.u = b; return c.f;
#endif
}

template <typename Real> struct KeyT;
template <> struct KeyT<float> { using type = uint32_t; static constexpr int bits = 32; };
template <> struct KeyT<double> { using type = uint64_t; static constexpr int bits = 64; };

// ---- per-block partial moments (written without atomics, reduced in fixed order) ----
struct BlockPartial {
    double sum_d;      // sum (x - s0)
    double sum_d2;     // sum (x - s0)^2
    double vmin, vmax;
    unsigned long long n;
    unsigned long long n_above;
};

// ---- stage-0 packed words (per rank) --------------------------------------------
// [0] n  [1] n_above  [2] sum_d  [3] sum_d2  [4] min  [5] max  [6..] digit-0 histogram
constexpr int STAGE0_HEADER = 6;

// ---- merged result block kept on the device, copied out by mc_fetch_result --------
struct DeviceResult {
    unsigned long long n, n_above;
    double mean, std, vmin, vmax;
    double pct[MC_MAX_PCTS], pct_lo[MC_MAX_PCTS], pct_hi[MC_MAX_PCTS];
    double tail_threshold;
    unsigned long long tail_count;
    double tail_sum;
    double hist_lo, hist_hi;
    double center;        // s0: moments are accumulated about it
};

// ---- radix-select bookkeeping -------------------------------------------------------
// One SelectPlan per problem (terminal values, or the rows of the stored paths).
struct SelectPlan {
    int n_targets;                        // sorted, distinct global ranks
    int n_pcts;
    int n_passes;
    int key_bits;
    int pass_bits[MAX_PASSES];
    int pass_shift[MAX_PASSES];
    unsigned long long rank[MAX_TARGETS]; // global 0-based order-statistic ranks
    int pct_lo_idx[MC_MAX_PCTS];          // target index of floor(h) and floor(h)+1
    int pct_hi_idx[MC_MAX_PCTS];
    double pct_gamma[MC_MAX_PCTS];
};

// per-row mutable state of the select
struct RowState {
    unsigned long long prefix[MAX_TARGETS];     // key bits decided so far (right-aligned)
    unsigned long long remaining[MAX_TARGETS];  // rank within the elements sharing the prefix
    int group[MAX_TARGETS];                     // index into group_prefix
    int n_groups;
    int pad;
    unsigned long long group_prefix[MAX_TARGETS];
};

}  // namespace mc
