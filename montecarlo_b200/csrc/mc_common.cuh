// mc_common.cuh -- shared device-side types of the simulator, the reductions and
// the radix select.  Host/device plain structs only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/mc_b200.h"

namespace mc {

constexpr int SIM_BLOCK = 256;
constexpr int MAX_TARGETS = 2 * MC_MAX_PCTS;    // two order statistics per percentile
constexpr int MAX_IV = MAX_TARGETS;             // key intervals tracked per row
constexpr int SEL_BIN_BITS = 10;
constexpr int SEL_BINS = 1 << SEL_BIN_BITS;     // sub-bins per interval per narrowing pass
constexpr int SEL_CAND = 1024;                  // an interval holding <= this many values is collected and sorted
constexpr int SEL_SAMPLE_MAX = 32768;           // pilot sample per row (all ranks together), fp32 keys
constexpr int SEL_SAMPLE_MAX_F64 = 16384;       // ... fp64 keys (sorted in shared memory)
constexpr int SEL_MAX_PASSES = 12;              // hard stop of the narrowing loop (never reached: 64 bits / 10 + miss)

// ---- order-preserving integer keys ------------------------------------------------
// key(a) < key(b)  <=>  a < b  for all non-NaN floats (negative portfolio values included);
// -0.0 is folded onto +0.0 (v + 0.0 in round-to-nearest) so that key order and IEEE comparison
// agree everywhere -- the fp32 selection passes compare floats, the scans compare keys.
constexpr uint32_t KEY32_NEG_INF = 0x007FFFFFu;   // key_of(-inf)
constexpr uint32_t KEY32_POS_INF = 0xFF800000u;   // key_of(+inf)
__host__ __device__ __forceinline__ uint32_t key_of(float v) {
    v = v + 0.0f;
#ifdef __CUDA_ARCH__
    const uint32_t b = __float_as_uint(v);
#else
    union { float f; uint32_t u; } c; c.f = v; const uint32_t b = c.u;
#endif
    return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__host__ __device__ __forceinline__ uint64_t key_of(double v) {
    v = v + 0.0;
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
// [0] n  [1] n_above  [2] sum_d  [3] sum_d2  [4] min  [5] max  [6] unused  [7] unused
// then the pilot samples: terminal problem, then the rows of the fan-chart problem
constexpr int STAGE0_HEADER = 8;

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
    unsigned long long status;   // fused closing (mc_fused.cuh): 0 ok, else the reductions must be redone by the staged selection
};

// numpy/lib/_function_base_impl.py:_lerp -- explicit _rn ops: no FMA contraction, bit-identical to np.percentile
__device__ __forceinline__ double numpy_lerp(double a, double b, double t) {
    const double diff = __dsub_rn(b, a);
    double r = __dadd_rn(a, __dmul_rn(diff, t));
    if (t >= 0.5) r = __dsub_rn(b, __dmul_rn(diff, __dsub_rn(1.0, t)));
    return r;
}

// ---- exact selection by interval narrowing ------------------------------------------
// One SelectPlan per problem (terminal values, or the rows of the stored paths).
struct SelectPlan {
    int n_targets;                        // sorted, distinct global ranks
    int n_pcts;
    int key_bits;
    int pad;
    unsigned long long n_global;          // values per row over all ranks
    unsigned long long rank[MAX_TARGETS]; // global 0-based order-statistic ranks
    int pct_lo_idx[MC_MAX_PCTS];          // target index of floor(h) and floor(h)+1
    int pct_hi_idx[MC_MAX_PCTS];
    double pct_gamma[MC_MAX_PCTS];
};

enum { IV_HIST = 0, IV_COLLECT = 1 };

// per-row state: a sorted list of disjoint key intervals [lo, hi], each known to hold some
// of the row's target ranks.  A pass histograms (or, for small populations, collects) the
// values inside each interval; the scan then shrinks every target's interval.
struct RowSel {
    int n_iv;                                   // 0: every target of this row is decided
    int below_known;                            // 0 before the first pass (counts below lo come from it)
    unsigned long long lo[MAX_IV], hi[MAX_IV];  // inclusive key bounds
    unsigned long long below[MAX_IV];           // # values (all ranks) with key < lo
    int shift[MAX_IV];                          // bin = (key - lo) >> shift  (< SEL_BINS)
    int mode[MAX_IV];                           // IV_HIST / IV_COLLECT
    int target_iv[MAX_TARGETS];                 // interval of target j, -1 when decided
    unsigned long long key[MAX_TARGETS];        // decided key of target j
};

// what the per-interval scan blocks hand to the per-row assembly (mc_select.cu)
struct ScanScratch {
    unsigned long long below[MAX_IV], total[MAX_IV];
    unsigned long long nlo[MAX_TARGETS], nhi[MAX_TARGETS], nbelow[MAX_TARGETS], npop[MAX_TARGETS];
    int code[MAX_TARGETS];
};

}  // namespace mc
