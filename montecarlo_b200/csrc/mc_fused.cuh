// mc_fused.cuh -- the lean closing of a run without a fan chart: every reduction of app.py:636-656 on the
// terminal values in FOUR launches, and -- when paths are sharded over GPUs -- with two exchanges that the
// kernels themselves carry out through peer-mapped memory (no host round trip, no NCCL launch):
//
//   K1  simulation.  sim32_kernel<..., SINK> bins every terminal value it produces into a KEY HISTOGRAM kept in
//       shared memory (8192 bins over a window of order-preserving keys chosen from the model's own scale) and,
//       for <= 2048 bins, into the caller's fixed-bin np.histogram; one shared atomic per PATH, flushed once
//       per persistent block.  Other simulation kernels are followed by K1b, a pass over their terminals.
//   X1  [ranks > 1] each rank pushes its key histogram (32 KB) into every peer's inbox and raises a flag there.
//   K3  plan: sum of the histograms, prefix sums, the bin of every order statistic np.percentile needs
//       (method='linear': ranks floor(h), floor(h) + 1), exact populations below each such bin.
//   K4  ONE pass over the terminal values: moments about s0, exceedance count, min / max, tail sums below the
//       CVaR bin, the fixed-bin histogram when K1 did not take it, and the values of the target bins
//       ("candidates", ~1 % of the data) appended to per-bin lists in this rank's buffer.  The last block pushes
//       the rank's header (moments, tail sums, list lengths) and fixed-bin counts to every peer and raises the
//       flags of X2.
//   K5  one block per target bin reads that bin's candidates from EVERY rank's list -- the peers' over NVLink --
//       into shared memory and selects its order statistics there (4096-bin sub-histogram, then a bitonic sort of
//       <= 1024 keys); the last block to finish applies NumPy's _lerp, merges moments, tail and histogram over
//       ranks in rank order and writes the result to the device AND to pinned host memory.
//
// Every rank computes the same plan from the same summed histogram, so all ranks end with identical results.
// A window that missed, a candidate list that overflowed or a peer that never arrived set a status the host
// reads with the result; the reductions are then redone by the general staged selection (mc_select.cu) on the
// terminal values already there (MC_FLAG_TERMINAL_GIVEN) -- slower, never wrong.
#pragma once
#include "mc_common.cuh"

namespace mc {

constexpr int FUSED_KB = 8192;                  // key-histogram bins; 0 and KB-1 are open-ended
constexpr int FUSED_MAX_SLOTS = MAX_TARGETS;    // distinct target bins
constexpr int FUSED_MAX_RANKS = 16;
constexpr int FUSED_SINK_NPH_BINS = 2048;       // fixed-bin histogram taken inside sim32_kernel up to this many bins
constexpr int FUSED_MAX_NPH_BINS = 16384;       // ... and by the fused closing at all (more: general path)
constexpr int FUSED_HDR_WORDS = 64;
constexpr int FUSED_SORT_CAP = 4096;

enum { FUSED_OK = 0, FUSED_FALLBACK = 1, FUSED_PEER_TIMEOUT = 2, FUSED_INCONSISTENT = 3 };

// 32-bit order-preserving key of a terminal value (fp64: the upper half of the 64-bit key) -> histogram bin:
//     bin(k) = floor(clamp(k, a, b) - a) * mult / 2^32)
// four integer instructions (two VIMNMX, IADD, IMAD.HI), monotone non-decreasing in the value.  Keys below a share
// bin 0 with the first keys of the window, keys above b the last bin: the end bins are open-ended.
struct KeyWindow { uint32_t a, b, mult; };

__host__ __device__ __forceinline__ uint32_t key32_of(float v) { return key_of(v); }
__host__ __device__ __forceinline__ uint32_t key32_of(double v) { return (uint32_t)(key_of(v) >> 32); }

__host__ __device__ __forceinline__ uint32_t window_bin(const KeyWindow& w, uint32_t k32) {
    const uint32_t c = k32 < w.a ? w.a : (k32 > w.b ? w.b : k32);
#ifdef __CUDA_ARCH__
    return __umulhi(c - w.a, w.mult);
#else
    return (uint32_t)(((unsigned long long)(c - w.a) * w.mult) >> 32);
#endif
}
// a window over [klo, khi] (32-bit keys) with all FUSED_KB bins in use
inline KeyWindow make_window(uint32_t klo, uint32_t khi) {
    KeyWindow w;
    if (khi < klo) khi = klo;
    w.a = klo ? klo - 1u : 0u;
    w.b = khi < 0xFFFFFFFFu ? khi + 1u : khi;
    const unsigned long long span = (unsigned long long)w.b - w.a;
    const unsigned long long m = span ? (((unsigned long long)FUSED_KB << 32) - 1ull) / span : 0xFFFFFFFFull;
    w.mult = (uint32_t)(m > 0xFFFFFFFFull ? 0xFFFFFFFFull : m);
    return w;
}

// np.histogram(x, bins, range=(lo, hi)) bin of one value (numpy/lib/_histograms_impl.py, uniform-bin path), or -1
// when the value lies outside [lo, hi].  `edge` = np.linspace(lo, hi, bins + 1) -- for fp32 data rounded so that
// every comparison with the float64 edge is exact in fp32 (hist_edges_kernel, mc_select.cu).
template <typename Real>
__device__ __forceinline__ int np_hist_bin(Real v, const Real* __restrict__ edge, int bins, Real lo, Real top, Real inv_w) {
    if (!(v >= lo && v <= top)) return -1;
    const int last = bins - 1;
    int idx = (int)((v - lo) * inv_w);
    idx = idx < 0 ? 0 : (idx > last ? last : idx);
    while (idx > 0 && v < edge[idx]) --idx;
    while (idx < last && v >= edge[idx + 1]) ++idx;
    return idx;
}

// what sim32_kernel<..., SINK> needs beside its own arguments
struct SinkArgs {
    KeyWindow win;
    unsigned int* g_khist;          // [FUSED_KB] this rank's key histogram (zeroed before the launch)
    unsigned int* g_nphist;         // [nph_bins] this rank's fixed-bin counts (zeroed), or null: K4 takes the histogram
    const float* nph_edges;         // [nph_bins + 1]
    unsigned int nph_bins;
    unsigned int claim_paths;       // paths a warp claims at a time (multiple of 32)
    unsigned long long* claim;      // next unclaimed local path id (zeroed before the launch)
    unsigned int smem_off;          // byte offset of the sink's shared memory behind the kernel's own table
};
inline size_t sink_smem_bytes(unsigned int nph_bins) {
    return (size_t)FUSED_KB * 4 + (nph_bins ? (size_t)nph_bins * 4 + ((size_t)nph_bins + 1) * 4 : 0);
}

// ---- the plan K3 writes and K4 / K5 read (device memory) ----------------------------------------------
struct FusedPlan {
    int status;
    int n_slots;
    int tail_slot_lo, tail_slot_hi;           // slots of the two order statistics of the CVaR percentile (-1: none)
    unsigned int tail_bin;                    // every value in a bin below this one is <= the CVaR threshold
    unsigned int overflow;                    // this rank: a candidate list overflowed (K4)
    unsigned int ticket4, ticket5;
    unsigned long long tail_below;            // global count of the values in bins < tail_bin
    unsigned int bin[FUSED_MAX_SLOTS];
    unsigned int cap[FUSED_MAX_SLOTS];        // per-rank capacity (entries) of the slot's candidate list
    unsigned int off[FUSED_MAX_SLOTS];        // per-rank offset (entries) of the slot's list
    unsigned int cand_count[FUSED_MAX_SLOTS]; // this rank's candidates per slot (K4)
    unsigned long long below[FUSED_MAX_SLOTS];
    unsigned long long pop[FUSED_MAX_SLOTS];
    int target_slot[MAX_TARGETS];
    unsigned long long key[MAX_TARGETS];      // decided keys (K5)
    // CVaR (K5): candidates at or below the lower order statistic of the tail percentile, and how often the upper one occurs
    unsigned long long tail_lo_cnt, tail_hi_mult;
    double tail_lo_sum;
};

// ---- the exchange buffer: one per handle, identical layout on every rank ---------------------------------
// rank r's contributions land in slot r of every section of every rank's buffer (its own included)
struct XchgLayout {
    unsigned int n_ranks;
    unsigned int cand_bytes;                  // of this rank's candidate lists
    size_t off_flags1, off_flags2, off_hist, off_hdr, off_nph, off_cand, total;
};
inline XchgLayout xchg_layout(unsigned int n_ranks, unsigned int cand_bytes) {
    XchgLayout L{};
    L.n_ranks = n_ranks;
    L.cand_bytes = cand_bytes;
    size_t o = 0;
    L.off_flags1 = o; o += FUSED_MAX_RANKS * 8;
    L.off_flags2 = o; o += FUSED_MAX_RANKS * 8;
    o = (o + 255) & ~(size_t)255;
    L.off_hist = o; o += (size_t)n_ranks * FUSED_KB * 4;
    L.off_hdr = o; o += (size_t)n_ranks * FUSED_HDR_WORDS * 8;
    L.off_nph = o; o += (size_t)n_ranks * FUSED_MAX_NPH_BINS * 4;
    L.off_cand = o; o += (size_t)cand_bytes;                 // this rank's own candidate lists; peers read them in place
    L.total = o;
    return L;
}

struct Peers {
    unsigned char* base[FUSED_MAX_RANKS];     // exchange buffer of rank r as mapped into THIS process (own: local)
    unsigned int n_ranks, me;
    unsigned long long seq;                   // this run's sequence number (flags carry it)
};

struct ClosePartial {
    double sum_d, sum_d2, vmin, vmax, tail_sum;
    unsigned long long n, n_above, tail_cnt;
};

struct FusedArgs {
    const void* x;                            // terminal values of this rank
    int dtype;
    unsigned long long n;                     // ... how many
    double center;                            // s0
    KeyWindow win;
    XchgLayout lay;
    Peers peers;
    const SelectPlan* d_plan;                 // ranks of the order statistics (global)
    int n_targets, n_pcts, tail_idx;          // host copies
    FusedPlan* plan;                          // device
    unsigned char* member;                    // device [FUSED_KB]: slot + 1 of a target bin, else 0
    ClosePartial* partials;                   // device [>= grid of K4]
    unsigned int nph_bins;                    // fixed-bin histogram (0: none)
    int nph_in_sim;                           // 1: K1 already filled this rank's counts
    const double* edges;                      // [bins + 1] fp64, then [bins + 1] fp32 (hist_edges_kernel)
    double hist_lo, hist_hi;
    DeviceResult* res;                        // device result block
    unsigned long long* hist_total;           // device [bins]
    unsigned char* host_out;                  // pinned host: DeviceResult, then bins x u64
    unsigned long long* claim;                // the simulation kernel's path-claim counter (K5 leaves it at zero)
};

cudaError_t launch_terminal_hist(const FusedArgs& a, int sm_count, cudaStream_t s);      // K1b
cudaError_t launch_fused_plan(const FusedArgs& a, cudaStream_t s);                       // X1 + K3
cudaError_t launch_fused_closing(const FusedArgs& a, int sm_count, int* grid, cudaStream_t s);   // K4 + X2
cudaError_t launch_fused_select(const FusedArgs& a, cudaStream_t s);                     // K5
// BlockPartial per block from terminal values already in memory (MC_FLAG_TERMINAL_GIVEN: the staged selection
// without a simulation in front of it); returns the number of partials written
cudaError_t launch_terminal_partials(const void* x, int dtype, unsigned long long n, double center, BlockPartial* out,
                                     int sm_count, int* n_partials, cudaStream_t s);
// fp64 and (ceil-rounded) fp32 edges of np.histogram's uniform bins; writes res->hist_lo/hi when res is given
cudaError_t launch_hist_edges(unsigned int bins, double lo, double hi, DeviceResult* d_res, double* d_edges, cudaStream_t s);

}  // namespace mc
