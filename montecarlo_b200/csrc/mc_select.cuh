// mc_select.cuh -- exact order statistics, tail sums and the fixed-bin histogram
// (launch interface of mc_select.cu).
#pragma once
#include "mc_common.cuh"

namespace mc {

struct SelectProblem;
constexpr int SIDE_WARPS = 8;             // warps of a pass block (FAST_THREADS / 32)
struct SideLists {
    float* values;            // [rows][chunks][SIDE_WARPS][cap]
    unsigned int* counts;     // [rows][chunks][SIDE_WARPS]   (> cap: that list overflowed)
    uint32_t* brackets;       // [rows][1 + 2 * MAX_IV]: n, lo keys, hi keys of the pass-1 brackets
    unsigned int cap;         // floats per warp list
    unsigned int chunks;      // the grid.x the lists were laid out for
};
// chunks (grid.x) of the fp32 pass kernels for this problem, and the side-list geometry that goes with it
unsigned long long select_chunks_f32(const SelectProblem& sp, int sm_count);
unsigned int side_list_cap(unsigned long long n, unsigned long long chunks);
inline size_t side_list_floats(unsigned long long rows, unsigned long long chunks, unsigned int cap) {
    return (size_t)rows * chunks * SIDE_WARPS * cap;
}
constexpr int SIDE_BRACKET_WORDS = 1 + 2 * MAX_IV;

// One selection problem: `rows` independent rows of `n` local values each
// (row r starts at x + r*ld).  Terminal percentiles: rows = 1.  Fan chart: rows = T+1.
struct SelectProblem {
    const void* x;
    int dtype;
    unsigned long long rows, n, ld;
    unsigned int m_slot;      // pilot-sample slots per row in this rank's stage-0 buffer
    unsigned int m_local;     // pilot samples this rank contributes per row (<= m_slot, <= n)
    int sample_contiguous;    // 1: first m_local values of the row (iid columns), 0: strided
    int pilot_in_place;       // 1 (single rank): the bracket kernel reads the pilot straight from the rows -- nothing is
                              //   copied into a stage buffer (m_slot = 0), and the pilot can be four times larger
    SelectPlan plan;          // host copy
    SelectPlan* d_plan;       // device copy
    RowSel* d_states;         // [rows]
    ScanScratch* d_scratch;   // [rows]
    // fp32 only (null: off).  Pass 1 keeps every value it finds inside a bracket -- a few percent of the row --
    // in a side list per (row, chunk, warp); later passes read the side lists instead of streaming the rows
    // again (the intervals they narrow lie inside the brackets; a block whose list overflowed, or whose row has
    // an interval outside the brackets, streams its chunk of the row as before).
    SideLists side;
};

// Build the plan for np.percentile(method='linear') over n_global values.
// Returns false on invalid input (pcts outside [0,100], not ascending, n_global == 0).
bool build_select_plan(const double* pcts, int n_pcts, unsigned long long n_global, int dtype,
                       SelectPlan* plan);

// ---- stage buffer geometry (uint32 units unless noted) ------------------------------
inline int key_units(int dtype) { return dtype == MC_F64 ? 2 : 1; }                  // uint32 per key
inline size_t iv_slot_u32(int dtype) { return 2 + (size_t)SEL_CAND * key_units(dtype); }
inline size_t row_slot_u32(int dtype, int n_targets) {                                // per row, per pass
    return (size_t)n_targets * (2 + iv_slot_u32(dtype));
}
inline size_t pass_words(const SelectProblem& sp) {                                   // 8-byte words
    return (size_t)sp.rows * row_slot_u32(sp.dtype, sp.plan.n_targets) / 2;
}
inline size_t sample_words(const SelectProblem& sp) {                                 // 8-byte words
    return ((size_t)sp.rows * sp.m_slot * key_units(sp.dtype) + 1) / 2;
}
// pilot-sample slots per row for one rank of n_ranks
unsigned int sample_slot(int dtype, unsigned int n_ranks);
// pilot size of a single-rank problem read in place; 0 restores the default
unsigned int pilot_in_place_size(bool contiguous);
void set_pilot_in_place_size(unsigned int m);

// copy this rank's pilot sample (as keys) into its stage-0 buffer
cudaError_t launch_select_sample(const SelectProblem& sp, void* d_sample_out, cudaStream_t stream);
// sort the pilot samples of all ranks (rank r's block at d_samples + r*rank_stride_words) and
// derive the starting intervals of every row; m_local of rank r is header word `hdr_word`
cudaError_t launch_select_brackets(const SelectProblem& sp, const unsigned long long* d_gathered,
                                   size_t sample_off_words, int hdr_word, unsigned int n_ranks,
                                   size_t rank_stride_words, int max_smem_optin, cudaStream_t stream);
// one narrowing pass over the local values: per row and interval, counts / histogram /
// collected candidates into d_out (layout row_slot_u32; zeroed by the callee)
cudaError_t launch_select_pass(const SelectProblem& sp, bool first_pass, void* d_out, int sm_count,
                               cudaStream_t stream, bool zero_out = true);
// adds to *d_blocks the number of (row, chunk) blocks of the next fp32 pass that cannot work from their side
// lists (they would stream the rows); adds rows * chunks when the problem has no side lists at all
cudaError_t launch_side_check(const SelectProblem& sp, int sm_count, unsigned int* d_blocks, cudaStream_t stream);
// merge the pass output of all ranks and shrink every target's interval; adds the number of
// rows that still have undecided targets to *d_pending
cudaError_t launch_select_scan(const SelectProblem& sp, const void* d_gathered_pass, unsigned int n_ranks,
                               size_t rank_stride_words, unsigned int* d_pending, cudaStream_t stream);
// keys -> values, np.percentile's _lerp; d_out[rows][n_pcts] (and lo/hi order statistics if non-null)
cudaError_t launch_select_finish(const SelectProblem& sp, double* d_out, double* d_lo, double* d_hi,
                                 cudaStream_t stream);

// moments: merge the stage-0 headers of all ranks into the device result
cudaError_t launch_merge_moments(const unsigned long long* d_gathered, unsigned int n_ranks,
                                 size_t rank_stride_words, double center, DeviceResult* d_res,
                                 cudaStream_t stream);
// copies pct/lo/hi of the terminal problem into the result and sets the tail threshold
cudaError_t launch_publish_percentiles(const double* d_pct, const double* d_lo, const double* d_hi,
                                       int n_pcts, int tail_idx, DeviceResult* d_res, cudaStream_t stream);

// ONE pass over the terminal values for both closing reductions: count and sum of
// x <= res->tail_threshold -> words[0..1] (deterministic), and np.histogram(x, bins, range)
// -> words[2..2+bins).  Edges from (lo,hi) or, if lo >= hi, from res->(vmin, vmax).
cudaError_t launch_tail_hist(const void* x, int dtype, unsigned long long n, bool want_tail, unsigned int bins,
                             double lo, double hi, DeviceResult* d_res, double* d_edges /*[bins+1]*/,
                             double* d_block_scratch /*[2*1024]*/, unsigned long long* d_words, int sm_count,
                             cudaStream_t stream);
// final merge: tail (count,sum) and histogram over ranks
cudaError_t launch_merge_final(const unsigned long long* d_gathered, unsigned int n_ranks,
                               size_t rank_stride_words, unsigned int bins, DeviceResult* d_res,
                               unsigned long long* d_hist_total, cudaStream_t stream);

}  // namespace mc
