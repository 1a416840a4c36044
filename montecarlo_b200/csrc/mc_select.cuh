// mc_select.cuh -- exact order statistics, tail sums and the fixed-bin histogram
// (launch interface of mc_select.cu).
#pragma once
#include "mc_common.cuh"

namespace mc {

// One selection problem: `rows` independent rows of `n` local values each
// (row r starts at x + r*ld).  Terminal percentiles: rows = 1.  Fan chart: rows = T+1.
struct SelectProblem {
    const void* x;
    int dtype;
    unsigned long long rows, n, ld;
    SelectPlan plan;          // host copy
    SelectPlan* d_plan;       // device copy
    RowState* d_states;       // [rows]
};

// Build the plan for np.percentile(method='linear') over n_global values.
// Returns false on invalid input (pcts outside [0,100], not ascending, n_global == 0).
bool build_select_plan(const double* pcts, int n_pcts, unsigned long long n_global, int dtype,
                       SelectPlan* plan);

inline size_t hist_words_pass0(unsigned long long rows) { return (size_t)rows * DIGIT0_BINS / 2; }
inline size_t hist_words_pass(unsigned long long rows, int n_targets) {
    return (size_t)rows * n_targets * DIGIT_BINS / 2;
}

// pass-p digit histogram of the rows' local values into d_hist (uint32, zeroed by callee)
cudaError_t launch_select_hist(const SelectProblem& sp, int pass, unsigned int* d_hist, int sm_count,
                               cudaStream_t stream);
// consume the pass-p histograms of all ranks (rank r at d_hist + r*rank_stride_u32) and
// narrow every target by one digit
cudaError_t launch_select_scan(const SelectProblem& sp, int pass, const unsigned int* d_hist,
                               unsigned int n_ranks, size_t rank_stride_u32, cudaStream_t stream);
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

// count and sum of x <= res->tail_threshold -> words[0..1] (count, sum bits); deterministic
cudaError_t launch_tail(const void* x, int dtype, unsigned long long n, const DeviceResult* d_res,
                        double* d_block_scratch /*[2*1024]*/, unsigned long long* d_words, int sm_count,
                        cudaStream_t stream);
// np.histogram(x, bins, range): edges from (lo,hi) or, if lo >= hi, from res->(vmin, vmax)
cudaError_t launch_value_hist(const void* x, int dtype, unsigned long long n, unsigned int bins, double lo,
                              double hi, DeviceResult* d_res, double* d_edges /*[bins+1]*/,
                              unsigned long long* d_counts /*[bins], zeroed by callee*/, int sm_count,
                              cudaStream_t stream);
// final merge: tail (count,sum) and histogram over ranks
cudaError_t launch_merge_final(const unsigned long long* d_gathered, unsigned int n_ranks,
                               size_t rank_stride_words, unsigned int bins, DeviceResult* d_res,
                               unsigned long long* d_hist_total, cudaStream_t stream);

}  // namespace mc
