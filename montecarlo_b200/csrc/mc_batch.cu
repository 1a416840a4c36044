// mc_batch.cu -- per-history closing reductions of mc_run_batch: what app.py:652-656 compute
// from the terminal prices, for every history of a batch in one launch.  One block per history; the
// terminal matrix [n_series, n] was just written by batch_sim_kernel and is read from L2.
#include <cuda_runtime.h>

#include <cstdint>

#include "mc_sim.cuh"

namespace mc {

constexpr int STATS_THREADS = 256;

struct RowAcc {
    double sum_d, sum_d2, vmin, vmax, tail_sum;
    unsigned long long n_above, tail_count;
};

__device__ __forceinline__ RowAcc merge(RowAcc a, const RowAcc& b) {
    a.sum_d += b.sum_d; a.sum_d2 += b.sum_d2; a.tail_sum += b.tail_sum;
    a.vmin = fmin(a.vmin, b.vmin); a.vmax = fmax(a.vmax, b.vmax);
    a.n_above += b.n_above; a.tail_count += b.tail_count;
    return a;
}

__device__ __forceinline__ RowAcc shuffle_down(const RowAcc& a, int d) {
    RowAcc o;
    o.sum_d = __shfl_down_sync(0xffffffffu, a.sum_d, d);
    o.sum_d2 = __shfl_down_sync(0xffffffffu, a.sum_d2, d);
    o.tail_sum = __shfl_down_sync(0xffffffffu, a.tail_sum, d);
    o.vmin = __shfl_down_sync(0xffffffffu, a.vmin, d);
    o.vmax = __shfl_down_sync(0xffffffffu, a.vmax, d);
    o.n_above = __shfl_down_sync(0xffffffffu, a.n_above, d);
    o.tail_count = __shfl_down_sync(0xffffffffu, a.tail_count, d);
    return o;
}

__global__ void __launch_bounds__(STATS_THREADS)
batch_stats_kernel(const double* __restrict__ terminal, unsigned long long n, unsigned long long ld,
                   const mc_series_inputs* __restrict__ inputs, const mc_garch_fit* __restrict__ fit, uint32_t min_history,
                   const double* __restrict__ pct,
                   const double* __restrict__ lo, const double* __restrict__ hi, int n_pcts, int tail_idx,
                   mc_result* __restrict__ results) {
    __shared__ RowAcc s_warp[STATS_THREADS / 32];
    const uint32_t b = blockIdx.x;
    mc_result* res = results + b;
    const mc_series_inputs in = inputs[b];
    if (in.pool_len < min_history || in.flags != 0 || (fit && fit[b].valid == 0)) {   // skipped history: n = 0 marks it
        if (threadIdx.x == 0) {
            res->n = 0; res->n_above = 0; res->mean = 0.0; res->std = 0.0; res->min = 0.0; res->max = 0.0;
            res->tail_count = 0; res->tail_sum = 0.0;
        }
        return;
    }
    const double center = in.s0;                                        // moments about s0, like the single run
    const double threshold = tail_idx >= 0 ? pct[(size_t)b * n_pcts + tail_idx] : -INFINITY;
    const double* x = terminal + (size_t)b * ld;
    RowAcc acc{0.0, 0.0, INFINITY, -INFINITY, 0.0, 0ull, 0ull};
    for (unsigned long long i = threadIdx.x; i < n; i += STATS_THREADS) {
        const double v = x[i];
        const double d = v - center;
        acc.sum_d += d;
        acc.sum_d2 = fma(d, d, acc.sum_d2);
        acc.vmin = fmin(acc.vmin, v);
        acc.vmax = fmax(acc.vmax, v);
        acc.n_above += v > center ? 1ull : 0ull;                        // app.py:654
        if (v <= threshold) { acc.tail_count += 1; acc.tail_sum += v; } // app.py:656
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) acc = merge(acc, shuffle_down(acc, d));
    if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < STATS_THREADS / 32; ++w) acc = merge(acc, s_warp[w]);     // fixed order
        const double inv = 1.0 / (double)n;
        const double md = acc.sum_d * inv;
        double var = acc.sum_d2 * inv - md * md;
        if (var < 0.0) var = 0.0;
        res->n = n;
        res->n_above = acc.n_above;
        res->mean = center + md;                                        // app.py:652
        res->std = sqrt(var);                                           // app.py:653 (ddof = 0)
        res->min = acc.vmin;
        res->max = acc.vmax;
        res->tail_count = acc.tail_count;
        res->tail_sum = acc.tail_sum;
        res->hist_lo = 0.0; res->hist_hi = 0.0;
    }
    if ((int)threadIdx.x < n_pcts) {
        const size_t k = (size_t)b * n_pcts + threadIdx.x;
        res->pct[threadIdx.x] = pct[k];
        res->pct_lo[threadIdx.x] = lo[k];
        res->pct_hi[threadIdx.x] = hi[k];
    }
}

cudaError_t launch_batch_stats(const double* d_terminal, unsigned long long n, unsigned long long ld,
                               uint32_t n_series, const mc_series_inputs* d_inputs, const mc_garch_fit* d_fit,
                               uint32_t min_history, const double* d_pct, const double* d_lo, const double* d_hi, int n_pcts,
                               int tail_idx, mc_result* d_results, cudaStream_t stream) {
    if (n_series == 0) return cudaSuccess;
    batch_stats_kernel<<<n_series, STATS_THREADS, 0, stream>>>(d_terminal, n, ld, d_inputs, d_fit, min_history, d_pct, d_lo,
                                                               d_hi, n_pcts, tail_idx, d_results);
    return cudaGetLastError();
}

}  // namespace mc
