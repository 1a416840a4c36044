// mc_inputs.cu -- the routine's inputs from raw close prices, many price histories per launch
// (SURVEY 8f N3).  Reference producers, all fp64:
//   log_ret        = log(close / close.shift(1))                              app.py:224
//   rvol_20        = log_ret.rolling(20).std() * sqrt(252)   (ddof = 1)        app.py:233
//   rows kept      = [first_row, n)  (the 200-day average leaves 199 NaN rows)  app.py:252, 392-393
//   returns        = log_ret of the kept rows                                  app.py:592
//   current_var    = rvol_20[last]^2 / 252                                     app.py:606
//   std_returns    = returns / (np.std(returns) + 1e-10)   (ddof = 0)          app.py:609
//   current_price  = close[last]                                               app.py:1213
// One block per history: three strided sweeps over <= a few thousand prices (mean, squared deviations,
// scaled pool) with fp64 block reductions; the prices stay in L1/L2 between sweeps.  Byte-bound and tiny:
// 8 B read + 4..8 B written per observation.
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/mc_b200.h"
#include "mc_inputs.cuh"

namespace mc {

constexpr int INPUT_THREADS = 256;
constexpr int RVOL_WINDOW = 20;

__device__ __forceinline__ double block_sum(double v, double* s_warp) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();                       // s_warp may still be read from the previous reduction
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    double t = 0.0;
#pragma unroll
    for (int w = 0; w < INPUT_THREADS / 32; ++w) t += s_warp[w];     // same order in every thread
    return t;
}

// log return of frame row i (i >= 1); app.py:392 turns +-inf into 0
__device__ __forceinline__ double log_return(const double* c, uint32_t i, unsigned int* bad) {
    double r = log(c[i] / c[i - 1]);
    if (isinf(r)) r = 0.0;
    if (r != r) *bad = 1u;                 // the reference would drop the row; not representable in a fixed layout
    return r;
}

template <typename T>
__global__ void __launch_bounds__(INPUT_THREADS)
prepare_inputs_kernel(const double* __restrict__ close, const uint32_t* __restrict__ lengths, uint32_t n_obs,
                      unsigned long long close_ld, uint32_t first_row, T* __restrict__ pool,
                      unsigned long long pool_ld, mc_series_inputs* __restrict__ out) {
    __shared__ double s_warp[INPUT_THREADS / 32];
    const uint32_t b = blockIdx.x;
    const uint32_t len = lengths ? (lengths[b] < n_obs ? lengths[b] : n_obs) : n_obs;
    const double* c = close + (size_t)b * close_ld;
    // rows before RVOL_WINDOW have no rvol_20 and go with the same dropna (app.py:393)
    uint32_t start = first_row > (uint32_t)RVOL_WINDOW ? first_row : (uint32_t)RVOL_WINDOW;
    const uint32_t n = len > start ? len - start : 0u;
    unsigned int bad = 0u;
    if (n == 0) {
        if (threadIdx.x == 0) {
            out[b].s0 = len ? c[len - 1] : 0.0;
            out[b].var0 = 0.0; out[b].pool_std = 0.0; out[b].pool_len = 0; out[b].flags = 0;
        }
        return;
    }
    double acc = 0.0;
    for (uint32_t i = start + threadIdx.x; i < len; i += INPUT_THREADS) acc += log_return(c, i, &bad);
    const double mean = block_sum(acc, s_warp) / (double)n;
    acc = 0.0;
    for (uint32_t i = start + threadIdx.x; i < len; i += INPUT_THREADS) {
        const double d = log_return(c, i, &bad) - mean;
        acc += d * d;
    }
    const double sd = sqrt(block_sum(acc, s_warp) / (double)n);          // np.std, ddof = 0
    const double denom = sd + 1e-10;
    T* p = pool + (size_t)b * pool_ld;
    for (uint32_t i = start + threadIdx.x; i < len; i += INPUT_THREADS) p[i - start] = (T)(log_return(c, i, &bad) / denom);

    // the last rolling window (warp 0): sample standard deviation of the final 20 log returns
    double var0 = 0.0;
    if (threadIdx.x < 32) {
        const bool in = threadIdx.x < RVOL_WINDOW;
        const double x = in ? log_return(c, len - RVOL_WINDOW + threadIdx.x, &bad) : 0.0;
        double s = x;
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        const double m = s / RVOL_WINDOW;
        double d = in ? (x - m) * (x - m) : 0.0;
        for (int o = 16; o > 0; o >>= 1) d += __shfl_xor_sync(0xffffffffu, d, o);
        const double rvol = sqrt(d / (RVOL_WINDOW - 1)) * sqrt(252.0);   // app.py:233
        var0 = rvol * rvol / 252.0;                                       // app.py:606
    }
    const int any_bad = __syncthreads_or((int)bad);
    if (threadIdx.x == 0) {
        out[b].s0 = c[len - 1];
        out[b].var0 = var0;
        out[b].pool_std = sd;
        out[b].pool_len = n;
        out[b].flags = any_bad ? MC_INPUT_NAN : 0u;
    }
}

cudaError_t launch_prepare_inputs(const double* d_close, const uint32_t* d_lengths, uint32_t n_series, uint32_t n_obs,
                                  unsigned long long close_ld, uint32_t first_row, int dtype, void* d_pool,
                                  unsigned long long pool_ld, mc_series_inputs* d_out, cudaStream_t stream) {
    if (n_series == 0) return cudaSuccess;
    if (dtype == MC_F64)
        prepare_inputs_kernel<double><<<n_series, INPUT_THREADS, 0, stream>>>(
            d_close, d_lengths, n_obs, close_ld, first_row, static_cast<double*>(d_pool), pool_ld, d_out);
    else
        prepare_inputs_kernel<float><<<n_series, INPUT_THREADS, 0, stream>>>(
            d_close, d_lengths, n_obs, close_ld, first_row, static_cast<float*>(d_pool), pool_ld, d_out);
    return cudaGetLastError();
}

}  // namespace mc
