// mc_calibrate.cu -- GARCH(1,1) parameters per price history by Gaussian quasi-maximum likelihood
// (SURVEY 8f N4).  The reference hard-codes omega = 1e-5, alpha = 0.10, beta = 0.85 (app.py:601-603) for its
// recursion var' = omega + alpha r^2 + beta var (app.py:632-633); this replaces the constants by the values
// that maximise the likelihood of each history's own returns under that same recursion.  Opt-in: it changes
// the results.
//
// Objective (to minimise), on the standardised returns z_t = r_t / (std + 1e-10) that mc_prepare_inputs
// already produced (the model is scale-equivariant: alpha and beta carry over, omega_r = c^2 omega_z):
//     l(alpha, beta) = sum_t [ log h_t + z_t^2 / h_t ],   h_0 = s2 = mean(z^2),
//     h_{t+1} = omega + alpha z_t^2 + beta h_t,           omega = s2 (1 - alpha - beta)   (variance targeting)
// Search: deterministic grid refinement in (alpha, persistence = alpha + beta): 16 x 16 candidates per level,
// the box shrinks 4x around the best candidate each level.  One block per history, one candidate per thread,
// each thread runs the whole recursion (serial in t, fp64); z^2 is staged in shared memory.  ~8 levels x H
// steps x (1 log + 1 div + 3 fma) per candidate: microseconds of work, latency-bound by the dependent chain.
#include <cuda_runtime.h>

#include <cstdint>

#include "mc_inputs.cuh"

namespace mc {

constexpr int FIT_GRID = 16;
constexpr int FIT_THREADS = FIT_GRID * FIT_GRID;
constexpr double FIT_ALPHA_MAX = 0.5;
constexpr double FIT_PERSIST_MIN = 0.5;
constexpr double FIT_PERSIST_MAX = 0.9995;

struct Best { double value; int index; };

__device__ __forceinline__ Best better(Best a, Best b) {
    // smaller objective wins; NaN never wins; ties go to the lower candidate index
    const bool take_b = (b.value < a.value) || (b.value == a.value && b.index < a.index) || (a.value != a.value && b.value == b.value);
    return take_b ? b : a;
}

__global__ void __launch_bounds__(FIT_THREADS)
calibrate_garch_kernel(const double* __restrict__ pool, unsigned long long pool_ld,
                       const mc_series_inputs* __restrict__ inputs, uint32_t min_history, uint32_t levels,
                       uint32_t smem_len, mc_garch_fit* __restrict__ fit) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    double* s_z2 = reinterpret_cast<double*>(smem_dyn);
    __shared__ double s_red[FIT_THREADS / 32];
    __shared__ Best s_best[FIT_THREADS / 32];
    __shared__ double s_box[4];                       // alpha_lo, alpha_step, persist_lo, persist_step
    const uint32_t b = blockIdx.x;
    const mc_series_inputs in = inputs[b];
    const uint32_t n = (uint32_t)in.pool_len;
    if (n < min_history || in.flags != 0) {
        if (threadIdx.x == 0) { fit[b].omega = 0.0; fit[b].alpha = 0.0; fit[b].beta = 0.0; fit[b].objective = 0.0; fit[b].valid = 0; }
        return;
    }
    const double* z = pool + (size_t)b * pool_ld;
    const bool staged = n <= smem_len;
    double acc = 0.0;
    for (uint32_t i = threadIdx.x; i < n; i += FIT_THREADS) {
        const double v = z[i] * z[i];
        if (staged) s_z2[i] = v;
        acc += v;
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    if (threadIdx.x == 0) {
        s_box[0] = 0.0;
        s_box[1] = FIT_ALPHA_MAX / (FIT_GRID - 1);
        s_box[2] = FIT_PERSIST_MIN;
        s_box[3] = (FIT_PERSIST_MAX - FIT_PERSIST_MIN) / (FIT_GRID - 1);
    }
    __syncthreads();
    double s2 = 0.0;
#pragma unroll
    for (int w = 0; w < FIT_THREADS / 32; ++w) s2 += s_red[w];
    s2 /= (double)n;

    const int ia = threadIdx.x % FIT_GRID, ip = threadIdx.x / FIT_GRID;
    double best_alpha = 0.0, best_persist = 0.0, best_value = 0.0;
    for (uint32_t level = 0; level < levels; ++level) {
        const double alpha = s_box[0] + ia * s_box[1];
        const double persist = s_box[2] + ip * s_box[3];
        const double beta = persist - alpha;
        double value = INFINITY;
        if (alpha >= 0.0 && beta >= 0.0 && persist <= FIT_PERSIST_MAX + 1e-12 && persist > 0.0) {
            const double omega = s2 * (1.0 - persist);
            double h = s2, l = 0.0;
            for (uint32_t t = 0; t < n; ++t) {
                const double x = staged ? s_z2[t] : z[t] * z[t];
                l += log(h) + x / h;
                h = omega + alpha * x + beta * h;
            }
            value = l;
        }
        Best mine{value, (int)threadIdx.x};
        for (int o = 16; o > 0; o >>= 1) {
            Best other;
            other.value = __shfl_xor_sync(0xffffffffu, mine.value, o);
            other.index = __shfl_xor_sync(0xffffffffu, mine.index, o);
            mine = better(mine, other);
        }
        __syncthreads();                              // everyone has read s_box / s_best of the last level
        if ((threadIdx.x & 31) == 0) s_best[threadIdx.x >> 5] = mine;
        __syncthreads();
        Best win = s_best[0];
#pragma unroll
        for (int w = 1; w < FIT_THREADS / 32; ++w) win = better(win, s_best[w]);
        const int wa = win.index % FIT_GRID, wp = win.index / FIT_GRID;
        best_alpha = s_box[0] + wa * s_box[1];
        best_persist = s_box[2] + wp * s_box[3];
        best_value = win.value;
        __syncthreads();
        if (threadIdx.x == 0) {
            // next box: +-2 old steps around the winner -> the step shrinks 4 x (15 intervals over 4 old steps = 3.75 x)
            const double sa = s_box[1], sp = s_box[3];
            s_box[0] = best_alpha - 2.0 * sa;
            s_box[1] = 4.0 * sa / (FIT_GRID - 1);
            s_box[2] = best_persist - 2.0 * sp;
            s_box[3] = 4.0 * sp / (FIT_GRID - 1);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double c = in.pool_std + 1e-10;         // r = c z  (app.py:609)
        fit[b].alpha = best_alpha;
        fit[b].beta = best_persist - best_alpha;
        fit[b].omega = s2 * (1.0 - best_persist) * c * c;
        fit[b].objective = best_value;
        fit[b].valid = 1;
    }
}

cudaError_t launch_calibrate_garch(const double* d_pool, unsigned long long pool_ld, const mc_series_inputs* d_inputs,
                                   uint32_t n_series, uint32_t min_history, uint32_t levels, mc_garch_fit* d_fit,
                                   cudaStream_t stream) {
    if (n_series == 0) return cudaSuccess;
    size_t smem = (size_t)pool_ld * sizeof(double);
    uint32_t smem_len = (uint32_t)pool_ld;
    if (smem > 40 * 1024) { smem = 0; smem_len = 0; }            // long histories: z^2 recomputed from L1/L2
    calibrate_garch_kernel<<<n_series, FIT_THREADS, smem, stream>>>(d_pool, pool_ld, d_inputs, min_history, levels,
                                                                   smem_len, d_fit);
    return cudaGetLastError();
}

}  // namespace mc
