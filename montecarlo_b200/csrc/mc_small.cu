// mc_small.cu -- the reference's own call sizes (app.py:1189 5..60 days, app.py:1195 500..5000 simulations; BASELINE
// configs[0] 10,000 x 252) in TWO launches: the simulation kernel, then one block per row of the stored paths that
// sorts its row in shared memory and reads the fan-chart percentiles off it (app.py:643-645); the block of the last
// row -- the terminal values, app.py:636 -- also takes the nine terminal percentiles (app.py:638-640), VaR / CVaR
// (app.py:655-656) and the moments (app.py:652-654) and writes the result block straight into pinned host memory.
// The general staged selection needs ~21 launches for the same answer; at these sizes launch latency is all there is.
//
// Same results as the general path: order statistics, interpolation (NumPy's _lerp), counts, minimum / maximum and --
// because the per-block partials of the simulation kernel are added in the very order reduce_partials_kernel /
// merge_moments_kernel use -- mean and std are bit-identical; the CVaR tail sum agrees to rounding (other order).
#include <math.h>

#include "mc_launch.cuh"
#include "mc_small.cuh"

namespace mc {

constexpr int SMALL_THREADS = 1024;

template <typename Key>
__device__ void small_bitonic_sort(Key* s, int P) {
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P; i += SMALL_THREADS) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const Key x = s[i], y = s[ixj];
                    const bool up = (i & k) == 0;
                    if ((x > y) == up) { s[i] = y; s[ixj] = x; }
                }
            }
            __syncthreads();
        }
    }
}

template <typename Key> struct SmallValue;
template <> struct SmallValue<uint32_t> { static __device__ __forceinline__ double get(uint32_t k) { return (double)value_of(k); } };
template <> struct SmallValue<uint64_t> { static __device__ __forceinline__ double get(uint64_t k) { return value_of(k); } };

__device__ __forceinline__ void combine_partial(BlockPartial& a, const BlockPartial& b) {
    a.sum_d += b.sum_d;
    a.sum_d2 += b.sum_d2;
    a.vmin = fmin(a.vmin, b.vmin);
    a.vmax = fmax(a.vmax, b.vmax);
    a.n += b.n;
    a.n_above += b.n_above;
}

template <typename Real>
__global__ void __launch_bounds__(SMALL_THREADS) small_rows_kernel(const SmallArgs a) {
    using Key = typename KeyT<Real>::type;
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    Key* s_keys = reinterpret_cast<Key*>(smem_dyn);                                  // [P]
    __shared__ double s_pct[MC_MAX_PCTS];
    __shared__ double s_red[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int row = blockIdx.x;
    const int n = (int)a.n;
    int P = 1;
    while (P < n) P <<= 1;
    const Real* x = static_cast<const Real*>(a.paths) + (size_t)row * a.ld;
    for (int i = tid; i < P; i += SMALL_THREADS) s_keys[i] = i < n ? key_of(x[i]) : ~(Key)0;
    __syncthreads();
    small_bitonic_sort(s_keys, P);
    unsigned long long* host_words = reinterpret_cast<unsigned long long*>(a.host_out);
    constexpr int RW = (int)(sizeof(DeviceResult) / 8);
    if (tid < a.n_path_pcts) {              // app.py:643-645
        const SelectPlan& pl = *a.plan_f;
        const double va = SmallValue<Key>::get(s_keys[pl.rank[pl.pct_lo_idx[tid]]]);
        const double vb = SmallValue<Key>::get(s_keys[pl.rank[pl.pct_hi_idx[tid]]]);
        const double pc = numpy_lerp(va, vb, pl.pct_gamma[tid]);
        a.fan[(size_t)row * a.n_path_pcts + tid] = pc;
        host_words[RW + (size_t)tid * a.rows + row] = (unsigned long long)__double_as_longlong(pc);    // [pct][row] for the caller
    }
    if (row != a.rows - 1) return;
    // ---- the terminal row: percentiles, tail, moments -> result block
    DeviceResult& res = *a.res;
    if (tid < a.n_pcts) {                   // app.py:638-640
        const SelectPlan& pl = *a.plan_t;
        const double va = SmallValue<Key>::get(s_keys[pl.rank[pl.pct_lo_idx[tid]]]);
        const double vb = SmallValue<Key>::get(s_keys[pl.rank[pl.pct_hi_idx[tid]]]);
        const double pc = numpy_lerp(va, vb, pl.pct_gamma[tid]);
        s_pct[tid] = pc;
        res.pct[tid] = pc; res.pct_lo[tid] = va; res.pct_hi[tid] = vb;
    }
    __syncthreads();
    // app.py:656: values <= the interpolated threshold are a prefix of the sorted row
    const bool want_tail = a.tail_idx >= 0 && a.n_pcts > 0;
    const double thr = want_tail ? s_pct[a.tail_idx] : -INFINITY;
    int cnt = 0;
    if (want_tail) {
        int lo = 0, hi = n;                 // first index with value > thr
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (SmallValue<Key>::get(s_keys[mid]) <= thr) lo = mid + 1; else hi = mid; }
        cnt = lo;
    }
    double tsum = 0.0;
    for (int i = tid; i < cnt; i += SMALL_THREADS) tsum += SmallValue<Key>::get(s_keys[i]);
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) tsum += __shfl_down_sync(0xffffffffu, tsum, d);
    if (lane == 0) s_red[warp] = tsum;
    __syncthreads();
    // moments: the simulation kernel's per-block partials, added in reduce_partials_kernel's order -- thread t takes
    // partials t, t + 1024, ...; then the tree (t, t + w), w = 512 .. 1 (slots beyond the partials are neutral elements,
    // so the smaller tree over the next power of two gives the same bits)
    BlockPartial* s_part = reinterpret_cast<BlockPartial*>(smem_dyn);          // the keys are done with
    __syncthreads();
    {
        BlockPartial acc;
        acc.sum_d = 0.0; acc.sum_d2 = 0.0; acc.vmin = INFINITY; acc.vmax = -INFINITY; acc.n = 0; acc.n_above = 0;
        for (int i = tid; i < a.n_partials; i += SMALL_THREADS) combine_partial(acc, a.partials[i]);
        s_part[tid] = acc;
    }
    __syncthreads();
    for (int w = SMALL_THREADS / 2; w > 0; w >>= 1) {
        if (tid < w) combine_partial(s_part[tid], s_part[tid + w]);
        __syncthreads();
    }
    if (tid == 0) {
        for (int w = 1; w < SMALL_THREADS / 32; ++w) tsum += s_red[w];
        const BlockPartial t = s_part[0];
        double sd = 0.0, sd2 = 0.0;         // merge_moments_kernel with one rank
        sd += t.sum_d; sd2 += t.sum_d2;
        const double inv = t.n ? 1.0 / (double)t.n : 0.0;
        const double md = sd * inv;
        double var = sd2 * inv - md * md;
        if (var < 0.0) var = 0.0;
        res.n = t.n; res.n_above = t.n_above; res.mean = a.center + md; res.std = sqrt(var);
        res.vmin = t.vmin; res.vmax = t.vmax;
        res.center = a.center;
        res.tail_threshold = thr;
        res.tail_count = (unsigned long long)cnt;
        res.tail_sum = want_tail ? tsum : 0.0;
        res.hist_lo = 0.0; res.hist_hi = 0.0;
        res.status = 0ull;
    }
    __syncthreads();
    const unsigned long long* res_words = reinterpret_cast<const unsigned long long*>(a.res);
    for (int i = tid; i < RW; i += SMALL_THREADS) host_words[i] = res_words[i];
}

size_t small_rows_smem(int dtype, unsigned long long n) {
    size_t P = 1;
    while (P < n) P <<= 1;
    const size_t keys = P * (dtype == MC_F64 ? 8 : 4), part = (size_t)SMALL_THREADS * sizeof(BlockPartial);
    return keys > part ? keys : part;
}

cudaError_t launch_small_rows(const SmallArgs& a, cudaStream_t s) {
    const size_t smem = small_rows_smem(a.dtype, a.n);
    cudaError_t e;
    if (a.dtype == MC_F64) {
        if ((e = ensure_dynamic_smem(small_rows_kernel<double>, smem)) != cudaSuccess) return e;
        small_rows_kernel<double><<<a.rows, SMALL_THREADS, smem, s>>>(a);
    } else {
        if ((e = ensure_dynamic_smem(small_rows_kernel<float>, smem)) != cudaSuccess) return e;
        small_rows_kernel<float><<<a.rows, SMALL_THREADS, smem, s>>>(a);
    }
    return cudaGetLastError();
}

}  // namespace mc
