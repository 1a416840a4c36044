// mc_select.cu -- exact percentiles by multi-target radix select, CVaR tail sums and
// the fixed-bin histogram, for sm_100a.
//
// Replaces the reference's `np.percentile(final_prices, p)` x9 (app.py:638-640) and
// `np.percentile(paths, p, axis=1)` x5 (app.py:643-645), each of which re-partitions
// its whole input on the host.  Here all order statistics of a row are found together:
// values are mapped to order-preserving integer keys and every target rank is narrowed
// one radix digit per pass from digit histograms.  The histograms are the ONLY thing
// that crosses GPUs when paths are sharded (a few KB per pass), so the result is the
// exact global order statistic with no data movement.  The first digit of the terminal
// values comes for free from the simulation kernel.
//
// Interpolation follows NumPy's method='linear' exactly (lib/_function_base_impl.py
// `_lerp`): a + (b-a)*t, and b - (b-a)*(1-t) when t >= 0.5, in fp64 without FMA
// contraction so results are bit-identical to np.percentile on the same values.
#include <math.h>

#include "mc_select.cuh"

namespace mc {

// ------------------------------------------------------------------ plan (host)
bool build_select_plan(const double* pcts, int n_pcts, unsigned long long n_global, int dtype,
                       SelectPlan* plan) {
    if (n_pcts < 0 || n_pcts > (int)MC_MAX_PCTS || n_global == 0) return false;
    SelectPlan& sp = *plan;
    sp = SelectPlan{};
    sp.n_pcts = n_pcts;
    sp.key_bits = dtype == MC_F64 ? 64 : 32;
    int left = sp.key_bits;
    int np_ = 0;
    while (left > 0) {
        const int want = np_ == 0 ? DIGIT0_BITS : DIGIT_BITS;
        const int b = left < want ? left : want;
        left -= b;
        sp.pass_bits[np_] = b;
        sp.pass_shift[np_] = left;
        ++np_;
    }
    sp.n_passes = np_;
    // np.percentile: q = p/100; virtual index (n-1)*q; floor and floor+1 (clamped); gamma
    unsigned long long lo[MC_MAX_PCTS], hi[MC_MAX_PCTS];
    for (int k = 0; k < n_pcts; ++k) {
        if (!(pcts[k] >= 0.0 && pcts[k] <= 100.0)) return false;
        if (k && pcts[k] < pcts[k - 1]) return false;
        const double q = pcts[k] / 100.0;
        const double virt = (double)(n_global - 1) * q;
        double prev = floor(virt);
        sp.pct_gamma[k] = virt - prev;
        if (virt >= (double)(n_global - 1)) {
            lo[k] = hi[k] = n_global - 1;
        } else {
            lo[k] = (unsigned long long)prev;
            hi[k] = lo[k] + 1;
        }
    }
    // sorted distinct target ranks (ranks are non-decreasing in k already; merge lo/hi)
    int nt = 0;
    auto push = [&](unsigned long long r) -> int {
        for (int j = 0; j < nt; ++j)
            if (sp.rank[j] == r) return j;
        sp.rank[nt] = r;
        return nt++;
    };
    for (int k = 0; k < n_pcts; ++k) { push(lo[k]); push(hi[k]); }
    // insertion sort (nt <= 32)
    for (int a = 1; a < nt; ++a) {
        const unsigned long long v = sp.rank[a];
        int b = a - 1;
        while (b >= 0 && sp.rank[b] > v) { sp.rank[b + 1] = sp.rank[b]; --b; }
        sp.rank[b + 1] = v;
    }
    sp.n_targets = nt;
    for (int k = 0; k < n_pcts; ++k) {
        for (int j = 0; j < nt; ++j) {
            if (sp.rank[j] == lo[k]) sp.pct_lo_idx[k] = j;
            if (sp.rank[j] == hi[k]) sp.pct_hi_idx[k] = j;
        }
    }
    return true;
}

// ------------------------------------------------------------------ digit histograms
template <typename Real>
__global__ void __launch_bounds__(256) select_hist_kernel(const Real* __restrict__ x, unsigned long long ld,
                                                          unsigned long long n, const SelectPlan* __restrict__ plan,
                                                          int pass, const RowState* __restrict__ states,
                                                          unsigned int* __restrict__ hist) {
    using Key = typename KeyT<Real>::type;
    extern __shared__ unsigned int s_bins[];
    __shared__ unsigned long long s_gp[MAX_TARGETS];
    const unsigned long long row = blockIdx.y;
    const int bits = plan->pass_bits[pass], shift = plan->pass_shift[pass];
    const int nb = 1 << bits;
    const int g_cap = pass == 0 ? 1 : plan->n_targets;
    const int ng = pass == 0 ? 1 : states[row].n_groups;
    if (pass > 0 && threadIdx.x < ng) s_gp[threadIdx.x] = states[row].group_prefix[threadIdx.x];
    for (int i = threadIdx.x; i < ng * nb; i += blockDim.x) s_bins[i] = 0u;
    __syncthreads();

    const Real* xr = x + row * ld;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const Key key = key_of(xr[i]);
        int g = 0;
        if (pass > 0) {
            const unsigned long long pfx = (unsigned long long)(key >> (shift + bits));
            // group prefixes are strictly increasing: binary search
            int lo = 0, hi = ng - 1;
            g = -1;
            while (lo <= hi) {
                const int mid = (lo + hi) >> 1;
                const unsigned long long v = s_gp[mid];
                if (v == pfx) { g = mid; break; }
                if (v < pfx) lo = mid + 1; else hi = mid - 1;
            }
            if (g < 0) continue;
        }
        atomicAdd(&s_bins[g * nb + (int)((key >> shift) & (Key)(nb - 1))], 1u);
    }
    __syncthreads();
    unsigned int* out = hist + row * (unsigned long long)g_cap * nb;
    for (int i = threadIdx.x; i < ng * nb; i += blockDim.x) {
        const unsigned int c = s_bins[i];
        if (c) atomicAdd(&out[i], c);
    }
}

cudaError_t launch_select_hist(const SelectProblem& sp, int pass, unsigned int* d_hist, int sm_count,
                               cudaStream_t stream) {
    const int nb = 1 << sp.plan.pass_bits[pass];
    const int g_cap = pass == 0 ? 1 : sp.plan.n_targets;
    const size_t bytes = (size_t)sp.rows * g_cap * nb * sizeof(unsigned int);
    cudaError_t e = cudaMemsetAsync(d_hist, 0, bytes, stream);
    if (e != cudaSuccess) return e;
    if (sp.n == 0) return cudaSuccess;
    const size_t smem = (size_t)g_cap * nb * sizeof(unsigned int);
    unsigned long long chunks = (sp.n + 256ull * 8 - 1) / (256ull * 8);
    unsigned long long budget = (unsigned long long)sm_count * 8 / sp.rows;
    if (budget < 1) budget = 1;
    if (chunks > budget) chunks = budget;
    if (sp.rows > 65535) return cudaErrorInvalidValue;
    dim3 grid((unsigned)chunks, (unsigned)sp.rows);
    if (sp.dtype == MC_F64) {
        auto k = select_hist_kernel<double>;
        if (smem > 48 * 1024) {
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        k<<<grid, 256, smem, stream>>>(static_cast<const double*>(sp.x), sp.ld, sp.n, sp.d_plan, pass,
                                       sp.d_states, d_hist);
    } else {
        auto k = select_hist_kernel<float>;
        if (smem > 48 * 1024) {
            e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return e;
        }
        k<<<grid, 256, smem, stream>>>(static_cast<const float*>(sp.x), sp.ld, sp.n, sp.d_plan, pass,
                                       sp.d_states, d_hist);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ scan: narrow every target by one digit
// One block per row.  For each group of targets sharing a key prefix, the block forms the
// inclusive cumulative count of that group's digit histogram (summed over ranks) and each
// target binary-searches the digit holding its rank.
constexpr int SCAN_THREADS = 256;

__global__ void __launch_bounds__(SCAN_THREADS) select_scan_kernel(const SelectPlan* __restrict__ plan, int pass,
                                                                   const unsigned int* __restrict__ hist,
                                                                   unsigned int n_ranks, size_t rank_stride_u32,
                                                                   RowState* __restrict__ states) {
    __shared__ unsigned long long s_cum[DIGIT0_BINS];
    __shared__ unsigned long long s_warp[SCAN_THREADS / 32];
    __shared__ unsigned long long s_prefix[MAX_TARGETS];
    const unsigned long long row = blockIdx.x;
    RowState& st = states[row];
    const int nt = plan->n_targets;
    const int bits = plan->pass_bits[pass];
    const int nb = 1 << bits;
    const int g_cap = pass == 0 ? 1 : nt;
    const int ng = pass == 0 ? 1 : st.n_groups;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // my target (threads 0..nt-1)
    unsigned long long my_prefix = 0, my_rem = 0;
    int my_group = 0;
    if (tid < nt) {
        if (pass == 0) { my_prefix = 0; my_rem = plan->rank[tid]; my_group = 0; }
        else { my_prefix = st.prefix[tid]; my_rem = st.remaining[tid]; my_group = st.group[tid]; }
    }
    const int per = nb / SCAN_THREADS;            // 16 (pass 0), 4 (10-bit digits), >=1 always? see below
    for (int g = 0; g < ng; ++g) {
        const unsigned int* h = hist + (row * g_cap + g) * (unsigned long long)nb;
        // thread-local inclusive sums over a contiguous segment of bins
        unsigned long long local[DIGIT0_BINS / SCAN_THREADS];
        unsigned long long run = 0;
        const int seg = per > 0 ? per : 1;
        const int base = tid * seg;
#pragma unroll
        for (int k = 0; k < DIGIT0_BINS / SCAN_THREADS; ++k) {
            if (k < seg && base + k < nb) {
                unsigned long long c = 0;
                for (unsigned int r = 0; r < n_ranks; ++r) c += h[r * rank_stride_u32 + base + k];
                run += c;
            }
            local[k] = run;
        }
        // exclusive scan of `run` over threads
        unsigned long long incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        unsigned long long warp_off = 0;
        for (int w = 0; w < warp; ++w) warp_off += s_warp[w];
        const unsigned long long excl = warp_off + incl - run;
#pragma unroll
        for (int k = 0; k < DIGIT0_BINS / SCAN_THREADS; ++k)
            if (k < seg && base + k < nb) s_cum[base + k] = excl + local[k];
        __syncthreads();
        if (tid < nt && my_group == g) {
            // smallest digit d with cum[d] > my_rem
            int lo = 0, hi = nb - 1;
            while (lo < hi) {
                const int mid = (lo + hi) >> 1;
                if (s_cum[mid] > my_rem) hi = mid; else lo = mid + 1;
            }
            my_rem -= lo ? s_cum[lo - 1] : 0ull;
            my_prefix = (my_prefix << bits) | (unsigned long long)lo;
        }
        __syncthreads();
    }
    if (tid < nt) {
        st.prefix[tid] = my_prefix;
        st.remaining[tid] = my_rem;
        s_prefix[tid] = my_prefix;
    }
    __syncthreads();
    if (tid == 0) {
        // targets are sorted by rank, hence prefixes are non-decreasing: group equal neighbours
        int g = -1;
        unsigned long long last = 0;
        for (int j = 0; j < nt; ++j) {
            if (j == 0 || s_prefix[j] != last) { ++g; st.group_prefix[g] = s_prefix[j]; last = s_prefix[j]; }
            st.group[j] = g;
        }
        st.n_groups = g + 1;
    }
}

cudaError_t launch_select_scan(const SelectProblem& sp, int pass, const unsigned int* d_hist,
                               unsigned int n_ranks, size_t rank_stride_u32, cudaStream_t stream) {
    if (sp.plan.n_targets == 0) return cudaSuccess;
    select_scan_kernel<<<(unsigned)sp.rows, SCAN_THREADS, 0, stream>>>(sp.d_plan, pass, d_hist, n_ranks,
                                                                       rank_stride_u32, sp.d_states);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ finish: keys -> values -> percentiles
__device__ __forceinline__ double numpy_lerp(double a, double b, double t) {
    // numpy/lib/_function_base_impl.py:_lerp -- explicit _rn ops: no FMA contraction
    const double diff = __dsub_rn(b, a);
    double r = __dadd_rn(a, __dmul_rn(diff, t));
    if (t >= 0.5) r = __dsub_rn(b, __dmul_rn(diff, __dsub_rn(1.0, t)));
    return r;
}

__global__ void select_finish_kernel(const SelectPlan* __restrict__ plan, const RowState* __restrict__ states,
                                     unsigned long long rows, double* __restrict__ out, double* __restrict__ out_lo,
                                     double* __restrict__ out_hi) {
    const unsigned long long idx = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int np_ = plan->n_pcts;
    if (idx >= rows * (unsigned long long)np_) return;
    const unsigned long long row = idx / np_;
    const int k = (int)(idx % np_);
    const RowState& st = states[row];
    const unsigned long long ka = st.prefix[plan->pct_lo_idx[k]], kb = st.prefix[plan->pct_hi_idx[k]];
    double a, b;
    if (plan->key_bits == 32) { a = (double)value_of((uint32_t)ka); b = (double)value_of((uint32_t)kb); }
    else { a = value_of((uint64_t)ka); b = value_of((uint64_t)kb); }
    out[idx] = numpy_lerp(a, b, plan->pct_gamma[k]);
    if (out_lo) out_lo[idx] = a;
    if (out_hi) out_hi[idx] = b;
}

cudaError_t launch_select_finish(const SelectProblem& sp, double* d_out, double* d_lo, double* d_hi,
                                 cudaStream_t stream) {
    const unsigned long long total = sp.rows * (unsigned long long)sp.plan.n_pcts;
    if (total == 0) return cudaSuccess;
    select_finish_kernel<<<(unsigned)((total + 127) / 128), 128, 0, stream>>>(sp.d_plan, sp.d_states, sp.rows, d_out,
                                                                              d_lo, d_hi);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ moments merge (app.py:652-654)
__global__ void merge_moments_kernel(const unsigned long long* __restrict__ g, unsigned int n_ranks, size_t stride,
                                     double center, DeviceResult* __restrict__ res) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long n = 0, above = 0;
    double sd = 0.0, sd2 = 0.0, mn = INFINITY, mx = -INFINITY;
    for (unsigned int r = 0; r < n_ranks; ++r) {       // rank order: deterministic
        const unsigned long long* w = g + r * stride;
        n += w[0];
        above += w[1];
        sd += __longlong_as_double((long long)w[2]);
        sd2 += __longlong_as_double((long long)w[3]);
        mn = fmin(mn, __longlong_as_double((long long)w[4]));
        mx = fmax(mx, __longlong_as_double((long long)w[5]));
    }
    const double inv = n ? 1.0 / (double)n : 0.0;
    const double md = sd * inv;
    double var = sd2 * inv - md * md;
    if (var < 0.0) var = 0.0;
    res->n = n;
    res->n_above = above;
    res->mean = center + md;
    res->std = sqrt(var);
    res->vmin = mn;
    res->vmax = mx;
    res->center = center;
    res->tail_count = 0;
    res->tail_sum = 0.0;
    res->tail_threshold = -INFINITY;
    res->hist_lo = 0.0;
    res->hist_hi = 0.0;
}

cudaError_t launch_merge_moments(const unsigned long long* d_gathered, unsigned int n_ranks, size_t stride,
                                 double center, DeviceResult* d_res, cudaStream_t stream) {
    merge_moments_kernel<<<1, 32, 0, stream>>>(d_gathered, n_ranks, stride, center, d_res);
    return cudaGetLastError();
}

__global__ void publish_percentiles_kernel(const double* pct, const double* lo, const double* hi, int n_pcts,
                                           int tail_idx, DeviceResult* res) {
    const int k = threadIdx.x;
    if (k < n_pcts) { res->pct[k] = pct[k]; res->pct_lo[k] = lo[k]; res->pct_hi[k] = hi[k]; }
    if (k == 0) res->tail_threshold = tail_idx >= 0 ? pct[tail_idx] : -INFINITY;
}

cudaError_t launch_publish_percentiles(const double* d_pct, const double* d_lo, const double* d_hi, int n_pcts,
                                       int tail_idx, DeviceResult* d_res, cudaStream_t stream) {
    publish_percentiles_kernel<<<1, 32, 0, stream>>>(d_pct, d_lo, d_hi, n_pcts, tail_idx, d_res);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ tail: mean(final[final <= p]) (app.py:656)
template <typename Real>
__global__ void __launch_bounds__(256) tail_kernel(const Real* __restrict__ x, unsigned long long n,
                                                   const DeviceResult* __restrict__ res, double* __restrict__ scratch) {
    __shared__ double s_sum[8];
    __shared__ unsigned long long s_cnt[8];
    const double thr = res->tail_threshold;
    double sum = 0.0;
    unsigned long long cnt = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double v = (double)x[i];
        if (v <= thr) { sum += v; ++cnt; }
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        sum += __shfl_down_sync(0xffffffffu, sum, d);
        cnt += __shfl_down_sync(0xffffffffu, cnt, d);
    }
    if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = sum; s_cnt[threadIdx.x >> 5] = cnt; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ts = 0.0; unsigned long long tc = 0;
        for (int w = 0; w < 8; ++w) { ts += s_sum[w]; tc += s_cnt[w]; }
        scratch[2 * blockIdx.x] = ts;
        scratch[2 * blockIdx.x + 1] = __longlong_as_double((long long)tc);
    }
}

__global__ void __launch_bounds__(256) tail_reduce_kernel(const double* scratch, int n_blocks, unsigned long long* words) {
    __shared__ double s_sum[256];
    __shared__ unsigned long long s_cnt[256];
    double ts = 0.0; unsigned long long tc = 0;
    for (int i = threadIdx.x; i < n_blocks; i += 256) {
        ts += scratch[2 * i];
        tc += (unsigned long long)__double_as_longlong(scratch[2 * i + 1]);
    }
    s_sum[threadIdx.x] = ts; s_cnt[threadIdx.x] = tc;
    __syncthreads();
    for (int w = 128; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) { s_sum[threadIdx.x] += s_sum[threadIdx.x + w]; s_cnt[threadIdx.x] += s_cnt[threadIdx.x + w]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { words[0] = s_cnt[0]; words[1] = (unsigned long long)__double_as_longlong(s_sum[0]); }
}

cudaError_t launch_tail(const void* x, int dtype, unsigned long long n, const DeviceResult* d_res, double* d_scratch,
                        unsigned long long* d_words, int sm_count, cudaStream_t stream) {
    unsigned long long blocks = (n + 256ull * 8 - 1) / (256ull * 8);
    if (blocks < 1) blocks = 1;
    if (blocks > 1024) blocks = 1024;
    if (blocks > (unsigned long long)sm_count * 6) blocks = (unsigned long long)sm_count * 6;
    if (dtype == MC_F64) tail_kernel<double><<<(unsigned)blocks, 256, 0, stream>>>(static_cast<const double*>(x), n, d_res, d_scratch);
    else tail_kernel<float><<<(unsigned)blocks, 256, 0, stream>>>(static_cast<const float*>(x), n, d_res, d_scratch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    tail_reduce_kernel<<<1, 256, 0, stream>>>(d_scratch, (int)blocks, d_words);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ fixed-bin histogram
// Semantics: np.histogram(x.astype(float64), bins, range=(lo, hi)) -- numpy/lib/_histograms_impl.py
// uniform-bin path: values outside [lo, hi] dropped; estimate the bin by scaling, then correct it
// against edges = linspace(lo, hi, bins+1); last bin closed on the right.
__global__ void hist_edges_kernel(unsigned int bins, double lo, double hi, DeviceResult* res, double* edges) {
    if (lo >= hi) { lo = res->vmin; hi = res->vmax; }       // range=None: (a.min(), a.max())
    if (lo == hi) { lo -= 0.5; hi += 0.5; }                 // _get_outer_edges
    if (blockIdx.x == 0 && threadIdx.x == 0) { res->hist_lo = lo; res->hist_hi = hi; }
    const double step = __ddiv_rn(__dsub_rn(hi, lo), (double)bins);      // np.linspace
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i <= bins; i += gridDim.x * blockDim.x)
        edges[i] = (i == bins) ? hi : __dadd_rn(__dmul_rn((double)i, step), lo);
}

template <typename Real>
__global__ void __launch_bounds__(256) value_hist_kernel(const Real* __restrict__ x, unsigned long long n,
                                                         unsigned int bins, const double* __restrict__ edges,
                                                         unsigned long long* __restrict__ counts) {
    extern __shared__ unsigned int s_counts[];
    for (unsigned int i = threadIdx.x; i < bins; i += blockDim.x) s_counts[i] = 0u;
    __syncthreads();
    const double lo = edges[0], hi = edges[bins];
    const double denom = __dsub_rn(hi, lo);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const double v = (double)x[i];
        if (!(v >= lo && v <= hi)) continue;
        const double f = __dmul_rn(__ddiv_rn(__dsub_rn(v, lo), denom), (double)bins);
        long long idx = (long long)f;
        if (idx >= (long long)bins) idx = bins - 1;
        if (idx < 0) idx = 0;
        if (v < edges[idx]) --idx;
        else if (v >= edges[idx + 1] && idx != (long long)bins - 1) ++idx;
        atomicAdd(&s_counts[idx], 1u);
    }
    __syncthreads();
    for (unsigned int i = threadIdx.x; i < bins; i += blockDim.x) {
        const unsigned int c = s_counts[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
    }
}

cudaError_t launch_value_hist(const void* x, int dtype, unsigned long long n, unsigned int bins, double lo, double hi,
                              DeviceResult* d_res, double* d_edges, unsigned long long* d_counts, int sm_count,
                              cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(d_counts, 0, (size_t)bins * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    hist_edges_kernel<<<(bins + 256) / 256, 256, 0, stream>>>(bins, lo, hi, d_res, d_edges);
    e = cudaGetLastError();
    if (e != cudaSuccess || n == 0) return e;
    unsigned long long blocks = (n + 256ull * 8 - 1) / (256ull * 8);
    if (blocks > (unsigned long long)sm_count * 8) blocks = (unsigned long long)sm_count * 8;
    const size_t smem = (size_t)bins * sizeof(unsigned int);
    if (dtype == MC_F64) {
        auto k = value_hist_kernel<double>;
        if (smem > 48 * 1024 && (e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        k<<<(unsigned)blocks, 256, smem, stream>>>(static_cast<const double*>(x), n, bins, d_edges, d_counts);
    } else {
        auto k = value_hist_kernel<float>;
        if (smem > 48 * 1024 && (e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)) != cudaSuccess) return e;
        k<<<(unsigned)blocks, 256, smem, stream>>>(static_cast<const float*>(x), n, bins, d_edges, d_counts);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ final merge over ranks
// final-stage words per rank: [0] tail count  [1] tail sum bits  [2..2+bins) histogram counts
__global__ void merge_final_kernel(const unsigned long long* __restrict__ g, unsigned int n_ranks, size_t stride,
                                   unsigned int bins, DeviceResult* res, unsigned long long* hist_total) {
    const unsigned int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < bins) {
        unsigned long long c = 0;
        for (unsigned int r = 0; r < n_ranks; ++r) c += g[r * stride + 2 + i];
        hist_total[i] = c;
    }
    if (i == 0) {
        unsigned long long tc = 0; double ts = 0.0;
        for (unsigned int r = 0; r < n_ranks; ++r) {
            tc += g[r * stride];
            ts += __longlong_as_double((long long)g[r * stride + 1]);
        }
        res->tail_count = tc;
        res->tail_sum = ts;
    }
}

cudaError_t launch_merge_final(const unsigned long long* d_gathered, unsigned int n_ranks, size_t stride,
                               unsigned int bins, DeviceResult* d_res, unsigned long long* d_hist_total,
                               cudaStream_t stream) {
    const unsigned int threads = bins > 0 ? bins : 1;
    merge_final_kernel<<<(threads + 255) / 256, 256, 0, stream>>>(d_gathered, n_ranks, stride, bins, d_res, d_hist_total);
    return cudaGetLastError();
}

}  // namespace mc
