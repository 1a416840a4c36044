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
__device__ __forceinline__ int bit_length_u64(unsigned long long v) { return v ? 64 - __clzll((long long)v) : 0; }

// ---- order statistics of a row held in shared memory ----------------------------------------------------
// Short rows are sorted outright.  Longer ones are narrowed: a 4096-bin histogram over the key range that still holds
// undecided ranks, each rank moves into its bin, bins with <= 1024 keys are collected and sorted (the same work-list
// scheme as the candidate selection of the fused closing, mc_fused.cu) -- three sweeps over the row instead of the
// ~100 compare-exchange rounds of a bitonic sort of 16384 keys.
constexpr int SMALL_SORT_MAX = 2048;
constexpr int SMALL_SUB = 4096, SMALL_COLLECT = 1024, SMALL_ITEMS = 96;
struct SmallItem { unsigned long long lo, hi, before; unsigned int pop; int t0, nt; };
template <typename Key>
struct SmallScratch {
    unsigned int sub[SMALL_SUB];
    Key sort[SMALL_COLLECT];
    SmallItem items[SMALL_ITEMS];
    unsigned long long red[2 * 32];
    unsigned int warp[32];
    unsigned int n_collected;
    int n_items, bad;
};

// keys[n] unsorted, ranks[nt] ascending -> out[nt]; all threads of the block; returns false on an inconsistency
template <typename Key>
__device__ bool small_select(const Key* keys, int n, const unsigned long long* ranks, int nt, unsigned long long* out,
                             SmallScratch<Key>& sc) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (nt == 0) return true;
    unsigned long long kmin = ~0ull, kmax = 0ull;
    for (int i = tid; i < n; i += SMALL_THREADS) { const unsigned long long k = keys[i]; kmin = k < kmin ? k : kmin; kmax = k > kmax ? k : kmax; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const unsigned long long x = __shfl_down_sync(0xffffffffu, kmin, d), y = __shfl_down_sync(0xffffffffu, kmax, d);
        kmin = x < kmin ? x : kmin; kmax = y > kmax ? y : kmax;
    }
    if (lane == 0) { sc.red[warp] = kmin; sc.red[32 + warp] = kmax; }
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < SMALL_THREADS / 32; ++w) { kmin = sc.red[w] < kmin ? sc.red[w] : kmin; kmax = sc.red[32 + w] > kmax ? sc.red[32 + w] : kmax; }
        sc.items[0] = SmallItem{kmin, kmax, 0ull, (unsigned int)n, 0, nt};
        sc.n_items = 1;
        sc.bad = 0;
    }
    __syncthreads();
    for (int head = 0; head < sc.n_items; ++head) {
        const SmallItem it = sc.items[head];
        const Key lo = (Key)it.lo, hi = (Key)it.hi;
        if (lo == hi) {
            if (tid < it.nt) out[it.t0 + tid] = it.lo;
            continue;
        }
        if (it.pop <= (unsigned int)SMALL_COLLECT) {
            if (tid == 0) sc.n_collected = 0u;
            __syncthreads();
            for (int i = tid; i < n; i += SMALL_THREADS) {
                const Key k = keys[i];
                if (k >= lo && k <= hi) { const unsigned int p = atomicAdd(&sc.n_collected, 1u); if (p < (unsigned int)SMALL_COLLECT) sc.sort[p] = k; }
            }
            __syncthreads();
            if (sc.n_collected != it.pop) return false;
            int Pw = 1;
            while (Pw < (int)it.pop) Pw <<= 1;
            for (int i = (int)it.pop + tid; i < Pw; i += SMALL_THREADS) sc.sort[i] = ~(Key)0;
            __syncthreads();
            small_bitonic_sort(sc.sort, Pw);
            if (tid < it.nt) out[it.t0 + tid] = (unsigned long long)sc.sort[ranks[it.t0 + tid] - it.before];
            __syncthreads();
            continue;
        }
        const int shift = max(0, bit_length_u64((unsigned long long)(hi - lo)) - 12);
        for (int i = tid; i < SMALL_SUB; i += SMALL_THREADS) sc.sub[i] = 0u;
        __syncthreads();
        for (int i = tid; i < n; i += SMALL_THREADS) {
            const Key k = keys[i];
            if (k >= lo && k <= hi) atomicAdd(&sc.sub[(unsigned int)((k - lo) >> shift)], 1u);
        }
        __syncthreads();
        unsigned int c4[4], run = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { run += sc.sub[tid * 4 + k]; c4[k] = run; }
        unsigned int incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned int o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) sc.warp[warp] = incl;
        __syncthreads();
        unsigned int offs = incl - run;
        for (int w = 0; w < warp; ++w) offs += sc.warp[w];
#pragma unroll
        for (int k = 0; k < 4; ++k) sc.sub[tid * 4 + k] = offs + c4[k];
        __syncthreads();
        if (tid == 0) {
            int prev = -1;
            if (sc.sub[SMALL_SUB - 1] != it.pop) sc.bad = 1;
            for (int q = 0; q < it.nt && !sc.bad; ++q) {
                const unsigned long long want = ranks[it.t0 + q] - it.before;
                int l2 = 0, h2 = SMALL_SUB - 1;
                while (l2 < h2) { const int mid = (l2 + h2) >> 1; if ((unsigned long long)sc.sub[mid] > want) h2 = mid; else l2 = mid + 1; }
                if (l2 == prev) { sc.items[sc.n_items - 1].nt += 1; continue; }
                prev = l2;
                if (sc.n_items >= SMALL_ITEMS) { sc.bad = 1; break; }
                const unsigned int before = l2 ? sc.sub[l2 - 1] : 0u;
                const Key nlo = lo + ((Key)l2 << shift);
                const Key span = shift ? (((Key)1 << shift) - 1) : 0;
                const Key nhi = (hi - nlo < span) ? hi : nlo + span;
                sc.items[sc.n_items++] = SmallItem{(unsigned long long)nlo, (unsigned long long)nhi, it.before + before,
                                                   sc.sub[l2] - before, it.t0 + q, 1};
            }
        }
        __syncthreads();
        if (sc.bad) return false;
    }
    __syncthreads();
    return true;
}

template <typename Real>
__global__ void __launch_bounds__(SMALL_THREADS) small_rows_kernel(const SmallArgs a) {
    using Key = typename KeyT<Real>::type;
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    Key* s_keys = reinterpret_cast<Key*>(smem_dyn);                                  // [P] (sorted path) or [n]
    __shared__ double s_pct[MC_MAX_PCTS];
    __shared__ double s_red[32];
    __shared__ unsigned long long s_redc[32];
    __shared__ unsigned long long s_key_f[MAX_TARGETS], s_key_t[MAX_TARGETS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int row = blockIdx.x;
    const int n = (int)a.n;
    const bool sorted = n <= SMALL_SORT_MAX;
    const bool last = row == a.rows - 1;
    int P = 1;
    while (P < n) P <<= 1;
    const Real* x = static_cast<const Real*>(a.paths) + (size_t)row * a.ld;
    const SelectPlan& pf = *a.plan_f;
    const SelectPlan& pt = *a.plan_t;
    bool ok = true;
    if (sorted) {
        for (int i = tid; i < P; i += SMALL_THREADS) s_keys[i] = i < n ? key_of(x[i]) : ~(Key)0;
        __syncthreads();
        small_bitonic_sort(s_keys, P);
        if (tid < pf.n_targets) s_key_f[tid] = (unsigned long long)s_keys[pf.rank[tid]];
        if (last && tid < pt.n_targets) s_key_t[tid] = (unsigned long long)s_keys[pt.rank[tid]];
    } else {
        for (int i = tid; i < n; i += SMALL_THREADS) s_keys[i] = key_of(x[i]);
        __syncthreads();
        size_t off = ((size_t)n * sizeof(Key) + 15) & ~(size_t)15;
        SmallScratch<Key>& sc = *reinterpret_cast<SmallScratch<Key>*>(smem_dyn + off);
        ok = small_select<Key>(s_keys, n, pf.rank, pf.n_targets, s_key_f, sc);
        if (last && ok) ok = small_select<Key>(s_keys, n, pt.rank, a.n_pcts ? pt.n_targets : 0, s_key_t, sc);
    }
    __syncthreads();
    unsigned long long* host_words = reinterpret_cast<unsigned long long*>(a.host_out);
    constexpr int RW = (int)(sizeof(DeviceResult) / 8);
    if (tid < a.n_path_pcts) {              // app.py:643-645
        const double va = SmallValue<Key>::get((Key)s_key_f[pf.pct_lo_idx[tid]]), vb = SmallValue<Key>::get((Key)s_key_f[pf.pct_hi_idx[tid]]);
        const double pc = numpy_lerp(va, vb, pf.pct_gamma[tid]);
        a.fan[(size_t)row * a.n_path_pcts + tid] = pc;
        host_words[RW + (size_t)tid * a.rows + row] = (unsigned long long)__double_as_longlong(pc);    // [pct][row] for the caller
    }
    if (!ok) *reinterpret_cast<volatile unsigned int*>(a.bad) = 1u;
    if (!last) return;
    // ---- the terminal row: percentiles, tail, moments -> result block
    DeviceResult& res = *a.res;
    if (tid < a.n_pcts) {                   // app.py:638-640
        const double va = SmallValue<Key>::get((Key)s_key_t[pt.pct_lo_idx[tid]]), vb = SmallValue<Key>::get((Key)s_key_t[pt.pct_hi_idx[tid]]);
        const double pc = numpy_lerp(va, vb, pt.pct_gamma[tid]);
        s_pct[tid] = pc;
        res.pct[tid] = pc; res.pct_lo[tid] = va; res.pct_hi[tid] = vb;
    }
    __syncthreads();
    // app.py:656: count and sum of the values at or below the interpolated threshold
    const bool want_tail = a.tail_idx >= 0 && a.n_pcts > 0;
    const double thr = want_tail ? s_pct[a.tail_idx] : -INFINITY;
    double tsum = 0.0;
    unsigned long long tcnt = 0;
    if (want_tail)
        for (int i = tid; i < n; i += SMALL_THREADS) { const double v = SmallValue<Key>::get(s_keys[i]); if (v <= thr) { tsum += v; tcnt += 1; } }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) { tsum += __shfl_down_sync(0xffffffffu, tsum, d); tcnt += __shfl_down_sync(0xffffffffu, tcnt, d); }
    if (lane == 0) { s_red[warp] = tsum; s_redc[warp] = tcnt; }
    __syncthreads();
    // moments: the simulation kernel's per-block partials, added in reduce_partials_kernel's order -- thread t takes
    // partials t, t + 1024, ...; then the tree (t, t + w), w = 512 .. 1
    BlockPartial* s_part = reinterpret_cast<BlockPartial*>(smem_dyn);          // the keys are done with
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
        for (int w = 1; w < SMALL_THREADS / 32; ++w) { tsum += s_red[w]; tcnt += s_redc[w]; }
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
        res.tail_count = tcnt;
        res.tail_sum = want_tail ? tsum : 0.0;
        res.hist_lo = 0.0; res.hist_hi = 0.0;
        res.status = 0ull;
    }
    __syncthreads();
    const unsigned long long* res_words = reinterpret_cast<const unsigned long long*>(a.res);
    for (int i = tid; i < RW; i += SMALL_THREADS) host_words[i] = res_words[i];
}

size_t small_rows_smem(int dtype, unsigned long long n) {
    const size_t ksz = dtype == MC_F64 ? 8 : 4;
    size_t P = 1;
    while (P < n) P <<= 1;
    size_t bytes;
    if (n <= (unsigned long long)SMALL_SORT_MAX) bytes = P * ksz;
    else bytes = (((size_t)n * ksz + 15) & ~(size_t)15) + (dtype == MC_F64 ? sizeof(SmallScratch<uint64_t>) : sizeof(SmallScratch<uint32_t>));
    const size_t part = (size_t)SMALL_THREADS * sizeof(BlockPartial);
    return bytes > part ? bytes : part;
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
