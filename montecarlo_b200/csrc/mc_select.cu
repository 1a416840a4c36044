// mc_select.cu -- exact percentiles by pilot-bracketed interval narrowing, CVaR tail sums
// and the fixed-bin histogram, for sm_100a.
//
// Replaces the reference's `np.percentile(final_prices, p)` x9 (app.py:638-640) and
// `np.percentile(paths, p, axis=1)` x5 (app.py:643-645), each of which re-partitions
// its whole input on the host (61-72 % of the reference's run time).
//
// All order statistics of a row are found together, exactly, in (normally) two streaming
// reads of the row:
//   pilot    a small sample of the row (the first columns: paths are iid) is histogrammed
//            and gives, for every percentile, a key bracket that holds its order statistics
//            with probability 1 - 1e-5 (4.5 sigma of the sample-quantile error);
//   pass 1   every value is compared with the bracket bounds IN REGISTERS (exact counts of
//            the values below each bracket); only the few percent inside a bracket touch
//            shared memory (a 1024-bin histogram of the bracket);
//   scan     the target's bin becomes its new interval: 1/1024 of the bracket;
//   pass 2   intervals holding <= 1024 values are collected and sorted -> exact keys;
//            wider ones are histogrammed again (x1024 narrower each time).
// A bracket that missed (or a heavy tie) just costs that row extra passes; the result is
// always the exact order statistic.  Counts, histograms and candidates are the ONLY thing
// that crosses GPUs when paths are sharded, so no path data moves.
//
// Interpolation follows NumPy's method='linear' exactly (lib/_function_base_impl.py
// `_lerp`): a + (b-a)*t, and b - (b-a)*(1-t) when t >= 0.5, in fp64 without FMA
// contraction so results are bit-identical to np.percentile on the same values.
#include <math.h>

#include <atomic>

#include "mc_launch.cuh"
#include "mc_select.cuh"

namespace mc {

constexpr double SEL_Z = 4.5;            // bracket half-width in sample-quantile standard errors
constexpr int PILOT_BIN_BITS = 12;
constexpr int PILOT_BINS = 1 << PILOT_BIN_BITS;

// ------------------------------------------------------------------ plan (host)
bool build_select_plan(const double* pcts, int n_pcts, unsigned long long n_global, int dtype,
                       SelectPlan* plan) {
    if (n_pcts < 0 || n_pcts > (int)MC_MAX_PCTS || n_global == 0) return false;
    SelectPlan& sp = *plan;
    sp = SelectPlan{};
    sp.n_pcts = n_pcts;
    sp.key_bits = dtype == MC_F64 ? 64 : 32;
    sp.n_global = n_global;
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
    auto push = [&](unsigned long long r) {
        for (int j = 0; j < nt; ++j)
            if (sp.rank[j] == r) return;
        sp.rank[nt++] = r;
    };
    for (int k = 0; k < n_pcts; ++k) { push(lo[k]); push(hi[k]); }
    for (int a = 1; a < nt; ++a) {      // insertion sort (nt <= 32)
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

unsigned int sample_slot(int dtype, unsigned int n_ranks) {
    const unsigned int m = dtype == MC_F64 ? SEL_SAMPLE_MAX_F64 : SEL_SAMPLE_MAX;
    const unsigned int s = m / (n_ranks ? n_ranks : 1);
    return s < 64 ? 64 : s;
}

// Single-rank problems read their pilot in place.  Contiguous (iid columns: the first m values of a row): 131072 --
// the brackets come out half as wide as with the 32768 a stage buffer carries, pass 1 sets aside half as many values
// (4.5 % of a row instead of 9 %: less queue / ring / side-list work, and pass 2 reads half as much).  Strided
// (arbitrary data): 32768, one 32-byte sector per sample.
static std::atomic<unsigned int> g_pilot_direct{131072u};
unsigned int pilot_in_place_size(bool contiguous) { return contiguous ? g_pilot_direct.load() : (unsigned int)SEL_SAMPLE_MAX; }
void set_pilot_in_place_size(unsigned int m) { g_pilot_direct.store(m >= 1024u && m <= (1u << 22) ? m : 131072u); }

template <typename Key> __device__ __forceinline__ Key key_max();
template <> __device__ __forceinline__ uint32_t key_max<uint32_t>() { return 0xFFFFFFFFu; }
template <> __device__ __forceinline__ uint64_t key_max<uint64_t>() { return 0xFFFFFFFFFFFFFFFFull; }

__device__ __forceinline__ int bit_length(unsigned long long v) { return v ? 64 - __clzll((long long)v) : 0; }

// ------------------------------------------------------------------ pilot sample
template <typename Real>
__global__ void __launch_bounds__(256) select_sample_kernel(const Real* __restrict__ x, unsigned long long ld,
                                                            unsigned long long n, unsigned int m_local,
                                                            unsigned int m_slot, unsigned long long stride,
                                                            typename KeyT<Real>::type* __restrict__ out) {
    const unsigned long long row = blockIdx.y;
    const Real* xr = x + row * ld;
    for (unsigned int s = blockIdx.x * blockDim.x + threadIdx.x; s < m_local; s += gridDim.x * blockDim.x)
        out[row * m_slot + s] = key_of(xr[(unsigned long long)s * stride]);
}

cudaError_t launch_select_sample(const SelectProblem& sp, void* d_out, cudaStream_t stream) {
    if (sp.m_local == 0 || sp.rows == 0 || sp.pilot_in_place) return cudaSuccess;
    if (sp.rows > 65535) return cudaErrorInvalidValue;
    const unsigned long long stride = sp.sample_contiguous ? 1ull : (sp.n / sp.m_local ? sp.n / sp.m_local : 1ull);
    dim3 grid((sp.m_local + 255) / 256, (unsigned)sp.rows);
    if (grid.x > 32) grid.x = 32;
    if (sp.dtype == MC_F64)
        select_sample_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(sp.x), sp.ld, sp.n, sp.m_local,
                                                               sp.m_slot, stride, static_cast<uint64_t*>(d_out));
    else
        select_sample_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(sp.x), sp.ld, sp.n, sp.m_local,
                                                              sp.m_slot, stride, static_cast<uint32_t*>(d_out));
    return cudaGetLastError();
}

// ------------------------------------------------------------------ brackets from the pilot sample
// One block per row.  The sample (all ranks' contributions) is histogrammed over its own key
// range (4096 linear bins: float keys are log-spaced, so this resolves ~0.1 % of a typical
// row) and each percentile's bracket is rounded OUTWARD to bin edges.
// brackets of one row from the (summed) 4096-bin histogram of its pilot: all 256 threads of the block
template <typename Key>
// `clamp_top` >= 0: samples outside the binned key range were clamped into bin 0 / bin clamp_top, so a bracket that
// starts in bin 0 opens down to key 0 and one that ends in bin clamp_top (or beyond) opens up to the largest key
__device__ void brackets_from_pilot_hist(const SelectPlan* __restrict__ plan, unsigned int* s_hist, unsigned int* s_warp,
                                         unsigned long long* s_lo, unsigned long long* s_hi, unsigned long long kmin,
                                         int sh, unsigned int m_total, RowSel& st, int clamp_top = -1) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // inclusive scan of the 4096 bins, 16 per thread
    unsigned int local[PILOT_BINS / 256], run = 0;
#pragma unroll
    for (int k = 0; k < PILOT_BINS / 256; ++k) { run += s_hist[tid * (PILOT_BINS / 256) + k]; local[k] = run; }
    unsigned int incl = run;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned int o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned int off = incl - run;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
#pragma unroll
    for (int k = 0; k < PILOT_BINS / 256; ++k) s_hist[tid * (PILOT_BINS / 256) + k] = off + local[k];
    __syncthreads();

    const int np_ = plan->n_pcts;
    const unsigned long long KMAX = (unsigned long long)key_max<Key>();
    if (tid < np_) {
        const double m = (double)m_total, ng = (double)plan->n_global;
        const double qa = (double)plan->rank[plan->pct_lo_idx[tid]] / ng;
        const double qb = ((double)plan->rank[plan->pct_hi_idx[tid]] + 1.0) / ng;
        const double fa = floor(qa * m - SEL_Z * sqrt(m * qa * (1.0 - qa)) - 2.0);
        const double fb = ceil(qb * m + SEL_Z * sqrt(m * qb * (1.0 - qb)) + 2.0);
        unsigned long long lo = 0ull, hi = KMAX;
        auto bin_of_rank = [&](unsigned int r) {       // first bin with cum > r
            int a = 0, b = PILOT_BINS - 1;
            while (a < b) { const int mid = (a + b) >> 1; if (s_hist[mid] > r) b = mid; else a = mid + 1; }
            return a;
        };
        if (fa >= 0.0 && m_total) {
            const int b = bin_of_rank((unsigned int)fa);
            if (!(clamp_top >= 0 && b == 0)) lo = kmin + ((unsigned long long)b << sh);
        }
        if (fb < m && m_total) {
            const int b = bin_of_rank((unsigned int)fb);
            if (!(clamp_top >= 0 && b >= clamp_top)) {
                const unsigned long long span = ((unsigned long long)(b + 1) << sh) - 1ull;
                hi = span > KMAX - kmin ? KMAX : kmin + span;
            }
        }
        s_lo[tid] = lo; s_hi[tid] = hi;
    }
    __syncthreads();
    if (tid == 0) {
        int iv_of_pct[MC_MAX_PCTS];
        int n_iv = 0;
        for (int k = 0; k < np_; ++k) {
            if (n_iv && s_lo[k] <= st.hi[n_iv - 1]) {             // overlaps the previous bracket: merge
                if (s_hi[k] > st.hi[n_iv - 1]) st.hi[n_iv - 1] = s_hi[k];
            } else {
                st.lo[n_iv] = s_lo[k]; st.hi[n_iv] = s_hi[k];
                ++n_iv;
            }
            iv_of_pct[k] = n_iv - 1;
        }
        for (int gix = 0; gix < n_iv; ++gix) {
            st.shift[gix] = max(0, bit_length(st.hi[gix] - st.lo[gix]) - SEL_BIN_BITS);
            st.mode[gix] = IV_HIST;
            st.below[gix] = 0ull;
        }
        for (int k = 0; k < np_; ++k) {
            st.target_iv[plan->pct_lo_idx[k]] = iv_of_pct[k];
            st.target_iv[plan->pct_hi_idx[k]] = iv_of_pct[k];
        }
        st.n_iv = n_iv;
        st.below_known = 0;
    }
}

template <typename Key>
__global__ void __launch_bounds__(256) select_bracket_kernel(const SelectPlan* __restrict__ plan,
                                                             const unsigned long long* __restrict__ g,
                                                             size_t sample_off_words, int hdr_word,
                                                             unsigned int n_ranks, size_t stride_words,
                                                             unsigned int m_slot, RowSel* __restrict__ states) {
    __shared__ unsigned int s_hist[PILOT_BINS];
    __shared__ unsigned long long s_min[8], s_max[8];
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned long long s_lo[MC_MAX_PCTS], s_hi[MC_MAX_PCTS];
    const unsigned long long row = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    unsigned long long kmin = ~0ull, kmax = 0ull;
    unsigned int m_total = 0;
    for (unsigned int r = 0; r < n_ranks; ++r) {
        const unsigned int m_r = (unsigned int)g[r * stride_words + hdr_word];
        const Key* src = reinterpret_cast<const Key*>(g + r * stride_words + sample_off_words) + row * m_slot;
        for (unsigned int i = tid; i < m_r; i += 256) {
            const unsigned long long k = src[i];
            kmin = k < kmin ? k : kmin;
            kmax = k > kmax ? k : kmax;
        }
        m_total += m_r;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const unsigned long long a = __shfl_down_sync(0xffffffffu, kmin, d), b = __shfl_down_sync(0xffffffffu, kmax, d);
        kmin = a < kmin ? a : kmin;
        kmax = b > kmax ? b : kmax;
    }
    if (lane == 0) { s_min[warp] = kmin; s_max[warp] = kmax; }
    for (int i = tid; i < PILOT_BINS; i += 256) s_hist[i] = 0u;
    __syncthreads();
    kmin = s_min[0]; kmax = s_max[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) { kmin = s_min[w] < kmin ? s_min[w] : kmin; kmax = s_max[w] > kmax ? s_max[w] : kmax; }
    if (m_total == 0) { kmin = 0; kmax = 0; }
    const int sh = max(0, bit_length(kmax - kmin) - PILOT_BIN_BITS);
    for (unsigned int r = 0; r < n_ranks; ++r) {
        const unsigned int m_r = (unsigned int)g[r * stride_words + hdr_word];
        const Key* src = reinterpret_cast<const Key*>(g + r * stride_words + sample_off_words) + row * m_slot;
        for (unsigned int i = tid; i < m_r; i += 256) atomicAdd(&s_hist[(unsigned int)(((unsigned long long)src[i] - kmin) >> sh)], 1u);
    }
    __syncthreads();
    brackets_from_pilot_hist<Key>(plan, s_hist, s_warp, s_lo, s_hi, kmin, sh, m_total, states[row]);
}

// the same from a pilot read IN PLACE: the first m values of the row (stride 1) or every stride-th value.
// ONE_PASS (iid columns: the first values of a row are a random sample of it): the binned key range comes from the
// first 8192 samples, widened by a quarter of its width on either side; the few samples beyond it are clamped into the
// end bins, which then open to the ends of the key range.  The pilot is read once (131072 samples x 1025 rows do not
// fit the L2: the two-pass form read 1.07 GB from DRAM and spent 225 us, ncu `profiles/r02/ncu_full_brc5.txt`).
constexpr unsigned int PILOT_RANGE_SAMPLES = 8192;

template <typename Real, bool ONE_PASS>
__global__ void __launch_bounds__(256) select_bracket_in_place_kernel(const SelectPlan* __restrict__ plan,
                                                                      const Real* __restrict__ x, unsigned long long ld,
                                                                      unsigned int m, unsigned long long stride,
                                                                      RowSel* __restrict__ states) {
    using Key = typename KeyT<Real>::type;
    __shared__ unsigned int s_hist[PILOT_BINS];
    __shared__ Key s_min[8], s_max[8];
    __shared__ unsigned int s_warp[8];
    __shared__ unsigned long long s_lo[MC_MAX_PCTS], s_hi[MC_MAX_PCTS];
    const unsigned long long row = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const Real* xr = x + row * ld;
    constexpr int PU = 16;                               // loads in flight per thread: the block is alone with its row's latency
    const unsigned int m_range = ONE_PASS ? (m < PILOT_RANGE_SAMPLES ? m : PILOT_RANGE_SAMPLES) : m;
    Key kmin = key_max<Key>(), kmax = 0;
    unsigned int i = tid;
    for (; i + (PU - 1) * 256 < m_range; i += PU * 256) {
        Real v[PU];
#pragma unroll
        for (int u = 0; u < PU; ++u) v[u] = xr[(unsigned long long)(i + u * 256) * stride];
#pragma unroll
        for (int u = 0; u < PU; ++u) {
            const Key k = key_of(v[u]);
            kmin = k < kmin ? k : kmin;
            kmax = k > kmax ? k : kmax;
        }
    }
    for (; i < m_range; i += 256) {
        const Key k = key_of(xr[(unsigned long long)i * stride]);
        kmin = k < kmin ? k : kmin;
        kmax = k > kmax ? k : kmax;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        const Key a = __shfl_down_sync(0xffffffffu, kmin, d), b = __shfl_down_sync(0xffffffffu, kmax, d);
        kmin = a < kmin ? a : kmin;
        kmax = b > kmax ? b : kmax;
    }
    if (lane == 0) { s_min[warp] = kmin; s_max[warp] = kmax; }
    for (int k = tid; k < PILOT_BINS; k += 256) s_hist[k] = 0u;
    __syncthreads();
    kmin = s_min[0]; kmax = s_max[0];
#pragma unroll
    for (int w = 1; w < 8; ++w) { kmin = s_min[w] < kmin ? s_min[w] : kmin; kmax = s_max[w] > kmax ? s_max[w] : kmax; }
    if (m == 0) { kmin = 0; kmax = 0; }
    int clamp_top = -1;
    if (ONE_PASS && m > m_range) {       // widen, stay inside the key type
        const Key quarter = (Key)((kmax - kmin) >> 2) + 1;
        kmin = kmin > quarter ? kmin - quarter : 0;
        kmax = key_max<Key>() - kmax > quarter ? kmax + quarter : key_max<Key>();
    }
    const int sh = max(0, bit_length((unsigned long long)(kmax - kmin)) - PILOT_BIN_BITS);
    if (ONE_PASS && m > m_range) clamp_top = (int)((unsigned long long)(kmax - kmin) >> sh);
    auto bin_of = [&](Real v) {
        Key k = key_of(v);
        if (ONE_PASS) { k = k < kmin ? kmin : k; k = k > kmax ? kmax : k; }
        return (unsigned int)((Key)(k - kmin) >> sh);
    };
    i = tid;
    if (kmin == kmax) {          // a constant row (row 0 of the paths is all s0, app.py:613): every sample in bin 0
        if (tid == 0) s_hist[0] = m;
        i = m;
        clamp_top = -1;
    }
    for (; i + (PU - 1) * 256 < m; i += PU * 256) {
        Real v[PU];
#pragma unroll
        for (int u = 0; u < PU; ++u) v[u] = xr[(unsigned long long)(i + u * 256) * stride];
#pragma unroll
        for (int u = 0; u < PU; ++u) atomicAdd(&s_hist[bin_of(v[u])], 1u);
    }
    for (; i < m; i += 256) atomicAdd(&s_hist[bin_of(xr[(unsigned long long)i * stride])], 1u);
    __syncthreads();
    brackets_from_pilot_hist<Key>(plan, s_hist, s_warp, s_lo, s_hi, (unsigned long long)kmin, sh, m, states[row], clamp_top);
}

cudaError_t launch_select_brackets(const SelectProblem& sp, const unsigned long long* d_gathered,
                                   size_t sample_off_words, int hdr_word, unsigned int n_ranks,
                                   size_t rank_stride_words, int /*max_smem_optin*/, cudaStream_t stream) {
    if (sp.plan.n_targets == 0 || sp.rows == 0) return cudaSuccess;
    if (sp.pilot_in_place) {
        const unsigned long long stride = sp.sample_contiguous ? 1ull : (sp.m_local && sp.n / sp.m_local ? sp.n / sp.m_local : 1ull);
        const bool one_pass = sp.sample_contiguous != 0;
        if (sp.dtype == MC_F64) {
            auto k = one_pass ? select_bracket_in_place_kernel<double, true> : select_bracket_in_place_kernel<double, false>;
            k<<<(unsigned)sp.rows, 256, 0, stream>>>(sp.d_plan, static_cast<const double*>(sp.x), sp.ld, sp.m_local, stride, sp.d_states);
        } else {
            auto k = one_pass ? select_bracket_in_place_kernel<float, true> : select_bracket_in_place_kernel<float, false>;
            k<<<(unsigned)sp.rows, 256, 0, stream>>>(sp.d_plan, static_cast<const float*>(sp.x), sp.ld, sp.m_local, stride, sp.d_states);
        }
        return cudaGetLastError();
    }
    if (sp.dtype == MC_F64)
        select_bracket_kernel<uint64_t><<<(unsigned)sp.rows, 256, 0, stream>>>(sp.d_plan, d_gathered, sample_off_words,
                                                                               hdr_word, n_ranks, rank_stride_words,
                                                                               sp.m_slot, sp.d_states);
    else
        select_bracket_kernel<uint32_t><<<(unsigned)sp.rows, 256, 0, stream>>>(sp.d_plan, d_gathered, sample_off_words,
                                                                               hdr_word, n_ranks, rank_stride_words,
                                                                               sp.m_slot, sp.d_states);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ narrowing pass
// grid (chunks, rows), 256 threads.  NIV > 0: first pass -- the (<= NIV) bracket lower bounds
// live in registers and every value updates `ge[g] += key >= lo[g]` (2 instructions per
// bracket, no memory); the same predicates give the candidate bracket.  NIV == 0: later
// passes -- `below` is already known, intervals are found by a shared-memory binary search.
constexpr int PASS_UNROLL = 8;

template <typename Key, int NIV>
struct PassCtx {
    const Key* s_lo; const Key* s_hi; const int* s_shift; const int* s_mode;
    unsigned int* s_bins; unsigned int* out_iv; size_t iv_slot; int n_iv;
    Key lo_r[NIV > 0 ? NIV : 1];
    unsigned int ge[NIV > 0 ? NIV : 1];
};

template <typename Key, int NIV>
__device__ __forceinline__ void pass_one(PassCtx<Key, NIV>& c, Key key) {
    int j = 0;
    if (NIV > 0) {
#pragma unroll
        for (int g = 0; g < NIV; ++g) {
            const unsigned int p = key >= c.lo_r[g] ? 1u : 0u;
            c.ge[g] += p;
            j += (int)p;
        }
    } else {
        int a = 0, b = c.n_iv;                 // j = #(lo[g] <= key), lo sorted ascending
        while (a < b) { const int mid = (a + b) >> 1; if (c.s_lo[mid] <= key) a = mid + 1; else b = mid; }
        j = a;
    }
    const int g = j - 1;
    if (g >= 0 && g < c.n_iv && key <= c.s_hi[g]) {
        const Key off = key - c.s_lo[g];
        if (c.s_mode[g] == IV_HIST) {
            atomicAdd(&c.s_bins[g * SEL_BINS + (unsigned int)(off >> c.s_shift[g])], 1u);
        } else {
            unsigned int* slot = c.out_iv + (size_t)g * c.iv_slot;
            const unsigned int idx = atomicAdd(slot, 1u);
            if (idx < (unsigned int)SEL_CAND) reinterpret_cast<Key*>(slot + 2)[idx] = key;
        }
    }
}

template <typename Real, int NIV>
__global__ void __launch_bounds__(256) select_pass_kernel(const Real* __restrict__ x, unsigned long long ld,
                                                          unsigned long long n, const RowSel* __restrict__ states,
                                                          unsigned int* __restrict__ out, size_t row_slot,
                                                          size_t iv_slot, int n_iv_cap) {
    using Key = typename KeyT<Real>::type;
    extern __shared__ unsigned int s_bins[];
    __shared__ Key s_lo[MAX_IV], s_hi[MAX_IV];
    __shared__ int s_shift[MAX_IV], s_mode[MAX_IV];
    const unsigned long long row = blockIdx.y;
    const RowSel& st = states[row];
    const int n_iv = st.n_iv;
    if (n_iv == 0) return;                                   // every target of this row is decided
    const int tid = threadIdx.x;
    if (tid < n_iv) {
        s_lo[tid] = (Key)st.lo[tid]; s_hi[tid] = (Key)st.hi[tid];
        s_shift[tid] = st.shift[tid]; s_mode[tid] = st.mode[tid];
    }
    for (int i = tid; i < n_iv * SEL_BINS; i += 256) s_bins[i] = 0u;
    __syncthreads();

    unsigned int* out_row = out + row * row_slot;
    PassCtx<Key, NIV> c;
    c.s_lo = s_lo; c.s_hi = s_hi; c.s_shift = s_shift; c.s_mode = s_mode; c.s_bins = s_bins;
    c.out_iv = out_row + 2 * (size_t)n_iv_cap; c.iv_slot = iv_slot; c.n_iv = n_iv;
    if (NIV > 0) {
#pragma unroll
        for (int g = 0; g < (NIV > 0 ? NIV : 1); ++g) { c.lo_r[g] = g < n_iv ? s_lo[g] : key_max<Key>(); c.ge[g] = 0u; }
    }
    const Real* xr = x + row * ld;
    const unsigned long long per = (n + gridDim.x - 1) / gridDim.x;
    const unsigned long long beg = (unsigned long long)blockIdx.x * per;
    const unsigned long long end = beg + per < n ? beg + per : n;
    unsigned long long i = beg + tid;
    for (; i + (PASS_UNROLL - 1) * 256ull < end; i += PASS_UNROLL * 256ull) {
        Real v[PASS_UNROLL];
#pragma unroll
        for (int u = 0; u < PASS_UNROLL; ++u) v[u] = xr[i + u * 256ull];
#pragma unroll
        for (int u = 0; u < PASS_UNROLL; ++u) pass_one<Key, NIV>(c, key_of(v[u]));
    }
    for (; i < end; i += 256ull) pass_one<Key, NIV>(c, key_of(xr[i]));

    if (NIV > 0) {
        unsigned long long* out_ge = reinterpret_cast<unsigned long long*>(out_row);
#pragma unroll
        for (int g = 0; g < (NIV > 0 ? NIV : 1); ++g) {
            unsigned int v = c.ge[g];
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) v += __shfl_down_sync(0xffffffffu, v, d);
            if ((tid & 31) == 0 && g < n_iv && v) atomicAdd(&out_ge[g], (unsigned long long)v);
        }
    }
    __syncthreads();
    for (int i2 = tid; i2 < n_iv * SEL_BINS; i2 += 256) {
        const unsigned int cnt = s_bins[i2];
        if (cnt) atomicAdd(&c.out_iv[(size_t)(i2 >> SEL_BIN_BITS) * iv_slot + 2 + (i2 & (SEL_BINS - 1))], cnt);
    }
}

// ------------------------------------------------------------------ fp32 narrowing passes
// The generic pass above spends ~40 issue slots and one shared-memory read per value (2.0 TB/s on the
// 17 GB fan-chart problem of BASELINE config 5).  fp32 rows get two leaner kernels whose hot loop has
// no memory operand besides the streaming load:
//   pass 1  bracket bounds live in registers AS FLOATS.  Per value and bracket: one compare, one
//           predicated count (the exact population at or above the bracket's lower bound) and one select
//           that carries the upper bound of the last bracket entered; `v <= that bound` then means "inside
//           a bracket" and only those values (a few percent) touch shared memory.  IEEE order equals key
//           order because key_of folds -0.0 onto +0.0.
//   pass 2  `below` is known and the intervals are 1/1024 of a bracket wide: a byte map over <= 2048
//           coarse key bins (one LDS.U8) rejects ~98 % of the values; the rest take the exact path.
constexpr int FAST_THREADS = 256;
constexpr int MAP_BITS_LOG2 = 14;                 // pass 2: a bit per hashed key bin
constexpr int MAP_BITS = 1 << MAP_BITS_LOG2;
constexpr int MAP_WORDS = MAP_BITS / 32;
constexpr uint32_t MAP_FILL_MAX = 2048;           // an interval wider than this many bins switches the filter off

struct FastRow {
    const uint32_t* s_lo; const uint32_t* s_hi; const int* s_shift; const int* s_mode;
    unsigned int* s_bins; unsigned int* out_iv; size_t iv_slot; int n_iv;
};

// exact treatment of one value that may lie inside an interval
__device__ __forceinline__ void fast_member(const FastRow& c, uint32_t key) {
    int a = 0, b = c.n_iv;                     // j = #(lo[g] <= key), lo ascending
    while (a < b) { const int mid = (a + b) >> 1; if (c.s_lo[mid] <= key) a = mid + 1; else b = mid; }
    const int g = a - 1;
    if (g < 0 || key > c.s_hi[g]) return;
    const uint32_t off = key - c.s_lo[g];
    if (c.s_mode[g] == IV_HIST) {
        atomicAdd(&c.s_bins[g * SEL_BINS + (off >> c.s_shift[g])], 1u);
    } else {
        unsigned int* slot = c.out_iv + (size_t)g * c.iv_slot;
        const unsigned int idx = atomicAdd(slot, 1u);
        if (idx < (unsigned int)SEL_CAND) slot[2 + idx] = key;
    }
}

// stream one block's share [beg, end) of a row through f(value), calling batch_done() after every 16
// values of a thread (warp-converged); 16-byte loads when the row allows it
template <bool PREFETCH, typename F, typename G>
__device__ __forceinline__ void fast_stream(const float* __restrict__ xr, unsigned long long beg, unsigned long long end,
                                            bool vec, F f, G batch_done) {
    const int tid = threadIdx.x;
    if (vec) {       // beg is a multiple of 4 and xr + beg is 16-byte aligned
        const float4* x4 = reinterpret_cast<const float4*>(xr + beg);
        const unsigned long long n4 = (end - beg) >> 2;
        const unsigned long long n4_even = n4 - n4 % (4ull * FAST_THREADS);     // whole batches: no lane idles
        unsigned long long i = tid;
        // the loads of the next batch are issued before this batch is worked on: a warp always has 64 bytes per
        // lane in flight (without this the pass stalls on its own loads at 50 % occupancy, ncu: long scoreboard)
        float4 q[4];
        if (PREFETCH && i < n4_even) {
#pragma unroll
            for (int u = 0; u < 4; ++u) q[u] = __ldcs(x4 + i + (unsigned long long)u * FAST_THREADS);
        }
        while (i < n4_even) {
            const unsigned long long nxt = i + 4ull * FAST_THREADS;
            float4 qn[4];
            if (!PREFETCH) {
#pragma unroll
                for (int u = 0; u < 4; ++u) q[u] = __ldcs(x4 + i + (unsigned long long)u * FAST_THREADS);
            } else if (nxt < n4_even) {
#pragma unroll
                for (int u = 0; u < 4; ++u) qn[u] = __ldcs(x4 + nxt + (unsigned long long)u * FAST_THREADS);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if constexpr (PREFETCH) f(q[u]);       // the functor of the prefetching form also takes four values at once
                else { f(q[u].x); f(q[u].y); f(q[u].z); f(q[u].w); }
            }
            batch_done();
            if (PREFETCH) {
#pragma unroll
                for (int u = 0; u < 4; ++u) q[u] = qn[u];
            }
            i = nxt;
        }
        for (; i < n4; i += FAST_THREADS) { const float4 q = __ldcs(x4 + i); f(q.x); f(q.y); f(q.z); f(q.w); }
        for (unsigned long long k = beg + (n4 << 2) + tid; k < end; k += FAST_THREADS) f(xr[k]);
    } else {
        const unsigned long long n_even = beg + (end - beg) / (16ull * FAST_THREADS) * (16ull * FAST_THREADS);
        unsigned long long i = beg + tid;
        for (; i < n_even; i += 16ull * FAST_THREADS) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float q[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) q[u] = xr[i + (unsigned long long)(8 * h + u) * FAST_THREADS];
#pragma unroll
                for (int u = 0; u < 8; ++u) f(q[u]);
            }
            batch_done();
        }
        for (; i < end; i += FAST_THREADS) f(xr[i]);
    }
    __syncwarp();
    batch_done();       // at most 3 x 4 + 1 (or 15) values per thread since the last whole batch
}

__device__ __forceinline__ void fast_block_range(unsigned long long n, unsigned long long* beg, unsigned long long* end) {
    unsigned long long per = (n + gridDim.x - 1) / gridDim.x;
    per = (per + 1023ull) & ~1023ull;                         // keeps every block's start 16-byte aligned
    *beg = (unsigned long long)blockIdx.x * per;
    *end = *beg + per < n ? *beg + per : n;
    if (*beg > n) *beg = n;
}

constexpr int QUEUE_DEPTH = 20;      // >= values a thread sees between two drains (16 per batch; <= 17 in the ragged end)

template <int NB>
__global__ void __launch_bounds__(FAST_THREADS) select_pass1_f32_kernel(const float* __restrict__ x, unsigned long long ld,
                                                                        unsigned long long n, int vec,
                                                                        const RowSel* __restrict__ states,
                                                                        unsigned int* __restrict__ out, size_t row_slot,
                                                                        size_t iv_slot, int n_iv_cap, SideLists side) {
    extern __shared__ unsigned int s_bins[];
    __shared__ uint32_t s_lo[MAX_IV];
    __shared__ int s_shift[MAX_IV];
    // Values inside a bracket (a few percent per bracket) are parked in the thread's own column of this
    // queue by PREDICATED stores -- the streaming loop has no branch -- and histogrammed after each batch,
    // when most lanes of a warp have something to do (doing it in place would send ~every warp down the
    // slow path for every value: 12.5 ms instead of 8.5 ms on the config-5 fan chart, measured).
    __shared__ float s_queue[QUEUE_DEPTH * FAST_THREADS];
    const unsigned long long row = blockIdx.y;
    const RowSel& st = states[row];
    const int n_iv = st.n_iv;
    if (n_iv == 0) return;
    const int tid = threadIdx.x;
    if (tid < n_iv) { s_lo[tid] = (uint32_t)st.lo[tid]; s_shift[tid] = st.shift[tid]; }
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) s_bins[i] = 0u;
    // side list of this warp: the parked values, in the order they are drained
    const size_t side_slot = (row * gridDim.x + blockIdx.x) * SIDE_WARPS + (tid >> 5);
    float* side_w = side.values ? side.values + side_slot * side.cap : nullptr;
    unsigned int side_n = 0;                 // warp-uniform
    if (side.values && blockIdx.x == 0) {    // remember the brackets the lists cover
        uint32_t* br = side.brackets + row * SIDE_BRACKET_WORDS;
        if (tid == 0) br[0] = (uint32_t)n_iv;
        if (tid < n_iv) { br[1 + tid] = (uint32_t)st.lo[tid]; br[1 + MAX_IV + tid] = (uint32_t)st.hi[tid]; }
    }
    __syncthreads();

    unsigned int* out_row = out + row * row_slot;
    unsigned int* out_iv = out_row + 2 * (size_t)n_iv_cap;
    float lo_f[NB], hi_f[NB], cnt[NB];      // per-thread counts stay below 2^24: exact in fp32
#pragma unroll
    for (int g = 0; g < NB; ++g) {
        const bool used = g < n_iv;
        const uint32_t kl = used ? (uint32_t)st.lo[g] : 0u, kh = used ? (uint32_t)st.hi[g] : 0u;
        lo_f[g] = used ? value_of(kl < KEY32_NEG_INF ? KEY32_NEG_INF : kl) : __int_as_float(0x7FC00000);   // NaN: never >=
        hi_f[g] = value_of(kh > KEY32_POS_INF ? KEY32_POS_INF : (kh < KEY32_NEG_INF ? KEY32_NEG_INF : kh));
        cnt[g] = 0.0f;
    }
    const uint32_t q_base = (uint32_t)__cvta_generic_to_shared(s_queue) + 4u * tid;
    uint32_t q_top = q_base;                 // byte address of this thread's next free queue slot
    unsigned long long beg, end;
    fast_block_range(n, &beg, &end);
    fast_stream<false>(x + row * ld, beg, end, vec != 0,
        [&](float v) {
            float top = -INFINITY;           // upper bound of the last bracket whose lower bound v reached
#pragma unroll
            for (int g = 0; g < NB; ++g) {
                // compare once; count on the FMA pipe (predicated FADD), carry the bound on the ALU pipe (FSEL)
                asm("{\n\t.reg .pred p;\n\tsetp.ge.f32 p, %2, %3;\n\t@p add.f32 %0, %0, 0f3F800000;\n\tselp.f32 %1, %4, %1, p;\n\t}"
                    : "+f"(cnt[g]), "+f"(top) : "f"(v), "f"(lo_f[g]), "f"(hi_f[g]));
            }
            asm volatile("{\n\t.reg .pred q;\n\tsetp.le.f32 q, %1, %2;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %3;\n\t}"
                         : "+r"(q_top) : "f"(v), "f"(top), "n"(4 * FAST_THREADS) : "memory");
        },
        [&]() {
            for (;;) {
                const bool has = q_top != q_base;
                const unsigned int who = __ballot_sync(0xffffffffu, has);
                if (!who) break;
                if (has) {
                    q_top -= 4u * FAST_THREADS;
                    float v;
                    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(q_top) : "memory");
                    int g = -1;
#pragma unroll
                    for (int k = 0; k < NB; ++k) g += v >= lo_f[k] ? 1 : 0;
                    if (g >= 0)   // always: the value was parked because it lies in bracket g
                        atomicAdd(&s_bins[g * SEL_BINS + ((key_of(v) - s_lo[g]) >> s_shift[g])], 1u);
                    if (side_w) {
                        const unsigned int pos = side_n + __popc(who & ((1u << (tid & 31)) - 1u));
                        if (pos < side.cap) side_w[pos] = v;
                    }
                }
                side_n += __popc(who);
            }
        });
    if (side_w && (tid & 31) == 0) side.counts[side_slot] = side_n;       // > cap: overflowed, not usable

    unsigned long long* out_ge = reinterpret_cast<unsigned long long*>(out_row);
#pragma unroll
    for (int g = 0; g < NB; ++g) {
        const unsigned int total = __reduce_add_sync(0xffffffffu, (unsigned int)cnt[g]);
        if ((tid & 31) == 0 && g < n_iv && total) atomicAdd(&out_ge[g], (unsigned long long)total);
    }
    __syncthreads();
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) {
        const unsigned int v = s_bins[i];
        if (v) atomicAdd(&out_iv[(size_t)(i >> SEL_BIN_BITS) * iv_slot + 2 + (i & (SEL_BINS - 1))], v);
    }
}

// Pass 1 for <= 5 brackets (the fan chart, app.py:643-645) with the classification of a value taken from a
// table instead of a compare chain.  The brackets of a row are sorted and disjoint, so a value lies in one
// of n_iv + 1 GAPS between them or inside a bracket.  q(v) = rn(2047 sat(v qs + qb)) is monotone in v (two
// correctly rounded FMAs and a clamp), so with q evaluated at the bracket bounds by the very same
// instructions, every table bin that lies strictly between q(hi[j-1]) and q(lo[j]) holds values of gap j
// only -- whatever the rounding did.  Those bins carry a one-hot 5-bit field (a thread sees <= 16 values
// between two flushes); all other bins carry 0 and their values (in a bracket, or in one of the <= 2 n_iv
// bins that straddle a bound) take the exact path through the queue.  Per value: 2 FFMA, LEA, LDS, IADD and the
// predicated park -- against 3 instructions PER BRACKET before.
constexpr int LUT_BINS = 2048;
constexpr float LUT_TOP_F = (float)(LUT_BINS - 1);

// byte address of a value's table word, straight out of two FMAs: t = sat(v * qs + qb) in [0, 1] (NaN -> 0), then
//     u = t * ((LUT_BINS - 1) * sizeof(Word)) + (2^23 + table address)
// has  table address + rn(t * (LUT_BINS - 1) * sizeof(Word))  in its mantissa; one AND drops the exponent and the low
// bits, which leaves the address of word floor(rn(.) / sizeof(Word)) -- monotone in v.  (The earlier form put the bin
// number into the mantissa and needed a shift-add and an add per value to make an address of it: 20.8 -> 19.8
// instructions per value.  The pass is bound by shared-memory wavefronts -- 3.4 per table lookup, 78 % of the pipe,
// profiles/r02/ncu_full_p1c5.txt -- so the time did not move.)
struct LutAddr {
    float qs, qb, scale, kb;          // kb = 2^23 + shared-window address of the table (exact: the address is < 2^18)
    uint32_t mask, base;
};
template <typename Word>
__device__ __forceinline__ LutAddr make_lut_addr(float qs, float qb, uint32_t table_addr) {
    LutAddr a;
    a.qs = qs; a.qb = qb;
    a.scale = LUT_TOP_F * (float)sizeof(Word);
    a.kb = 8388608.0f + (float)table_addr;
    a.mask = 0x007FFFFFu & ~(uint32_t)(sizeof(Word) - 1);
    a.base = table_addr;
    return a;
}
__device__ __forceinline__ uint32_t lut_addr(const LutAddr& a, float v) {
    float t, u;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(t) : "f"(v), "f"(a.qs), "f"(a.qb));
    asm("fma.rn.f32 %0, %1, %2, %3;" : "=f"(u) : "f"(t), "f"(a.scale), "f"(a.kb));
    return __float_as_uint(u) & a.mask;
}
// two values: the clamp FMAs stay scalar (no .sat on the packed form), the second FMA is one FFMA2 -- each half the
// same correctly rounded FMA as in lut_addr, so the words are the ones the table was built for
__device__ __forceinline__ void lut_addr2(const LutAddr& a, float v0, float v1, uint32_t& a0, uint32_t& a1) {
    float t0, t1;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(t0) : "f"(v0), "f"(a.qs), "f"(a.qb));
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(t1) : "f"(v1), "f"(a.qs), "f"(a.qb));
    unsigned long long t2, m2, k2, u2;
    asm("mov.b64 %0, {%1, %2};" : "=l"(t2) : "f"(t0), "f"(t1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(a.scale));
    asm("mov.b64 %0, {%1, %1};" : "=l"(k2) : "f"(a.kb));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(u2) : "l"(t2), "l"(m2), "l"(k2));
    uint32_t b0, b1;
    asm("mov.b64 {%0, %1}, %2;" : "=r"(b0), "=r"(b1) : "l"(u2));
    a0 = b0 & a.mask; a1 = b1 & a.mask;
}

template <int NB> struct LutWord { using type = uint32_t; };
template <> struct LutWord<9> { using type = uint64_t; };
template <typename W> __device__ __forceinline__ W lut_word(uint32_t addr);
template <> __device__ __forceinline__ uint32_t lut_word<uint32_t>(uint32_t addr) {
    uint32_t w;
    asm("ld.shared.u32 %0, [%1];" : "=r"(w) : "r"(addr));
    return w;
}
template <> __device__ __forceinline__ uint64_t lut_word<uint64_t>(uint32_t addr) {
    uint64_t w;
    asm("ld.shared.u64 %0, [%1];" : "=l"(w) : "r"(addr));
    return w;
}
// park a value in the thread's queue column when its table word is 0 ("look closer"): predicated, no branch
__device__ __forceinline__ void park1(uint32_t& q_top, float v, uint32_t w) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.eq.u32 q, %2, 0;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %3;\n\t}"
                 : "+r"(q_top) : "f"(v), "r"(w), "n"(4 * FAST_THREADS) : "memory");
}
__device__ __forceinline__ void park1(uint32_t& q_top, float v, uint64_t w) {
    asm volatile("{\n\t.reg .pred q;\n\tsetp.eq.u64 q, %2, 0;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %3;\n\t}"
                 : "+r"(q_top) : "f"(v), "l"(w), "n"(4 * FAST_THREADS) : "memory");
}
__device__ __forceinline__ void park4(uint32_t& q_top, float4 v, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
    asm volatile("{\n\t.reg .pred q;\n\t"
                 "setp.eq.u32 q, %5, 0;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u32 q, %6, 0;\n\t@q st.shared.f32 [%0], %2;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u32 q, %7, 0;\n\t@q st.shared.f32 [%0], %3;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u32 q, %8, 0;\n\t@q st.shared.f32 [%0], %4;\n\t@q add.u32 %0, %0, %9;\n\t}"
                 : "+r"(q_top) : "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "r"(w0), "r"(w1), "r"(w2), "r"(w3),
                   "n"(4 * FAST_THREADS) : "memory");
}
__device__ __forceinline__ void park4(uint32_t& q_top, float4 v, uint64_t w0, uint64_t w1, uint64_t w2, uint64_t w3) {
    asm volatile("{\n\t.reg .pred q;\n\t"
                 "setp.eq.u64 q, %5, 0;\n\t@q st.shared.f32 [%0], %1;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u64 q, %6, 0;\n\t@q st.shared.f32 [%0], %2;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u64 q, %7, 0;\n\t@q st.shared.f32 [%0], %3;\n\t@q add.u32 %0, %0, %9;\n\t"
                 "setp.eq.u64 q, %8, 0;\n\t@q st.shared.f32 [%0], %4;\n\t@q add.u32 %0, %0, %9;\n\t}"
                 : "+r"(q_top) : "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w), "l"(w0), "l"(w1), "l"(w2), "l"(w3),
                   "n"(4 * FAST_THREADS) : "memory");
}

template <int NB>
__global__ void __launch_bounds__(FAST_THREADS, NB <= 5 ? 3 : 2) select_pass1_lut_f32_kernel(const float* __restrict__ x, unsigned long long ld,
                                                                            unsigned long long n, int vec,
                                                                            const RowSel* __restrict__ states,
                                                                            unsigned int* __restrict__ out, size_t row_slot,
                                                                            size_t iv_slot, int n_iv_cap, SideLists side) {
    static_assert(NB == 5 || NB == 9, "gap counters: six 5-bit fields in 32 bits, or ten 6-bit fields in 64");
    using Word = typename LutWord<NB>::type;
    constexpr int FIELD = NB <= 5 ? 5 : 6, WSHIFT = NB <= 5 ? 2 : 3;
    extern __shared__ unsigned int s_bins[];
    __shared__ uint4 s_br[NB];               // per bracket: lower key, shift, upper bound (float bits), address of its bins
    __shared__ Word s_lut[LUT_BINS];
    __shared__ float s_queue[QUEUE_DEPTH * FAST_THREADS];
    __shared__ float s_ring[SIDE_WARPS][64];
    const unsigned long long row = blockIdx.y;
    const RowSel& st = states[row];
    const int n_iv = st.n_iv;
    if (n_iv == 0) return;
    const int tid = threadIdx.x;
    if (tid < NB) {
        const uint32_t kh = tid < n_iv ? (uint32_t)st.hi[tid] : 0u;
        const float hf = value_of(kh > KEY32_POS_INF ? KEY32_POS_INF : (kh < KEY32_NEG_INF ? KEY32_NEG_INF : kh));
        s_br[tid] = make_uint4(tid < n_iv ? (uint32_t)st.lo[tid] : 0u, tid < n_iv ? (uint32_t)st.shift[tid] : 0u,
                               __float_as_uint(hf),
                               (uint32_t)__cvta_generic_to_shared(s_bins) + (uint32_t)tid * SEL_BINS * 4u);
    }
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) s_bins[i] = 0u;
    const size_t side_slot = (row * gridDim.x + blockIdx.x) * SIDE_WARPS + (tid >> 5);
    float* side_w = side.values ? side.values + side_slot * side.cap : nullptr;
    unsigned int side_n = 0;                 // warp-uniform
    if (side.values && blockIdx.x == 0) {    // remember the brackets the lists cover
        uint32_t* br = side.brackets + row * SIDE_BRACKET_WORDS;
        if (tid == 0) br[0] = (uint32_t)n_iv;
        if (tid < n_iv) { br[1 + tid] = (uint32_t)st.lo[tid]; br[1 + MAX_IV + tid] = (uint32_t)st.hi[tid]; }
    }

    float lo_f[NB], hi_f[NB], cnt[NB];      // cnt: parked values at or above lo (per thread < 2^24: exact in fp32)
    float fmin_ = INFINITY, fmax_ = -INFINITY;
#pragma unroll
    for (int g = 0; g < NB; ++g) {
        const bool used = g < n_iv;
        const uint32_t kl = used ? (uint32_t)st.lo[g] : 0u, kh = used ? (uint32_t)st.hi[g] : 0u;
        lo_f[g] = used ? value_of(kl < KEY32_NEG_INF ? KEY32_NEG_INF : kl) : __int_as_float(0x7FC00000);   // NaN: never >=
        hi_f[g] = value_of(kh > KEY32_POS_INF ? KEY32_POS_INF : (kh < KEY32_NEG_INF ? KEY32_NEG_INF : kh));
        cnt[g] = 0.0f;
        if (used && fabsf(lo_f[g]) < INFINITY) { fmin_ = fminf(fmin_, lo_f[g]); fmax_ = fmaxf(fmax_, lo_f[g]); }
        if (used && fabsf(hi_f[g]) < INFINITY) { fmin_ = fminf(fmin_, hi_f[g]); fmax_ = fmaxf(fmax_, hi_f[g]); }
    }
    // the table spans the finite bracket bounds plus 1/64 on either side
    float qs = 0.0f, qb = 0.0f;
    {
        float w = fmax_ - fmin_;
        // one finite bound only (a bracket that reaches to an infinity: percentile 0 or 100; a constant row): any
        // scale around it does -- the infinite bounds quantise to the first / last bin by saturation
        if (!(w > 0.0f) && fabsf(fmin_) < INFINITY) w = fmaxf(fabsf(fmin_) * 0.015625f, 1e-30f);
        const float m = w * (1.0f / 64.0f);
        const float s = 1.0f / (w + 2.0f * m), b = -(fmin_ - m) * s;
        if (w > 0.0f && s > 0.0f && s < INFINITY && fabsf(b) < INFINITY) { qs = s; qb = b; }      // else: every bin parks
    }
    const LutAddr la = make_lut_addr<Word>(qs, qb, (uint32_t)__cvta_generic_to_shared(s_lut));
    {
        uint32_t qlo[NB], qhi[NB];
#pragma unroll
        for (int g = 0; g < NB; ++g) {
            qlo[g] = (lut_addr(la, lo_f[g]) - la.base) >> WSHIFT;
            qhi[g] = (lut_addr(la, hi_f[g]) - la.base) >> WSHIFT;
        }
        for (uint32_t k = tid; k < (uint32_t)LUT_BINS; k += FAST_THREADS) {
            int gap = 0;
            bool straddles = qs == 0.0f;
#pragma unroll
            for (int g = 0; g < NB; ++g) {
                if (g < n_iv) {
                    gap += qhi[g] < k ? 1 : 0;
                    straddles |= qlo[g] <= k && k <= qhi[g];
                }
            }
            s_lut[k] = straddles ? Word(0) : Word(1) << (FIELD * gap);
        }
    }
    __syncthreads();

    unsigned int* out_row = out + row * row_slot;
    unsigned int* out_iv = out_row + 2 * (size_t)n_iv_cap;
    Word acc = 0;                            // the gap counters of the current batch: NB + 1 fields of FIELD bits
    unsigned int gapc[NB + 1];
#pragma unroll
    for (int j = 0; j <= NB; ++j) gapc[j] = 0u;
    const uint32_t q_base = (uint32_t)__cvta_generic_to_shared(s_queue) + 4u * tid;
    uint32_t q_top = q_base;                 // byte address of this thread's next free queue slot
    if (side_w) asm volatile("mov.u64 %0, %0;" : "+l"(side_w));      // keep the pointer in registers (it was rebuilt per store)
    float* ring = s_ring[tid >> 5];
    const uint32_t ring_addr = (uint32_t)__cvta_generic_to_shared(ring);
    uint32_t ring_top = ring_addr;           // fill level of the ring, as an address (warp-uniform)
    const unsigned int lane = tid & 31, lane_lt = (1u << lane) - 1u;
    // exact treatment of one parked value per lane (NaN: idle lane): counts at or above each lower bound, and --
    // inside a bracket -- the bracket's histogram and the warp's side list
    uint32_t br_addr;
    asm volatile("mov.u32 %0, %1;" : "=r"(br_addr) : "r"((uint32_t)__cvta_generic_to_shared(s_br)));
    auto exact = [&](float v) {
        float gf = 0.0f;
#pragma unroll
        for (int k = 0; k < NB; ++k) {
            float ge;
            asm("set.ge.f32.f32 %0, %1, %2;" : "=f"(ge) : "f"(v), "f"(lo_f[k]));     // 1.0f / 0.0f
            cnt[k] += ge;
            gf += ge;
        }
        const int g = __float2int_rn(gf) - 1;
        uint4 br;
        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(br.x), "=r"(br.y), "=r"(br.z), "=r"(br.w)
                     : "r"(br_addr + 16u * (uint32_t)(g < 0 ? 0 : g)) : "memory");
        const bool inb = g >= 0 && v <= __uint_as_float(br.z);
        if (inb) {
            const uint32_t a = br.w + (((key_of(v) - br.x) >> br.y) << 2);
            asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(a) : "memory");
        }
        const unsigned int who = __ballot_sync(0xffffffffu, inb);
        if (side_w && inb) {
            const unsigned int pos = side_n + __popc(who & lane_lt);
            if (pos < side.cap) side_w[pos] = v;
        }
        side_n += __popc(who);
    };
    unsigned long long beg, end;
    fast_block_range(n, &beg, &end);
    fast_stream<true>(x + row * ld, beg, end, vec != 0,
        [&](auto v) {
            if constexpr (sizeof(v) == sizeof(float4)) {       // four values: one register for the queue top throughout
                uint32_t a0, a1, a2, a3;
                lut_addr2(la, v.x, v.y, a0, a1);
                lut_addr2(la, v.z, v.w, a2, a3);
                const Word w0 = lut_word<Word>(a0);
                const Word w1 = lut_word<Word>(a1);
                const Word w2 = lut_word<Word>(a2);
                const Word w3 = lut_word<Word>(a3);
                acc += (w0 + w1) + (w2 + w3);
                park4(q_top, v, w0, w1, w2, w3);
            } else {
                const Word w = lut_word<Word>(lut_addr(la, v));
                acc += w;
                park1(q_top, v, w);
            }
        },
        [&]() {
#pragma unroll
            for (int j = 0; j <= NB; ++j) gapc[j] += (unsigned int)(acc >> (FIELD * j)) & ((1u << FIELD) - 1u);
            acc = 0;
            // the lanes hold 0..16 parked values each (1.5 on average).  They are packed into the warp's ring --
            // a lane's values go behind those of the lanes below it -- and treated 32 at a time, every lane busy
            {
                // one value per lane and turn; shared-memory addresses are kept as 32-bit registers and the
                // predicated part is spelled out (the C++ form of this loop compiled to 40 instructions a turn)
                for (;;) {
                    const unsigned int who = __ballot_sync(0xffffffffu, q_top != q_base);
                    if (!who) break;
                    asm volatile("{\n\t.reg .pred p;\n\t.reg .b32 t;\n\t.reg .f32 v;\n\t"
                                 "setp.ne.u32 p, %0, %2;\n\t"
                                 "@p sub.u32 %0, %0, %3;\n\t"
                                 "@p ld.shared.f32 v, [%0];\n\t"
                                 "and.b32 t, %4, %5;\n\t"
                                 "popc.b32 t, t;\n\t"
                                 "shl.b32 t, t, 2;\n\t"
                                 "add.u32 t, t, %1;\n\t"
                                 "@p st.shared.f32 [t], v;\n\t}"
                                 : "+r"(q_top) : "r"(ring_top), "r"(q_base), "n"(4 * FAST_THREADS), "r"(who), "r"(lane_lt)
                                 : "memory");
                    ring_top += 4u * __popc(who);
                    __syncwarp();
                    if (ring_top >= ring_addr + 128u) {
                        ring_top -= 128u;
                        float v;
                        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(ring_top + 4u * lane) : "memory");
                        exact(v);
                        __syncwarp();
                    }
                }
            }
        });
    const unsigned int n_ring = (ring_top - ring_addr) >> 2;
    if (n_ring) exact(lane < n_ring ? ring[lane] : __int_as_float(0x7FC00000));
    if (side_w && (tid & 31) == 0) side.counts[side_slot] = side_n;       // > cap: overflowed, not usable

    unsigned long long* out_ge = reinterpret_cast<unsigned long long*>(out_row);
    unsigned int above = 0u;                 // values of this thread in the gaps above bracket g
#pragma unroll
    for (int g = NB - 1; g >= 0; --g) {
        above += gapc[g + 1];
        const unsigned int total = __reduce_add_sync(0xffffffffu, (unsigned int)cnt[g] + above);
        if ((tid & 31) == 0 && g < n_iv && total) atomicAdd(&out_ge[g], (unsigned long long)total);
    }
    __syncthreads();
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) {
        const unsigned int v = s_bins[i];
        if (v) atomicAdd(&out_iv[(size_t)(i >> SEL_BIN_BITS) * iv_slot + 2 + (i & (SEL_BINS - 1))], v);
    }
}

__global__ void __launch_bounds__(FAST_THREADS) select_pass2_f32_kernel(const float* __restrict__ x, unsigned long long ld,
                                                                        unsigned long long n, int vec,
                                                                        const RowSel* __restrict__ states,
                                                                        unsigned int* __restrict__ out, size_t row_slot,
                                                                        size_t iv_slot, int n_iv_cap, SideLists side) {
    extern __shared__ unsigned int s_bins[];
    __shared__ uint32_t s_lo[MAX_IV], s_hi[MAX_IV];
    __shared__ int s_shift[MAX_IV], s_mode[MAX_IV];
    __shared__ uint32_t s_map[MAP_WORDS];
    __shared__ int s_use_side;
    const unsigned long long row = blockIdx.y;
    const RowSel& st = states[row];
    const int n_iv = st.n_iv;
    if (n_iv == 0) return;
    const int tid = threadIdx.x;
    if (tid < n_iv) {
        s_lo[tid] = (uint32_t)st.lo[tid]; s_hi[tid] = (uint32_t)st.hi[tid];
        s_shift[tid] = st.shift[tid]; s_mode[tid] = st.mode[tid];
    }
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) s_bins[i] = 0u;
    for (int i = tid; i < MAP_WORDS; i += FAST_THREADS) s_map[i] = 0u;
    __syncthreads();
    // prefilter: bit (key >> cs) mod 2^14 is set for every key inside an interval.  cs makes the narrowest interval
    // about four bins wide, so a handful of the 16384 bits are set and a value outside the intervals gets past
    // with probability ~1e-3 (the exact test behind it diverges the warp: with the former 2048 bins over the
    // whole span of the intervals ~4 % of the parked values got past and most warps took the slow path)
    int min_bl = 32;
    for (int i = 0; i < n_iv; ++i) min_bl = min(min_bl, bit_length((unsigned long long)(s_hi[i] - s_lo[i])));
    const int cs = max(0, min_bl - 2);
    if (tid < n_iv) {
        const uint32_t b0 = s_lo[tid] >> cs, nb = (s_hi[tid] >> cs) - b0 + 1u;
        if (nb > MAP_FILL_MAX) {
            for (int i = 0; i < MAP_WORDS; ++i) atomicOr(&s_map[i], 0xFFFFFFFFu);
        } else {
            for (uint32_t b = 0; b < nb; ++b) {
                const uint32_t idx = (b0 + b) & (uint32_t)(MAP_BITS - 1);
                atomicOr(&s_map[idx >> 5], 1u << (idx & 31u));
            }
        }
    }
    __syncthreads();

    // the side lists of pass 1 hold every value of this chunk that lies in a pass-1 bracket: enough when each
    // interval of the row is inside one of those brackets and none of this block's eight lists overflowed
    const size_t side_slot = (row * gridDim.x + blockIdx.x) * SIDE_WARPS;
    {
        // one thread per interval and one per list, all at once (a single thread doing this alone was a chain of ~20
        // dependent global loads, ~10 us per block with the other 255 threads at the barrier)
        const bool have = side.values != nullptr && side.chunks == gridDim.x;
        int ok = 1;
        if (have && tid < n_iv) {
            const uint32_t* br = side.brackets + row * SIDE_BRACKET_WORDS;
            const int nb = (int)br[0];
            int inside = 0;
            for (int b = 0; b < nb; ++b) inside |= (br[1 + b] <= s_lo[tid] && s_hi[tid] <= br[1 + MAX_IV + b]) ? 1 : 0;
            ok = inside;
        }
        if (have && tid >= 32 && tid < 32 + SIDE_WARPS) ok = side.counts[side_slot + (tid - 32)] <= side.cap;
        const int all_ok = __syncthreads_and(ok);
        if (tid == 0) s_use_side = have && all_ok;
    }
    __syncthreads();

    unsigned int* out_row = out + row * row_slot;
    FastRow c{s_lo, s_hi, s_shift, s_mode, s_bins, out_row + 2 * (size_t)n_iv_cap, iv_slot, n_iv};
    auto visit = [&](float v) {
        const uint32_t key = key_of(v);
        const uint32_t idx = (key >> cs) & (uint32_t)(MAP_BITS - 1);
        if ((s_map[idx >> 5] >> (idx & 31u)) & 1u) fast_member(c, key);
    };
    if (s_use_side) {
        const int warp = tid >> 5, lane = tid & 31;
        const float* list = side.values + (side_slot + warp) * side.cap;
        const unsigned int cnt = side.counts[side_slot + warp];
        // four 16-byte loads in flight per lane (the lists start on 128-byte boundaries): 512 values per warp and round
        // (8 scalar loads per lane left the pass waiting on its own loads: long scoreboard 5.6, ncu ncu_full_p2c5.txt)
        const float4* list4 = reinterpret_cast<const float4*>(list);
        const unsigned int cnt4 = (reinterpret_cast<uintptr_t>(list) & 15u) == 0 ? cnt >> 2 : 0u;     // (a test knob may set an odd capacity)
        unsigned int i = lane;
        for (; i + 96 < cnt4; i += 128) {
            float4 a[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = __ldcs(list4 + i + 32 * u);
#pragma unroll
            for (int u = 0; u < 4; ++u) { visit(a[u].x); visit(a[u].y); visit(a[u].z); visit(a[u].w); }
        }
        for (; i < cnt4; i += 32) { const float4 q = __ldcs(list4 + i); visit(q.x); visit(q.y); visit(q.z); visit(q.w); }
        for (unsigned int k = (cnt4 << 2) + lane; k < cnt; k += 32) visit(__ldcs(list + k));
    } else {
        unsigned long long beg, end;
        fast_block_range(n, &beg, &end);
        fast_stream<false>(x + row * ld, beg, end, vec != 0, visit, [] {});
    }
    __syncthreads();
    for (int i = tid; i < n_iv * SEL_BINS; i += FAST_THREADS) {
        const unsigned int v = s_bins[i];
        if (v) atomicAdd(&c.out_iv[(size_t)(i >> SEL_BIN_BITS) * iv_slot + 2 + (i & (SEL_BINS - 1))], v);
    }
}

template <typename K>
static cudaError_t launch_fast(K k, const SelectProblem& sp, void* d_out, int n_iv_smem, dim3 grid, cudaStream_t stream) {
    const size_t smem = (size_t)n_iv_smem * SEL_BINS * sizeof(unsigned int);
    cudaError_t e = ensure_dynamic_smem(k, smem);      // the kernels also hold ~21 KB of static shared memory
    if (e != cudaSuccess) return e;
    const float* x = static_cast<const float*>(sp.x);
    const int vec = (sp.ld % 4 == 0 && ((uintptr_t)x & 15u) == 0) ? 1 : 0;
    SideLists side = sp.side;
    if (side.values && side.chunks != grid.x) side.values = nullptr;         // laid out for another grid: off
    k<<<grid, FAST_THREADS, smem, stream>>>(x, sp.ld, sp.n, vec, sp.d_states, static_cast<unsigned int*>(d_out),
                                            row_slot_u32(sp.dtype, sp.plan.n_targets), iv_slot_u32(sp.dtype),
                                            sp.plan.n_targets, side);
    return cudaGetLastError();
}

unsigned long long select_chunks_f32(const SelectProblem& sp, int sm_count) {
    // one row: two blocks per SM stream it; many rows: ~8 resident blocks per SM in total, whole chunks of 4 KB
    unsigned long long chunks = sp.rows == 1 ? 2ull * sm_count : ((unsigned long long)sm_count * 8 + sp.rows - 1) / sp.rows;
    const unsigned long long max_chunks = (sp.n + 4095ull) / 4096ull;
    if (chunks > max_chunks) chunks = max_chunks;
    if (chunks < 1) chunks = 1;
    return chunks;
}

unsigned int side_list_cap(unsigned long long n, unsigned long long chunks) {
    // a quarter of the chunk, split over the block's warps (the brackets of 5..9 percentiles hold 9..13 % of a
    // row when the pilot sample has 32768 values); at least one 128-byte line per list
    unsigned long long per = (n + chunks - 1) / chunks;
    per = (per + 1023ull) & ~1023ull;                       // fast_block_range's chunk length
    unsigned long long cap = per / (4ull * SIDE_WARPS);
    cap = (cap + 31ull) & ~31ull;
    return (unsigned int)(cap < 32 ? 32 : cap);
}

// the decision select_pass2_f32_kernel takes per block, taken for all blocks ahead of the launch
__global__ void side_check_kernel(const RowSel* __restrict__ states, unsigned long long rows, SideLists side,
                                  unsigned int* __restrict__ blocks) {
    const unsigned long long row = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    const RowSel& st = states[row];
    if (st.n_iv == 0) return;
    const uint32_t* br = side.brackets + row * SIDE_BRACKET_WORDS;
    const int nb = (int)br[0];
    bool ok = true;
    for (int i = 0; i < st.n_iv && ok; ++i) {
        bool inside = false;
        for (int b = 0; b < nb; ++b) inside |= br[1 + b] <= (uint32_t)st.lo[i] && (uint32_t)st.hi[i] <= br[1 + MAX_IV + b];
        ok = inside;
    }
    unsigned int bad = 0;
    for (unsigned int c = 0; c < side.chunks; ++c) {
        bool fits = ok;
        for (int w = 0; w < SIDE_WARPS && fits; ++w) fits = side.counts[(row * side.chunks + c) * SIDE_WARPS + w] <= side.cap;
        bad += fits ? 0u : 1u;
    }
    if (bad) atomicAdd(blocks, bad);
}

cudaError_t launch_side_check(const SelectProblem& sp, int sm_count, unsigned int* d_blocks, cudaStream_t stream) {
    if (sp.plan.n_targets == 0 || sp.rows == 0 || sp.n == 0) return cudaSuccess;
    if (sp.dtype != MC_F32 || !sp.side.values || sp.side.chunks != select_chunks_f32(sp, sm_count)) {
        const unsigned int all = 1u;        // no lists: every block streams its rows
        return cudaMemcpyAsync(d_blocks, &all, sizeof(all), cudaMemcpyHostToDevice, stream);
    }
    side_check_kernel<<<(unsigned)((sp.rows + 127) / 128), 128, 0, stream>>>(sp.d_states, sp.rows, sp.side, d_blocks);
    return cudaGetLastError();
}

static cudaError_t launch_pass_f32(const SelectProblem& sp, bool first, void* d_out, int sm_count, cudaStream_t s) {
    const unsigned long long chunks = select_chunks_f32(sp, sm_count);
    dim3 grid((unsigned)chunks, (unsigned)sp.rows);
    const int np_ = sp.plan.n_pcts;
    if (!first) return launch_fast(select_pass2_f32_kernel, sp, d_out, sp.plan.n_targets, grid, s);
    if (np_ <= 5) return launch_fast(select_pass1_lut_f32_kernel<5>, sp, d_out, np_, grid, s);
    if (np_ <= 9) return launch_fast(select_pass1_lut_f32_kernel<9>, sp, d_out, np_, grid, s);
    return launch_fast(select_pass1_f32_kernel<16>, sp, d_out, np_, grid, s);
}

template <typename Real, int NIV>
static cudaError_t launch_pass_t(const SelectProblem& sp, void* d_out, int n_iv_smem, dim3 grid, cudaStream_t stream) {
    auto k = select_pass_kernel<Real, NIV>;
    const size_t smem = (size_t)n_iv_smem * SEL_BINS * sizeof(unsigned int);
    cudaError_t e = ensure_dynamic_smem(k, smem);
    if (e != cudaSuccess) return e;
    k<<<grid, 256, smem, stream>>>(static_cast<const Real*>(sp.x), sp.ld, sp.n, sp.d_states,
                                   static_cast<unsigned int*>(d_out), row_slot_u32(sp.dtype, sp.plan.n_targets),
                                   iv_slot_u32(sp.dtype), sp.plan.n_targets);
    return cudaGetLastError();
}

template <typename Real>
static cudaError_t launch_pass_real(const SelectProblem& sp, bool first, void* d_out, dim3 grid, cudaStream_t s) {
    const int np_ = sp.plan.n_pcts;
    if (!first) return launch_pass_t<Real, 0>(sp, d_out, sp.plan.n_targets, grid, s);
    if (np_ <= 5) return launch_pass_t<Real, 5>(sp, d_out, np_, grid, s);
    if (np_ <= 9) return launch_pass_t<Real, 9>(sp, d_out, np_, grid, s);
    return launch_pass_t<Real, 16>(sp, d_out, np_, grid, s);
}

cudaError_t launch_select_pass(const SelectProblem& sp, bool first_pass, void* d_out, int sm_count,
                               cudaStream_t stream, bool zero_out) {
    if (sp.plan.n_targets == 0 || sp.rows == 0) return cudaSuccess;
    if (sp.rows > 65535) return cudaErrorInvalidValue;
    // everything a pass writes is additive (counts, histogram bins, appended candidates): several problems of one
    // GPU may accumulate into one buffer, which only the first of them clears
    cudaError_t e = zero_out ? cudaMemsetAsync(d_out, 0, pass_words(sp) * sizeof(unsigned long long), stream) : cudaSuccess;
    if (e != cudaSuccess || sp.n == 0) return e;
    if (sp.dtype == MC_F32) return launch_pass_f32(sp, first_pass, d_out, sm_count, stream);
    unsigned long long chunks = ((unsigned long long)sm_count * 16 + sp.rows - 1) / sp.rows;
    const unsigned long long max_chunks = (sp.n + 256ull * PASS_UNROLL - 1) / (256ull * PASS_UNROLL);
    if (chunks > max_chunks) chunks = max_chunks;
    if (chunks < 1) chunks = 1;
    dim3 grid((unsigned)chunks, (unsigned)sp.rows);
    return sp.dtype == MC_F64 ? launch_pass_real<double>(sp, first_pass, d_out, grid, stream)
                              : launch_pass_real<float>(sp, first_pass, d_out, grid, stream);
}

// ------------------------------------------------------------------ scan: shrink every target's interval
// Two kernels.  (1) One block of 1024 threads per (row, interval): a histogrammed interval is prefix-summed
// over its 1024 bins (summed over ranks), a collected one is sorted; each target of the interval then moves
// to the bin that holds its rank, is decided outright (single-key bin, or position in the sorted
// candidates), or -- when its bracket missed -- is flagged for the gap beside the interval.  (2) One thread
// per row assembles the next RowSel: gaps get their exact population below from the neighbours' totals.
constexpr int SCAN_THREADS = 256;           // four bins per thread: eight blocks per SM instead of two (10250 blocks for the
constexpr int SCAN_PER = SEL_BINS / SCAN_THREADS;   // 1025 rows x 10 intervals of a config-5 fan chart: 90 -> 35 us)
constexpr unsigned long long POP_UNKNOWN = ~0ull;
enum { SC_NONE = 0, SC_BIN = 1, SC_DONE = 2, SC_GAP_LOW = 3, SC_GAP_HIGH = 4 };

template <typename Key>
__global__ void __launch_bounds__(SCAN_THREADS) select_scan_iv_kernel(const SelectPlan* __restrict__ plan,
                                                                      const unsigned int* __restrict__ g,
                                                                      unsigned int n_ranks, size_t rank_stride_u32,
                                                                      size_t row_slot, size_t iv_slot, int n_iv_cap,
                                                                      const RowSel* __restrict__ states,
                                                                      ScanScratch* __restrict__ scratch) {
    __shared__ unsigned long long s_cum[SEL_BINS];
    __shared__ unsigned long long s_keys[SEL_CAND];
    __shared__ unsigned long long s_warp[SCAN_THREADS / 32];
    const unsigned long long row = blockIdx.x;
    const int gi = blockIdx.y;
    const RowSel& st = states[row];
    if (gi >= st.n_iv) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nt = plan->n_targets;
    const unsigned int* grow = g + row * row_slot;
    unsigned long long below = st.below[gi];
    if (!st.below_known) {           // first scan: the pass counted the values at or above each lower bound
        unsigned long long ge = 0;
        for (unsigned int r = 0; r < n_ranks; ++r)
            ge += reinterpret_cast<const unsigned long long*>(grow + r * rank_stride_u32)[gi];
        below = plan->n_global - ge;
    }
    const int mode = st.mode[gi];
    const unsigned int* slot0 = grow + 2 * (size_t)n_iv_cap + (size_t)gi * iv_slot;
    unsigned long long total;
    if (mode == IV_HIST) {
        unsigned long long c[SCAN_PER], run = 0;
#pragma unroll
        for (int k = 0; k < SCAN_PER; ++k) c[k] = 0;
        for (unsigned int r = 0; r < n_ranks; ++r) {
            // bins tid * 4 .. tid * 4 + 3 of rank r: the slot starts 2 words into an 8-byte aligned row: 8-byte loads
            const uint2* src = reinterpret_cast<const uint2*>(slot0 + r * rank_stride_u32 + 2 + tid * SCAN_PER);
            static_assert(SCAN_PER == 4, "two 8-byte loads per thread and rank");
            const uint2 a = src[0], b = src[1];
            c[0] += a.x; c[1] += a.y; c[2] += b.x; c[3] += b.y;
        }
#pragma unroll
        for (int k = 0; k < SCAN_PER; ++k) { run += c[k]; c[k] = run; }
        unsigned long long incl = run;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
        }
        if (lane == 31) s_warp[warp] = incl;
        __syncthreads();
        unsigned long long off = incl - run;
        for (int w = 0; w < warp; ++w) off += s_warp[w];
#pragma unroll
        for (int k = 0; k < SCAN_PER; ++k) s_cum[tid * SCAN_PER + k] = off + c[k];
        __syncthreads();
        total = s_cum[SEL_BINS - 1];
    } else {
        unsigned int base = 0;
        for (unsigned int r = 0; r < n_ranks; ++r) {
            const unsigned int* slot = slot0 + r * rank_stride_u32;
            unsigned int cr = slot[0];
            if (cr > (unsigned int)SEL_CAND) cr = SEL_CAND;
            for (unsigned int t = tid; t < cr; t += SCAN_THREADS)
                if (base + t < (unsigned int)SEL_CAND) s_keys[base + t] = (unsigned long long)reinterpret_cast<const Key*>(slot + 2)[t];
            base += cr;
        }
        const unsigned int cnt_all = base < (unsigned int)SEL_CAND ? base : SEL_CAND;
        int P = 1;
        while (P < (int)cnt_all) P <<= 1;
        for (int t = (int)cnt_all + tid; t < P; t += SCAN_THREADS) s_keys[t] = ~0ull;
        __syncthreads();
        for (int k = 2; k <= P; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int t = tid; t < P; t += SCAN_THREADS) {
                    const int ixj = t ^ j;
                    if (ixj > t) {
                        const unsigned long long a = s_keys[t], b2 = s_keys[ixj];
                        const bool up = (t & k) == 0;
                        if ((a > b2) == up) { s_keys[t] = b2; s_keys[ixj] = a; }
                    }
                }
                __syncthreads();
            }
        }
        total = cnt_all;
    }
    ScanScratch& sc = scratch[row];
    if (tid == 0) { sc.below[gi] = below; sc.total[gi] = total; }
    if (tid < nt && st.target_iv[tid] == gi) {
        const unsigned long long rank = plan->rank[tid];
        const unsigned long long lo = st.lo[gi], hi = st.hi[gi];
        int code;
        unsigned long long nlo = 0, nhi = 0, nbelow = 0, npop = POP_UNKNOWN;
        if (rank < below) {
            code = SC_GAP_LOW;
        } else if (rank - below >= total) {
            code = SC_GAP_HIGH;
        } else if (mode == IV_HIST) {
            const unsigned long long r = rank - below;
            int a = 0, b2 = SEL_BINS - 1;          // first bin with cum > r
            while (a < b2) { const int mid = (a + b2) >> 1; if (s_cum[mid] > r) b2 = mid; else a = mid + 1; }
            const int sh = st.shift[gi];
            const unsigned long long before = a ? s_cum[a - 1] : 0ull;
            nlo = lo + ((unsigned long long)a << sh);
            const unsigned long long span = (1ull << sh) - 1ull;
            nhi = (hi - nlo < span) ? hi : nlo + span;
            nbelow = below + before;
            npop = s_cum[a] - before;
            code = nlo == nhi ? SC_DONE : SC_BIN;
        } else {
            nlo = nhi = s_keys[rank - below];
            code = SC_DONE;
        }
        sc.nlo[tid] = nlo; sc.nhi[tid] = nhi; sc.nbelow[tid] = nbelow; sc.npop[tid] = npop; sc.code[tid] = code;
    }
}

__global__ void __launch_bounds__(128) select_scan_assemble_kernel(const SelectPlan* __restrict__ plan,
                                                                   unsigned long long rows, unsigned long long key_max_v,
                                                                   RowSel* __restrict__ states,
                                                                   const ScanScratch* __restrict__ scratch,
                                                                   unsigned int* __restrict__ pending) {
    const unsigned long long row = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= rows) return;
    RowSel& st = states[row];
    const int n_iv = st.n_iv;
    if (n_iv == 0) return;
    const ScanScratch& sc = scratch[row];
    const int nt = plan->n_targets;
    RowSel nw;
    int n_new = 0;
    for (int j = 0; j < nt; ++j) {
        nw.key[j] = st.key[j];
        const int gi = st.target_iv[j];
        if (gi < 0) { nw.target_iv[j] = -1; continue; }
        const int code = sc.code[j];
        if (code == SC_DONE) { nw.key[j] = sc.nlo[j]; nw.target_iv[j] = -1; continue; }
        unsigned long long nlo, nhi, nbelow, npop = POP_UNKNOWN;
        if (code == SC_GAP_LOW) {            // the bracket missed: the value lies in the gap below it ...
            nlo = gi == 0 ? 0ull : st.hi[gi - 1] + 1ull;
            nhi = st.lo[gi] - 1ull;
            nbelow = gi == 0 ? 0ull : sc.below[gi - 1] + sc.total[gi - 1];
        } else if (code == SC_GAP_HIGH) {    // ... or above it
            nlo = st.hi[gi] + 1ull;
            nhi = gi + 1 < n_iv ? st.lo[gi + 1] - 1ull : key_max_v;
            nbelow = sc.below[gi] + sc.total[gi];
        } else {
            nlo = sc.nlo[j]; nhi = sc.nhi[j]; nbelow = sc.nbelow[j]; npop = sc.npop[j];
        }
        if (!(n_new && nw.lo[n_new - 1] == nlo && nw.hi[n_new - 1] == nhi)) {
            nw.lo[n_new] = nlo; nw.hi[n_new] = nhi; nw.below[n_new] = nbelow;
            nw.shift[n_new] = max(0, bit_length(nhi - nlo) - SEL_BIN_BITS);
            nw.mode[n_new] = npop <= (unsigned long long)SEL_CAND ? IV_COLLECT : IV_HIST;
            ++n_new;
        }
        nw.target_iv[j] = n_new - 1;
    }
    st.n_iv = n_new;
    st.below_known = 1;
    for (int i = 0; i < n_new; ++i) {
        st.lo[i] = nw.lo[i]; st.hi[i] = nw.hi[i]; st.below[i] = nw.below[i];
        st.shift[i] = nw.shift[i]; st.mode[i] = nw.mode[i];
    }
    for (int j = 0; j < nt; ++j) { st.key[j] = nw.key[j]; st.target_iv[j] = nw.target_iv[j]; }
    if (n_new) atomicAdd(pending, 1u);
}

cudaError_t launch_select_scan(const SelectProblem& sp, const void* d_gathered_pass, unsigned int n_ranks,
                               size_t rank_stride_words, unsigned int* d_pending, cudaStream_t stream) {
    if (sp.plan.n_targets == 0 || sp.rows == 0) return cudaSuccess;
    if (sp.rows > 65535 * 32ull || !sp.d_scratch) return cudaErrorInvalidValue;
    const unsigned int* g = static_cast<const unsigned int*>(d_gathered_pass);
    const size_t rs = row_slot_u32(sp.dtype, sp.plan.n_targets), is = iv_slot_u32(sp.dtype);
    const dim3 grid((unsigned)sp.rows, (unsigned)sp.plan.n_targets);
    if (sp.dtype == MC_F64)
        select_scan_iv_kernel<uint64_t><<<grid, SCAN_THREADS, 0, stream>>>(
            sp.d_plan, g, n_ranks, rank_stride_words * 2, rs, is, sp.plan.n_targets, sp.d_states, sp.d_scratch);
    else
        select_scan_iv_kernel<uint32_t><<<grid, SCAN_THREADS, 0, stream>>>(
            sp.d_plan, g, n_ranks, rank_stride_words * 2, rs, is, sp.plan.n_targets, sp.d_states, sp.d_scratch);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
    const unsigned long long kmax = sp.dtype == MC_F64 ? ~0ull : 0xFFFFFFFFull;
    select_scan_assemble_kernel<<<(unsigned)((sp.rows + 127) / 128), 128, 0, stream>>>(sp.d_plan, sp.rows, kmax, sp.d_states,
                                                                                      sp.d_scratch, d_pending);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ finish: keys -> values -> percentiles
__global__ void select_finish_kernel(const SelectPlan* __restrict__ plan, const RowSel* __restrict__ states,
                                     unsigned long long rows, double* __restrict__ out, double* __restrict__ out_lo,
                                     double* __restrict__ out_hi) {
    const unsigned long long idx = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int np_ = plan->n_pcts;
    if (idx >= rows * (unsigned long long)np_) return;
    const unsigned long long row = idx / np_;
    const int k = (int)(idx % np_);
    const RowSel& st = states[row];
    const unsigned long long ka = st.key[plan->pct_lo_idx[k]], kb = st.key[plan->pct_hi_idx[k]];
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

// ------------------------------------------------------------------ closing pass over the terminal values
// tail: mean(final[final <= p]) (app.py:656) and the fixed-bin histogram in ONE read.
// Histogram semantics: np.histogram(x.astype(float64), bins, range=(lo, hi)) -- numpy/lib/_histograms_impl.py
// uniform-bin path: values outside [lo, hi] dropped; the bin is the one whose edges (np.linspace(lo, hi,
// bins+1)) bracket the value, last bin closed on the right.  For fp32 data every comparison "v < edge" with a
// float64 edge is done exactly in fp32 against ceil_f32(edge), so the kernel needs no fp64 arithmetic.
__global__ void hist_edges_kernel(unsigned int bins, double lo, double hi, DeviceResult* res, double* edges,
                                  float* edges_f) {
    if (lo >= hi) { lo = res->vmin; hi = res->vmax; }       // range=None: (a.min(), a.max())
    if (lo == hi) { lo -= 0.5; hi += 0.5; }                 // _get_outer_edges
    if (blockIdx.x == 0 && threadIdx.x == 0) { res->hist_lo = lo; res->hist_hi = hi; }
    const double step = __ddiv_rn(__dsub_rn(hi, lo), (double)bins);      // np.linspace
    for (unsigned int i = blockIdx.x * blockDim.x + threadIdx.x; i <= bins; i += gridDim.x * blockDim.x) {
        const double e = (i == bins) ? hi : __dadd_rn(__dmul_rn((double)i, step), lo);
        edges[i] = e;
        float f = (float)e;
        if (i == bins) { if ((double)f > e) f = nextafterf(f, -INFINITY); }   // v <= hi  <=>  v <= floor_f32(hi)
        else if ((double)f < e) f = nextafterf(f, INFINITY);                  // v <  e   <=>  v <  ceil_f32(e)
        edges_f[i] = f;
    }
}

template <typename Real> struct EdgeSrc;
template <> struct EdgeSrc<float> { static __device__ const float* get(const double*, const float* f) { return f; } };
template <> struct EdgeSrc<double> { static __device__ const double* get(const double* d, const float*) { return d; } };

template <typename Real>
__global__ void __launch_bounds__(256) tail_hist_kernel(const Real* __restrict__ x, unsigned long long n,
                                                        const DeviceResult* __restrict__ res, int want_tail,
                                                        unsigned int bins, const double* __restrict__ edges,
                                                        const float* __restrict__ edges_f,
                                                        double* __restrict__ scratch,
                                                        unsigned long long* __restrict__ counts, int staged) {
    // staged: edges and per-block counts live in shared memory; else (more bins than fit) the edges are read
    // through L1 and every value adds to the global counts directly
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ double s_sum[8];
    __shared__ unsigned long long s_cnt[8];
    Real* s_edge_buf = reinterpret_cast<Real*>(smem_raw);                           // [bins + 1]
    unsigned int* s_counts = reinterpret_cast<unsigned int*>(s_edge_buf + (bins ? bins + 2 : 0));   // [bins]
    const Real* esrc = EdgeSrc<Real>::get(edges, edges_f);
    if (staged) {
        for (unsigned int i = threadIdx.x; i < bins; i += blockDim.x) s_counts[i] = 0u;
        for (unsigned int i = threadIdx.x; bins && i <= bins; i += blockDim.x) s_edge_buf[i] = esrc[i];
    }
    const Real* s_edge = staged ? s_edge_buf : esrc;
    __syncthreads();
    // v <= thr (a double) for a float v  <=>  v <= floor_f32(thr)
    Real thr = (Real)(-INFINITY);
    if (want_tail) {
        const double t = res->tail_threshold;
        thr = (Real)t;
        if ((double)thr > t) thr = sizeof(Real) == 4 ? (Real)nextafterf((float)thr, -INFINITY) : thr;
    }
    Real lo = Real(0), top = Real(0), inv_w = Real(0);
    if (bins) { lo = s_edge[0]; top = s_edge[bins]; inv_w = (Real)bins / (top - lo); }
    const int last = (int)bins - 1;
    double sum = 0.0;
    unsigned long long cnt = 0;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    auto visit = [&](Real v) {
        if (v <= thr) { sum += (double)v; ++cnt; }
        if (bins && v >= lo && v <= top) {
            int idx = (int)((v - lo) * inv_w);
            idx = idx < 0 ? 0 : (idx > last ? last : idx);
            while (idx > 0 && v < s_edge[idx]) --idx;
            while (idx < last && v >= s_edge[idx + 1]) ++idx;
            if (staged) atomicAdd(&s_counts[idx], 1u);
            else atomicAdd(&counts[idx], 1ull);
        }
    };
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n; i += 4 * stride) {        // four loads in flight per thread
        const Real a = x[i], b = x[i + stride], c = x[i + 2 * stride], d = x[i + 3 * stride];
        visit(a); visit(b); visit(c); visit(d);
    }
    for (; i < n; i += stride) visit(x[i]);
    if (want_tail) {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) {
            sum += __shfl_down_sync(0xffffffffu, sum, d);
            cnt += __shfl_down_sync(0xffffffffu, cnt, d);
        }
        if ((threadIdx.x & 31) == 0) { s_sum[threadIdx.x >> 5] = sum; s_cnt[threadIdx.x >> 5] = cnt; }
    }
    __syncthreads();
    if (want_tail && threadIdx.x == 0) {
        double ts = 0.0; unsigned long long tc = 0;
        for (int w = 0; w < 8; ++w) { ts += s_sum[w]; tc += s_cnt[w]; }
        scratch[2 * blockIdx.x] = ts;
        scratch[2 * blockIdx.x + 1] = __longlong_as_double((long long)tc);
    }
    for (unsigned int i = threadIdx.x; staged && i < bins; i += blockDim.x) {
        const unsigned int c = s_counts[i];
        if (c) atomicAdd(&counts[i], (unsigned long long)c);
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

cudaError_t launch_hist_edges(unsigned int bins, double lo, double hi, DeviceResult* d_res, double* d_edges, cudaStream_t s) {
    if (!bins) return cudaSuccess;
    hist_edges_kernel<<<(bins + 256) / 256, 256, 0, s>>>(bins, lo, hi, d_res, d_edges, reinterpret_cast<float*>(d_edges + bins + 1));
    return cudaGetLastError();
}

cudaError_t launch_tail_hist(const void* x, int dtype, unsigned long long n, bool want_tail, unsigned int bins,
                             double lo, double hi, DeviceResult* d_res, double* d_edges, double* d_scratch,
                             unsigned long long* d_words, int sm_count, cudaStream_t stream) {
    cudaError_t e = cudaMemsetAsync(d_words, 0, (2 + (size_t)bins) * sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    float* d_edges_f = reinterpret_cast<float*>(d_edges + bins + 1);
    if (bins) {
        hist_edges_kernel<<<(bins + 256) / 256, 256, 0, stream>>>(bins, lo, hi, d_res, d_edges, d_edges_f);
        if ((e = cudaGetLastError()) != cudaSuccess) return e;
    }
    if (n == 0 || (!want_tail && !bins)) return cudaSuccess;
    unsigned long long blocks = (n + 256ull * 8 - 1) / (256ull * 8);
    if (blocks < 1) blocks = 1;
    if (blocks > 1024) blocks = 1024;
    if (blocks > (unsigned long long)sm_count * 6) blocks = (unsigned long long)sm_count * 6;
    const size_t esz = dtype == MC_F64 ? 8 : 4;
    size_t smem = bins ? ((size_t)bins + 2) * esz + (size_t)bins * 4 : 0;
    const int staged = smem <= 96 * 1024 ? 1 : 0;           // 8192 fp64 / 12288 fp32 bins; two blocks per SM
    if (!staged) smem = 0;
    if (dtype == MC_F64) {
        auto k = tail_hist_kernel<double>;
        if ((e = ensure_dynamic_smem(k, smem)) != cudaSuccess) return e;
        k<<<(unsigned)blocks, 256, smem, stream>>>(static_cast<const double*>(x), n, d_res, want_tail ? 1 : 0, bins,
                                                   d_edges, d_edges_f, d_scratch, d_words + 2, staged);
    } else {
        auto k = tail_hist_kernel<float>;
        if ((e = ensure_dynamic_smem(k, smem)) != cudaSuccess) return e;
        k<<<(unsigned)blocks, 256, smem, stream>>>(static_cast<const float*>(x), n, d_res, want_tail ? 1 : 0, bins,
                                                   d_edges, d_edges_f, d_scratch, d_words + 2, staged);
    }
    if ((e = cudaGetLastError()) != cudaSuccess) return e;
    if (want_tail) {
        tail_reduce_kernel<<<1, 256, 0, stream>>>(d_scratch, (int)blocks, d_words);
        e = cudaGetLastError();
    }
    return e;
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
