// mc_fused.cu -- kernels of the lean closing (see mc_fused.cuh): key histogram of terminal values that were not
// binned by their simulation kernel (K1b), plan (K3), closing pass (K4), candidate selection + result (K5), and
// the two peer-memory exchanges that ride inside K3 and K4.
//
// Replaces, for runs without a fan chart, the reference's reductions app.py:636-640 and 652-656: nine
// np.percentile calls (each a partition of all final prices on the host), mean / std / exceedance / VaR / CVaR.
#include <math.h>

#include "mc_fused.cuh"
#include "mc_launch.cuh"

namespace mc {

// ------------------------------------------------------------------ peer flags
constexpr unsigned long long PEER_TIMEOUT_NS = 20ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ unsigned long long now_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void flag_raise(unsigned long long* remote, unsigned long long seq) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" :: "l"(remote), "l"(seq) : "memory");
}
__device__ __forceinline__ bool flag_wait(const unsigned long long* local, unsigned long long seq) {
    const unsigned long long t0 = now_ns();
    for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(local) : "memory");
        if (v >= seq) return true;
        if (now_ns() - t0 > PEER_TIMEOUT_NS) return false;
        __nanosleep(100);
    }
}
__device__ __forceinline__ unsigned long long* flags_of(const FusedArgs& a, unsigned int owner, int which) {
    return reinterpret_cast<unsigned long long*>(a.peers.base[owner] + (which == 1 ? a.lay.off_flags1 : a.lay.off_flags2));
}
__device__ __forceinline__ unsigned int* hist_in(const FusedArgs& a, unsigned int owner, unsigned int from) {
    return reinterpret_cast<unsigned int*>(a.peers.base[owner] + a.lay.off_hist) + (size_t)from * FUSED_KB;
}
__device__ __forceinline__ unsigned long long* hdr_in(const FusedArgs& a, unsigned int owner, unsigned int from) {
    return reinterpret_cast<unsigned long long*>(a.peers.base[owner] + a.lay.off_hdr) + (size_t)from * FUSED_HDR_WORDS;
}
__device__ __forceinline__ unsigned int* nph_in(const FusedArgs& a, unsigned int owner, unsigned int from) {
    return reinterpret_cast<unsigned int*>(a.peers.base[owner] + a.lay.off_nph) + (size_t)from * FUSED_MAX_NPH_BINS;
}
// rank `owner`'s candidate lists live in its own buffer only; the peers read them there
template <typename Key>
__device__ __forceinline__ Key* cand_of(const FusedArgs& a, unsigned int owner) {
    return reinterpret_cast<Key*>(a.peers.base[owner] + a.lay.off_cand);
}
// header words a rank publishes with X2
enum { HW_N = 0, HW_ABOVE, HW_SUM_D, HW_SUM_D2, HW_MIN, HW_MAX, HW_TAIL_CNT, HW_TAIL_SUM, HW_OVERFLOW, HW_STATUS, HW_CAND = 10 };
static_assert(HW_CAND + FUSED_MAX_SLOTS <= FUSED_HDR_WORDS, "header too small");

__device__ __forceinline__ double dbl(unsigned long long w) { return __longlong_as_double((long long)w); }
__device__ __forceinline__ unsigned long long bits(double v) { return (unsigned long long)__double_as_longlong(v); }
__device__ __forceinline__ int bit_length64(unsigned long long v) { return v ? 64 - __clzll((long long)v) : 0; }

// ------------------------------------------------------------------ K1b: key histogram of stored terminals
template <typename Real>
__global__ void __launch_bounds__(256) terminal_hist_kernel(const Real* __restrict__ x, unsigned long long n, KeyWindow win,
                                                            unsigned int* __restrict__ g_khist) {
    __shared__ unsigned int s_h[FUSED_KB];
    for (int i = threadIdx.x; i < FUSED_KB; i += 256) s_h[i] = 0u;
    __syncthreads();
    const unsigned long long stride = (unsigned long long)gridDim.x * 256;
    unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x;
    for (; i + 3 * stride < n; i += 4 * stride) {
        const Real v0 = x[i], v1 = x[i + stride], v2 = x[i + 2 * stride], v3 = x[i + 3 * stride];
        atomicAdd(&s_h[window_bin(win, key32_of(v0))], 1u);
        atomicAdd(&s_h[window_bin(win, key32_of(v1))], 1u);
        atomicAdd(&s_h[window_bin(win, key32_of(v2))], 1u);
        atomicAdd(&s_h[window_bin(win, key32_of(v3))], 1u);
    }
    for (; i < n; i += stride) atomicAdd(&s_h[window_bin(win, key32_of(x[i]))], 1u);
    __syncthreads();
    for (int b = threadIdx.x; b < FUSED_KB; b += 256) {
        const unsigned int c = s_h[b];
        if (c) atomicAdd(&g_khist[b], c);
    }
}

static unsigned int pass_blocks(unsigned long long n, int sm_count, unsigned int per_block_min) {
    unsigned long long b = (n + per_block_min - 1) / per_block_min;
    if (b > 2ull * sm_count) b = 2ull * sm_count;
    return (unsigned int)(b < 1 ? 1 : b);
}

cudaError_t launch_terminal_hist(const FusedArgs& a, int sm_count, cudaStream_t s) {
    if (a.n == 0) return cudaSuccess;
    unsigned int* g = reinterpret_cast<unsigned int*>(a.peers.base[a.peers.me] + a.lay.off_hist) + (size_t)a.peers.me * FUSED_KB;
    const unsigned int blocks = pass_blocks(a.n, sm_count, 8192);
    if (a.dtype == MC_F64)
        terminal_hist_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(a.x), a.n, a.win, g);
    else
        terminal_hist_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(a.x), a.n, a.win, g);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ X1 + K3: the plan
__global__ void __launch_bounds__(1024) fused_plan_kernel(const FusedArgs a) {
    extern __shared__ unsigned long long s_cum[];          // [FUSED_KB] inclusive prefix sums of the summed histogram
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned int s_tbin[MAX_TARGETS];
    __shared__ int s_status;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int R = a.peers.n_ranks, me = a.peers.me;
    if (tid == 0) s_status = FUSED_OK;
    __syncthreads();
    if (R > 1) {            // X1: my histogram into every peer's inbox, then wait for theirs
        // read once, store to every peer (a load per peer would chain 2 (R - 1) L2 latencies in front of the stores)
        const uint4* src = reinterpret_cast<const uint4*>(hist_in(a, me, me));
        static_assert(FUSED_KB / 4 == 2 * 1024, "two 16-byte pieces per thread");
        const uint4 q0 = __ldcg(src + tid), q1 = __ldcg(src + tid + 1024);
        for (unsigned int r = 0; r < R; ++r) {
            if (r == me) continue;
            uint4* dst = reinterpret_cast<uint4*>(hist_in(a, r, me));
            dst[tid] = q0;
            dst[tid + 1024] = q1;
        }
        __threadfence_system();
        __syncthreads();
        if ((unsigned)tid < R && (unsigned)tid != me) {
            flag_raise(flags_of(a, tid, 1) + me, a.peers.seq);
            if (!flag_wait(flags_of(a, me, 1) + tid, a.peers.seq)) s_status = FUSED_PEER_TIMEOUT;
        }
        __syncthreads();
    }
    constexpr int PER = FUSED_KB / 1024;
    unsigned long long c[PER], run = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        unsigned long long v = 0;
        for (unsigned int r = 0; r < R; ++r) v += __ldcg(hist_in(a, me, r) + tid * PER + k);      // peers wrote these: not through L1
        run += v;
        c[k] = run;
    }
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
    for (int k = 0; k < PER; ++k) s_cum[tid * PER + k] = off + c[k];
    for (int i = tid; i < FUSED_KB / 4; i += 1024) reinterpret_cast<uint32_t*>(a.member)[i] = 0u;
    __syncthreads();
    const unsigned long long total = s_cum[FUSED_KB - 1];
    const int nt = a.n_targets;
    if (tid < nt) {
        const unsigned long long rank = a.d_plan->rank[tid];
        int lo = 0, hi = FUSED_KB - 1;             // first bin with cum > rank
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (s_cum[mid] > rank) hi = mid; else lo = mid + 1; }
        s_tbin[tid] = (unsigned int)lo;
    }
    __syncthreads();
    if (warp != 0) return;
    // one warp builds the plan: lane j = target j (targets are sorted by rank, so their bins are non-decreasing)
    FusedPlan& P = *a.plan;
    static_assert(MAX_TARGETS <= 32, "one lane per target");
    const bool has = lane < nt;
    const unsigned int b = has ? s_tbin[lane] : 0u;
    const bool first = has && (lane == 0 || s_tbin[lane - 1] != b);
    const unsigned int firsts = __ballot_sync(0xffffffffu, first);
    const int slot = __popc(firsts & ((2u << lane) - 1u)) - 1;      // slot of this lane's target
    const int n_slots = __popc(firsts);
    unsigned long long below = 0, pop = 0, cap = 0;
    if (first) {
        below = b ? s_cum[b - 1] : 0ull;
        pop = s_cum[b] - below;
        cap = pop;
        if (R > 1) {          // iid paths: a rank holds pop / R of the bin, give or take a few sqrt
            const double share = (double)pop / (double)R;
            cap = (unsigned long long)(share + 8.0 * sqrt(share) + 64.0);
            if (cap > pop) cap = pop;
        }
        cap = (cap + 3ull) & ~3ull;
        if (cap > 0x7FFFFFF0ull) cap = 0x7FFFFFF0ull;
    }
    unsigned long long used = cap;                                  // inclusive scan of the capacities
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long o = __shfl_up_sync(0xffffffffu, used, d);
        if (lane >= d) used += o;
    }
    const unsigned long long cap_total = a.lay.cand_bytes / (a.dtype == MC_F64 ? 8u : 4u);
    const unsigned long long all_used = __shfl_sync(0xffffffffu, used, 31);
    int status = s_status;
    if (status == FUSED_OK && total != a.d_plan->n_global) status = FUSED_INCONSISTENT;
    if (status == FUSED_OK && all_used > cap_total) status = FUSED_FALLBACK;
    if (first) {
        P.bin[slot] = b; P.below[slot] = below; P.pop[slot] = pop; P.cap[slot] = (unsigned int)cap;
        P.off[slot] = (unsigned int)((used - cap) < cap_total ? (used - cap) : 0ull);
        if (status == FUSED_OK) a.member[b] = (unsigned char)(slot + 1);
    }
    if (has) { P.target_slot[lane] = slot; P.key[lane] = 0ull; }
    P.cand_count[lane] = 0u;
    const int tlo = a.tail_idx >= 0 && nt ? a.d_plan->pct_lo_idx[a.tail_idx] : -1;
    const int thi = a.tail_idx >= 0 && nt ? a.d_plan->pct_hi_idx[a.tail_idx] : -1;
    const int slot_lo = __shfl_sync(0xffffffffu, slot, tlo < 0 ? 0 : tlo), slot_hi = __shfl_sync(0xffffffffu, slot, thi < 0 ? 0 : thi);
    const unsigned int bin_lo = __shfl_sync(0xffffffffu, b, tlo < 0 ? 0 : tlo);
    if (lane == 0) {
        P.n_slots = n_slots;
        P.tail_slot_lo = tlo >= 0 ? slot_lo : -1;
        P.tail_slot_hi = thi >= 0 ? slot_hi : -1;
        P.tail_bin = tlo >= 0 ? bin_lo : 0u;
        P.tail_below = tlo >= 0 && bin_lo ? s_cum[bin_lo - 1] : 0ull;
        P.tail_lo_cnt = 0ull; P.tail_hi_mult = 0ull; P.tail_lo_sum = 0.0;
        P.status = status;
        P.overflow = 0u; P.ticket4 = 0u; P.ticket5 = 0u;
    }
}

cudaError_t launch_fused_plan(const FusedArgs& a, cudaStream_t s) {
    const size_t smem = (size_t)FUSED_KB * sizeof(unsigned long long);
    cudaError_t e = ensure_dynamic_smem(fused_plan_kernel, smem);
    if (e != cudaSuccess) return e;
    fused_plan_kernel<<<1, 1024, smem, s>>>(a);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ K4 + X2: the closing pass
constexpr int CLOSE_THREADS = 256;
constexpr int CLOSE_TILE = 4 * CLOSE_THREADS;
constexpr int CLOSE_STG = 2048;                 // staged candidates per block between two flushes (>= 2 tiles)
constexpr unsigned int CLOSE_NPH_SMEM = 8192;   // fixed-bin histogram staged in shared memory up to this many bins

__device__ __forceinline__ void combine(ClosePartial& a, const ClosePartial& b) {
    a.sum_d += b.sum_d; a.sum_d2 += b.sum_d2;
    a.vmin = fmin(a.vmin, b.vmin); a.vmax = fmax(a.vmax, b.vmax);
    a.tail_sum += b.tail_sum;
    a.n += b.n; a.n_above += b.n_above; a.tail_cnt += b.tail_cnt;
}
__device__ __forceinline__ ClosePartial empty_close() {
    ClosePartial p;
    p.sum_d = 0.0; p.sum_d2 = 0.0; p.vmin = INFINITY; p.vmax = -INFINITY; p.tail_sum = 0.0;
    p.n = 0; p.n_above = 0; p.tail_cnt = 0;
    return p;
}
__device__ __forceinline__ ClosePartial warp_reduce(ClosePartial p) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        ClosePartial o;
        o.sum_d = __shfl_down_sync(0xffffffffu, p.sum_d, d);
        o.sum_d2 = __shfl_down_sync(0xffffffffu, p.sum_d2, d);
        o.vmin = __shfl_down_sync(0xffffffffu, p.vmin, d);
        o.vmax = __shfl_down_sync(0xffffffffu, p.vmax, d);
        o.tail_sum = __shfl_down_sync(0xffffffffu, p.tail_sum, d);
        o.n = __shfl_down_sync(0xffffffffu, p.n, d);
        o.n_above = __shfl_down_sync(0xffffffffu, p.n_above, d);
        o.tail_cnt = __shfl_down_sync(0xffffffffu, p.tail_cnt, d);
        combine(p, o);
    }
    return p;
}
// fixed order: lanes, then warps 0..7 -- valid in thread 0
__device__ __forceinline__ ClosePartial block_reduce(ClosePartial p, ClosePartial* s_part /*[CLOSE_THREADS / 32]*/) {
    p = warp_reduce(p);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = p;
    __syncthreads();
    if (threadIdx.x == 0)
        for (int w = 1; w < CLOSE_THREADS / 32; ++w) combine(p, s_part[w]);
    return p;
}

// What a thread keeps per value.  fp32 data: minimum, maximum and the exceedance test stay in fp32 (exact: the
// comparison v > s0 against a double s0 equals v > floor_f32(s0)); only the two sums are fp64.
template <typename Real>
struct CloseAcc {
    double sum_d = 0.0, sum_d2 = 0.0, tail_sum = 0.0;
    Real vmin, vmax;
    unsigned int n_above = 0u, tail_cnt = 0u, n = 0u;     // per thread: far below 2^32
};

template <typename Real>
__global__ void __launch_bounds__(CLOSE_THREADS, 4) fused_closing_kernel(const FusedArgs a) {
    using Key = typename KeyT<Real>::type;
    extern __shared__ __align__(16) unsigned char smem_dyn[];      // fixed-bin histogram: edges, counts (when staged)
    __shared__ __align__(16) unsigned char s_member[FUSED_KB];
    __shared__ Key s_stg_key[CLOSE_STG];
    __shared__ unsigned short s_stg_pos[CLOSE_STG];
    __shared__ unsigned char s_stg_slot[CLOSE_STG];
    __shared__ unsigned int s_stg_n, s_slot_cnt[FUSED_MAX_SLOTS], s_slot_base[FUSED_MAX_SLOTS];
    __shared__ ClosePartial s_part[CLOSE_THREADS / 32];
    __shared__ int s_last;
    const int tid = threadIdx.x;
    const unsigned int R = a.peers.n_ranks, me = a.peers.me;
    FusedPlan& P = *a.plan;
    const bool live = P.status == FUSED_OK;
    const unsigned int tail_bin = P.tail_bin;
    const Real* x = static_cast<const Real*>(a.x);

    for (int i = tid; i < FUSED_KB / 16; i += CLOSE_THREADS)
        reinterpret_cast<uint4*>(s_member)[i] = reinterpret_cast<const uint4*>(a.member)[i];
    if (tid < FUSED_MAX_SLOTS) s_slot_cnt[tid] = 0u;
    if (tid == 0) s_stg_n = 0u;
    // fixed-bin histogram (np.histogram semantics), unless the simulation kernel already took it
    const int bins = (live && a.nph_bins && !a.nph_in_sim) ? (int)a.nph_bins : 0;
    const bool staged = bins && (unsigned)bins <= CLOSE_NPH_SMEM;
    const Real* g_edge = sizeof(Real) == 8 ? reinterpret_cast<const Real*>(a.edges)
                                           : reinterpret_cast<const Real*>(a.edges + a.nph_bins + 1);
    Real* s_edge = reinterpret_cast<Real*>(smem_dyn);
    unsigned int* s_cnt = reinterpret_cast<unsigned int*>(s_edge + (bins + 2));
    if (staged) {
        for (int i = tid; i <= bins; i += CLOSE_THREADS) s_edge[i] = g_edge[i];
        for (int i = tid; i < bins; i += CLOSE_THREADS) s_cnt[i] = 0u;
    }
    const Real* edge = staged ? s_edge : g_edge;
    __syncthreads();
    Real h_lo = Real(0), h_top = Real(0), h_inv = Real(0);
    if (bins) { h_lo = edge[0]; h_top = edge[bins]; h_inv = (Real)bins / (h_top - h_lo); }
    unsigned int* my_nph = nph_in(a, me, me);

    // append the staged candidates to this rank's lists (one per slot)
    auto flush = [&]() {
        const unsigned int n_stg = s_stg_n;
        if (n_stg == 0u) return;                   // uniform: read behind a barrier
        for (unsigned int i = tid; i < n_stg; i += CLOSE_THREADS)
            s_stg_pos[i] = (unsigned short)atomicAdd(&s_slot_cnt[s_stg_slot[i]], 1u);
        __syncthreads();
        if (tid < FUSED_MAX_SLOTS) {
            const unsigned int c = s_slot_cnt[tid];
            s_slot_base[tid] = c ? atomicAdd(&P.cand_count[tid], c) : 0u;
        }
        __syncthreads();
        for (unsigned int i = tid; i < n_stg; i += CLOSE_THREADS) {
            const unsigned int sl = s_stg_slot[i];
            const unsigned int idx = s_slot_base[sl] + s_stg_pos[i];
            if (idx < P.cap[sl]) {
                const size_t e = (size_t)P.off[sl] + idx;
                cand_of<Key>(a, me)[e] = s_stg_key[i];       // peers read the list over NVLink in K5
            } else {
                atomicOr(&P.overflow, 1u);
            }
        }
        __syncthreads();
        if (tid < FUSED_MAX_SLOTS) s_slot_cnt[tid] = 0u;
        if (tid == 0) s_stg_n = 0u;
        __syncthreads();
    };

    CloseAcc<Real> acc;
    acc.vmin = (Real)INFINITY; acc.vmax = (Real)(-INFINITY);
    // v > center (a double) for a float v  <=>  v > floor_f32(center)
    Real center_r = (Real)a.center;
    if (sizeof(Real) == 4 && (double)center_r > a.center) center_r = (Real)nextafterf((float)center_r, -INFINITY);
    const KeyWindow win = a.win;
    auto visit = [&](Real v) {
        const double dv = (double)v, d = dv - a.center;
        acc.sum_d += d;
        acc.sum_d2 = fma(d, d, acc.sum_d2);
        acc.vmin = v < acc.vmin ? v : acc.vmin;
        acc.vmax = v > acc.vmax ? v : acc.vmax;
        acc.n_above += v > center_r ? 1u : 0u;                      // app.py:654
        const unsigned int bin = window_bin(win, key32_of(v));
        if (bin < tail_bin) { acc.tail_sum += dv; acc.tail_cnt += 1u; }
        const unsigned int m = s_member[bin];
        if (m) {
            const unsigned int pos = atomicAdd(&s_stg_n, 1u);
            s_stg_key[pos] = key_of(v);
            s_stg_slot[pos] = (unsigned char)(m - 1u);
        }
        if (bins) {
            const int b = np_hist_bin<Real>(v, edge, bins, h_lo, h_top, h_inv);
            if (b >= 0) { if (staged) atomicAdd(&s_cnt[b], 1u); else atomicAdd(&my_nph[b], 1u); }
        }
    };
    if (live) {
        unsigned long long per = (a.n + gridDim.x - 1) / gridDim.x;
        per = (per + CLOSE_TILE - 1) / CLOSE_TILE * CLOSE_TILE;
        unsigned long long beg = (unsigned long long)blockIdx.x * per;
        if (beg > a.n) beg = a.n;
        const unsigned long long end = beg + per < a.n ? beg + per : a.n;
        // whole tiles: the loads of the NEXT tile are issued before this one is worked on (a barrier closes every tile,
        // so without this a block waits out a full L2 / HBM latency per 4 values a thread: issue 39 %, ncu)
        const unsigned long long n_full = (end - beg) / CLOSE_TILE;
        Real v[4], vn[4];
        if (n_full) {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = x[beg + tid + (unsigned long long)k * CLOSE_THREADS];
        }
        for (unsigned long long q = 0; q < n_full; ++q) {
            const unsigned long long nxt = beg + (q + 1) * CLOSE_TILE;
            if (q + 1 < n_full) {
#pragma unroll
                for (int k = 0; k < 4; ++k) vn[k] = x[nxt + tid + (unsigned long long)k * CLOSE_THREADS];
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) visit(v[k]);
            acc.n += 4u;
            // every thread looks at the fill level BEFORE the barrier and the barrier ORs what they saw: the last
            // one to arrive sees the final level, and nobody appends again before all have looked
            if (__syncthreads_or(*reinterpret_cast<volatile unsigned int*>(&s_stg_n) > (unsigned int)(CLOSE_STG - CLOSE_TILE))) flush();
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = vn[k];
        }
        for (unsigned long long i = beg + n_full * CLOSE_TILE + tid; i < end; i += CLOSE_THREADS) { visit(x[i]); acc.n += 1u; }   // < one tile
        __syncthreads();
        flush();
    }
    if (staged) {
        __syncthreads();
        for (int i = tid; i < bins; i += CLOSE_THREADS) {
            const unsigned int c = s_cnt[i];
            if (c) atomicAdd(&my_nph[i], c);
        }
    }
    ClosePartial mine;
    mine.sum_d = acc.sum_d; mine.sum_d2 = acc.sum_d2; mine.vmin = (double)acc.vmin; mine.vmax = (double)acc.vmax;
    mine.tail_sum = acc.tail_sum; mine.n = acc.n; mine.n_above = acc.n_above; mine.tail_cnt = acc.tail_cnt;
    const ClosePartial blockp = block_reduce(mine, s_part);
    if (tid == 0) a.partials[blockIdx.x] = blockp;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = atomicAdd(&P.ticket4, 1u) == gridDim.x - 1 ? 1 : 0;
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    // the last block: this rank's partials in fixed order, then the header and the fixed-bin counts go out (X2)
    ClosePartial tot = empty_close();
    for (unsigned int i = tid; i < gridDim.x; i += CLOSE_THREADS) {
        const ulonglong2* src = reinterpret_cast<const ulonglong2*>(a.partials + i);
        static_assert(sizeof(ClosePartial) == 64, "four 16-byte loads");
        const ulonglong2 q0 = __ldcg(src), q1 = __ldcg(src + 1), q2 = __ldcg(src + 2), q3 = __ldcg(src + 3);
        ClosePartial p;
        p.sum_d = dbl(q0.x); p.sum_d2 = dbl(q0.y); p.vmin = dbl(q1.x); p.vmax = dbl(q1.y); p.tail_sum = dbl(q2.x);
        p.n = q2.y; p.n_above = q3.x; p.tail_cnt = q3.y;
        combine(tot, p);
    }
    __syncthreads();
    tot = block_reduce(tot, s_part);
    const unsigned int my_overflow = *reinterpret_cast<volatile unsigned int*>(&P.overflow);
    const unsigned int my_cand = tid < FUSED_MAX_SLOTS ? *reinterpret_cast<volatile unsigned int*>(&P.cand_count[tid]) : 0u;
    const int my_status = P.status;
    for (unsigned int r = 0; r < R; ++r) {
        unsigned long long* h = hdr_in(a, r, me);
        if (tid == 0) {
            h[HW_N] = tot.n; h[HW_ABOVE] = tot.n_above; h[HW_SUM_D] = bits(tot.sum_d); h[HW_SUM_D2] = bits(tot.sum_d2);
            h[HW_MIN] = bits(tot.vmin); h[HW_MAX] = bits(tot.vmax); h[HW_TAIL_CNT] = tot.tail_cnt;
            h[HW_TAIL_SUM] = bits(tot.tail_sum);
            h[HW_OVERFLOW] = my_overflow;
            h[HW_STATUS] = (unsigned long long)my_status;
        }
        if (tid < FUSED_MAX_SLOTS) h[HW_CAND + tid] = my_cand;
    }
    // this rank's fixed-bin counts to every peer: each piece is read once and stored R - 1 times
    for (unsigned int i0 = 0; R > 1 && i0 < a.nph_bins; i0 += 8u * CLOSE_THREADS) {
        unsigned int c[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { const unsigned int i = i0 + tid + k * CLOSE_THREADS; c[k] = i < a.nph_bins ? __ldcg(my_nph + i) : 0u; }
        for (unsigned int r = 0; r < R; ++r) {
            if (r == me) continue;
            unsigned int* dst = nph_in(a, r, me);
#pragma unroll
            for (int k = 0; k < 8; ++k) { const unsigned int i = i0 + tid + k * CLOSE_THREADS; if (i < a.nph_bins) dst[i] = c[k]; }
        }
    }
    if (R > 1) {
        __threadfence_system();
        __syncthreads();
        if ((unsigned)tid < R && (unsigned)tid != me) flag_raise(flags_of(a, tid, 2) + me, a.peers.seq);
    }
}

cudaError_t launch_fused_closing(const FusedArgs& a, int sm_count, int* grid, cudaStream_t s) {
    const unsigned int bins = (a.nph_bins && !a.nph_in_sim && a.nph_bins <= CLOSE_NPH_SMEM) ? a.nph_bins : 0u;
    const size_t esz = a.dtype == MC_F64 ? 8 : 4;
    const size_t smem = bins ? ((size_t)bins + 2) * esz + (size_t)bins * 4 : 0;
    cudaError_t e;
    // ONE wave of blocks: as many as are resident at once (each starts by copying the 8 KB membership table)
    int per_sm = 0;
    if (a.dtype == MC_F64) {
        if ((e = ensure_dynamic_smem(fused_closing_kernel<double>, smem)) != cudaSuccess) return e;
        e = resident_blocks_per_sm(fused_closing_kernel<double>, CLOSE_THREADS, smem, &per_sm);
    } else {
        if ((e = ensure_dynamic_smem(fused_closing_kernel<float>, smem)) != cudaSuccess) return e;
        e = resident_blocks_per_sm(fused_closing_kernel<float>, CLOSE_THREADS, smem, &per_sm);
    }
    if (e != cudaSuccess) return e;
    if (per_sm < 1) per_sm = 1;
    unsigned long long nb = (a.n + 2ull * CLOSE_TILE - 1) / (2ull * CLOSE_TILE);
    if (nb > (unsigned long long)per_sm * sm_count) nb = (unsigned long long)per_sm * sm_count;
    const unsigned int blocks = (unsigned int)(nb < 1 ? 1 : nb);
    *grid = (int)blocks;
    if (a.dtype == MC_F64) fused_closing_kernel<double><<<blocks, CLOSE_THREADS, smem, s>>>(a);
    else fused_closing_kernel<float><<<blocks, CLOSE_THREADS, smem, s>>>(a);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ K5: selection among the candidates + the result
constexpr int SEL_THREADS = 1024;
constexpr int SUB_BINS = 4096;
constexpr int SEL_SORT = 1024;                   // a key range holding <= this many candidates is sorted
constexpr size_t SEL_CAND_SMEM = 160 * 1024;     // a slot's candidates are held in shared memory when they fit here

// key range of a bin of the window (32-bit keys), widened to the key type
template <typename Key>
__device__ __forceinline__ void bin_key_range(const KeyWindow& w, unsigned int b, Key* lo, Key* hi) {
    const unsigned long long span = (unsigned long long)w.b - w.a;
    const unsigned int jmax = __umulhi(w.b - w.a, w.mult);
    const unsigned long long j = b;
    unsigned long long dl = ((j << 32) + w.mult - 1ull) / w.mult, dh = (((j + 1ull) << 32) + w.mult - 1ull) / w.mult - 1ull;
    if (dl > span) dl = span;
    if (dh > span) dh = span;
    const uint32_t l32 = b == 0u ? 0u : w.a + (uint32_t)dl;
    const uint32_t h32 = b >= jmax ? 0xFFFFFFFFu : w.a + (uint32_t)dh;
    if (sizeof(Key) == 8) { *lo = (Key)((unsigned long long)l32 << 32); *hi = (Key)(((unsigned long long)h32 << 32) | 0xFFFFFFFFull); }
    else { *lo = (Key)l32; *hi = (Key)h32; }
}

template <typename Key> struct ValueOf;
template <> struct ValueOf<uint32_t> { static __device__ __forceinline__ double get(uint32_t k) { return (double)value_of(k); } };
template <> struct ValueOf<uint64_t> { static __device__ __forceinline__ double get(uint64_t k) { return value_of(k); } };

template <typename Key>
__device__ void bitonic_sort(Key* s, int P) {                // P a power of two; all threads of the block
    for (int k = 2; k <= P; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < P; i += SEL_THREADS) {
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

// a key range of one histogram bin that still holds undecided order statistics
struct SelItem {
    unsigned long long lo, hi;       // inclusive
    unsigned long long before;       // candidates of the bin below lo
    unsigned int pop;                // candidates inside
    int t0, nt;                      // its targets: s_tl[t0 .. t0 + nt)
};
constexpr int SEL_ITEMS = 192;

template <typename Key>
__global__ void __launch_bounds__(SEL_THREADS) fused_select_kernel(const FusedArgs a) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    unsigned int* s_sub = reinterpret_cast<unsigned int*>(smem_dyn);                  // [SUB_BINS]
    Key* s_keys = reinterpret_cast<Key*>(s_sub + SUB_BINS);                           // [SEL_SORT]
    Key* s_cand = s_keys + SEL_SORT;                                                  // [SEL_CAND_SMEM / sizeof(Key)]
    __shared__ unsigned int s_cnt[FUSED_MAX_RANKS];
    __shared__ unsigned int s_warp[32];
    __shared__ unsigned int s_n;
    __shared__ SelItem s_items[SEL_ITEMS];
    __shared__ int s_tl[MAX_TARGETS];
    __shared__ int s_nitems;
    __shared__ int s_flag;
    __shared__ double s_pct[MC_MAX_PCTS];
    __shared__ double s_red[32];
    __shared__ unsigned long long s_redc[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned int R = a.peers.n_ranks, me = a.peers.me;
    FusedPlan& P = *a.plan;
    const int slot = blockIdx.x;

    if (tid == 0) s_flag = FUSED_OK;
    __syncthreads();
    if (R > 1) {                                    // X2: every rank's candidates and header have landed here
        if ((unsigned)tid < R && (unsigned)tid != me && !flag_wait(flags_of(a, me, 2) + tid, a.peers.seq)) s_flag = FUSED_PEER_TIMEOUT;
        __syncthreads();
    }
    if (tid < (int)R) {
        const unsigned long long* h = hdr_in(a, me, tid);
        if (__ldcg(h + HW_OVERFLOW) != 0ull && s_flag == FUSED_OK) s_flag = FUSED_FALLBACK;
        const unsigned long long st = __ldcg(h + HW_STATUS);
        if (st != FUSED_OK && s_flag == FUSED_OK) s_flag = (int)st;
        s_cnt[tid] = slot < FUSED_MAX_SLOTS ? (unsigned int)__ldcg(h + HW_CAND + slot) : 0u;
    }
    __syncthreads();
    int status = P.status != FUSED_OK ? P.status : s_flag;
    const bool active = status == FUSED_OK && slot < P.n_slots;
    if (active) {
        unsigned int total = 0;
        for (unsigned int r = 0; r < R; ++r) total += s_cnt[r];
        const size_t off = P.off[slot];
        // The slot's candidates are copied into shared memory when they fit: whole keys, or -- fp32 keys of one bin of the
        // window differ in their low bits only (a bin is ~4000 - 12000 consecutive keys) -- as 16-bit offsets from the
        // bin's first key, which doubles what fits.  At 8 ranks x 10M paths a central bin holds ~50,000 values: beyond the
        // 40,960 whole keys, and every later look at them was another sweep over NVLink (K5 was 64 us at 8 GPUs against 25 us on one; now 33 us).
        Key bin_lo, bin_hi;
        bin_key_range<Key>(a.win, P.bin[slot], &bin_lo, &bin_hi);
        const bool in_smem = (size_t)total * sizeof(Key) <= SEL_CAND_SMEM;
        const bool in_smem16 = !in_smem && sizeof(Key) == 4 && (unsigned long long)(bin_hi - bin_lo) <= 0xFFFFull &&
                               (size_t)total * 2 <= SEL_CAND_SMEM;
        unsigned short* s_cand16 = reinterpret_cast<unsigned short*>(s_cand);
        // every candidate of the slot, eight loads in flight per thread: from the ranks' lists, or -- after the
        // first such sweep has copied them -- from shared memory.
        // Warps are dealt out to the ranks (warp w reads rank w mod R's list), so the loads from every peer are in flight
        // together -- rank after rank, each list cost its own NVLink round trip: 51 us instead of 23 at 8 GPUs -- and a
        // candidate's address is its list plus an index.  (One flat index space over all lists did the same, but finding
        // the list of position g took ~80 instructions per candidate: 16 us of K5 at 8 ranks x 10M paths.)
        auto sweep_lists = [&](auto f) {
            const unsigned int r = (unsigned int)warp % R, sub = (unsigned int)warp / R;
            const unsigned int n_sub = ((unsigned int)(SEL_THREADS / 32) - 1u - r) / R + 1u;       // warps on rank r's list
            const bool idle = R > (unsigned int)(SEL_THREADS / 32);                                 // (R <= 16: never)
            const Key* list = cand_of<Key>(a, (int)r) + off;
            unsigned int base = 0;                          // where this rank's candidates start in the concatenation
            for (unsigned int q = 0; q < r; ++q) base += s_cnt[q];
            const unsigned int cnt = idle ? 0u : s_cnt[r], step = 32u * n_sub;
            // 16-byte loads (the lists start on 16-byte boundaries: capacities are multiples of four entries), four in
            // flight per lane: 64 KB per block on the wire instead of 32 KB in 4-byte pieces -- a sweep over 8 ranks'
            // lists is a handful of NVLink round trips
            constexpr unsigned int VEC = 16u / (unsigned int)sizeof(Key);
            constexpr int U = 4;
            const uint4* list4 = reinterpret_cast<const uint4*>(list);
            const unsigned int n_vec = cnt / VEC;
            auto give = [&](const uint4& q, unsigned int at) {       // at: index of the vector's first entry
                if constexpr (sizeof(Key) == 4) {
                    f((Key)q.x, base + at); f((Key)q.y, base + at + 1u); f((Key)q.z, base + at + 2u); f((Key)q.w, base + at + 3u);
                } else {
                    f((Key)(((unsigned long long)q.y << 32) | q.x), base + at);
                    f((Key)(((unsigned long long)q.w << 32) | q.z), base + at + 1u);
                }
            };
            for (unsigned int j = sub * 32u + lane; j < n_vec; j += (unsigned int)U * step) {
                uint4 q[U];
#pragma unroll
                for (int u = 0; u < U; ++u) q[u] = j + u * step < n_vec ? __ldcg(list4 + j + u * step) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                for (int u = 0; u < U; ++u) if (j + u * step < n_vec) give(q[u], (j + u * step) * VEC);
            }
            for (unsigned int j = n_vec * VEC + sub * 32u + lane; j < cnt; j += step) f(__ldcg(list + j), base + j);
        };
        if (tid == 0) {                 // work list: key ranges of this bin that still hold undecided targets
            int m = 0;
            for (int t = 0; t < a.n_targets; ++t) if (P.target_slot[t] == slot) s_tl[m++] = t;
            s_items[0] = SelItem{(unsigned long long)bin_lo, (unsigned long long)bin_hi, 0ull, total, 0, m};
            s_nitems = 1;
        }
        __syncthreads();
        if (in_smem) sweep_lists([&](Key k, unsigned int pos) { s_cand[pos] = k; });
        else if (in_smem16) sweep_lists([&](Key k, unsigned int pos) { s_cand16[pos] = (unsigned short)(k - bin_lo); });
        __syncthreads();
        auto for_each = [&](auto f) {
            if (in_smem) { for (unsigned int j = tid; j < total; j += SEL_THREADS) f(s_cand[j]); }
            else if (in_smem16) { for (unsigned int j = tid; j < total; j += SEL_THREADS) f((Key)(bin_lo + (Key)s_cand16[j])); }
            else sweep_lists([&](Key k, unsigned int) { f(k); });
        };
        if ((unsigned long long)total != P.pop[slot]) status = FUSED_INCONSISTENT;
        for (int head = 0; status == FUSED_OK && head < s_nitems; ++head) {
            const SelItem it = s_items[head];
            const Key lo = (Key)it.lo, hi = (Key)it.hi;
            if (lo == hi) {
                if (tid < it.nt) P.key[s_tl[it.t0 + tid]] = it.lo;
                continue;
            }
            if (it.pop <= (unsigned int)SEL_SORT) {      // few enough: collect them, sort, read the targets off
                if (tid == 0) s_n = 0u;
                __syncthreads();
                for_each([&](Key k) {
                    if (k >= lo && k <= hi) { const unsigned int p = atomicAdd(&s_n, 1u); if (p < (unsigned int)SEL_SORT) s_keys[p] = k; }
                });
                __syncthreads();
                if (s_n != it.pop) { status = FUSED_INCONSISTENT; break; }
                int Pw = 1;
                while (Pw < (int)it.pop) Pw <<= 1;
                for (int i = (int)it.pop + tid; i < Pw; i += SEL_THREADS) s_keys[i] = ~(Key)0;
                __syncthreads();
                bitonic_sort(s_keys, Pw);
                if (tid < it.nt) {
                    const int t = s_tl[it.t0 + tid];
                    P.key[t] = (unsigned long long)s_keys[a.d_plan->rank[t] - P.below[slot] - it.before];
                }
                __syncthreads();
                continue;
            }
            // too many to sort: 4096 sub-bins over the range, every target moves into the sub-bin that holds its rank
            const int shift = max(0, bit_length64((unsigned long long)(hi - lo)) - 12);
            for (int i = tid; i < SUB_BINS; i += SEL_THREADS) s_sub[i] = 0u;
            __syncthreads();
            for_each([&](Key k) { if (k >= lo && k <= hi) atomicAdd(&s_sub[(unsigned int)((k - lo) >> shift)], 1u); });
            __syncthreads();
            unsigned int c4[4], run = 0;                   // inclusive scan, four sub-bins per thread
#pragma unroll
            for (int k = 0; k < 4; ++k) { run += s_sub[tid * 4 + k]; c4[k] = run; }
            unsigned int incl = run;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned int o = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += o;
            }
            if (lane == 31) s_warp[warp] = incl;
            __syncthreads();
            unsigned int offs = incl - run;
            for (int w = 0; w < warp; ++w) offs += s_warp[w];
#pragma unroll
            for (int k = 0; k < 4; ++k) s_sub[tid * 4 + k] = offs + c4[k];
            __syncthreads();
            if (tid == 0) {
                int prev_sb = -1;
                if (s_sub[SUB_BINS - 1] != it.pop) s_flag = FUSED_INCONSISTENT;
                for (int q = 0; q < it.nt && s_flag == FUSED_OK; ++q) {
                    const int t = s_tl[it.t0 + q];
                    const unsigned long long want = a.d_plan->rank[t] - P.below[slot] - it.before;
                    int l2 = 0, h2 = SUB_BINS - 1;         // first sub-bin with cum > want
                    while (l2 < h2) { const int mid = (l2 + h2) >> 1; if ((unsigned long long)s_sub[mid] > want) h2 = mid; else l2 = mid + 1; }
                    if (l2 == prev_sb) { s_items[s_nitems - 1].nt += 1; continue; }
                    prev_sb = l2;
                    if (s_nitems >= SEL_ITEMS) { s_flag = FUSED_INCONSISTENT; break; }
                    const unsigned int before = l2 ? s_sub[l2 - 1] : 0u;
                    const Key nlo = lo + ((Key)l2 << shift);
                    const Key span = shift ? (((Key)1 << shift) - 1) : 0;
                    const Key nhi = (hi - nlo < span) ? hi : nlo + span;
                    s_items[s_nitems++] = SelItem{(unsigned long long)nlo, (unsigned long long)nhi, it.before + before,
                                                  s_sub[l2] - before, it.t0 + q, 1};
                }
            }
            __syncthreads();
            if (s_flag != FUSED_OK) status = s_flag;
        }
        // CVaR: this bin's share of the tail -- the candidates at or below the lower order statistic of the tail
        // percentile (count, sum), and the multiplicity of the upper one (it joins the tail only when the
        // interpolated threshold reaches it)
        __syncthreads();
        const bool is_lo = status == FUSED_OK && a.tail_idx >= 0 && slot == P.tail_slot_lo;
        const bool is_hi = status == FUSED_OK && a.tail_idx >= 0 && slot == P.tail_slot_hi;
        if (is_lo || is_hi) {
            const Key klo = (Key)*reinterpret_cast<volatile unsigned long long*>(&P.key[a.d_plan->pct_lo_idx[a.tail_idx]]);
            const Key khi = (Key)*reinterpret_cast<volatile unsigned long long*>(&P.key[a.d_plan->pct_hi_idx[a.tail_idx]]);
            double tsum = 0.0;
            unsigned long long tcnt = 0, mult = 0;
            for_each([&](Key k) {
                if (is_lo && k <= klo) { tsum += ValueOf<Key>::get(k); tcnt += 1; }
                if (is_hi && k == khi) mult += 1;
            });
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                tsum += __shfl_down_sync(0xffffffffu, tsum, d);
                tcnt += __shfl_down_sync(0xffffffffu, tcnt, d);
                mult += __shfl_down_sync(0xffffffffu, mult, d);
            }
            if (lane == 0) { s_red[warp] = tsum; s_redc[warp] = tcnt; s_warp[warp] = (unsigned int)mult; }
            __syncthreads();
            if (tid == 0) {
                unsigned long long m2 = 0;
                tsum = 0.0; tcnt = 0;
                for (int w = 0; w < SEL_THREADS / 32; ++w) { tsum += s_red[w]; tcnt += s_redc[w]; m2 += s_warp[w]; }
                if (is_lo) { P.tail_lo_cnt = tcnt; P.tail_lo_sum = tsum; }
                if (is_hi) P.tail_hi_mult = m2;
            }
        }
    }
    if (status != FUSED_OK && tid == 0) atomicMax(&P.status, status);
    __threadfence();
    __syncthreads();
    if (tid == 0) s_flag = atomicAdd(&P.ticket5, 1u) == gridDim.x - 1 ? 1 : 0;
    __syncthreads();
    if (!s_flag) return;
    __threadfence();

    // ---- the last block: percentiles, moments, tail, histogram -> device result and pinned host memory
    const SelectPlan& sp = *a.d_plan;
    const int final_status = *reinterpret_cast<volatile int*>(&P.status);
    const bool want_tail = a.tail_idx >= 0 && a.n_targets > 0 && final_status == FUSED_OK;
    if (tid < a.n_pcts) {
        const unsigned long long ka = *reinterpret_cast<volatile unsigned long long*>(&P.key[sp.pct_lo_idx[tid]]);
        const unsigned long long kb = *reinterpret_cast<volatile unsigned long long*>(&P.key[sp.pct_hi_idx[tid]]);
        const double va = ValueOf<Key>::get((Key)ka), vb = ValueOf<Key>::get((Key)kb);
        const double pc = numpy_lerp(va, vb, sp.pct_gamma[tid]);
        s_pct[tid] = pc;
        a.res->pct[tid] = pc; a.res->pct_lo[tid] = va; a.res->pct_hi[tid] = vb;
        if (want_tail && tid == a.tail_idx) { s_red[0] = vb; s_redc[0] = ka == kb ? 1ull : 0ull; }
    }
    __syncthreads();
    if (tid == 0) {
        unsigned long long n = 0, above = 0, below_cnt = 0;
        double sd = 0.0, sd2 = 0.0, mn = INFINITY, mx = -INFINITY, below_sum = 0.0;
        for (unsigned int r = 0; r < R; ++r) {              // rank order: every rank adds the same numbers the same way
            const unsigned long long* h = hdr_in(a, me, r);
            n += __ldcg(h + HW_N); above += __ldcg(h + HW_ABOVE);
            sd += dbl(__ldcg(h + HW_SUM_D)); sd2 += dbl(__ldcg(h + HW_SUM_D2));
            mn = fmin(mn, dbl(__ldcg(h + HW_MIN))); mx = fmax(mx, dbl(__ldcg(h + HW_MAX)));
            below_cnt += __ldcg(h + HW_TAIL_CNT); below_sum += dbl(__ldcg(h + HW_TAIL_SUM));
        }
        int st = final_status;
        if (st == FUSED_OK && (n != sp.n_global || (want_tail && below_cnt != P.tail_below))) st = FUSED_INCONSISTENT;
        const double inv = n ? 1.0 / (double)n : 0.0;
        const double md = sd * inv;
        double var = sd2 * inv - md * md;
        if (var < 0.0) var = 0.0;
        // values <= threshold: everything below the tail bin, the tail bin's candidates up to the lower order
        // statistic, and -- when the threshold reaches it -- every occurrence of the upper one
        const double thr = want_tail ? s_pct[a.tail_idx] : -INFINITY;
        unsigned long long tcnt = 0;
        double tsum = 0.0;
        if (want_tail) {
            tcnt = below_cnt + *reinterpret_cast<volatile unsigned long long*>(&P.tail_lo_cnt);
            tsum = below_sum + *reinterpret_cast<volatile double*>(&P.tail_lo_sum);
            const double v_hi = s_red[0];
            if (s_redc[0] == 0ull && v_hi <= thr) {         // distinct keys and the lerp landed on the upper one
                const unsigned long long m = *reinterpret_cast<volatile unsigned long long*>(&P.tail_hi_mult);
                tcnt += m;
                tsum += (double)m * v_hi;
            }
        }
        DeviceResult& res = *a.res;
        res.n = n; res.n_above = above; res.mean = a.center + md; res.std = sqrt(var); res.vmin = mn; res.vmax = mx;
        res.center = a.center;
        res.tail_threshold = thr;
        res.tail_count = tcnt;
        res.tail_sum = tsum;
        res.hist_lo = a.hist_lo; res.hist_hi = a.hist_hi;
        res.status = (unsigned long long)st;
    }
    __syncthreads();
    unsigned long long* host_words = reinterpret_cast<unsigned long long*>(a.host_out);
    const unsigned long long* res_words = reinterpret_cast<const unsigned long long*>(a.res);
    constexpr int RW = (int)(sizeof(DeviceResult) / 8);
    for (int i = tid; i < RW; i += SEL_THREADS) host_words[i] = res_words[i];
    for (unsigned int i = tid; i < a.nph_bins; i += SEL_THREADS) {
        unsigned long long c = 0;
        for (unsigned int r = 0; r < R; ++r) c += __ldcg(nph_in(a, me, r) + i);
        a.hist_total[i] = c;
        host_words[RW + i] = c;
        nph_in(a, me, me)[i] = 0u;          // this rank's accumulators start the next run at zero
    }
    unsigned int* kh = hist_in(a, me, me);
    for (int i = tid; i < FUSED_KB; i += SEL_THREADS) kh[i] = 0u;
    if (tid == 0) *a.claim = 0ull;
}

cudaError_t launch_fused_select(const FusedArgs& a, cudaStream_t s) {
    const size_t ksz = a.dtype == MC_F64 ? 8 : 4;
    const size_t smem = (size_t)SUB_BINS * 4 + (size_t)SEL_SORT * ksz + SEL_CAND_SMEM;
    // one block per possible slot; with no targets a single block writes the result
    const unsigned int blocks = a.n_targets ? (unsigned int)(a.n_targets < FUSED_MAX_SLOTS ? a.n_targets : FUSED_MAX_SLOTS) : 1u;
    cudaError_t e;
    if (a.dtype == MC_F64) {
        if ((e = ensure_dynamic_smem(fused_select_kernel<uint64_t>, smem)) != cudaSuccess) return e;
        fused_select_kernel<uint64_t><<<blocks, SEL_THREADS, smem, s>>>(a);
    } else {
        if ((e = ensure_dynamic_smem(fused_select_kernel<uint32_t>, smem)) != cudaSuccess) return e;
        fused_select_kernel<uint32_t><<<blocks, SEL_THREADS, smem, s>>>(a);
    }
    return cudaGetLastError();
}

// ------------------------------------------------------------------ BlockPartial from stored terminal values
template <typename Real>
__global__ void __launch_bounds__(256) terminal_partials_kernel(const Real* __restrict__ x, unsigned long long n, double center,
                                                                BlockPartial* __restrict__ out) {
    __shared__ BlockPartial s_w[8];
    BlockPartial acc;
    acc.sum_d = 0.0; acc.sum_d2 = 0.0; acc.vmin = INFINITY; acc.vmax = -INFINITY; acc.n = 0; acc.n_above = 0;
    unsigned long long per = (n + gridDim.x - 1) / gridDim.x;
    const unsigned long long beg = (unsigned long long)blockIdx.x * per < n ? (unsigned long long)blockIdx.x * per : n;
    const unsigned long long end = beg + per < n ? beg + per : n;
    for (unsigned long long i = beg + threadIdx.x; i < end; i += 256) {
        const double v = (double)x[i], d = v - center;
        acc.sum_d += d; acc.sum_d2 = fma(d, d, acc.sum_d2);
        acc.vmin = fmin(acc.vmin, v); acc.vmax = fmax(acc.vmax, v);
        acc.n += 1; acc.n_above += v > center ? 1ull : 0ull;
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        acc.sum_d += __shfl_down_sync(0xffffffffu, acc.sum_d, d);
        acc.sum_d2 += __shfl_down_sync(0xffffffffu, acc.sum_d2, d);
        acc.vmin = fmin(acc.vmin, __shfl_down_sync(0xffffffffu, acc.vmin, d));
        acc.vmax = fmax(acc.vmax, __shfl_down_sync(0xffffffffu, acc.vmax, d));
        acc.n += __shfl_down_sync(0xffffffffu, acc.n, d);
        acc.n_above += __shfl_down_sync(0xffffffffu, acc.n_above, d);
    }
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) {
            acc.sum_d += s_w[w].sum_d; acc.sum_d2 += s_w[w].sum_d2;
            acc.vmin = fmin(acc.vmin, s_w[w].vmin); acc.vmax = fmax(acc.vmax, s_w[w].vmax);
            acc.n += s_w[w].n; acc.n_above += s_w[w].n_above;
        }
        out[blockIdx.x] = acc;
    }
}

cudaError_t launch_terminal_partials(const void* x, int dtype, unsigned long long n, double center, BlockPartial* out,
                                     int sm_count, int* n_partials, cudaStream_t s) {
    const unsigned int blocks = pass_blocks(n, sm_count, 4096);
    *n_partials = (int)blocks;
    if (dtype == MC_F64) terminal_partials_kernel<double><<<blocks, 256, 0, s>>>(static_cast<const double*>(x), n, center, out);
    else terminal_partials_kernel<float><<<blocks, 256, 0, s>>>(static_cast<const float*>(x), n, center, out);
    return cudaGetLastError();
}

}  // namespace mc
