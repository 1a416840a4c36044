// mc_sim32.cuh -- the fp32 path-stepping kernel with shocks generated on the device (included by
// mc_sim.cu after the block-reduction helpers).  This is the kernel every BASELINE config runs:
//   NORMAL + CONST,  no store : fused GBM (configs 2, 4) -- log-price = sigma * sum of the normals
//   NORMAL + CONST,  store    : materialised GBM paths (config 5), four paths per thread, st.global.v4
//   BOOTSTRAP + GARCH         : the reference's own model (app.py:618-633)
// Replaces app.py:618-633, which streams three [N] float64 arrays through memory per step.
//
// Shared memory holds the one table the shock source gathers from:
//   NORMAL    : 4096 x (cos, sin) of the Box-Muller directions (32 KB, mc_rng.cuh)
//   BOOTSTRAP : the pool as (z, alpha z^2 + beta) pairs, replicated 2^lc times so that the lanes of a
//               half-warp (one LDS.64 wavefront) read different banks whatever indices they drew.
//               With g = alpha z^2 + beta precomputed, the reference's step (app.py:623-633)
//                   r = z sqrt(v);  y += r;  v = omega + alpha r^2 + beta v
//               is   y = fma(z, sqrt(v), y);  v = fma(g, v, omega)   -- one MUFU and two FFMA.
#pragma once

namespace mc {

struct Sim32Args {
    unsigned long long n_paths, path_offset, ld_paths;
    uint32_t n_steps, pool_len;
    int table_lc;                  // BOOTSTRAP: log2(copies) of the shared table
    float idx_scale, idx_bias;     // BOOTSTRAP: H 2^-23 and 2^23 - H + 1 (mc_rng.cuh: boot_index_bits)
    RoundKeys rk;
    const float2* lut;             // NORMAL: [DIR_COUNT] (cos, sin), built once per handle
    const float* pool;             // BOOTSTRAP: [pool_len]
    float* paths;
    float* terminal;
    float s0, var0, omega, alpha, beta, drift, sig;   // kernel units (log2)
    double center;
    BlockPartial* partials;
    SinkArgs sink;                 // SINK kernels only (mc_fused.cuh)
};

__device__ __forceinline__ float price_of(const Sim32Args& a, uint32_t steps_done, float y) {
    return a.s0 * fast_ex2(fmaf((float)steps_done, a.drift, y));                 // app.py:629 in log2 space
}

// One Philox call's worth of steps for ONE path: K = 8 (normal) or 5 (bootstrap).  `limit` < K only in
// the ragged last call (TAIL).  emit(k, y) sees the log-price after step k of this call.
enum { TAB_GLOBAL = 0, TAB_SMEM = 1, TAB_SMEM8 = 2 };     // where the bootstrap table lives / how it is addressed

template <int SHOCK, int VAR, int TAB, bool TAIL, typename Emit>
__device__ __forceinline__ void advance(const Sim32Args& a, const u32x4& w, uint32_t table_base, int table_shift,
                                        float sigr, int limit, float& y, float& v, Emit emit) {
    if (SHOCK == MC_SHOCK_NORMAL || SHOCK == SHOCK_NORMAL_FULL) {
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        constexpr int PAIRS = SHOCK == MC_SHOCK_NORMAL ? 4 : 2;
#pragma unroll
        for (int k = 0; k < PAIRS; ++k) {
            if (TAIL && 2 * k >= limit) break;
            float c, s, rho;
            if (SHOCK == MC_SHOCK_NORMAL) { direction_of(ww[k], table_base, c, s); rho = rho20(ww[k]); }
            else pair_full(ww[2 * k], ww[2 * k + 1], rho, c, s);
            const float rs = rho * sigr;                   // sigr = sqrt(2 ln 2) [* sigma]
            if (VAR == MC_VAR_CONST) {
                y = fmaf(rs, c, y);
                emit(2 * k, y);
                if (!TAIL || 2 * k + 1 < limit) { y = fmaf(rs, s, y); emit(2 * k + 1, y); }
            } else {
                float r = rs * c * fast_sqrt(v);                                  // app.py:623-626
                y += r;
                v = fmaf(a.alpha, r * r, fmaf(a.beta, v, a.omega));               // app.py:632-633
                emit(2 * k, y);
                if (!TAIL || 2 * k + 1 < limit) {
                    r = rs * s * fast_sqrt(v);
                    y += r;
                    v = fmaf(a.alpha, r * r, fmaf(a.beta, v, a.omega));
                    emit(2 * k + 1, y);
                }
            }
        }
    } else {
        const float gf[5] = {boot_g(w.x), boot_g(w.y), boot_g(w.z), boot_g(w.w), boot_g4(w)};
        uint32_t tbs[5];                           // bits(2^23 + idx + 1): four of the five floors by two packed FMAs
        boot_index_bits2(gf[0], gf[1], a.idx_scale, a.idx_bias, tbs[0], tbs[1]);
        boot_index_bits2(gf[2], gf[3], a.idx_scale, a.idx_bias, tbs[2], tbs[3]);
        tbs[4] = boot_index_bits(gf[4], a.idx_scale, a.idx_bias);
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            if (TAIL && k >= limit) break;
            const uint32_t tb = tbs[k];
            float z, g;
            if (TAB == TAB_SMEM8) {          // eight copies: the shift is an immediate and folds into one LEA
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(z), "=f"(g) : "r"((tb << 6) + table_base));
            } else if (TAB == TAB_SMEM) {
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(z), "=f"(g) : "r"((tb << table_shift) + table_base));
            } else {
                z = __ldg(a.pool + ((tb & 0x007FFFFFu) - 1u));
                g = fmaf(a.alpha, z * z, a.beta);
            }
            if (VAR == MC_VAR_CONST) {
                y = fmaf(a.sig, z, y);
            } else {
                y = fmaf(z, fast_sqrt(v), y);                                     // app.py:623-629
                v = fmaf(g, v, a.omega);                                          // app.py:632-633
            }
            emit(k, y);
        }
    }
}

// V paths per thread: 1, or 4 consecutive paths whose prices leave as one 16-byte store per step
// (a warp then writes 512 contiguous bytes of a time-major row: measured 6.0 TB/s against 4.4 TB/s
// for 4-byte stores at the config-5 shape, profiles/r01/microbench2_sm100a_v3.txt).
// SINK (runs without a fan chart, mc_fused.cuh): the block is persistent, its warps claim paths from a global
// counter (warps of a block do not advance at the same pace -- the issue arbiter is not fair -- so a static
// split would leave SMs idle at the end), and every terminal value is binned on the spot into a key histogram
// and, optionally, the caller's fixed-bin histogram, both in shared memory: one shared atomic per PATH each,
// flushed once per block.  Moments are then taken by the closing pass, not here.
template <int SHOCK, int VAR, bool STORE, int V, int TAB, bool SINK>
__global__ void __launch_bounds__(SHOCK != MC_SHOCK_BOOTSTRAP ? 512 : 1024) sim32_kernel(const __grid_constant__ Sim32Args a) {
    static_assert(!SINK || (!STORE && V == 1), "the sink rides on the kernels that keep no paths");
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    __shared__ BlockPartial s_warp[SINK ? 1 : 32];
    constexpr int K = SHOCK == MC_SHOCK_NORMAL ? 8 : (SHOCK == SHOCK_NORMAL_FULL ? 4 : 5);
    constexpr bool FUSED_SUM = SHOCK == MC_SHOCK_NORMAL && VAR == MC_VAR_CONST && !STORE;

    uint32_t table_base = (uint32_t)__cvta_generic_to_shared(smem_dyn);
    int table_shift = 0;
    float2* s_tab = reinterpret_cast<float2*>(smem_dyn);
    if (SHOCK == MC_SHOCK_NORMAL) {
        for (int i = threadIdx.x; i < DIR_COUNT; i += blockDim.x) s_tab[i] = a.lut[i];
    } else if (SHOCK == SHOCK_NORMAL_FULL) {
        // no table: directions through MUFU
    } else if (TAB != TAB_GLOBAL) {
        const int lc = TAB == TAB_SMEM8 ? 3 : a.table_lc;
        for (uint32_t i = threadIdx.x; i < (a.pool_len << lc); i += blockDim.x) {
            const float z = a.pool[i >> lc];
            s_tab[i] = make_float2(z, fmaf(a.alpha, z * z, a.beta));
        }
        // entry idx of copy c sits at ((idx << lc) + c) * 8; (bits(t) << (lc+3)) = wrap + ((idx + 1) << (lc+3))
        table_shift = lc + 3;
        const uint32_t wrap = (uint32_t)(((unsigned long long)0x4B000000u << table_shift) & 0xFFFFFFFFull);
        table_base += (threadIdx.x & ((1u << lc) - 1u)) * 8u - (1u << table_shift) - wrap;
    }
    unsigned int* s_kh = nullptr;          // SINK: key histogram [FUSED_KB], fixed-bin counts [nph], edges [nph + 1]
    unsigned int* s_nph = nullptr;
    float* s_edge = nullptr;
    int nph = 0;
    if constexpr (SINK) {
        s_kh = reinterpret_cast<unsigned int*>(smem_dyn + a.sink.smem_off);
        for (int i = threadIdx.x; i < FUSED_KB; i += blockDim.x) s_kh[i] = 0u;
        nph = a.sink.g_nphist ? (int)a.sink.nph_bins : 0;
        if (nph) {
            s_nph = s_kh + FUSED_KB;
            s_edge = reinterpret_cast<float*>(s_nph + nph);
            for (int i = threadIdx.x; i < nph; i += blockDim.x) s_nph[i] = 0u;
            for (int i = threadIdx.x; i <= nph; i += blockDim.x) s_edge[i] = a.sink.nph_edges[i];
        }
    }
    __syncthreads();
    float h_lo = 0.0f, h_top = 0.0f, h_inv = 0.0f;
    if (SINK && nph) { h_lo = s_edge[0]; h_top = s_edge[nph]; h_inv = (float)nph / (h_top - h_lo); }

    BlockPartial acc = empty_partial();
    const float sigr = (VAR == MC_VAR_CONST && !FUSED_SUM) ? a.sig * BM_RADIUS_SCALE_F
                                                          : (FUSED_SUM ? 1.0f : BM_RADIUS_SCALE_F);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * V;
    const uint32_t n_full = a.n_steps / K;
    const int n_tail = (int)(a.n_steps - n_full * K);

    auto run_paths = [&](const unsigned long long i0) {       // the V paths i0 .. i0 + V - 1, all steps
        const bool full = V == 1 || i0 + V <= a.n_paths;          // all V paths of this thread exist
        float y[V], v[V];
#pragma unroll
        for (int p = 0; p < V; ++p) { y[p] = 0.0f; v[p] = a.var0; }

        auto store_row = [&](uint32_t row, const float (&px)[V]) {
            float* dst = a.paths + (unsigned long long)row * a.ld_paths + i0;
            if (V == 4 && full) {
                *reinterpret_cast<float4*>(dst) = make_float4(px[0], px[1], px[2], px[3]);
            } else {
#pragma unroll
                for (int p = 0; p < V; ++p)
                    if (i0 + p < a.n_paths) dst[p] = px[p];
            }
        };
        if (STORE) {
            float px[V];
#pragma unroll
            for (int p = 0; p < V; ++p) px[p] = a.s0;
            store_row(0u, px);                                     // app.py:613
        }

        uint32_t t = 0;
        if (FUSED_SUM) {
            // constant sigma, nothing stored: y_T = sigma sqrt(2 ln 2) * sum(rho * dir); one FFMA per normal
            const unsigned long long gid = a.path_offset + i0;
            const uint32_t c0 = (uint32_t)gid, c1 = (uint32_t)(gid >> 32);
            unsigned long long s2 = 0ull;       // (sum of the cosine terms, sum of the sine terms)
#pragma unroll 2
            for (uint32_t j = 0; j < n_full; ++j) s2 = normals8_sum(philox4x32_10(c0, c1, j, 0u, a.rk), table_base, s2);
            float s_cos, s_sin;
            asm("mov.b64 {%0, %1}, %2;" : "=f"(s_cos), "=f"(s_sin) : "l"(s2));
            y[0] = a.sig * BM_RADIUS_SCALE_F * (s_cos + s_sin);
            t = n_full * K;
        } else {
            // two calls per trip when nothing is stored: the round keys and multipliers (uniform-register loads) are
            // then fetched once per ten steps of the reference's model instead of once per five
#pragma unroll (STORE ? 1 : 2)
            for (uint32_t j = 0; j < n_full; ++j) {
                float px[STORE ? K : 1][V];
#pragma unroll
                for (int p = 0; p < V; ++p) {
                    const unsigned long long gid = a.path_offset + i0 + p;
                    const u32x4 w = philox4x32_10((uint32_t)gid, (uint32_t)(gid >> 32), j, 0u, a.rk);
                    advance<SHOCK, VAR, TAB, false>(a, w, table_base, table_shift, sigr, K, y[p], v[p],
                                                    [&](int k, float yk) { if constexpr (STORE) px[k][p] = price_of(a, t + k + 1, yk); });
                }
                if (STORE) {
#pragma unroll
                    for (int k = 0; k < K; ++k) store_row(t + k + 1, px[k]);
                }
                t += K;
            }
        }
        if (n_tail) {      // ragged last call: n_steps not a multiple of K
            float px[STORE ? K : 1][V];
            const float sigr_tail = (VAR == MC_VAR_CONST) ? a.sig * BM_RADIUS_SCALE_F : BM_RADIUS_SCALE_F;
#pragma unroll
            for (int p = 0; p < V; ++p) {
                const unsigned long long gid = a.path_offset + i0 + p;
                const u32x4 w = philox4x32_10((uint32_t)gid, (uint32_t)(gid >> 32), n_full, 0u, a.rk);
                advance<SHOCK, VAR, TAB, true>(a, w, table_base, table_shift, sigr_tail, n_tail, y[p], v[p],
                                               [&](int k, float yk) { if constexpr (STORE) px[k][p] = price_of(a, t + k + 1, yk); });
            }
            if (STORE) {
#pragma unroll
                for (int k = 0; k < K - 1; ++k)
                    if (k < n_tail) store_row(t + k + 1, px[k]);
            }
        }
        float fin[V];
#pragma unroll
        for (int p = 0; p < V; ++p) fin[p] = price_of(a, a.n_steps, y[p]);
        if (V == 4 && full) {
            *reinterpret_cast<float4*>(a.terminal + i0) = make_float4(fin[0], fin[1], fin[2], fin[3]);   // app.py:636
        } else {
#pragma unroll
            for (int p = 0; p < V; ++p)
                if (i0 + p < a.n_paths) a.terminal[i0 + p] = fin[p];
        }
        if constexpr (SINK) {
            atomicAdd(&s_kh[window_bin(a.sink.win, key_of(fin[0]))], 1u);
            if (nph) {
                const int b = np_hist_bin<float>(fin[0], s_edge, nph, h_lo, h_top, h_inv);
                if (b >= 0) atomicAdd(&s_nph[b], 1u);
            }
        } else {
#pragma unroll
            for (int p = 0; p < V; ++p)
                if (i0 + p < a.n_paths) account<float>(acc, fin[p], a.center);
        }
    };
    if constexpr (SINK) {
        const unsigned int lane = threadIdx.x & 31u;
        // the claim of the NEXT batch is in flight while this one is simulated: the atomic's round trip is never waited for
        unsigned long long next = 0ull;
        if (lane == 0u) next = atomicAdd(a.sink.claim, (unsigned long long)a.sink.claim_paths);
        for (;;) {
            const unsigned long long base = __shfl_sync(0xffffffffu, next, 0);
            if (base >= a.n_paths) break;
            if (lane == 0u) next = atomicAdd(a.sink.claim, (unsigned long long)a.sink.claim_paths);
            const unsigned long long lim = base + a.sink.claim_paths < a.n_paths ? base + a.sink.claim_paths : a.n_paths;
            for (unsigned long long i0 = base + lane; i0 < lim; i0 += 32ull) run_paths(i0);
        }
        __syncthreads();
        for (int i = threadIdx.x; i < FUSED_KB; i += blockDim.x) {
            const unsigned int c = s_kh[i];
            if (c) atomicAdd(&a.sink.g_khist[i], c);
        }
        for (int i = threadIdx.x; i < nph; i += blockDim.x) {
            const unsigned int c = s_nph[i];
            if (c) atomicAdd(&a.sink.g_nphist[i], c);
        }
    } else {
        for (unsigned long long i0 = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * V; i0 < a.n_paths; i0 += stride)
            run_paths(i0);
        finish_block(acc, s_warp, a.partials);
    }
}

// (cos, sin) of the 4096 Box-Muller directions, rounded once from fp64
__global__ void build_direction_table_kernel(float2* lut) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= DIR_COUNT) return;
    double s, c;
    sincospi(2.0 * ((double)i / DIR_COUNT - 0.5), &s, &c);
    lut[i] = make_float2((float)c, (float)s);
}

}  // namespace mc
