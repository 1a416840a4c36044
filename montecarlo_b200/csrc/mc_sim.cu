// mc_sim.cu -- path-stepping kernels for sm_100a.
//
// Replaces the reference's time loop (app.py:618-633), which streams three
// [N] float64 arrays through memory per step (shocks, paths[t], variances[t]).
// Here one thread owns one path: its log-price and variance live in registers
// for all T steps, shocks are generated in registers (mc_rng.cuh), and the only
// HBM traffic is the terminal value (4 B per PATH, not per step) unless the
// caller asks for the full time-major paths matrix (app.py:612 layout), in which
// case each step adds one coalesced store per path.
//
// The terminal reductions of app.py:636-656 start here too: each thread
// accumulates moments about s0 in fp64 (one partial per block, reduced in fixed order).
//
// Two kernel families: sim32_kernel (fp32, shocks generated on the device: the fast path of
// every BASELINE config) and the generic sim_kernel (fp64, and injected shocks in either
// precision: the parity path).
//
// Units: the fp32 kernels accumulate the log-price in log2 units (the host
// pre-multiplies drift, omega and var0 by log2(e) or its square; GARCH is
// homogeneous so the recursion is unchanged) so that price = s0 * ex2(y) is a
// single MUFU.EX2.  The fp64 kernels work in natural units with exp().
#include <atomic>

#include "mc_launch.cuh"
#include "mc_rng.cuh"
#include "mc_sim.cuh"

namespace mc {

template <typename Real>
struct SimArgs {
    unsigned long long n_paths, path_offset, ld_paths, ld_shocks;
    uint32_t n_steps, pool_len;
    int pool_log2_copies;          // -1: pool stays in global memory
    RoundKeys rk;
    const Real* pool;
    const Real* shocks;
    Real* paths;
    Real* terminal;
    Real s0, var0, omega, alpha, beta, drift;   // kernel units
    double center;
    BlockPartial* partials;
};

__device__ __forceinline__ float exp_units(float y) { return fast_ex2(y); }
__device__ __forceinline__ double exp_units(double y) { return exp(y); }
__device__ __forceinline__ float sqrt_units(float v) { return fast_sqrt(v); }
__device__ __forceinline__ double sqrt_units(double v) { return sqrt(v); }
__device__ __forceinline__ float fma_units(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ double fma_units(double a, double b, double c) { return fma(a, b, c); }

// ------------------------------------------------------------------ block reduction
__device__ __forceinline__ double shfl_down_f64(double v, int d) { return __shfl_down_sync(0xffffffffu, v, d); }
__device__ __forceinline__ unsigned long long shfl_down_u64(unsigned long long v, int d) {
    return __shfl_down_sync(0xffffffffu, v, d);
}

__device__ __forceinline__ void combine(BlockPartial& a, const BlockPartial& b) {
    a.sum_d += b.sum_d;
    a.sum_d2 += b.sum_d2;
    a.vmin = fmin(a.vmin, b.vmin);
    a.vmax = fmax(a.vmax, b.vmax);
    a.n += b.n;
    a.n_above += b.n_above;
}

__device__ __forceinline__ BlockPartial warp_reduce(BlockPartial p) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
        BlockPartial o;
        o.sum_d = shfl_down_f64(p.sum_d, d);
        o.sum_d2 = shfl_down_f64(p.sum_d2, d);
        o.vmin = shfl_down_f64(p.vmin, d);
        o.vmax = shfl_down_f64(p.vmax, d);
        o.n = shfl_down_u64(p.n, d);
        o.n_above = shfl_down_u64(p.n_above, d);
        combine(p, o);
    }
    return p;
}

// all threads of the block must call; result valid in thread 0
__device__ __forceinline__ BlockPartial block_reduce(BlockPartial p, BlockPartial* s_warp /*[32]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int n_warps = (blockDim.x + 31) >> 5;
    p = warp_reduce(p);
    if (lane == 0) s_warp[warp] = p;
    __syncthreads();
    if (warp == 0) {
        BlockPartial q;
        if (lane < n_warps) q = s_warp[lane];
        else { q.sum_d = 0; q.sum_d2 = 0; q.vmin = INFINITY; q.vmax = -INFINITY; q.n = 0; q.n_above = 0; }
        p = warp_reduce(q);
    }
    return p;
}

__device__ __forceinline__ BlockPartial empty_partial() {
    BlockPartial p;
    p.sum_d = 0.0; p.sum_d2 = 0.0; p.vmin = INFINITY; p.vmax = -INFINITY; p.n = 0; p.n_above = 0;
    return p;
}

template <typename Real>
__device__ __forceinline__ void account(BlockPartial& acc, Real value, double center) {
    const double d = (double)value - center;
    acc.sum_d += d;
    acc.sum_d2 = fma(d, d, acc.sum_d2);
    acc.vmin = fmin(acc.vmin, (double)value);
    acc.vmax = fmax(acc.vmax, (double)value);
    acc.n += 1;
    acc.n_above += ((double)value > center) ? 1ull : 0ull;       // app.py:654 (compared in fp64)
}

__device__ __forceinline__ void finish_block(BlockPartial acc, BlockPartial* s_warp, BlockPartial* partials) {
    acc = block_reduce(acc, s_warp);
    if (threadIdx.x == 0) partials[blockIdx.x] = acc;
}

// ------------------------------------------------------------------ shock sources (generic kernel)
// The generic kernel draws its own shocks only in fp64: 4 normals or 4 bootstrap indices per Philox
// call (one full 32-bit word each).  Injected shocks (either precision) are read in groups of 4.
// Step t of a path uses shock (t mod 4) of call (t div 4).  fp32 device shocks: sim32_kernel below.
template <typename Real, int SHOCK>
struct ShocksPerCall { static constexpr int value = 4; };

template <typename Real, int SHOCK>
__device__ __forceinline__ void draw(const SimArgs<Real>& a, const Real* s_pool, unsigned long long i,
                                     uint32_t c0, uint32_t c1, uint32_t call, uint32_t t, Real (&z)[8]) {
#pragma unroll
    for (int k = 4; k < 8; ++k) z[k] = Real(0);
    if (SHOCK == MC_SHOCK_INJECTED) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
            z[k] = (t + k < a.n_steps) ? a.shocks[(unsigned long long)(t + k) * a.ld_shocks + i] : Real(0);
        return;
    }
    const u32x4 w = philox4x32_10(c0, c1, call, 0u, a.rk);
    if (SHOCK == MC_SHOCK_NORMAL) {
        draw_normals(w, 0u, z);
    } else {
        const uint32_t u[4] = {w.x, w.y, w.z, w.w};
        if (a.pool_log2_copies >= 0) {
            // replicated pool: lane l reads copy (l mod copies), so with 32 copies every
            // lane stays in its own bank no matter which index it drew
            const uint32_t copies_mask = (1u << a.pool_log2_copies) - 1u;
            const uint32_t sub = threadIdx.x & copies_mask;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                z[k] = s_pool[(bootstrap_index(u[k], a.pool_len) << a.pool_log2_copies) + sub];
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) z[k] = __ldg(&a.pool[bootstrap_index(u[k], a.pool_len)]);
        }
    }
}

// one step of app.py:623-633 in log space; y excludes the deterministic drift
template <typename Real, int VAR>
__device__ __forceinline__ void step(const SimArgs<Real>& a, Real z, Real sig, Real& y, Real& v) {
    if (VAR == MC_VAR_CONST) {
        y = fma_units(sig, z, y);
    } else {
        const Real r = z * sqrt_units(v);                         // app.py:623-626
        y += r;                                                   // app.py:629
        v = fma_units(a.alpha, r * r, fma_units(a.beta, v, a.omega));   // app.py:632-633
    }
}

template <typename Real, int SHOCK, int VAR, bool STORE, int MAXB>
__global__ void __launch_bounds__(MAXB) sim_kernel(const __grid_constant__ SimArgs<Real> a) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    __shared__ BlockPartial s_warp[32];
    Real* s_pool = reinterpret_cast<Real*>(smem_dyn);
    constexpr int K = ShocksPerCall<Real, SHOCK>::value;

    if (SHOCK == MC_SHOCK_BOOTSTRAP && a.pool_log2_copies >= 0) {
        const uint32_t total = a.pool_len << a.pool_log2_copies;
        for (uint32_t i = threadIdx.x; i < total; i += blockDim.x) s_pool[i] = a.pool[i >> a.pool_log2_copies];
    }
    __syncthreads();

    BlockPartial acc = empty_partial();
    const Real sig = (VAR == MC_VAR_CONST) ? sqrt_units(a.var0) : Real(0);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const uint32_t n_full = a.n_steps / K;

    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_paths;
         i += stride) {
        const unsigned long long gid = a.path_offset + i;
        const uint32_t c0 = (uint32_t)gid, c1 = (uint32_t)(gid >> 32);
        Real y = Real(0), v = a.var0;
        if (STORE) a.paths[i] = a.s0;                              // app.py:613

        uint32_t t = 0;
#pragma unroll 2
        for (uint32_t j = 0; j < n_full; ++j) {
            Real z[8];
            draw<Real, SHOCK>(a, s_pool, i, c0, c1, j, t, z);
#pragma unroll
            for (int k = 0; k < K; ++k) {
                step<Real, VAR>(a, z[k], sig, y, v);
                if (STORE)
                    a.paths[(unsigned long long)(t + k + 1) * a.ld_paths + i] =
                        a.s0 * exp_units(fma_units((Real)(t + k + 1), a.drift, y));
            }
            t += K;
        }
        if (t < a.n_steps) {   // ragged tail: n_steps not a multiple of K
            Real z[8];
            draw<Real, SHOCK>(a, s_pool, i, c0, c1, n_full, t, z);
#pragma unroll
            for (int k = 0; k < K - 1; ++k) {
                if (t + k < a.n_steps) {
                    step<Real, VAR>(a, z[k], sig, y, v);
                    if (STORE)
                        a.paths[(unsigned long long)(t + k + 1) * a.ld_paths + i] =
                            a.s0 * exp_units(fma_units((Real)(t + k + 1), a.drift, y));
                }
            }
        }
        const Real final_price = a.s0 * exp_units(fma_units((Real)a.n_steps, a.drift, y));
        a.terminal[i] = final_price;                               // app.py:636
        account<Real>(acc, final_price, a.center);
    }
    finish_block(acc, s_warp, a.partials);
}

// ------------------------------------------------------------------ many price histories, one launch
// the reference's model (bootstrap + GARCH, fp64) for n_series histories whose pools and
// scalars were produced on the device (mc_inputs.cu).  blockIdx.y = history; the per-history scalars are read
// from device memory, so nothing goes through the host between the producers and this kernel.  History b
// uses Philox key seed + b and the arithmetic of sim_kernel<double, BOOTSTRAP, GARCH> (same draw/step
// functions): its paths are bit-identical to a single call with that seed.
template <bool STORE>
__global__ void __launch_bounds__(BATCH_BLOCK) batch_sim_kernel(const BatchSimArgs b) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    double* s_pool = reinterpret_cast<double*>(smem_dyn);
    const uint32_t series = blockIdx.y;
    const mc_series_inputs in = b.inputs[series];
    bool live = in.pool_len >= b.min_history && in.flags == 0;
    double omega = b.omega, alpha = b.alpha, beta = b.beta;
    if (b.fit) {                        // calibrated per history (mc_calibrate_garch)
        const mc_garch_fit f = b.fit[series];
        live = live && f.valid != 0;
        omega = f.omega; alpha = f.alpha; beta = f.beta;
    }
    double* terminal = b.terminal + (size_t)series * b.ld_terminal;
    double* paths = STORE ? b.paths + (size_t)series * (b.n_steps + 1ull) * b.ld_paths : nullptr;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const unsigned long long first = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (!live) {       // the reference returns None (app.py:594-595): rows of zeros keep the reductions defined
        for (unsigned long long i = first; i < b.n_paths; i += stride) {
            terminal[i] = 0.0;
            if (STORE) for (uint32_t t = 0; t <= b.n_steps; ++t) paths[(unsigned long long)t * b.ld_paths + i] = 0.0;
        }
        return;
    }
    SimArgs<double> a;
    const unsigned long long key = b.seed + series;
    a.rk = expand_key((uint32_t)key, (uint32_t)(key >> 32));
    a.pool = b.pool + (size_t)series * b.pool_ld;
    a.pool_len = (uint32_t)in.pool_len;
    a.pool_log2_copies = b.pool_in_smem ? 0 : -1;
    a.n_steps = b.n_steps;
    a.omega = omega; a.alpha = alpha; a.beta = beta;
    if (b.pool_in_smem) {
        for (uint32_t i = threadIdx.x; i < a.pool_len; i += blockDim.x) s_pool[i] = a.pool[i];
        __syncthreads();
    }
    const double s0 = in.s0;
    const uint32_t n_full = b.n_steps / 4;
    for (unsigned long long i = first; i < b.n_paths; i += stride) {
        const uint32_t c0 = (uint32_t)i, c1 = (uint32_t)(i >> 32);
        double y = 0.0, v = in.var0;
        if (STORE) paths[i] = s0;                                   // app.py:613
        uint32_t t = 0;
        for (uint32_t j = 0; j <= n_full; ++j) {
            if (t >= b.n_steps) break;
            double z[8];
            draw<double, MC_SHOCK_BOOTSTRAP>(a, s_pool, i, c0, c1, j, t, z);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (t + k < b.n_steps) {
                    step<double, MC_VAR_GARCH>(a, z[k], 0.0, y, v);
                    if (STORE) paths[(unsigned long long)(t + k + 1) * b.ld_paths + i] = s0 * exp(y);
                }
            }
            t += 4;
        }
        terminal[i] = s0 * exp(y);                                  // app.py:636
    }
}

cudaError_t launch_batch_simulation(const BatchSimArgs& args, bool store, int sm_count, cudaStream_t stream) {
    if (args.n_series == 0 || args.n_paths == 0) return cudaSuccess;
    BatchSimArgs b = args;
    const size_t smem = (size_t)args.pool_ld * sizeof(double);
    b.pool_in_smem = smem <= 40 * 1024 ? 1 : 0;
    unsigned long long gx = (args.n_paths + BATCH_BLOCK - 1) / BATCH_BLOCK;
    const unsigned long long cap = (unsigned long long)sm_count * 32;      // beyond that, threads loop over paths
    if (gx * args.n_series > cap) gx = (cap + args.n_series - 1) / args.n_series;
    if (gx < 1) gx = 1;
    const dim3 grid((unsigned)gx, args.n_series);
    if (store) batch_sim_kernel<true><<<grid, BATCH_BLOCK, b.pool_in_smem ? smem : 0, stream>>>(b);
    else       batch_sim_kernel<false><<<grid, BATCH_BLOCK, b.pool_in_smem ? smem : 0, stream>>>(b);
    return cudaGetLastError();
}

}  // namespace mc
#include "mc_sim32.cuh"
namespace mc {

// ------------------------------------------------------------------ correlated portfolio
// Mode P (north_star extension; nothing like it in the reference): d assets driven by
// y = L z with L the Cholesky factor of the correlation matrix.  L (pre-scaled by
// sigma_i sqrt(dt)) arrives as a kernel PARAMETER, i.e. in the constant bank, so the
// 136 FFMAs of the d=16 triangular product read it as an instruction operand and the
// kernel keeps no mutable global state (concurrent handles cannot race).
template <typename Real, int D>
struct PortArgs {
    unsigned long long n_paths, path_offset, ld_paths, ld_shocks;
    uint32_t n_steps, calls_per_step, n_assets;
    RoundKeys rk;
    const Real* shocks;
    const float2* lut;                // fp32 device shocks: Box-Muller direction table (mc_rng.cuh)
    Real* paths;
    Real* terminal;
    Real v0;                          // portfolio value at t=0
    double center;
    BlockPartial* partials;
    // Cholesky factor (times sigma_i per step), kernel units.  fp64: packed lower triangle, row-major,
    // chol[r (r + 1) / 2 + c].  fp32: ROW PAIRS for the packed FFMA2 of sm_100 -- pair p = rows (2p, 2p + 1),
    // columns 0 .. 2p + 1, entry (p, c) = (L[2p][c], L[2p+1][c]) at chol[2 (p (p + 1) + c)] (L[2p][2p+1] = 0)
    Real chol[D * (D + 2) / 2];
    Real drift[D];                    // per step, kernel units
    Real s0w[D];                      // w_i * S0_i
};

template <typename Real, int D>
__device__ __forceinline__ Real portfolio_value(const PortArgs<Real, D>& a, const Real (&x)[D], Real steps_done) {
    Real v = Real(0);
#pragma unroll
    for (int i = 0; i < D; ++i) v = fma_units(a.s0w[i], exp_units(fma_units(steps_done, a.drift[i], x[i])), v);
    return v;
}

template <typename Real, int D, bool INJECTED, bool STORE>
__global__ void __launch_bounds__(SIM_BLOCK) portfolio_kernel(const __grid_constant__ PortArgs<Real, D> a) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    __shared__ BlockPartial s_warp[32];
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(smem_dyn);
    if (!INJECTED && sizeof(Real) == 4) {
        float2* s_lut = reinterpret_cast<float2*>(smem_dyn);
        for (int i = threadIdx.x; i < DIR_COUNT; i += blockDim.x) s_lut[i] = a.lut[i];
    }
    __syncthreads();

    BlockPartial acc = empty_partial();
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_paths;
         i += stride) {
        const unsigned long long gid = a.path_offset + i;
        const uint32_t c0 = (uint32_t)gid, c1 = (uint32_t)(gid >> 32);
        Real x[D];
#pragma unroll
        for (int k = 0; k < D; ++k) x[k] = Real(0);
        if (STORE) a.paths[i] = a.v0;
        uint32_t call = 0;
        for (uint32_t t = 0; t < a.n_steps; ++t) {
            Real z[D];
            if (INJECTED) {
#pragma unroll
                for (int k = 0; k < D; ++k) {
                    const unsigned long long row = (unsigned long long)t * a.n_assets + k;
                    z[k] = (k < (int)a.n_assets) ? a.shocks[row * a.ld_shocks + i] : Real(0);
                }
            } else {
                constexpr int NPC = NormalsPerCall<Real>::value;
                constexpr int QMAX = (D + NPC - 1) / NPC;
#pragma unroll
                for (int q = 0; q < QMAX; ++q) {
                    Real zq[8] = {Real(0), Real(0), Real(0), Real(0), Real(0), Real(0), Real(0), Real(0)};
                    if (q < (int)a.calls_per_step) {
                        const u32x4 w = philox4x32_10(c0, c1, call + q, 0u, a.rk);
                        draw_normals<false>(w, lut_base, zq);      // fp32: sqrt(2 ln 2) is folded into a.chol
                    }
#pragma unroll
                    for (int k = 0; k < NPC; ++k)
                        if (NPC * q + k < D) z[NPC * q + k] = zq[k];
                }
                call += a.calls_per_step;
            }
            if constexpr (sizeof(Real) == 4) {
                // y = L z, two rows per instruction: fma.rn.f32x2 (FFMA2) takes the pair of L from a uniform
                // register and z[c] as a broadcast operand; 72 of them replace 136 FFMA at D = 16, each row's
                // sum in the same order as before (the extra term of an even row is fma(0, z, x) = x)
#pragma unroll
                for (int pr = 0; pr < D / 2; ++pr) {
                    unsigned long long acc;
                    asm("mov.b64 %0, {%1, %2};" : "=l"(acc) : "f"(x[2 * pr]), "f"(x[2 * pr + 1]));
#pragma unroll
                    for (int c = 0; c <= 2 * pr + 1; ++c) {
                        unsigned long long l2, z2;
                        asm("mov.b64 %0, {%1, %2};" : "=l"(l2) : "f"(a.chol[2 * (pr * (pr + 1) + c)]), "f"(a.chol[2 * (pr * (pr + 1) + c) + 1]));
                        asm("mov.b64 %0, {%1, %1};" : "=l"(z2) : "f"(z[c]));
                        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(l2), "l"(z2));
                    }
                    asm("mov.b64 {%0, %1}, %2;" : "=f"(x[2 * pr]), "=f"(x[2 * pr + 1]) : "l"(acc));
                }
            } else {
#pragma unroll
                for (int r = 0; r < D; ++r) {
                    Real s = x[r];
#pragma unroll
                    for (int c = 0; c <= r; ++c) s = fma_units(a.chol[r * (r + 1) / 2 + c], z[c], s);
                    x[r] = s;
                }
            }
            if (STORE)
                a.paths[(unsigned long long)(t + 1) * a.ld_paths + i] = portfolio_value<Real, D>(a, x, (Real)(t + 1));
        }
        const Real v_final = portfolio_value<Real, D>(a, x, (Real)a.n_steps);
        a.terminal[i] = v_final;
        account<Real>(acc, v_final, a.center);
    }
    finish_block(acc, s_warp, a.partials);
}

// ------------------------------------------------------------------ debug / test kernels
template <typename Real, int SHOCK>
__global__ void debug_shocks_kernel(const SimArgs<Real> a, Real* out) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    constexpr int K = ShocksPerCall<Real, SHOCK>::value;
    const uint32_t n_calls = (a.n_steps + K - 1) / K;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_paths;
         i += stride) {
        const unsigned long long gid = a.path_offset + i;
        for (uint32_t j = 0; j < n_calls; ++j) {
            Real z[8];
            draw<Real, SHOCK>(a, nullptr, i, (uint32_t)gid, (uint32_t)(gid >> 32), j, K * j, z);
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (K * j + k < a.n_steps) out[(unsigned long long)(K * j + k) * a.n_paths + i] = z[k];
        }
    }
}

// the shocks sim32_kernel consumes: normals (8 per call) or pool[index] (5 per call)
template <int SHOCK>
__global__ void debug_shocks32_kernel(const Sim32Args a, float* out) {
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(smem_dyn);
    if (SHOCK == MC_SHOCK_NORMAL) {
        float2* s_lut = reinterpret_cast<float2*>(smem_dyn);
        for (int i = threadIdx.x; i < DIR_COUNT; i += blockDim.x) s_lut[i] = a.lut[i];
    }
    __syncthreads();
    constexpr int K = SHOCK == MC_SHOCK_NORMAL ? 8 : (SHOCK == SHOCK_NORMAL_FULL ? 4 : 5);
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    const uint32_t n_calls = (a.n_steps + K - 1) / K;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n_paths;
         i += stride) {
        const unsigned long long gid = a.path_offset + i;
        for (uint32_t j = 0; j < n_calls; ++j) {
            const u32x4 w = philox4x32_10((uint32_t)gid, (uint32_t)(gid >> 32), j, 0u, a.rk);
            float z[8];
            if (SHOCK == MC_SHOCK_NORMAL) {
                normals8(w, lut_base, z);
            } else if (SHOCK == SHOCK_NORMAL_FULL) {
                normals4_full(w, z);
            } else {
                const float gf[5] = {boot_g(w.x), boot_g(w.y), boot_g(w.z), boot_g(w.w), boot_g4(w)};
#pragma unroll
                for (int k = 0; k < 5; ++k)
                    z[k] = a.pool[(boot_index_bits(gf[k], a.idx_scale, a.idx_bias) & 0x007FFFFFu) - 1u];
            }
#pragma unroll
            for (int k = 0; k < K; ++k)
                if (K * j + k < a.n_steps) out[(unsigned long long)(K * j + k) * a.n_paths + i] = z[k];
        }
    }
}

__global__ void debug_philox_kernel(const uint32_t* ctr, uint32_t n, uint32_t k0, uint32_t k1, uint32_t* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32x4 w = philox4x32_10(ctr[4 * i], ctr[4 * i + 1], ctr[4 * i + 2], ctr[4 * i + 3], expand_key(k0, k1));
    out[4 * i] = w.x; out[4 * i + 1] = w.y; out[4 * i + 2] = w.z; out[4 * i + 3] = w.w;
}

// per-block partials -> stage-0 header; one block, fixed summation order => deterministic
constexpr int REDUCE_THREADS = 1024;
__global__ void __launch_bounds__(REDUCE_THREADS) reduce_partials_kernel(const BlockPartial* partials, int n,
                                                                         unsigned long long* words, unsigned long long w6,
                                                                         unsigned long long w7) {
    __shared__ BlockPartial s[REDUCE_THREADS];
    BlockPartial acc = empty_partial();
    int i = threadIdx.x;
    for (; i + 3 * REDUCE_THREADS < n; i += 4 * REDUCE_THREADS) {      // four loads in flight per thread
        const BlockPartial a = partials[i], b = partials[i + REDUCE_THREADS], c = partials[i + 2 * REDUCE_THREADS],
                           d = partials[i + 3 * REDUCE_THREADS];
        combine(acc, a); combine(acc, b); combine(acc, c); combine(acc, d);
    }
    for (; i < n; i += REDUCE_THREADS) combine(acc, partials[i]);
    s[threadIdx.x] = acc;
    __syncthreads();
    for (int w = REDUCE_THREADS / 2; w > 0; w >>= 1) {
        if ((int)threadIdx.x < w) combine(s[threadIdx.x], s[threadIdx.x + w]);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        words[0] = s[0].n;
        words[1] = s[0].n_above;
        words[2] = (unsigned long long)__double_as_longlong(s[0].sum_d);
        words[3] = (unsigned long long)__double_as_longlong(s[0].sum_d2);
        words[4] = (unsigned long long)__double_as_longlong(s[0].vmin);
        words[5] = (unsigned long long)__double_as_longlong(s[0].vmax);
        words[6] = w6;       // pilot samples this rank contributes per row: terminal / fan-chart problem
        words[7] = w7;
    }
}

// ------------------------------------------------------------------ host side
static const double LOG2E = 1.4426950408889634074;

template <typename Real> static double unit_scale() { return sizeof(Real) == 4 ? LOG2E : 1.0; }

template <typename Real>
static SimArgs<Real> make_args(const mc_params& p, const SimBuffers& b) {
    const double u = unit_scale<Real>();
    SimArgs<Real> a;
    a.n_paths = p.n_paths;
    a.path_offset = p.path_offset;
    a.ld_paths = p.paths_ld ? p.paths_ld : p.n_paths;
    a.ld_shocks = p.n_paths;
    a.n_steps = p.n_steps;
    a.rk = expand_key((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
    a.pool_len = p.pool_len;
    a.pool_log2_copies = -1;
    a.pool = static_cast<const Real*>(b.d_pool);
    a.shocks = static_cast<const Real*>(b.d_shocks);
    a.paths = static_cast<Real*>(b.d_paths);
    a.terminal = static_cast<Real*>(b.d_terminal);
    a.s0 = (Real)p.s0;
    a.var0 = (Real)(p.var0 * u * u);
    a.omega = (Real)(p.omega * u * u);
    a.alpha = (Real)p.alpha;
    a.beta = (Real)p.beta;
    a.drift = (Real)(p.drift * u);
    a.center = p.s0;
    a.partials = b.d_partials;
    return a;
}

static std::atomic<int> g_tune_waves{24}, g_tune_nblock{512};
void set_sim_tuning(int waves, int nblock) {
    g_tune_waves.store(waves >= 1 ? waves : 24);
    g_tune_nblock.store((nblock == 128 || nblock == 256 || nblock == 512) ? nblock : 512);
}

template <typename Kernel>
static cudaError_t pick_grid(Kernel kernel, int block, size_t smem, const DeviceInfo& dev,
                             unsigned long long n, int* grid) {
    cudaError_t e = ensure_dynamic_smem(kernel, smem);
    if (e != cudaSuccess) return e;
    int per_sm = 0;
    e = resident_blocks_per_sm(kernel, block, smem, &per_sm);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorLaunchOutOfResources;
    const unsigned long long resident = (unsigned long long)dev.sm_count * per_sm;
    unsigned long long need = (n + block - 1) / block;
    if (need < 1) need = 1;
    unsigned long long g = need;
    if (need > resident) {
        // More blocks than fit at once: let the hardware block scheduler balance the load.  Warps of a
        // resident block do not finish together (the issue arbiter is not fair), so a persistent grid
        // drains unevenly; ~24 waves of short blocks keep every SM full until the last few percent.
        const unsigned long long waves = (unsigned long long)g_tune_waves.load();
        unsigned long long per_thread = n / ((unsigned long long)block * resident * waves);
        if (per_thread < 1) per_thread = 1;
        g = (n + block * per_thread - 1) / (block * per_thread);
        if (g > MAX_SIM_GRID) g = MAX_SIM_GRID;
    }
    *grid = (int)g;
    return cudaSuccess;
}

template <typename Real, int SHOCK, int VAR, bool STORE>
static cudaError_t launch_single(const mc_params& p, const SimBuffers& b, const DeviceInfo& dev, int* grid,
                                 cudaStream_t stream) {
    constexpr int MAXB = sizeof(Real) == 4 ? 1024 : 256;
    SimArgs<Real> a = make_args<Real>(p, b);
    int block = SIM_BLOCK;
    size_t smem = 0;
    if (SHOCK == MC_SHOCK_BOOTSTRAP) {
        block = MAXB;
        // static smem: 1.5 KB warp partials
        const size_t budget = (MAXB == 1024) ? (size_t)dev.max_smem_optin - 4 * 1024 : 40 * 1024;
        int lg = 5;
        while (lg >= 0 && ((size_t)p.pool_len << lg) * sizeof(Real) > budget) --lg;
        a.pool_log2_copies = lg;               // -1 => gather from global/L1
        smem = lg >= 0 ? ((size_t)p.pool_len << lg) * sizeof(Real) : 0;
        if (p.n_paths < (unsigned long long)block * dev.sm_count) {
            // small request: spread over more SMs with narrower blocks and a narrower replication
            block = 128;
            int lg2 = 5;
            while (lg2 >= 0 && ((size_t)p.pool_len << lg2) * sizeof(Real) > 24 * 1024) --lg2;
            a.pool_log2_copies = lg2;
            smem = lg2 >= 0 ? ((size_t)p.pool_len << lg2) * sizeof(Real) : 0;
        }
    }
    auto kernel = sim_kernel<Real, SHOCK, VAR, STORE, MAXB>;
    cudaError_t e = pick_grid(kernel, block, smem, dev, p.n_paths, grid);
    if (e != cudaSuccess) return e;
    kernel<<<*grid, block, smem, stream>>>(a);
    return cudaGetLastError();
}

template <typename Real, int SHOCK, int VAR>
static cudaError_t dispatch_store(const mc_params& p, const SimBuffers& b, const DeviceInfo& dev, int* grid,
                                  cudaStream_t s) {
    return (p.flags & MC_FLAG_STORE_PATHS) ? launch_single<Real, SHOCK, VAR, true>(p, b, dev, grid, s)
                                           : launch_single<Real, SHOCK, VAR, false>(p, b, dev, grid, s);
}
template <typename Real, int SHOCK>
static cudaError_t dispatch_var(const mc_params& p, const SimBuffers& b, const DeviceInfo& dev, int* grid,
                                cudaStream_t s) {
    return p.var_model == MC_VAR_CONST ? dispatch_store<Real, SHOCK, MC_VAR_CONST>(p, b, dev, grid, s)
                                       : dispatch_store<Real, SHOCK, MC_VAR_GARCH>(p, b, dev, grid, s);
}
// generic kernel: fp64 with any shock source; fp32 only with injected shocks
template <typename Real>
static cudaError_t dispatch_shock(const mc_params& p, const SimBuffers& b, const DeviceInfo& dev, int* grid,
                                  cudaStream_t s) {
    if (p.shock == MC_SHOCK_INJECTED) return dispatch_var<Real, MC_SHOCK_INJECTED>(p, b, dev, grid, s);
    if constexpr (sizeof(Real) == 8) {
        if (p.shock == MC_SHOCK_BOOTSTRAP) return dispatch_var<Real, MC_SHOCK_BOOTSTRAP>(p, b, dev, grid, s);
        if (p.shock == MC_SHOCK_NORMAL) return dispatch_var<Real, MC_SHOCK_NORMAL>(p, b, dev, grid, s);
    }
    return cudaErrorInvalidValue;
}

// ---- fp32 with device shocks: sim32_kernel -------------------------------------------------
static bool boot32_eligible(const mc_params& p) { return p.pool_len >= 1 && p.pool_len <= BOOT32_MAX_POOL; }

static Sim32Args make_args32(const mc_params& p, const SimBuffers& b) {
    Sim32Args a{};
    a.n_paths = p.n_paths;
    a.path_offset = p.path_offset;
    a.ld_paths = p.paths_ld ? p.paths_ld : p.n_paths;
    a.n_steps = p.n_steps;
    a.pool_len = p.pool_len;
    a.table_lc = 0;
    a.idx_scale = (float)p.pool_len * 1.1920928955078125e-07f;     // H 2^-23: exact
    a.idx_bias = 8388608.0f - (float)p.pool_len + 1.0f;            // exact (an integer below 2^24)
    a.rk = expand_key((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
    a.lut = b.d_lut;
    a.pool = static_cast<const float*>(b.d_pool);
    a.paths = static_cast<float*>(b.d_paths);
    a.terminal = static_cast<float*>(b.d_terminal);
    a.s0 = (float)p.s0;
    a.var0 = (float)(p.var0 * LOG2E * LOG2E);
    a.omega = (float)(p.omega * LOG2E * LOG2E);
    a.alpha = (float)p.alpha;
    a.beta = (float)p.beta;
    a.drift = (float)(p.drift * LOG2E);
    a.sig = (float)(sqrt(p.var0) * LOG2E);
    a.center = p.s0;
    a.partials = b.d_partials;
    if (b.sink && !(p.flags & MC_FLAG_STORE_PATHS)) a.sink = *b.sink;
    return a;
}

template <int SHOCK, int VAR, bool STORE, int V, int TAB, bool SINK = false>
static cudaError_t launch32_t(Sim32Args& a, const DeviceInfo& dev, int* grid, cudaStream_t stream) {
    auto kernel = sim32_kernel<SHOCK, VAR, STORE, V, TAB, SINK>;
    const unsigned long long items = (a.n_paths + V - 1) / V;      // threads' worth of work
    int block;
    size_t smem;
    if (SHOCK == MC_SHOCK_NORMAL || SHOCK == SHOCK_NORMAL_FULL) {
        smem = SHOCK == MC_SHOCK_NORMAL ? (size_t)DIR_COUNT * sizeof(float2) : 0;
        block = items >= 512ull * dev.sm_count ? g_tune_nblock.load() : 128;
    } else {
        block = items >= 1024ull * dev.sm_count ? 1024 : 256;
        smem = TAB != TAB_GLOBAL ? ((size_t)a.pool_len << a.table_lc) * sizeof(float2) : 0;
    }
    if (SINK) {
        // persistent blocks: as many as stay resident, their warps claim paths as they go
        smem = (smem + 15) & ~(size_t)15;
        a.sink.smem_off = (unsigned int)smem;
        smem += sink_smem_bytes(a.sink.g_nphist ? a.sink.nph_bins : 0u);
        cudaError_t e = ensure_dynamic_smem(kernel, smem);
        if (e != cudaSuccess) return e;
        int per_sm = 0;
        if ((e = resident_blocks_per_sm(kernel, block, smem, &per_sm)) != cudaSuccess) return e;
        if (per_sm < 1) return cudaErrorLaunchOutOfResources;
        const unsigned long long resident = (unsigned long long)dev.sm_count * per_sm;
        a.sink.claim_paths = a.n_paths >= 16ull * resident * block ? 64u : 32u;
        const unsigned long long warps = (a.n_paths + a.sink.claim_paths - 1) / a.sink.claim_paths;
        unsigned long long need = (warps + block / 32 - 1) / (block / 32);
        if (need < 1) need = 1;
        *grid = (int)(need < resident ? need : resident);
        kernel<<<*grid, block, smem, stream>>>(a);
        return cudaGetLastError();
    }
    cudaError_t e = pick_grid(kernel, block, smem, dev, items, grid);
    if (e != cudaSuccess) return e;
    kernel<<<*grid, block, smem, stream>>>(a);
    return cudaGetLastError();
}

template <int SHOCK, int VAR, int TAB>
static cudaError_t launch32_store(Sim32Args& a, const mc_params& p, const DeviceInfo& dev, int* grid, cudaStream_t s) {
    if (!(p.flags & MC_FLAG_STORE_PATHS))
        return a.sink.g_khist ? launch32_t<SHOCK, VAR, false, 1, TAB, true>(a, dev, grid, s)
                              : launch32_t<SHOCK, VAR, false, 1, TAB>(a, dev, grid, s);
    // 16-byte stores need rows and base pointers that keep every group of four paths aligned
    const bool vec = a.ld_paths % 4 == 0 && ((uintptr_t)a.paths & 15u) == 0 && ((uintptr_t)a.terminal & 15u) == 0 &&
                     a.n_paths >= 4;
    return vec ? launch32_t<SHOCK, VAR, true, 4, TAB>(a, dev, grid, s)
               : launch32_t<SHOCK, VAR, true, 1, TAB>(a, dev, grid, s);
}

template <int SHOCK, int TAB>
static cudaError_t launch32_var(Sim32Args& a, const mc_params& p, const DeviceInfo& dev, int* grid, cudaStream_t s) {
    return p.var_model == MC_VAR_CONST ? launch32_store<SHOCK, MC_VAR_CONST, TAB>(a, p, dev, grid, s)
                                       : launch32_store<SHOCK, MC_VAR_GARCH, TAB>(a, p, dev, grid, s);
}

static cudaError_t launch_sim32(const mc_params& p, const SimBuffers& b, const DeviceInfo& dev, int* grid,
                                cudaStream_t s) {
    Sim32Args a = make_args32(p, b);
    if (p.shock == MC_SHOCK_NORMAL) {
        if (p.flags & MC_FLAG_FULL_RES_NORMALS) return launch32_var<SHOCK_NORMAL_FULL, TAB_GLOBAL>(a, p, dev, grid, s);
        if (!b.d_lut) return cudaErrorInvalidValue;
        return launch32_var<MC_SHOCK_NORMAL, TAB_SMEM>(a, p, dev, grid, s);
    }
    // bootstrap: widest replication (<= 8 copies: a half-warp's LDS.64 then rarely collides) that fits
    const size_t budget = (size_t)dev.max_smem_optin - 4 * 1024 -
                          (a.sink.g_khist ? sink_smem_bytes(a.sink.g_nphist ? a.sink.nph_bins : 0u) : 0);
    int lc = 3;
    while (lc >= 0 && ((size_t)p.pool_len << lc) * sizeof(float2) > budget) --lc;
    if (lc < 0) return launch32_var<MC_SHOCK_BOOTSTRAP, TAB_GLOBAL>(a, p, dev, grid, s);   // gather from global/L1
    if (p.n_paths < 1024ull * dev.sm_count)        // small request: cheaper table fill, more blocks per SM
        while (lc > 0 && ((size_t)p.pool_len << lc) * sizeof(float2) > 40 * 1024) --lc;
    a.table_lc = lc;
    return lc == 3 ? launch32_var<MC_SHOCK_BOOTSTRAP, TAB_SMEM8>(a, p, dev, grid, s)
                   : launch32_var<MC_SHOCK_BOOTSTRAP, TAB_SMEM>(a, p, dev, grid, s);
}

template <typename Real, int D, bool INJECTED, bool STORE>
static cudaError_t launch_portfolio_impl(const mc_params& p, const mc_portfolio& pf, const SimBuffers& b,
                                         const DeviceInfo& dev, int* grid, cudaStream_t stream) {
    const double u = unit_scale<Real>();
    const int d = (int)p.n_assets;
    static PortArgs<Real, D> proto;      // zero-initialised template; copied, never mutated
    PortArgs<Real, D> a = proto;
    a.n_paths = p.n_paths;
    a.path_offset = p.path_offset;
    a.ld_paths = p.paths_ld ? p.paths_ld : p.n_paths;
    a.ld_shocks = p.n_paths;
    a.n_steps = p.n_steps;
    a.rk = expand_key((uint32_t)p.seed, (uint32_t)(p.seed >> 32));
    a.calls_per_step = (uint32_t)((d + NormalsPerCall<Real>::value - 1) / NormalsPerCall<Real>::value);
    a.n_assets = (uint32_t)d;
    a.shocks = static_cast<const Real*>(b.d_shocks);
    a.lut = b.d_lut;
    a.paths = static_cast<Real*>(b.d_paths);
    a.terminal = static_cast<Real*>(b.d_terminal);
    double v0 = 0.0;
    for (int i = 0; i < d; ++i) {
        for (int c = 0; c <= i; ++c) {
            // fp32 device normals come without their constant sqrt(2 ln 2) (normals8<false>): it goes in here
            const double zs = (sizeof(Real) == 4 && !INJECTED) ? (double)BM_RADIUS_SCALE_F : 1.0;
            const Real l = (Real)(pf.h_chol[i * d + c] * pf.h_sigma_step[i] * u * zs);
            if (sizeof(Real) == 4) a.chol[2 * ((i / 2) * (i / 2 + 1) + c) + (i & 1)] = l;      // row pairs (see PortArgs)
            else a.chol[i * (i + 1) / 2 + c] = l;
        }
        a.drift[i] = (Real)(pf.h_drift_step[i] * u);
        a.s0w[i] = (Real)(pf.h_s0[i] * pf.h_weights[i]);
        v0 += pf.h_s0[i] * pf.h_weights[i];
    }
    a.v0 = (Real)v0;
    a.center = p.s0;
    a.partials = b.d_partials;
    auto kernel = portfolio_kernel<Real, D, INJECTED, STORE>;
    const size_t smem = (!INJECTED && sizeof(Real) == 4) ? (size_t)DIR_COUNT * sizeof(float2) : 0;
    if (smem && !b.d_lut) return cudaErrorInvalidValue;
    cudaError_t e = pick_grid(kernel, SIM_BLOCK, smem, dev, p.n_paths, grid);
    if (e != cudaSuccess) return e;
    kernel<<<*grid, SIM_BLOCK, smem, stream>>>(a);
    return cudaGetLastError();
}

template <typename Real, int D>
static cudaError_t launch_portfolio_d(const mc_params& p, const mc_portfolio& pf, const SimBuffers& b,
                                      const DeviceInfo& dev, int* grid, cudaStream_t s) {
    const bool inj = p.shock == MC_SHOCK_INJECTED, st = (p.flags & MC_FLAG_STORE_PATHS) != 0;
    if (inj) return st ? launch_portfolio_impl<Real, D, true, true>(p, pf, b, dev, grid, s)
                       : launch_portfolio_impl<Real, D, true, false>(p, pf, b, dev, grid, s);
    return st ? launch_portfolio_impl<Real, D, false, true>(p, pf, b, dev, grid, s)
              : launch_portfolio_impl<Real, D, false, false>(p, pf, b, dev, grid, s);
}

template <typename Real>
static cudaError_t launch_portfolio(const mc_params& p, const mc_portfolio& pf, const SimBuffers& b,
                                    const DeviceInfo& dev, int* grid, cudaStream_t s) {
    if (p.shock == MC_SHOCK_BOOTSTRAP || p.var_model != MC_VAR_CONST) return cudaErrorInvalidValue;
    const int d = (int)p.n_assets;
    if (d <= 4) return launch_portfolio_d<Real, 4>(p, pf, b, dev, grid, s);
    if (d <= 8) return launch_portfolio_d<Real, 8>(p, pf, b, dev, grid, s);
    if (d <= 16) return launch_portfolio_d<Real, 16>(p, pf, b, dev, grid, s);
    if (d <= 32) return launch_portfolio_d<Real, 32>(p, pf, b, dev, grid, s);
    return cudaErrorInvalidValue;
}

bool sim_takes_sink(const mc_params& p) {
    return p.n_assets == 1 && p.dtype == MC_F32 && !(p.flags & MC_FLAG_STORE_PATHS) &&
           (p.shock == MC_SHOCK_NORMAL || (p.shock == MC_SHOCK_BOOTSTRAP && boot32_eligible(p)));
}

cudaError_t launch_simulation(const mc_params& p, const mc_portfolio* pf, const SimBuffers& b,
                              const DeviceInfo& dev, int* grid, cudaStream_t stream) {
    if (p.n_assets > 1) {
        if (!pf) return cudaErrorInvalidValue;
        return p.dtype == MC_F64 ? launch_portfolio<double>(p, *pf, b, dev, grid, stream)
                                 : launch_portfolio<float>(p, *pf, b, dev, grid, stream);
    }
    if (p.dtype == MC_F64) return dispatch_shock<double>(p, b, dev, grid, stream);
    if (p.shock == MC_SHOCK_NORMAL || (p.shock == MC_SHOCK_BOOTSTRAP && boot32_eligible(p)))
        return launch_sim32(p, b, dev, grid, stream);
    if (p.shock == MC_SHOCK_BOOTSTRAP) return cudaErrorNotSupported;   // fp32 pools beyond BOOT32_MAX_POOL: use fp64
    return dispatch_shock<float>(p, b, dev, grid, stream);
}

cudaError_t launch_build_direction_table(float2* d_lut, cudaStream_t stream) {
    build_direction_table_kernel<<<(DIR_COUNT + 255) / 256, 256, 0, stream>>>(d_lut);
    return cudaGetLastError();
}

template <typename Real>
static cudaError_t debug_shocks_t(const mc_params& p, const void* d_pool, void* d_out, cudaStream_t stream) {
    SimBuffers b{};
    b.d_pool = d_pool;
    SimArgs<Real> a = make_args<Real>(p, b);
    const int grid = (int)((p.n_paths + 255) / 256 < 4096 ? (p.n_paths + 255) / 256 : 4096);
    if (p.shock == MC_SHOCK_NORMAL)
        debug_shocks_kernel<Real, MC_SHOCK_NORMAL><<<grid ? grid : 1, 256, 0, stream>>>(a, static_cast<Real*>(d_out));
    else if (p.shock == MC_SHOCK_BOOTSTRAP)
        debug_shocks_kernel<Real, MC_SHOCK_BOOTSTRAP><<<grid ? grid : 1, 256, 0, stream>>>(a, static_cast<Real*>(d_out));
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

cudaError_t launch_debug_shocks(const mc_params& p, const void* d_pool, const float2* d_lut, void* d_out,
                                cudaStream_t stream) {
    if (p.dtype == MC_F64) return debug_shocks_t<double>(p, d_pool, d_out, stream);
    SimBuffers b{};
    b.d_pool = d_pool;
    b.d_lut = d_lut;
    const Sim32Args a = make_args32(p, b);
    int grid = (int)((p.n_paths + 255) / 256 < 4096 ? (p.n_paths + 255) / 256 : 4096);
    if (grid < 1) grid = 1;
    if (p.shock == MC_SHOCK_NORMAL && (p.flags & MC_FLAG_FULL_RES_NORMALS)) {
        debug_shocks32_kernel<SHOCK_NORMAL_FULL><<<grid, 256, 0, stream>>>(a, static_cast<float*>(d_out));
    } else if (p.shock == MC_SHOCK_NORMAL) {
        if (!d_lut) return cudaErrorInvalidValue;
        debug_shocks32_kernel<MC_SHOCK_NORMAL><<<grid, 256, DIR_COUNT * sizeof(float2), stream>>>(a, static_cast<float*>(d_out));
    } else if (p.shock == MC_SHOCK_BOOTSTRAP && boot32_eligible(p)) {
        debug_shocks32_kernel<MC_SHOCK_BOOTSTRAP><<<grid, 256, 0, stream>>>(a, static_cast<float*>(d_out));
    } else {
        return cudaErrorInvalidValue;
    }
    return cudaGetLastError();
}

cudaError_t launch_debug_philox(const uint32_t* d_ctr, uint32_t n, uint32_t k0, uint32_t k1, uint32_t* d_out,
                                cudaStream_t stream) {
    if (n == 0) return cudaSuccess;
    debug_philox_kernel<<<(n + 127) / 128, 128, 0, stream>>>(d_ctr, n, k0, k1, d_out);
    return cudaGetLastError();
}

cudaError_t launch_reduce_partials(const BlockPartial* d_partials, int n_partials, unsigned long long* d_words,
                                   unsigned long long w6, unsigned long long w7, cudaStream_t stream) {
    reduce_partials_kernel<<<1, REDUCE_THREADS, 0, stream>>>(d_partials, n_partials, d_words, w6, w7);
    return cudaGetLastError();
}

}  // namespace mc
