// mc_sim.cuh -- launch interface of the path-stepping kernels (mc_sim.cu).
#pragma once
#include "mc_common.cuh"
#include "mc_fused.cuh"

namespace mc {

struct SimBuffers {
    const void* d_pool;
    const void* d_shocks;
    void* d_paths;
    void* d_terminal;
    BlockPartial* d_partials;     // [max_grid]
    const float2* d_lut;          // Box-Muller direction table (fp32 normal shocks), [4096]
    const SinkArgs* sink;         // non-null: bin the terminal values inside the kernel (only if sim_takes_sink(p))
};

// true when the kernel launch_simulation picks for `p` can bin its terminal values itself (sim32_kernel<..., SINK>:
// fp32, one asset, shocks generated on the device, no stored paths); it then leaves NO per-block moments behind
bool sim_takes_sink(const mc_params& p);

// Test / tuning knobs (mc_debug_set_tuning): waves of blocks per launch of the non-persistent kernels (>= 1),
// threads per block of the fused GBM kernel (128, 256 or 512).  0 restores the default.  Process-wide.
void set_sim_tuning(int waves, int nblock);

constexpr int MAX_SIM_GRID = 1 << 16;     // per-block partials are reduced in fixed order afterwards

struct DeviceInfo {
    int sm_count;
    int max_smem_optin;
};

// Enqueue the simulation of p.n_paths paths; returns the grid size used in *grid.
cudaError_t launch_simulation(const mc_params& p, const mc_portfolio* pf, const SimBuffers& b,
                              const DeviceInfo& dev, int* grid, cudaStream_t stream);

// z[t, i] exactly as the simulator draws them (tests)
cudaError_t launch_debug_shocks(const mc_params& p, const void* d_pool, const float2* d_lut, void* d_out,
                                cudaStream_t stream);
// fill the handle's 4096-entry (cos, sin) table of Box-Muller directions (once, at mc_create)
cudaError_t launch_build_direction_table(float2* d_lut, cudaStream_t stream);
cudaError_t launch_debug_philox(const uint32_t* d_ctr, uint32_t n, uint32_t k0, uint32_t k1,
                                uint32_t* d_out, cudaStream_t stream);

// fixed-order reduction of the per-block partials into the stage-0 header words [0..5];
// header words [6], [7] are set to w6, w7 (n_partials may be 0: header only)
cudaError_t launch_reduce_partials(const BlockPartial* d_partials, int n_partials,
                                   unsigned long long* d_stage_words, unsigned long long w6,
                                   unsigned long long w7, cudaStream_t stream);

// ---- many price histories in one launch (mc_run_batch) ---------------------------------
constexpr int BATCH_BLOCK = 128;
struct BatchSimArgs {
    unsigned long long n_paths, ld_terminal, ld_paths, pool_ld, seed;
    uint32_t n_steps, n_series, min_history;
    int pool_in_smem;                     // set by the launcher
    double omega, alpha, beta;
    const double* pool;                   // [n_series, pool_ld]
    const mc_series_inputs* inputs;       // [n_series] (device)
    const mc_garch_fit* fit;              // [n_series] per-history omega/alpha/beta (device) or NULL
    double* paths;                        // [n_series, n_steps + 1, ld_paths] or NULL
    double* terminal;                     // [n_series, ld_terminal]
};
cudaError_t launch_batch_simulation(const BatchSimArgs& args, bool store, int sm_count, cudaStream_t stream);

// per history: moments about s0, exceedance count, tail (count, sum) at the percentile `tail_idx`, and the
// percentiles themselves, written as one mc_result per history (device memory)
cudaError_t launch_batch_stats(const double* d_terminal, unsigned long long n, unsigned long long ld,
                               uint32_t n_series, const mc_series_inputs* d_inputs, const mc_garch_fit* d_fit,
                               uint32_t min_history, const double* d_pct, const double* d_lo, const double* d_hi, int n_pcts,
                               int tail_idx, mc_result* d_results, cudaStream_t stream);

}  // namespace mc
