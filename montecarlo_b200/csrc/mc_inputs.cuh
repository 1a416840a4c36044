// mc_inputs.cuh -- launcher of the input producers (mc_inputs.cu)
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "../../include/mc_b200.h"

namespace mc {

cudaError_t launch_prepare_inputs(const double* d_close, const uint32_t* d_lengths, uint32_t n_series, uint32_t n_obs,
                                  unsigned long long close_ld, uint32_t first_row, int dtype, void* d_pool,
                                  unsigned long long pool_ld, mc_series_inputs* d_out, cudaStream_t stream);

cudaError_t launch_calibrate_garch(const double* d_pool, unsigned long long pool_ld, const mc_series_inputs* d_inputs,
                                   uint32_t n_series, uint32_t min_history, uint32_t levels, mc_garch_fit* d_fit,
                                   cudaStream_t stream);

}  // namespace mc
