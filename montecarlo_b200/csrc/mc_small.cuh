// mc_small.cuh -- launch interface of the two-launch path for the reference's own call sizes (mc_small.cu).
#pragma once
#include "mc_common.cuh"

namespace mc {

constexpr unsigned long long SMALL_MAX_PATHS_F64 = 16384, SMALL_MAX_PATHS_F32 = 32768;   // a row must fit in shared memory
constexpr int SMALL_MAX_PARTIALS = 1024;

struct SmallArgs {
    const void* paths;            // [rows, ld] time-major, row rows-1 = the terminal values
    int dtype;
    unsigned long long n, ld;
    unsigned int rows;            // n_steps + 1
    const SelectPlan* plan_t;     // device: terminal percentiles
    const SelectPlan* plan_f;     // device: fan-chart percentiles
    int n_pcts, n_path_pcts, tail_idx;
    const BlockPartial* partials; // the simulation kernel's per-block moments
    int n_partials;
    double center;
    DeviceResult* res;            // device result block
    double* fan;                  // device [rows][n_path_pcts]
    unsigned char* host_out;      // pinned host: DeviceResult, then the fan chart as [n_path_pcts][rows] doubles
    unsigned int* bad;            // pinned host flag: a row's selection found an inconsistency (the general path redoes the job)
};
size_t small_rows_smem(int dtype, unsigned long long n);
cudaError_t launch_small_rows(const SmallArgs& a, cudaStream_t s);

}  // namespace mc
