// microbench.cu -- per-instruction throughput on sm_100a, to pin the denominators the
// simulation kernels are judged against (SURVEY 8d asks for this before trusting the
// roofline): MUFU variants, IMAD.WIDE, LOP3, FFMA, one Philox-4x32-10 call, Box-Muller.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "rng_variants.cuh"

using namespace mc;
constexpr int ITERS = 4096;
constexpr int ILP = 8;

#define MUFU_KERNEL(NAME, ASM)                                                         \
    __global__ void k_##NAME(float* out, float seed) {                                 \
        float x[ILP];                                                                   \
        for (int i = 0; i < ILP; ++i) x[i] = seed + 0.001f * (threadIdx.x + i);         \
        for (int it = 0; it < ITERS; ++it) {                                            \
            _Pragma("unroll") for (int i = 0; i < ILP; ++i)                             \
                asm volatile(ASM : "=f"(x[i]) : "f"(x[i]));                             \
        }                                                                               \
        float s = 0; for (int i = 0; i < ILP; ++i) s += x[i];                           \
        if (s == 123.456f) out[0] = s;                                                  \
    }
MUFU_KERNEL(lg2, "lg2.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(ex2, "ex2.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(sqrt, "sqrt.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(rsqrt, "rsqrt.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(rcp, "rcp.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(sin, "sin.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(cos, "cos.approx.ftz.f32 %0, %1;")
MUFU_KERNEL(tanh, "tanh.approx.f32 %0, %1;")

__global__ void k_ffma(float* out, float seed) {
    float x[ILP];
    for (int i = 0; i < ILP; ++i) x[i] = seed + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = fmaf(x[i], 1.0001f, 0.5f);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}
__global__ void k_ffma3(float* out, float seed) {   // three register operands
    float x[ILP], a = seed, b = seed * 0.5f;
    for (int i = 0; i < ILP; ++i) x[i] = seed + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = fmaf(x[i], a, b);
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}
__global__ void k_imadwide(float* out, uint32_t seed) {
    uint32_t x[ILP];
    for (int i = 0; i < ILP; ++i) x[i] = seed + i + threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            const uint64_t p = (uint64_t)PHILOX_M0 * x[i];
            x[i] = (uint32_t)(p >> 32) ^ (uint32_t)p;     // 1 IMAD.WIDE + 1 LOP3
        }
    }
    uint32_t s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 123456u) out[0] = s;
}
__global__ void k_lop3(float* out, uint32_t seed) {
    uint32_t x[ILP];
    for (int i = 0; i < ILP; ++i) x[i] = seed + i + threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) x[i] = (x[i] ^ 0x9E3779B9u) & (x[(i + 1) % ILP] | 0x55u);
    }
    uint32_t s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 123456u) out[0] = s;
}
__global__ void k_philox(float* out, uint32_t seed, const __grid_constant__ RoundKeys rk) {
    uint32_t acc = 0, c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
    for (int it = 0; it < ITERS / 4; ++it) {
#pragma unroll 2
        for (int j = 0; j < 4; ++j) {
            const u32x4 w = philox4x32_10(c0, 0u, it * 4 + j, 0u, rk);
            acc ^= w.x ^ w.y ^ w.z ^ w.w;
        }
    }
    if (acc == 123456u) out[0] = acc;
}
__global__ void k_boxmuller_only(float* out, uint32_t seed) {   // transform cost without Philox
    uint32_t a = seed + threadIdx.x, b = seed * 3u + threadIdx.x;
    float s = 0.f;
    for (int it = 0; it < ITERS; ++it) {
        a = a * 1664525u + 1013904223u;
        b = b * 22695477u + 1u;
        const float r = bm_rho(a), th = bm_theta(b);
        s = fmaf(r, fast_cos(th), s);
        s = fmaf(r, fast_sin(th), s);
    }
    if (s == 123.456f) out[0] = s;
}
__global__ void k_normals(float* out, uint32_t seed, const __grid_constant__ RoundKeys rk) {   // the fused inner loop
    float s = 0.f;
    const uint32_t c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
#pragma unroll 2
    for (int j = 0; j < ITERS; ++j) s = normals6_sum(philox4x32_10(c0, 0u, j, 0u, rk), s);
    if (s == 123.456f) out[0] = s;
}

template <typename F>
static void run(const char* name, double ops_per_thread, int block, int blocks_per_sm, F launch) {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    const int grid = prop.multiProcessorCount * blocks_per_sm;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch(grid, block);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(a); launch(grid, block); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    const double total = ops_per_thread * (double)grid * block;
    const double per_s = total / (best * 1e-3);
    printf("%-16s block=%4d x%d/SM  %8.3f ms  %9.2f Gops/s  %6.2f ops/clk/SM @%d MHz (nominal max)\n", name, block,
           blocks_per_sm, best, per_s / 1e9, per_s / (prop.multiProcessorCount * (double)khz * 1e3), khz / 1000);
}

int main() {
    float* out; cudaMalloc(&out, 64);
    const RoundKeys rk = expand_key(0x12345678u, 0x9abcdef0u);
    const double n = (double)ITERS * ILP;
#define R(NAME) run(#NAME, n, 256, 8, [&](int g, int b) { k_##NAME<<<g, b>>>(out, 1.5f); });
    R(lg2) R(ex2) R(sqrt) R(rsqrt) R(rcp) R(sin) R(cos) R(tanh) R(ffma) R(ffma3)
    run("imadwide+lop3", n, 256, 8, [&](int g, int b) { k_imadwide<<<g, b>>>(out, 7u); });
    run("lop3 x2", 2 * n, 256, 8, [&](int g, int b) { k_lop3<<<g, b>>>(out, 7u); });
    for (int bps : {2, 4, 8})
        run("philox call", (double)ITERS, 256, bps, [&](int g, int b) { k_philox<<<g, b>>>(out, 7u, rk); });
    run("boxmuller pair", (double)ITERS, 256, 8, [&](int g, int b) { k_boxmuller_only<<<g, b>>>(out, 7u); });
    for (int bps : {2, 4, 6, 8})
        run("6 normals/call", (double)ITERS, 256, bps, [&](int g, int b) { k_normals<<<g, b>>>(out, 7u, rk); });
    run("6 normals b1024", (double)ITERS, 1024, 2, [&](int g, int b) { k_normals<<<g, b>>>(out, 7u, rk); });
    run("6 normals b128", (double)ITERS, 128, 16, [&](int g, int b) { k_normals<<<g, b>>>(out, 7u, rk); });
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
