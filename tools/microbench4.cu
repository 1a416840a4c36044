// microbench4.cu -- packed fp32 (FFMA2 / FMUL2, sm_100's fma/mul.rn.f32x2) issue rates on sm_100a, alone and next to
// the instructions they share a pipe with in the simulation kernels (IMAD.WIDE, FFMA).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench4 microbench4.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;
static int g_sms, g_khz;
enum { OP_FFMA, OP_FFMA2, OP_FMUL2, OP_FFMA2_RM, OP_FFMA2_PLUS_WIDE, OP_2FFMA_PLUS_WIDE, OP_FFMA2_PLUS_FFMA, OP_WIDE };

template <int OP>
__global__ void k_op(uint32_t* out, uint32_t seed) {
    uint32_t y[ILP], z[ILP];
    float f[ILP], g[ILP];
    unsigned long long p[ILP];
    const float m = 1.0000001f, c = 0.5f;
    unsigned long long m2, c2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(m));
    asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(c));
    for (int i = 0; i < ILP; ++i) {
        y[i] = seed + i + threadIdx.x; z[i] = seed * 3 + i; f[i] = 1.0f + i; g[i] = 2.0f + i;
        asm("mov.b64 %0, {%1, %2};" : "=l"(p[i]) : "f"(f[i]), "f"(g[i]));
    }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == OP_FFMA || OP == OP_2FFMA_PLUS_WIDE || OP == OP_FFMA2_PLUS_FFMA)
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(f[i]) : "f"(m), "f"(c));
            if (OP == OP_2FFMA_PLUS_WIDE)
                asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(g[i]) : "f"(m), "f"(c));
            if (OP == OP_FFMA2 || OP == OP_FFMA2_PLUS_WIDE || OP == OP_FFMA2_PLUS_FFMA)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
            if (OP == OP_FFMA2_RM)
                asm volatile("fma.rm.f32x2 %0, %0, %1, %2;" : "+l"(p[i]) : "l"(m2), "l"(c2));
            if (OP == OP_FMUL2)
                asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(m2));
            if (OP == OP_WIDE || OP == OP_FFMA2_PLUS_WIDE || OP == OP_2FFMA_PLUS_WIDE) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 q; mul.wide.u32 q, %2, %3; mov.b64 {%0,%1}, q;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(0xD2511F53u));
                y[i] = hi; z[i] = lo;
            }
        }
    }
    uint32_t acc = 0;
    for (int i = 0; i < ILP; ++i) {
        float a, b;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(p[i]));
        acc ^= y[i] ^ z[i] ^ __float_as_uint(f[i]) ^ __float_as_uint(g[i]) ^ __float_as_uint(a) ^ __float_as_uint(b);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP>
static void run(const char* name, double groups_per_iter, uint32_t* d_out) {
    const int blocks = g_sms * 4, threads = 512;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_op<OP><<<blocks, threads>>>(d_out, 1u);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_op<OP><<<blocks, threads>>>(d_out, 2u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double n = (double)blocks * threads * ITERS * ILP * groups_per_iter;
    printf("%-44s %8.3f ms  %9.3f lane-groups per clk per SM @%d MHz\n", name, ms, n / (ms * 1e-3) / g_sms / (g_khz * 1e3), g_khz / 1000);
}

int main() {
    cudaDeviceProp pr;
    cudaGetDeviceProperties(&pr, 0);
    g_sms = pr.multiProcessorCount;
    cudaDeviceGetAttribute(&g_khz, cudaDevAttrClockRate, 0);
    uint32_t* d_out;
    cudaMalloc(&d_out, (size_t)g_sms * 4 * 512 * 4);
    printf("# one 'lane-group' = the listed instruction(s) for one thread\n");
    run<OP_FFMA>("fma.rn.f32 (FFMA)", 1, d_out);
    run<OP_FFMA2>("fma.rn.f32x2 (FFMA2)", 1, d_out);
    run<OP_FFMA2_RM>("fma.rm.f32x2 (FFMA2.RM)", 1, d_out);
    run<OP_FMUL2>("mul.rn.f32x2 (FMUL2)", 1, d_out);
    run<OP_WIDE>("mul.wide.u32 (IMAD.WIDE)", 1, d_out);
    run<OP_FFMA2_PLUS_WIDE>("FFMA2 + IMAD.WIDE (pairs)", 1, d_out);
    run<OP_2FFMA_PLUS_WIDE>("2 FFMA + IMAD.WIDE (triples)", 1, d_out);
    run<OP_FFMA2_PLUS_FFMA>("FFMA2 + FFMA (pairs)", 1, d_out);
    return 0;
}
