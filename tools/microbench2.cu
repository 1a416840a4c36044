// microbench2.cu -- design-space probe for the shock generators on sm_100a (round 1, v3):
//   * do MUFU and IMAD.WIDE overlap (independent chains in one loop)?
//   * normals per Philox call: 6 (v2) vs 8, with 0..4 of the 4 pairs taking sin/cos from a polynomial
//   * the materialising loop (normal -> log-price -> ex2 -> coalesced store), same variants
//   * bootstrap + GARCH step: 4 indices per call by IMAD.HI (v2) vs 5 per call by FFMA floor
//   * accuracy of the eight-normal transforms against fp64
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench2 microbench2.cu
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
#include "rng_variants.cuh"

using namespace mc;
constexpr int ITERS = 4096;
constexpr int ILP = 8;

__global__ void k_mufu_only(float* out, float seed) {
    float x[ILP];
    for (int i = 0; i < ILP; ++i) x[i] = seed + 0.001f * (threadIdx.x + i);
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(x[i]) : "f"(x[i]));
    }
    float s = 0; for (int i = 0; i < ILP; ++i) s += x[i];
    if (s == 123.456f) out[0] = s;
}
__global__ void k_imad_only(float* out, uint32_t seed) {
    uint32_t y[ILP];
    for (int i = 0; i < ILP; ++i) y[i] = seed + i + threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            const uint64_t p = (uint64_t)PHILOX_M0 * y[i];
            y[i] = (uint32_t)(p >> 32) ^ (uint32_t)p;
        }
    }
    uint32_t s = 0; for (int i = 0; i < ILP; ++i) s += y[i];
    if (s == 123456u) out[0] = s;
}
// NM MUFU and NI IMAD.WIDE(+LOP3) per iteration, independent chains
template <int NM, int NI>
__global__ void k_mix(float* out, float seed) {
    float x[NM > 0 ? NM : 1];
    uint32_t y[NI > 0 ? NI : 1];
    for (int i = 0; i < NM; ++i) x[i] = seed + 0.001f * (threadIdx.x + i);
    for (int i = 0; i < NI; ++i) y[i] = (uint32_t)seed + i + threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < (NM > NI ? NM : NI); ++i) {
            if (i < NM) asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(x[i]) : "f"(x[i]));
            if (i < NI) {
                const uint64_t p = (uint64_t)PHILOX_M0 * y[i];
                y[i] = (uint32_t)(p >> 32) ^ (uint32_t)p;
            }
        }
    }
    float s = 0; for (int i = 0; i < NM; ++i) s += x[i];
    uint32_t q = 0; for (int i = 0; i < NI; ++i) q += y[i];
    if (s == 123.456f || q == 123456u) out[0] = s + q;
}
// NF FFMA and NI IMAD.WIDE per iteration
template <int NF, int NI>
__global__ void k_mix_ffma(float* out, float seed) {
    float x[NF];
    uint32_t y[NI > 0 ? NI : 1];
    for (int i = 0; i < NF; ++i) x[i] = seed + 0.001f * (threadIdx.x + i);
    for (int i = 0; i < NI; ++i) y[i] = (uint32_t)seed + i + threadIdx.x;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < (NF > NI ? NF : NI); ++i) {
            if (i < NF) x[i] = fmaf(x[i], 1.0001f, 0.5f);
            if (i < NI) {
                const uint64_t p = (uint64_t)PHILOX_M0 * y[i];
                y[i] = (uint32_t)(p >> 32) ^ (uint32_t)p;
            }
        }
    }
    float s = 0; for (int i = 0; i < NF; ++i) s += x[i];
    uint32_t q = 0; for (int i = 0; i < NI; ++i) q += y[i];
    if (s == 123.456f || q == 123456u) out[0] = s + q;
}

__global__ void k_n6(float* out, uint32_t seed, const __grid_constant__ RoundKeys rk) {
    float s = 0.f;
    const uint32_t c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
#pragma unroll 2
    for (int j = 0; j < ITERS; ++j) s = normals6_sum(philox4x32_10(c0, 0u, j, 0u, rk), s);
    if (s == 123.456f) out[0] = s;
}
template <int PP, int UNR>
__global__ void k_n8(float* out, uint32_t seed, const __grid_constant__ RoundKeys rk) {
    float s = 0.f;
    const uint32_t c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
#pragma unroll UNR
    for (int j = 0; j < ITERS; ++j) s = normals8_sum<PP>(philox4x32_10(c0, 0u, j, 0u, rk), s);
    if (s == 123.456f) out[0] = s;
}


// LUT variant: L of the 4 pairs read (cos, sin) of a 12-bit angle from a 32 KB shared-memory table
template <int L, int PP>
__global__ void k_n8lut(float* out, uint32_t seed, const float2* __restrict__ g_lut, const __grid_constant__ RoundKeys rk) {
    extern __shared__ float2 s_lut[];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) s_lut[i] = g_lut[i];
    __syncthreads();
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(s_lut);
    float acc = 0.f;
    const uint32_t c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
#pragma unroll 2
    for (int j = 0; j < ITERS; ++j) {
        const u32x4 w = philox4x32_10(c0, 0u, j, 0u, rk);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float r = rho20(ww[k]);
            float c, s;
            if (k < L) {
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c), "=f"(s) : "r"(lut_base + ((ww[k] >> 17) & 0x7FF8u)));
            } else if (k < L + PP) sincos_turn_poly(mantissa_float(ww[k]), c, s);
            else sincos_turn_mufu(mantissa_float(ww[k]), c, s);
            acc = fmaf(r, c, acc);
            acc = fmaf(r, s, acc);
        }
    }
    if (acc == 123.456f) out[0] = acc;
}


// all four pairs from the LUT, LC = log2(copies of the table); lane l reads copy (l mod copies)
template <int LC>
__global__ void k_n8lutc(float* out, uint32_t seed, const float2* __restrict__ g_lut, const __grid_constant__ RoundKeys rk) {
    extern __shared__ float2 s_lut[];
    for (int i = threadIdx.x; i < (4096 << LC); i += blockDim.x) s_lut[i] = g_lut[i >> LC];
    __syncthreads();
    const uint32_t lut_base = (uint32_t)__cvta_generic_to_shared(s_lut);
    const uint32_t lane_off = (threadIdx.x & ((1u << LC) - 1u)) << 3;
    float acc = 0.f;
    const uint32_t c0 = seed + threadIdx.x + blockIdx.x * blockDim.x;
#pragma unroll 2
    for (int j = 0; j < ITERS; ++j) {
        const u32x4 w = philox4x32_10(c0, 0u, j, 0u, rk);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float r = rho20(ww[k]);
            float c, s;
            const uint32_t off = ((ww[k] >> (17 - LC)) & (0x7FF8u << LC)) | lane_off;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c), "=f"(s) : "r"(lut_base + off));
            acc = fmaf(r, c, acc);
            acc = fmaf(r, s, acc);
        }
    }
    if (acc == 123.456f) out[0] = acc;
}

// bootstrap + GARCH with a (z, alpha z^2 + beta) table: v' = fma(g, v, omega), y = fma(z, sqrt(v), y)
template <int LC>
__global__ void __launch_bounds__(1024) k_boot2(float* out, const float* __restrict__ pool, uint32_t pool_len, uint32_t n_calls,
                                                float var0, float omega, float alpha, float beta,
                                                const __grid_constant__ RoundKeys rk) {
    extern __shared__ float2 s_tab[];
    for (uint32_t i = threadIdx.x; i < (pool_len << LC); i += blockDim.x) {
        const float z = pool[i >> LC];
        s_tab[i] = make_float2(z, fmaf(alpha, z * z, beta));
    }
    __syncthreads();
    const uint32_t c0 = threadIdx.x + blockIdx.x * blockDim.x;
    const uint32_t sub = threadIdx.x & ((1u << LC) - 1u);
    // (tb << (LC+3)) wraps: 0x4B000000 << (LC+3) mod 2^32
    const uint32_t wrap = (uint32_t)(((unsigned long long)0x4B000000u << (LC + 3)) & 0xFFFFFFFFull);
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(s_tab) + sub * 8u - (8u << LC) - wrap;
    float y = 0.f, v = var0;
    const float cH = (float)pool_len * 1.1920928955078125e-07f;
    const float cK = 8388608.0f - (float)pool_len + 0.5f;
#pragma unroll 2
    for (uint32_t j = 0; j < n_calls; ++j) {
        const u32x4 w = philox4x32_10(c0, 0u, j, 0u, rk);
        const uint32_t f[5] = {w.x, w.y, w.z, w.w, __byte_perm(__byte_perm(w.x, w.y, 0x0073), w.z, 0x7710)};
#pragma unroll
        for (int k = 0; k < 5; ++k) {
            const float g = __uint_as_float((f[k] & 0x007FFFFFu) | 0x4B000000u);
            const uint32_t tb = __float_as_uint(fmaf(g, cH, cK));
            float z, gg;
            asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(z), "=f"(gg) : "r"((tb << (LC + 3)) + sbase));
            y = fmaf(z, fast_sqrt(v), y);
            v = fmaf(gg, v, omega);
        }
    }
    if (y == 123.456f) out[0] = y + v;
}

// materialising loop: one thread per path, time-major coalesced store per step
template <int NPC, int PP>
__global__ void __launch_bounds__(256) k_store(float* __restrict__ paths, unsigned long long n_paths, uint32_t n_steps,
                                               float sig, float drift, float s0,
                                               const __grid_constant__ RoundKeys rk) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_paths; i += stride) {
        float y = 0.f;
        paths[i] = s0;
        uint32_t t = 0;
        for (uint32_t j = 0; j < n_steps / NPC; ++j) {
            float z[8];
            const u32x4 w = philox4x32_10((uint32_t)i, 0u, j, 0u, rk);
            if (NPC == 6) { float z6[6]; normals6(w, z6); for (int k = 0; k < 6; ++k) z[k] = z6[k]; }
            else normals8<PP>(w, z);
#pragma unroll
            for (int k = 0; k < NPC; ++k) {
                y = fmaf(sig, z[k], y);
                paths[(unsigned long long)(t + k + 1) * n_paths + i] = s0 * fast_ex2(fmaf((float)(t + k + 1), drift, y));
            }
            t += NPC;
        }
    }
}
// pure store of the same shape (upper bound for the store pattern)
__global__ void __launch_bounds__(256) k_store_only(float* __restrict__ paths, unsigned long long n_paths, uint32_t n_steps) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_paths; i += stride) {
        float y = (float)i;
        for (uint32_t t = 0; t <= n_steps; ++t) { paths[(unsigned long long)t * n_paths + i] = y; y += 1.0f; }
    }
}


// store-pattern probes at the C5 shape: V = paths per thread (1: st.32, 4: st.128), CS = streaming hint
template <int V, bool CS>
__global__ void __launch_bounds__(256) k_store_pat(float* __restrict__ paths, unsigned long long n_paths, uint32_t n_steps) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x * V;
    for (unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * V; i < n_paths; i += stride) {
        float y = (float)i;
        for (uint32_t t = 0; t <= n_steps; ++t) {
            float* p = paths + (unsigned long long)t * n_paths + i;
            if (V == 4) { const float4 v = make_float4(y, y + 1, y + 2, y + 3); if (CS) __stcs(reinterpret_cast<float4*>(p), v); else *reinterpret_cast<float4*>(p) = v; }
            else { if (CS) __stcs(p, y); else *p = y; }
            y += 1.0f;
        }
    }
}
__global__ void __launch_bounds__(256) k_linear_fill(float4* __restrict__ p, unsigned long long n4) {
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride)
        p[i] = make_float4(1.f, 2.f, 3.f, 4.f);
}

// bootstrap + GARCH(1,1) step (app.py:620-633), replicated pool in shared memory
template <int MODE>   // 0: 4 idx/call by mulhi   1: 5 idx/call, 23-bit fields, FFMA floor
__global__ void __launch_bounds__(1024) k_boot(float* out, const float* __restrict__ pool, uint32_t pool_len, uint32_t n_calls,
                                               float var0, float omega, float alpha, float beta,
                                               const __grid_constant__ RoundKeys rk) {
    extern __shared__ float s_pool[];
    for (uint32_t i = threadIdx.x; i < pool_len * 32; i += blockDim.x) s_pool[i] = pool[i >> 5];
    __syncthreads();
    const uint32_t c0 = threadIdx.x + blockIdx.x * blockDim.x;
    const uint32_t sub = threadIdx.x & 31u;
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(s_pool) + sub * 4u - 128u - 0x80000000u;
    float y = 0.f, v = var0;
    const float cH = (float)pool_len * 1.1920928955078125e-07f;              // H * 2^-23, exact
    const float cK = 8388608.0f - (float)pool_len + 0.5f;                    // exact: below 2^23 the ulp is 1/2
#pragma unroll 2
    for (uint32_t j = 0; j < n_calls; ++j) {
        const u32x4 w = philox4x32_10(c0, 0u, j, 0u, rk);
        if (MODE == 0) {
            const uint32_t u[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float z = s_pool[(__umulhi(u[k], pool_len) << 5) + sub];
                const float r = z * fast_sqrt(v);
                y += r;
                v = fmaf(alpha, r * r, fmaf(beta, v, omega));
            }
        } else {
            const uint32_t f[5] = {w.x, w.y, w.z, w.w, __byte_perm(__byte_perm(w.x, w.y, 0x0073), w.z, 0x7710)};
#pragma unroll
            for (int k = 0; k < 5; ++k) {
                const float g = __uint_as_float((f[k] & 0x007FFFFFu) | 0x4B000000u);   // 2^23 + field, exact
                const uint32_t tb = __float_as_uint(fmaf(g, cH, cK));                 // bits(2^23 + floor(field*H/2^23) + 1)
                float z;                                                               // (tb << 7) wraps to 2^31 + 128*(idx+1)
                asm volatile("ld.shared.f32 %0, [%1];" : "=f"(z) : "r"((tb << 7) + sbase));
                const float r = z * fast_sqrt(v);
                y += r;
                v = fmaf(alpha, r * r, fmaf(beta, v, omega));
            }
        }
    }
    if (y == 123.456f) out[0] = y + v;
}

// accuracy: max |z - z_fp64| over a sample of words for normals8<0> and normals8<4>
__global__ void k_accuracy(double* err, uint32_t n, const __grid_constant__ RoundKeys rk) {
    double e0 = 0.0, e4 = 0.0, s1 = 0.0, s2 = 0.0;
    for (uint32_t i = threadIdx.x + blockIdx.x * blockDim.x; i < n; i += gridDim.x * blockDim.x) {
        const u32x4 w = philox4x32_10(i, 0u, 0u, 7u, rk);
        float a[8], b[8];
        normals8<0>(w, a);
        normals8<4>(w, b);
        const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
        for (int k = 0; k < 4; ++k) {
            const double v1 = 1.0 - (double)(ww[k] & 0xFFFFFu) * 9.5367431640625e-07;
            const double m = (double)__uint_as_float((ww[k] >> 9) | 0x3F800000u);
            const double th = 6.283185307179586476925 * (m - 1.5);
            const double r = sqrt(-2.0 * log(v1));
            const double zc = r * cos(th), zs = r * sin(th);
            e0 = fmax(e0, fmax(fabs(a[2 * k] - zc), fabs(a[2 * k + 1] - zs)));
            e4 = fmax(e4, fmax(fabs(b[2 * k] - zc), fabs(b[2 * k + 1] - zs)));
            s1 += zc + zs; s2 += zc * zc + zs * zs;
        }
    }
    // block-level max via atomics on bit patterns (values are non-negative)
    atomicMax(reinterpret_cast<unsigned long long*>(err), (unsigned long long)__double_as_longlong(e0));
    atomicMax(reinterpret_cast<unsigned long long*>(err) + 1, (unsigned long long)__double_as_longlong(e4));
    atomicAdd(err + 2, s1);
    atomicAdd(err + 3, s2);
}

static int g_sms = 0, g_khz = 0;
template <typename F>
static float time_best(F launch, int reps = 5) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < reps; ++rep) {
        cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    return best;
}
static void report(const char* name, double units, float ms) {
    const double per_s = units / (ms * 1e-3);
    printf("%-34s %9.3f ms  %10.2f G/s  %7.3f per clk per SM @%d MHz\n", name, ms, per_s / 1e9,
           per_s / (g_sms * (double)g_khz * 1e3), g_khz / 1000);
}

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    g_sms = prop.multiProcessorCount;
    cudaDeviceGetAttribute(&g_khz, cudaDevAttrClockRate, 0);
    float* out; cudaMalloc(&out, 64);
    const RoundKeys rk = expand_key(0x12345678u, 0x9abcdef0u);
    const int grid = g_sms * 8, block = 256;
    const double thr = (double)grid * block;

    report("mufu x8", thr * ITERS * 8, time_best([&] { k_mufu_only<<<grid, block>>>(out, 1.5f); }));
    report("imad.wide x8", thr * ITERS * 8, time_best([&] { k_imad_only<<<grid, block>>>(out, 7u); }));
    report("mix mufu4+imad8 (iters)", thr * ITERS, time_best([&] { k_mix<4, 8><<<grid, block>>>(out, 1.5f); }));
    report("mix mufu4+imad0 (iters)", thr * ITERS, time_best([&] { k_mix<4, 0><<<grid, block>>>(out, 1.5f); }));
    report("mix mufu0+imad8 (iters)", thr * ITERS, time_best([&] { k_mix<0, 8><<<grid, block>>>(out, 1.5f); }));
    report("mix mufu8+imad8 (iters)", thr * ITERS, time_best([&] { k_mix<8, 8><<<grid, block>>>(out, 1.5f); }));
    report("mix ffma16+imad8 (iters)", thr * ITERS, time_best([&] { k_mix_ffma<16, 8><<<grid, block>>>(out, 1.5f); }));
    report("mix ffma16+imad0 (iters)", thr * ITERS, time_best([&] { k_mix_ffma<16, 0><<<grid, block>>>(out, 1.5f); }));
    report("mix ffma32+imad8 (iters)", thr * ITERS, time_best([&] { k_mix_ffma<32, 8><<<grid, block>>>(out, 1.5f); }));

    report("normals: 6/call (v2)", thr * ITERS * 6, time_best([&] { k_n6<<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 0/4", thr * ITERS * 8, time_best([&] { k_n8<0, 2><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 1/4", thr * ITERS * 8, time_best([&] { k_n8<1, 2><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 2/4", thr * ITERS * 8, time_best([&] { k_n8<2, 2><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 3/4", thr * ITERS * 8, time_best([&] { k_n8<3, 2><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 4/4", thr * ITERS * 8, time_best([&] { k_n8<4, 2><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 0/4 unroll1", thr * ITERS * 8, time_best([&] { k_n8<0, 1><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 1/4 unroll1", thr * ITERS * 8, time_best([&] { k_n8<1, 1><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 0/4 unroll4", thr * ITERS * 8, time_best([&] { k_n8<0, 4><<<grid, block>>>(out, 7u, rk); }));
    report("normals: 8/call poly 1/4 unroll4", thr * ITERS * 8, time_best([&] { k_n8<1, 4><<<grid, block>>>(out, 7u, rk); }));
    for (int bps : {4, 6}) {
        const int g2 = g_sms * bps;
        char nm[64]; snprintf(nm, sizeof nm, "normals: 8/call poly 1/4 x%d/SM", bps);
        report(nm, (double)g2 * block * ITERS * 8, time_best([&] { k_n8<1, 2><<<g2, block>>>(out, 7u, rk); }));
    }


    {
        float2* g_lut; cudaMalloc(&g_lut, 4096 * sizeof(float2));
        float2 h[4096];
        for (int i = 0; i < 4096; ++i) { const double th = 6.283185307179586476925 * (i / 4096.0 - 0.5); h[i] = make_float2((float)cos(th), (float)sin(th)); }
        cudaMemcpy(g_lut, h, sizeof h, cudaMemcpyHostToDevice);
#define LUTRUN(L, PP, B, BPS) { cudaFuncSetAttribute(k_n8lut<L, PP>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768); \
        const int g2 = g_sms * BPS; char nm[64]; snprintf(nm, sizeof nm, "normals: 8/call lut %d poly %d b%d x%d", L, PP, B, BPS); \
        report(nm, (double)g2 * B * ITERS * 8, time_best([&] { k_n8lut<L, PP><<<g2, B, 32768>>>(out, 7u, g_lut, rk); })); }
        LUTRUN(1, 0, 512, 4) LUTRUN(2, 0, 512, 4) LUTRUN(3, 0, 512, 4) LUTRUN(4, 0, 512, 4)
        LUTRUN(1, 1, 512, 4) LUTRUN(2, 1, 512, 4) LUTRUN(3, 1, 512, 4)
        LUTRUN(2, 0, 1024, 2) LUTRUN(3, 0, 1024, 2) LUTRUN(2, 0, 256, 6) LUTRUN(3, 0, 256, 6)

#define LUTC(LC, B, BPS) { cudaFuncSetAttribute(k_n8lutc<LC>, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768 << LC); \
        const int g2 = g_sms * BPS; char nm[64]; snprintf(nm, sizeof nm, "normals: 8/call all-lut copies %d b%d x%d", 1 << LC, B, BPS); \
        report(nm, (double)g2 * B * ITERS * 8, time_best([&] { k_n8lutc<LC><<<g2, B, 32768 << LC>>>(out, 7u, g_lut, rk); })); }
        LUTC(0, 512, 4) LUTC(0, 1024, 2) LUTC(1, 512, 3) LUTC(1, 1024, 2) LUTC(2, 1024, 1) LUTC(2, 512, 1) LUTC(1, 1024, 1) LUTC(0, 1024, 1)
        cudaFree(g_lut);
    }

    // materialising loop at the C5 shape
    {
        const unsigned long long n_paths = 4194304ull; const uint32_t n_steps = 1024;   // 6 | 1020 only: use 1008/1024
        float* paths;
        if (cudaMalloc(&paths, (size_t)(n_steps + 1) * n_paths * 4) == cudaSuccess) {
            const int gs = g_sms * 8;
            const double bytes = (double)(n_steps + 1) * n_paths * 4;
            float ms = time_best([&] { k_store_only<<<gs, 256>>>(paths, n_paths, n_steps); }, 3);
            printf("%-34s %9.3f ms  %10.2f GB/s\n", "store only (C5 shape)", ms, bytes / ms / 1e6);
            ms = time_best([&] { k_store<6, 0><<<gs, 256>>>(paths, n_paths, 1020, 0.0126f, 2e-4f, 100.f, rk); }, 3);
            printf("%-34s %9.3f ms  %10.2f GB/s\n", "store 6/call (v2), 1020 steps", ms, 1021.0 * n_paths * 4 / ms / 1e6);
#define STORE8(PP) ms = time_best([&] { k_store<8, PP><<<gs, 256>>>(paths, n_paths, n_steps, 0.0126f, 2e-4f, 100.f, rk); }, 3); \
            printf("store 8/call poly %d/4               %9.3f ms  %10.2f GB/s\n", PP, ms, bytes / ms / 1e6);
            STORE8(0) STORE8(1) STORE8(2) STORE8(3) STORE8(4)

            ms = time_best([&] { cudaMemsetAsync(paths, 0, (size_t)bytes); }, 3);
            printf("%-34s %9.3f ms  %10.2f GB/s\n", "cudaMemset", ms, bytes / ms / 1e6);
            ms = time_best([&] { k_linear_fill<<<g_sms * 16, 256>>>(reinterpret_cast<float4*>(paths), (unsigned long long)(bytes / 16)); }, 3);
            printf("%-34s %9.3f ms  %10.2f GB/s\n", "linear float4 fill", ms, bytes / ms / 1e6);
#define PAT(V, CS, BPS, B) ms = time_best([&] { k_store_pat<V, CS><<<g_sms * BPS, B>>>(paths, n_paths, n_steps); }, 3); \
            printf("store pattern V=%d cs=%d b%d x%d        %9.3f ms  %10.2f GB/s\n", V, (int)CS, B, BPS, ms, bytes / ms / 1e6);
            PAT(1, false, 8, 256) PAT(1, true, 8, 256) PAT(4, false, 8, 256) PAT(4, true, 8, 256) PAT(4, true, 4, 256) PAT(4, true, 2, 256)
            PAT(1, true, 16, 128) PAT(4, true, 16, 128) PAT(1, true, 4, 256) PAT(1, true, 2, 1024) 
            cudaFree(paths);
        } else printf("store test skipped: allocation failed\n");
    }

    // bootstrap + GARCH
    {
        const uint32_t H = 1250;
        float* pool; cudaMalloc(&pool, H * 4);
        float hp[1250]; for (uint32_t i = 0; i < H; ++i) hp[i] = 0.01f * (float)((int)(i % 17) - 8);
        cudaMemcpy(pool, hp, H * 4, cudaMemcpyHostToDevice);
        const size_t smem = (size_t)H * 32 * 4;
        cudaFuncSetAttribute(k_boot<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(k_boot<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        const int g = g_sms, b = 1024;
        const uint32_t steps = 20000;
        report("bootstrap+garch 4/call mulhi", (double)g * b * (steps / 4) * 4,
               time_best([&] { k_boot<0><<<g, b, smem>>>(out, pool, H, steps / 4, 1.6e-4f, 1e-5f, 0.1f, 0.85f, rk); }));
        report("bootstrap+garch 5/call ffma", (double)g * b * (steps / 5) * 5,
               time_best([&] { k_boot<1><<<g, b, smem>>>(out, pool, H, steps / 5, 1.6e-4f, 1e-5f, 0.1f, 0.85f, rk); }));
    


        cudaFuncSetAttribute(k_boot2<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)H * 16 * 8));
        report("bootstrap+garch 5/call f2 tab x16", (double)g * b * (steps / 5) * 5,
               time_best([&] { k_boot2<4><<<g, b, (size_t)H * 16 * 8>>>(out, pool, H, steps / 5, 1.6e-4f, 1e-5f, 0.1f, 0.85f, rk); }));
        cudaFuncSetAttribute(k_boot2<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((size_t)H * 8 * 8));
        report("bootstrap+garch 5/call f2 tab x8", (double)g * b * (steps / 5) * 5,
               time_best([&] { k_boot2<3><<<g, b, (size_t)H * 8 * 8>>>(out, pool, H, steps / 5, 1.6e-4f, 1e-5f, 0.1f, 0.85f, rk); }));
        report("bootstrap+garch f2 tab x8 2blk/SM", (double)g * 2 * b * (steps / 5) * 5,
               time_best([&] { k_boot2<3><<<g * 2, b, (size_t)H * 8 * 8>>>(out, pool, H, steps / 5, 1.6e-4f, 1e-5f, 0.1f, 0.85f, rk); }));
    }
    // accuracy
    {
        double* err; cudaMalloc(&err, 32); cudaMemset(err, 0, 32);
        const uint32_t n = 1u << 24;
        k_accuracy<<<g_sms * 4, 256>>>(err, n, rk);
        double h[4]; cudaMemcpy(h, err, 32, cudaMemcpyDeviceToHost);
        printf("accuracy over %u calls: max|z - z64| mufu pairs %.3e  poly pairs %.3e   mean %.3e  var %.6f\n", n, h[0], h[1],
               h[2] / (8.0 * n), h[3] / (8.0 * n));
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
