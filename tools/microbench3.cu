// microbench3.cu -- integer-multiply issue rates on sm_100a: is there a cheaper way to get the two halves
// of a 32x32 product than IMAD.WIDE (the instruction that binds Philox-4x32-10)?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench3 microbench3.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

constexpr int ITERS = 4096;
constexpr int ILP = 8;
constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
static int g_sms, g_khz;
// x*M >> 32 for 32-bit x: fma.rz(2^52 + x, M*2^-16, 2^68 - 2^36*M) = 2^68 + floor(x*M / 2^32) * 2^16 exactly
// (the product is exact inside the FMA; rounding toward zero at ulp 2^16 is the floor), so the low mantissa
// word of the result is the high word of the product.
#define MD(M) ((double)(M) * 0x1p-16)
#define AD(M) (0x1p68 - (double)(M) * 0x1p36)
constexpr double MD0 = MD(0xD2511F53u), AD0 = AD(0xD2511F53u), MD1 = MD(0xCD9E8D57u), AD1 = AD(0xCD9E8D57u);

enum { OP_WIDE, OP_HI, OP_LO, OP_HI_LO, OP_MAD_LO, OP_WIDE16, OP_MUL24, OP_DFMA, OP_WIDE_MAD, OP_LOP3, OP_IADD3, OP_SHF,
       OP_WIDE_PLUS_LO, OP_WIDE_PLUS_FFMA_IMM, OP_WIDE_PLUS_FMUL,
       OP_WIDE_PLUS_DFMA, OP_LO_PLUS_DFMA, OP_DFMA_HI };

template <int OP>
__global__ void k_op(uint32_t* out, uint32_t seed) {
    uint32_t y[ILP], z[ILP];
    double d[ILP];
    float f[ILP];
    for (int i = 0; i < ILP; ++i) { y[i] = seed + i + threadIdx.x; z[i] = seed * 3 + i; d[i] = 1.0 + y[i]; f[i] = 1.0f + i; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < ILP; ++i) {
            if (OP == OP_WIDE) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p; mul.wide.u32 p, %2, %3; mov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
            } else if (OP == OP_HI) {
                asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(y[i]) : "r"(y[i]), "r"(M0));
            } else if (OP == OP_LO) {
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(y[i]) : "r"(y[i]), "r"(M0));
            } else if (OP == OP_HI_LO) {
                uint32_t lo, hi;
                asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(hi) : "r"(y[i]), "r"(M0));
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(y[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
            } else if (OP == OP_MAD_LO) {
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(y[i]) : "r"(y[i]), "r"(M0), "r"(z[i]));
            } else if (OP == OP_WIDE16) {
                asm volatile("{.reg .u16 h; cvt.u16.u32 h, %1; mul.wide.u16 %0, h, 0x8D57;}" : "=r"(y[i]) : "r"(y[i]));
            } else if (OP == OP_MUL24) {
                asm volatile("mul24.lo.u32 %0, %1, %2;" : "=r"(y[i]) : "r"(y[i]), "r"(0x511F53u));
            } else if (OP == OP_DFMA) {
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            } else if (OP == OP_WIDE_MAD) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p, q; mov.b64 q, {%3,%3}; mad.wide.u32 p, %2, %4, q; mov.b64 {%0,%1}, p;}"
                             : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(z[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
            } else if (OP == OP_LOP3) {
                asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[i]) : "r"(z[i]), "r"(seed));
            } else if (OP == OP_IADD3) {
                asm volatile("add.u32 %0, %0, %1;" : "+r"(y[i]) : "r"(z[i]));
            } else if (OP == OP_SHF) {
                asm volatile("shf.r.wrap.b32 %0, %0, %1, 9;" : "+r"(y[i]) : "r"(z[i]));
            } else if (OP == OP_WIDE_PLUS_LO) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p; mul.wide.u32 p, %2, %3; mov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(M0));
                y[i] = hi;
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(z[i]) : "r"(z[i]), "r"(M1), "r"(lo));
            } else if (OP == OP_WIDE_PLUS_FFMA_IMM) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p; mul.wide.u32 p, %2, %3; mov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
                asm volatile("fma.rn.f32 %0, %0, 0f3F800100, 0f3A000000;" : "+f"(f[i]));
            } else if (OP == OP_WIDE_PLUS_DFMA) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p; mul.wide.u32 p, %2, %3; mov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            } else if (OP == OP_LO_PLUS_DFMA) {
                asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(y[i]) : "r"(y[i]), "r"(M0), "r"(z[i]));
                asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[i]) : "d"(1.0000001), "d"(0.5));
            } else if (OP == OP_DFMA_HI) {
                // both halves of y*M0: high word by one DFMA.RZ, low word by IMAD
                const double xd = __hiloint2double(0x43300000, (int)y[i]);
                double r;
                asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(xd), "d"(MD0), "d"(AD0));
                uint32_t lo;
                asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(y[i]), "r"(M0));
                y[i] = (uint32_t)__double2loint(r) ^ z[i]; z[i] = lo;
            } else if (OP == OP_WIDE_PLUS_FMUL) {
                uint32_t lo, hi;
                asm volatile("{.reg .u64 p; mul.wide.u32 p, %2, %3; mov.b64 {%0,%1}, p;}" : "=r"(lo), "=r"(hi) : "r"(y[i]), "r"(M0));
                y[i] = hi; z[i] = lo;
                asm volatile("mul.rn.f32 %0, %0, 0f3F800100;" : "+f"(f[i]));
            }
        }
    }
    uint32_t s = 0; double ds = 0; float fs = 0;
    for (int i = 0; i < ILP; ++i) { s += y[i] ^ z[i]; ds += d[i]; fs += f[i]; }
    if (s == 123456u || ds == 1.25 || fs == 1.25f) out[0] = s;
}

// Philox-4x32-10, products by IMAD.WIDE (W=0) or IMAD.HI + IMAD (W=1)
template <int W>
__device__ __forceinline__ void mulhilo(uint32_t a, uint32_t m, uint32_t& hi, uint32_t& lo) {
    if (W == 0) {
        const uint64_t p = (uint64_t)a * m; hi = (uint32_t)(p >> 32); lo = (uint32_t)p;
    } else if (W == 1) {
        asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(hi) : "r"(a), "r"(m));
        asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(lo) : "r"(a), "r"(m));
    } else {
        const double xd = __hiloint2double(0x43300000, (int)a);
        double r;
        if (m == M0) asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(xd), "d"(MD0), "d"(AD0));
        else         asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(r) : "d"(xd), "d"(MD1), "d"(AD1));
        hi = (uint32_t)__double2loint(r);
        lo = a * m;
    }
}
// which of the 20 multiplies of a call go through the FP64 pipe: PAT 0 none, 1 all, 2 the M0 ones, 3 two of three
template <int PAT> __device__ __forceinline__ constexpr int use_dfma(int r, int which) {
    return PAT == 0 ? 0 : PAT == 1 ? 1 : PAT == 2 ? (which == 0) : PAT == 3 ? ((2 * r + which) % 3 != 0) : ((r + which) % 2 == 0);
}
template <int PAT>
__global__ void k_philox_mix(uint32_t* out, uint32_t seed, uint32_t* check) {
    uint32_t acc = 0;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int it = 0; it < ITERS / 4; ++it) {
        uint32_t c0 = tid * 2654435761u + it, c1 = 0, c2 = it * 40503u + tid, c3 = 0, k0 = seed, k1 = seed ^ 0x5555u;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t h0, l0, h1, l1;
            if (use_dfma<PAT>(r, 0)) mulhilo<2>(c0, M0, h0, l0); else mulhilo<0>(c0, M0, h0, l0);
            if (use_dfma<PAT>(r, 1)) mulhilo<2>(c2, M1, h1, l1); else mulhilo<0>(c2, M1, h1, l1);
            c0 = h1 ^ c1 ^ k0; c1 = l1; c2 = h0 ^ c3 ^ k1; c3 = l0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        acc ^= c0 ^ c1 ^ c2 ^ c3;
    }
    if (check) atomicXor(check, acc);
    if (acc == 123456u) out[0] = acc;
}
template <int W>
__global__ void k_philox(uint32_t* out, uint32_t seed) {
    uint32_t acc = 0;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int it = 0; it < ITERS / 4; ++it) {
        uint32_t c0 = tid, c1 = 0, c2 = it, c3 = 0, k0 = seed, k1 = seed ^ 0x5555u;
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            uint32_t h0, l0, h1, l1;
            mulhilo<W>(c0, M0, h0, l0);
            mulhilo<W>(c2, M1, h1, l1);
            c0 = h1 ^ c1 ^ k0; c1 = l1; c2 = h0 ^ c3 ^ k1; c3 = l0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        acc ^= c0 ^ c1 ^ c2 ^ c3;
    }
    if (acc == 123456u) out[0] = acc;
}

// Philox with the words that feed the multiplies carried as doubles (c, 0x44300000) = 2^68 + c*2^16: the high word
// of the pair survives every step (fma.rz(D, M*2^-32, 2^68 - 2^36*M) = 2^68 + hi32(c*M)*2^16), so the xor of the next
// round can be written onto the low word in place and no move is needed to form the next operand.
constexpr double MS0 = (double)0xD2511F53u * 0x1p-32, MS1 = (double)0xCD9E8D57u * 0x1p-32;
template <int PAT>
__global__ void k_philox_carry(uint32_t* out, uint32_t seed, uint32_t* check) {
    uint32_t acc = 0;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int it = 0; it < ITERS / 4; ++it) {
        uint32_t c1 = 0, c3 = 0, k0 = seed, k1 = seed ^ 0x5555u;
        double d0 = __hiloint2double(0x44300000, (int)(tid * 2654435761u + it));
        double d2 = __hiloint2double(0x44300000, (int)(it * 40503u + tid));
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            // the xor is applied in place to the low word of the FMA result (same virtual register in and out)
            double q0, q1; uint32_t l0, l1;
            const uint32_t c0 = (uint32_t)__double2loint(d0), c2 = (uint32_t)__double2loint(d2);
            if (use_dfma<PAT>(r, 0)) {
                asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(q0) : "d"(d0), "d"(MS0), "d"(AD0)); l0 = c0 * M0;
            } else { uint32_t h; mulhilo<0>(c0, M0, h, l0); q0 = __hiloint2double(0x44300000, (int)h); }
            if (use_dfma<PAT>(r, 1)) {
                asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(q1) : "d"(d2), "d"(MS1), "d"(AD1)); l1 = c2 * M1;
            } else { uint32_t h; mulhilo<0>(c2, M1, h, l1); q1 = __hiloint2double(0x44300000, (int)h); }
            asm volatile("{ .reg .b32 lo, hi; mov.b64 {lo,hi}, %0; lop3.b32 lo, lo, %1, %2, 0x96; mov.b64 %0, {lo,hi}; }" : "+d"(q1) : "r"(c1), "r"(k0));
            asm volatile("{ .reg .b32 lo, hi; mov.b64 {lo,hi}, %0; lop3.b32 lo, lo, %1, %2, 0x96; mov.b64 %0, {lo,hi}; }" : "+d"(q0) : "r"(c3), "r"(k1));
            d0 = q1; c1 = l1; d2 = q0; c3 = l0;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        acc ^= (uint32_t)__double2loint(d0) ^ c1 ^ (uint32_t)__double2loint(d2) ^ c3;
    }
    if (check) atomicXor(check, acc);
    if (acc == 123456u) out[0] = acc;
}

// one of the two interleaved flows of Philox on the FP64 pipe (carried as a double), the other on IMAD.WIDE (plain word)
#define XOR_LOW_IN_PLACE(q, a, b) asm volatile("{ .reg .b32 lo, hi; mov.b64 {lo,hi}, %0; lop3.b32 lo, lo, %1, %2, 0x96; mov.b64 %0, {lo,hi}; }" : "+d"(q) : "r"(a), "r"(b))
__global__ void k_philox_flow(uint32_t* out, uint32_t seed, uint32_t* check) {
    uint32_t acc = 0;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    for (int it = 0; it < ITERS / 4; ++it) {
        uint32_t c1 = 0, c3 = 0, k0 = seed, k1 = seed ^ 0x5555u;
        double dA = __hiloint2double(0x44300000, (int)(tid * 2654435761u + it));   // word 0
        uint32_t xB = it * 40503u + tid;                                              // word 2
#pragma unroll
        for (int r = 0; r < 10; ++r) {
            const uint32_t xA = (uint32_t)__double2loint(dA);
            double qA; uint32_t lA, hB, lB;
            if (r % 2 == 0) {      // A holds word 0 (x M0), B holds word 2 (x M1)
                asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(qA) : "d"(dA), "d"(MS0), "d"(AD0)); lA = xA * M0;
                mulhilo<0>(xB, M1, hB, lB);
                xB = hB ^ c1 ^ k0; c1 = lB;                    // new word 0 <- B
                XOR_LOW_IN_PLACE(qA, c3, k1); c3 = lA;         // new word 2 <- A
            } else {               // B holds word 0 (x M0), A holds word 2 (x M1)
                asm volatile("fma.rz.f64 %0, %1, %2, %3;" : "=d"(qA) : "d"(dA), "d"(MS1), "d"(AD1)); lA = xA * M1;
                mulhilo<0>(xB, M0, hB, lB);
                XOR_LOW_IN_PLACE(qA, c1, k0); c1 = lA;         // new word 0 <- A
                xB = hB ^ c3 ^ k1; c3 = lB;                    // new word 2 <- B
            }
            dA = qA;
            k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
        }
        // after 10 rounds A holds word 0 again
        acc ^= (uint32_t)__double2loint(dA) ^ c1 ^ xB ^ c3;
    }
    if (check) atomicXor(check, acc);
    if (acc == 123456u) out[0] = acc;
}

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
#define RUN(name, OP) report(name, thr * ITERS * ILP, time_best([&] { k_op<OP><<<grid, block>>>(out, 7u); }))

int main() {
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    g_sms = prop.multiProcessorCount;
    cudaDeviceGetAttribute(&g_khz, cudaDevAttrClockRate, 0);
    uint32_t* out; cudaMalloc(&out, 64);
    const int grid = g_sms * 8, block = 256;
    const double thr = (double)grid * block;
    RUN("mul.wide.u32 (IMAD.WIDE)", OP_WIDE);
    RUN("mul.hi.u32 (IMAD.HI)", OP_HI);
    RUN("mul.lo.u32 (IMAD)", OP_LO);
    RUN("mul.hi + mul.lo", OP_HI_LO);
    RUN("mad.lo.u32", OP_MAD_LO);
    RUN("mul.wide.u16", OP_WIDE16);
    RUN("mul24.lo", OP_MUL24);
    RUN("fma.f64 (DFMA)", OP_DFMA);
    RUN("mad.wide.u32", OP_WIDE_MAD);
    RUN("lop3", OP_LOP3);
    RUN("add.u32", OP_IADD3);
    RUN("shf", OP_SHF);
    RUN("mul.wide + mad.lo (pairs)", OP_WIDE_PLUS_LO);
    RUN("mul.wide + ffma imm (pairs)", OP_WIDE_PLUS_FFMA_IMM);
    RUN("mul.wide + fmul imm (pairs)", OP_WIDE_PLUS_FMUL);
    report("philox call, IMAD.WIDE", thr * (ITERS / 4), time_best([&] { k_philox<0><<<grid, block>>>(out, 7u); }));
    report("philox call, IMAD.HI + IMAD", thr * (ITERS / 4), time_best([&] { k_philox<1><<<grid, block>>>(out, 7u); }));
    RUN("mul.wide + dfma (pairs)", OP_WIDE_PLUS_DFMA);
    RUN("mad.lo + dfma (pairs)", OP_LO_PLUS_DFMA);
    RUN("hi by dfma.rz + lo by imad", OP_DFMA_HI);
    uint32_t* chk; cudaMalloc(&chk, 16); cudaMemset(chk, 0, 16);
    report("philox mix: none on fp64", thr * (ITERS / 4), time_best([&] { k_philox_mix<0><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox mix: all on fp64", thr * (ITERS / 4), time_best([&] { k_philox_mix<1><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox mix: M0 on fp64", thr * (ITERS / 4), time_best([&] { k_philox_mix<2><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox mix: 2 of 3 on fp64", thr * (ITERS / 4), time_best([&] { k_philox_mix<3><<<grid, block>>>(out, 7u, nullptr); }));
    k_philox_mix<0><<<grid, block>>>(out, 7u, chk + 0); k_philox_mix<1><<<grid, block>>>(out, 7u, chk + 1);
    k_philox_mix<2><<<grid, block>>>(out, 7u, chk + 2); k_philox_mix<3><<<grid, block>>>(out, 7u, chk + 3);
    report("philox carry: all on fp64", thr * (ITERS / 4), time_best([&] { k_philox_carry<1><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox carry: M0 on fp64", thr * (ITERS / 4), time_best([&] { k_philox_carry<2><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox carry: 2 of 3 on fp64", thr * (ITERS / 4), time_best([&] { k_philox_carry<3><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox carry: one flow on fp64", thr * (ITERS / 4), time_best([&] { k_philox_carry<4><<<grid, block>>>(out, 7u, nullptr); }));
    report("philox flow: A fp64, B imad.wide", thr * (ITERS / 4), time_best([&] { k_philox_flow<<<grid, block>>>(out, 7u, nullptr); }));
    cudaMemset(chk, 0, 16);
    k_philox_mix<0><<<grid, block>>>(out, 7u, chk + 0); k_philox_carry<1><<<grid, block>>>(out, 7u, chk + 1);
    k_philox_carry<4><<<grid, block>>>(out, 7u, chk + 2); k_philox_flow<<<grid, block>>>(out, 7u, chk + 3);
    uint32_t h[4]; cudaMemcpy(h, chk, 16, cudaMemcpyDeviceToHost);
    printf("xor of all outputs: %08x %08x %08x %08x (%s)\n", h[0], h[1], h[2], h[3],
           (h[0] == h[1] && h[0] == h[2] && h[0] == h[3]) ? "identical" : "DIFFERENT");
    printf("status: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
