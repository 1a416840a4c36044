// mc_rng.cuh -- counter-based generator and shock transforms, all in registers.
//
// Replaces the reference's per-step `np.random.choice(std_returns, size=n_sims)`
// (app.py:620): instead of a sequential Mersenne Twister feeding an [N] array per
// step through memory, every path derives its own draws from (seed, global path
// id, call index), so no random number ever touches HBM and a path's values do
// not depend on how paths are sharded over GPUs.
//
// Conventions (mirrored bit-for-bit by oracle/philox.py):
//   ctr = (path_lo, path_hi, call, stream), key = (seed_lo, seed_hi)
//   one call -> 4 words -> 8 fp32 normals / 5 fp32 bootstrap indices, or 4 of either in fp64
#pragma once
#include <stdint.h>

namespace mc {

constexpr uint32_t PHILOX_M0 = 0xD2511F53u;
constexpr uint32_t PHILOX_M1 = 0xCD9E8D57u;
constexpr uint32_t PHILOX_W0 = 0x9E3779B9u;
constexpr uint32_t PHILOX_W1 = 0xBB67AE85u;

struct u32x4 { uint32_t x, y, z, w; };

// Philox-4x32-10.  The key schedule depends only on the seed, so the host expands it
// once into the ten round keys (RoundKeys travels as a kernel parameter, i.e. in the
// constant bank, and LOP3 reads it as an operand): per call the per-thread cost is
// exactly 20 IMAD.WIDE + 20 LOP3.
struct RoundKeys { uint32_t k[20]; };

__host__ __device__ __forceinline__ RoundKeys expand_key(uint32_t k0, uint32_t k1) {
    RoundKeys rk;
    for (int r = 0; r < 10; ++r) {
        rk.k[2 * r] = k0;
        rk.k[2 * r + 1] = k1;
        k0 += PHILOX_W0;
        k1 += PHILOX_W1;
    }
    return rk;
}

__device__ __forceinline__ u32x4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               const RoundKeys& rk) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)PHILOX_M0 * c0;
        const uint64_t p1 = (uint64_t)PHILOX_M1 * c2;
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ rk.k[2 * r];
        const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ rk.k[2 * r + 1];
        c1 = (uint32_t)p1;
        c3 = (uint32_t)p0;
        c0 = n0;
        c2 = n2;
    }
    return u32x4{c0, c1, c2, c3};
}

// bits(0x3f800000 | (u >> 9)): one funnel shift, no I2F (which would queue on the
// 16-lane XU pipe next to the MUFU ops).
__device__ __forceinline__ float mantissa_float(uint32_t u) {
    return __uint_as_float(__funnelshift_r(u, 0x7Fu, 9));
}

__device__ __forceinline__ float fast_lg2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_sqrt(float x) {
    float y;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_sin(float x) {
    float y;
    asm("sin.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float fast_cos(float x) {
    float y;
    asm("cos.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

constexpr float TWO_PI_F = 6.28318530717958647692f;
constexpr float THREE_PI_F = 9.42477796076937971538f;
constexpr float LN2_F = 0.69314718055994530942f;

constexpr float BM_RADIUS_SCALE_F = 1.17741002251547469101f;   // sqrt(2 ln 2)

__device__ __forceinline__ void box_muller(uint32_t ua, uint32_t ub, double& z0, double& z1) {
    const double v1 = ((double)ua + 1.0) * 2.3283064365386963e-10;       // (0, 1]
    const double th = 6.283185307179586476925 * ((double)ub * 2.3283064365386963e-10)
                      - 3.141592653589793238463;
    const double r = sqrt(-2.0 * log(v1));
    double s, c;
    sincos(th, &s, &c);
    z0 = r * c;
    z1 = r * s;
}

// ---- fp32: EIGHT normals from one 128-bit Philox call, directions from a shared-memory table ----
// Measured on sm_100a (tools/microbench2.cu, profiles/r01/microbench2_sm100a_v3.txt):
//   * IMAD.WIDE (the Philox multiply) issues at 32 lanes/clk/SM, a quarter of the FP32 rate, and MUFU at
//     16 lanes/clk/SM; with six normals per call and MUFU.SIN/COS (round-1 v2) both pipes and the issue
//     port sat within 15 % of each other and the loop reached 6.0 normals/clk/SM;
//   * one Box-Muller pair per 32-bit word (8 normals per call) cuts the Philox cost per normal by a
//     quarter: 6.6;  a polynomial sin/cos on the FMA pipes does not pay (FFMAs queue behind IMAD.WIDE
//     on the heavy FMA pipe): 6.0 - 6.7;
//   * reading (cos, sin) of the 12-bit direction from a 4096-entry table in shared memory (one LDS.64,
//     random banks, ~6 wavefronts) takes the two MUFU.SIN/COS off the XU pipe altogether: 8.4 - 8.5.
// Word layout:  radius uniform = word[31:12] (20 bits)  v1 = 1 - k 2^-20 in (0, 1],  |z| <= 5.27
//               direction      = word[11:0]  (12 bits)  theta = 2 pi (d / 4096 - 1/2)
// (the radius field sits on top so that ONE funnel shift drops it into a float's mantissa)
// The fields are disjoint, so radius and direction are independent; 4096 equispaced directions integrate
// every smooth functional of (z0, z1) exactly (trapezoid rule on a periodic integrand).
constexpr int DIR_BITS = 12;
constexpr int DIR_COUNT = 1 << DIR_BITS;                 // float2 (cos, sin) entries: 32 KB

__device__ __forceinline__ float radius_uniform20(uint32_t w) {           // (0, 1], multiples of 2^-20
    return fmaf(__uint_as_float(__funnelshift_r(w, 0x3F8u, 12)), -8.0f, 9.0f);     // bits 0x3F800000 | w[31:12]
}
// Box-Muller radius without its constant: rho = sqrt(-lg2 v1), r = sqrt(2 ln 2) rho
__device__ __forceinline__ float rho20(uint32_t w) { return fast_sqrt(-fast_lg2(radius_uniform20(w))); }

// (cos, sin) of the word's direction; lut_base = shared-window byte address of the table
__device__ __forceinline__ void direction_of(uint32_t w, uint32_t lut_base, float& c, float& s) {
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(c), "=f"(s) : "r"(lut_base + ((w << 3) & 0x7FF8u)));
}

// (cos, sin) as one 64-bit register pair, for the packed fp32 instructions of sm_100 (FMUL2 / FFMA2)
__device__ __forceinline__ unsigned long long direction_pair_of(uint32_t w, uint32_t lut_base) {
    unsigned long long cs;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(cs) : "r"(lut_base + ((w << 3) & 0x7FF8u)));
    return cs;
}

// SCALED: eight N(0,1) draws.  !SCALED: the draws divided by sqrt(2 ln 2) -- for callers that fold the constant
// into what multiplies the draws (the portfolio kernel's Cholesky factor)
template <bool SCALED = true>
__device__ __forceinline__ void normals8(const u32x4& w, uint32_t lut_base, float (&z)[8]) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
    float v[4];                     // the four radius uniforms (radius_uniform20) by two packed FMAs
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        unsigned long long b2, m2, k2, u2;
        asm("mov.b64 %0, {%1, %2};" : "=l"(b2) : "r"(__funnelshift_r(ww[2 * h], 0x3F8u, 12)), "r"(__funnelshift_r(ww[2 * h + 1], 0x3F8u, 12)));
        asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(-8.0f));
        asm("mov.b64 %0, {%1, %1};" : "=l"(k2) : "f"(9.0f));
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(u2) : "l"(b2), "l"(m2), "l"(k2));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(v[2 * h]), "=f"(v[2 * h + 1]) : "l"(u2));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float rho = fast_sqrt(-fast_lg2(v[k]));
        const float r = SCALED ? BM_RADIUS_SCALE_F * rho : rho;
        unsigned long long rr, zz;
        asm("mov.b64 %0, {%1, %1};" : "=l"(rr) : "f"(r));
        asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(zz) : "l"(rr), "l"(direction_pair_of(ww[k], lut_base)));   // (r cos, r sin)
        asm("mov.b64 {%0, %1}, %2;" : "=f"(z[2 * k]), "=f"(z[2 * k + 1]) : "l"(zz));
    }
}
// acc + (the eight normals) / sqrt(2 ln 2), cosine terms in the low half and sine terms in the high half of a
// packed pair: one FFMA2 per Box-Muller pair; the caller adds the halves and applies the constant
__device__ __forceinline__ unsigned long long normals8_sum(const u32x4& w, uint32_t lut_base, unsigned long long acc) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
    // the four radius uniforms (radius_uniform20) by two packed FMAs
    float v[4];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        unsigned long long b2, m2, k2, u2;
        asm("mov.b64 %0, {%1, %2};" : "=l"(b2) : "r"(__funnelshift_r(ww[2 * h], 0x3F8u, 12)), "r"(__funnelshift_r(ww[2 * h + 1], 0x3F8u, 12)));
        asm("mov.b64 %0, {%1, %1};" : "=l"(m2) : "f"(-8.0f));
        asm("mov.b64 %0, {%1, %1};" : "=l"(k2) : "f"(9.0f));
        asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(u2) : "l"(b2), "l"(m2), "l"(k2));
        asm("mov.b64 {%0, %1}, %2;" : "=f"(v[2 * h]), "=f"(v[2 * h + 1]) : "l"(u2));
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float r = fast_sqrt(-fast_lg2(v[k]));
        unsigned long long rr;
        asm("mov.b64 %0, {%1, %1};" : "=l"(rr) : "f"(r));
        asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(rr), "l"(direction_pair_of(ww[k], lut_base)));
    }
    return acc;
}

// ---- fp32, full resolution (MC_FLAG_FULL_RES_NORMALS): FOUR normals per call -------------------------
// The eight-per-call normals above are exact to |z| ~ 4 (P(|z| > 4) is 0.4 % low) and thin out beyond: P(|z| > 5)
// is 27 % low and nothing lies beyond 5.27.  Irrelevant for a sum of 252 steps, visible in
// the one-step marginal of a stored path.  Here a pair takes two whole words: the radius from all 32 bits of the
// first (v1 = (u + 1) 2^-32, converted with one I2F: |z| <= 6.66, the float keeps 24 significant bits of v1 -- all
// of them where the tail is), the direction from the top 23 bits of the second through MUFU.SIN / MUFU.COS.
// Five XU operations per pair instead of two: ~6 normals per SM-clock instead of 8 (measured: bench `c2full`).
constexpr int SHOCK_NORMAL_FULL = 3;            // kernel-template value of "MC_SHOCK_NORMAL + the flag"
__device__ __forceinline__ float radius_uniform32(uint32_t w) {            // (0, 1]
    return fmaf(__uint2float_rn(w), 2.3283064365386963e-10f, 2.3283064365386963e-10f);
}
__device__ __forceinline__ void pair_full(uint32_t wr, uint32_t wd, float& rho, float& c, float& s) {
    rho = fast_sqrt(-fast_lg2(radius_uniform32(wr)));                      // r = sqrt(2 ln 2) rho
    const float th = fmaf(mantissa_float(wd), TWO_PI_F, -THREE_PI_F);      // [-pi, pi)
    c = fast_cos(th);
    s = fast_sin(th);
}
__device__ __forceinline__ void normals4_full(const u32x4& w, float (&z)[8]) {
    float rho, c, s;
    pair_full(w.x, w.y, rho, c, s);
    z[0] = BM_RADIUS_SCALE_F * rho * c; z[1] = BM_RADIUS_SCALE_F * rho * s;
    pair_full(w.z, w.w, rho, c, s);
    z[2] = BM_RADIUS_SCALE_F * rho * c; z[3] = BM_RADIUS_SCALE_F * rho * s;
    z[4] = 0.0f; z[5] = 0.0f; z[6] = 0.0f; z[7] = 0.0f;
}

// ---- fp64: four normals per call from full 32-bit uniforms ------------------------------
__device__ __forceinline__ void normals4(const u32x4& w, double (&z)[8]) {
    box_muller(w.x, w.y, z[0], z[1]);
    box_muller(w.z, w.w, z[2], z[3]);
    z[4] = 0.0; z[5] = 0.0; z[6] = 0.0; z[7] = 0.0;
}

template <typename Real> struct NormalsPerCall { static constexpr int value = sizeof(Real) == 4 ? 8 : 4; };
template <bool SCALED = true>
__device__ __forceinline__ void draw_normals(const u32x4& w, uint32_t lut_base, float (&z)[8]) { normals8<SCALED>(w, lut_base, z); }
template <bool SCALED = true>
__device__ __forceinline__ void draw_normals(const u32x4& w, uint32_t, double (&z)[8]) { normals4(w, z); }

// ---- bootstrap indices (the reference's np.random.choice(std_returns, size=n_sims), app.py:620) ----
// fp64 kernels: FOUR per call, idx = floor(word * H / 2^32) by IMAD.HI; non-uniformity H 2^-32.
__device__ __forceinline__ uint32_t bootstrap_index(uint32_t u, uint32_t pool_len) {
    return __umulhi(u, pool_len);
}
// fp32 kernels: FIVE per call from 23-bit fields, idx = floor(field * H / 2^23), computed on the FMA
// pipe (IMAD.HI is quarter rate like IMAD.WIDE and the Philox multiplies already saturate that pipe):
//     g  = bits(0x4B000000 | field)              = 2^23 + field                    (exact)
//     t  = fma(g, H 2^-23, 2^23 - H + 1/2)       = 2^23 + field H 2^-23 + 1/2      (ONE rounding, ulp 1)
//        = 2^23 + idx + 1            (a tie -- field H 2^-23 an exact integer -- rounds to even)
// so bits(t) - 0x4B000000 = idx + 1, and (bits(t) << s) wraps the exponent bits out of a byte offset.
// Fields 0-3 are word[31:9] (one funnel shift builds g); field 4 is the low bytes of words 0, 1, 2 (23 of
// those 24 bits).  Per-index probabilities differ by at most H 2^-23 (1.5e-4 at the reference's H ~ 1250);
// pools longer than 65536 use the fp64 kernels' IMAD.HI path.
constexpr uint32_t BOOT32_MAX_POOL = 65536u;
__device__ __forceinline__ float boot_g(uint32_t word) {                       // fields 0-3
    return __uint_as_float(__funnelshift_r(word, 0x96u, 9));                   // bits 0x4B000000 | word[31:9]
}
__device__ __forceinline__ float boot_g4(const u32x4& w) {                      // field 4
    const uint32_t b = __byte_perm(__byte_perm(w.x, w.y, 0x0040), w.z, 0x7410);   // [w0.b0, w1.b0, w2.b0, -]
    return __uint_as_float((b & 0x007FFFFFu) | 0x4B000000u);
}
// bits(2^23 + floor(field * H / 2^23) + 1) from g = 2^23 + field, scale = H 2^-23, bias = 2^23 - H + 1:
// the FMA is exact inside and ROUNDS DOWN at ulp 1 (2^23 <= result < 2^24), i.e. it takes the floor.  (Round to
// nearest with bias ... + 1/2 is wrong on the ties field * H = k 2^23: half of them land one index low, and
// field = 0 lands on index -1.)
__device__ __forceinline__ uint32_t boot_index_bits(float g, float scale, float bias) {
    return __float_as_uint(__fmaf_rd(g, scale, bias));
}
// two at once: FFMA2.RM (sm_100 packed fp32), each half the same round-down FMA
__device__ __forceinline__ void boot_index_bits2(float g0, float g1, float scale, float bias, uint32_t& t0, uint32_t& t1) {
    unsigned long long g2, s2, b2, r2;
    asm("mov.b64 %0, {%1, %2};" : "=l"(g2) : "f"(g0), "f"(g1));
    asm("mov.b64 %0, {%1, %1};" : "=l"(s2) : "f"(scale));
    asm("mov.b64 %0, {%1, %1};" : "=l"(b2) : "f"(bias));
    asm("fma.rm.f32x2 %0, %1, %2, %3;" : "=l"(r2) : "l"(g2), "l"(s2), "l"(b2));
    asm("mov.b64 {%0, %1}, %2;" : "=r"(t0), "=r"(t1) : "l"(r2));
}

}  // namespace mc
