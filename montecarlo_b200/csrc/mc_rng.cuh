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
//   one call -> 4 words -> 4 normals (steps 4j..4j+3) or 4 bootstrap indices
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

// Box-Muller radius with the factor sqrt(2 ln 2) left out: rho = sqrt(-lg2(v1)),
// r = sqrt(2 ln2) * rho.  Callers that only need sigma * sum(z) fold the constant
// into sigma and save one multiply per pair.
__device__ __forceinline__ float bm_rho(uint32_t ua) {
    const float v1 = 2.0f - mantissa_float(ua);       // (0, 1]
    return fast_sqrt(-fast_lg2(v1));
}
__device__ __forceinline__ float bm_theta(uint32_t ub) {
    return fmaf(mantissa_float(ub), TWO_PI_F, -THREE_PI_F);   // [-pi, pi)
}
constexpr float BM_RADIUS_SCALE_F = 1.17741002251547469101f;   // sqrt(2 ln 2)

__device__ __forceinline__ void box_muller(uint32_t ua, uint32_t ub, float& z0, float& z1) {
    const float r = BM_RADIUS_SCALE_F * bm_rho(ua);
    const float th = bm_theta(ub);
    z0 = r * fast_cos(th);
    z1 = r * fast_sin(th);
}

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

// ---- fp32: SIX normals from one 128-bit Philox call -------------------------------------
// The Philox multiply (IMAD.WIDE) runs at a quarter of the FP32 rate on sm_100a (measured:
// tools/microbench.cu), so the generator -- not the transform -- bounds the fused kernels.
// A float only holds 23 mantissa bits, so 4 x 32 bits are cut into three Box-Muller pairs
// instead of two:   radius uniform k = word_k[22:0]            (23 bits, k = 0,1,2)
//                   angle A = word_3[31:9]                      (23 bits)
//                   angle B = word_0[31:24] : word_1[31:24]     (16 bits)
//                   angle C = word_2[31:24] : word_3[7:0]       (16 bits)
// All fields are disjoint, so the six uniforms are independent.  (The low mantissa bits of the
// 16-bit angles repeat the field's own top byte: 65,536 distinct, near-equispaced angles.)
__device__ __forceinline__ float radius_uniform(uint32_t w) {           // (0, 1]
    return 2.0f - __uint_as_float((w & 0x007FFFFFu) | 0x3F800000u);
}
__device__ __forceinline__ float rho_of(uint32_t w) { return fast_sqrt(-fast_lg2(radius_uniform(w))); }
__device__ __forceinline__ float theta_of(uint32_t field_at_top) {       // [-pi, pi)
    return fmaf(mantissa_float(field_at_top), TWO_PI_F, -THREE_PI_F);
}
__device__ __forceinline__ uint32_t angle_b_bits(const u32x4& w) { return __byte_perm(w.x, w.y, 0x3733); }
__device__ __forceinline__ uint32_t angle_c_bits(const u32x4& w) { return __byte_perm(w.z, w.w, 0x3433); }

__device__ __forceinline__ void normals6(const u32x4& w, float (&z)[6]) {
    const float ra = BM_RADIUS_SCALE_F * rho_of(w.x), rb = BM_RADIUS_SCALE_F * rho_of(w.y),
                rc = BM_RADIUS_SCALE_F * rho_of(w.z);
    const float ta = theta_of(w.w), tb = theta_of(angle_b_bits(w)), tc = theta_of(angle_c_bits(w));
    z[0] = ra * fast_cos(ta); z[1] = ra * fast_sin(ta);
    z[2] = rb * fast_cos(tb); z[3] = rb * fast_sin(tb);
    z[4] = rc * fast_cos(tc); z[5] = rc * fast_sin(tc);
}
// sum of the six normals / sqrt(2 ln 2), one FFMA per normal (constant folded in by the caller)
__device__ __forceinline__ float normals6_sum(const u32x4& w, float s) {
    const float ra = rho_of(w.x), rb = rho_of(w.y), rc = rho_of(w.z);
    const float ta = theta_of(w.w), tb = theta_of(angle_b_bits(w)), tc = theta_of(angle_c_bits(w));
    s = fmaf(ra, fast_cos(ta), s); s = fmaf(ra, fast_sin(ta), s);
    s = fmaf(rb, fast_cos(tb), s); s = fmaf(rb, fast_sin(tb), s);
    s = fmaf(rc, fast_cos(tc), s); s = fmaf(rc, fast_sin(tc), s);
    return s;
}

// ---- fp64: four normals per call from full 32-bit uniforms ------------------------------
__device__ __forceinline__ void normals4(const u32x4& w, double (&z)[6]) {
    box_muller(w.x, w.y, z[0], z[1]);
    box_muller(w.z, w.w, z[2], z[3]);
    z[4] = 0.0; z[5] = 0.0;
}

template <typename Real> struct NormalsPerCall { static constexpr int value = sizeof(Real) == 4 ? 6 : 4; };
__device__ __forceinline__ void draw_normals(const u32x4& w, float (&z)[6]) { normals6(w, z); }
__device__ __forceinline__ void draw_normals(const u32x4& w, double (&z)[6]) { normals4(w, z); }

// bootstrap index: floor(u * H / 2^32), exact, unbiased to 2^-32 * H
__device__ __forceinline__ uint32_t bootstrap_index(uint32_t u, uint32_t pool_len) {
    return __umulhi(u, pool_len);
}

}  // namespace mc
