// rng_variants.cuh -- shock-transform variants that were measured and NOT adopted (tools/microbench*.cu
// time them against the product's mc_rng.cuh): six normals per Philox call with MUFU sin/cos (round-1 v2),
// eight per call with MUFU or polynomial sin/cos.  Results: profiles/r01/microbench2_sm100a_v3.txt.
#pragma once
#include "../montecarlo_b200/csrc/mc_rng.cuh"

namespace mc {

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

__device__ __forceinline__ void box_muller(uint32_t ua, uint32_t ub, float& z0, float& z1) {
    const float r = BM_RADIUS_SCALE_F * bm_rho(ua);
    const float th = bm_theta(ub);
    z0 = r * fast_cos(th);
    z1 = r * fast_sin(th);
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

// ---- fp32: EIGHT normals from one 128-bit Philox call -----------------------------------
// Measured on sm_100a (tools/microbench2.cu): with six normals per call the 16-lane XU pipe
// (4 MUFU per pair), the quarter-rate IMAD.WIDE of Philox and the issue port are all within
// 15 % of each other, and the kernel reaches ~71 % of the XU bound.  Cutting every 32-bit word
// into ONE Box-Muller pair lowers the Philox cost per normal by a quarter and leaves the XU
// pipe as the single limiter:
//     radius uniform  = word[19:0]   20 bits  ->  v1 = 1 - k * 2^-20 in (0, 1]   (|z| <= 5.27)
//     angle           = word[31:20]  12 bits  ->  4096 equispaced directions; the float's lower
//                       mantissa bits carry word[19:9] along, which only shifts the 4096-point
//                       grid by a radius-dependent sub-step (for every radius the directions stay
//                       equispaced, so the trapezoid sum over them is exact for smooth functionals)
// PP of the four pairs (compile-time) take sin/cos from a polynomial on the FMA pipes instead
// of MUFU.SIN/COS: sin(pi t), cos(pi t) for the half angle t in [-1/2, 1/2) followed by the
// double-angle identities, max abs error 5e-7 (the MUFU pair: 2^-20.9 = 5e-7).
// cos(2 pi t), sin(2 pi t) for t = m - 1.5, m in [1, 2)
__device__ __forceinline__ void sincos_turn_mufu(float m, float& c, float& s) {
    const float th = fmaf(m, TWO_PI_F, -THREE_PI_F);
    c = fast_cos(th);
    s = fast_sin(th);
}
__device__ __forceinline__ void sincos_turn_poly(float m, float& c, float& s) {
    const float t = m - 1.5f, u = t * t;
    // sqrt(2) sin(pi t) = t * P(u),  sqrt(2) cos(pi t) = Q(u)   (Chebyshev fits on u in [0, 1/4])
    float p = fmaf(u, 0.10982452f, -0.84611164f);
    p = fmaf(u, p, 3.60635410f);
    p = fmaf(u, p, -7.30824600f);
    p = fmaf(u, p, 4.44288306f);
    float q = fmaf(u, 0.31068802f, -1.88355783f);
    q = fmaf(u, q, 5.73946005f);
    q = fmaf(u, q, -6.97885118f);
    q = fmaf(u, q, 1.41421348f);
    const float sh = p * t;
    c = fmaf(-sh, sh, 1.0f);          // 1 - 2 sin^2(pi t)
    s = sh * q;                       // 2 sin(pi t) cos(pi t)
}

template <int PP>
__device__ __forceinline__ void normals8(const u32x4& w, float (&z)[8]) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float r = BM_RADIUS_SCALE_F * rho20(ww[k]);
        float c, s;
        if (k < PP) sincos_turn_poly(mantissa_float(ww[k]), c, s);
        else sincos_turn_mufu(mantissa_float(ww[k]), c, s);
        z[2 * k] = r * c;
        z[2 * k + 1] = r * s;
    }
}
// sum of the eight normals / sqrt(2 ln 2), one FFMA per normal (constant folded in by the caller)
template <int PP>
__device__ __forceinline__ float normals8_sum(const u32x4& w, float acc) {
    const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const float r = rho20(ww[k]);
        float c, s;
        if (k < PP) sincos_turn_poly(mantissa_float(ww[k]), c, s);
        else sincos_turn_mufu(mantissa_float(ww[k]), c, s);
        acc = fmaf(r, c, acc);
        acc = fmaf(r, s, acc);
    }
    return acc;
}


}  // namespace mc
