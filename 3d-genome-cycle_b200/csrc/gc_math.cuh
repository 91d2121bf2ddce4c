// gc_math.cuh -- FP32 device restatements of the closed-form potentials the
// stage drivers use, plus Philox4x32-10 and small helpers.  The reference
// evaluates all of these in FP64 (md::scalar = double, MD/basic_types.hpp:21-28);
// parity is <= 1e-5 relative on per-bead forces (tests/test_gpu_parity.py).
#pragma once

#include <cstdint>

#include <cuda_runtime.h>

namespace gc {

struct f3 {
    float x, y, z;
};

__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return {s * a.x, s * a.y, s * a.z}; }
__device__ __forceinline__ float dot(f3 a, f3 b) { return fmaf(a.x, b.x, fmaf(a.y, b.y, a.z * b.z)); }
__device__ __forceinline__ f3 make_f3(float4 p) { return {p.x, p.y, p.z}; }

// ---- order-preserving float <-> int encodings for atomicMin/Max ----------
__device__ __forceinline__ int float_to_ordered(float f)
{
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __host__ __forceinline__ float ordered_to_float(int i)
{
    int b = i >= 0 ? i : i ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

// ---- pair potentials ------------------------------------------------------
// Pre-digested pair coefficients for one step (core_scale folded in).
struct PairCoef {
    float kA2, kB2;   // 1 / (sigma s)^2
    float cA, cB;     // 6 eps_A kA2, 24 eps_B kB2
    float eA, eB;     // eps_A, eps_B
    float rc2;        // (max sigma s)^2
};

// a * softcore<2,3> + b * softcore<8,3> force factor: F = fac * r_vec
// (MD/potential/softcore_potential.hpp:22-62 with P,Q = 2,3 and 8,3;
//  SIM/stage_interphase/simulation_driver_forcefield.cpp:30-52)
//   A: P Q eps k2 g^2 v,  v = 1,     g = 1 - u2
//   B: P Q eps k2 g^2 v,  v = u2^3,  g = 1 - u2^4
__device__ __forceinline__ float pair_force_factor(const PairCoef &c, float a, float b, float r2)
{
    float uA = r2 * c.kA2;
    float gA = 1.0f - uA;
    float fA = gA < 0.0f ? 0.0f : c.cA * gA * gA;
    float uB = r2 * c.kB2;
    float uB3 = uB * uB * uB;
    float gB = 1.0f - uB3 * uB;
    float fB = gB < 0.0f ? 0.0f : c.cB * gB * gB * uB3;
    return fmaf(a, fA, b * fB);
}

__device__ __forceinline__ float pair_energy(const PairCoef &c, float a, float b, float r2)
{
    float uA = r2 * c.kA2;
    float gA = 1.0f - uA;
    float eA = gA < 0.0f ? 0.0f : c.eA * gA * gA * gA;
    float uB = r2 * c.kB2;
    float uB2 = uB * uB;
    float gB = 1.0f - uB2 * uB2;
    float eB = gB < 0.0f ? 0.0f : c.eB * gB * gB * gB;
    return fmaf(a, eA, b * eB);
}

// softwell<6> with hard cutoff (MD/potential/softwell_potential.hpp:25-51,
// cutoff_potential.hpp:18-53; r2 >= cutoff2 -> 0, exclusive)
__device__ __forceinline__ float well_force_factor(float eps, float k2, float cut2, float r2)
{
    if (r2 >= cut2) return 0.0f;
    float u2 = k2 * r2;
    float u4 = u2 * u2;
    float v1 = 1.0f + u4 * u2;
    return (-6.0f * eps * k2) * u4 / (v1 * v1);
}
__device__ __forceinline__ float well_energy(float eps, float k2, float cut2, float r2)
{
    if (r2 >= cut2) return 0.0f;
    float u2 = k2 * r2;
    return -eps / (1.0f + u2 * u2 * u2);
}

// semispring / harmonic / spring (MD/potential/{semispring,harmonic,spring}_potential.hpp)
// kind: 0 semispring, 1 harmonic, 2 spring.  Returns fac with F = fac * r_vec.
__device__ __forceinline__ float bond_force_factor(int kind, float K, float l, float r2)
{
    if (kind == 1) return -K;
    // rest length 0 (every interphase bond): K * 0 / r - K = -K exactly, and r <= 0 only at
    // r = 0 -- the same bits without the square root and the division
#ifndef GC_NO_BOND_SHORTCUT
    if (l == 0.0f) return r2 > 0.0f ? -K : 0.0f;
#endif
    float r1 = sqrtf(r2);
    if (kind == 0 ? (r1 <= l) : (r1 == 0.0f)) return 0.0f;
    return K * l / r1 - K;
}
__device__ __forceinline__ float bond_energy(int kind, float K, float l, float r2)
{
    if (kind == 1) return 0.5f * K * r2;
    float u = sqrtf(r2) - l;
    if (kind == 0 && u <= 0.0f) return 0.0f;
    return 0.5f * K * u * u;
}

// force_flux_potential (SIM/common/potentials/force_flux_potential.hpp:19-38)
__device__ __forceinline__ float flux_force_factor(float f, float b, float r2)
{
    float r1 = sqrtf(r2);
    float b2 = b * b;
    float den = b2 * r1 + r1 * r2;
    return den == 0.0f ? 0.0f : f * b2 / den;
}
__device__ __forceinline__ float flux_energy(float f, float b, float r2)
{
    return f * b * atan2f(b, sqrtf(r2));
}

// cosine bending (MD/potential/cosine_bending_potential.hpp:23-62).
// role 0/1/2 selects the force on bead i / j / k of the triple.
__device__ __forceinline__ f3 bending_force(float e, f3 rij, f3 rjk, int role)
{
    float dij2 = dot(rij, rij);
    float djk2 = dot(rjk, rjk);
    float dd = sqrtf(dij2 * djk2);
    if (dd == 0.0f) return {0.0f, 0.0f, 0.0f};
    float d = dot(rij, rjk);
    float ed = e / dd;
    f3 fij = ed * (rjk - (d / dij2) * rij);
    f3 fjk = ed * (rij - (d / djk2) * rjk);
    if (role == 0) return fij;
    if (role == 1) return fjk - fij;
    return {-fjk.x, -fjk.y, -fjk.z};
}
__device__ __forceinline__ float bending_energy(float e, f3 rij, f3 rjk)
{
    float dij2 = dot(rij, rij);
    float djk2 = dot(rjk, rjk);
    if (dij2 * djk2 == 0.0f) return 0.0f;
    return e - e * dot(rij, rjk) / sqrtf(dij2 * djk2);
}

// ---- Philox4x32-10 (Salmon et al., SC'11), counter-based ------------------
__device__ __host__ __forceinline__ void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2,
                                                       uint32_t c3, uint32_t k0, uint32_t k1,
                                                       uint32_t out[4])
{
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
    const uint32_t W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; r++) {
        uint64_t p0 = (uint64_t) M0 * c0;
        uint64_t p1 = (uint64_t) M1 * c2;
        uint32_t n0 = (uint32_t) (p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t) p1;
        uint32_t n2 = (uint32_t) (p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t) p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// Three standard normals for (bead, step) of a replica: Box-Muller on the four
// Philox words.  u in (0, 1): 24 random bits + half-ulp offset.
// `epoch` (counter word 3) counts the logical runs of a context: every restart of gc_run
// draws a fresh stream, as the reference's drivers seed every simulate_* call with the
// next output of their mt19937_64 (SIM/stage_interphase/simulation_driver_interphase.cpp:61).
__device__ __forceinline__ f3 normal3(unsigned long long seed, uint32_t bead, long long step,
                                      uint32_t epoch)
{
    uint32_t w[4];
    philox4x32_10(bead, (uint32_t) step, (uint32_t) ((unsigned long long) step >> 32), epoch,
                  (uint32_t) seed, (uint32_t) (seed >> 32), w);
    const float S = 1.0f / 16777216.0f;
    float u0 = ((float) (w[0] >> 8) + 0.5f) * S;
    float u1 = ((float) (w[1] >> 8) + 0.5f) * S;
    float u2 = ((float) (w[2] >> 8) + 0.5f) * S;
    float u3 = ((float) (w[3] >> 8) + 0.5f) * S;
#ifdef GC_SLOW_BM
    {
        float r0 = sqrtf(-2.0f * logf(u0));
        float r1 = sqrtf(-2.0f * logf(u2));
        float s0, c0, s1, c1;
        sincospif(2.0f * u1, &s0, &c0);
        sincospif(2.0f * u3, &s1, &c1);
        (void) s1;
        return {r0 * c0, r0 * s0, r1 * c1};
    }
#endif
    // hardware lg2 / sin / cos (absolute error ~1e-6: far below the sampling
    // noise of any observable; the stream's moments and KS statistic are tested)
    const float NEG_2LN2 = -1.3862943611198906f;
    float r0 = sqrtf(fmaxf(NEG_2LN2 * __log2f(u0), 0.0f));
    float r1 = sqrtf(fmaxf(NEG_2LN2 * __log2f(u2), 0.0f));
    const float TWO_PI = 6.283185307179586f;
    float a0 = TWO_PI * (u1 - 0.5f), a1 = TWO_PI * (u3 - 0.5f);   // (-pi, pi): no range reduction
    return {r0 * __cosf(a0), r0 * __sinf(a0), r1 * __cosf(a1)};
}

} // namespace gc
