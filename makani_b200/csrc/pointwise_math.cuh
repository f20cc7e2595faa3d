// GELU arithmetic of the fused instance-norm kernels, host + device (the host build is what
// tests/cpp/gelu_host_test.cpp checks against std::erf in double precision).
//
// libdevice's erff (a branchy ~40-instruction routine) makes the GELU kernels issue-bound; pw_erf_fast is the rational
// approximation 7.1.26 of Abramowitz & Stegun,
//     erf(x) = 1 - (a1 t + a2 t^2 + a3 t^3 + a4 t^4 + a5 t^5) exp(-x^2),  t = 1 / (1 + p x),  x >= 0,
// |error| <= 1.5e-7 in exact arithmetic (about 3e-7 in fp32, below one bf16 / fp16 ulp of any GELU value by orders of
// magnitude and inside the fp32 tolerance of the parity tests): one reciprocal, one exp2, six FMAs, branch-free, and
// the exp(-x^2) it computes is the Gaussian the derivative needs anyway.
// Measured on B200 (round 2, profiles/r2_gelu_erf_ab.txt; 384 x 721 x 1440 bf16, fraction of the HBM peak): norm + GELU
// 0.59 -> 0.64, its backward statistics 0.44 -> 0.57, backward apply 0.57 -> 0.63, bias + GELU 0.58 -> 0.64, its
// backward 0.54 -> 0.62; all GPU parity tests pass with it (fp32 I/O at 2e-6).  It is therefore the default;
// -DSFNO_FAST_ERF=0 selects erff (kept for the host accuracy test, which checks both).
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define PW_HD __host__ __device__ __forceinline__
#else
#define PW_HD inline
#endif

#ifndef SFNO_FAST_ERF
#define SFNO_FAST_ERF 1
#endif

namespace sfno {

// returns erf(x); *gauss = exp(-x^2)
PW_HD float pw_erf_fast(float x, float* gauss) {
  const float ax = fabsf(x);
#if defined(__CUDA_ARCH__)
  // MUFU.RCP / MUFU.EX2 (1 - 2 ulp each) instead of the IEEE division and exp with their slow-path branches; the
  // denominator lies in [1, inf) and the exponent in (-inf, 0], so neither needs range fix-ups
  const float t = __fdividef(1.0f, fmaf(0.3275911f, ax, 1.0f));
  const float e = __expf(-ax * ax);
#else
  const float t = 1.0f / fmaf(0.3275911f, ax, 1.0f);
  const float e = expf(-ax * ax);
#endif
  float poly = 1.061405429f;
  poly = fmaf(poly, t, -1.453152027f);
  poly = fmaf(poly, t, 1.421413741f);
  poly = fmaf(poly, t, -0.284496736f);
  poly = fmaf(poly, t, 0.254829592f);
  const float r = 1.0f - poly * t * e;
  *gauss = e;
  return copysignf(r, x);
}

PW_HD float pw_gelu(float z) {
#if SFNO_FAST_ERF
  float g;
  return 0.5f * z * (1.f + pw_erf_fast(z * 0.70710678118654752f, &g));
#else
  return 0.5f * z * (1.f + erff(z * 0.70710678118654752f));
#endif
}

// d/dz [ z Phi(z) ] = Phi(z) + z phi(z)
PW_HD float pw_gelu_grad(float z) {
#if SFNO_FAST_ERF
  float g;                                                     // g = exp(-z^2 / 2)
  const float er = pw_erf_fast(z * 0.70710678118654752f, &g);
  return fmaf(z * 0.3989422804014327f, g, 0.5f * (1.f + er));
#else
#if defined(__CUDA_ARCH__)
  return 0.5f * (1.f + erff(z * 0.70710678118654752f)) + z * 0.3989422804014327f * __expf(-0.5f * z * z);
#else
  return 0.5f * (1.f + erff(z * 0.70710678118654752f)) + z * 0.3989422804014327f * expf(-0.5f * z * z);
#endif
#endif
}

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ >= 1000) && SFNO_FAST_ERF
// Two elements at a time with the packed fp32 instructions of sm_100 (FMUL2 / FFMA2): the GELU kernels are bound by
// issue slots (~12 instructions per bf16 element is what an SM issues at the HBM rate), and the polynomial, the scalings
// and the final products are the same operation on both elements.  Reciprocal, exp2 and the sign transfer stay scalar.
// Same arithmetic as pw_erf_fast (the packed forms round like their scalar counterparts).
#define SFNO_PW_PACKED 1
namespace pw2 {
__device__ __forceinline__ unsigned long long pk(float x, float y) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y));
  return r;
}
__device__ __forceinline__ void up(unsigned long long v, float& x, float& y) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v));
}
__device__ __forceinline__ unsigned long long mul(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma(unsigned long long a, unsigned long long b, unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ unsigned long long bc(float s) { return pk(s, s); }
__device__ __forceinline__ float ex2(float x) {                      // MUFU.EX2, what __expf uses after its multiply
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// erf of (u0, u1) -> er; g = exp(-u^2) of both
__device__ __forceinline__ void erf_pair(float u0, float u1, unsigned long long& er, unsigned long long& g) {
  const float a0 = fabsf(u0), a1 = fabsf(u1);
  const unsigned long long ax = pk(a0, a1);
  float d0, d1;
  up(fma(ax, bc(0.3275911f), bc(1.0f)), d0, d1);
  const unsigned long long t = pk(__fdividef(1.0f, d0), __fdividef(1.0f, d1));
  float m0, m1;
  up(mul(mul(ax, ax), bc(-1.4426950408889634f)), m0, m1);            // -x^2 log2(e)
  g = pk(ex2(m0), ex2(m1));
  unsigned long long poly = bc(1.061405429f);
  poly = fma(poly, t, bc(-1.453152027f));
  poly = fma(poly, t, bc(1.421413741f));
  poly = fma(poly, t, bc(-0.284496736f));
  poly = fma(poly, t, bc(0.254829592f));
  float r0, r1;
  up(fma(mul(mul(poly, t), g), bc(-1.0f), bc(1.0f)), r0, r1);       // 1 - poly t e
  er = pk(copysignf(r0, u0), copysignf(r1, u1));
}
}  // namespace pw2

// z0, z1 -> gelu(z0), gelu(z1)
__device__ __forceinline__ void pw_gelu_pair(float& z0, float& z1) {
  const unsigned long long z = pw2::pk(z0, z1);
  float u0, u1;
  pw2::up(pw2::mul(z, pw2::bc(0.70710678118654752f)), u0, u1);
  unsigned long long er, g;
  pw2::erf_pair(u0, u1, er, g);
  pw2::up(pw2::mul(z, pw2::fma(er, pw2::bc(0.5f), pw2::bc(0.5f))), z0, z1);
}
// z0, z1 -> gelu'(z0), gelu'(z1) = Phi(z) + z phi(z)
__device__ __forceinline__ void pw_gelu_grad_pair(float& z0, float& z1) {
  const unsigned long long z = pw2::pk(z0, z1);
  float u0, u1;
  pw2::up(pw2::mul(z, pw2::bc(0.70710678118654752f)), u0, u1);
  unsigned long long er, g;                                            // g = exp(-z^2 / 2)
  pw2::erf_pair(u0, u1, er, g);
  const unsigned long long cdf = pw2::fma(er, pw2::bc(0.5f), pw2::bc(0.5f));
  pw2::up(pw2::fma(pw2::mul(z, pw2::bc(0.3989422804014327f)), g, cdf), z0, z1);
}
#else
#define SFNO_PW_PACKED 0
#endif

}  // namespace sfno
