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

}  // namespace sfno
