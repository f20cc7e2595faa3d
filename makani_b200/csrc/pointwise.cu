// Fused instance-norm kernels for the layers that wrap the spectral filter inside a FourierNeuralOperatorBlock
// (makani/models/networks/sfnonet.py:239-267 with nn.InstanceNorm2d(eps, affine) of sfnonet.py:375-380 and nn.GELU):
//
//     x = act(norm0(filter(x)))                      ->  stats + apply<GELU>
//     x = norm1(mlp(x)) + outer_skip(residual)       ->  stats + apply<none, skip>
//
// SURVEY 8(f) rank 1.  Pure HBM-bandwidth work on [B*C] planes of H*W contiguous elements (bf16 / fp16 / fp32 in and
// out, fp32 arithmetic): 16-byte vector access, every plane cut into kSeg segments so that the grid is
// planes x kSeg CTAs (3072 at C = 384: ~20 per SM) and the per-plane reduction is a fixed-order sum of kSeg
// partials - deterministic, no atomics.
//
//   forward   pass 1  instnorm_stats     partial[plane][seg] = ( sum (x-K), sum (x-K)^2 ),  K = first element of the plane
//             (tiny)  instnorm_finalize  mean = K + s1/n, var = s2/n - (s1/n)^2 ;  stats[plane] = (mean, rsqrt(var+eps))
//             pass 2  instnorm_apply     out = act( gamma_c (x - mean) rstd + beta_c [+ skip] )
//   backward  pass 1  instnorm_bwd_stats partial[plane][seg] = ( sum dz, sum dz xh ),  dz = g act'(z), xh = (x-mean) rstd
//             pass 2  instnorm_bwd_apply dx = gamma_c rstd ( dz - (S1 + xh S2) / n_total )
//   (the per-plane totals S1, S2 are summed by the caller: that is also where a spatially sharded instance all-reduces
//    them, and d gamma_c = sum_b S2, d beta_c = sum_b S1)
//
// Traffic per field element: forward 2 reads + 1 write, backward 4 reads + 1 write; the unfused torch sequence
// (instance_norm, gelu and their backward kernels) moves 5 and 8.
#include "common.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../../include/makani_b200.h"
#include "pointwise_math.cuh"

namespace sfno {

constexpr int kSeg = SFNO_INSTNORM_SEGMENTS;
constexpr int kPwThreads = 256;

template <typename T> struct PwVec { static constexpr int n = 16 / sizeof(T); };

template <typename T> __device__ __forceinline__ float pw_to_float(T v);
template <> __device__ __forceinline__ float pw_to_float<float>(float v) { return v; }
template <> __device__ __forceinline__ float pw_to_float<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float pw_to_float<__half>(__half v) { return __half2float(v); }

template <typename T> __device__ __forceinline__ T pw_from_float(float v);
template <> __device__ __forceinline__ float pw_from_float<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 pw_from_float<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half pw_from_float<__half>(float v) { return __float2half_rn(v); }

// 16-byte vector <-> PwVec<T>::n floats
template <typename T>
__device__ __forceinline__ void pw_unpack(const uint4& q, float (&f)[PwVec<T>::n]) {
  const T* e = reinterpret_cast<const T*>(&q);
#pragma unroll
  for (int j = 0; j < PwVec<T>::n; ++j) f[j] = pw_to_float<T>(e[j]);
}
template <typename T>
__device__ __forceinline__ uint4 pw_pack(const float (&f)[PwVec<T>::n]) {
  uint4 q;
  T* e = reinterpret_cast<T*>(&q);
#pragma unroll
  for (int j = 0; j < PwVec<T>::n; ++j) e[j] = pw_from_float<T>(f[j]);
  return q;
}

__device__ __forceinline__ uint4 pw_ldg(const void* p) {            // streaming read: no reuse inside the kernel
  return __ldcs(reinterpret_cast<const uint4*>(p));
}

// exact (erf) GELU, the default of nn.GELU: arithmetic in pointwise_math.cuh (shared with the host accuracy test)
__device__ __forceinline__ float gelu_f(float z) { return pw_gelu(z); }
__device__ __forceinline__ float gelu_grad_f(float z) { return pw_gelu_grad(z); }

// GELU / its derivative over a vector of V values, two at a time where the packed path exists (pointwise_math.cuh)
template <int V>
__device__ __forceinline__ void gelu_vec(float (&z)[V]) {
#if SFNO_PW_PACKED
#pragma unroll
  for (int j = 0; j + 1 < V; j += 2) pw_gelu_pair(z[j], z[j + 1]);
  if (V & 1) z[V - 1] = gelu_f(z[V - 1]);
#else
#pragma unroll
  for (int j = 0; j < V; ++j) z[j] = gelu_f(z[j]);
#endif
}
template <int V>
__device__ __forceinline__ void gelu_grad_vec(float (&z)[V]) {
#if SFNO_PW_PACKED
#pragma unroll
  for (int j = 0; j + 1 < V; j += 2) pw_gelu_grad_pair(z[j], z[j + 1]);
  if (V & 1) z[V - 1] = gelu_grad_f(z[V - 1]);
#else
#pragma unroll
  for (int j = 0; j < V; ++j) z[j] = gelu_grad_f(z[j]);
#endif
}

// Segment `seg` of a plane of n elements, in units of `unit` elements: [lo, hi)
// Segment boundaries lie on multiples of 8 units (128 bytes for 16-byte vectors) so that no warp store starts in
// mid-sector, and results are written with streaming stores (they are not read again by the kernel).  Measured on
// B200 at 384 x 721 x 1440 bf16 (profiles/r2_pointwise_store_ab.txt): bias add in place 0.65 -> 0.83 of the HBM peak,
// norm + skip 0.79 -> 0.89; the GELU kernels stay at 0.65 - 0.69 (issue-bound: ~12 instructions per bf16 element is all
// an SM can issue at the HBM rate, the erf alone takes more).
__device__ __forceinline__ void pw_range(long long n_units, int seg, long long& lo, long long& hi) {
  lo = (n_units * seg / kSeg) & ~7LL;
  hi = n_units * (seg + 1) / kSeg;
  if (seg + 1 < kSeg) hi &= ~7LL;
}
__device__ __forceinline__ void pw_stg(void* p, const uint4& q) { __stcs(reinterpret_cast<uint4*>(p), q); }

// Software-pipelined walk over the 16-byte vectors [lo, hi) of one or two streams: the loads of the next kPwDepth
// vectors are issued before the current ones are processed, so that a warp keeps bytes in flight while it computes
// (with ~100 instructions of GELU arithmetic per vector a plain load - compute - store loop leaves the memory pipe idle
// for 40 % of a warp's time).  Measured at 384 x 721 x 1440 bf16 (profiles/r2u_pointwise_pipelined.txt): bias + GELU
// 365 -> 299 us (0.81 of the HBM peak), its backward 523 -> 399 us (0.91), norm + GELU 368 -> 306 us (0.80), its backward
// apply 450 -> 399 us (0.91), norm + skip 410 -> 393 us (0.93).
constexpr int kPwDepth = 2;
template <typename T, class F>
__device__ __forceinline__ void pw_stream1(const T* __restrict__ p, long long lo, long long hi, F fn) {
  constexpr int V = PwVec<T>::n;
  long long i = lo + threadIdx.x;
  uint4 cur[kPwDepth];
#pragma unroll
  for (int u = 0; u < kPwDepth; ++u)
    if (i + u * kPwThreads < hi) cur[u] = pw_ldg(p + (i + u * kPwThreads) * V);
  for (; i < hi; i += kPwDepth * kPwThreads) {
    uint4 nxt[kPwDepth];
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u) {
      const long long k = i + (kPwDepth + u) * kPwThreads;
      if (k < hi) nxt[u] = pw_ldg(p + k * V);
    }
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u)
      if (i + u * kPwThreads < hi) fn(i + u * kPwThreads, cur[u]);
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u) cur[u] = nxt[u];
  }
}
template <typename T, class F>
__device__ __forceinline__ void pw_stream2(const T* __restrict__ pa, const T* __restrict__ pb, long long lo, long long hi,
                                           F fn) {
  constexpr int V = PwVec<T>::n;
  long long i = lo + threadIdx.x;
  uint4 ca[kPwDepth], cb[kPwDepth];
#pragma unroll
  for (int u = 0; u < kPwDepth; ++u)
    if (i + u * kPwThreads < hi) {
      ca[u] = pw_ldg(pa + (i + u * kPwThreads) * V);
      cb[u] = pw_ldg(pb + (i + u * kPwThreads) * V);
    }
  for (; i < hi; i += kPwDepth * kPwThreads) {
    uint4 na[kPwDepth], nb[kPwDepth];
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u) {
      const long long k = i + (kPwDepth + u) * kPwThreads;
      if (k < hi) {
        na[u] = pw_ldg(pa + k * V);
        nb[u] = pw_ldg(pb + k * V);
      }
    }
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u)
      if (i + u * kPwThreads < hi) fn(i + u * kPwThreads, ca[u], cb[u]);
#pragma unroll
    for (int u = 0; u < kPwDepth; ++u) {
      ca[u] = na[u];
      cb[u] = nb[u];
    }
  }
}

// block-wide sum of two values; result valid on thread 0
__device__ __forceinline__ float2 pw_block_sum2(float a, float b) {
  __shared__ float2 red[kPwThreads / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_xor_sync(0xffffffffu, a, o);
    b += __shfl_xor_sync(0xffffffffu, b, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = make_float2(a, b);
  __syncthreads();
  float2 r = make_float2(0.f, 0.f);
  if (threadIdx.x == 0) {
#pragma unroll
    for (int w = 0; w < kPwThreads / 32; ++w) {
      r.x += red[w].x;
      r.y += red[w].y;
    }
  }
  return r;
}

// ------------------------------------------------------------------------------------------------ forward
template <typename T>
__global__ void __launch_bounds__(kPwThreads)
instnorm_stats_kernel(const T* __restrict__ x, long long n, int vec_ok, float2* __restrict__ partial) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const T* px = x + plane * n;
  const float K = pw_to_float<T>(px[0]);
  float s1 = 0.f, s2 = 0.f;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
#pragma unroll 4
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      float f[V];
      pw_unpack<T>(pw_ldg(px + i * V), f);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float d = f[j] - K;
        s1 += d;
        s2 = fmaf(d, d, s2);
      }
    }
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      const float d = pw_to_float<T>(px[i]) - K;
      s1 += d;
      s2 = fmaf(d, d, s2);
    }
  }
  const float2 r = pw_block_sum2(s1, s2);
  if (threadIdx.x == 0) partial[blockIdx.x] = r;
}

template <typename T>
__global__ void instnorm_finalize_kernel(const T* __restrict__ x, long long n, int planes, float eps,
                                         const float2* __restrict__ partial, float2* __restrict__ mean_var,
                                         float2* __restrict__ stats) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= planes) return;
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int s = 0; s < kSeg; ++s) {
    const float2 v = partial[(long long)p * kSeg + s];
    s1 += v.x;
    s2 += v.y;
  }
  const float K = pw_to_float<T>(x[(long long)p * n]);
  const float inv_n = 1.f / (float)n;
  const float m1 = s1 * inv_n;
  const float var = fmaxf(s2 * inv_n - m1 * m1, 0.f);
  const float mean = K + m1;
  if (mean_var) mean_var[p] = make_float2(mean, var);
  if (stats) stats[p] = make_float2(mean, rsqrtf(var + eps));
}

template <typename T, int ACT, bool SKIP>
__global__ void __launch_bounds__(kPwThreads)
instnorm_apply_kernel(const T* __restrict__ x, const T* __restrict__ skip, const float2* __restrict__ stats,
                      const float* __restrict__ gamma, const float* __restrict__ beta, int C, long long n, int vec_ok,
                      T* __restrict__ out) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const int c = (int)(plane % C);
  const float2 st = stats[plane];
  const float a = st.y * (gamma ? gamma[c] : 1.f);
  const float b = beta ? beta[c] : 0.f;
  const float mean = st.x;                     // z = a (x - mean) + b: subtract first, a large mean must not cancel late
  const T* px = x + plane * n;
  const T* ps = SKIP ? skip + plane * n : nullptr;
  T* po = out + plane * n;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
    if (SKIP) {
      pw_stream2<T>(px, ps, lo, hi, [&](long long i, const uint4& qx, const uint4& qs) {
        float f[V], k[V];
        pw_unpack<T>(qx, f);
        pw_unpack<T>(qs, k);
#pragma unroll
        for (int j = 0; j < V; ++j) f[j] = fmaf(a, f[j] - mean, b) + k[j];
        pw_stg(po + i * V, pw_pack<T>(f));
      });
    } else {
      pw_stream1<T>(px, lo, hi, [&](long long i, const uint4& qx) {
        float f[V];
        pw_unpack<T>(qx, f);
#pragma unroll
        for (int j = 0; j < V; ++j) f[j] = fmaf(a, f[j] - mean, b);
        if (ACT == 1) gelu_vec<V>(f);
        pw_stg(po + i * V, pw_pack<T>(f));
      });
    }
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      float z = fmaf(a, pw_to_float<T>(px[i]) - mean, b);
      if (SKIP) z += pw_to_float<T>(ps[i]);
      po[i] = pw_from_float<T>(ACT == 1 ? gelu_f(z) : z);
    }
  }
}

// ------------------------------------------------------------------------------------------------ backward
template <typename T, int ACT>
__global__ void __launch_bounds__(kPwThreads)
instnorm_bwd_stats_kernel(const T* __restrict__ g, const T* __restrict__ x, const float2* __restrict__ stats,
                          const float* __restrict__ gamma, const float* __restrict__ beta, int C, long long n,
                          int vec_ok, float2* __restrict__ partial) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const int c = (int)(plane % C);
  const float2 st = stats[plane];
  const float ga = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
  const T* px = x + plane * n;
  const T* pg = g + plane * n;
  float s1 = 0.f, s2 = 0.f;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
    pw_stream2<T>(px, pg, lo, hi, [&](long long, const uint4& qx, const uint4& qg) {
      float f[V], d[V], gp[V];
      pw_unpack<T>(qx, f);
      pw_unpack<T>(qg, d);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        f[j] = (f[j] - st.x) * st.y;                                 // xhat
        gp[j] = fmaf(ga, f[j], be);
      }
      if (ACT == 1) gelu_grad_vec<V>(gp);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float dz = ACT == 1 ? d[j] * gp[j] : d[j];
        s1 += dz;
        s2 = fmaf(dz, f[j], s2);
      }
    });
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      const float xh = (pw_to_float<T>(px[i]) - st.x) * st.y;
      const float d = pw_to_float<T>(pg[i]);
      const float dz = ACT == 1 ? d * gelu_grad_f(fmaf(ga, xh, be)) : d;
      s1 += dz;
      s2 = fmaf(dz, xh, s2);
    }
  }
  const float2 r = pw_block_sum2(s1, s2);
  if (threadIdx.x == 0) partial[blockIdx.x] = r;
}

template <typename T, int ACT>
__global__ void __launch_bounds__(kPwThreads)
instnorm_bwd_apply_kernel(const T* __restrict__ g, const T* __restrict__ x, const float2* __restrict__ stats,
                          const float* __restrict__ gamma, const float* __restrict__ beta,
                          const float2* __restrict__ bsum, float inv_n_total, int C, long long n, int vec_ok,
                          T* __restrict__ dx) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const int c = (int)(plane % C);
  const float2 st = stats[plane];
  const float ga = gamma ? gamma[c] : 1.f, be = beta ? beta[c] : 0.f;
  const float a = st.y * ga;
  const float2 tot = bsum[plane];
  const float m1 = tot.x * inv_n_total, m2 = tot.y * inv_n_total;
  const T* px = x + plane * n;
  const T* pg = g + plane * n;
  T* po = dx + plane * n;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
    pw_stream2<T>(px, pg, lo, hi, [&](long long i, const uint4& qx, const uint4& qg) {
      float f[V], d[V], gp[V];
      pw_unpack<T>(qx, f);
      pw_unpack<T>(qg, d);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        f[j] = (f[j] - st.x) * st.y;                                 // xhat
        gp[j] = fmaf(ga, f[j], be);
      }
      if (ACT == 1) gelu_grad_vec<V>(gp);
#pragma unroll
      for (int j = 0; j < V; ++j) {
        const float dz = ACT == 1 ? d[j] * gp[j] : d[j];
        f[j] = a * (dz - m1 - f[j] * m2);
      }
      pw_stg(po + i * V, pw_pack<T>(f));
    });
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      const float xh = (pw_to_float<T>(px[i]) - st.x) * st.y;
      const float d = pw_to_float<T>(pg[i]);
      const float dz = ACT == 1 ? d * gelu_grad_f(fmaf(ga, xh, be)) : d;
      po[i] = pw_from_float<T>(a * (dz - m1 - xh * m2));
    }
  }
}

// ------------------------------------------------------------------------------------------------ bias + activation
// The elementwise work around the 1x1 convolutions of the MLP / encoder / decoder (makani/models/common/layers.py:
// 86-214: Conv2d(1x1, bias) -> GELU -> Conv2d(1x1)), on [B*C] planes like the kernels above:
//   bias_act      out = act(x + bias_c)                                  1 read + 1 write (in place allowed)
//   bias_act_bwd  dx = g * act'(x + bias_c),  partial[plane][seg] = sum dx   2 reads + 1 write; d bias_c = sum_b sum_seg
// (dx == nullptr with ACT = none: only the sums - the bias gradient of a convolution without activation)
template <typename T, int ACT>
__global__ void __launch_bounds__(kPwThreads)
bias_act_kernel(const T* __restrict__ x, const float* __restrict__ bias, int C, long long n, int vec_ok,
                T* __restrict__ out) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const float b = bias ? bias[(int)(plane % C)] : 0.f;
  const T* px = x + plane * n;
  T* po = out + plane * n;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
    pw_stream1<T>(px, lo, hi, [&](long long i, const uint4& q) {
      float f[V];
      pw_unpack<T>(q, f);
#pragma unroll
      for (int j = 0; j < V; ++j) f[j] += b;
      if (ACT == 1) gelu_vec<V>(f);
      pw_stg(po + i * V, pw_pack<T>(f));
    });
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      const float z = pw_to_float<T>(px[i]) + b;
      po[i] = pw_from_float<T>(ACT == 1 ? gelu_f(z) : z);
    }
  }
}

template <typename T, int ACT>
__global__ void __launch_bounds__(kPwThreads)
bias_act_bwd_kernel(const T* __restrict__ g, const T* __restrict__ x, const float* __restrict__ bias, int C, long long n,
                    int vec_ok, T* __restrict__ dx, float* __restrict__ partial) {
  constexpr int V = PwVec<T>::n;
  const long long plane = blockIdx.x / kSeg;
  const int seg = blockIdx.x % kSeg;
  const float b = bias ? bias[(int)(plane % C)] : 0.f;
  const T* pg = g + plane * n;
  const T* px = ACT == 1 ? x + plane * n : nullptr;
  T* po = dx ? dx + plane * n : nullptr;
  float s = 0.f;
  long long lo, hi;
  if (vec_ok) {
    pw_range(n / V, seg, lo, hi);
    auto finish = [&](long long i, float (&d)[V]) {
      if (po) {
        const uint4 q = pw_pack<T>(d);
        pw_stg(po + i * V, q);
        if (sizeof(T) == 2) pw_unpack<T>(q, d);          // sum what was stored (the rounded values feed the GEMMs)
      }
#pragma unroll
      for (int j = 0; j < V; ++j) s += d[j];
    };
    if (ACT == 1) {
      pw_stream2<T>(pg, px, lo, hi, [&](long long i, const uint4& qg, const uint4& qx) {
        float d[V], f[V];
        pw_unpack<T>(qg, d);
        pw_unpack<T>(qx, f);
#pragma unroll
        for (int j = 0; j < V; ++j) f[j] += b;
        gelu_grad_vec<V>(f);
#pragma unroll
        for (int j = 0; j < V; ++j) d[j] *= f[j];
        finish(i, d);
      });
    } else {
      pw_stream1<T>(pg, lo, hi, [&](long long i, const uint4& qg) {
        float d[V];
        pw_unpack<T>(qg, d);
        finish(i, d);
      });
    }
  } else {
    pw_range(n, seg, lo, hi);
    for (long long i = lo + threadIdx.x; i < hi; i += kPwThreads) {
      float d = pw_to_float<T>(pg[i]);
      if (ACT == 1) d *= gelu_grad_f(pw_to_float<T>(px[i]) + b);
      if (po) {
        const T r = pw_from_float<T>(d);
        po[i] = r;
        d = pw_to_float<T>(r);
      }
      s += d;
    }
  }
  const float2 r = pw_block_sum2(s, 0.f);
  if (threadIdx.x == 0) partial[blockIdx.x] = r.x;
}

static bool pw_vec_ok(const void* p0, const void* p1, const void* p2, long long n, int esize) {
  const int V = 16 / esize;
  auto al = [](const void* p) { return p == nullptr || (reinterpret_cast<uintptr_t>(p) & 15u) == 0; };
  return (n % V == 0) && al(p0) && al(p1) && al(p2);
}

}  // namespace sfno

using namespace sfno;

#define SFNO_PW_DISPATCH(dtype, ...)                                   \
  switch (dtype) {                                                     \
    case SFNO_F32: { using T = float; __VA_ARGS__; } break;            \
    case SFNO_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;   \
    case SFNO_F16: { using T = __half; __VA_ARGS__; } break;           \
    default: SFNO_CHECK(false, "unsupported dtype");                   \
  }

extern "C" {

int sfno_instnorm_stats(const void* x, int dtype, int planes, long long hw, float eps, float* partial_ws,
                        float* mean_var, float* stats, void* stream) {
  SFNO_CHECK(x && partial_ws && (mean_var || stats), "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && (long long)planes * kSeg < (1LL << 31), "bad sizes");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(x, nullptr, nullptr, hw, esize) ? 1 : 0;
  SFNO_PW_DISPATCH(dtype, instnorm_stats_kernel<T><<<planes * kSeg, kPwThreads, 0, s>>>(
                              (const T*)x, hw, vec, (float2*)partial_ws));
  SFNO_LAUNCHED();
  SFNO_PW_DISPATCH(dtype, instnorm_finalize_kernel<T><<<(planes + 127) / 128, 128, 0, s>>>(
                              (const T*)x, hw, planes, eps, (const float2*)partial_ws, (float2*)mean_var,
                              (float2*)stats));
  SFNO_LAUNCHED();
  return 0;
}

int sfno_instnorm_apply(const void* x, const void* skip, int dtype, int planes, int C, long long hw,
                        const float* stats, const float* gamma, const float* beta, int act, void* out, void* stream) {
  SFNO_CHECK(x && stats && out, "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && C > 0 && planes % C == 0, "bad sizes");
  SFNO_CHECK(act == SFNO_ACT_NONE || act == SFNO_ACT_GELU, "unknown activation");
  SFNO_CHECK(!(skip && act != SFNO_ACT_NONE), "a skip input is only supported without activation");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(x, skip, out, hw, esize) ? 1 : 0;
  const int grid = planes * kSeg;
#define SFNO_PW_APPLY(ACT, SKIP)                                                                          \
  SFNO_PW_DISPATCH(dtype, instnorm_apply_kernel<T, ACT, SKIP><<<grid, kPwThreads, 0, s>>>(               \
                              (const T*)x, (const T*)skip, (const float2*)stats, gamma, beta, C, hw, vec, (T*)out))
  if (act == SFNO_ACT_GELU) { SFNO_PW_APPLY(1, false); }
  else if (skip) { SFNO_PW_APPLY(0, true); }
  else { SFNO_PW_APPLY(0, false); }
#undef SFNO_PW_APPLY
  SFNO_LAUNCHED();
  return 0;
}

int sfno_instnorm_bwd_stats(const void* g, const void* x, int dtype, int planes, int C, long long hw,
                            const float* stats, const float* gamma, const float* beta, int act, float* partial,
                            void* stream) {
  SFNO_CHECK(g && x && stats && partial, "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && C > 0 && planes % C == 0, "bad sizes");
  SFNO_CHECK(act == SFNO_ACT_NONE || act == SFNO_ACT_GELU, "unknown activation");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(x, g, nullptr, hw, esize) ? 1 : 0;
  const int grid = planes * kSeg;
  if (act == SFNO_ACT_GELU) {
    SFNO_PW_DISPATCH(dtype, instnorm_bwd_stats_kernel<T, 1><<<grid, kPwThreads, 0, s>>>(
                                (const T*)g, (const T*)x, (const float2*)stats, gamma, beta, C, hw, vec,
                                (float2*)partial));
  } else {
    SFNO_PW_DISPATCH(dtype, instnorm_bwd_stats_kernel<T, 0><<<grid, kPwThreads, 0, s>>>(
                                (const T*)g, (const T*)x, (const float2*)stats, gamma, beta, C, hw, vec,
                                (float2*)partial));
  }
  SFNO_LAUNCHED();
  return 0;
}

int sfno_instnorm_bwd_apply(const void* g, const void* x, int dtype, int planes, int C, long long hw,
                            long long hw_total, const float* stats, const float* gamma, const float* beta,
                            const float* bsum, int act, void* dx, void* stream) {
  SFNO_CHECK(g && x && stats && bsum && dx, "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && hw_total >= hw && C > 0 && planes % C == 0, "bad sizes");
  SFNO_CHECK(act == SFNO_ACT_NONE || act == SFNO_ACT_GELU, "unknown activation");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(x, g, dx, hw, esize) ? 1 : 0;
  const int grid = planes * kSeg;
  const float inv = 1.f / (float)hw_total;
  if (act == SFNO_ACT_GELU) {
    SFNO_PW_DISPATCH(dtype, instnorm_bwd_apply_kernel<T, 1><<<grid, kPwThreads, 0, s>>>(
                                (const T*)g, (const T*)x, (const float2*)stats, gamma, beta, (const float2*)bsum, inv,
                                C, hw, vec, (T*)dx));
  } else {
    SFNO_PW_DISPATCH(dtype, instnorm_bwd_apply_kernel<T, 0><<<grid, kPwThreads, 0, s>>>(
                                (const T*)g, (const T*)x, (const float2*)stats, gamma, beta, (const float2*)bsum, inv,
                                C, hw, vec, (T*)dx));
  }
  SFNO_LAUNCHED();
  return 0;
}

int sfno_bias_act(const void* x, int dtype, int planes, int C, long long hw, const float* bias, int act, void* out,
                  void* stream) {
  SFNO_CHECK(x && out, "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && C > 0 && planes % C == 0 && (long long)planes * kSeg < (1LL << 31), "bad sizes");
  SFNO_CHECK(act == SFNO_ACT_NONE || act == SFNO_ACT_GELU, "unknown activation");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(x, out, nullptr, hw, esize) ? 1 : 0;
  const int grid = planes * kSeg;
  if (act == SFNO_ACT_GELU) {
    SFNO_PW_DISPATCH(dtype, bias_act_kernel<T, 1><<<grid, kPwThreads, 0, s>>>((const T*)x, bias, C, hw, vec, (T*)out));
  } else {
    SFNO_PW_DISPATCH(dtype, bias_act_kernel<T, 0><<<grid, kPwThreads, 0, s>>>((const T*)x, bias, C, hw, vec, (T*)out));
  }
  SFNO_LAUNCHED();
  return 0;
}

int sfno_bias_act_bwd(const void* g, const void* x, int dtype, int planes, int C, long long hw, const float* bias,
                      int act, void* dx, float* partial, void* stream) {
  SFNO_CHECK(g && partial, "null argument");
  SFNO_CHECK(planes > 0 && hw > 0 && C > 0 && planes % C == 0 && (long long)planes * kSeg < (1LL << 31), "bad sizes");
  SFNO_CHECK(act == SFNO_ACT_NONE || act == SFNO_ACT_GELU, "unknown activation");
  SFNO_CHECK(act == SFNO_ACT_NONE || (x && dx), "the GELU gradient needs the pre-activation x and an output dx");
  cudaStream_t s = (cudaStream_t)stream;
  const int esize = dtype == SFNO_F32 ? 4 : 2;
  const int vec = pw_vec_ok(g, act == SFNO_ACT_GELU ? x : nullptr, dx, hw, esize) ? 1 : 0;
  const int grid = planes * kSeg;
  if (act == SFNO_ACT_GELU) {
    SFNO_PW_DISPATCH(dtype, bias_act_bwd_kernel<T, 1><<<grid, kPwThreads, 0, s>>>((const T*)g, (const T*)x, bias, C, hw,
                                                                                  vec, (T*)dx, partial));
  } else {
    SFNO_PW_DISPATCH(dtype, bias_act_bwd_kernel<T, 0><<<grid, kPwThreads, 0, s>>>((const T*)g, (const T*)x, bias, C, hw,
                                                                                  vec, (T*)dx, partial));
  }
  SFNO_LAUNCHED();
  return 0;
}

}  // extern "C"
