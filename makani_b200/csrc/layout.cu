// Bandwidth kernels around the tensor-core path: layout conversions at the module boundary
// (spectral-internal <-> complex64 tensors of the reference API), dhconv weight (un)packing, bias add,
// and the "diagonal" channel contraction (weights streamed once; no reuse, so no tensor cores).
// All global accesses are coalesced on both sides through 32x32 shared-memory tile transposes.
#include "common.cuh"
#include "plan.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <algorithm>
#include <type_traits>

#include <cstdlib>

namespace sfno {



// ------------------------------------------------------------------ SP <-> complex64 [B][C][L][M]
// grid (ceil(M/32), ceil(C/32), L*B), block (32, 8)
__global__ void sp_to_c64_kernel(const float* __restrict__ sp, float2* __restrict__ out, int B, int C, int L, int M,
                                 int l_off, int m_off) {
  __shared__ float tr[32][33], ti[32][33];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int l = blockIdx.z / B, b = blockIdx.z % B;
  const int m0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int m = m0 + j, c = c0 + threadIdx.x;
    float re = 0.f, im = 0.f;
    if (m < M && c < C && (m + m_off) <= (l + l_off)) {
      const long long base = ((long long)l * M + m) * RS + (long long)(b * 2) * Cp + c;
      re = sp[base];
      im = sp[base + Cp];
    }
    tr[j][threadIdx.x] = re;
    ti[j][threadIdx.x] = im;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, m = m0 + threadIdx.x;
    if (c < C && m < M) out[((long long)(b * C + c) * L + l) * M + m] = make_float2(tr[threadIdx.x][j], ti[threadIdx.x][j]);
  }
}

__global__ void c64_to_sp_kernel(const float2* __restrict__ in, float* __restrict__ sp, int B, int C, int L, int M,
                                 int l_off, int m_off, int round_out) {
  __shared__ float tr[32][33], ti[32][33];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int l = blockIdx.z / B, b = blockIdx.z % B;
  const int m0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int c = c0 + j, m = m0 + threadIdx.x;
    float2 v = make_float2(0.f, 0.f);
    // spectral data vanish for m > l: whatever the caller holds there is not read (the SP contract wants exact zeros)
    if (c < C && m < M && (m + m_off) <= (l + l_off)) v = in[((long long)(b * C + c) * L + l) * M + m];
    tr[j][threadIdx.x] = v.x;
    ti[j][threadIdx.x] = v.y;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int m = m0 + j, c = c0 + threadIdx.x;
    if (m < M && c < Cp) {      // pad channels are written as zeros so the whole SP tensor is defined
      const long long base = ((long long)l * M + m) * RS + (long long)(b * 2) * Cp + c;
      const float vr = tr[threadIdx.x][j], vi = ti[threadIdx.x][j];
      sp[base] = (c < C) ? (round_out ? tf32_rn(vr) : vr) : 0.f;     // operands of TF32 GEMMs: pre-rounded
      sp[base + Cp] = (c < C) ? (round_out ? tf32_rn(vi) : vi) : 0.f;
    }
  }
}

// ------------------------------------------------------------------ XM <-> complex64 [B][C][K][M]
// grid (ceil(M/32), ceil(K/32), B*C), block (32, 8)
__global__ void xm_to_c64_kernel(const float* __restrict__ xm, float2* __restrict__ out, int B, int C, int K, int Kp,
                                 int M) {
  __shared__ float tr[32][33], ti[32][33];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int b = blockIdx.z / C, c = blockIdx.z % C;
  const int m0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int m = m0 + j, k = k0 + threadIdx.x;
    float re = 0.f, im = 0.f;
    if (m < M && k < K) {
      const long long base = ((long long)m * RS + (long long)(b * 2) * Cp + c) * Kp + k;
      re = xm[base];
      im = xm[base + (long long)Cp * Kp];
    }
    tr[j][threadIdx.x] = re;
    ti[j][threadIdx.x] = im;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int k = k0 + j, m = m0 + threadIdx.x;
    if (k < K && m < M) out[((long long)(b * C + c) * K + k) * M + m] = make_float2(tr[threadIdx.x][j], ti[threadIdx.x][j]);
  }
}

__global__ void c64_to_xm_kernel(const float2* __restrict__ in, float* __restrict__ xm, int B, int C, int K, int Kp,
                                 int M) {
  __shared__ float tr[32][33], ti[32][33];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int b = blockIdx.z / C, c = blockIdx.z % C;
  const int m0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int k = k0 + j, m = m0 + threadIdx.x;
    float2 v = make_float2(0.f, 0.f);
    if (k < K && m < M) v = in[((long long)(b * C + c) * K + k) * M + m];
    tr[j][threadIdx.x] = v.x;
    ti[j][threadIdx.x] = v.y;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += 8) {
    const int m = m0 + j, k = k0 + threadIdx.x;
    if (m < M && k < Kp) {
      const long long base = ((long long)m * RS + (long long)(b * 2) * Cp + c) * Kp + k;
      xm[base] = (k < K) ? tf32_rn(tr[threadIdx.x][j]) : 0.f;
      xm[base + (long long)Cp * Kp] = (k < K) ? tf32_rn(ti[threadIdx.x][j]) : 0.f;
    }
  }
}

// ------------------------------------------------------------------ dhconv weights  w[Cin][Cout][L] c64 <-> WP[L][2][Coutp][Cinp]
// A 283 MB transpose of the slowest and the fastest dimension (per block of the model, every step).  One CTA moves a
// 64 (i) x 64 (l) tile of one output channel o through shared memory: the complex side is accessed in 512-byte runs
// (64 consecutive degrees), the planar side in 256-byte runs (64 consecutive input channels, 16-byte vectors), and
// every thread keeps 16 independent 8-byte (or 8 x 16-byte) loads in flight.
// grid (ceil(L/64), ceil(Cinp/64), Cout), block 256
constexpr int kWT = 64;
// shared tile element (i, l) lives at [i][l ^ (i >> 2)]: conflict free both for lanes along l (fixed i) and for
// lanes along i in steps of 4 (fixed l)
__device__ __forceinline__ int wsw(int i, int l) { return l ^ (i >> 2); }

// The unpack kernel is persistent: a CTA walks over tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... (degree tiles
// fastest, then input-channel tiles, then output channels) and requests the global loads of tile t+1 into registers
// BEFORE it stores tile t from shared memory.  With one tile per CTA every CTA of a wave first only reads and then only
// writes, in step with all others (measured: 58 us reading + 56 us writing = 115 us in total); overlapping the two
// phases and dropping the per-load slice lookup bring the unpack from 132 to 103 us (profiles/r2y_weight_layout_pipelined.txt).
struct WTile {
  int l0, i0, o;
};
__device__ __forceinline__ WTile wtile(int t, int nlt, int nit) {
  WTile r;
  r.l0 = (t % nlt) * kWT;
  t /= nlt;
  r.i0 = (t % nit) * kWT;
  r.o = t / nit;
  return r;
}

// One tile per CTA: grid (ceil(L/64), ceil(Cinp/64), Cout).  (A persistent variant with register prefetch like the
// unpack below needs 116 registers for its 16 + 16 eight-byte loads in flight and was slower: 103 - 113 us against 95.)
__global__ void __launch_bounds__(256)
pack_weight_kernel(const float2* __restrict__ w, float* __restrict__ wp, int Cin, int Cout, int L, int round_out) {
  __shared__ float tr[kWT][kWT], ti[kWT][kWT];
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const int o = blockIdx.z;
  const int l0 = blockIdx.x * kWT, i0 = blockIdx.y * kWT;
  const int tid = threadIdx.x;
  {
    const int ll = tid & 63, ib = tid >> 6;           // 64 degrees x 4 rows per pass
    float2 v[16];
#pragma unroll
    for (int p = 0; p < 16; ++p) {
      const int i = i0 + ib + 4 * p, l = l0 + ll;
      v[p] = (i < Cin && l < L) ? w[((long long)i * Cout + o) * L + l] : make_float2(0.f, 0.f);
    }
#pragma unroll
    for (int p = 0; p < 16; ++p) {
      // operands of TF32 GEMMs: pre-rounded (pads are exact zeros); round_out = 0 keeps the fp32 values (packed master copy)
      tr[ib + 4 * p][wsw(ib + 4 * p, ll)] = round_out ? tf32_rn(v[p].x) : v[p].x;
      ti[ib + 4 * p][wsw(ib + 4 * p, ll)] = round_out ? tf32_rn(v[p].y) : v[p].y;
    }
  }
  __syncthreads();
  {
    const int iq = (tid & 15) * 4, lb = tid >> 4;     // 16 lanes x float4 = 64 channels, 16 (l, re/im) rows per pass
    const bool iok = i0 + iq < Cinp;                  // Cinp is a multiple of 4: a vector is entirely in or out
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int row = lb + 16 * p;                    // 0..127: row = 2 * l_local + reim
      const int ll = row >> 1, reim = row & 1;
      const int l = l0 + ll;
      float (*t)[kWT] = reim ? ti : tr;
      const int cl = wsw(iq, ll);                     // iq .. iq+3 share i >> 2
      const float4 v = make_float4(t[iq][cl], t[iq + 1][cl], t[iq + 2][cl], t[iq + 3][cl]);
      if (iok && l < L)
        *reinterpret_cast<float4*>(wp + (((long long)l * 2 + reim) * Coutp + o) * Cinp + i0 + iq) = v;
    }
  }
}

// Source of the packed gradient: one buffer, or (gradient exchange over peer memory) slice k of the flat WP index
// space lives in bases[k] - this rank's own averaged slice locally, the others in the peers' buffers, read over
// NVLink.  slice is a multiple of 4 floats, so a 16-byte vector never straddles two slices.
struct WpSource {
  const float* bases[8];
  long long slice;          // floats per slice; 0: everything in bases[0]
  __device__ __forceinline__ const float* at(long long off) const {
    return (slice ? bases[off / slice] : bases[0]) + off;
  }
};

// GATHER = false: one buffer, addressed through a plain pointer (the per-load slice lookup of `at` - a 64-bit division
// and an indexed read of the parameter array - cost 16 of 132 us)
template <bool GATHER>
__global__ void __launch_bounds__(256)
unpack_wgrad_kernel(const WpSource src, float2* __restrict__ gw, int Cin, int Cout, int L, int ntiles) {
  const float* __restrict__ base0 = src.bases[0];
  __shared__ float tr[kWT][kWT], ti[kWT][kWT];
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const int nlt = (L + kWT - 1) / kWT, nit = (Cin + kWT - 1) / kWT;
  const int tid = threadIdx.x;
  const int iq = (tid & 15) * 4, lb = tid >> 4;       // loads: 16 lanes x float4 = 64 channels, 16 (l, re/im) rows per pass
  float4 v[8];
  auto load = [&](int t) {
    const WTile c = wtile(t, nlt, nit);
    const bool iok = c.i0 + iq < Cinp;
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int row = lb + 16 * p;
      const int l = c.l0 + (row >> 1), reim = row & 1;
      const long long off = (((long long)l * 2 + reim) * Coutp + c.o) * Cinp + c.i0 + iq;
      v[p] = (iok && l < L) ? *reinterpret_cast<const float4*>(GATHER ? src.at(off) : base0 + off)
                            : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  int t = blockIdx.x;
  if (t < ntiles) load(t);
  for (; t < ntiles; t += gridDim.x) {
    const WTile c = wtile(t, nlt, nit);
#pragma unroll
    for (int p = 0; p < 8; ++p) {
      const int row = lb + 16 * p;
      const int lr = row >> 1, reim = row & 1;
      float (*tt)[kWT] = reim ? ti : tr;
      const int cl = wsw(iq, lr);
      tt[iq][cl] = v[p].x;
      tt[iq + 1][cl] = v[p].y;
      tt[iq + 2][cl] = v[p].z;
      tt[iq + 3][cl] = v[p].w;
    }
    __syncthreads();
    if (t + (int)gridDim.x < ntiles) load(t + gridDim.x);         // in flight while this tile is stored
    if ((L & 1) == 0) {
      // two degrees per thread: 16-byte stores, a warp writes 512 contiguous bytes of one (i, o) row
      const int lp = (tid & 31) * 2, ib = tid >> 5;
#pragma unroll
      for (int p = 0; p < 8; ++p) {
        const int il = ib + 8 * p;
        const int i = c.i0 + il, l = c.l0 + lp;
        if (i < Cin && l < L) {                         // L even and l even: l + 1 < L as well
          const int c0 = wsw(il, lp), c1 = wsw(il, lp + 1);
          *reinterpret_cast<float4*>(gw + ((long long)i * Cout + c.o) * L + l) =
              make_float4(tr[il][c0], ti[il][c0], tr[il][c1], ti[il][c1]);
        }
      }
    } else {
      const int ll = tid & 63, ib = tid >> 6;
#pragma unroll
      for (int p = 0; p < 16; ++p) {
        const int i = c.i0 + ib + 4 * p, l = c.l0 + ll;
        if (i < Cin && l < L) gw[((long long)i * Cout + c.o) * L + l] =
              make_float2(tr[ib + 4 * p][wsw(ib + 4 * p, ll)], ti[ib + 4 * p][wsw(ib + 4 * p, ll)]);
      }
    }
    __syncthreads();                                  // the tile is overwritten by the next iteration
  }
}

// ------------------------------------------------------------------ bias
template <typename T>
__global__ void add_bias_kernel(T* __restrict__ y, const float* __restrict__ bias, int kind, long long n, int C,
                                long long HW) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long c = (i / HW) % C;
    const float bv = kind == 0 ? bias[c] : bias[c * HW + i % HW];
    float v;
    if constexpr (sizeof(T) == 4) {
      v = (float)y[i] + bv;
      y[i] = v;
    } else if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      y[i] = __float2bfloat16_rn(__bfloat162float(y[i]) + bv);
    } else {
      y[i] = __float2half_rn(__half2float(y[i]) + bv);
    }
  }
}

// ------------------------------------------------------------------ per-degree energy of an SP tensor (spectral loss)
// P[b][c][l] = sum_{m <= l} w_m |SP[l][m][b][c]|^2 with w_m = 1 at the global m = 0 and 2 elsewhere: what the
// reference's GeometricH1Loss forms from the complex64 coefficients (makani/utils/losses.py:299-303, 327-337:
// coeffs[...,0]^2 + coeffs[...,1]^2, then column 0 + 2 * sum of the others), read here straight from the SP layout.
// grid (ceil(C/32), L, B), block (32, 8): threadIdx.y strides over m, the eight partial sums meet in shared memory.
__global__ void __launch_bounds__(256) sp_degree_power_kernel(const float* __restrict__ sp, float* __restrict__ out,
                                                              int B, int C, int L, int M, int l_off, int m_off) {
  __shared__ float part[8][33];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int l = blockIdx.y, b = blockIdx.z, c = blockIdx.x * 32 + threadIdx.x;
  const int mend = min(M, l + l_off - m_off + 1);              // defined entries: m + m_off <= l + l_off
  float acc = 0.f;
  if (c < C) {
    const float* p = sp + (long long)l * M * RS + (long long)(2 * b) * Cp + c;
    for (int m = threadIdx.y; m < mend; m += 8) {
      const float re = p[m * RS], im = p[m * RS + Cp];
      const float w = (m + m_off == 0) ? 1.f : 2.f;
      acc = fmaf(w * re, re, acc);
      acc = fmaf(w * im, im, acc);
    }
  }
  part[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < C) {
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += part[j][threadIdx.x];
    out[((long long)b * C + c) * L + l] = s;
  }
}

// its adjoint: gSP[l][m][b][re|im][c] = 2 w_m g[b][c][l] SP[l][m][b][re|im][c] for m <= l and exact zeros in the rest of
// the tensor (m > l, pad channels), as the SP contract wants.  grid (L, B), block (64, 4)
__global__ void __launch_bounds__(256) sp_degree_power_bwd_kernel(const float* __restrict__ sp, const float* __restrict__ g,
                                                                  float* __restrict__ gsp, int B, int C, int L, int M,
                                                                  int l_off, int m_off, int round_out) {
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int l = blockIdx.x, b = blockIdx.y;
  const int mend = min(M, l + l_off - m_off + 1);
  const long long row0 = (long long)l * M * RS + (long long)(2 * b) * Cp;
  for (int j = threadIdx.x; j < 2 * Cp; j += 64) {
    const int c = j < Cp ? j : j - Cp;
    const float gl = c < C ? 2.f * g[((long long)b * C + c) * L + l] : 0.f;
    for (int m = threadIdx.y; m < M; m += 4) {
      const long long off = row0 + m * RS + j;
      float v = 0.f;
      if (m < mend && c < C) {
        v = ((m + m_off == 0) ? gl : 2.f * gl) * sp[off];
        if (round_out) v = tf32_rn(v);
      }
      gsp[off] = v;
    }
  }
}

// dst += src over n floats (n % 4 == 0, 16-byte aligned): sum of two SP gradients on the h-sharded path
__global__ void __launch_bounds__(256) add_inplace_kernel(float4* __restrict__ dst, const float4* __restrict__ src, long long n4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 a = dst[i];
    const float4 b = src[i];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    dst[i] = a;
  }
}

// dst = (dst + sum_k src[k]) * scale over n floats: the reduction step of the peer-memory gradient exchange (this
// rank's slice of the packed weight gradient plus the same slice pulled from every other rank)
constexpr int kMaxAvgSrc = 7;
struct AvgArgs {
  const float4* src[kMaxAvgSrc];
};
__global__ void __launch_bounds__(256) slices_average_kernel(float4* __restrict__ dst, const AvgArgs args, int nsrc,
                                                             long long n4, float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 a = dst[i];
#pragma unroll
    for (int k = 0; k < kMaxAvgSrc; ++k) {
      if (k < nsrc) {
        const float4 b = args.src[k][i];
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      }
    }
    a.x *= scale; a.y *= scale; a.z *= scale; a.w *= scale;
    dst[i] = a;
  }
}

// constant bias as the m = 0 Fourier coefficient: irfft(norm="forward") adds Re XM[0][b][c][k] to every longitude of
// latitude k, so  y + bias[c]  ==  irfft(XM with XM[0].re += bias[c])  - the add happens in fp32 before the cast to the
// output dtype, as the reference does (spectral_convolution.py:143-146), and costs B*C*nlat elements instead of a pass
// over the field.
__global__ void xm_add_dc_kernel(float* __restrict__ xm, const float* __restrict__ bias, int B, int C, int K, int Kp) {
  const int Cp = (C + 3) & ~3;
  const int n = B * C * K;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int k = i % K, c = (i / K) % C, b = i / (K * C);
    xm[((long long)(b * 2) * Cp + c) * Kp + k] += bias[c];
  }
}

// ------------------------------------------------------------------ diagonal contraction (contractions.py:122-127)
//   MODE 0: y[b,o,l,m]  = sum_i x[b,i,l,m] w[i,o,l,m]          (a = SP x,  out = SP y)
//   MODE 1: dx[b,i,l,m] = sum_o g[b,o,l,m] conj(w[i,o,l,m])    (a = SP g,  out = SP dx)
//   MODE 2: dw[i,o,l,m] = sum_b conj(x[b,i,l,m]) g[b,o,l,m]    (a = SP x, b2 = SP g, out = complex64 dw)
// One CTA per (l, 32 consecutive m, chunk): the SP tiles are staged in shared memory (channel-contiguous reads), the
// weight is streamed with m across lanes (contiguous).  block (32, 8).
// Chunking keeps the staged tiles inside shared memory for any batch / channel count (the first version staged all
// batch entries and all channels and failed for B*(Cin+Cout) > 799):
//   MODE 0/1: blockIdx.z = (batch chunk of Bc <= kDiagBc entries) x (chunk of Cc result channels); the Bc results of
//             one weight element are accumulated in registers, so the weight is read once per batch chunk
//   MODE 2  : blockIdx.z = chunk of Cc output channels o; the batch is walked in chunks of Bc inside the kernel and
//             the partial sums go through the (L2-resident) result
constexpr int kDiagBc = 4;
constexpr int kDiagCi = 128;   // MODE 2: input channels per CTA
template <int MODE>
__global__ void diag_kernel(const float* __restrict__ spa, const void* __restrict__ wb, void* __restrict__ outp,
                            int B, int Cin, int Cout, int L, int M, int l_off, int m_off, int Bc, int Cc) {
  extern __shared__ float smd[];
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const int RSi = 2 * B * Cinp, RSo = 2 * B * Coutp;
  const int l = blockIdx.y, m0 = blockIdx.x * 32;
  const int mloc = threadIdx.x, m = m0 + mloc;
  const int tid = threadIdx.y * 32 + threadIdx.x;
  const int mhi_l = max(0, min(M, ((l + l_off + 16) & ~15) - m_off));   // defined extent of SP row l
  const long long LM = (long long)L * M;
  // stage rows [r0, r0 + n) of the SP tile of (l, m0..m0+31) into t[32][n + 1]; zero outside the defined / triangular region
  auto stage = [&](float* t, const float* sp, int RS, int r0, int n) {
    for (int idx = tid; idx < 32 * n; idx += 256) {
      const int mm = idx / n, r = idx - mm * n;
      float v = 0.f;
      if (m0 + mm < mhi_l && (m0 + mm + m_off) <= (l + l_off)) v = sp[((long long)l * M + m0 + mm) * RS + r0 + r];
      t[mm * (n + 1) + r] = v;
    }
  };
  if (MODE == 0 || MODE == 1) {
    const int Cred = (MODE == 0) ? Cin : Cout, Cres = (MODE == 0) ? Cout : Cin;
    const int Credp = (MODE == 0) ? Cinp : Coutp, Cresp = (MODE == 0) ? Coutp : Cinp;
    const int RSa = (MODE == 0) ? RSi : RSo, RSb = (MODE == 0) ? RSo : RSi;
    const int ncc = (Cresp + Cc - 1) / Cc;
    const int b0 = (blockIdx.z / ncc) * Bc, nb = min(Bc, B - b0);
    const int c0 = (blockIdx.z % ncc) * Cc;
    const int ncw = min(Cc, Cresp - c0);                // channels written by this CTA (pad channels: zeros)
    const int nco = max(0, min(Cc, Cres - c0));         // channels computed
    const int na = 2 * nb * Credp, nt = 2 * nb * Cc;
    float* ta = smd;                                    // [32][na + 1]   x (or g) of this batch chunk, all channels
    float* tb = smd + 32 * (na + 1);                    // [32][nt + 1]   results of this batch chunk, this channel chunk
    stage(ta, spa, RSa, b0 * 2 * Credp, na);
    __syncthreads();
    const float2* w = static_cast<const float2*>(wb);
    const float* xrow = ta + mloc * (na + 1);
    for (int co = threadIdx.y; co < ncw; co += 8) {
      float re[kDiagBc], im[kDiagBc];
#pragma unroll
      for (int bb = 0; bb < kDiagBc; ++bb) re[bb] = im[bb] = 0.f;
      if (m < M && co < nco) {
        for (int ci = 0; ci < Cred; ++ci) {
          const int i = (MODE == 0) ? ci : c0 + co, o = (MODE == 0) ? c0 + co : ci;
          const float2 wv = w[((long long)i * Cout + o) * LM + (long long)l * M + m];
#pragma unroll
          for (int bb = 0; bb < kDiagBc; ++bb) {
            if (bb < nb) {
              const float xr = xrow[(bb * 2) * Credp + ci], xi = xrow[(bb * 2 + 1) * Credp + ci];
              if (MODE == 0) {
                re[bb] += xr * wv.x - xi * wv.y;
                im[bb] += xr * wv.y + xi * wv.x;
              } else {
                re[bb] += xr * wv.x + xi * wv.y;
                im[bb] += xi * wv.x - xr * wv.y;
              }
            }
          }
        }
      }
#pragma unroll
      for (int bb = 0; bb < kDiagBc; ++bb) {
        if (bb < nb) {
          tb[mloc * (nt + 1) + (bb * 2) * Cc + co] = re[bb];
          tb[mloc * (nt + 1) + (bb * 2 + 1) * Cc + co] = im[bb];
        }
      }
    }
    __syncthreads();
    float* out = static_cast<float*>(outp);
    const int nw = 2 * nb * ncw;
    for (int idx = tid; idx < 32 * nw; idx += 256) {
      const int mm = idx / nw, r = idx - mm * nw;
      const int j = r / ncw, c = r - j * ncw;           // j = bb*2 + reim
      if (m0 + mm < mhi_l)
        out[((long long)l * M + m0 + mm) * RSb + (long long)(b0 * 2 + j) * Cresp + c0 + c] = tb[mm * (nt + 1) + j * Cc + c];
    }
  } else {
    // blockIdx.z = (chunk of kDiagCi input channels) x (chunk of Cc output channels); both staged tiles hold one batch
    // chunk of their channel chunk only
    const int nic = (Cin + kDiagCi - 1) / kDiagCi;
    const int i0 = (blockIdx.z % nic) * kDiagCi, nci = min(kDiagCi, Cin - i0);
    const int o0 = (blockIdx.z / nic) * Cc, nco = min(Cc, Cout - o0);
    float2* gw = static_cast<float2*>(outp);
    const float* spg = static_cast<const float*>(wb);
    float* ta = smd;                                    // [32][2*Bc*kDiagCi + 1]  x of this batch / channel chunk
    float* tb = smd + 32 * (2 * Bc * kDiagCi + 1);      // [32][2*Bc*Cc + 1]       g of this batch / channel chunk
    // stage channels [c0, c0 + nc) of batch entries [b0, b0 + nb): t[mm][(bb*2 + reim)*nc + c]
    auto stage_ch = [&](float* t, const float* sp, int RS, int Cp, int b0, int nb, int c0, int nc) {
      const int n = 2 * nb * nc;
      for (int idx = tid; idx < 32 * n; idx += 256) {
        const int mm = idx / n, r = idx - mm * n;
        const int j = r / nc, c = r - j * nc;
        float v = 0.f;
        if (m0 + mm < mhi_l && (m0 + mm + m_off) <= (l + l_off))
          v = sp[((long long)l * M + m0 + mm) * RS + (long long)(b0 * 2 + j) * Cp + c0 + c];
        t[mm * (n + 1) + r] = v;
      }
    };
    for (int b0 = 0; b0 < B; b0 += Bc) {
      const int nb = min(Bc, B - b0);
      const int na = 2 * nb * nci, nt = 2 * nb * nco;
      __syncthreads();
      stage_ch(ta, spa, RSi, Cinp, b0, nb, i0, nci);
      stage_ch(tb, spg, RSo, Coutp, b0, nb, o0, nco);
      __syncthreads();
      for (int idx = threadIdx.y; idx < nci * nco; idx += 8) {
        const int i = idx / nco, o = idx - i * nco;
        float re = 0.f, im = 0.f;
        for (int bb = 0; bb < nb; ++bb) {
          const float xr = ta[mloc * (na + 1) + (bb * 2) * nci + i], xi = ta[mloc * (na + 1) + (bb * 2 + 1) * nci + i];
          const float gr = tb[mloc * (nt + 1) + (bb * 2) * nco + o], gi = tb[mloc * (nt + 1) + (bb * 2 + 1) * nco + o];
          re += xr * gr + xi * gi;
          im += xr * gi - xi * gr;
        }
        if (m < M) {
          float2* dst = gw + ((long long)(i0 + i) * Cout + o0 + o) * LM + (long long)l * M + m;
          if (b0 > 0) {                                 // same thread wrote it in the previous pass
            const float2 prev = *dst;
            re += prev.x;
            im += prev.y;
          }
          *dst = make_float2(re, im);
        }
      }
    }
  }
}

}  // namespace sfno

// ====================================================================== C ABI
using namespace sfno;

// ------------------------------------------------------------------ row scatter (distributed transposition)
constexpr int kMaxScatterDesc = 8;
struct ScatterArgs {
  sfno_copy_desc d[kMaxScatterDesc];
};

// one warp per row; blockIdx.y selects the descriptor (= destination rank).  Rows are short (a latitude slab or a
// channel slab, 0.4 - 0.7 KB), both sides contiguous; stores to a peer-mapped destination travel over NVLink.
// one warp per row; blockIdx.y selects the descriptor (= destination rank).  Rows are short (a latitude slab or a
// channel slab, 0.2 - 1.5 KB), both sides contiguous; stores to a peer-mapped destination travel over NVLink.
// VEC: everything is a multiple of 4 floats -> 16-byte accesses, two loads in flight per lane before the stores.
// (A flat copy over 16-byte chunks with four chunks per thread in flight measured 10 % slower on 2 and 4 GPUs.)
template <bool VEC>
__global__ void __launch_bounds__(256) rows_scatter_kernel(const float* __restrict__ src, const ScatterArgs args) {
  const sfno_copy_desc& d = args.d[blockIdx.y];
  const int n12 = d.n1 * d.n2;
  const int rows = d.n0 * n12;                                  // < 2^31 (checked by the launcher)
  const int lane = threadIdx.x & 31;
  const int nwarps = gridDim.x * 8;
  float* __restrict__ dst = static_cast<float*>(d.dst);
  for (int row = blockIdx.x * 8 + (threadIdx.x >> 5); row < rows; row += nwarps) {
    const int i0 = row / n12, t = row - i0 * n12;
    const int i1 = t / d.n2, i2 = t - i1 * d.n2;
    if (d.tri_l0 >= 0 && i1 >= ((d.tri_l0 + i0 + 16) & ~15)) continue;
    const float* s = src + d.src_off + i0 * d.ss0 + i1 * d.ss1 + i2 * d.ss2;
    float* o = dst + d.dst_off + i0 * d.ds0 + i1 * d.ds1 + i2 * d.ds2;
    if (VEC) {
      for (int e = lane * 4; e < d.len; e += 256) {
        const bool two = e + 128 < d.len;
        const float4 a = *reinterpret_cast<const float4*>(s + e);
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (two) b = *reinterpret_cast<const float4*>(s + e + 128);
        *reinterpret_cast<float4*>(o + e) = a;
        if (two) *reinterpret_cast<float4*>(o + e + 128) = b;
      }
    } else {
      for (int e = lane; e < d.len; e += 32) o[e] = s[e];
    }
  }
}

// ------------------------------------------------------------------ peer-memory barrier
struct BarrierArgs {
  int* flags[kMaxScatterDesc];
};
// epoch > 0: the caller numbers the barriers (legacy).  epoch <= 0: the barrier number lives in device memory - slot
// kEpochSlot of this rank's own flag array is incremented by the kernel itself - so that a CUDA graph that captured the
// launch replays with fresh numbers (a host-side epoch would be baked into the graph: from the second replay on every
// flag already holds it and the barrier would fall through).  All ranks run the same sequence of barriers, so their
// counters stay equal.
constexpr int kEpochSlot = 32;
__global__ void peer_barrier_kernel(const BarrierArgs args, int rank, int size, int epoch) {
  __shared__ int s_epoch;
  if (threadIdx.x == 0) {
    if (epoch <= 0) {
      int* ctr = args.flags[rank] + kEpochSlot;
      epoch = *ctr + 1;
      *ctr = epoch;
    }
    s_epoch = epoch;
  }
  __syncthreads();
  epoch = s_epoch;
  const int p = threadIdx.x;
  if (p >= size) return;
  // everything this GPU did before this kernel (stream order) is published with the flag: system-scope release
  __threadfence_system();
  int* remote = args.flags[p] + rank;
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
  const int* mine = args.flags[rank] + p;
  const long long t0 = clock64();
  int v;
  do {
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
    if (v >= epoch) break;
    __nanosleep(100);
    if (clock64() - t0 > 20000000000LL) __trap();      // ~10 s: a rank never arrived
  } while (true);
  __threadfence_system();
}

extern "C" {

int sfno_sp_to_c64(const float* sp, int B, int C, int lmax, int mmax, int l_offset, int m_offset, void* out_c64,
                   void* stream) {
  dim3 grid((mmax + 31) / 32, (C + 31) / 32, lmax * B), block(32, 8);
  sp_to_c64_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(sp, (float2*)out_c64, B, C, lmax, mmax, l_offset, m_offset);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_c64_to_sp(const void* in_c64, int B, int C, int lmax, int mmax, int l_offset, int m_offset, int round_tf32,
                   float* sp, void* stream) {
  const int Cp = (C + 3) & ~3;
  dim3 grid((mmax + 31) / 32, (Cp + 31) / 32, lmax * B), block(32, 8);
  c64_to_sp_kernel<<<grid, block, 0, (cudaStream_t)stream>>>((const float2*)in_c64, sp, B, C, lmax, mmax, l_offset, m_offset, round_tf32);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_xm_to_c64(const float* xm, int B, int C, int nlat, int mmax, void* out_c64, void* stream) {
  const int Kp = round_up(nlat, 4);
  dim3 grid((mmax + 31) / 32, (nlat + 31) / 32, B * C), block(32, 8);
  xm_to_c64_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(xm, (float2*)out_c64, B, C, nlat, Kp, mmax);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_c64_to_xm(const void* in_c64, int B, int C, int nlat, int mmax, float* xm, void* stream) {
  const int Kp = round_up(nlat, 4);
  dim3 grid((mmax + 31) / 32, (Kp + 31) / 32, B * C), block(32, 8);
  c64_to_xm_kernel<<<grid, block, 0, (cudaStream_t)stream>>>((const float2*)in_c64, xm, B, C, nlat, Kp, mmax);
  SFNO_LAUNCHED();
  return 0;
}
// persistent grids of the weight-layout kernels: SFNO_WLAYOUT_CTAS CTAs per SM (default 4; A/B runs)
static int wlayout_grid(int ntiles) {
  static int per_sm = -1;
  if (per_sm < 0) {
    const char* env = getenv("SFNO_WLAYOUT_CTAS");
    per_sm = env ? std::max(1, atoi(env)) : 4;
  }
  return std::min(ntiles, per_sm * sm_count_current_device());
}
int sfno_dhconv_pack_weight_ex(const void* w_c64, int Cin, int Cout, int L, float* wp, int round_tf32, void* stream) {
  SFNO_CHECK(w_c64 && wp && Cin > 0 && Cout > 0 && L > 0, "sfno_dhconv_pack_weight: bad arguments");
  const int Cinp = (Cin + 3) & ~3;
  dim3 grid((L + kWT - 1) / kWT, (Cinp + kWT - 1) / kWT, Cout);
  pack_weight_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>((const float2*)w_c64, wp, Cin, Cout, L, round_tf32);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_dhconv_pack_weight(const void* w_c64, int Cin, int Cout, int L, float* wp, void* stream) {
  return sfno_dhconv_pack_weight_ex(w_c64, Cin, Cout, L, wp, 1, stream);
}
int sfno_dhconv_unpack_wgrad(const float* gwp, int Cin, int Cout, int L, void* gw_c64, void* stream) {
  SFNO_CHECK(gwp && gw_c64 && Cin > 0 && Cout > 0 && L > 0, "sfno_dhconv_unpack_wgrad: bad arguments");
  const long long nt = (long long)((L + kWT - 1) / kWT) * ((Cin + kWT - 1) / kWT) * Cout;
  SFNO_CHECK(nt < (1LL << 31), "sfno_dhconv_unpack_wgrad: too many tiles");
  WpSource src = {};
  src.bases[0] = gwp;
  unpack_wgrad_kernel<false><<<wlayout_grid((int)nt), 256, 0, (cudaStream_t)stream>>>(src, (float2*)gw_c64, Cin, Cout, L,
                                                                                      (int)nt);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_dhconv_unpack_wgrad_gather(const float* const* gwp_slices_host, int nslices, long long slice_floats, int Cin,
                                    int Cout, int L, void* gw_c64, void* stream) {
  SFNO_CHECK(gwp_slices_host && gw_c64 && nslices >= 1 && nslices <= 8 && slice_floats > 0 && slice_floats % 4 == 0,
             "sfno_dhconv_unpack_wgrad_gather: bad arguments (1..8 slices of a multiple of 4 floats)");
  SFNO_CHECK(slice_floats * nslices >= sfno_wp_elems(L, Cin, Cout), "sfno_dhconv_unpack_wgrad_gather: slices do not cover WP");
  WpSource src = {};
  for (int k = 0; k < 8; ++k) src.bases[k] = gwp_slices_host[k < nslices ? k : 0];
  src.slice = slice_floats;
  const long long nt = (long long)((L + kWT - 1) / kWT) * ((Cin + kWT - 1) / kWT) * Cout;
  SFNO_CHECK(nt < (1LL << 31), "sfno_dhconv_unpack_wgrad_gather: too many tiles");
  // one tile per CTA here: with loads that cross NVLink the persistent form (99 registers with the slice lookup, two
  // CTAs per SM) was slower - 16.3 against 14.7 ms per data-parallel step at 2 GPUs
  unpack_wgrad_kernel<true><<<(int)nt, 256, 0, (cudaStream_t)stream>>>(src, (float2*)gw_c64, Cin, Cout, L, (int)nt);
  SFNO_LAUNCHED();
  return 0;
}
int sfno_add_bias(void* y, int y_dtype, const float* bias, int bias_kind, int B, int C, int H, int W, void* stream) {
  const long long HW = (long long)H * W, n = (long long)B * C * HW;
  const int grid = (int)std::min<long long>((n + 255) / 256, 148 * 16);
  cudaStream_t s = (cudaStream_t)stream;
  switch (y_dtype) {
    case SFNO_F32: add_bias_kernel<float><<<grid, 256, 0, s>>>((float*)y, bias, bias_kind, n, C, HW); break;
    case SFNO_BF16: add_bias_kernel<__nv_bfloat16><<<grid, 256, 0, s>>>((__nv_bfloat16*)y, bias, bias_kind, n, C, HW); break;
    case SFNO_F16: add_bias_kernel<__half><<<grid, 256, 0, s>>>((__half*)y, bias, bias_kind, n, C, HW); break;
    default: SFNO_CHECK(false, "unsupported dtype");
  }
  SFNO_LAUNCHED();
  return 0;
}

int sfno_sp_degree_power(const float* sp, int B, int C, int lmax, int mmax, int l_offset, int m_offset, float* out,
                         void* stream) {
  SFNO_CHECK(sp && out && B > 0 && C > 0 && lmax > 0 && mmax > 0, "sfno_sp_degree_power: bad arguments");
  SFNO_CHECK(lmax <= 65535 && B <= 65535, "sfno_sp_degree_power: lmax and B must fit a grid dimension");
  dim3 grid((C + 31) / 32, lmax, B), block(32, 8);
  sp_degree_power_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(sp, out, B, C, lmax, mmax, l_offset, m_offset);
  SFNO_LAUNCHED();
  return 0;
}

int sfno_sp_degree_power_bwd(const float* sp, const float* g, int B, int C, int lmax, int mmax, int l_offset, int m_offset,
                             int round_tf32, float* gsp, void* stream) {
  SFNO_CHECK(sp && g && gsp && B > 0 && C > 0 && lmax > 0 && mmax > 0, "sfno_sp_degree_power_bwd: bad arguments");
  SFNO_CHECK(B <= 65535, "sfno_sp_degree_power_bwd: B must fit a grid dimension");
  dim3 grid(lmax, B), block(64, 4);
  sp_degree_power_bwd_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(sp, g, gsp, B, C, lmax, mmax, l_offset, m_offset,
                                                                       round_tf32);
  SFNO_LAUNCHED();
  return 0;
}

int sfno_add_inplace(float* dst, const float* src, long long n, void* stream) {
  SFNO_CHECK(dst && src && n % 4 == 0, "sfno_add_inplace: null argument or n % 4 != 0");
  SFNO_CHECK(((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0, "sfno_add_inplace: alignment");
  if (n == 0) return 0;
  const long long n4 = n / 4;
  const int grid = (int)std::min<long long>((n4 + 255) / 256, 148 * 16);
  add_inplace_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(dst), reinterpret_cast<const float4*>(src), n4);
  SFNO_LAUNCHED();
  return 0;
}

int sfno_slices_average(float* dst, const float* const* srcs_host, int nsrc, long long n, float scale, void* stream) {
  SFNO_CHECK(dst && (nsrc == 0 || srcs_host) && nsrc >= 0 && nsrc <= kMaxAvgSrc && n % 4 == 0,
             "sfno_slices_average: bad arguments (nsrc <= 7, n % 4 == 0)");
  if (n == 0) return 0;
  AvgArgs args;
  uintptr_t align = reinterpret_cast<uintptr_t>(dst);
  for (int k = 0; k < kMaxAvgSrc; ++k) {
    args.src[k] = reinterpret_cast<const float4*>(srcs_host[k < nsrc ? k : 0]);
    if (nsrc) align |= reinterpret_cast<uintptr_t>(args.src[k]);
  }
  SFNO_CHECK((align & 15) == 0, "sfno_slices_average: 16-byte alignment");
  const long long n4 = n / 4;
  // a modest grid: this kernel runs on the communication stream next to the backward kernels of the main stream
  const int grid = (int)std::min<long long>((n4 + 255) / 256, 32 * 8);
  slices_average_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<float4*>(dst), args, nsrc, n4, scale);
  SFNO_LAUNCHED();
  return 0;
}

// device-to-device copy that may cross NVLink (src or dst mapped with sfno_ipc_open): runs on a copy engine, no SMs
int sfno_peer_copy(void* dst, const void* src, size_t bytes, void* stream) {
  SFNO_CHECK(dst && src, "null argument");
  if (bytes == 0) return 0;
  SFNO_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, (cudaStream_t)stream));
  return 0;
}

int sfno_xm_add_dc(float* xm, int B, int C, int nlat, const float* bias, void* stream) {
  SFNO_CHECK(xm && bias, "null argument");
  const int n = B * C * nlat;
  xm_add_dc_kernel<<<(n + 255) / 256, 256, 0, (cudaStream_t)stream>>>(xm, bias, B, C, nlat, round_up(nlat, 4));
  SFNO_LAUNCHED();
  return 0;
}

int sfno_diag_contract(int mode, const float* sp_a, const void* w_or_spb, void* out, int B, int Cin, int Cout,
                       int lmax, int mmax, int l_offset, int m_offset, void* stream) {
  SFNO_CHECK(mode >= 0 && mode <= 2, "bad mode");
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const size_t cap = 200 * 1024;
  // chunk sizes (see diag_kernel): as many batch entries per pass as fit (<= kDiagBc), then as many channels
  int Bc = std::min(B, kDiagBc), Cc = 0, nz = 0;
  size_t smem = 0;
  if (mode == 2) {
    Cc = 64;                                       // output channels per CTA: 32*(2*4*128 + 2*4*64 + 2)*4 B = 197 KB
    smem = (size_t)32 * (2 * Bc * kDiagCi + 2 * Bc * Cc + 2) * 4;
    nz = ((Cin + kDiagCi - 1) / kDiagCi) * ((Cout + Cc - 1) / Cc);
  } else {
    const int Credp = mode == 0 ? Cinp : Coutp, Cresp = mode == 0 ? Coutp : Cinp;
    for (; Bc >= 1 && !Cc; --Bc) {
      for (int c = Cresp; c >= 4; c = (c > 8 ? ((c / 2 + 3) & ~3) : c - 4)) {
        smem = (size_t)32 * (2 * Bc * Credp + 2 * Bc * c + 2) * 4;
        if (smem <= cap) { Cc = c; break; }
      }
      if (Cc) break;
    }
    SFNO_CHECK(Cc > 0, "diagonal contraction: channel count too large for the shared-memory tile (> 1590)");
    nz = ((B + Bc - 1) / Bc) * ((Cresp + Cc - 1) / Cc);
  }
  SFNO_CHECK(nz <= 65535, "diagonal contraction: too many chunks");
  dim3 grid((mmax + 31) / 32, lmax, nz), block(32, 8);
  cudaStream_t s = (cudaStream_t)stream;
#define SFNO_DIAG(MODE)                                                                                     \
  SFNO_CUDA(cudaFuncSetAttribute(diag_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)cap)); \
  diag_kernel<MODE><<<grid, block, smem, s>>>(sp_a, w_or_spb, out, B, Cin, Cout, lmax, mmax, l_offset, m_offset, Bc, Cc)
  if (mode == 0) { SFNO_DIAG(0); }
  else if (mode == 1) { SFNO_DIAG(1); }
  else { SFNO_DIAG(2); }
#undef SFNO_DIAG
  SFNO_LAUNCHED();
  return 0;
}

int sfno_ipc_alloc(size_t bytes, void** ptr, unsigned char handle_out[64]) {
  SFNO_CHECK(ptr && handle_out && bytes > 0, "null argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  SFNO_CUDA(cudaMalloc(&p, bytes));
  cudaError_t e = cudaMemset(p, 0, bytes);
  cudaIpcMemHandle_t h;
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    cudaFree(p);
    SFNO_CUDA(e);
  }
  memcpy(handle_out, &h, 64);
  *ptr = p;
  return 0;
}
int sfno_ipc_open(const unsigned char handle[64], void** ptr) {
  SFNO_CHECK(ptr && handle, "null argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  SFNO_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
int sfno_ipc_close(void* ptr) {
  if (ptr) SFNO_CUDA(cudaIpcCloseMemHandle(ptr));
  return 0;
}
int sfno_ipc_free(void* ptr) {
  if (ptr) SFNO_CUDA(cudaFree(ptr));
  return 0;
}

int sfno_rows_scatter(const float* src, const sfno_copy_desc* descs_host, int ndesc, void* stream) {
  SFNO_CHECK(src && descs_host, "null argument");
  SFNO_CHECK(ndesc >= 1 && ndesc <= kMaxScatterDesc, "sfno_rows_scatter: 1..8 descriptors");
  ScatterArgs args;
  long long max_rows = 0;
  bool vec = (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
  for (int i = 0; i < ndesc; ++i) {
    const sfno_copy_desc& d = descs_host[i];
    SFNO_CHECK(d.dst && d.n0 > 0 && d.n1 > 0 && d.n2 > 0 && d.len > 0, "sfno_rows_scatter: bad descriptor");
    args.d[i] = d;
    max_rows = std::max(max_rows, (long long)d.n0 * d.n1 * d.n2);
    SFNO_CHECK((long long)d.n0 * d.n1 * d.n2 < (1LL << 31) - (1LL << 22), "sfno_rows_scatter: too many rows");
    const long long all = d.src_off | d.dst_off | d.ss0 | d.ss1 | d.ss2 | d.ds0 | d.ds1 | d.ds2 | d.len;
    vec = vec && (all & 3) == 0 && (reinterpret_cast<uintptr_t>(d.dst) & 15u) == 0;
  }
  for (int i = ndesc; i < kMaxScatterDesc; ++i) args.d[i] = descs_host[0];
  const int gx = (int)std::max<long long>(1, std::min<long long>((max_rows + 7) / 8, 148LL * 8));
  dim3 grid(gx, ndesc);
  if (vec)
    rows_scatter_kernel<true><<<grid, 256, 0, (cudaStream_t)stream>>>(src, args);
  else
    rows_scatter_kernel<false><<<grid, 256, 0, (cudaStream_t)stream>>>(src, args);
  SFNO_LAUNCHED();
  return 0;
}

// Same copy as sfno_rows_scatter (descriptors without the triangular skip) on the COPY ENGINES: one strided 3-D
// cudaMemcpy3DAsync per descriptor, rows of `len` floats, n2 rows per slice, n0 * n1 slices.  No SM is used, so the
// transfer can overlap kernels of other streams; whether a copy engine moves short rows over NVLink as fast as the
// SM kernel does is measured, not assumed (tools/benchlib/spatial.py).
int sfno_rows_copy_ce(const float* src, const sfno_copy_desc* descs_host, int ndesc, void* stream) {
  SFNO_CHECK(src && descs_host, "null argument");
  SFNO_CHECK(ndesc >= 1 && ndesc <= kMaxScatterDesc, "sfno_rows_copy_ce: 1..8 descriptors");
  for (int i = 0; i < ndesc; ++i) {
    const sfno_copy_desc& d = descs_host[i];
    SFNO_CHECK(d.dst && d.n0 > 0 && d.n1 > 0 && d.n2 > 0 && d.len > 0 && d.tri_l0 < 0, "sfno_rows_copy_ce: bad descriptor");
    SFNO_CHECK(d.ss0 == d.n1 * d.ss1 && d.ds0 == d.n1 * d.ds1 && d.ss1 % d.ss2 == 0 && d.ds1 % d.ds2 == 0 &&
               d.ss2 >= d.len && d.ds2 >= d.len, "sfno_rows_copy_ce: strides do not form a pitched 3-D box");
    cudaMemcpy3DParms p = {};
    p.srcPtr = make_cudaPitchedPtr(const_cast<float*>(src) + d.src_off, (size_t)d.ss2 * 4, (size_t)d.ss2 * 4,
                                   (size_t)(d.ss1 / d.ss2));
    p.dstPtr = make_cudaPitchedPtr(static_cast<float*>(d.dst) + d.dst_off, (size_t)d.ds2 * 4, (size_t)d.ds2 * 4,
                                   (size_t)(d.ds1 / d.ds2));
    p.extent = make_cudaExtent((size_t)d.len * 4, (size_t)d.n2, (size_t)d.n0 * d.n1);
    p.kind = cudaMemcpyDefault;
    SFNO_CUDA(cudaMemcpy3DAsync(&p, (cudaStream_t)stream));
  }
  return 0;
}

int sfno_peer_barrier(int* const* flags_host, int rank, int size, int epoch, void* stream) {
  SFNO_CHECK(flags_host, "null argument");
  SFNO_CHECK(size >= 1 && size <= kMaxScatterDesc && rank >= 0 && rank < size, "sfno_peer_barrier: bad rank / size");
  BarrierArgs args;
  for (int i = 0; i < kMaxScatterDesc; ++i) args.flags[i] = flags_host[i < size ? i : 0];
  peer_barrier_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(args, rank, size, epoch);
  SFNO_LAUNCHED();
  return 0;
}

}  // extern "C"
