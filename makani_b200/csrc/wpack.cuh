// dhconv weight pack / gradient unpack with a 64 (i) x TL (l) shared tile, TL in {64, 32}.
//
//   w [Cin][Cout][L] complex64   <->   WP [L][2][Coutp][Cinp] planar fp32 (TF32-rounded on the way in)
//
// The kernels in layout.cu use a 64 x 64 tile = 32 KB of static shared memory.  That is fine when they run alone, but
// they are pure streaming kernels that sit between latency-bound persistent kernels (the tcgen05 GEMMs occupy ~198 KB
// of an SM's 228 KB and leave DRAM 40 % idle): to run beside them on a side stream a CTA has to fit into what is left.
// TL = 32 halves the tile to 16 KB (complex side: 256-byte runs instead of 512; planar side unchanged).
// Selected at run time with SFNO_WPACK_TL=32 (default: the layout.cu kernels) - to be measured, DESIGN.md 7.1 item 4.
//
// The per-thread bodies are host + device functions (WP_HD) over a tile passed by reference, so that
// tests/cpp/wpack_host_test.cpp can run the exact index arithmetic on the CPU (256 "threads" per tile, phase 1 for all
// of them, then phase 2 - the kernels have exactly one barrier) against a plain double loop.
#pragma once
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#include <cuda_runtime.h>
#define WP_HD __host__ __device__ __forceinline__
#else
#include <vector_types.h>
#include <vector_functions.h>
#define WP_HD inline
#endif

namespace sfno {
namespace wpack {

constexpr int kTI = 64;        // input channels per tile

template <int TL>
struct Tile {
  float tr[kTI][TL];
  float ti[kTI][TL];
};

// tile element (i, l) lives in column l ^ (i >> 2): conflict free for lanes along l (fixed i) and for lanes along i in
// steps of 4 (fixed l).  i >> 2 < 16 <= TL, so the column stays inside the row for TL = 32 and 64.
WP_HD int col(int i, int l) { return l ^ (i >> 2); }

WP_HD float round_tf32(float x) {
#if defined(__CUDA_ARCH__)
  uint32_t r;                                   // the destination of cvt.rna.tf32.f32 is a .b32 register
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
#else
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7f800000u) == 0x7f800000u) return x;
  u = (u + 0x1000u) & 0xffffe000u;              // round to nearest, ties away: what cvt.rna does for finite values
  memcpy(&x, &u, 4);
  return x;
#endif
}

// ---------------------------------------------------------------------------------------------- pack
// phase 1: 256 threads read a 64 (i) x TL (l) tile of output channel o from w (l contiguous: TL * 8-byte runs)
template <int TL>
WP_HD void pack_load(int tid, Tile<TL>& t, const float2* w, int o, int l0, int i0, int Cin, int Cout, int L) {
  constexpr int RPP = 256 / TL;                 // tile rows (i) covered per pass
  constexpr int NP = kTI / RPP;
  const int ll = tid % TL, ib = tid / TL;
  float2 v[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int i = i0 + ib + RPP * p, l = l0 + ll;
    v[p] = (i < Cin && l < L) ? w[((long long)i * Cout + o) * L + l] : make_float2(0.f, 0.f);
  }
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int il = ib + RPP * p;
    t.tr[il][col(il, ll)] = round_tf32(v[p].x);  // operands of TF32 GEMMs: pre-rounded (pads are exact zeros)
    t.ti[il][col(il, ll)] = round_tf32(v[p].y);
  }
}
// phase 2: write 2 * TL rows (l, re/im) of 64 channels each (16 lanes x float4 = 256 bytes per row)
template <int TL>
WP_HD void pack_store(int tid, const Tile<TL>& t, float* wp, int o, int l0, int i0, int Cinp, int Coutp, int L) {
  constexpr int NP = 2 * TL / 16;
  const int iq = (tid & 15) * 4, lb = tid >> 4;
  const bool iok = i0 + iq < Cinp;              // Cinp is a multiple of 4: a vector is entirely in or out
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int row = lb + 16 * p;                // row = 2 * l_local + reim
    const int ll = row >> 1, reim = row & 1;
    const int l = l0 + ll;
    const float (*a)[TL] = reim ? t.ti : t.tr;
    const int cl = col(iq, ll);                 // iq .. iq+3 share i >> 2
    const float4 v = make_float4(a[iq][cl], a[iq + 1][cl], a[iq + 2][cl], a[iq + 3][cl]);
    if (iok && l < L)
      *reinterpret_cast<float4*>(wp + (((long long)l * 2 + reim) * Coutp + o) * Cinp + i0 + iq) = v;
  }
}

// ---------------------------------------------------------------------------------------------- unpack
template <int TL>
WP_HD void unpack_load(int tid, Tile<TL>& t, const float* gwp, int o, int l0, int i0, int Cinp, int Coutp, int L) {
  constexpr int NP = 2 * TL / 16;
  const int iq = (tid & 15) * 4, lb = tid >> 4;
  const bool iok = i0 + iq < Cinp;
  float4 v[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int row = lb + 16 * p;
    const int l = l0 + (row >> 1), reim = row & 1;
    v[p] = (iok && l < L)
               ? *reinterpret_cast<const float4*>(gwp + (((long long)l * 2 + reim) * Coutp + o) * Cinp + i0 + iq)
               : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    const int row = lb + 16 * p;
    const int ll = row >> 1, reim = row & 1;
    float (*a)[TL] = reim ? t.ti : t.tr;
    const int cl = col(iq, ll);
    a[iq][cl] = v[p].x;
    a[iq + 1][cl] = v[p].y;
    a[iq + 2][cl] = v[p].z;
    a[iq + 3][cl] = v[p].w;
  }
}
template <int TL>
WP_HD void unpack_store(int tid, const Tile<TL>& t, float2* gw, int o, int l0, int i0, int Cin, int Cout, int L) {
  if ((L & 1) == 0) {
    // two degrees per thread: 16-byte stores, TL / 2 lanes write TL * 8 contiguous bytes of one (i, o) row
    constexpr int LPR = TL / 2;                 // lanes per row
    constexpr int RPP = 256 / LPR;
    constexpr int NP = kTI / RPP;
    const int lp = (tid % LPR) * 2, ib = tid / LPR;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int il = ib + RPP * p;
      const int i = i0 + il, l = l0 + lp;
      if (i < Cin && l < L) {                   // L even and l even: l + 1 < L as well
        const int c0 = col(il, lp), c1 = col(il, lp + 1);
        *reinterpret_cast<float4*>(gw + ((long long)i * Cout + o) * L + l) =
            make_float4(t.tr[il][c0], t.ti[il][c0], t.tr[il][c1], t.ti[il][c1]);
      }
    }
  } else {
    constexpr int RPP = 256 / TL;
    constexpr int NP = kTI / RPP;
    const int ll = tid % TL, ib = tid / TL;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const int il = ib + RPP * p;
      const int i = i0 + il, l = l0 + ll;
      if (i < Cin && l < L)
        gw[((long long)i * Cout + o) * L + l] = make_float2(t.tr[il][col(il, ll)], t.ti[il][col(il, ll)]);
    }
  }
}

#if defined(__CUDACC__)
// grid (ceil(L / TL), ceil(Cinp / 64), Cout), block 256
template <int TL>
__global__ void __launch_bounds__(256)
pack_weight_tl_kernel(const float2* __restrict__ w, float* __restrict__ wp, int Cin, int Cout, int L) {
  __shared__ Tile<TL> t;
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const int o = blockIdx.z, l0 = blockIdx.x * TL, i0 = blockIdx.y * kTI;
  pack_load<TL>(threadIdx.x, t, w, o, l0, i0, Cin, Cout, L);
  __syncthreads();
  pack_store<TL>(threadIdx.x, t, wp, o, l0, i0, Cinp, Coutp, L);
}
template <int TL>
__global__ void __launch_bounds__(256)
unpack_wgrad_tl_kernel(const float* __restrict__ gwp, float2* __restrict__ gw, int Cin, int Cout, int L) {
  __shared__ Tile<TL> t;
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  const int o = blockIdx.z, l0 = blockIdx.x * TL, i0 = blockIdx.y * kTI;
  unpack_load<TL>(threadIdx.x, t, gwp, o, l0, i0, Cinp, Coutp, L);
  __syncthreads();
  unpack_store<TL>(threadIdx.x, t, gw, o, l0, i0, Cin, Cout, L);
}
#endif

}  // namespace wpack
}  // namespace sfno
