// K1 / K5: longitude DFT kernels (real input -> first mmax bins; first mmax bins -> real output).
//
// Layout contract (see include/makani_b200.h):
//   x  : [B*C][nlat][nlon]  (fp32 / bf16 / fp16, longitude contiguous)
//   XM : [mmax][RS][Kp]     RS = 2*B*Cp, row r = (b*2+reim)*Cp + c, latitude contiguous
// A CTA owns 16 consecutive "flat" positions f = c*Kp + lat of one batch entry b, so that for every
// (m, reim) it reads / writes 16 consecutive floats of XM (the transpose between the longitude-major
// field and the latitude-major Legendre operand happens in shared memory).
//
// Two families:
//   * generic  - direct O(W*M) DFT with a shared-memory twiddle table; any nlon (used for odd sizes)
//   * fft480   - register/shuffle FFT for nlon = 480*R, R in {1,3} (the SFNO grids 240x480 and 721x1440)
#include "common.cuh"
#include "plan.h"
#include "fft_lane.cuh"
#include "fft_pass.cuh"

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <map>
#include <mutex>
#include <vector>
#include <cmath>
#include <type_traits>

namespace sfno {

// ------------------------------------------------------------------ twiddle tables (device cache)
static std::map<std::pair<int, int>, float2*> g_tw_cache;
static std::mutex g_tw_mutex;

static int get_twiddles(int W, const float2** out) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(g_tw_mutex);
  auto key = std::make_pair(dev, W);
  auto it = g_tw_cache.find(key);
  if (it != g_tw_cache.end()) {
    *out = it->second;
    return 0;
  }
  std::vector<float2> h(W);
  for (int j = 0; j < W; ++j) {
    double a = 2.0 * M_PI * (double)j / (double)W;
    double c = cos(a), s = sin(a);
    if (4 * j == W) { c = 0; s = 1; }
    if (2 * j == W) { c = -1; s = 0; }
    if (4 * j == 3 * W) { c = 0; s = -1; }
    h[j] = make_float2((float)c, (float)s);
  }
  float2* d = nullptr;
  SFNO_CUDA(cudaMalloc(&d, W * sizeof(float2)));
  SFNO_CUDA(cudaMemcpy(d, h.data(), W * sizeof(float2), cudaMemcpyHostToDevice));
  g_tw_cache[key] = d;
  *out = d;
  return 0;
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }

constexpr int kRows = 16;   // flat positions per CTA tile
constexpr int kGenericSmemMax = 227 * 1024;   // dynamic shared memory the generic kernels may ask for

// ------------------------------------------------------------------ generic forward
template <typename T>
__global__ void __launch_bounds__(256)
dft_fwd_generic_kernel(const T* __restrict__ x, float* __restrict__ xm, const float2* __restrict__ tw,
                       const float* __restrict__ mscale, int B, int C, int K, int Kp, int W, int M, int round_out, long long xbs) {
  extern __shared__ float sm[];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int b = blockIdx.x / tiles_per_b;
  const int f0 = (blockIdx.x % tiles_per_b) * kRows;
  float* xs = sm;                                  // [kRows][W+1]
  float2* tws = reinterpret_cast<float2*>(xs + kRows * (W + 1) + ((kRows * (W + 1)) & 1));   // [W]
  float* outs = reinterpret_cast<float*>(tws + W);  // [2][M][kRows+1]

  for (int j = threadIdx.x; j < W; j += blockDim.x) tws[j] = tw[j];
  for (int r = 0; r < kRows; ++r) {
    const int f = f0 + r;
    const int c = f / Kp, lat = f % Kp;
    const bool valid = (f < flat) && (lat < K);
    const T* row = x + b * xbs + ((long long)c * K + lat) * W;
    for (int w = threadIdx.x; w < W; w += blockDim.x) xs[r * (W + 1) + w] = valid ? to_f32<T>(row[w]) : 0.f;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < kRows * M; idx += blockDim.x) {
    const int r = idx % kRows, m = idx / kRows;
    const float* xr = xs + r * (W + 1);
    float re = 0.f, im = 0.f;
    int t = 0;
    for (int w = 0; w < W; ++w) {
      const float2 cs = tws[t];
      const float v = xr[w];
      re = fmaf(v, cs.x, re);
      im = fmaf(-v, cs.y, im);
      t += m;
      if (t >= W) t -= W;
    }
    const float s = mscale[m];
    outs[(0 * M + m) * (kRows + 1) + r] = re * s;
    outs[(1 * M + m) * (kRows + 1) + r] = im * s;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 2 * M * kRows; idx += blockDim.x) {
    const int r = idx % kRows, pm = idx / kRows;
    const int reim = pm / M, m = pm % M;
    if (f0 + r < flat) {
      const float v = outs[pm * (kRows + 1) + r];
      xm[((long long)m * RS + (long long)(b * 2 + reim) * Cp) * Kp + f0 + r] = round_out ? tf32_rn(v) : v;
    }
  }
}

// ------------------------------------------------------------------ generic inverse
template <typename T>
__global__ void __launch_bounds__(256)
dft_inv_generic_kernel(const float* __restrict__ xm, T* __restrict__ y, const float2* __restrict__ tw,
                       const float* __restrict__ mscale, int B, int C, int K, int Kp, int W, int M, int m_valid, long long xbs) {
  extern __shared__ float sm[];
  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int b = blockIdx.x / tiles_per_b;
  const int f0 = (blockIdx.x % tiles_per_b) * kRows;
  float2* tws = reinterpret_cast<float2*>(sm);       // [W]
  float* ins = reinterpret_cast<float*>(tws + W);    // [2][M][kRows+1]

  for (int j = threadIdx.x; j < W; j += blockDim.x) tws[j] = tw[j];
  for (int idx = threadIdx.x; idx < 2 * m_valid * kRows; idx += blockDim.x) {
    const int r = idx % kRows, pm = idx / kRows;
    const int reim = pm / m_valid, m = pm % m_valid;
    float v = 0.f;
    if (f0 + r < flat) v = xm[((long long)m * RS + (long long)(b * 2 + reim) * Cp) * Kp + f0 + r];
    // irfft weights: the Hermitian pair of bin m doubles every bin except m = 0 and the Nyquist bin
    const float cm = (m == 0 || 2 * m == W) ? 1.f : 2.f;
    ins[(reim * M + m) * (kRows + 1) + r] = v * mscale[m] * cm;
  }
  __syncthreads();
  for (int r = 0; r < kRows; ++r) {
    const int f = f0 + r;
    const int c = f / Kp, lat = f % Kp;
    if (f >= flat || lat >= K) continue;
    T* row = y + b * xbs + ((long long)c * K + lat) * W;
    for (int w = threadIdx.x; w < W; w += blockDim.x) {
      float acc = 0.f;
      int t = 0;
      for (int m = 0; m < m_valid; ++m) {
        const float2 cs = tws[t];
        acc = fmaf(ins[(0 * M + m) * (kRows + 1) + r], cs.x, acc);
        acc = fmaf(-ins[(1 * M + m) * (kRows + 1) + r], cs.y, acc);
        t += w;
        if (t >= W) t -= W;
      }
      row[w] = from_f32<T>(acc);
    }
  }
}

// ------------------------------------------------------------------ fft480 family (nlon = 480 * R)
struct DeviceWarp16 {
  __device__ __forceinline__ int lane16() const { return threadIdx.x & 15; }
  __device__ __forceinline__ float shfl_xor(float v, int mask) const { return __shfl_xor_sync(0xffffffffu, v, mask); }
  __device__ __forceinline__ float shfl_idx(float v, int src) const {
    return __shfl_sync(0xffffffffu, v, (threadIdx.x & 16) | src);
  }
};

// device tables for the fft480 family: tw240[240] | tw480[241] | tw1440[241]   (fl::cf each)
constexpr int kTwCount = 240 + 241 + 241;
static std::map<int, fl::cf*> g_fft_tw_cache;

static int get_fft_tables(const fl::cf** out) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(g_tw_mutex);
  auto it = g_fft_tw_cache.find(dev);
  if (it != g_fft_tw_cache.end()) {
    *out = it->second;
    return 0;
  }
  std::vector<fl::cf> h(kTwCount);
  for (int b = 0; b < 16; ++b)
    for (int k1 = 0; k1 < 15; ++k1) {
      double a = -2.0 * M_PI * (double)(b * k1) / 240.0;
      h[b * 15 + k1] = fl::mk((float)cos(a), (float)sin(a));
    }
  for (int k = 0; k < 241; ++k) {
    double a = -2.0 * M_PI * (double)k / 480.0;
    h[240 + k] = fl::mk((float)cos(a), (float)sin(a));
    double c = -2.0 * M_PI * (double)k / 1440.0;
    h[240 + 241 + k] = fl::mk((float)cos(c), (float)sin(c));
  }
  h[240 + 120] = fl::mk(0.f, -1.f);
  h[240 + 240] = fl::mk(-1.f, 0.f);
  fl::cf* d = nullptr;
  SFNO_CUDA(cudaMalloc(&d, kTwCount * sizeof(fl::cf)));
  SFNO_CUDA(cudaMemcpy(d, h.data(), kTwCount * sizeof(fl::cf), cudaMemcpyHostToDevice));
  g_fft_tw_cache[dev] = d;
  *out = d;
  return 0;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

__device__ __forceinline__ void lane_butterfly_twiddles(fl::cf (&bf)[4]) {
  const int lane = threadIdx.x & 15;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int d = 1 << s;
    float sn, cs;
    sincospif(-(float)(lane & (d - 1)) / (float)d, &sn, &cs);      // e^{-2 pi i j / (2d)}
    bf[s] = fl::mk(cs, sn);
  }
}

constexpr int kStagePitch = 20;   // floats per (m, re/im) staging row: 16 latitudes + 4 pad (16 B aligned, skews banks)
constexpr int kStageRows = 2 * 241;

template <int R, typename T>
struct FftSmem {
  static constexpr int W = 480 * R;
  static constexpr int kRowStride = W + 16 / (int)sizeof(T);         // elements; keeps 16 B alignment, skews banks
  static constexpr int kRawBytes = kRows * kRowStride * (int)sizeof(T);
  static constexpr int kStageBytes = kStageRows * kStagePitch * 4;
  static constexpr int kBufBytes = ((kRawBytes > kStageBytes ? kRawBytes : kStageBytes) + 15) & ~15;
  static constexpr int kBytes = kBufBytes + kTwCount * 8 + 244 * 4 + kRows * 8 + 16;
};

// packed complex sample z_r[n] = (x[2 R n + r], x[2 R n + R + r]) of a row held in shared memory
template <int R, typename T>
__device__ __forceinline__ fl::cf load_pair(const T* row, int n, int r) {
  if constexpr (R == 1) {
    if constexpr (sizeof(T) == 4) {
      const float2 v = *reinterpret_cast<const float2*>(row + 2 * n);
      return fl::mk(v.x, v.y);
    } else if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      const float2 v = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(row + 2 * n));
      return fl::mk(v.x, v.y);
    } else {
      const float2 v = __half22float2(*reinterpret_cast<const __half2*>(row + 2 * n));
      return fl::mk(v.x, v.y);
    }
  } else {
    return fl::mk(to_f32<T>(row[2 * R * n + r]), to_f32<T>(row[2 * R * n + R + r]));
  }
}
template <int R, typename T>
__device__ __forceinline__ void store_pair(T* row, int n, int r, fl::cf z) {
  if constexpr (R == 1) {
    if constexpr (sizeof(T) == 4) {
      *reinterpret_cast<float2*>(row + 2 * n) = make_float2(z.x, z.y);
    } else if constexpr (std::is_same<T, __nv_bfloat16>::value) {
      *reinterpret_cast<__nv_bfloat162*>(row + 2 * n) = __floats2bfloat162_rn(z.x, z.y);
    } else {
      *reinterpret_cast<__half2*>(row + 2 * n) = __floats2half2_rn(z.x, z.y);
    }
  } else {
    row[2 * R * n + r] = from_f32<T>(z.x);
    row[2 * R * n + R + r] = from_f32<T>(z.y);
  }
}

// forward: x rows -> XM.  256 threads = 16 rows x 16 lanes; persistent over tiles of 16 flat positions.
template <int R, typename T>
__global__ void __launch_bounds__(256, 2)
dft_fwd_fft480_kernel(const T* __restrict__ x, float* __restrict__ xm, const fl::cf* __restrict__ tables,
                      const float* __restrict__ mscale, int B, int C, int K, int Kp, int M, int round_out, long long xbs) {
  using S = FftSmem<R, T>;
  constexpr int W = S::W;
  extern __shared__ __align__(16) uint8_t smraw[];
  T* raw = reinterpret_cast<T*>(smraw);
  float* stage = reinterpret_cast<float*>(smraw);                     // aliases raw (used after the FFT)
  fl::cf* tw = reinterpret_cast<fl::cf*>(smraw + S::kBufBytes);
  float* msc = reinterpret_cast<float*>(tw + kTwCount);
  long long* rowoff = reinterpret_cast<long long*>(msc + 244);

  for (int j = threadIdx.x; j < kTwCount; j += blockDim.x) tw[j] = tables[j];
  for (int j = threadIdx.x; j < 241; j += blockDim.x) msc[j] = j < M ? mscale[j] : 0.f;
  fl::cf bf[4];
  lane_butterfly_twiddles(bf);
  const int lane = threadIdx.x & 15;
  const fl::LaneConsts lc = fl::make_lane_consts(lane, bf);

  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = threadIdx.x >> 4;
  const int k2 = fl::bitrev4(lane);
  constexpr int kChunks = W * (int)sizeof(T) / 16;                    // 16-byte chunks per row
  DeviceWarp16 warp;

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    if (threadIdx.x < kRows) {
      const int f = f0 + threadIdx.x;
      const int c = f / Kp, lat = f - c * Kp;
      rowoff[threadIdx.x] = (f < flat && lat < K) ? b * xbs + ((long long)c * K + lat) * W : -1;
    }
    __syncthreads();
    // ---- 1. rows -> smem (coalesced 16 B async copies); padded / out-of-range rows are zero filled
    for (int idx = threadIdx.x; idx < kRows * kChunks; idx += blockDim.x) {
      const int rr = idx / kChunks, ch = idx - rr * kChunks;
      uint8_t* dst = reinterpret_cast<uint8_t*>(raw + rr * S::kRowStride) + ch * 16;
      const long long off = rowoff[rr];
      if (off >= 0) cp_async16(dst, reinterpret_cast<const uint8_t*>(x + off) + ch * 16);
      else *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
    }
    cp_async_wait_all();
    __syncthreads();
    // ---- 2. per-row FFT in registers / shuffles
    fl::cf X[15], x240;
    const T* myrow = raw + r * S::kRowStride;
    fl::rfft_row_fwd<R>(warp, [&](int n, int rr) { return load_pair<R, T>(myrow, n, rr); }, X, x240, tw, tw + 240,
                        tw + 481, lc);
    __syncthreads();
    // ---- 3. transpose through smem: stage[reim][m][row]
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      const int m = k1 + 15 * k2;
      const float sc = msc[m];
      stage[m * kStagePitch + r] = X[k1].x * sc;
      stage[(241 + m) * kStagePitch + r] = X[k1].y * sc;
    }
    if (lane == 0) {
      const float sc = msc[240];
      stage[240 * kStagePitch + r] = x240.x * sc;
      stage[(241 + 240) * kStagePitch + r] = x240.y * sc;
    }
    __syncthreads();
    const long long obase = ((long long)(b * 2) * Cp) * Kp + f0;
    for (int idx = threadIdx.x; idx < kStageRows * 4; idx += blockDim.x) {
      const int pm = idx >> 2, q = idx & 3;
      const int reim = pm >= 241 ? 1 : 0;
      const int m = pm - reim * 241;
      if (m < M && f0 + 4 * q < flat) {
        float4 v = *reinterpret_cast<const float4*>(stage + pm * kStagePitch + 4 * q);
        if (round_out) { v.x = tf32_rn(v.x); v.y = tf32_rn(v.y); v.z = tf32_rn(v.z); v.w = tf32_rn(v.w); }
        *reinterpret_cast<float4*>(xm + ((long long)m * RS + (long long)reim * Cp) * Kp + obase + 4 * q) = v;
      }
    }
    __syncthreads();
  }
}

// inverse: XM -> y rows.
template <int R, typename T>
__global__ void __launch_bounds__(256, 2)
dft_inv_fft480_kernel(const float* __restrict__ xm, T* __restrict__ y, const fl::cf* __restrict__ tables,
                      const float* __restrict__ mscale, int B, int C, int K, int Kp, int M, int m_valid, long long xbs) {
  using S = FftSmem<R, T>;
  constexpr int W = S::W;
  extern __shared__ __align__(16) uint8_t smraw[];
  T* raw = reinterpret_cast<T*>(smraw);
  float* stage = reinterpret_cast<float*>(smraw);
  fl::cf* tw = reinterpret_cast<fl::cf*>(smraw + S::kBufBytes);
  float* msc = reinterpret_cast<float*>(tw + kTwCount);
  long long* rowoff = reinterpret_cast<long long*>(msc + 244);

  for (int j = threadIdx.x; j < kTwCount; j += blockDim.x) tw[j] = tables[j];
  for (int j = threadIdx.x; j < 241; j += blockDim.x) msc[j] = j < m_valid ? mscale[j] : 0.f;
  fl::cf bf[4];
  lane_butterfly_twiddles(bf);
  const int lane = threadIdx.x & 15;
  const fl::LaneConsts lc = fl::make_lane_consts(lane, bf);

  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = threadIdx.x >> 4;
  const int k2 = fl::bitrev4(lane);
  constexpr int kChunks = W * (int)sizeof(T) / 16;
  DeviceWarp16 warp;

  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    if (threadIdx.x < kRows) {
      const int f = f0 + threadIdx.x;
      const int c = f / Kp, lat = f - c * Kp;
      rowoff[threadIdx.x] = (f < flat && lat < K) ? b * xbs + ((long long)c * K + lat) * W : -1;
    }
    // ---- 1. XM -> stage[reim][m][16 latitudes]: 16 B async copies, all in flight at once
    const long long ibase = ((long long)(b * 2) * Cp) * Kp + f0;
    for (int idx = threadIdx.x; idx < kStageRows * 4; idx += blockDim.x) {
      const int pm = idx >> 2, q = idx & 3;
      const int reim = pm >= 241 ? 1 : 0;
      const int m = pm - reim * 241;
      float* dst = stage + pm * kStagePitch + 4 * q;
      if (m < m_valid && f0 + 4 * q < flat)
        cp_async16(dst, xm + ((long long)m * RS + (long long)reim * Cp) * Kp + ibase + 4 * q);
      else
        *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    cp_async_wait_all();
    __syncthreads();
    // ---- 2. gather this lane's bins (scaled) into registers, then the stage buffer may be overwritten
    fl::cf H[15], HP[15];
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      const int k = k1 + 15 * k2, kp = 240 - k;
      const float s0 = msc[k], s1 = msc[kp];
      H[k1] = fl::mk(stage[k * kStagePitch + r] * s0, stage[(241 + k) * kStagePitch + r] * s0);
      HP[k1] = fl::mk(stage[kp * kStagePitch + r] * s1, stage[(241 + kp) * kStagePitch + r] * s1);
    }
    __syncthreads();
    T* myrow = raw + r * S::kRowStride;
    fl::irfft_row<R>(warp, H, HP, [&](int n, int rr, fl::cf z) { store_pair<R, T>(myrow, n, rr, z); }, tw, tw + 240,
                     tw + 481, lc);
    __syncthreads();
    // ---- 3. rows -> global, 16 B vectors
    for (int idx = threadIdx.x; idx < kRows * kChunks; idx += blockDim.x) {
      const int rr = idx / kChunks, ch = idx - rr * kChunks;
      const long long off = rowoff[rr];
      if (off >= 0) {
        const uint4 v = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(raw + rr * S::kRowStride) + ch * 16);
        *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(y + off) + ch * 16) = v;
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------ host dispatch
static int sm_count() { return sm_count_current_device(); }

}  // namespace sfno
#include "dft_p2.cuh"
namespace sfno {

template <int R, typename T>
static int launch_fwd_fft(const void* x, int B, int C, int Ctot, int K, int Kp, int M, const float* mscale, float* xm,
                          int round_out, cudaStream_t stream) {
  using S = FftSmem<R, T>;
  const fl::cf* tables;
  if (get_fft_tables(&tables)) return 1;
  static PerDeviceOnce attr;
  if (attr.needed()) {
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_fft480_kernel<R, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr.mark();
  }
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  const int grid = std::min(ntiles, 2 * sm_count());
  dft_fwd_fft480_kernel<R, T><<<grid, 256, S::kBytes, stream>>>(static_cast<const T*>(x), xm, tables, mscale, B, C, K, Kp, M, round_out, (long long)Ctot * K * S::W);
  SFNO_LAUNCHED();
  return 0;
}
template <int R, typename T>
static int launch_inv_fft(const float* xm, int B, int C, int Ctot, int K, int Kp, int M, int m_valid, const float* mscale,
                          void* y, cudaStream_t stream) {
  using S = FftSmem<R, T>;
  const fl::cf* tables;
  if (get_fft_tables(&tables)) return 1;
  static PerDeviceOnce attr;
  if (attr.needed()) {
    SFNO_CUDA(cudaFuncSetAttribute(dft_inv_fft480_kernel<R, T>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr.mark();
  }
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  const int grid = std::min(ntiles, 2 * sm_count());
  dft_inv_fft480_kernel<R, T><<<grid, 256, S::kBytes, stream>>>(xm, static_cast<T*>(y), tables, mscale, B, C, K, Kp, M, m_valid, (long long)Ctot * K * S::W);
  SFNO_LAUNCHED();
  return 0;
}

template <typename T>
static int launch_fwd_generic(const void* x, int B, int C, int Ctot, int K, int Kp, int W, int M, const float* mscale, float* xm,
                              int round_out, cudaStream_t stream) {
  const float2* tw;
  if (get_twiddles(W, &tw)) return 1;
  const size_t smem = (size_t)(kRows * (W + 1) + 2 * W + 2 * M * (kRows + 1)) * 4 + 16;
  // 227 KB is the sm_100 limit per CTA: nlon = 1440 with all 721 modes (the spectral loss's transform) needs 202 KB
  SFNO_CHECK(smem <= kGenericSmemMax, "generic DFT: nlon too large for the shared-memory kernel");
  SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGenericSmemMax));
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  dft_fwd_generic_kernel<T><<<ntiles, 256, smem, stream>>>(static_cast<const T*>(x), xm, tw, mscale, B, C, K, Kp, W, M, round_out, (long long)Ctot * K * W);
  SFNO_LAUNCHED();
  return 0;
}
template <typename T>
static int launch_inv_generic(const float* xm, int B, int C, int Ctot, int K, int Kp, int W, int M, int m_valid,
                              const float* mscale, void* y, cudaStream_t stream) {
  const float2* tw;
  if (get_twiddles(W, &tw)) return 1;
  const size_t smem = (size_t)(2 * W + 2 * M * (kRows + 1)) * 4 + 16;
  SFNO_CHECK(smem <= kGenericSmemMax, "generic inverse DFT: sizes too large for the shared-memory kernel");
  SFNO_CUDA(cudaFuncSetAttribute(dft_inv_generic_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGenericSmemMax));
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  dft_inv_generic_kernel<T><<<ntiles, 256, smem, stream>>>(xm, static_cast<T*>(y), tw, mscale, B, C, K, Kp, W, M, m_valid, (long long)Ctot * K * W);
  SFNO_LAUNCHED();
  return 0;
}

// SFNO_DFT_IMPL=lane selects the older 16-lane shuffle FFT kernels (kept for A/B measurements)
static bool use_lane_fft() {
  static int v = -1;
  if (v < 0) {
    const char* env = getenv("SFNO_DFT_IMPL");
    v = (env && std::string(env) == "lane") ? 1 : 0;
  }
  return v == 1;
}

static bool use_fft480(int W, int M, const void* p) {
  const char* env = getenv("SFNO_FORCE_GENERIC_DFT");
  if (env && env[0] == '1') return false;
  return (W == 480 || W == 1440) && M <= 241 && (reinterpret_cast<uintptr_t>(p) & 15) == 0;
}

int dft_fwd(const void* x, int x_dtype, int B, int C, int Ctot, int nlat, int nlon, int mmax, const float* mscale,
            float* xm, int round_out, cudaStream_t stream) {
  SFNO_CHECK(mmax <= nlon / 2 + 1, "mmax exceeds nlon/2+1");
  SFNO_CHECK(Ctot >= C, "channel window larger than the field");
  const int Kp = round_up(nlat, 4);
  const bool fast = use_fft480(nlon, mmax, x);
  const bool lane_fft = use_lane_fft();
#define SFNO_DISPATCH_FWD(T)                                                                           \
  do {                                                                                                 \
    if (fast && lane_fft && nlon == 480) return launch_fwd_fft<1, T>(x, B, C, Ctot, nlat, Kp, mmax, mscale, xm, round_out, stream); \
    if (fast && lane_fft && nlon == 1440) return launch_fwd_fft<3, T>(x, B, C, Ctot, nlat, Kp, mmax, mscale, xm, round_out, stream);\
    if (fast && nlon == 480) return launch_fwd_p2<1, T>(x, B, C, Ctot, nlat, Kp, mmax, mscale, xm, round_out, stream); \
    if (fast && nlon == 1440) return launch_fwd_p2<3, T>(x, B, C, Ctot, nlat, Kp, mmax, mscale, xm, round_out, stream);\
    return launch_fwd_generic<T>(x, B, C, Ctot, nlat, Kp, nlon, mmax, mscale, xm, round_out, stream);                   \
  } while (0)
  switch (x_dtype) {
    case SFNO_F32: SFNO_DISPATCH_FWD(float);
    case SFNO_BF16: SFNO_DISPATCH_FWD(__nv_bfloat16);
    case SFNO_F16: SFNO_DISPATCH_FWD(__half);
    default: SFNO_CHECK(false, "unsupported input dtype");
  }
#undef SFNO_DISPATCH_FWD
}

int dft_inv(const float* xm, int B, int C, int Ctot, int nlat, int nlon, int mmax, int m_valid, const float* mscale,
            void* y, int y_dtype, cudaStream_t stream) {
  SFNO_CHECK(mmax <= nlon / 2 + 1, "mmax exceeds nlon/2+1");
  SFNO_CHECK(Ctot >= C, "channel window larger than the field");
  const int Kp = round_up(nlat, 4);
  m_valid = std::min(m_valid, mmax);
  const bool fast = use_fft480(nlon, mmax, y);
  const bool lane_fft = use_lane_fft();
#define SFNO_DISPATCH_INV(T)                                                                                    \
  do {                                                                                                          \
    if (fast && lane_fft && nlon == 480) return launch_inv_fft<1, T>(xm, B, C, Ctot, nlat, Kp, mmax, m_valid, mscale, y, stream); \
    if (fast && lane_fft && nlon == 1440) return launch_inv_fft<3, T>(xm, B, C, Ctot, nlat, Kp, mmax, m_valid, mscale, y, stream);\
    if (fast && nlon == 480) return launch_inv_p2<1, T>(xm, B, C, Ctot, nlat, Kp, mmax, m_valid, mscale, y, stream); \
    if (fast && nlon == 1440) return launch_inv_p2<3, T>(xm, B, C, Ctot, nlat, Kp, mmax, m_valid, mscale, y, stream);\
    return launch_inv_generic<T>(xm, B, C, Ctot, nlat, Kp, nlon, mmax, m_valid, mscale, y, stream);                   \
  } while (0)
  switch (y_dtype) {
    case SFNO_F32: SFNO_DISPATCH_INV(float);
    case SFNO_BF16: SFNO_DISPATCH_INV(__nv_bfloat16);
    case SFNO_F16: SFNO_DISPATCH_INV(__half);
    default: SFNO_CHECK(false, "unsupported output dtype");
  }
#undef SFNO_DISPATCH_INV
}

// ---- nlon = 1440 with more than 241 modes (the spectral loss keeps all 721, makani/utils/losses.py:291) ----
// The tiled kernels deliver bins k = 0..240 of a row from three length-480 transforms X_r:  X[k] = sum_r t(k)^r X_r[k],
// t(k) = e^{-2 pi i k / 1440}.  X_r has period 480 and is Hermitian, so with w = e^{-2 pi i / 3}
//     X[480 + k] = sum_r (w t)^r X_r[k]                     band 2: the same kernel with its t-table multiplied by w
//     X[480 - k] = conj( sum_r (conj(w) t)^r X_r[k] )       band 1: table times conj(w), then mirrored and conjugated
// and, by linearity, the inverse is the sum of three inverse launches on the rows of each band (the caller passes the
// weights: bin 480 enters band 2 as its k = 0, which the kernel counts once - so twice there - and the Nyquist bin 720
// as k = 240, which the kernel doubles - so half).  `xm_rows` points at the first bin of the band, `mcount` bins.
int dft_fwd_band(const void* x, int x_dtype, int B, int C, int Ctot, int nlat, int band, int mcount, const float* mscale,
                 float* xm_rows, int round_out, cudaStream_t stream) {
  SFNO_CHECK(band >= 0 && band <= 2 && mcount >= 1 && mcount <= 241, "dft band: band 0..2, 1..241 bins");
  SFNO_CHECK(Ctot >= C, "channel window larger than the field");
  SFNO_CHECK(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(xm_rows)) & 15) == 0, "dft band: alignment");
  const int Kp = round_up(nlat, 4);
  switch (x_dtype) {
    case SFNO_F32: return launch_fwd_p2<3, float>(x, B, C, Ctot, nlat, Kp, mcount, mscale, xm_rows, round_out, stream, band);
    case SFNO_BF16: return launch_fwd_p2<3, __nv_bfloat16>(x, B, C, Ctot, nlat, Kp, mcount, mscale, xm_rows, round_out, stream, band);
    case SFNO_F16: return launch_fwd_p2<3, __half>(x, B, C, Ctot, nlat, Kp, mcount, mscale, xm_rows, round_out, stream, band);
    default: SFNO_CHECK(false, "unsupported input dtype");
  }
}

int dft_inv_band(const float* xm_rows, int B, int C, int Ctot, int nlat, int band, int mcount, const float* mscale, void* y,
                 int y_dtype, cudaStream_t stream) {
  SFNO_CHECK(band >= 0 && band <= 2 && mcount >= 1 && mcount <= 241, "dft band: band 0..2, 1..241 bins");
  SFNO_CHECK(Ctot >= C, "channel window larger than the field");
  SFNO_CHECK(((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(xm_rows)) & 15) == 0, "dft band: alignment");
  const int Kp = round_up(nlat, 4);
  switch (y_dtype) {
    case SFNO_F32: return launch_inv_p2<3, float>(xm_rows, B, C, Ctot, nlat, Kp, mcount, mcount, mscale, y, stream, band);
    case SFNO_BF16: return launch_inv_p2<3, __nv_bfloat16>(xm_rows, B, C, Ctot, nlat, Kp, mcount, mcount, mscale, y, stream, band);
    case SFNO_F16: return launch_inv_p2<3, __half>(xm_rows, B, C, Ctot, nlat, Kp, mcount, mcount, mscale, y, stream, band);
    default: SFNO_CHECK(false, "unsupported output dtype");
  }
}

// dst row i = conj(src row nrows-1-i) * scale[i]: rows are bins of an XM tensor ([2B planes][plane_floats], odd planes
// hold the imaginary parts).  grid (chunks, nrows), 256 threads, float4
__global__ void __launch_bounds__(256) xm_mirror_conj_kernel(const float4* __restrict__ src, float4* __restrict__ dst,
                                                             int nrows, int nplanes, long long plane4,
                                                             const float* __restrict__ scale, int round_out) {
  const int i = blockIdx.y;
  const long long row4 = plane4 * nplanes;
  const float s = scale ? scale[i] : 1.f;
  const float4* sr = src + (long long)(nrows - 1 - i) * row4;
  float4* dr = dst + (long long)i * row4;
  for (long long e = blockIdx.x * 256LL + threadIdx.x; e < row4; e += gridDim.x * 256LL) {
    const float f = ((e / plane4) & 1) ? -s : s;
    float4 v = sr[e];
    v.x *= f; v.y *= f; v.z *= f; v.w *= f;
    if (round_out) { v.x = tf32_rn(v.x); v.y = tf32_rn(v.y); v.z = tf32_rn(v.z); v.w = tf32_rn(v.w); }
    dr[e] = v;
  }
}

int xm_mirror_conj(const float* src, float* dst, int nrows, int nplanes, long long plane_floats, const float* scale,
                   int round_out, cudaStream_t stream) {
  SFNO_CHECK(src && dst && nrows >= 0 && nplanes > 0 && plane_floats > 0 && plane_floats % 4 == 0,
             "xm_mirror_conj: bad arguments (plane_floats % 4 == 0)");
  SFNO_CHECK(((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0, "xm_mirror_conj: alignment");
  if (nrows == 0) return 0;
  const long long row4 = plane_floats / 4 * nplanes;
  const int chunks = (int)std::min<long long>((row4 + 255) / 256, 64);
  xm_mirror_conj_kernel<<<dim3(chunks, nrows), 256, 0, stream>>>(reinterpret_cast<const float4*>(src),
                                                                 reinterpret_cast<float4*>(dst), nrows, nplanes,
                                                                 plane_floats / 4, scale, round_out);
  SFNO_LAUNCHED();
  return 0;
}

// ---- h-sharded transforms: the DFT kernels address the channel-sharded XM_ch buffers of all ranks (peer memory)
bool dft_peer_supported(int nlon, int mmax) {
  const char* env = getenv("SFNO_FORCE_GENERIC_DFT");
  if (env && env[0] == '1') return false;
  return (nlon == 480 || nlon == 1440) && mmax <= 241 && !use_lane_fft();
}

int dft_fwd_peer(const void* x, int x_dtype, int B, int C, int nlat_loc, int nlon, int mmax, const float* mscale,
                 void* const* xm_ch, const int* chan_shapes, int nranks, int Kg, int lat_off, int lat_span,
                 int round_out, cudaStream_t stream) {
  SFNO_CHECK(dft_peer_supported(nlon, mmax), "peer DFT: only the nlon = 480 / 1440 kernels (mmax <= 241)");
  SFNO_CHECK((reinterpret_cast<uintptr_t>(x) & 15) == 0, "peer DFT: input rows must be 16-byte aligned");
  PeerMaps pm;
  if (make_peer_maps(&pm, xm_ch, chan_shapes, nranks, B, Kg, lat_off, lat_span, mmax)) return 1;
  SFNO_CHECK(pm.c0[nranks] == C, "peer DFT: channel slabs do not add up to C");
  const int pitch = round_up(nlat_loc, 16);
#define SFNO_DISPATCH_FWD_PEER(T)                                                                                   \
  do {                                                                                                              \
    if (nlon == 480) return launch_fwd_p2_maps<1, T, PeerMaps>(x, B, C, C, nlat_loc, pitch, mmax, mscale, pm, round_out, stream); \
    return launch_fwd_p2_maps<3, T, PeerMaps>(x, B, C, C, nlat_loc, pitch, mmax, mscale, pm, round_out, stream);       \
  } while (0)
  switch (x_dtype) {
    case SFNO_F32: SFNO_DISPATCH_FWD_PEER(float);
    case SFNO_BF16: SFNO_DISPATCH_FWD_PEER(__nv_bfloat16);
    case SFNO_F16: SFNO_DISPATCH_FWD_PEER(__half);
    default: SFNO_CHECK(false, "unsupported input dtype");
  }
#undef SFNO_DISPATCH_FWD_PEER
}

int dft_inv_peer(void* const* xm_ch, const int* chan_shapes, int nranks, int Kg, int lat_off, int lat_span, int B,
                 int C, int nlat_loc, int nlon, int mmax, int m_valid, const float* mscale, void* y, int y_dtype,
                 cudaStream_t stream) {
  SFNO_CHECK(dft_peer_supported(nlon, mmax), "peer DFT: only the nlon = 480 / 1440 kernels (mmax <= 241)");
  SFNO_CHECK((reinterpret_cast<uintptr_t>(y) & 15) == 0, "peer DFT: output rows must be 16-byte aligned");
  m_valid = std::max(1, std::min(m_valid, mmax));
  PeerMaps pm;
  if (make_peer_maps(&pm, xm_ch, chan_shapes, nranks, B, Kg, lat_off, lat_span, m_valid)) return 1;
  SFNO_CHECK(pm.c0[nranks] == C, "peer DFT: channel slabs do not add up to C");
  const int pitch = round_up(nlat_loc, 16);
#define SFNO_DISPATCH_INV_PEER(T)                                                                                   \
  do {                                                                                                              \
    if (nlon == 480) return launch_inv_p2_maps<1, T, PeerMaps>(pm, B, C, C, nlat_loc, pitch, m_valid, mscale, y, stream); \
    return launch_inv_p2_maps<3, T, PeerMaps>(pm, B, C, C, nlat_loc, pitch, m_valid, mscale, y, stream);               \
  } while (0)
  switch (y_dtype) {
    case SFNO_F32: SFNO_DISPATCH_INV_PEER(float);
    case SFNO_BF16: SFNO_DISPATCH_INV_PEER(__nv_bfloat16);
    case SFNO_F16: SFNO_DISPATCH_INV_PEER(__half);
    default: SFNO_CHECK(false, "unsupported output dtype");
  }
#undef SFNO_DISPATCH_INV_PEER
}

}  // namespace sfno
