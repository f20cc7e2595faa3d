// Longitude DFT kernels for nlon = 480*R built on the two-pass FFT of fft_pass.cuh  (included by dft.cu).
//
//   tile      16 consecutive flat positions f = c*Kp + lat of one batch entry; 256 threads = 16 rows x 16 lanes
//   rows      move with 1-D bulk async copies (cp.async.bulk, one per row, completion on an mbarrier),
//             double buffered: the rows of tile i+1 (i+2) are in flight while tile i is transformed
//   XM side   goes through a swizzled [2*241][16] staging tile so that both the per-lane accesses of the FFT
//             (lane = bin) and the 16-byte global accesses (4 consecutive latitudes) are (nearly) conflict free
//   work      the FFT exchange buffer aliases the staging tile (separated by __syncthreads)
#pragma once

namespace sfno {

struct DeviceWarp16p {
  __device__ __forceinline__ int lane16() const { return threadIdx.x & 15; }
  __device__ __forceinline__ float shfl_idx(float v, int src) const {
    return __shfl_sync(0xffffffffu, v, (threadIdx.x & 16) | src);
  }
  __device__ __forceinline__ void sync() const { __syncwarp(); }
};

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// staging tile: float index of (pm = reim*241 + m, row r).  The 16-byte chunk (4 rows) is XOR-swizzled with pm >> 1:
// 16 lanes with consecutive pm then spread over 2 (pm & 1: 64-byte row parity) x 4 (chunk) bank groups, i.e. the
// per-lane 4-byte accesses of the FFT side are 2-way conflicted (4-way with a pm & 3 swizzle, 16-way without).
__device__ __forceinline__ int stage_idx(int pm, int r) {
  return pm * 16 + ((((r >> 2) ^ (pm >> 1)) & 3) << 2) + (r & 3);
}
__device__ __forceinline__ int stage_chunk(int pm, int q) { return pm * 16 + (((q ^ (pm >> 1)) & 3) << 2); }

constexpr int kStageTileBytes = 2 * 241 * 16 * 4;      // 30848 >= 16 rows x 240 complex (30720)

// tables for the p2 kernels: twA[240] | tw480[241] | tw1440[241]
static std::map<int, fl::cf*> g_p2_tw_cache;

static int get_p2_tables(const fl::cf** out) {
  int dev = 0;
  cudaGetDevice(&dev);
  std::lock_guard<std::mutex> lock(g_tw_mutex);
  auto it = g_p2_tw_cache.find(dev);
  if (it != g_p2_tw_cache.end()) {
    *out = it->second;
    return 0;
  }
  std::vector<fl::cf> h(kTwCount);
  for (int k1 = 0; k1 < 16; ++k1)
    for (int n2 = 0; n2 < 15; ++n2) {
      double a = -2.0 * M_PI * (double)(n2 * k1) / 240.0;
      h[k1 * 15 + n2] = fl::mk((float)cos(a), (float)sin(a));
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
  g_p2_tw_cache[dev] = d;
  *out = d;
  return 0;
}

// ================================================================================ forward
template <int R, typename T, int NBUF>
struct P2FwdSmem {
  static constexpr int W = 480 * R;
  static constexpr int kRowBytes = W * (int)sizeof(T);
  static constexpr int kRawBytes = kRows * kRowBytes;
  static constexpr int kOffU = NBUF * kRawBytes;
  static constexpr int kOffTw = kOffU + kStageTileBytes;
  static constexpr int kOffSc = kOffTw + kTwCount * 8;
  static constexpr int kOffBar = kOffSc + 244 * 4;
  static constexpr int kBytes = kOffBar + 64;
  static constexpr int kFit = (227 * 1024) / (kBytes + 1024);
  static constexpr int kCtasPerSm = kFit >= 3 ? 3 : (kFit >= 2 ? 2 : 1);
};

template <int R, typename T, int NBUF, bool ROUND, int OCC>
__global__ void __launch_bounds__(256, OCC)
dft_fwd_p2_kernel(const T* __restrict__ x, float* __restrict__ xm, const fl::cf* __restrict__ tables,
                  const float* __restrict__ mscale, int B, int C, int K, int Kp, int M) {
  using S = P2FwdSmem<R, T, NBUF>;
  constexpr int W = S::W;
  extern __shared__ __align__(128) uint8_t smp2[];
  uint8_t* raw = smp2;
  fl::cf* work = reinterpret_cast<fl::cf*>(smp2 + S::kOffU);
  float* stage = reinterpret_cast<float*>(smp2 + S::kOffU);
  fl::cf* tw = reinterpret_cast<fl::cf*>(smp2 + S::kOffTw);
  float* hs = reinterpret_cast<float*>(smp2 + S::kOffSc);      // half of the output scale of bin m (0 for m >= M)
  uint64_t* full = reinterpret_cast<uint64_t*>(smp2 + S::kOffBar);

  const int tid = threadIdx.x;
  for (int j = tid; j < kTwCount; j += 256) tw[j] = tables[j];
  for (int j = tid; j < 244; j += 256) hs[j] = j < M ? 0.5f * mscale[j] : 0.f;
  if (tid == 0) {
    for (int b = 0; b < NBUF; ++b) mbar_init(&full[b], 1);
    fence_barrier_init();
  }
  __syncthreads();

  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = tid >> 4;
  const int lane = tid & 15;
  DeviceWarp16p warp;
  // staging addresses of this thread: bin m = lane + 16 k2 sits 256 floats after bin m - 16 (the chunk swizzle only
  // depends on (m >> 1) & 3, unchanged by + 16), so every store below is base + immediate.
  // The staging tile holds TWICE the unscaled spectrum; scale (and TF32 rounding) happen on the way out.
  float* st_re = stage + stage_idx(lane, r);
  float* st_im = stage + stage_idx(241 + lane, r);

  // rows of `tile` -> raw[buf]   (executed by all 32 lanes of warp 0)
  auto issue = [&](int tile, int buf) {
    const int wl = tid & 31;
    const int b = tile / tiles_per_b;
    const int f = (tile - b * tiles_per_b) * kRows + wl;
    const int c = f / Kp, lat = f - c * Kp;
    const bool valid = wl < kRows && f < flat && lat < K;
    uint8_t* dst = raw + buf * S::kRawBytes + wl * S::kRowBytes;
    if (wl < kRows && !valid) {
      for (int i = 0; i < S::kRowBytes; i += 16) *reinterpret_cast<uint4*>(dst + i) = make_uint4(0, 0, 0, 0);
    }
    const unsigned mask = __ballot_sync(0xffffffffu, valid);
    if (wl == 0) mbar_arrive_expect_tx(&full[buf], (uint32_t)__popc(mask) * S::kRowBytes);
    __syncwarp();
    if (valid) bulk_g2s(dst, x + ((long long)(b * C + c) * K + lat) * W, S::kRowBytes, &full[buf]);
  };

  if (tid < 32) {
    for (int j = 0; j < NBUF; ++j) {
      const int t = blockIdx.x + j * gridDim.x;
      if (t < ntiles) issue(t, j);
    }
  }

  int it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const int buf = it % NBUF;
    const uint32_t parity = (uint32_t)(it / NBUF) & 1u;
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    mbar_wait(&full[buf], parity);
    const T* myrow = reinterpret_cast<const T*>(raw + buf * S::kRawBytes + r * S::kRowBytes);

    if constexpr (R == 1) {
      fl::cf Z[15];
      fp::fft240p_fwd(warp, [&](int n) { return load_pair<1, T>(myrow, n, 0); }, work + r * 240, tw, Z);
      __syncthreads();                          // raw[buf] consumed by every row; work no longer read
      if (tid < 32) {
        const int nt = tile + NBUF * gridDim.x;
        if (nt < ntiles) issue(nt, buf);
      }
      const float nyq = fp::rfft480_split(warp, Z, [&](int k2, fl::cf v) {
        st_re[k2 * 256] = v.x;
        st_im[k2 * 256] = v.y;
      }, tw + 240);
      if (lane == 0) {
        stage[stage_idx(240, r)] = nyq;
        stage[stage_idx(481, r)] = 0.f;
      }
    } else {
      fl::cf X[15], x240;
      fp::rfft_row_fwd<R>(warp, [&](int n, int rr) { return load_pair<R, T>(myrow, n, rr); }, work + r * 240, X, x240,
                          tw, tw + 240, tw + 481);
      __syncthreads();
      if (tid < 32) {
        const int nt = tile + NBUF * gridDim.x;
        if (nt < ntiles) issue(nt, buf);
      }
#pragma unroll
      for (int k2 = 0; k2 < 15; ++k2) {
        st_re[k2 * 256] = X[k2].x;
        st_im[k2 * 256] = X[k2].y;
      }
      if (lane == 0) {
        stage[stage_idx(240, r)] = x240.x;
        stage[stage_idx(481, r)] = x240.y;
      }
    }
    __syncthreads();
    // ---- staging tile -> XM: 16-byte vectors, 4 lanes cover the 16 latitudes of one (m, re/im)
    {
      const int q = tid & 3;
      const bool fok = f0 + 4 * q < flat;
      const long long gstep = 64LL * RS * Kp;
#pragma unroll
      for (int reim = 0; reim < 2; ++reim) {
        float* gp = xm + ((long long)(tid >> 2) * RS + (long long)(b * 2 + reim) * Cp) * Kp + f0 + 4 * q;
        for (int m = tid >> 2; m < M; m += 64, gp += gstep) {
          float4 v = *reinterpret_cast<const float4*>(stage + stage_chunk(reim * 241 + m, q));
          const float sc = hs[m];
          v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
          if (ROUND) { v.x = tf32_rn(v.x); v.y = tf32_rn(v.y); v.z = tf32_rn(v.z); v.w = tf32_rn(v.w); }
          if (fok) *reinterpret_cast<float4*>(gp) = v;
        }
      }
    }
    __syncthreads();
  }
}

// ================================================================================ inverse
// per-bin constants of the inverse pre-processing, k in [0, 240): scales of bins k and 240-k, and e^{-2 pi i k / 480}
struct InvC {
  float s0, s1, wx, wy;
};

template <int R, typename T, int NST>
struct P2InvSmem {
  static constexpr int W = 480 * R;
  static constexpr int kRowBytes = W * (int)sizeof(T);
  static constexpr int kRowsBytes = kRows * kRowBytes;
  static constexpr int kOffRows = NST * kStageTileBytes;
  static constexpr int kOffTw = kOffRows + kRowsBytes;
  static constexpr int kOffC = kOffTw + kTwCount * 8;
  static constexpr int kBytes = kOffC + 240 * 16;
  static constexpr int kFit = (227 * 1024) / (kBytes + 1024);
  static constexpr int kCtasPerSm = kFit >= 2 ? 2 : 1;
};

template <int R, typename T, int NST>
__global__ void __launch_bounds__(256, P2InvSmem<R, T, NST>::kCtasPerSm)
dft_inv_p2_kernel(const float* __restrict__ xm, T* __restrict__ y, const fl::cf* __restrict__ tables,
                  const float* __restrict__ mscale, int B, int C, int K, int Kp, int M, int m_valid) {
  using S = P2InvSmem<R, T, NST>;
  constexpr int W = S::W;
  extern __shared__ __align__(128) uint8_t smp2[];
  uint8_t* rows = smp2 + S::kOffRows;
  fl::cf* tw = reinterpret_cast<fl::cf*>(smp2 + S::kOffTw);
  InvC* ic = reinterpret_cast<InvC*>(smp2 + S::kOffC);

  const int tid = threadIdx.x;
  for (int j = tid; j < kTwCount; j += 256) tw[j] = tables[j];
  for (int j = tid; j < 240; j += 256) {
    const fl::cf w480 = tables[240 + j];
    InvC c;
    c.s0 = j < m_valid ? mscale[j] : 0.f;
    c.s1 = (240 - j) < m_valid ? mscale[240 - j] : 0.f;
    c.wx = w480.x;
    c.wy = w480.y;
    ic[j] = c;
  }

  const int Cp = (C + 3) & ~3;
  const long long RS = 2LL * B * Cp;
  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = tid >> 4;
  const int lane = tid & 15;
  DeviceWarp16p warp;
  // float offsets inside a staging tile of this thread's bins: k = lane + 16 k2 -> + 256 k2 ; 240 - k -> - 256 k2
  const int o_re = stage_idx(lane, r), o_im = stage_idx(241 + lane, r);
  const int o_pre = stage_idx(240 - lane, r), o_pim = stage_idx(481 - lane, r);

  // XM slices of `tile` -> staging tile `buf` (all threads; 16-byte async copies)
  auto prefetch = [&](int tile, int buf) {
    float* st = reinterpret_cast<float*>(smp2 + buf * kStageTileBytes);
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    const int q = tid & 3;
    const bool fok = f0 + 4 * q < flat;
    const long long gstep = 64LL * RS * Kp;
#pragma unroll
    for (int reim = 0; reim < 2; ++reim) {
      const float* gp = xm + ((long long)(tid >> 2) * RS + (long long)(b * 2 + reim) * Cp) * Kp + f0 + 4 * q;
      for (int m = tid >> 2; m < 241; m += 64, gp += gstep) {
        float* dst = st + stage_chunk(reim * 241 + m, q);
        if (m < m_valid && fok) cp_async16(dst, gp);
        else *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
      }
    }
    cp_async_commit();
  };

  if (blockIdx.x < ntiles) prefetch(blockIdx.x, 0);
  __syncthreads();                               // tables visible

  int it = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
    const int cur = (NST == 2) ? (it & 1) : 0;
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    const int nt = tile + gridDim.x;
    cp_async_wait0();
    if (tid < 32) bulk_wait_read0();             // row buffer of the previous tile has been read by the bulk stores
    __syncthreads();
    if (NST == 2 && nt < ntiles) prefetch(nt, cur ^ 1);
    const float* st = reinterpret_cast<const float*>(smp2 + cur * kStageTileBytes);
    fl::cf* work = reinterpret_cast<fl::cf*>(smp2 + cur * kStageTileBytes);
    T* myrow = reinterpret_cast<T*>(rows + r * S::kRowBytes);
    auto storez = [&](int n, int rr, fl::cf z) { store_pair<R, T>(myrow, n, rr, z); };

    if constexpr (R == 1) {
      fl::cf V[15];
#pragma unroll
      for (int k2 = 0; k2 < 15; ++k2) {
        const InvC c = ic[lane + 16 * k2];
        fl::cf h = fl::mk(st[o_re + 256 * k2] * c.s0, st[o_im + 256 * k2] * c.s0);
        fl::cf hp = fl::mk(st[o_pre - 256 * k2] * c.s1, st[o_pim - 256 * k2] * c.s1);
        if (k2 == 0) fp::irfft_twist<1>(0, lane, h, hp, tw + 481);
        V[k2] = fp::irfft480_prep(h, hp, fl::mk(c.wx, c.wy));
      }
      __syncthreads();                           // staging tile consumed: it now serves as the FFT work buffer
      fp::irfft_row<1>(warp, [&](int, int k2) { return V[k2]; }, work + r * 240, storez, tw);
    } else {
      fl::cf H[15], HP[15];
#pragma unroll
      for (int k2 = 0; k2 < 15; ++k2) {
        const InvC c = ic[lane + 16 * k2];
        H[k2] = fl::mk(st[o_re + 256 * k2] * c.s0, st[o_im + 256 * k2] * c.s0);
        HP[k2] = fl::mk(st[o_pre - 256 * k2] * c.s1, st[o_pim - 256 * k2] * c.s1);
      }
      __syncthreads();
      fp::irfft_row<R>(warp, [&](int rr, int k2) {
        const int k = lane + 16 * k2;
        fl::cf h = H[k2], hp = HP[k2];
        fp::irfft_twist<R>(rr, k, h, hp, tw + 481);
        return fp::irfft480_prep(h, hp, tw[240 + k]);
      }, work + r * 240, storez, tw);
    }
    fence_proxy_async();                         // generic-proxy row writes -> visible to the bulk (async proxy) stores
    __syncthreads();
    if (NST == 1 && nt < ntiles) prefetch(nt, 0);
    if (tid < 32) {
      const int f = f0 + tid;
      const int c = f / Kp, lat = f - c * Kp;
      if (tid < kRows && f < flat && lat < K)
        bulk_s2g(y + ((long long)(b * C + c) * K + lat) * W, rows + tid * S::kRowBytes, S::kRowBytes);
      bulk_commit();
    }
  }
  cp_async_wait0();
  if (tid < 32) bulk_wait_read0();
}

// ================================================================================ host launchers
template <int R, typename T>
static int launch_fwd_p2(const void* x, int B, int C, int K, int Kp, int M, const float* mscale, float* xm,
                         int round_out, cudaStream_t stream) {
  // 16-bit rows at R = 3: a single row buffer lets two CTAs share an SM (16 warps instead of 8)
  constexpr int NBUF = (R == 3 && sizeof(T) == 2) ? 1 : 2;
  using S = P2FwdSmem<R, T, NBUF>;
  static_assert(S::kBytes <= 227 * 1024, "forward DFT tile does not fit shared memory");
  const fl::cf* tables;
  if (get_p2_tables(&tables)) return 1;
  constexpr int OCC = S::kCtasPerSm;
  constexpr int OCC2 = OCC > 1 ? OCC - 1 : 1;        // experiment knob: SFNO_P2_OCC=low trades occupancy for registers
  static bool attr = false;
  if (!attr) {
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_p2_kernel<R, T, NBUF, true, OCC>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_p2_kernel<R, T, NBUF, false, OCC>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_p2_kernel<R, T, NBUF, true, OCC2>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr = true;
  }
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  const char* env = getenv("SFNO_P2_OCC");
  const bool low = env && env[0] == 'l';
  const int grid = std::min(ntiles, (low ? OCC2 : OCC) * sm_count());
  if (low && round_out)
    dft_fwd_p2_kernel<R, T, NBUF, true, OCC2><<<grid, 256, S::kBytes, stream>>>(static_cast<const T*>(x), xm, tables, mscale, B, C, K, Kp, M);
  else if (round_out)
    dft_fwd_p2_kernel<R, T, NBUF, true, OCC><<<grid, 256, S::kBytes, stream>>>(static_cast<const T*>(x), xm, tables, mscale, B, C, K, Kp, M);
  else
    dft_fwd_p2_kernel<R, T, NBUF, false, OCC><<<grid, 256, S::kBytes, stream>>>(static_cast<const T*>(x), xm, tables, mscale, B, C, K, Kp, M);
  SFNO_LAUNCHED();
  return 0;
}

template <int R, typename T>
static int launch_inv_p2(const float* xm, int B, int C, int K, int Kp, int M, int m_valid, const float* mscale,
                         void* y, cudaStream_t stream) {
  constexpr int NST = (R == 3 && sizeof(T) == 2) ? 1 : 2;
  using S = P2InvSmem<R, T, NST>;
  static_assert(S::kBytes <= 227 * 1024, "inverse DFT tile does not fit shared memory");
  const fl::cf* tables;
  if (get_p2_tables(&tables)) return 1;
  static bool attr = false;
  if (!attr) {
    SFNO_CUDA(cudaFuncSetAttribute(dft_inv_p2_kernel<R, T, NST>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr = true;
  }
  const int ntiles = ((C * Kp + kRows - 1) / kRows) * B;
  const int grid = std::min(ntiles, S::kCtasPerSm * sm_count());
  dft_inv_p2_kernel<R, T, NST><<<grid, 256, S::kBytes, stream>>>(xm, static_cast<T*>(y), tables, mscale, B, C, K, Kp, M, m_valid);
  SFNO_LAUNCHED();
  return 0;
}

}  // namespace sfno
