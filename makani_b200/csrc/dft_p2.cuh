// Longitude DFT kernels for nlon = 480*R built on the two-pass FFT of fft_pass.cuh  (included by dft.cu).
//
//   tile      16 consecutive flat positions f = c*Kp + lat of one batch entry; 256 threads = 16 rows x 16 lanes
//   rows      move with 1-D bulk async copies (cp.async.bulk, one per row); the rows of tile i+1 are requested as
//             soon as every warp has read tile i's for the last time
//   XM side   goes through a swizzled [2*241][16] staging tile so that both the per-lane accesses of the FFT
//             (lane = bin) and the 16-byte global accesses (4 consecutive latitudes) are (nearly) conflict free
//   work      private FFT exchange row per 16-lane group; the tile loop has no CTA-wide barrier (mbarrier hand-offs)
#pragma once

namespace sfno {

struct DeviceWarp16p {
  __device__ __forceinline__ int lane16() const { return threadIdx.x & 15; }
  __device__ __forceinline__ float shfl_idx(float v, int src) const {
    return __shfl_sync(0xffffffffu, v, (threadIdx.x & 16) | src);
  }
  __device__ __forceinline__ void sync() const { __syncwarp(); }
};

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes, uint64_t policy) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
               ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait0() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// Staging tile between the FFT lanes and XM: two planes (re, im) of [248][16] floats; plane p, bin m, row r at
//     p*3968 + m*16 + (((r >> 2) ^ (m >> 1)) & 3)*4 + (r & 3)
// This is exactly the TMA SWIZZLE_64B layout of a box {16 latitudes, 1, 241 bins} (16-byte chunk index XOR address
// bits [7,9)), so one tensor copy per plane moves the whole tile to / from XM[m][r][k]; on the FFT side 16 lanes with
// consecutive m and two adjacent rows hit 16 different banks (2-way conflict; a dense tile would be 16-way).
constexpr int kPlaneFloats = 248 * 16;                     // 15872 B: planes start 512-byte aligned (swizzle period)
constexpr int kStageTileBytes = 2 * kPlaneFloats * 4;      // 31744 >= 16 rows x 240 complex (30720)
__device__ __forceinline__ int stage_idx(int m, int r) { return m * 16 + ((((r >> 2) ^ (m >> 1)) & 3) << 2) + (r & 3); }

__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2,
                                             uint64_t policy) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3, %4}], [%1], %5;"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
               : "memory");
}

__device__ __forceinline__ void tma_store_4d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
               : "memory");
}

// h-sharded transforms (distributed/p2p.py): the XM side of the DFT kernels is not the local XM tensor but the
// channel-sharded XM_ch buffers of ALL ranks, in peer memory - channel c belongs to rank p with c0[p] <= c < c0[p+1],
// whose buffer is [m][2B][Cp_p][Kg]; map[p] views the columns of THIS rank's latitude slab in it as
// (lat, c - c0[p], j, m).  The forward kernel stores its staging tiles there (the latitude -> channel exchange rides
// on the tensor stores), the inverse kernel loads its tiles from there (channel -> latitude).
constexpr int kMaxPeers = 8;
struct PeerMaps {
  CUtensorMap map[kMaxPeers];
  int c0[kMaxPeers + 1];
  int n;
};
struct LocalMap {
  CUtensorMap map;
};

// tile (16 latitudes lat0.. of channel c / flat position f0 = c*pitch + lat0, batch b, plane reim) <-> XM
template <class MAPS>
struct XmTile;
template <>
struct XmTile<LocalMap> {
  static constexpr bool kDist = false;
  __device__ static void prefetch(const LocalMap& m) { tma_prefetch_desc(&m.map); }
  __device__ static void store(const LocalMap& m, const void* src, int f0, int, int j, uint64_t pol) {
    tma_store_3d(&m.map, src, f0, j, 0, pol);
  }
  __device__ static void load(const LocalMap& m, void* dst, uint64_t* bar, int f0, int, int j, uint64_t pol) {
    tma_load_3d_hint(dst, &m.map, bar, f0, j, 0, pol);
  }
};
template <>
struct XmTile<PeerMaps> {
  static constexpr bool kDist = true;
  __device__ static void prefetch(const PeerMaps& m) {
    for (int p = 0; p < m.n; ++p) tma_prefetch_desc(&m.map[p]);
  }
  __device__ static int owner(const PeerMaps& m, int c) {
    int p = 0;
    while (p + 1 < m.n && c >= m.c0[p + 1]) ++p;
    return p;
  }
  __device__ static void store(const PeerMaps& m, const void* src, int f0, int pitch, int j, uint64_t) {
    const int c = f0 / pitch, p = owner(m, c);
    tma_store_4d(&m.map[p], src, f0 - c * pitch, c - m.c0[p], j, 0);
  }
  __device__ static void load(const PeerMaps& m, void* dst, uint64_t* bar, int f0, int pitch, int j, uint64_t) {
    const int c = f0 / pitch, p = owner(m, c);
    tma_load_4d(dst, &m.map[p], bar, f0 - c * pitch, c - m.c0[p], j, 0);
  }
};

// XM viewed as [m][j = b*2 + reim][f = c*Kp + lat]; box = 16 flat positions x 241 bins of one (b, re/im) plane.
// mdim bounds the bins (loads beyond it read zeros, stores beyond it are dropped).
static int xm_stage_map(CUtensorMap* map, const float* xm, int B, int C, int Kp, int mdim) {
  const int Cp = (C + 3) & ~3;
  uint64_t dims[3] = {(uint64_t)C * Kp, (uint64_t)2 * B, (uint64_t)mdim};
  uint64_t str[2] = {(uint64_t)Cp * Kp * 4, (uint64_t)2 * B * Cp * Kp * 4};
  uint32_t box[3] = {16, 1, 241};
  return encode_tensor_map(map, xm, 3, dims, str, box, 3);
}

// tables for the p2 kernels: twA[240] | tw480[241] | tw1440[241]
// band (nlon = 1440 with more than 241 modes, see dft_fwd_band in dft.cu): the kernels combine the three length-480
// sub-transforms with tw1440[k]^r; with the table multiplied by a cube root of unity the very same kernels produce
// (consume) bins 480 + k (band 2, factor e^{-2 pi i / 3}) and the conjugates of bins 480 - k (band 1, e^{+2 pi i / 3}).
static std::map<int, fl::cf*> g_p2_tw_cache;

static int get_p2_tables(const fl::cf** out, int band = 0) {
  int dev = 0;
  cudaGetDevice(&dev);
  dev = dev * 4 + band;
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
    double c = -2.0 * M_PI * ((double)k / 1440.0 + (band == 2 ? 1.0 / 3.0 : (band == 1 ? -1.0 / 3.0 : 0.0)));
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
  static constexpr int kOffRaw = kStageTileBytes;           // staging tile first (1024-byte aligned base)
  static constexpr int kOffTw = kOffRaw + NBUF * kRawBytes;
  static constexpr int kOffSc = kOffTw + kTwCount * 8;
  static constexpr int kOffBar = kOffSc + 244 * 4;
  static constexpr int kBytes = kOffBar + 64 + 1024;        // + slack for the manual 1024-byte alignment
  static constexpr int kFit = (227 * 1024) / (kBytes + 1024);
  static constexpr int kCtasPerSm = kFit >= 3 ? 3 : (kFit >= 2 ? 2 : 1);
};

template <int R, typename T, int NBUF, bool ROUND, int OCC, class MAPS>
__global__ void __launch_bounds__(256, OCC)
dft_fwd_p2_kernel(const T* __restrict__ x, const __grid_constant__ MAPS xm_map,
                  const fl::cf* __restrict__ tables, const float* __restrict__ mscale, int B, int C, int K, int Kp,
                  int M, int* __restrict__ sched, long long xbs, unsigned long long pol_x, unsigned long long pol_xm) {
  using S = P2FwdSmem<R, T, NBUF>;
  constexpr int W = S::W;
  // 1024-byte aligned base, computed as an OFFSET from the shared array so that every access below stays in the
  // shared address space (an integer round-trip of the pointer turns them into slower generic LD / ST)
  extern __shared__ __align__(1024) uint8_t smp2_raw[];
  uint8_t* smp2 = smp2_raw + ((1024u - (smem_u32(smp2_raw) & 1023u)) & 1023u);
  fl::cf* work = reinterpret_cast<fl::cf*>(smp2);
  float* stage = reinterpret_cast<float*>(smp2);
  uint8_t* raw = smp2 + S::kOffRaw;
  fl::cf* tw = reinterpret_cast<fl::cf*>(smp2 + S::kOffTw);
  float* hs = reinterpret_cast<float*>(smp2 + S::kOffSc);      // half of the output scale of bin m (0 for m >= M)
  uint64_t* full = reinterpret_cast<uint64_t*>(smp2 + S::kOffBar);
  volatile int* tq = reinterpret_cast<volatile int*>(full + 2);   // ring of claimed tile indices (-1: no more work)

  const int tid = threadIdx.x;
  for (int j = tid; j < kTwCount; j += 256) tw[j] = tables[j];
  for (int j = tid; j < 244; j += 256) hs[j] = j < M ? 0.5f * mscale[j] : 0.f;
  if (tid == 0) {
    XmTile<MAPS>::prefetch(xm_map);
    for (int b = 0; b < NBUF; ++b) mbar_init(&full[b], 1);
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();                       // set-up above overlapped the previous kernel's tail (launch_pdl)
  pdl_wait();

  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = tid >> 4;
  const int lane = tid & 15;
  DeviceWarp16p warp;
  // staging addresses of this thread: bin m = lane + 16 k2 sits 256 floats after bin m - 16 (the chunk swizzle only
  // depends on (m >> 1) & 3, unchanged by + 16), so every store below is base + immediate
  float* st_re = stage + stage_idx(lane, r);
  float* st_im = st_re + kPlaneFloats;
  auto fin = [](float v, float sc) { return ROUND ? tf32_rn(v * sc) : v * sc; };

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
    if (valid) bulk_g2s(dst, x + b * xbs + ((long long)c * K + lat) * W, S::kRowBytes, &full[buf], pol_x);
  };

  // tiles are claimed from a global counter (a CTA that starts late - SMs shared with NCCL kernels - just takes fewer);
  // warp 0 claims tile it+NBUF while tile it is transformed and starts its row copies
  auto claim = [&](int slot, int buf) {          // all 32 lanes of warp 0
    int t = 0;
    if ((tid & 31) == 0) t = atomicAdd(sched, 1);
    t = __shfl_sync(0xffffffffu, t, 0);
    if (t >= ntiles) t = -1;
    if ((tid & 31) == 0) tq[slot] = t;
    if (t >= 0) issue(t, buf);
  };
  if (tid < 32) {
    for (int j = 0; j < NBUF; ++j) claim(j, j);
  }
  __syncthreads();

  for (int it = 0;; ++it) {
    const int tile = tq[it & 3];
    if (tile < 0) break;
    const int buf = it % NBUF;
    const uint32_t parity = (uint32_t)(it / NBUF) & 1u;
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    mbar_wait(&full[buf], parity);
    if (it > 0) {                                // the staging tile doubles as the FFT work buffer: the tensor stores
      if (tid == 0) bulk_wait_read0();           // of the previous tile must have read it (overlaps the row wait above)
      __syncthreads();
    }
    const T* myrow = reinterpret_cast<const T*>(raw + buf * S::kRawBytes + r * S::kRowBytes);

    if constexpr (R == 1) {
      fl::cf Z[15];
      fp::fft240p_fwd<fp::tw_regs<1>::value>(warp, [&](int n) { return load_pair<1, T>(myrow, n, 0); },
                                             work + r * 240, tw, Z);
      __syncthreads();                          // raw[buf] consumed by every row; work no longer read
      if (tid < 32) claim((it + NBUF) & 3, buf);
      const float nyq = fp::rfft480_split<fp::tw_regs<1>::value>(warp, Z, [&](int k2, fl::cf v) {
        const float sc = hs[lane + 16 * k2];
        st_re[k2 * 256] = fin(v.x, sc);
        st_im[k2 * 256] = fin(v.y, sc);
      }, tw + 240);
      if (lane == 0) {
        stage[3840 + r] = fin(nyq, hs[240]);
        stage[kPlaneFloats + 3840 + r] = 0.f;
      }
    } else {
      fl::cf X[15], x240;
      fp::rfft_row_fwd<R>(warp, [&](int n, int rr) { return load_pair<R, T>(myrow, n, rr); }, work + r * 240, X, x240,
                          tw, tw + 240, tw + 481);
      __syncthreads();
      if (tid < 32) claim((it + NBUF) & 3, buf);
#pragma unroll
      for (int k2 = 0; k2 < 15; ++k2) {
        const float sc = hs[lane + 16 * k2];
        st_re[k2 * 256] = fin(X[k2].x, sc);
        st_im[k2 * 256] = fin(X[k2].y, sc);
      }
      if (lane == 0) {
        stage[3840 + r] = fin(x240.x, hs[240]);
        stage[kPlaneFloats + 3840 + r] = fin(x240.y, hs[240]);
      }
    }
    fence_proxy_async();                         // staging tile (generic-proxy writes) -> visible to the tensor stores
    __syncthreads();
    // ---- staging tile -> XM: one tensor store per plane (16 latitudes x 241 bins; bins >= M and positions beyond
    //      C*Kp are clipped by the tensor map)
    if (tid == 0) {
      XmTile<MAPS>::store(xm_map, stage, f0, Kp, 2 * b, pol_xm);
      XmTile<MAPS>::store(xm_map, stage + kPlaneFloats, f0, Kp, 2 * b + 1, pol_xm);
      bulk_commit();
    }
  }
  if (tid == 0) {
    bulk_wait_read0();                           // shared memory must outlive the last tensor stores' reads
    __threadfence();
    if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {   // last CTA re-arms the scheduler for the next launch
      sched[0] = 0;
      sched[1] = 0;
    }
  }
}

// ================================================================================ inverse
// shared memory: NST staging tiles | 16 output rows | 16 FFT exchange rows | twiddles | per-bin scale | mbarriers + ring.
// Every region except the staging tiles is private to one 16-lane group, so the tile loop needs no CTA-wide barrier.
template <int R, typename T, int NST>
struct P2InvSmem {
  static constexpr int W = 480 * R;
  static constexpr int kRowBytes = W * (int)sizeof(T);
  static constexpr int kOffRows = NST * kStageTileBytes;     // staging tiles first (swizzle: 512-byte aligned base)
  static constexpr int kOffWork = kOffRows + kRows * kRowBytes;
  static constexpr int kTwUsed = (R == 1) ? 481 : kTwCount;  // tw1440 only for R = 3
  static constexpr int kOffTw = kOffWork + kRows * 240 * 8;
  static constexpr int kOffSc = kOffTw + kTwUsed * 8;
  static constexpr int kOffBar = kOffSc + 244 * 4;
  static constexpr int kBytes = kOffBar + 128;               // 2 x NST mbarriers (<= 6) + ring of 8 tile indices
  static constexpr bool kTwoCtas = 2 * (kBytes + 1024) <= 228 * 1024;
  static constexpr int kCtasPerSm = kTwoCtas ? 2 : 1;
};
// Staging depth: the kernel is bound by the bytes it keeps in flight (a tile is ~480 scattered 64-byte pieces of XM,
// latency ~2 us under load), so take as many staging tiles (<= 3) as still let two CTAs share an SM.
template <int R, typename T>
struct P2InvDepth {
  static constexpr int value =
      P2InvSmem<R, T, 3>::kTwoCtas ? 3 : (P2InvSmem<R, T, 2>::kTwoCtas ? 2 : 1);
};

// hand-offs of one tile:
//   full      thread 0 -> all     both planes of the staging tile have landed (tensor loads) and tq[it & 3] is valid
//   consumed  8 warps -> thread 0  every warp holds its bins in registers -> thread 0 claims tile it+1 and starts its
//                                 tensor loads, which then overlap the whole transform of tile it
// The output rows are private to their 16-lane group: lane 0 of the group stores its row with one bulk copy and waits
// for that copy's shared-memory read just before the group writes the row again (one tile later).
template <int R, typename T, int NST, class MAPS>
__global__ void __launch_bounds__(256, P2InvSmem<R, T, NST>::kCtasPerSm)
dft_inv_p2_kernel(const __grid_constant__ MAPS xm_map, T* __restrict__ y, const fl::cf* __restrict__ tables,
                  const float* __restrict__ mscale, int B, int C, int K, int Kp, int m_valid, int* __restrict__ sched,
                  long long xbs, unsigned long long pol_y, unsigned long long pol_xm) {
  using S = P2InvSmem<R, T, NST>;
  constexpr int W = S::W;
  extern __shared__ __align__(1024) uint8_t smp2[];
  fl::cf* tw = reinterpret_cast<fl::cf*>(smp2 + S::kOffTw);
  float* msc = reinterpret_cast<float*>(smp2 + S::kOffSc);
  uint64_t* full = reinterpret_cast<uint64_t*>(smp2 + S::kOffBar);   // [NST]
  uint64_t* consumed = full + 3;                                     // [NST]
  volatile int* tq = reinterpret_cast<volatile int*>(full + 10);     // ring of 8 claimed tile indices (-1: no more work)

  const int tid = threadIdx.x;
  for (int j = tid; j < S::kTwUsed; j += 256) tw[j] = tables[j];
  for (int j = tid; j < 244; j += 256) msc[j] = j < m_valid ? mscale[j] : 0.f;
  if (tid == 0) {
    if (smem_u32(smp2) & 511u) __trap();
    XmTile<MAPS>::prefetch(xm_map);
    for (int j = 0; j < NST; ++j) {
      mbar_init(full + j, 1);
      mbar_init(consumed + j, 8);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  pdl_wait();

  const int flat = C * Kp;
  const int tiles_per_b = (flat + kRows - 1) / kRows;
  const int ntiles = tiles_per_b * B;
  const int r = tid >> 4;
  const int lane = tid & 15;
  const int psrc = (16 - lane) & 15;
  DeviceWarp16p warp;
  const int o_re = stage_idx(lane, r);           // bin k = lane + 16 k2 -> + 256 floats per k2
  fl::cf* work = reinterpret_cast<fl::cf*>(smp2 + S::kOffWork) + r * 240;
  T* myrow = reinterpret_cast<T*>(smp2 + S::kOffRows + r * S::kRowBytes);
  auto storez = [&](int n, int rr, fl::cf z) { store_pair<R, T>(myrow, n, rr, z); };

  // thread 0: claim the next tile and start one tensor load per plane.  The tensor map is bounded by m_valid bins and
  // C*Kp positions; everything outside is zero filled, and the full box always counts for the mbarrier.
  auto claim = [&](int slot, int buf) {
    int t = atomicAdd(sched, 1);
    if (t >= ntiles) t = -1;
    tq[slot] = t;
    if (t < 0) {
      mbar_arrive(full + buf);                   // releases the waiters, who then read tq = -1
      return;
    }
    float* dst = reinterpret_cast<float*>(smp2 + buf * kStageTileBytes);
    const int b = t / tiles_per_b;
    const int f0 = (t - b * tiles_per_b) * kRows;
    mbar_arrive_expect_tx(full + buf, 2u * 241u * 64u);
    XmTile<MAPS>::load(xm_map, dst, full + buf, f0, Kp, 2 * b, pol_xm);
    XmTile<MAPS>::load(xm_map, dst + kPlaneFloats, full + buf, f0, Kp, 2 * b + 1, pol_xm);
  };
  if (tid == 0) {
    for (int j = 0; j < NST; ++j) claim(j, j);
  }

  for (int it = 0;; ++it) {
    const int buf = it % NST;
    const uint32_t bph = (uint32_t)(it / NST) & 1u;
    mbar_wait(full + buf, bph);
    const int tile = tq[it & 7];
    if (tile < 0) break;
    const int b = tile / tiles_per_b;
    const int f0 = (tile - b * tiles_per_b) * kRows;
    const float* st = reinterpret_cast<const float*>(smp2 + buf * kStageTileBytes);

    // this lane's bins k = lane + 16 k2 (scaled); the partner bins 240 - k come from lane 16 - k1 by shuffle
    // (register 14 - k2), lane 0 pairs its own registers (k2 <-> 15 - k2) and bin 0 with the extra bin 240
    fl::cf H[15], HP[15];
#pragma unroll
    for (int k2 = 0; k2 < 15; ++k2) {
      const float s0 = msc[lane + 16 * k2];
      H[k2] = fl::mk(st[o_re + 256 * k2] * s0, st[kPlaneFloats + o_re + 256 * k2] * s0);
    }
    const float s240 = msc[240];
    const fl::cf h240 = fl::mk(st[3840 + r] * s240, st[kPlaneFloats + 3840 + r] * s240);
    __syncwarp();
    if ((tid & 31) == 0) mbar_arrive(consumed + buf);
    if (tid == 0) {
      mbar_wait(consumed + buf, bph);
      claim((it + NST) & 7, buf);
    }
#pragma unroll
    for (int k2 = 0; k2 < 15; ++k2) {
      const fl::cf hs = fl::mk(warp.shfl_idx(H[14 - k2].x, psrc), warp.shfl_idx(H[14 - k2].y, psrc));
      const fl::cf h0 = (k2 == 0) ? h240 : H[15 - k2];
      HP[k2] = (lane == 0) ? h0 : hs;
    }
    if (lane == 0) bulk_wait_read0();            // the previous bulk store of this row has read it
    __syncwarp();
    fp::irfft_row<R>(warp, H, HP, work, storez, tw, tw + 240, tw + 481);
    fence_proxy_async();                         // generic-proxy row writes -> visible to the bulk (async proxy) store
    __syncwarp();
    if (lane == 0) {
      const int f = f0 + r;
      const int c = f / Kp, lat = f - c * Kp;
      if (f < flat && lat < K) bulk_s2g(y + b * xbs + ((long long)c * K + lat) * W, myrow, S::kRowBytes, pol_y);
      bulk_commit();
    }
  }
  if (lane == 0) bulk_wait_read0();
  if (tid == 0) {
    __threadfence();
    if (atomicAdd(sched + 1, 1) == (int)gridDim.x - 1) {
      sched[0] = 0;
      sched[1] = 0;
    }
  }
}

// ================================================================================ host launchers
template <class Kern>
static int p2_prepare(Kern kern, int smem_bytes) {
  SFNO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  // two CTAs per SM only fit with the largest shared-memory carve-out
  SFNO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  return 0;
}

// maps: LocalMap (XM tensor of this GPU, pitch = Kp) or PeerMaps (channel-sharded XM_ch buffers of all ranks,
// pitch = round_up(K, 16): tiles never straddle two channels there)
template <int R, typename T, class MAPS>
static int launch_fwd_p2_maps(const void* x, int B, int C, int Ctot, int K, int pitch, int M, const float* mscale,
                              const MAPS& maps, int round_out, cudaStream_t stream, int band = 0) {
  // 16-bit rows at R = 3: a single row buffer lets two CTAs share an SM (16 warps instead of 8)
  // (re-measured after the packed arithmetic, -DSFNO_R3_NBUF=2: 679 us against 464 us - the kernel lives on warps per SM)
#ifndef SFNO_R3_NBUF
#define SFNO_R3_NBUF 1
#endif
  constexpr int NBUF = (R == 3 && sizeof(T) == 2) ? SFNO_R3_NBUF : 2;
  using S = P2FwdSmem<R, T, NBUF>;
  static_assert(S::kBytes <= 227 * 1024, "forward DFT tile does not fit shared memory");
  constexpr int OCC = S::kCtasPerSm;
  const fl::cf* tables;
  if (get_p2_tables(&tables, band)) return 1;
  static PerDeviceOnce attr;
  if (attr.needed()) {
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_p2_kernel<R, T, NBUF, true, OCC, MAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    SFNO_CUDA(cudaFuncSetAttribute(dft_fwd_p2_kernel<R, T, NBUF, false, OCC, MAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::kBytes));
    attr.mark();
  }
  const int ntiles = ((C * pitch + kRows - 1) / kRows) * B;
  const int grid = std::min(ntiles, OCC * sm_count());
  int* sched = sched_counters();
  SFNO_CHECK(sched != nullptr, "could not allocate the tile scheduler counters");
  // channel window (Ctot > C): the field rows stream through L2, the XM window is meant to stay there
  const bool win = Ctot != C && window_hints();
  const unsigned long long pol_x = win ? kL2EvictFirst : kL2EvictNormal, pol_xm = win ? kL2EvictLast : kL2EvictNormal;
  const long long xbs = (long long)Ctot * K * S::W;
  if (round_out)
    SFNO_CUDA(launch_pdl(dft_fwd_p2_kernel<R, T, NBUF, true, OCC, MAPS>, dim3(grid), dim3(256), S::kBytes, stream,
                         static_cast<const T*>(x), maps, tables, mscale, B, C, K, pitch, M, sched, xbs, pol_x, pol_xm));
  else
    SFNO_CUDA(launch_pdl(dft_fwd_p2_kernel<R, T, NBUF, false, OCC, MAPS>, dim3(grid), dim3(256), S::kBytes, stream,
                         static_cast<const T*>(x), maps, tables, mscale, B, C, K, pitch, M, sched, xbs, pol_x, pol_xm));
  SFNO_LAUNCHED();
  return 0;
}

template <int R, typename T>
static int launch_fwd_p2(const void* x, int B, int C, int Ctot, int K, int Kp, int M, const float* mscale, float* xm,
                         int round_out, cudaStream_t stream, int band = 0) {
  LocalMap lm;
  if (xm_stage_map(&lm.map, xm, B, C, Kp, M)) return 1;
  return launch_fwd_p2_maps<R, T, LocalMap>(x, B, C, Ctot, K, Kp, M, mscale, lm, round_out, stream, band);
}

template <int R, typename T, class MAPS>
static int launch_inv_p2_maps(const MAPS& maps, int B, int C, int Ctot, int K, int pitch, int m_valid, const float* mscale,
                              void* y, cudaStream_t stream, int band = 0) {
  constexpr int NST = P2InvDepth<R, T>::value;
  using S = P2InvSmem<R, T, NST>;
  static_assert(S::kBytes <= 227 * 1024, "inverse DFT tile does not fit shared memory");
  const fl::cf* tables;
  if (get_p2_tables(&tables, band)) return 1;
  static PerDeviceOnce attr;
  if (attr.needed()) {
    if (p2_prepare(dft_inv_p2_kernel<R, T, NST, MAPS>, S::kBytes)) return 1;
    attr.mark();
  }
  const int ntiles = ((C * pitch + kRows - 1) / kRows) * B;
  const int grid = std::min(ntiles, S::kCtasPerSm * sm_count());
  int* sched = sched_counters();
  SFNO_CHECK(sched != nullptr, "could not allocate the tile scheduler counters");
  const bool win = Ctot != C && window_hints();
  const unsigned long long pol_y = win ? kL2EvictFirst : kL2EvictNormal, pol_xm = win ? kL2EvictLast : kL2EvictNormal;
  SFNO_CUDA(launch_pdl(dft_inv_p2_kernel<R, T, NST, MAPS>, dim3(grid), dim3(256), S::kBytes, stream, maps,
                       static_cast<T*>(y), tables, mscale, B, C, K, pitch, m_valid, sched,
                       (long long)Ctot * K * S::W, pol_y, pol_xm));
  SFNO_LAUNCHED();
  return 0;
}

template <int R, typename T>
static int launch_inv_p2(const float* xm, int B, int C, int Ctot, int K, int Kp, int M, int m_valid, const float* mscale,
                         void* y, cudaStream_t stream, int band = 0) {
  LocalMap lm;
  if (xm_stage_map(&lm.map, xm, B, C, Kp, std::max(1, std::min(M, m_valid)))) return 1;
  return launch_inv_p2_maps<R, T, LocalMap>(lm, B, C, Ctot, K, Kp, std::min(M, m_valid), mscale, y, stream, band);
}

// channel-sharded XM_ch buffers of all ranks as seen from this rank's latitude slab (see PeerMaps)
static int make_peer_maps(PeerMaps* pm, void* const* xm_ch, const int* chan_shapes, int nranks, int B, int Kg,
                          int lat_off, int lat_span, int mdim) {
  SFNO_CHECK(nranks >= 1 && nranks <= kMaxPeers, "peer DFT: 1..8 ranks");
  SFNO_CHECK(lat_off % 4 == 0 && lat_span > 0 && lat_off + lat_span <= Kg && Kg % 4 == 0, "peer DFT: bad latitude slab");
  pm->n = nranks;
  int c = 0;
  for (int p = 0; p < nranks; ++p) {
    pm->c0[p] = c;
    c += chan_shapes[p];
    const int Cpp = (chan_shapes[p] + 3) & ~3;
    if (chan_shapes[p] == 0) {
      pm->map[p] = pm->map[0];                    // never addressed (no channel maps to it)
      continue;
    }
    SFNO_CHECK(xm_ch[p] != nullptr, "peer DFT: null buffer");
    uint64_t dims[4] = {(uint64_t)lat_span, (uint64_t)chan_shapes[p], (uint64_t)2 * B, (uint64_t)mdim};
    uint64_t str[3] = {(uint64_t)Kg * 4, (uint64_t)Cpp * Kg * 4, (uint64_t)2 * B * Cpp * Kg * 4};
    uint32_t box[4] = {16, 1, 1, 241};
    if (encode_tensor_map(&pm->map[p], static_cast<const float*>(xm_ch[p]) + lat_off, 4, dims, str, box, 3)) return 1;
  }
  for (int p = nranks; p <= kMaxPeers; ++p) pm->c0[p] = c;
  for (int p = nranks; p < kMaxPeers; ++p) pm->map[p] = pm->map[0];
  return 0;
}

}  // namespace sfno
