// Host side of the tcgen05 batched GEMM: tile-list construction (triangular skipping, cost-sorted),
// TMA tensor-map encoding, launches.  See gemm_tc.cuh for the kernel and the five contraction modes.
#include "gemm_tc.cuh"
#include "plan.h"

#include <algorithm>
#include <map>
#include <mutex>
#include <tuple>
#include <vector>

namespace sfno {

// ------------------------------------------------------------------ TMA encode through the driver entry point
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode_fn() {
  static PFN_encodeTiled fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  });
  return fn;
}

// L2 fetch granularity of TMA loads; SFNO_L2_PROMO = 64 | 128 | 256 (default 128) for A/B measurements
static CUtensorMapL2promotion l2_promotion() {
  static int v = -1;
  if (v < 0) {
    const char* env = getenv("SFNO_L2_PROMO");
    v = env ? atoi(env) : 128;
  }
  return v == 256 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : (v == 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
}

int encode_tensor_map(CUtensorMap* map, const void* gptr, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, int swizzle) {
  PFN_encodeTiled fn = get_encode_fn();
  SFNO_CHECK(fn != nullptr, "cuTensorMapEncodeTiled entry point not available (no CUDA driver?)");
  cuuint64_t d[5];
  cuuint64_t s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    b[i] = box[i];
    e[i] = 1;
  }
  for (int i = 0; i + 1 < rank; ++i) {
    s[i] = strides_bytes[i];
    SFNO_CHECK(s[i] % 16 == 0, "TMA global stride must be a multiple of 16 bytes");
  }
  SFNO_CHECK((reinterpret_cast<uintptr_t>(gptr) & 15) == 0, "TMA global address must be 16-byte aligned");
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, (cuuint32_t)rank, const_cast<void*>(gptr), d, s, b, e,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_128B
                               : (swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                               : (swizzle == 3 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE)),
                  l2_promotion(), CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  SFNO_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with code " + std::to_string((int)r));
  return 0;
}

// ------------------------------------------------------------------ tile lists (cached on the device)
struct TileList {
  GemmTile* dev = nullptr;
  int count = 0;
};
typedef std::tuple<int, int, int, int, int, int, int, int, int> TileKey;   // device, mode, dims...
static std::map<TileKey, TileList> g_tile_cache;
static std::mutex g_tile_mutex;

static int floor_to(int x, int m) { return x / m * m; }

// valid local-m extent of row l of an SP tensor: global m < min(M, round_up(l_glob + 1, 16))
int sp_mhi(int l_local, int l_offset, int m_offset, int M_local) {
  int hi = round_up(l_local + l_offset + 1, 16) - m_offset;
  return std::max(0, std::min(M_local, hi));
}

static int upload_tiles(std::vector<GemmTile>& tiles, TileList* out) {
  // Longest-processing-time first: static round-robin over a cost-sorted list balances the triangular work.
  // The cost of a tile is taken as the largest cost inside its batch index g, so that all tiles of one g
  // (which share an operand: the table of order m, the weight of degree l) stay adjacent in the list and run
  // at the same time on neighbouring CTAs - the shared operand is then read from HBM once and hits L2 after.
  std::map<int, long long> gcost;
  for (const GemmTile& t : tiles) {
    long long c = (long long)t.nkk * t.n_mma;
    auto it = gcost.find(t.g);
    if (it == gcost.end() || it->second < c) gcost[t.g] = c;
  }
  std::stable_sort(tiles.begin(), tiles.end(), [&](const GemmTile& a, const GemmTile& b) {
    const long long ca = gcost[a.g], cb = gcost[b.g];
    if (ca != cb) return ca > cb;
    return a.g > b.g;
  });
  out->count = (int)tiles.size();
  if (tiles.empty()) return 0;
  SFNO_CUDA(cudaMalloc(&out->dev, tiles.size() * sizeof(GemmTile)));
  SFNO_CUDA(cudaMemcpy(out->dev, tiles.data(), tiles.size() * sizeof(GemmTile), cudaMemcpyHostToDevice));
  return 0;
}

static int get_tiles(const TileKey& key, const std::function<void(std::vector<GemmTile>&)>& build, TileList* out) {
  std::lock_guard<std::mutex> lock(g_tile_mutex);
  auto it = g_tile_cache.find(key);
  if (it != g_tile_cache.end()) {
    *out = it->second;
    return 0;
  }
  std::vector<GemmTile> tiles;
  build(tiles);
  TileList tl;
  if (upload_tiles(tiles, &tl)) return 1;
  g_tile_cache[key] = tl;
  *out = tl;
  return 0;
}

int sm_count_current_device() {
  static std::atomic<int> cache[64];
  const int dev = PerDeviceOnce::device();
  int n = cache[dev].load(std::memory_order_relaxed);
  if (!n) {
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}
static int num_sms() { return sm_count_current_device(); }

// SFNO_L2_HINTS=1 enables L2 eviction-priority hints (keep SP resident between producer and consumer kernels).
// Measured on B200: no gain (14.25 vs 14.20 ms/step), so they are off by default.
static bool l2_hints() {
  static int v = -1;
  if (v < 0) {
    const char* env = getenv("SFNO_L2_HINTS");
    v = (env && env[0] == '1') ? 1 : 0;
  }
  return v == 1;
}
// Channel-window mode (XM produced and consumed window by window through a small re-used buffer): keep the window and
// the Legendre table in L2 (evict_last), let the field rows stream (evict_first).  SFNO_WINDOW_HINTS=0 turns the
// hints off for A/B measurements.
bool window_hints() {
  static int v = -1;
  if (v < 0) {
    const char* env = getenv("SFNO_WINDOW_HINTS");
    v = (env && env[0] == '0') ? 0 : 1;
  }
  return v == 1;
}
// keep: the tensor is re-read soon by the next kernel of the chain and fits L2 (SP, 89 MB defined);
// stream: read or written once and larger than what L2 can usefully hold (XM, weights)
static void set_policies(GemmArgs& a, bool keep_a, bool keep_b, bool keep_out, bool stream_a, bool stream_b,
                         bool stream_out) {
  a.pol_a = a.pol_b = a.pol_out = kL2EvictNormal;
  if (!l2_hints()) return;
  if (keep_a) a.pol_a = kL2EvictLast;
  if (keep_b) a.pol_b = kL2EvictLast;
  if (keep_out) a.pol_out = kL2EvictLast;
  if (stream_a) a.pol_a = kL2EvictFirst;
  if (stream_b) a.pol_b = kL2EvictFirst;
  if (stream_out) a.pol_out = kL2EvictFirst;
}

// Dynamic tile scheduler state: a pool of {next tile, finished CTAs} counter pairs on the device, used round-robin by
// successive launches.  A kernel finds its pair zeroed and its last CTA zeroes it again, so no memset is enqueued
// (calls stay CUDA-graph capturable); 64 pairs keep launches that may overlap on different streams apart.
int* sched_counters() {
  static std::map<int, int*> pool;
  static std::map<int, unsigned> next;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  int dev = 0;
  cudaGetDevice(&dev);
  auto it = pool.find(dev);
  if (it == pool.end()) {
    int* p = nullptr;
    if (cudaMalloc(&p, 64 * 2 * sizeof(int)) != cudaSuccess) return nullptr;
    cudaMemset(p, 0, 64 * 2 * sizeof(int));
    it = pool.emplace(dev, p).first;
    next[dev] = 0;
  }
  return it->second + 2 * (next[dev]++ % 64);
}

template <int MODE>
static int launch(const CUtensorMap& ma, const CUtensorMap& mb, GemmArgs args, cudaStream_t stream) {
  using Cfg = GemmCfg<MODE>;
  if (args.num_tiles == 0) return 0;
  args.sched = sched_counters();
  SFNO_CHECK(args.sched != nullptr, "could not allocate the tile scheduler counters");
  SFNO_CHECK(args.out_col_stride > 0 && args.out_col_stride < (1LL << 31) / 32,
             "output column stride too large for the 32-bit epilogue addressing (reduce batch*channels)");
  static PerDeviceOnce attr_set;
  if (attr_set.needed()) {
    SFNO_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   Cfg::kSmemBytes));
    attr_set.mark();
  }
  int grid = std::min(args.num_tiles, num_sms());
  SFNO_CUDA(launch_pdl(gemm_tc_kernel<MODE>, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, stream, ma, mb, args));
  SFNO_LAUNCHED();
  return 0;
}

static int current_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev;
}

// ------------------------------------------------------------------ Legendre analysis  (K2)
// Channel window: XM holds channels [c0, c0 + C) of a field with Ctot channels (compact: [M][2B][round_up(C,4)][Kp]);
// SP is the full tensor over Ctot channels.  Ctot == C, c0 == 0: no window.
static int window_ok(int C, int Ctot, int c0) {
  if (Ctot == C && c0 == 0) return 0;
  SFNO_CHECK(c0 >= 0 && c0 + C <= Ctot && C % 32 == 0 && c0 % 4 == 0,
             "channel window: needs 0 <= c0, c0 + C <= Ctot, C % 32 == 0, c0 % 4 == 0");
  return 0;
}

int legendre_analysis(const sfno_plan* plan, const float* xm, int B, int C, int Ctot, int c0, float* sp, int round_out,
                      int accumulate, cudaStream_t stream) {
  const int K = plan->nlat, Kp = plan->Kp, L = plan->lmax, M = plan->mmax, moff = plan->m_offset;
  const int RS = 2 * B * round_up(C, 4);                 // rows of the XM operand (the window)
  const int RSt = 2 * B * round_up(Ctot, 4);             // rows of SP
  if (window_ok(C, Ctot, c0)) return 1;
  const bool win = !(Ctot == C && c0 == 0);
  CUtensorMap ma, mb;
  {
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)RS, (uint64_t)M};
    uint64_t str[2] = {(uint64_t)Kp * 4, (uint64_t)RS * Kp * 4};
    uint32_t box[3] = {32, 128, 1};
    if (encode_tensor_map(&ma, xm, 3, dims, str, box, 1)) return 1;
  }
  {
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)L, (uint64_t)M};
    uint64_t str[2] = {(uint64_t)Kp * 4, (uint64_t)L * Kp * 4};
    uint32_t box[3] = {32, 64, 1};
    if (encode_tensor_map(&mb, plan->table, 3, dims, str, box, 1)) return 1;
  }
  TileList tl;
  TileKey key{current_device(), LEG_ANA, K, L, M, moff, RS, 0, 0};
  if (get_tiles(key, [&](std::vector<GemmTile>& tiles) {
        for (int m = 0; m < M; ++m) {
          const int n0 = moff == SFNO_M_OFFSET_DENSE ? 0 : floor_to(m + moff, 16);   // dense table: no l >= m structure
          for (int c0 = n0; c0 < L; c0 += 256) {
            const int ncols = std::min(256, L - c0);
            for (int r0 = 0; r0 < RS; r0 += 128) {
              GemmTile t{m, r0, c0, round_up(ncols, 16), 0, (K + 7) / 8, c0 + ncols, 0};
              tiles.push_back(t);
            }
          }
        }
      }, &tl))
    return 1;
  GemmArgs args;
  args.out = sp;
  args.tiles = tl.dev;
  args.num_tiles = tl.count;
  args.row_limit = RS;
  args.out_g_stride = RSt;                      // SP[l][m][r]: g = m
  args.out_col_stride = (long long)M * RSt;     // column = l
  args.out_im_offset = 0;
  args.round_out = round_out;
  args.accumulate = accumulate;
  args.grp_rows = win ? round_up(C, 4) : 0;
  args.grp_stride = round_up(Ctot, 4);
  args.grp_off = c0;
  set_policies(args, false, false, true, true, false, false);     // XM streams through, SP stays for the channel contraction
  if (win && window_hints()) {                  // the XM window and the table are re-used from L2 (see window_hints)
    args.pol_a = kL2EvictLast;
    args.pol_b = kL2EvictLast;
  }
  return launch<LEG_ANA>(ma, mb, args, stream);
}

// ------------------------------------------------------------------ Legendre synthesis  (K4)
int legendre_synthesis(const sfno_plan* plan, const float* sp, int B, int C, int Ctot, int c0, float* xm,
                       cudaStream_t stream) {
  const int K = plan->nlat, Kp = plan->Kp, L = plan->lmax, M = plan->mmax, moff = plan->m_offset;
  const int RS = 2 * B * round_up(C, 4);                 // rows of the XM result (the window)
  const int RSt = 2 * B * round_up(Ctot, 4);             // rows of SP
  if (window_ok(C, Ctot, c0)) return 1;
  const bool win = !(Ctot == C && c0 == 0);
  CUtensorMap ma, mb;
  {
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)L, (uint64_t)M};
    uint64_t str[2] = {(uint64_t)Kp * 4, (uint64_t)L * Kp * 4};
    uint32_t box[3] = {32, 32, 1};
    if (encode_tensor_map(&ma, plan->table, 3, dims, str, box, 2)) return 1;
  }
  {
    uint64_t dims[3] = {(uint64_t)RSt, (uint64_t)M, (uint64_t)L};
    uint64_t str[2] = {(uint64_t)RSt * 4, (uint64_t)M * RSt * 4};
    uint32_t box[3] = {32, 1, 32};
    if (encode_tensor_map(&mb, sp, 3, dims, str, box, 2)) return 1;
  }
  TileList tl;
  TileKey key{current_device(), LEG_SYN, K, L, M, moff, RS, 0, 0};
  if (get_tiles(key, [&](std::vector<GemmTile>& tiles) {
        for (int m = 0; m < M; ++m) {
          const int l0 = moff == SFNO_M_OFFSET_DENSE ? 0 : std::min(L, floor_to(m + moff, 8));
          const int nkk = (L - l0 + 7) / 8;        // 0 => this m has no degrees: zeros are written
          for (int c0 = 0; c0 < RS; c0 += 256) {
            const int ncols = std::min(256, RS - c0);
            for (int r0 = 0; r0 < K; r0 += 128) {
              GemmTile t{m, r0, c0, round_up(ncols, 16), l0, nkk, c0 + ncols, 0};
              tiles.push_back(t);
            }
          }
        }
      }, &tl))
    return 1;
  GemmArgs args;
  args.out = xm;
  args.tiles = tl.dev;
  args.num_tiles = tl.count;
  args.row_limit = K;
  args.out_g_stride = (long long)RS * Kp;       // XM[m][r][k]: g = m
  args.out_col_stride = Kp;                     // column = r
  args.out_im_offset = 0;
  args.round_out = 0;                           // consumed by the fp32 inverse DFT
  args.accumulate = 0;
  args.grp_rows = win ? round_up(C, 4) : 0;
  args.grp_stride = round_up(Ctot, 4);
  args.grp_off = c0;
  set_policies(args, false, false, false, false, true, true);     // SP is consumed here; XM (178 MB) cannot stay in L2
  if (win && window_hints()) {                  // the table is re-read by every window, the XM window is consumed from L2
    args.pol_a = kL2EvictLast;
    args.pol_out = kL2EvictLast;
  }
  return launch<LEG_SYN>(ma, mb, args, stream);
}

// ------------------------------------------------------------------ dhconv  (K3 fwd / bwd-x / bwd-w)
// SP_in  : [L][M][B][2][Cinp]     SP_out : [L][M][B][2][Coutp]     WP : [L][2][Coutp][Cinp]
static int sp_map_kmajor(CUtensorMap* map, const float* sp, int C, int B, int L, int M) {
  const int Cp = round_up(C, 4);
  uint64_t dims[4] = {(uint64_t)C, 2, (uint64_t)M * B, (uint64_t)L};
  uint64_t str[3] = {(uint64_t)Cp * 4, (uint64_t)2 * Cp * 4, (uint64_t)M * B * 2 * Cp * 4};
  uint32_t box[4] = {32, 1, 64, 1};
  return encode_tensor_map(map, sp, 4, dims, str, box, 1);
}
static int sp_map_mnmajor(CUtensorMap* map, const float* sp, int C, int B, int L, int M) {
  const int Cp = round_up(C, 4);
  uint64_t dims[4] = {(uint64_t)C, 2, (uint64_t)M * B, (uint64_t)L};
  uint64_t str[3] = {(uint64_t)Cp * 4, (uint64_t)2 * Cp * 4, (uint64_t)M * B * 2 * Cp * 4};
  uint32_t box[4] = {32, 1, 32, 1};
  return encode_tensor_map(map, sp, 4, dims, str, box, 2);
}
static int wp_map(CUtensorMap* map, const float* wp, int Cin, int Cout, int L, bool mn_major) {
  const int Cinp = round_up(Cin, 4), Coutp = round_up(Cout, 4);
  uint64_t dims[4] = {(uint64_t)Cin, (uint64_t)Cout, 2, (uint64_t)L};
  uint64_t str[3] = {(uint64_t)Cinp * 4, (uint64_t)Coutp * Cinp * 4, (uint64_t)2 * Coutp * Cinp * 4};
  uint32_t box_k[4] = {32, 128, 1, 1};
  uint32_t box_mn[4] = {32, 32, 1, 1};
  return encode_tensor_map(map, wp, 4, dims, str, mn_major ? box_mn : box_k, mn_major ? 2 : 1);
}

int dhconv_fwd(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout, int L, int M,
               int l_offset, int m_offset, int round_out, cudaStream_t stream) {
  CUtensorMap ma, mb;
  if (wp_map(&ma, wp, Cin, Cout, L, false)) return 1;
  if (sp_map_kmajor(&mb, sp_in, Cin, B, L, M)) return 1;
  TileList tl;
  TileKey key{current_device(), DH_FWD, Cin, Cout, L, M, B, l_offset, m_offset};
  if (get_tiles(key, [&](std::vector<GemmTile>& tiles) {
        for (int l = 0; l < L; ++l) {
          const int nn = sp_mhi(l, l_offset, m_offset, M) * B;
          for (int c0 = 0; c0 < nn; c0 += 128) {
            const int ncols = std::min(128, nn - c0);
            for (int r0 = 0; r0 < Cout; r0 += 128) {
              GemmTile t{l, r0, c0, round_up(ncols, 16), 0, (Cin + 7) / 8, c0 + ncols, 0};
              tiles.push_back(t);
            }
          }
        }
      }, &tl))
    return 1;
  const int Coutp = round_up(Cout, 4);
  GemmArgs args;
  args.out = sp_out;
  args.tiles = tl.dev;
  args.num_tiles = tl.count;
  args.row_limit = Cout;
  args.out_g_stride = (long long)M * B * 2 * Coutp;
  args.out_col_stride = 2 * Coutp;              // column = n = m*B + b
  args.out_im_offset = Coutp;
  args.round_out = round_out;
  args.accumulate = args.grp_rows = args.grp_stride = args.grp_off = 0;
  set_policies(args, false, false, true, true, false, false);     // weights stream, the result feeds the synthesis
  return launch<DH_FWD>(ma, mb, args, stream);
}

int dhconv_bwd_x(const float* gsp_out, const float* wp, float* gsp_in, int B, int Cin, int Cout, int L, int M,
                 int l_offset, int m_offset, int round_out, cudaStream_t stream) {
  CUtensorMap ma, mb;
  if (wp_map(&ma, wp, Cin, Cout, L, true)) return 1;
  if (sp_map_kmajor(&mb, gsp_out, Cout, B, L, M)) return 1;
  TileList tl;
  TileKey key{current_device(), DH_BWDX, Cin, Cout, L, M, B, l_offset, m_offset};
  if (get_tiles(key, [&](std::vector<GemmTile>& tiles) {
        for (int l = 0; l < L; ++l) {
          const int nn = sp_mhi(l, l_offset, m_offset, M) * B;
          for (int c0 = 0; c0 < nn; c0 += 128) {
            const int ncols = std::min(128, nn - c0);
            for (int r0 = 0; r0 < Cin; r0 += 128) {
              GemmTile t{l, r0, c0, round_up(ncols, 16), 0, (Cout + 7) / 8, c0 + ncols, 0};
              tiles.push_back(t);
            }
          }
        }
      }, &tl))
    return 1;
  const int Cinp = round_up(Cin, 4);
  GemmArgs args;
  args.out = gsp_in;
  args.tiles = tl.dev;
  args.num_tiles = tl.count;
  args.row_limit = Cin;
  args.out_g_stride = (long long)M * B * 2 * Cinp;
  args.out_col_stride = 2 * Cinp;
  args.out_im_offset = Cinp;
  args.round_out = round_out;
  args.accumulate = args.grp_rows = args.grp_stride = args.grp_off = 0;
  set_policies(args, false, false, true, true, false, false);
  return launch<DH_BWDX>(ma, mb, args, stream);
}

int dhconv_bwd_w(const float* sp_in, const float* gsp_out, float* gwp, int B, int Cin, int Cout, int L, int M,
                 int l_offset, int m_offset, cudaStream_t stream) {
  CUtensorMap ma, mb;
  if (sp_map_mnmajor(&ma, sp_in, Cin, B, L, M)) return 1;
  if (sp_map_mnmajor(&mb, gsp_out, Cout, B, L, M)) return 1;
  TileList tl;
  TileKey key{current_device(), DH_BWDW, Cin, Cout, L, M, B, l_offset, m_offset};
  if (get_tiles(key, [&](std::vector<GemmTile>& tiles) {
        for (int l = 0; l < L; ++l) {
          const int nn = sp_mhi(l, l_offset, m_offset, M) * B;      // contraction extent (multiple of 8 or clipped by M*B)
          for (int c0 = 0; c0 < Cout; c0 += 128) {
            const int ncols = std::min(128, Cout - c0);
            for (int r0 = 0; r0 < Cin; r0 += 128) {
              GemmTile t{l, r0, c0, round_up(ncols, 16), 0, (nn + 7) / 8, c0 + ncols, 0};
              tiles.push_back(t);
            }
          }
        }
      }, &tl))
    return 1;
  const int Cinp = round_up(Cin, 4), Coutp = round_up(Cout, 4);
  GemmArgs args;
  args.out = gwp;
  args.tiles = tl.dev;
  args.num_tiles = tl.count;
  args.row_limit = Cin;
  args.out_g_stride = (long long)2 * Coutp * Cinp;   // WP[l][reim][o][i]: rows = i, columns = o
  args.out_col_stride = Cinp;
  args.out_im_offset = (long long)Coutp * Cinp;
  args.round_out = 0;
  args.accumulate = args.grp_rows = args.grp_stride = args.grp_off = 0;
  set_policies(args, false, true, false, true, false, true);      // dY is read again by the dX contraction
  return launch<DH_BWDW>(ma, mb, args, stream);
}

// ------------------------------------------------------------------ dhconv forward without any allocation
// sfno_dhconv_fwd_static (include/makani_b200.h): the contraction of dhconv_fwd for callers that may not allocate
// (CUDA-graph capture, serving loops).  What dhconv_fwd keeps in cudaMalloc'ed memory - the tile list of the shape and
// the {next tile, finished CTAs} pair of the dynamic scheduler - lives here in one file-scope device array:
//     bytes [0, 256)   : the scheduler pair (two ints, rest unused)
//     bytes [256, ...) : up to SFNO_STATIC_MAX_TILES tiles of 32 bytes, written by dhfwd_static_tiles_kernel
// Every GPU that loads the library holds the array (0.5 MB) until exit.  All calls share it and it keeps the previous
// call's contents: the tile kernel rewrites the list and zeroes the scheduler pair at the start of every call, in
// stream order before the GEMM.  Two calls that overlap on the GPU (different streams) race on it.
constexpr int kStaticMaxTiles = SFNO_STATIC_MAX_TILES;
constexpr int kStaticMaxL = SFNO_STATIC_MAX_L;
constexpr int kStaticTileOffset = 256;
constexpr size_t kStaticBytes = kStaticTileOffset + (size_t)kStaticMaxTiles * sizeof(GemmTile);
static __device__ __align__(256) unsigned char g_dhfwd_static_scratch[kStaticBytes];

// One block.  Tiles of degree l in the order dhconv_fwd builds them (columns outer, rows inner), degrees from the
// highest down - the order its cost sort produces for all but the few lowest degrees; the order only affects the
// load balance, every tile is computed independently.
__global__ void __launch_bounds__(256)
dhfwd_static_tiles_kernel(int L, int M, int B, int Cin, int Cout, int l_offset, int m_offset) {
  __shared__ int start[kStaticMaxL + 1];
  GemmTile* tiles = reinterpret_cast<GemmTile*>(g_dhfwd_static_scratch + kStaticTileOffset);
  int* sched = reinterpret_cast<int*>(g_dhfwd_static_scratch);
  const int row_tiles = (Cout + 127) / 128;
  auto columns = [&](int l) {                  // sp_mhi(l) * B
    int hi = (l + l_offset + 1 + 15) / 16 * 16 - m_offset;
    hi = hi < 0 ? 0 : (hi > M ? M : hi);
    return hi * B;
  };
  for (int p = threadIdx.x; p < L; p += blockDim.x) start[p + 1] = (columns(L - 1 - p) + 127) / 128 * row_tiles;
  if (threadIdx.x == 0) {
    start[0] = 0;
    sched[0] = 0;
    sched[1] = 0;
  }
  __syncthreads();
  if (threadIdx.x == 0)
    for (int p = 0; p < L; ++p) start[p + 1] += start[p];
  __syncthreads();
  for (int p = threadIdx.x; p < L; p += blockDim.x) {
    const int l = L - 1 - p, nn = columns(l);
    int t = start[p];
    for (int c0 = 0; c0 < nn; c0 += 128) {
      const int ncols = nn - c0 < 128 ? nn - c0 : 128;
      for (int r0 = 0; r0 < Cout; r0 += 128) {
        if (t < kStaticMaxTiles) tiles[t] = GemmTile{l, r0, c0, (ncols + 15) / 16 * 16, 0, (Cin + 7) / 8, c0 + ncols, 0};
        ++t;
      }
    }
  }
}

// device address of the array on the current device (the name itself is only a host-side handle of the symbol)
static int dhfwd_static_scratch(unsigned char** out) {
  static unsigned char* addr[64];
  static PerDeviceOnce looked_up;
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  const int dev = PerDeviceOnce::device();
  if (looked_up.needed()) {
    void* p = nullptr;
    SFNO_CUDA(cudaGetSymbolAddress(&p, g_dhfwd_static_scratch));
    SFNO_CHECK(p != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch array");
    addr[dev] = static_cast<unsigned char*>(p);
    looked_up.mark();
  }
  *out = addr[dev];
  return 0;
}

static int dhconv_fwd_static(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout, int L, int M,
                             int l_offset, int m_offset, int round_out, cudaStream_t stream) {
  using Cfg = GemmCfg<DH_FWD>;
  SFNO_CHECK(B > 0 && Cin > 0 && Cout > 0 && L > 0 && M > 0, "sfno_dhconv_fwd_static: non-positive size");
  SFNO_CHECK(L <= kStaticMaxL, "sfno_dhconv_fwd_static: lmax " + std::to_string(L) + " exceeds SFNO_STATIC_MAX_L = " +
                                   std::to_string(kStaticMaxL) + " (use sfno_dhconv_fwd)");
  long long count = 0;
  for (int l = 0; l < L; ++l) count += (long long)((sp_mhi(l, l_offset, m_offset, M) * B + 127) / 128) * ((Cout + 127) / 128);
  SFNO_CHECK(count <= kStaticMaxTiles,
             "sfno_dhconv_fwd_static: the shape needs " + std::to_string(count) + " tiles, the static scratch holds " +
                 "SFNO_STATIC_MAX_TILES = " + std::to_string(kStaticMaxTiles) + " (use sfno_dhconv_fwd)");
  if (count == 0) return 0;
  unsigned char* scratch = nullptr;
  if (dhfwd_static_scratch(&scratch)) return 1;
  CUtensorMap ma, mb;
  if (wp_map(&ma, wp, Cin, Cout, L, false)) return 1;
  if (sp_map_kmajor(&mb, sp_in, Cin, B, L, M)) return 1;
  const int Coutp = round_up(Cout, 4);
  GemmArgs args;
  args.out = sp_out;
  args.tiles = reinterpret_cast<const GemmTile*>(scratch + kStaticTileOffset);
  args.num_tiles = (int)count;
  args.row_limit = Cout;
  args.out_g_stride = (long long)M * B * 2 * Coutp;
  args.out_col_stride = 2 * Coutp;
  args.out_im_offset = Coutp;
  args.round_out = round_out;
  args.accumulate = args.grp_rows = args.grp_stride = args.grp_off = 0;
  set_policies(args, false, false, true, true, false, false);
  args.sched = reinterpret_cast<int*>(scratch);
  SFNO_CHECK(args.out_col_stride < (1LL << 31) / 32, "output column stride too large for the 32-bit epilogue addressing");
  static PerDeviceOnce attr_set;
  if (attr_set.needed()) {
    SFNO_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<DH_FWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set.mark();
  }
  dhfwd_static_tiles_kernel<<<1, 256, 0, stream>>>(L, M, B, Cin, Cout, l_offset, m_offset);
  SFNO_LAUNCHED();
  const int grid = std::min(args.num_tiles, num_sms());
  // the GEMM reads the list and the scheduler pair only after its griddepcontrol.wait, i.e. after the tile kernel
  SFNO_CUDA(launch_pdl(gemm_tc_kernel<DH_FWD>, dim3(grid), dim3(kGemmThreads), Cfg::kSmemBytes, stream, ma, mb, args));
  SFNO_LAUNCHED();
  return 0;
}

}  // namespace sfno

extern "C" int sfno_dhconv_fwd_static(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout,
                                      int lmax, int mmax, int l_offset, int m_offset, int round_tf32, void* stream) {
  SFNO_CHECK(sp_in && wp && sp_out, "null argument");
  return sfno::dhconv_fwd_static(sp_in, wp, sp_out, B, Cin, Cout, lmax, mmax, l_offset, m_offset, round_tf32,
                                 (cudaStream_t)stream);
}
