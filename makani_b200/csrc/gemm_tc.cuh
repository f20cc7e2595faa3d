// Persistent, warp-specialised, batched TF32 GEMM on tcgen05 tensor cores (sm_100a).
//
//   one kernel template, five "modes" = the five dense contractions of the SFNO spectral path
//     LEG_ANA   SP[l][m][r]  = sum_k XM[m][r][k] * T[m][l][k]              (per m; real)
//     LEG_SYN   XM[m][r][k]  = sum_l T[m][l][k]  * SP[l][m][r]             (per m; real; both operands MN-major)
//     DH_FWD    Y[l][n][o]   = sum_i W[l][o][i]  * X[l][n][i]              (per l; complex, planar re/im)
//     DH_BWDX   dX[l][n][i]  = sum_o conj(W[l][o][i]) * dY[l][n][o]        (per l; complex; A MN-major)
//     DH_BWDW   dW[l][o][i]  = sum_n conj(X[l][n][i]) * dY[l][n][o]        (per l; complex; both MN-major)
//
//   roles   warp 0: TMA producer (one lane)     warp 1: tcgen05.mma issuer (one lane)
//           warp 2: TMEM allocator              warps 4-7: epilogue (TMEM -> registers -> coalesced st.global)
//   pipes   smem ring  full[]/empty[]  (TMA <-> MMA),  TMEM double buffer  tmem_full[]/tmem_empty[] (MMA <-> epilogue)
//   tiles   host-built, cost-sorted tile list (triangular l >= m structure is skipped at 16/8 granularity), handed out
//           through a global atomic counter: the producer warp claims the next tile and publishes it to the MMA and
//           epilogue warps through a small ring, so CTAs that start late (SMs shared with NCCL) just take less
//
//   D is always stored "row contiguous": TMEM lane == output row == fastest memory index, so each
//   st.global of a warp writes 128 contiguous bytes without staging through shared memory.
#pragma once
#include "common.cuh"

namespace sfno {

enum GemmMode { LEG_ANA = 0, LEG_SYN = 1, DH_FWD = 2, DH_BWDX = 3, DH_BWDW = 4 };

struct GemmTile {     // 32 bytes
  int g;              // batch index (m for Legendre, l for dhconv) - local index into the tensors
  int row0;           // first output row of the tile
  int col0;           // first output column of the tile
  int n_mma;          // UMMA N for this tile (multiple of 16, <= BN)
  int k0;             // first contraction index
  int nkk;            // number of 8-wide contraction steps (0 => output tile is all zeros)
  int col_limit;      // columns >= col_limit are not stored
  int pad;
};

struct GemmArgs {
  float* out;
  const GemmTile* tiles;
  int num_tiles;
  int row_limit;            // rows >= row_limit are not stored
  long long out_g_stride;   // output offset per batch index g
  long long out_col_stride; // output offset per column
  long long out_im_offset;  // complex modes: offset of the imaginary plane
  int round_out;            // 1: store round-to-nearest TF32 values (the consumer is another TF32 GEMM)
  unsigned long long pol_a, pol_b, pol_out;   // L2 cache policies of the two operand streams and of the output
  int* sched;               // [0] next tile (dynamic scheduler), [1] finished CTAs; both zero at launch, reset by the last CTA
  // Channel-window view (Legendre modes): the XM operand / result is a compact tensor over a WINDOW of `grp_rows`
  // (padded) channels per (batch, re/im) group, while SP is the full tensor with `grp_stride` channels per group:
  //   SP row (LEG_ANA output row / LEG_SYN B column)  =  (r / grp_rows) * grp_stride + grp_off + r % grp_rows
  // grp_rows == 0: no window (r maps to itself).  LEG_SYN needs grp_rows % 32 == 0 (one TMA box never straddles groups).
  int grp_rows, grp_stride, grp_off;
  int accumulate;           // 1: add to what the output already holds (LEG_ANA: gradients of two consumers of one SP)
};

template <int MODE>
struct GemmCfg {
  static constexpr bool kComplex = (MODE >= DH_FWD);
  static constexpr bool kAMn = (MODE == LEG_SYN || MODE == DH_BWDX || MODE == DH_BWDW);
  static constexpr bool kBMn = (MODE == LEG_SYN || MODE == DH_BWDW);
  static constexpr int kBN = kComplex ? 128 : 256;           // max UMMA N per accumulator
  static constexpr int kStages = kComplex ? 3 : 4;
  static constexpr int kPlanes = kComplex ? 2 : 1;
  static constexpr int kABytes = 128 * 128;                  // 128 rows x 32 fp32
  static constexpr int kBBytes = kBN * 128;
  static constexpr int kStageBytes = kPlanes * (kABytes + kBBytes);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align*/ + 256 /*barriers*/;
  static constexpr int kAccCols = 256;                       // TMEM columns per accumulator buffer
};

constexpr int kGemmThreads = 256;
constexpr int kSchedDepth = 4;     // tiles the producer may claim ahead of the slowest consumer warp (power of two)

template <int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
               const GemmArgs args) {
  using Cfg = GemmCfg<MODE>;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);   // stays a shared-space pointer
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + Cfg::kStages * Cfg::kStageBytes);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + Cfg::kStages;
  uint64_t* tmem_full = bars + 2 * Cfg::kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* sched_full = tmem_empty + 2;
  uint64_t* sched_empty = sched_full + kSchedDepth;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(sched_empty + kSchedDepth);
  volatile int* sched_tile = reinterpret_cast<volatile int*>(tmem_ptr_smem + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Cfg::kStages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full[b], 1);
      mbar_init(&tmem_empty[b], 128);
    }
    for (int b = 0; b < kSchedDepth; ++b) {
      mbar_init(&sched_full[b], 1);
      mbar_init(&sched_empty[b], 5);             // MMA lane + one lane of each epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_ptr_smem, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr_smem;
  // everything above overlapped the tail of the previous kernel in the stream (programmatic dependent launch);
  // from here on global memory written by it is read / buffers it read are overwritten
  pdl_launch_dependents();
  pdl_wait();

  auto stage_a = [&](int s, int plane) -> uint8_t* { return smem + s * Cfg::kStageBytes + plane * Cfg::kABytes; };
  auto stage_b = [&](int s, int plane) -> uint8_t* {
    return smem + s * Cfg::kStageBytes + Cfg::kPlanes * Cfg::kABytes + plane * Cfg::kBBytes;
  };

  if (warp == 0) {
    // =========================================================== TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      // first tile: static (blockIdx.x < num_tiles, no atomic on the start-up path); later tiles: claimed from the
      // counter, one tile ahead so that the atomic's latency hides behind the current tile's loads
      int t_next = (int)blockIdx.x;
      for (uint32_t sn = 0;; ++sn) {
        const int slot = sn & (kSchedDepth - 1);
        int t = t_next;
        if (t >= args.num_tiles) t = -1;
        else t_next = (int)gridDim.x + atomicAdd(args.sched, 1);
        mbar_wait(&sched_empty[slot], ((sn / kSchedDepth) & 1) ^ 1);
        sched_tile[slot] = t;
        mbar_arrive(&sched_full[slot]);
        if (t < 0) break;
        const GemmTile tile = args.tiles[t];
        if (tile.nkk == 0) continue;
        const int nkb = (tile.nkk + 3) >> 2;
        const int nb_boxes = Cfg::kBMn ? ((tile.n_mma + 31) >> 5) : ((tile.n_mma + 63) >> 6);
        const uint32_t bytes = Cfg::kPlanes * (Cfg::kABytes + nb_boxes * (Cfg::kBMn ? 4096 : 8192));
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          mbar_arrive_expect_tx(&full_bar[stage], bytes);
          const int kc = tile.k0 + kb * 32;
#pragma unroll
          for (int p = 0; p < Cfg::kPlanes; ++p) {
            uint8_t* sa = stage_a(stage, p);
            uint8_t* sb = stage_b(stage, p);
            // ---- A operand
            if constexpr (MODE == LEG_ANA) {                 // XM (k, r, m)  K-major
              tma_load_3d_hint(sa, &map_a, &full_bar[stage], kc, tile.row0, tile.g, args.pol_a);
            } else if constexpr (MODE == LEG_SYN) {          // table (k_lat, l, m)  MN-major: 4 chunks of 32 rows
#pragma unroll
              for (int j = 0; j < 4; ++j)
                tma_load_3d_hint(sa + j * 4096, &map_a, &full_bar[stage], tile.row0 + 32 * j, kc, tile.g, args.pol_a);
            } else if constexpr (MODE == DH_FWD) {           // WP (i, o, reim, l)  K-major
              tma_load_4d_hint(sa, &map_a, &full_bar[stage], kc, tile.row0, p, tile.g, args.pol_a);
            } else if constexpr (MODE == DH_BWDX) {          // WP (i, o, reim, l)  MN-major in i
#pragma unroll
              for (int j = 0; j < 4; ++j)
                tma_load_4d_hint(sa + j * 4096, &map_a, &full_bar[stage], tile.row0 + 32 * j, kc, p, tile.g, args.pol_a);
            } else {                                         // DH_BWDW: SP_in (i, reim, n, l)  MN-major in i
#pragma unroll
              for (int j = 0; j < 4; ++j)
                tma_load_4d_hint(sa + j * 4096, &map_a, &full_bar[stage], tile.row0 + 32 * j, p, kc, tile.g, args.pol_a);
            }
            // ---- B operand
            for (int j = 0; j < nb_boxes; ++j) {
              if constexpr (MODE == LEG_ANA) {               // table (k, l, m) K-major, 64-row boxes
                tma_load_3d_hint(sb + j * 8192, &map_b, &full_bar[stage], kc, tile.col0 + 64 * j, tile.g, args.pol_b);
              } else if constexpr (MODE == LEG_SYN) {        // SP (r, m, l) MN-major, 32-col chunks
                int rcol = tile.col0 + 32 * j;
                if (args.grp_rows) rcol = (rcol / args.grp_rows) * args.grp_stride + args.grp_off + rcol % args.grp_rows;
                tma_load_3d_hint(sb + j * 4096, &map_b, &full_bar[stage], rcol, tile.g, kc, args.pol_b);
              } else if constexpr (MODE == DH_FWD || MODE == DH_BWDX) {   // SP (c, reim, n, l) K-major
                tma_load_4d_hint(sb + j * 8192, &map_b, &full_bar[stage], kc, p, tile.col0 + 64 * j, tile.g, args.pol_b);
              } else {                                       // DH_BWDW: gSP_out (o, reim, n, l) MN-major in o
                tma_load_4d_hint(sb + j * 4096, &map_b, &full_bar[stage], tile.col0 + 32 * j, p, kc, tile.g, args.pol_b);
              }
            }
          }
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // =========================================================== MMA issuer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t acc_it = 0;
      for (uint32_t sn = 0;; ++sn) {
        const int slot = sn & (kSchedDepth - 1);
        mbar_wait(&sched_full[slot], (sn / kSchedDepth) & 1);
        const int t = sched_tile[slot];
        mbar_arrive(&sched_empty[slot]);
        if (t < 0) break;
        const GemmTile tile = args.tiles[t];
        if (tile.nkk == 0) continue;
        const uint32_t buf = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
        mbar_wait(&tmem_empty[buf], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_re = tmem_base + buf * Cfg::kAccCols;
        const uint32_t d_im = d_re + 128;
        const uint32_t n = (uint32_t)tile.n_mma;
        const uint32_t id_pp = umma_idesc_tf32(n, Cfg::kAMn, Cfg::kBMn, 0, 0);
        const uint32_t id_np = umma_idesc_tf32(n, Cfg::kAMn, Cfg::kBMn, 1, 0);   // negated A
        const int nkb = (tile.nkk + 3) >> 2;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const int steps = min(4, tile.nkk - 4 * kb);
          const uint32_t a0 = smem_u32(stage_a(stage, 0));
          const uint32_t b0 = smem_u32(stage_b(stage, 0));
          for (int kk = 0; kk < steps; ++kk) {
            const uint32_t acc = (kb | kk) ? 1u : 0u;
            // K-major: advance 8 tf32 = 32 B inside the 128 B swizzle row.  MN-major: 8 k-rows = 1024 B.
            const uint32_t aoff = Cfg::kAMn ? kk * 1024 : kk * 32;
            const uint32_t boff = Cfg::kBMn ? kk * 1024 : kk * 32;
            // K-major  : SWIZZLE_128B, 8-row groups 1024 B apart (SBO), LBO unused.
            // MN-major : SWIZZLE_128B_BASE32B - 32-element MN chunks 4096 B apart (LBO), 4-row K atoms 512 B apart (SBO).
            const uint32_t albo = Cfg::kAMn ? 4096 : 0, blbo = Cfg::kBMn ? 4096 : 0;
            const uint32_t asbo = Cfg::kAMn ? 512 : 1024, bsbo = Cfg::kBMn ? 512 : 1024;
            const uint32_t alay = Cfg::kAMn ? kUmmaSw128Base32 : kUmmaSw128;
            const uint32_t blay = Cfg::kBMn ? kUmmaSw128Base32 : kUmmaSw128;
            const uint64_t da0 = umma_smem_desc(a0 + aoff, albo, asbo, alay);
            const uint64_t db0 = umma_smem_desc(b0 + boff, blbo, bsbo, blay);
            if constexpr (!Cfg::kComplex) {
              umma_tf32(d_re, da0, db0, id_pp, acc);
            } else {
              const uint64_t da1 = umma_smem_desc(a0 + Cfg::kABytes + aoff, albo, asbo, alay);   // imaginary planes
              const uint64_t db1 = umma_smem_desc(b0 + Cfg::kBBytes + boff, blbo, bsbo, blay);
              if constexpr (MODE == DH_FWD) {
                // Y = W X :  re = Wr Xr - Wi Xi ;  im = Wi Xr + Wr Xi
                umma_tf32(d_re, da0, db0, id_pp, acc);
                umma_tf32(d_re, da1, db1, id_np, 1u);
                umma_tf32(d_im, da1, db0, id_pp, acc);
                umma_tf32(d_im, da0, db1, id_pp, 1u);
              } else {
                // conj(A) B :  re = Ar Br + Ai Bi ;  im = Ar Bi - Ai Br      (A = W^T or X^T)
                umma_tf32(d_re, da0, db0, id_pp, acc);
                umma_tf32(d_re, da1, db1, id_pp, 1u);
                umma_tf32(d_im, da0, db1, id_pp, acc);
                umma_tf32(d_im, da1, db0, id_np, 1u);
              }
            }
          }
          umma_commit(&empty_bar[stage]);          // smem slot free once these MMAs retire
          if (++stage == Cfg::kStages) { stage = 0; phase ^= 1; }
        }
        umma_commit(&tmem_full[buf]);              // accumulator complete -> epilogue
        ++acc_it;
      }
    }
  } else if (warp >= 4) {
    // =========================================================== epilogue
    const int q = warp - 4;                        // TMEM lane quarter of this warp (== warp % 4)
    uint32_t acc_it = 0;
    for (uint32_t sn = 0;; ++sn) {
      const int slot = sn & (kSchedDepth - 1);
      mbar_wait(&sched_full[slot], (sn / kSchedDepth) & 1);
      const int t = sched_tile[slot];
      __syncwarp();
      if (lane == 0) mbar_arrive(&sched_empty[slot]);
      if (t < 0) break;
      const GemmTile tile = args.tiles[t];
      const int row = tile.row0 + q * 32 + lane;
      bool row_ok = row < args.row_limit;
      int orow = row;
      if constexpr (MODE == LEG_ANA) {
        if (args.grp_rows) orow = (row / args.grp_rows) * args.grp_stride + args.grp_off + row % args.grp_rows;
      }
      float* obase = args.out + (long long)tile.g * args.out_g_stride + orow;
      if (tile.nkk == 0) {
        for (int c = tile.col0; c < tile.col_limit; ++c) {
          if (row_ok) {
            obase[(long long)c * args.out_col_stride] = 0.f;
            if constexpr (Cfg::kComplex) obase[(long long)c * args.out_col_stride + args.out_im_offset] = 0.f;
          }
        }
        continue;
      }
      const uint32_t buf = acc_it & 1, acc_phase = (acc_it >> 1) & 1;
      mbar_wait(&tmem_full[buf], acc_phase);
      tc_fence_after();
      const uint32_t tbase = tmem_base + buf * Cfg::kAccCols + ((uint32_t)(q * 32) << 16);
      // Stores: lane = output row (contiguous in memory), one STG per column => 128 B per warp store.
      // The column stride fits 32 bits (asserted on the host), so every address is one IMAD.WIDE off a
      // per-chunk base pointer; full chunks take the unpredicated path.
      const int cstride = (int)args.out_col_stride;
      const int nvalid = tile.col_limit - tile.col0;
#pragma unroll 1
      for (int p = 0; p < Cfg::kPlanes; ++p) {
        float* op = obase + (p ? args.out_im_offset : 0) + (long long)tile.col0 * args.out_col_stride;
        uint32_t v[32];
        tmem_ld_32x32(tbase + p * 128, v);
#pragma unroll 1
        for (int c = 0; c < tile.n_mma; c += 32) {
          tmem_ld_wait();
          float w[32];
          if (args.round_out) {
#pragma unroll
            for (int j = 0; j < 32; ++j) w[j] = tf32_rn(__uint_as_float(v[j]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) w[j] = __uint_as_float(v[j]);
          }
          if (c + 32 < tile.n_mma) tmem_ld_32x32(tbase + p * 128 + c + 32, v);   // prefetch the next chunk
          if constexpr (MODE == LEG_ANA) {
            if (args.accumulate && row_ok) {
              const int nv = nvalid - c;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nv) w[j] += op[j * cstride];
              if (args.round_out) {
#pragma unroll
                for (int j = 0; j < 32; ++j) w[j] = tf32_rn(w[j]);
              }
            }
          }
          if (row_ok) {
            const int nv = nvalid - c;
            if (nv >= 32) {
#pragma unroll
              for (int j = 0; j < 32; ++j) st_global_hint(op + j * cstride, w[j], args.pol_out);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nv) st_global_hint(op + j * cstride, w[j], args.pol_out);
            }
          }
          op += 32 * (long long)cstride;
        }
      }
      tc_fence_before();
      mbar_arrive(&tmem_empty[buf]);
      ++acc_it;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
  if (threadIdx.x == 0) {                        // the last CTA to finish re-arms the scheduler for the next launch
    __threadfence();
    if (atomicAdd(args.sched + 1, 1) == (int)gridDim.x - 1) {
      args.sched[0] = 0;
      args.sched[1] = 0;
    }
  }
}

}  // namespace sfno
