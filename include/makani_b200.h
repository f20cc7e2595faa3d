/*
 * makani_b200 - C ABI of the B200-native SFNO spectral hot path.
 *
 * Drop-in boundary for the path  RealSHT -> SpectralConv (dhconv / diagonal) -> InverseRealSHT  of
 * choutilin/makani.  The reference has no FFI for this path (it is pure Python over cuFFT / cuBLAS);
 * each entry point below names the reference call it replaces (paths relative to the reference root).
 * The Python host modules in makani_b200/ bind these with ctypes (makani_b200/_ext.py); see
 * INTEGRATION.md for the stub a maintainer of the reference would add.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; sfno_last_error() gives the message
 *     (thread local);
 *   - all data pointers are DEVICE pointers unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; calls only enqueue work (no host sync, no allocation)
 *     so they can be captured in a CUDA graph (reference captures fwd+bwd, makani/utils/trainer.py:84-152);
 *   - inputs are borrowed and never modified; outputs are caller-allocated.
 *
 * Internal device layouts ("spectral-internal", fp32)
 *   RS   = 2 * B * Cp,  Cp = round_up(C, 4);  row index r = (b*2 + reim)*Cp + c
 *   Kp   = round_up(nlat, 4)
 *   XM   : [mmax][RS][Kp]      longitude-Fourier coefficients, latitude contiguous   (DFT <-> Legendre)
 *   SP   : [lmax][mmax][RS]    spherical-harmonic coefficients, channel contiguous    (Legendre <-> conv)
 *          only entries with m < min(mmax, round_up(l+1,16)) are defined; those with m > l are exact zeros
 *   WP   : [lmax][2][Coutp][Cinp]   dhconv weights, planar re/im, Cin contiguous
 */
#ifndef MAKANI_B200_H_
#define MAKANI_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SFNO_ABI_VERSION 3

/* element types of spatial tensors crossing the boundary */
enum { SFNO_F32 = 0, SFNO_BF16 = 1, SFNO_F16 = 2 };

typedef struct sfno_plan sfno_plan;

int sfno_abi_version(void);
const char* sfno_last_error(void);
/* number of kernels this library has launched in this process (bench.py "gpu_launches") */
long long sfno_launch_count(void);

/* ---- plan: one Legendre table [mmax][lmax][nlat] on the device --------------------------------
 * Replaces the `weights` / `pct` buffers that torch_harmonics.RealSHT / InverseRealSHT register
 * (constructed at makani/models/networks/sfnonet.py:536-539).  table_host is fp64, rounded once to
 * fp32 (the reference calls .float() on the module).  m_offset is the first global order m held by
 * this table (azimuth-sharded DistributedRealSHT keeps only its m-slab); 0 otherwise.            */
/* m_offset of a plan whose table has NO l >= m structure (every (l, m) entry is used): the latitude DFT matrices of
 * the FFT2 transforms (makani/models/common/layers.py:217-280).  The Legendre kernels then skip nothing; callers pass
 * l_offset = SFNO_DENSE_L_OFFSET to the SP <-> complex64 conversions so that they skip nothing either.          */
#define SFNO_M_OFFSET_DENSE (-1)
#define SFNO_DENSE_L_OFFSET (1 << 20)
int sfno_plan_create(int nlat, int nlon, int lmax, int mmax, int m_offset,
                     const double* table_host, sfno_plan** out);
int sfno_plan_destroy(sfno_plan* plan);

/* sizes (in floats) of the internal buffers for a given batch*channels */
long long sfno_xm_elems(int nlat, int mmax, int B, int C);
long long sfno_sp_elems(int lmax, int mmax, int B, int C);
long long sfno_wp_elems(int lmax, int Cin, int Cout);

/* ---- K1 longitude DFT:   x[B*C][nlat][nlon] -> XM     (torch.fft.rfft at RealSHT.forward; call site
 * makani/models/common/spectral_convolution.py:131).  mscale_dev[mmax] multiplies bin m
 * (forward SHT: 2*pi/nlon for all m; adjoint of the inverse DFT: 1,2,2,..(,1 at Nyquist)).       */
int sfno_dft_fwd(const void* x, int x_dtype, int B, int C, int nlat, int nlon, int mmax,
                 const float* mscale_dev, float* xm, int round_tf32, void* stream);

/* ---- K5 inverse longitude DFT:   XM -> y[B*C][nlat][nlon]   (torch.fft.irfft(n=nlon, norm="forward") at
 * InverseRealSHT.forward; call site spectral_convolution.py:133,141).  mscale_dev[mmax] pre-multiplies
 * bin m (inverse SHT: 1; adjoint of the forward DFT: pi/nlon * {2,1,1,...}).  m_valid: bins >= m_valid
 * are taken as zero.                                                                              */
int sfno_dft_inv(const float* xm, int B, int C, int nlat, int nlon, int mmax, int m_valid,
                 const float* mscale_dev, void* y, int y_dtype, void* stream);

/* ---- K2 / K4 Legendre contractions on tcgen05 tensor cores (TF32 in, fp32 accumulate) ---------
 * analysis : SP[l][m][r]  = sum_k XM[m][r][k] * table[m][l][k]   (einsum '...km,mlk->...lm')
 * synthesis: XM[m][r][k]  = sum_l SP[l][m][r] * table[m][l][k]   (einsum '...lm,mlk->...km')
 * The same plan serves both directions (a forward plan's synthesis is the adjoint needed by backward). */
/* round_tf32 = 1 stores round-to-nearest TF32 values (set it when the output feeds another TF32 contraction:
 * the tensor core otherwise truncates, a one-sided ~1e-3 bias); 0 keeps full fp32 (output leaves the path). */
int sfno_legendre_analysis(const sfno_plan* plan, const float* xm, int B, int C, float* sp, int round_tf32,
                           void* stream);
int sfno_legendre_synthesis(const sfno_plan* plan, const float* sp, int B, int C, float* xm, void* stream);

/* ---- channel-window variants of K1 / K2 / K4 / K5 -----------------------------------------------
 * The chain DFT -> Legendre (and Legendre -> inverse DFT) is independent per channel, so it can be run window by window
 * over the channels, with the Fourier coefficients of ONE window going through a small XM buffer that is re-used for
 * every window: the buffer (tens of MB) then lives in the 126 MB L2 instead of travelling to HBM and back (XM of the
 * whole field is 178 MB at 240x480 and 536 MB at 721x1440, C = 384).  This is how SpectralConv fuses across the
 * RFFT -> Legendre and Legendre -> inverse-RFFT boundaries of makani/models/common/spectral_convolution.py:131-141.
 *   x / y   : the WHOLE field [B][Ctot][nlat][nlon];   sp : the WHOLE SP tensor (Ctot channels)
 *   xm_win  : compact XM of the window, [mmax][2B][round_up(C,4)][Kp]  (sfno_xm_elems(nlat, mmax, B, C) floats)
 *   window  : channels [c0, c0 + C);  a true window (C < Ctot) needs C % 32 == 0 and c0 % 4 == 0
 * accumulate = 1 (analysis): SP += result (sum of the gradients of two consumers of one SP tensor).
 * The kernels give the window and the Legendre table L2 priority (evict_last) and let the field rows stream
 * (evict_first); SFNO_WINDOW_HINTS=0 in the environment disables the hints.                                  */
int sfno_dft_fwd_win(const void* x, int x_dtype, int B, int Ctot, int c0, int C, int nlat, int nlon, int mmax,
                     const float* mscale_dev, float* xm_win, int round_tf32, void* stream);
int sfno_dft_inv_win(const float* xm_win, int B, int Ctot, int c0, int C, int nlat, int nlon, int mmax, int m_valid,
                     const float* mscale_dev, void* y, int y_dtype, void* stream);
int sfno_legendre_analysis_win(const sfno_plan* plan, const float* xm_win, int B, int Ctot, int c0, int C, float* sp,
                               int round_tf32, int accumulate, void* stream);
int sfno_legendre_synthesis_win(const sfno_plan* plan, const float* sp, int B, int Ctot, int c0, int C, float* xm_win,
                                void* stream);

/* nlon = 1440 with more than 241 longitude modes - the transform of the spectral loss, which keeps all 721
 * (makani/utils/losses.py:291; torch-harmonics RealSHT: rfft over the longitudes, all mmax columns kept).  The tiled
 * kernels handle 241 bins per launch; three launches with their 1440-point twiddles multiplied by a cube root of unity
 * cover the spectrum:   band 0: bins k,   band 2: bins 480 + k,   band 1: conj of bins 480 - k   (k = 0 .. 240).
 *   xm_rows : first bin of the band inside an XM tensor (band 0: xm, band 2: xm + 480 rows) or, for band 1, a scratch
 *             of 241 rows that sfno_xm_mirror_conj turns around into rows 241 .. 479 (forward) / fills from them
 *             (inverse).  One row = 2B * round_up(C,4) * Kp floats.
 *   mscale_dev : mcount per-bin factors of THIS band.  Inverse: the kernel counts its bin 0 once and doubles its bin
 *             240, so band 2 wants 2 * s[480] first and s[720] / 2 last, and band 1 zeros at both ends; the three
 *             results are added by the caller (makani_b200/ops.py: _dft_inv_banded).
 * sfno_xm_mirror_conj: dst row i = conj(src row nrows-1-i) * scale[i] (scale may be NULL); planes alternate re / im. */
int sfno_dft_fwd_band(const void* x, int x_dtype, int B, int Ctot, int c0, int C, int nlat, int band, int mcount,
                      const float* mscale_dev, float* xm_rows, int round_tf32, void* stream);
int sfno_dft_inv_band(const float* xm_rows, int B, int Ctot, int c0, int C, int nlat, int band, int mcount,
                      const float* mscale_dev, void* y, int y_dtype, void* stream);
int sfno_xm_mirror_conj(const float* src, float* dst, int nrows, int nplanes, long long plane_floats,
                        const float* scale_dev, int round_tf32, void* stream);

/* ---- K3 dhconv channel contraction (makani/models/common/contractions.py:131-136,
 * einsum "bixy,iox->boxy" on complex64) as planar-complex TF32 tcgen05 GEMMs, batched over l.   */
int sfno_dhconv_pack_weight(const void* w_c64 /*[Cin][Cout][L] complex64*/, int Cin, int Cout, int L,
                            float* wp, void* stream);
/* round_tf32 = 0: the packed copy keeps the fp32 values (a packed MASTER copy that an optimizer updates in place,
 * SpectralConv.enable_packed_master(); the tensor cores then truncate instead of rounding to nearest);
 * round_tf32 = 1 is sfno_dhconv_pack_weight.                                                                    */
int sfno_dhconv_pack_weight_ex(const void* w_c64, int Cin, int Cout, int L, float* wp, int round_tf32, void* stream);
int sfno_dhconv_unpack_wgrad(const float* gwp, int Cin, int Cout, int L,
                             void* gw_c64 /*[Cin][Cout][L] complex64*/, void* stream);
/* l_offset / m_offset: global degree / order of local index 0 (non-zero only for the h/w-sharded
 * DistributedRealSHT layout); they place the triangular m <= l structure that the kernels skip. */
int sfno_dhconv_fwd(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout,
                    int lmax, int mmax, int l_offset, int m_offset, int round_tf32, void* stream);
/* dX[b,i,l,m] = sum_o dY[b,o,l,m] * conj(w[i,o,l]) */
int sfno_dhconv_bwd_x(const float* gsp_out, const float* wp, float* gsp_in, int B, int Cin, int Cout,
                      int lmax, int mmax, int l_offset, int m_offset, int round_tf32, void* stream);
/* dW[i,o,l] = sum_{b,m} conj(x[b,i,l,m]) * dY[b,o,l,m]   (packed layout WP) */
int sfno_dhconv_bwd_w(const float* sp_in, const float* gsp_out, float* gwp, int B, int Cin, int Cout,
                      int lmax, int mmax, int l_offset, int m_offset, void* stream);
/* The same contraction as sfno_dhconv_fwd for callers that may not allocate (code under CUDA graph capture, serving
 * loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * It calls no cudaMalloc / cudaFree and no allocating library routine; the two things sfno_dhconv_fwd
 * keeps in allocated device memory (the tile list of the shape, the tile scheduler's counter pair) live in one
 * file-scope device array of 256 + 32 * SFNO_STATIC_MAX_TILES bytes (0.5 MB) that every GPU loading the library
 * holds until exit.  Limits: lmax <= SFNO_STATIC_MAX_L and
 *     sum_l ceil(B * min(mmax, round_up(l + l_offset + 1, 16) - m_offset) / 128) * ceil(Cout / 128) <= SFNO_STATIC_MAX_TILES
 * (the named config needs 1 056 tiles at B = 1, 5 760 at B = 8, 11 520 at B = 16); beyond them the call returns 1 with a message.
 * All calls share the array: it is rewritten at the start of every call (in stream order), so calls on one stream
 * are safe back to back, calls that overlap on the GPU (different streams) race.  Results are bit-identical to
 * sfno_dhconv_fwd's.                                                                                              */
#define SFNO_STATIC_MAX_TILES 16384
#define SFNO_STATIC_MAX_L 1024
int sfno_dhconv_fwd_static(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout,
                           int lmax, int mmax, int l_offset, int m_offset, int round_tf32, void* stream);

/* ---- K3' diagonal contraction (contractions.py:121-127, einsum "bixy,ioxy->boxy"): weights are used once,
 * so this is a bandwidth kernel.  w is complex64 [Cin][Cout][lmax][mmax].
 *   mode 0: y  = x * w            sp_a = SP x,  w_or_spb = w,    out = SP y
 *   mode 1: dx = dy * conj(w)     sp_a = SP dy, w_or_spb = w,    out = SP dx
 *   mode 2: dw = conj(x) * dy     sp_a = SP x,  w_or_spb = SP dy, out = complex64 dw
 * The separable variants (contractions.py:139-154) are not provided: the reference's own einsum strings
 * for them ("bixy,ix->boxy") have an output index absent from the inputs and raise in torch.       */
int sfno_diag_contract(int mode, const float* sp_a, const void* w_or_spb, void* out, int B, int Cin, int Cout,
                       int lmax, int mmax, int l_offset, int m_offset, void* stream);

/* ---- layout conversion at the module boundary ------------------------------------------------
 * SP <-> complex64 [B][C][lmax][mmax] (what RealSHT returns / InverseRealSHT accepts).           */
int sfno_sp_to_c64(const float* sp, int B, int C, int lmax, int mmax, int l_offset, int m_offset, void* out_c64,
                   void* stream);
/* entries with (m + m_offset) > (l + l_offset) are not read: they become the exact zeros the SP layout requires.
 * round_tf32 = 1: values are stored round-to-nearest TF32 (the consumer is a tensor-core contraction: Legendre
 * synthesis, dhconv); 0: exact fp32 (the consumer is the fp32 diagonal contraction).                          */
int sfno_c64_to_sp(const void* in_c64, int B, int C, int lmax, int mmax, int l_offset, int m_offset, int round_tf32,
                   float* sp, void* stream);
/* XM <-> complex64 [B][C][nlat][mmax]  (needed by the distributed SHT between its all-to-alls)   */
int sfno_xm_to_c64(const float* xm, int B, int C, int nlat, int mmax, void* out_c64, void* stream);
int sfno_c64_to_xm(const void* in_c64, int B, int C, int nlat, int mmax, float* xm, void* stream);

/* y[b][c][h][w] (+)= bias   (spectral_convolution.py:143-144); bias_kind 0: [C], 1: [C][H][W]    */
int sfno_add_bias(void* y, int y_dtype, const float* bias, int bias_kind, int B, int C, int H, int W,
                  void* stream);

/* Per-degree energy of an SP tensor and its adjoint - the spectral loss (makani/utils/losses.py:299-303, 327-337:
 * coeffs.re^2 + coeffs.im^2, then column m = 0 plus twice the sum of the other columns):
 *   out[b][c][l] = sum_{m + m_offset <= l + l_offset} w_m |SP[l][m][b][c]|^2,   w_m = 1 at m + m_offset == 0, else 2
 *   gsp[l][m][b][re|im][c] = 2 w_m g[b][c][l] sp[l][m][b][re|im][c]   (exact zeros for m above the diagonal and in
 *                            pad channels; round_tf32: the consumer is the TF32 Legendre synthesis of the backward)
 * out / g: fp32 [B][C][lmax].  The complex64 coefficient tensor of the reference API is never formed.          */
int sfno_sp_degree_power(const float* sp, int B, int C, int lmax, int mmax, int l_offset, int m_offset, float* out,
                         void* stream);
int sfno_sp_degree_power_bwd(const float* sp, const float* g, int B, int C, int lmax, int mmax, int l_offset,
                             int m_offset, int round_tf32, float* gsp, void* stream);

/* dst[i] += src[i], i < n (n % 4 == 0, 16-byte aligned): the sum of two gradients in an internal layout (SP_l of the
 * h-sharded SpectralConv backward: filter branch + re-projected residual, spectral_convolution.py:132-134).   */
int sfno_add_inplace(float* dst, const float* src, long long n, void* stream);

/* Constant bias [C] folded into the m = 0 Fourier coefficient of an XM tensor (or XM window: pass bias + c0) before
 * the inverse DFT:  XM[0][b][re][c][k] += bias[c].  Equals `x = x + bias` in fp32 before the cast to the output dtype
 * (spectral_convolution.py:143-146) at the cost of B*C*nlat elements.                                        */
int sfno_xm_add_dc(float* xm, int B, int C, int nlat, const float* bias, void* stream);

/* ---- composite calls: one call per reference operator ------------------------------------------
 * For callers that do not want to sequence the kernels themselves (the C ABI list of SURVEY.md 8(b)).
 * They enqueue exactly the kernels above on `stream`, with the intermediates (XM, SP) in `ws`, a device
 * scratch buffer of at least sfno_workspace_bytes(...) bytes that the caller owns (no allocation inside,
 * so the calls can be captured in a CUDA graph).                                                       */
enum { SFNO_OP_SHT_FWD = 0, SFNO_OP_SHT_INV = 1, SFNO_OP_BLOCK_FWD = 2 };
/* plan_b is only read for SFNO_OP_BLOCK_FWD (the inverse transform's plan); C = max(Cin, Cout) there. */
size_t sfno_workspace_bytes(const sfno_plan* plan, const sfno_plan* plan_b, int op, int B, int C);
/* RealSHT.forward (spectral_convolution.py:131, sfnonet.py:596):
 *   x [B][C][nlat][nlon] (x_dtype) -> coeffs complex64 [B][C][lmax][mmax]                              */
int sfno_sht_fwd(const sfno_plan* plan, const void* x, int x_dtype, int B, int C, void* coeffs_c64,
                 void* ws, size_t ws_bytes, void* stream);
/* InverseRealSHT.forward (spectral_convolution.py:133,141, sfnonet.py:598,616):
 *   coeffs complex64 [B][C][lmax][mmax] -> y [B][C][nlat][nlon] (y_dtype)                              */
int sfno_sht_inv(const sfno_plan* plan, const void* coeffs_c64, int B, int C, void* y, int y_dtype,
                 void* ws, size_t ws_bytes, void* stream);
/* SpectralConv.forward, dhconv, same-grid case without bias (spectral_convolution.py:124-148):
 *   y = isht(dhconv(sht(x), w)),  x [B][Cin][nlat_f][nlon_f] -> y [B][Cout][nlat_i][nlon_i] (same dtype);
 * wp is the packed weight from sfno_dhconv_pack_weight.  The reference's second return value
 * (`residual`) is x itself unless the block changes resolution; callers that need the re-projected
 * residual call sfno_sht_fwd + sfno_sht_inv.  Both plans must agree in lmax and mmax.                  */
int sfno_block_fwd(const sfno_plan* plan_fwd, const sfno_plan* plan_inv, const void* x, int dtype, int B,
                   int Cin, int Cout, const float* wp, void* y, void* ws, size_t ws_bytes, void* stream);

/* ---- exchange step of the h-sharded (polar) distributed SHT over NVLink peer memory -------------
 * Replaces the split / contiguous / all_to_all / cat of thd.distributed_transpose_polar
 * (makani/mpu/layers.py:73-100 call form; torch-harmonics distributed/primitives.py) for the internal
 * layouts: one kernel reads the local XM / SP tensor and stores every row straight into the destination
 * layout of its owner - the local buffer for the own part, peer-mapped buffers (CUDA IPC) for the rest -
 * so the transposition costs one pass over the data and needs no packing on either side.
 *
 * sfno_ipc_alloc    : cudaMalloc + zero a buffer and export its IPC handle (64 bytes) for the other ranks
 * sfno_ipc_open     : map a buffer exported by another process of this node (enables peer access)
 * sfno_ipc_close / sfno_ipc_free : undo the above
 * sfno_rows_scatter : for each descriptor, copy rows (i0, i1, i2), i_k < n_k, of `len` floats
 *                       dst[dst_off + i0*ds0 + i1*ds1 + i2*ds2 + e] = src[src_off + i0*ss0 + i1*ss1 + i2*ss2 + e]
 *                     tri_l0 >= 0 marks an SP copy: i0 is a degree (l = tri_l0 + i0), i1 an order m, and the rows
 *                     with m >= round_up(l + 1, 16) - which the SP layout leaves undefined - are skipped.      */
typedef struct {
  void* dst;
  long long src_off, dst_off;
  int n0, n1, n2, len;
  long long ss0, ss1, ss2, ds0, ds1, ds2;
  int tri_l0;
} sfno_copy_desc;
int sfno_ipc_alloc(size_t bytes, void** ptr, unsigned char handle_out[64]);
int sfno_ipc_open(const unsigned char handle[64], void** ptr);
int sfno_ipc_close(void* ptr);
int sfno_ipc_free(void* ptr);
int sfno_rows_scatter(const float* src, const sfno_copy_desc* descs_host, int ndesc, void* stream);
/* sfno_rows_scatter's copy on the copy engines (descriptors with tri_l0 < 0 whose strides form a pitched 3-D box: ss0 ==
 * n1*ss1, ss1 % ss2 == 0, same for ds): one strided 3-D async copy per descriptor, no SMs.                      */
int sfno_rows_copy_ce(const float* src, const sfno_copy_desc* descs_host, int ndesc, void* stream);
/* Stream-ordered barrier between the `size` (<= 8) ranks of one node through peer memory: rank r stores the barrier
 * number into slot r of every rank's flag array (flags_host[p] = pointer to rank p's array of >= 64 ints, zero at
 * allocation, mapped with sfno_ipc_open) and waits until all slots of its own array have reached it.
 * Everything the ranks enqueued before the barrier (e.g. sfno_rows_scatter into peer buffers) is visible to what
 * they enqueue after it.
 * epoch <= 0 (recommended): the barrier number is kept in device memory (slot 32 of the rank's own array, incremented by
 * the kernel), so the call can be captured in a CUDA graph and replayed (makani/utils/trainer.py:84-152); every rank
 * must then issue the same sequence of barriers on these arrays.  epoch > 0: the caller numbers the barriers itself
 * (+1 per barrier on the same arrays) - NOT replayable from a captured graph.                                    */
int sfno_peer_barrier(int* const* flags_host, int rank, int size, int epoch, void* stream);

/* ---- data-parallel exchange of the weight gradients over NVLink peer memory --------------------------------
 * Replaces the NCCL all-reduce that DDP issues for the spectral weights (makani/mpu/mappings.py:104-172; 8 x 283 MB per
 * step) by copy-engine transfers between CUDA-IPC mapped buffers plus one small kernel, so that the exchange takes no
 * SMs away from the backward kernels it overlaps with (makani_b200/distributed/grad_exchange.py):
 *   reduce-scatter : own slice = (own slice + the same slice of every peer) * scale     (sfno_slices_average: the peers'
 *                    slices are read straight from their buffers over NVLink, or pulled first with sfno_peer_copy)
 *   all-gather     : folded into the layout change packed -> complex64 that backward needs anyway: it reads every
 *                    averaged slice from its owner                                    (sfno_dhconv_unpack_wgrad_gather)
 * ordered by sfno_peer_barrier on the communication stream.
 * sfno_peer_copy     : cudaMemcpyAsync(dst, src, bytes) where either side may be a buffer mapped with sfno_ipc_open
 * sfno_slices_average: dst[i] = (dst[i] + sum_k srcs_host[k][i]) * scale,  k < nsrc <= 7, n % 4 == 0, 16-byte aligned */
int sfno_peer_copy(void* dst, const void* src, size_t bytes, void* stream);
/* sfno_dhconv_unpack_wgrad reading slice k (flat WP offsets [k*slice_floats, (k+1)*slice_floats)) from
 * gwp_slices_host[k] - BASE pointers of WP-sized buffers, typically one per rank, mapped with sfno_ipc_open: the
 * all-gather of the averaged slices happens inside the layout change, nothing is gathered into a local packed copy. */
int sfno_dhconv_unpack_wgrad_gather(const float* const* gwp_slices_host, int nslices, long long slice_floats, int Cin,
                                    int Cout, int L, void* gw_c64, void* stream);
int sfno_slices_average(float* dst, const float* const* srcs_host, int nsrc, long long n, float scale, void* stream);

/* ---- longitude DFT fused with the latitude <-> channel exchange of the h-sharded transforms --------
 * xm_ch_host[p] (p < nranks <= 8) is rank p's channel-sharded XM buffer  [mmax][2B][round_up(chan_shapes[p],4)][Kg]
 * in peer memory (sfno_ipc_open); this rank's nlat_loc latitudes occupy columns [lat_off, lat_off + lat_span) of
 * every one of them (lat_off multiple of 4, lat_span >= nlat_loc).
 *   sfno_dft_fwd_peer : x [B][C][nlat_loc][nlon] -> every tile of spectrum is stored by the kernel's tensor stores
 *                       straight into the buffer of the rank that owns its channel (no local XM, no separate exchange)
 *   sfno_dft_inv_peer : the kernel's tensor loads fetch each tile from the owner of its channel -> y [B][C][nlat_loc][nlon]
 * Only for the nlon = 480 / 1440 kernels (sfno_dft_peer_supported); callers order the ranks with sfno_peer_barrier
 * (after the forward kernel / before the inverse kernel).                                                   */
int sfno_dft_peer_supported(int nlon, int mmax);
int sfno_dft_fwd_peer(const void* x, int x_dtype, int B, int C, int nlat_loc, int nlon, int mmax,
                      const float* mscale_dev, void* const* xm_ch_host, const int* chan_shapes_host, int nranks, int Kg,
                      int lat_off, int lat_span, int round_tf32, void* stream);
int sfno_dft_inv_peer(void* const* xm_ch_host, const int* chan_shapes_host, int nranks, int Kg, int lat_off,
                      int lat_span, int B, int C, int nlat_loc, int nlon, int mmax, int m_valid,
                      const float* mscale_dev, void* y, int y_dtype, void* stream);

/* ---- fused instance norm (+ GELU, + skip) around the spectral filter ----------
 * Replaces  act_layer0(norm0(x))  and  norm1(x) + outer_skip(residual)  of FourierNeuralOperatorBlock.forward
 * (makani/models/networks/sfnonet.py:239-267; nn.InstanceNorm2d(eps=1e-6, affine=True) of sfnonet.py:375-380, nn.GELU).
 * x, skip, g, out, dx: [planes = B*C][hw] contiguous, dtype SFNO_F32 / SFNO_BF16 / SFNO_F16; arithmetic in fp32.
 *   sfno_instnorm_stats     : per-plane statistics. partial_ws: planes * SFNO_INSTNORM_SEGMENTS * 2 floats of scratch;
 *                             mean_var [planes][2] = (mean, biased variance) and/or stats [planes][2] = (mean,
 *                             1/sqrt(var + eps)) - either may be NULL.  A spatially sharded instance merges the ranks'
 *                             (mean, var, count) itself and hands the merged `stats` to the calls below.
 *   sfno_instnorm_apply     : out = act( gamma[c] * (x - mean) * rstd + beta[c] (+ skip) ), c = plane % C.
 *                             gamma / beta may be NULL (affine=False); skip is only accepted with SFNO_ACT_NONE.
 *   sfno_instnorm_bwd_stats : partial [planes][SFNO_INSTNORM_SEGMENTS][2] = partial sums of (dz, dz * xhat) with
 *                             dz = g * act'(z), xhat = (x - mean) * rstd.  The caller sums the segments (and, when the
 *                             instance is sharded, all-reduces) into bsum [planes][2]; d gamma[c] = sum_b bsum[.,1],
 *                             d beta[c] = sum_b bsum[.,0].
 *   sfno_instnorm_bwd_apply : dx = gamma[c] * rstd * ( dz - (bsum0 + xhat * bsum1) / hw_total ); hw_total = elements of
 *                             the whole instance (= hw unless sharded).  The gradient of `skip` is g itself.       */
#define SFNO_INSTNORM_SEGMENTS 8
#define SFNO_ACT_NONE 0
#define SFNO_ACT_GELU 1
int sfno_instnorm_stats(const void* x, int dtype, int planes, long long hw, float eps, float* partial_ws,
                        float* mean_var, float* stats, void* stream);
int sfno_instnorm_apply(const void* x, const void* skip, int dtype, int planes, int C, long long hw,
                        const float* stats, const float* gamma, const float* beta, int act, void* out, void* stream);
int sfno_instnorm_bwd_stats(const void* g, const void* x, int dtype, int planes, int C, long long hw,
                            const float* stats, const float* gamma, const float* beta, int act, float* partial,
                            void* stream);
int sfno_instnorm_bwd_apply(const void* g, const void* x, int dtype, int planes, int C, long long hw,
                            long long hw_total, const float* stats, const float* gamma, const float* beta,
                            const float* bsum, int act, void* dx, void* stream);


/* ---- bias + activation around the 1x1 convolutions (the MLP of the block and the
 * encoder / decoder, makani/models/common/layers.py:86-214 - Conv2d(1x1, bias) -> GELU -> Conv2d(1x1)).  The
 * convolutions themselves are plain GEMMs on the NCHW tensor viewed as [C][H*W] (library GEMMs, no layout change);
 * these two calls are the elementwise passes between them, on [planes = B*C] planes of hw contiguous elements:
 *   sfno_bias_act     : out = act(x + bias[c]), c = plane % C; bias may be NULL; out may alias x.
 *   sfno_bias_act_bwd : dx = g * act'(x + bias[c]) and partial[plane][SFNO_INSTNORM_SEGMENTS] = partial sums of dx
 *                       (d bias[c] = sum over the batch and the segments).  act = NONE: x is not read and dx may be
 *                       NULL (only the sums: the bias gradient of a convolution without activation).            */
int sfno_bias_act(const void* x, int dtype, int planes, int C, long long hw, const float* bias, int act, void* out,
                  void* stream);
int sfno_bias_act_bwd(const void* g, const void* x, int dtype, int planes, int C, long long hw, const float* bias,
                      int act, void* dx, float* partial, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAKANI_B200_H_ */
