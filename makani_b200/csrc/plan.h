// Internal declarations shared by the translation units of libmakani_b200.so
#pragma once
#include <cuda_runtime.h>
#include <functional>
#include "../../include/makani_b200.h"

struct sfno_plan {
  int nlat, nlon, lmax, mmax, m_offset;
  int Kp;            // round_up(nlat, 4): latitude stride of the table and of XM
  float* table;      // device, [mmax][lmax][Kp] fp32 (pad entries zero)
  float* mscale;     // device, [2][nlon/2 + 1]: row 0 = 2*pi/nlon (RealSHT), row 1 = 1 (InverseRealSHT)
  int device;
};

namespace sfno {

int sp_mhi(int l_local, int l_offset, int m_offset, int M_local);
int* sched_counters();   // {next tile, finished CTAs} pair for one launch of a dynamically scheduled kernel

bool window_hints();     // L2 eviction hints of the channel-window mode (SFNO_WINDOW_HINTS=0 disables)

// XM holds the channel window [c0, c0 + C) of Ctot channels; SP is the full tensor (Ctot == C, c0 == 0: no window)
int legendre_analysis(const sfno_plan* plan, const float* xm, int B, int C, int Ctot, int c0, float* sp, int round_out,
                      int accumulate, cudaStream_t stream);
int legendre_synthesis(const sfno_plan* plan, const float* sp, int B, int C, int Ctot, int c0, float* xm,
                       cudaStream_t stream);
int dhconv_fwd(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout, int L, int M,
               int l_offset, int m_offset, int round_out, cudaStream_t stream);
int dhconv_bwd_x(const float* gsp_out, const float* wp, float* gsp_in, int B, int Cin, int Cout, int L, int M,
                 int l_offset, int m_offset, int round_out, cudaStream_t stream);
int dhconv_bwd_w(const float* sp_in, const float* gsp_out, float* gwp, int B, int Cin, int Cout, int L, int M,
                 int l_offset, int m_offset, cudaStream_t stream);

// x / y point at channel c0 of a field with Ctot channels per batch entry; C channels are transformed
int dft_fwd(const void* x, int x_dtype, int B, int C, int Ctot, int nlat, int nlon, int mmax, const float* mscale,
            float* xm, int round_out, cudaStream_t stream);
int dft_inv(const float* xm, int B, int C, int Ctot, int nlat, int nlon, int mmax, int m_valid, const float* mscale,
            void* y, int y_dtype, cudaStream_t stream);

// nlon = 1440 beyond 241 modes: one band (0: bins 0..240, 1: conjugates of bins 480-k, 2: bins 480+k) per launch
int dft_fwd_band(const void* x, int x_dtype, int B, int C, int Ctot, int nlat, int band, int mcount, const float* mscale,
                 float* xm_rows, int round_out, cudaStream_t stream);
int dft_inv_band(const float* xm_rows, int B, int C, int Ctot, int nlat, int band, int mcount, const float* mscale, void* y,
                 int y_dtype, cudaStream_t stream);
int xm_mirror_conj(const float* src, float* dst, int nrows, int nplanes, long long plane_floats, const float* scale,
                   int round_out, cudaStream_t stream);

bool dft_peer_supported(int nlon, int mmax);
int dft_fwd_peer(const void* x, int x_dtype, int B, int C, int nlat_loc, int nlon, int mmax, const float* mscale,
                 void* const* xm_ch, const int* chan_shapes, int nranks, int Kg, int lat_off, int lat_span,
                 int round_out, cudaStream_t stream);
int dft_inv_peer(void* const* xm_ch, const int* chan_shapes, int nranks, int Kg, int lat_off, int lat_span, int B,
                 int C, int nlat_loc, int nlon, int mmax, int m_valid, const float* mscale, void* y, int y_dtype,
                 cudaStream_t stream);

}  // namespace sfno
