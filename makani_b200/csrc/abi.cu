// extern "C" surface of libmakani_b200.so (declared in include/makani_b200.h): error plumbing, plans,
// and thin wrappers that forward to the kernels' host launchers.
#include "common.cuh"
#include "plan.h"

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

namespace sfno {

static thread_local std::string t_last_error;

// round-to-nearest (ties away) to TF32 on the host, same result as cvt.rna.tf32.f32 for finite values
static float tf32_rn_host(float x) {
  uint32_t u;
  memcpy(&u, &x, 4);
  if ((u & 0x7f800000u) == 0x7f800000u) return x;
  u = (u + 0x1000u) & 0xffffe000u;
  memcpy(&x, &u, 4);
  return x;
}
std::atomic<long long> g_launch_count{0};

void set_error(const std::string& msg) { t_last_error = msg; }

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* env = getenv("SFNO_PDL");
    v = (env && env[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

}  // namespace sfno

using namespace sfno;

extern "C" {

int sfno_abi_version(void) { return SFNO_ABI_VERSION; }
const char* sfno_last_error(void) { return t_last_error.c_str(); }
long long sfno_launch_count(void) { return g_launch_count.load(); }

long long sfno_xm_elems(int nlat, int mmax, int B, int C) {
  return (long long)mmax * (2LL * B * round_up(C, 4)) * round_up(nlat, 4);
}
long long sfno_sp_elems(int lmax, int mmax, int B, int C) {
  return (long long)lmax * mmax * (2LL * B * round_up(C, 4));
}
long long sfno_wp_elems(int lmax, int Cin, int Cout) {
  return (long long)lmax * 2 * round_up(Cout, 4) * round_up(Cin, 4);
}

int sfno_plan_create(int nlat, int nlon, int lmax, int mmax, int m_offset, const double* table_host,
                     sfno_plan** out) {
  SFNO_CHECK(out != nullptr && table_host != nullptr, "null argument");
  SFNO_CHECK(nlat > 0 && nlon > 0 && lmax > 0 && mmax > 0 && m_offset >= SFNO_M_OFFSET_DENSE, "bad sizes");
  sfno_plan* p = new sfno_plan();
  p->nlat = nlat;
  p->nlon = nlon;
  p->lmax = lmax;
  p->mmax = mmax;
  p->m_offset = m_offset;
  p->Kp = round_up(nlat, 4);
  p->table = nullptr;
  cudaGetDevice(&p->device);
  const size_t n = (size_t)mmax * lmax * p->Kp;
  std::vector<float> h(n, 0.f);
  for (int m = 0; m < mmax; ++m)
    for (int l = 0; l < lmax; ++l) {
      const double* src = table_host + ((size_t)m * lmax + l) * nlat;
      float* dst = h.data() + ((size_t)m * lmax + l) * p->Kp;
      for (int k = 0; k < nlat; ++k) dst[k] = tf32_rn_host((float)src[k]);
    }
  const int nbins = nlon / 2 + 1;
  std::vector<float> sc(2 * (size_t)nbins, 1.f);
  for (int m = 0; m < nbins; ++m) sc[m] = (float)(2.0 * M_PI / nlon);
  p->mscale = nullptr;
  cudaError_t e = cudaMalloc(&p->table, n * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->table, h.data(), n * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&p->mscale, sc.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->mscale, sc.data(), sc.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    if (p->table) cudaFree(p->table);
    if (p->mscale) cudaFree(p->mscale);
    delete p;
    set_error(std::string("sfno_plan_create: ") + cudaGetErrorString(e));
    return 1;
  }
  *out = p;
  return 0;
}

int sfno_plan_destroy(sfno_plan* plan) {
  if (!plan) return 0;
  if (plan->table) cudaFree(plan->table);
  if (plan->mscale) cudaFree(plan->mscale);
  delete plan;
  return 0;
}

int sfno_dft_fwd(const void* x, int x_dtype, int B, int C, int nlat, int nlon, int mmax, const float* mscale_dev,
                 float* xm, int round_tf32, void* stream) {
  SFNO_CHECK(x && xm && mscale_dev, "null argument");
  return dft_fwd(x, x_dtype, B, C, C, nlat, nlon, mmax, mscale_dev, xm, round_tf32, (cudaStream_t)stream);
}
int sfno_dft_inv(const float* xm, int B, int C, int nlat, int nlon, int mmax, int m_valid, const float* mscale_dev,
                 void* y, int y_dtype, void* stream) {
  SFNO_CHECK(xm && y && mscale_dev, "null argument");
  return dft_inv(xm, B, C, C, nlat, nlon, mmax, m_valid, mscale_dev, y, y_dtype, (cudaStream_t)stream);
}
int sfno_legendre_analysis(const sfno_plan* plan, const float* xm, int B, int C, float* sp, int round_tf32,
                           void* stream) {
  SFNO_CHECK(plan && xm && sp, "null argument");
  return legendre_analysis(plan, xm, B, C, C, 0, sp, round_tf32, 0, (cudaStream_t)stream);
}
int sfno_legendre_synthesis(const sfno_plan* plan, const float* sp, int B, int C, float* xm, void* stream) {
  SFNO_CHECK(plan && xm && sp, "null argument");
  return legendre_synthesis(plan, sp, B, C, C, 0, xm, (cudaStream_t)stream);
}

// ---- channel-window variants (see the header): XM is a compact tensor over the window, the field / SP are whole
static size_t dtype_size(int dt) { return dt == SFNO_F32 ? 4 : 2; }
int sfno_dft_fwd_win(const void* x, int x_dtype, int B, int Ctot, int c0, int C, int nlat, int nlon, int mmax,
                     const float* mscale_dev, float* xm_win, int round_tf32, void* stream) {
  SFNO_CHECK(x && xm_win && mscale_dev, "null argument");
  SFNO_CHECK(c0 >= 0 && C > 0 && c0 + C <= Ctot, "sfno_dft_fwd_win: bad channel window");
  const char* xw = static_cast<const char*>(x) + (size_t)c0 * nlat * nlon * dtype_size(x_dtype);
  return dft_fwd(xw, x_dtype, B, C, Ctot, nlat, nlon, mmax, mscale_dev, xm_win, round_tf32, (cudaStream_t)stream);
}
int sfno_dft_inv_win(const float* xm_win, int B, int Ctot, int c0, int C, int nlat, int nlon, int mmax, int m_valid,
                     const float* mscale_dev, void* y, int y_dtype, void* stream) {
  SFNO_CHECK(xm_win && y && mscale_dev, "null argument");
  SFNO_CHECK(c0 >= 0 && C > 0 && c0 + C <= Ctot, "sfno_dft_inv_win: bad channel window");
  char* yw = static_cast<char*>(y) + (size_t)c0 * nlat * nlon * dtype_size(y_dtype);
  return dft_inv(xm_win, B, C, Ctot, nlat, nlon, mmax, m_valid, mscale_dev, yw, y_dtype, (cudaStream_t)stream);
}
int sfno_dft_fwd_band(const void* x, int x_dtype, int B, int Ctot, int c0, int C, int nlat, int band, int mcount,
                      const float* mscale_dev, float* xm_rows, int round_tf32, void* stream) {
  SFNO_CHECK(x && xm_rows && mscale_dev, "null argument");
  SFNO_CHECK(c0 >= 0 && C > 0 && c0 + C <= Ctot, "sfno_dft_fwd_band: bad channel window");
  const char* xw = static_cast<const char*>(x) + (size_t)c0 * nlat * 1440 * dtype_size(x_dtype);
  return dft_fwd_band(xw, x_dtype, B, C, Ctot, nlat, band, mcount, mscale_dev, xm_rows, round_tf32, (cudaStream_t)stream);
}
int sfno_dft_inv_band(const float* xm_rows, int B, int Ctot, int c0, int C, int nlat, int band, int mcount,
                      const float* mscale_dev, void* y, int y_dtype, void* stream) {
  SFNO_CHECK(xm_rows && y && mscale_dev, "null argument");
  SFNO_CHECK(c0 >= 0 && C > 0 && c0 + C <= Ctot, "sfno_dft_inv_band: bad channel window");
  char* yw = static_cast<char*>(y) + (size_t)c0 * nlat * 1440 * dtype_size(y_dtype);
  return dft_inv_band(xm_rows, B, C, Ctot, nlat, band, mcount, mscale_dev, yw, y_dtype, (cudaStream_t)stream);
}
int sfno_xm_mirror_conj(const float* src, float* dst, int nrows, int nplanes, long long plane_floats,
                        const float* scale_dev, int round_tf32, void* stream) {
  return xm_mirror_conj(src, dst, nrows, nplanes, plane_floats, scale_dev, round_tf32, (cudaStream_t)stream);
}
int sfno_legendre_analysis_win(const sfno_plan* plan, const float* xm_win, int B, int Ctot, int c0, int C, float* sp,
                               int round_tf32, int accumulate, void* stream) {
  SFNO_CHECK(plan && xm_win && sp, "null argument");
  return legendre_analysis(plan, xm_win, B, C, Ctot, c0, sp, round_tf32, accumulate, (cudaStream_t)stream);
}
int sfno_legendre_synthesis_win(const sfno_plan* plan, const float* sp, int B, int Ctot, int c0, int C, float* xm_win,
                                void* stream) {
  SFNO_CHECK(plan && xm_win && sp, "null argument");
  return legendre_synthesis(plan, sp, B, C, Ctot, c0, xm_win, (cudaStream_t)stream);
}
int sfno_dhconv_fwd(const float* sp_in, const float* wp, float* sp_out, int B, int Cin, int Cout, int lmax, int mmax,
                    int l_offset, int m_offset, int round_tf32, void* stream) {
  SFNO_CHECK(sp_in && wp && sp_out, "null argument");
  return dhconv_fwd(sp_in, wp, sp_out, B, Cin, Cout, lmax, mmax, l_offset, m_offset, round_tf32, (cudaStream_t)stream);
}
int sfno_dhconv_bwd_x(const float* gsp_out, const float* wp, float* gsp_in, int B, int Cin, int Cout, int lmax,
                      int mmax, int l_offset, int m_offset, int round_tf32, void* stream) {
  SFNO_CHECK(gsp_out && wp && gsp_in, "null argument");
  return dhconv_bwd_x(gsp_out, wp, gsp_in, B, Cin, Cout, lmax, mmax, l_offset, m_offset, round_tf32,
                      (cudaStream_t)stream);
}
int sfno_dhconv_bwd_w(const float* sp_in, const float* gsp_out, float* gwp, int B, int Cin, int Cout, int lmax,
                      int mmax, int l_offset, int m_offset, void* stream) {
  SFNO_CHECK(sp_in && gsp_out && gwp, "null argument");
  return dhconv_bwd_w(sp_in, gsp_out, gwp, B, Cin, Cout, lmax, mmax, l_offset, m_offset, (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------- composite calls
namespace {
constexpr size_t kWsAlign = 1024;
size_t ws_round(size_t bytes) { return (bytes + kWsAlign - 1) / kWsAlign * kWsAlign; }
size_t xm_bytes(const sfno_plan* p, int B, int C) { return ws_round((size_t)sfno_xm_elems(p->nlat, p->mmax, B, C) * 4); }
size_t sp_bytes(const sfno_plan* p, int B, int C) { return ws_round((size_t)sfno_sp_elems(p->lmax, p->mmax, B, C) * 4); }
bool ws_ok(const void* ws, size_t have, size_t need) {
  return ws != nullptr && have >= need && (reinterpret_cast<uintptr_t>(ws) & 15u) == 0;
}
}  // namespace

size_t sfno_workspace_bytes(const sfno_plan* plan, const sfno_plan* plan_b, int op, int B, int C) {
  if (!plan || B <= 0 || C <= 0) return 0;
  switch (op) {
    case SFNO_OP_SHT_FWD:
    case SFNO_OP_SHT_INV: return xm_bytes(plan, B, C) + sp_bytes(plan, B, C);
    case SFNO_OP_BLOCK_FWD:
      if (!plan_b) return 0;
      // XM(in) and XM(out) share one region (the first is dead once the analysis has run); two SP buffers
      return std::max(xm_bytes(plan, B, C), xm_bytes(plan_b, B, C)) + 2 * sp_bytes(plan, B, C);
    default: return 0;
  }
}

int sfno_sht_fwd(const sfno_plan* plan, const void* x, int x_dtype, int B, int C, void* coeffs_c64, void* ws,
                 size_t ws_bytes, void* stream) {
  SFNO_CHECK(plan && x && coeffs_c64, "null argument");
  SFNO_CHECK(plan->m_offset == 0, "sfno_sht_fwd: azimuth-sharded plans go through the distributed modules");
  SFNO_CHECK(ws_ok(ws, ws_bytes, sfno_workspace_bytes(plan, nullptr, SFNO_OP_SHT_FWD, B, C)),
             "sfno_sht_fwd: workspace missing, misaligned or smaller than sfno_workspace_bytes()");
  cudaStream_t s = (cudaStream_t)stream;
  float* xm = static_cast<float*>(ws);
  float* sp = reinterpret_cast<float*>(static_cast<char*>(ws) + xm_bytes(plan, B, C));
  if (dft_fwd(x, x_dtype, B, C, C, plan->nlat, plan->nlon, plan->mmax, plan->mscale, xm, 1, s)) return 1;
  if (legendre_analysis(plan, xm, B, C, C, 0, sp, 0, 0, s)) return 1;
  return sfno_sp_to_c64(sp, B, C, plan->lmax, plan->mmax, 0, 0, coeffs_c64, stream);
}

int sfno_sht_inv(const sfno_plan* plan, const void* coeffs_c64, int B, int C, void* y, int y_dtype, void* ws,
                 size_t ws_bytes, void* stream) {
  SFNO_CHECK(plan && y && coeffs_c64, "null argument");
  SFNO_CHECK(plan->m_offset == 0, "sfno_sht_inv: azimuth-sharded plans go through the distributed modules");
  SFNO_CHECK(ws_ok(ws, ws_bytes, sfno_workspace_bytes(plan, nullptr, SFNO_OP_SHT_INV, B, C)),
             "sfno_sht_inv: workspace missing, misaligned or smaller than sfno_workspace_bytes()");
  cudaStream_t s = (cudaStream_t)stream;
  float* xm = static_cast<float*>(ws);
  float* sp = reinterpret_cast<float*>(static_cast<char*>(ws) + xm_bytes(plan, B, C));
  if (sfno_c64_to_sp(coeffs_c64, B, C, plan->lmax, plan->mmax, 0, plan->m_offset, 1, sp, stream)) return 1;
  if (legendre_synthesis(plan, sp, B, C, C, 0, xm, s)) return 1;
  return dft_inv(xm, B, C, C, plan->nlat, plan->nlon, plan->mmax, plan->mmax, plan->mscale + (plan->nlon / 2 + 1), y,
                 y_dtype, s);
}

int sfno_block_fwd(const sfno_plan* pf, const sfno_plan* pi, const void* x, int dtype, int B, int Cin, int Cout,
                   const float* wp, void* y, void* ws, size_t ws_bytes, void* stream) {
  SFNO_CHECK(pf && pi && x && wp && y, "null argument");
  SFNO_CHECK(pf->lmax == pi->lmax && pf->mmax == pi->mmax, "sfno_block_fwd: the two plans differ in lmax / mmax");
  SFNO_CHECK(pf->m_offset == 0 && pi->m_offset == 0, "sfno_block_fwd: azimuth-sharded plans are not supported");
  const int C = std::max(Cin, Cout);
  SFNO_CHECK(ws_ok(ws, ws_bytes, sfno_workspace_bytes(pf, pi, SFNO_OP_BLOCK_FWD, B, C)),
             "sfno_block_fwd: workspace missing, misaligned or smaller than sfno_workspace_bytes()");
  cudaStream_t s = (cudaStream_t)stream;
  char* base = static_cast<char*>(ws);
  float* xm = reinterpret_cast<float*>(base);
  float* sp_a = reinterpret_cast<float*>(base + std::max(xm_bytes(pf, B, C), xm_bytes(pi, B, C)));
  float* sp_b = reinterpret_cast<float*>(reinterpret_cast<char*>(sp_a) + sp_bytes(pf, B, C));
  // every intermediate feeds a TF32 contraction: store round-to-nearest TF32 values (see sfno_legendre_analysis)
  if (dft_fwd(x, dtype, B, Cin, Cin, pf->nlat, pf->nlon, pf->mmax, pf->mscale, xm, 1, s)) return 1;
  if (legendre_analysis(pf, xm, B, Cin, Cin, 0, sp_a, 1, 0, s)) return 1;
  if (dhconv_fwd(sp_a, wp, sp_b, B, Cin, Cout, pf->lmax, pf->mmax, 0, 0, 1, s)) return 1;
  if (legendre_synthesis(pi, sp_b, B, Cout, Cout, 0, xm, s)) return 1;
  return dft_inv(xm, B, Cout, Cout, pi->nlat, pi->nlon, pi->mmax, pi->mmax, pi->mscale + (pi->nlon / 2 + 1), y, dtype, s);
}

int sfno_dft_peer_supported(int nlon, int mmax) { return dft_peer_supported(nlon, mmax) ? 1 : 0; }
int sfno_dft_fwd_peer(const void* x, int x_dtype, int B, int C, int nlat_loc, int nlon, int mmax,
                      const float* mscale_dev, void* const* xm_ch_host, const int* chan_shapes_host, int nranks, int Kg,
                      int lat_off, int lat_span, int round_tf32, void* stream) {
  SFNO_CHECK(x && mscale_dev && xm_ch_host && chan_shapes_host, "null argument");
  return dft_fwd_peer(x, x_dtype, B, C, nlat_loc, nlon, mmax, mscale_dev, xm_ch_host, chan_shapes_host, nranks, Kg,
                      lat_off, lat_span, round_tf32, (cudaStream_t)stream);
}
int sfno_dft_inv_peer(void* const* xm_ch_host, const int* chan_shapes_host, int nranks, int Kg, int lat_off,
                      int lat_span, int B, int C, int nlat_loc, int nlon, int mmax, int m_valid,
                      const float* mscale_dev, void* y, int y_dtype, void* stream) {
  SFNO_CHECK(y && mscale_dev && xm_ch_host && chan_shapes_host, "null argument");
  return dft_inv_peer(xm_ch_host, chan_shapes_host, nranks, Kg, lat_off, lat_span, B, C, nlat_loc, nlon, mmax, m_valid,
                      mscale_dev, y, y_dtype, (cudaStream_t)stream);
}

}  // extern "C"
