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
  SFNO_CHECK(nlat > 0 && nlon > 0 && lmax > 0 && mmax > 0 && m_offset >= 0, "bad sizes");
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
  cudaError_t e = cudaMalloc(&p->table, n * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->table, h.data(), n * sizeof(float), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    if (p->table) cudaFree(p->table);
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
  delete plan;
  return 0;
}

int sfno_dft_fwd(const void* x, int x_dtype, int B, int C, int nlat, int nlon, int mmax, const float* mscale_dev,
                 float* xm, int round_tf32, void* stream) {
  SFNO_CHECK(x && xm && mscale_dev, "null argument");
  return dft_fwd(x, x_dtype, B, C, nlat, nlon, mmax, mscale_dev, xm, round_tf32, (cudaStream_t)stream);
}
int sfno_dft_inv(const float* xm, int B, int C, int nlat, int nlon, int mmax, int m_valid, const float* mscale_dev,
                 void* y, int y_dtype, void* stream) {
  SFNO_CHECK(xm && y && mscale_dev, "null argument");
  return dft_inv(xm, B, C, nlat, nlon, mmax, m_valid, mscale_dev, y, y_dtype, (cudaStream_t)stream);
}
int sfno_legendre_analysis(const sfno_plan* plan, const float* xm, int B, int C, float* sp, int round_tf32,
                           void* stream) {
  SFNO_CHECK(plan && xm && sp, "null argument");
  return legendre_analysis(plan, xm, B, C, sp, round_tf32, (cudaStream_t)stream);
}
int sfno_legendre_synthesis(const sfno_plan* plan, const float* sp, int B, int C, float* xm, void* stream) {
  SFNO_CHECK(plan && xm && sp, "null argument");
  return legendre_synthesis(plan, sp, B, C, xm, (cudaStream_t)stream);
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

}  // extern "C"
