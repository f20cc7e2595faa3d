// Host check of makani_b200/csrc/pointwise_math.cuh: the fast erf / GELU / GELU' against double-precision references on
// a dense grid.  Prints the maximum absolute errors; exit status 0 when they are inside the stated bounds.
//   g++ -O2 -std=c++17 -DSFNO_FAST_ERF=1 tests/cpp/gelu_host_test.cpp -o tests/cpp/gelu_host_test.bin
#include <cmath>
#include <cstdio>

#include "../../makani_b200/csrc/pointwise_math.cuh"

int main() {
  double e_erf = 0, e_gelu = 0, e_grad = 0, e_gauss = 0;
  for (int i = -120000; i <= 120000; ++i) {
    const float z = i * 1e-4f;                                  // [-12, 12]
    float g;
    const double er = sfno::pw_erf_fast(z, &g);
    e_erf = std::fmax(e_erf, std::fabs(er - std::erf((double)z)));
    e_gauss = std::fmax(e_gauss, std::fabs((double)g - std::exp(-(double)z * z)));
    const double zd = z;
    const double gelu = 0.5 * zd * (1.0 + std::erf(zd / std::sqrt(2.0)));
    const double grad = 0.5 * (1.0 + std::erf(zd / std::sqrt(2.0))) + zd * std::exp(-0.5 * zd * zd) / std::sqrt(2.0 * M_PI);
    e_gelu = std::fmax(e_gelu, std::fabs((double)sfno::pw_gelu(z) - gelu) / std::fmax(1.0, std::fabs(zd)));
    e_grad = std::fmax(e_grad, std::fabs((double)sfno::pw_gelu_grad(z) - grad));
  }
  std::printf("fast_erf=%d max|erf err|=%.3g max|gauss err|=%.3g max|gelu err|/max(1,|z|)=%.3g max|gelu' err|=%.3g\n",
              SFNO_FAST_ERF, e_erf, e_gauss, e_gelu, e_grad);
  const bool ok = e_erf < 5e-7 && e_gauss < 2e-7 && e_gelu < 5e-7 && e_grad < 1e-6;
  return ok ? 0 : 1;
}
