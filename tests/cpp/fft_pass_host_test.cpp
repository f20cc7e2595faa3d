// Host emulation of the two-pass (16 x 15) FFT in makani_b200/csrc/fft_pass.cuh: one std::thread per lane,
// shuffles through a shared slot array + barrier.  Checks forward and inverse row transforms for
// nlon = 480 (R = 1) and nlon = 1440 (R = 3) against a direct double-precision DFT.
// Build: g++ -O2 -std=c++17 -pthread tests/cpp/fft_lane_host_test.cpp -o /tmp/fft_lane_host_test
#include <pthread.h>
#include <cmath>
#include <complex>
#include <cstdio>
#include <cstdlib>
#include <thread>
#include <vector>

#include "../../makani_b200/csrc/fft_pass.cuh"

using namespace sfno::fl;
namespace fp = sfno::fp;

struct HostWarp {
  int lane;
  volatile float* slots;
  pthread_barrier_t* bar;
  int lane16() const { return lane; }
  float exchange(float v, int src) {
    slots[lane] = v;
    pthread_barrier_wait(bar);
    float r = slots[src];
    pthread_barrier_wait(bar);
    return r;
  }
  float shfl_xor(float v, int mask) { return exchange(v, lane ^ mask); }
  float shfl_idx(float v, int src) { return exchange(v, src); }
  void sync() { pthread_barrier_wait(bar); }
};

static std::vector<cf> make_tw(int count, double denom, int mul_rows = 0) {
  std::vector<cf> t(count);
  for (int j = 0; j < count; ++j) {
    double a = -2.0 * M_PI * j / denom;
    t[j] = mk((float)cos(a), (float)sin(a));
  }
  return t;
}

template <int R>
static double run(bool verbose) {
  const int W = 480 * R;
  std::vector<cf> twA(16 * 15), tw480 = make_tw(241, 480.0), tw1440 = make_tw(241, 1440.0);
  for (int k1 = 0; k1 < 16; ++k1)
    for (int n2 = 0; n2 < 15; ++n2) {
      double a = -2.0 * M_PI * n2 * k1 / 240.0;
      twA[k1 * 15 + n2] = mk((float)cos(a), (float)sin(a));
    }
  std::vector<cf> work(240);

  std::vector<float> x(W);
  srand(1234 + R);
  for (auto& v : x) v = (float)rand() / RAND_MAX - 0.5f;

  // reference forward
  std::vector<std::complex<double>> Xref(241);
  for (int m = 0; m <= 240; ++m) {
    std::complex<double> s = 0;
    for (int n = 0; n < W; ++n) s += (double)x[n] * std::polar(1.0, -2.0 * M_PI * m * n / W);
    Xref[m] = s;
  }

  // ---- forward through the lane code
  std::vector<std::complex<float>> X(241);
  volatile float slots[16];
  pthread_barrier_t bar;
  pthread_barrier_init(&bar, nullptr, 16);
  {
    std::vector<std::thread> th;
    for (int lane = 0; lane < 16; ++lane)
      th.emplace_back([&, lane] {
        HostWarp w{lane, slots, &bar};
        cf Xl[15], x240;
        fp::rfft_row_fwd<R, HostWarp>(w, [&](int n, int r) { return mk(x[2 * R * n + r], x[2 * R * n + R + r]); }, work.data(),
                            Xl, x240, twA.data(), tw480.data(), tw1440.data());
        for (int k2 = 0; k2 < 15; ++k2) X[lane + 16 * k2] = {0.5f * Xl[k2].x, 0.5f * Xl[k2].y};
        if (lane == 0) X[240] = {0.5f * x240.x, 0.5f * x240.y};
      });
    for (auto& t : th) t.join();
  }
  double err_f = 0, nrm = 0;
  for (int m = 0; m <= 240; ++m) {
    err_f = std::max(err_f, std::abs(std::complex<double>(X[m]) - Xref[m]));
    nrm = std::max(nrm, std::abs(Xref[m]));
  }
  if (verbose) printf("R=%d forward  max|err| = %.3e (max|X| = %.3f)\n", R, err_f, nrm);

  // ---- inverse: random spectrum (complex everywhere, like real use where imag of bin 0 / Nyquist is ignored)
  std::vector<std::complex<float>> S(241);
  for (auto& v : S) v = {(float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f};
  std::vector<double> yref(W);
  for (int n = 0; n < W; ++n) {
    double s = 0;
    for (int m = 0; m <= 240; ++m) {
      double cm = (m == 0 || 2 * m == W) ? 1.0 : 2.0;
      s += cm * (std::complex<double>(S[m]) * std::polar(1.0, 2.0 * M_PI * m * n / W)).real();
    }
    yref[n] = s;
  }
  std::vector<float> y(W, 0.f);
  {
    std::vector<std::thread> th;
    for (int lane = 0; lane < 16; ++lane)
      th.emplace_back([&, lane] {
        HostWarp w{lane, slots, &bar};
        cf H[15], HP[15];
        for (int k2 = 0; k2 < 15; ++k2) {
          const int k = lane + 16 * k2;
          H[k2] = mk(S[k].real(), S[k].imag());
          HP[k2] = mk(S[240 - k].real(), S[240 - k].imag());
        }
        fp::irfft_row<R>(w, H, HP, work.data(),
                         [&](int n, int r, cf z) { y[2 * R * n + r] = z.x; y[2 * R * n + R + r] = z.y; }, twA.data(),
                         tw480.data(), tw1440.data());
      });
    for (auto& t : th) t.join();
  }
  double err_i = 0, nrm_i = 0;
  for (int n = 0; n < W; ++n) {
    err_i = std::max(err_i, std::fabs(y[n] - yref[n]));
    nrm_i = std::max(nrm_i, std::fabs(yref[n]));
  }
  if (verbose) printf("R=%d inverse  max|err| = %.3e (max|y| = %.3f)\n", R, err_i, nrm_i);
  pthread_barrier_destroy(&bar);
  return std::max(err_f / nrm, err_i / nrm_i);
}

int main() {
  double e1 = run<1>(true);
  double e3 = run<3>(true);
  bool ok = e1 < 2e-6 && e3 < 2e-6;
  printf("%s\n", ok ? "FFT_PASS_OK" : "FFT_PASS_FAIL");
  return ok ? 0 : 1;
}
