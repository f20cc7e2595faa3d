// Host emulation of makani_b200/csrc/wpack.cuh: the per-thread pack / unpack bodies run for every CTA of the grid and
// every one of its 256 threads (phase 1 for all threads, then phase 2 - the kernels have exactly one barrier), for
// TL = 64 and 32, even and odd L, channel counts that are not multiples of 4 or 64.  Checked against a plain loop:
//   pack   : WP[l][reim][o][i] == tf32(w[i][o][l].{re,im}), pads zero where written
//   unpack : gw[i][o][l] == (GWP[l][0][o][i], GWP[l][1][o][i]) - bit exact
//   g++ -O2 -std=c++17 -I/usr/local/cuda/include tests/cpp/wpack_host_test.cpp -o tests/cpp/wpack_host_test.bin
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../makani_b200/csrc/wpack.cuh"

using namespace sfno::wpack;

static float frand() { return (float)rand() / RAND_MAX * 2.f - 1.f; }

template <int TL>
static int check(int Cin, int Cout, int L) {
  const int Cinp = (Cin + 3) & ~3, Coutp = (Cout + 3) & ~3;
  std::vector<float2> w((size_t)Cin * Cout * L), gw((size_t)Cin * Cout * L, make_float2(-7.f, -7.f));
  for (auto& v : w) v = make_float2(frand(), frand());
  const float SENT = 12345.f;
  std::vector<float> wp((size_t)L * 2 * Coutp * Cinp, SENT), gwp((size_t)L * 2 * Coutp * Cinp);
  for (auto& v : gwp) v = frand();
  const int gx = (L + TL - 1) / TL, gy = (Cinp + kTI - 1) / kTI;
  Tile<TL>* t = new Tile<TL>();
  for (int o = 0; o < Cout; ++o)
    for (int by = 0; by < gy; ++by)
      for (int bx = 0; bx < gx; ++bx) {
        for (int tid = 0; tid < 256; ++tid) pack_load<TL>(tid, *t, w.data(), o, bx * TL, by * kTI, Cin, Cout, L);
        for (int tid = 0; tid < 256; ++tid) pack_store<TL>(tid, *t, wp.data(), o, bx * TL, by * kTI, Cinp, Coutp, L);
        for (int tid = 0; tid < 256; ++tid) unpack_load<TL>(tid, *t, gwp.data(), o, bx * TL, by * kTI, Cinp, Coutp, L);
        for (int tid = 0; tid < 256; ++tid) unpack_store<TL>(tid, *t, gw.data(), o, bx * TL, by * kTI, Cin, Cout, L);
      }
  delete t;
  int bad = 0;
  for (int l = 0; l < L; ++l)
    for (int s = 0; s < 2; ++s)
      for (int o = 0; o < Coutp; ++o)
        for (int i = 0; i < Cinp; ++i) {
          const float got = wp[(((size_t)l * 2 + s) * Coutp + o) * Cinp + i];
          float want;
          if (o >= Cout) want = SENT;                               // rows of pad output channels are never written
          else if (i >= Cin) want = 0.f;                            // pad input channels: exact zeros
          else {
            const float2 v = w[((size_t)i * Cout + o) * L + l];
            want = round_tf32(s ? v.y : v.x);
          }
          if (got != want && bad++ < 5) std::printf("pack TL=%d l=%d s=%d o=%d i=%d got %g want %g\n", TL, l, s, o, i, got, want);
        }
  for (int i = 0; i < Cin; ++i)
    for (int o = 0; o < Cout; ++o)
      for (int l = 0; l < L; ++l) {
        const float2 got = gw[((size_t)i * Cout + o) * L + l];
        const float wr = gwp[(((size_t)l * 2 + 0) * Coutp + o) * Cinp + i];
        const float wi = gwp[(((size_t)l * 2 + 1) * Coutp + o) * Cinp + i];
        if ((got.x != wr || got.y != wi) && bad++ < 5) std::printf("unpack TL=%d i=%d o=%d l=%d\n", TL, i, o, l);
      }
  std::printf("TL=%d Cin=%d Cout=%d L=%d: %s\n", TL, Cin, Cout, L, bad ? "FAILED" : "ok");
  return bad;
}

int main() {
  srand(1);
  int bad = 0;
  const int cases[][3] = {{70, 3, 50}, {64, 2, 33}, {6, 5, 7}, {130, 2, 96}, {384, 1, 240}};
  for (auto& c : cases) {
    bad += check<64>(c[0], c[1], c[2]);
    bad += check<32>(c[0], c[1], c[2]);
  }
  return bad ? 1 : 0;
}
