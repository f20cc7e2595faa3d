// Two-pass (16 x 15) length-240 complex FFT for the longitude DFT kernels, exchange through shared memory.
//
// A row of 480 real samples is packed into 240 complex points z[n] = x[2n] + i x[2n+1] and transformed by the
// 16 lanes of a half-warp:
//     n = 15 n1 + n2   (n1 in [0,16), n2 in [0,15)),      k = k1 + 16 k2   (k1 in [0,16), k2 in [0,15))
//     Z[k1 + 16 k2] = sum_n2 w15^(n2 k2) * [ w240^(n2 k1) * sum_n1 w16^(n1 k1) z[15 n1 + n2] ]
//   pass A  lane n2 (< 15): 16-point FFT over n1 in registers, twiddle by w240^(n2 k1), store work[k1*15 + n2]
//   pass B  lane k1       : load work[k1*15 + n2], 15-point DFT over n2 in registers  ->  Z[k1 + 16 k2]
// Both shared-memory access patterns are conflict free for 8-byte elements (lane stride 1 / lane stride 15).
// Only the real-input split needs data of another lane (bin 240-k lives on lane 16-k1): two shuffles per bin.
// Compared with the 16-lane shuffle FFT of fft_lane.cuh this removes four shuffle butterfly stages
// (8 instructions per complex value and stage) - about half of the instructions per row.
//
// The inverse runs the transposed flow graph (15-point DFT first, then 16-point FFT).
//
// Written against a "Warp" policy (lane16 / shfl_idx / sync over a 16-lane group) so the same code runs on
// the device (DeviceWarp16p) and in a host emulation with one thread per lane (tests/cpp/fft_pass_host_test.cpp).
#pragma once
#include "fft_lane.cuh"

namespace sfno {
namespace fp {

using fl::cf;
using fl::mk;
using fl::cadd;
using fl::csub;
using fl::cmul;
using fl::cconj;
using fl::cscale;

// Twiddles formed in registers: e^{-2 pi i k / 480} and e^{-2 pi i k / 1440} at k = lane + 16 k2 as (table entry of
// the lane) x (compile-time constant of k2), and the twiddles of the 16 x 15 exchange as powers of one table entry per
// lane - two packed instructions each instead of an 8-byte shared-memory load (38 percent of the kernels'
// shared-memory wavefronts are twiddle loads).  Measured on B200 with the packed arithmetic of fft_lane.cuh
// (profiles/r2t_dft_f32x2_ab.txt): 240x480 forward / inverse 69.6 -> 67.9 / 63.9 -> 61.9 us, 721x1440 464 -> 472 /
// 424 -> 423 us (bound by latency at 16 warps per SM, not by either pipe).  So the R = 1 rows (nlon = 480) use the
// register form and the R = 3 rows keep the table loads:  SFNO_TW_REGS  0 = tables everywhere, 1 = registers at
// R = 1 only (default), 2 = registers everywhere.
#ifndef SFNO_TW_REGS
#define SFNO_TW_REGS 1
#endif
template <int R>
struct tw_regs { static constexpr bool value = SFNO_TW_REGS == 2 || (SFNO_TW_REGS == 1 && R == 1); };
#ifndef SFNO_TW_REGS_TWIST
#define SFNO_TW_REGS_TWIST 0
#endif
// e^{-2 pi i k2 / 30} = w480^(16 k2)  and  e^{-2 pi i k2 / 90} = w1440^(16 k2),  k2 = 0..14.  Switches, not arrays:
// every call site has a compile-time k2 after unrolling, so they fold to immediates (an array captured by a lambda or
// indexed in a large unrolled body ended up in local memory).
FL_HD cf w30c(int k2) {
  switch (k2) {
    case 1: return mk(0.9781476007338057f, -0.20791169081775931f);
    case 2: return mk(0.9135454576426009f, -0.40673664307580015f);
    case 3: return mk(0.8090169943749475f, -0.5877852522924731f);
    case 4: return mk(0.6691306063588582f, -0.7431448254773941f);
    case 5: return mk(0.5000000000000001f, -0.8660254037844386f);
    case 6: return mk(0.30901699437494745f, -0.9510565162951535f);
    case 7: return mk(0.10452846326765346f, -0.9945218953682733f);
    case 8: return mk(-0.10452846326765333f, -0.9945218953682734f);
    case 9: return mk(-0.30901699437494734f, -0.9510565162951536f);
    case 10: return mk(-0.4999999999999998f, -0.8660254037844387f);
    case 11: return mk(-0.6691306063588579f, -0.7431448254773945f);
    case 12: return mk(-0.8090169943749473f, -0.5877852522924732f);
    case 13: return mk(-0.913545457642601f, -0.40673664307580004f);
    case 14: return mk(-0.9781476007338057f, -0.20791169081775931f);
    default: return mk(1.f, 0.f);
  }
}
FL_HD cf w90c(int k2) {
  switch (k2) {
    case 1: return mk(0.9975640502598242f, -0.0697564737441253f);
    case 2: return mk(0.9902680687415704f, -0.13917310096006544f);
    case 3: return mk(0.9781476007338057f, -0.20791169081775931f);
    case 4: return mk(0.9612616959383189f, -0.27563735581699916f);
    case 5: return mk(0.9396926207859084f, -0.3420201433256687f);
    case 6: return mk(0.9135454576426009f, -0.40673664307580015f);
    case 7: return mk(0.882947592858927f, -0.4694715627858908f);
    case 8: return mk(0.848048096156426f, -0.5299192642332049f);
    case 9: return mk(0.8090169943749475f, -0.5877852522924731f);
    case 10: return mk(0.766044443118978f, -0.6427876096865393f);
    case 11: return mk(0.7193398003386512f, -0.6946583704589973f);
    case 12: return mk(0.6691306063588582f, -0.7431448254773941f);
    case 13: return mk(0.6156614753256583f, -0.788010753606722f);
    case 14: return mk(0.5591929034707468f, -0.8290375725550417f);
    default: return mk(1.f, 0.f);
  }
}

// 4-point DFT, natural order.  w = e^{SGN i pi/2} = SGN i
template <int SGN>
FL_HD void dft4(cf& a0, cf& a1, cf& a2, cf& a3) {
  const cf s02 = cadd(a0, a2), d02 = csub(a0, a2), s13 = cadd(a1, a3), d13 = csub(a1, a3);
  const cf t = fl::mul_i<SGN>(d13);
  a0 = cadd(s02, s13);
  a2 = csub(s02, s13);
  a1 = cadd(d02, t);
  a3 = csub(d02, t);
}

// 16-point FFT, natural order in and out (radix 4 x 4).  n = 4a + b, k = c + 4d.
template <int SGN>
FL_HD void fft16(cf (&v)[16]) {
  const float C1 = 0.9238795325112867f, S1 = 0.3826834323650898f, R2 = 0.7071067811865476f;
  cf y[4][4];
#pragma unroll
  for (int b = 0; b < 4; ++b) {
#pragma unroll
    for (int a = 0; a < 4; ++a) y[b][a] = v[4 * a + b];
    dft4<SGN>(y[b][0], y[b][1], y[b][2], y[b][3]);          // second index is now c
  }
  // twiddles w16^(b c), w16 = e^{SGN 2 pi i / 16}
  y[1][1] = cmul(y[1][1], mk(C1, SGN * S1));                // w^1
  y[1][2] = cmul(y[1][2], mk(R2, SGN * R2));                // w^2
  y[1][3] = cmul(y[1][3], mk(S1, SGN * C1));                // w^3
  y[2][1] = cmul(y[2][1], mk(R2, SGN * R2));                // w^2
  y[2][2] = fl::mul_i<SGN>(y[2][2]);                        // w^4 = SGN i
  y[2][3] = cmul(y[2][3], mk(-R2, SGN * R2));               // w^6
  y[3][1] = cmul(y[3][1], mk(S1, SGN * C1));                // w^3
  y[3][2] = cmul(y[3][2], mk(-R2, SGN * R2));               // w^6
  y[3][3] = cmul(y[3][3], mk(-C1, -SGN * S1));              // w^9 = -w^1
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    dft4<SGN>(y[0][c], y[1][c], y[2][c], y[3][c]);          // first index is now d
#pragma unroll
    for (int d = 0; d < 4; ++d) v[c + 4 * d] = y[d][c];
  }
}

// Twiddle tables (built once on the host in double precision):
//   twA[k1*15 + n2] = e^{-2 pi i n2 k1 / 240}     k1 in [0,16), n2 in [0,15)
//   tw480[k]        = e^{-2 pi i k / 480}          k in [0,241)
//   tw1440[m]       = e^{-2 pi i m / 1440}         m in [0,241)

// forward complex FFT-240.  loadz(n) returns z[n].  out: Z[k2] = Z[lane + 16 k2].
// `work` is a 240-element scratch row in shared memory owned by this 16-lane group.
template <bool TWR, class Warp, class LoadZ>
FL_HD void fft240p_fwd(Warp& w, LoadZ loadz, cf* work, const cf* twA, cf (&Z)[15]) {
  const int t = w.lane16();
  if (t < 15) {
    cf a[16];
#pragma unroll
    for (int n1 = 0; n1 < 16; ++n1) a[n1] = loadz(15 * n1 + t);
    fft16<-1>(a);
    work[t] = a[0];
    if constexpr (TWR) {
      const cf p = twA[15 + t];                                // e^{-2 pi i t / 240}; twA[k1*15 + t] = p^k1
      cf pw = p;
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) {
        work[k1 * 15 + t] = cmul(a[k1], pw);
        if (k1 < 15) pw = cmul(pw, p);
      }
    } else {
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) work[k1 * 15 + t] = cmul(a[k1], twA[k1 * 15 + t]);
    }
  }
  w.sync();
#pragma unroll
  for (int n2 = 0; n2 < 15; ++n2) Z[n2] = work[t * 15 + n2];
  fl::dft15<-1>(Z);
  w.sync();
}

// inverse (unnormalised, e^{+}) complex FFT-240.  in: V[k2] = V[lane + 16 k2] (destroyed);  storez(n, z[n]).
template <bool TWR, class Warp, class StoreZ>
FL_HD void fft240p_inv(Warp& w, cf (&V)[15], cf* work, const cf* twA, StoreZ storez) {
  const int t = w.lane16();
  fl::dft15<+1>(V);                                          // index is now n2
  work[t * 15] = V[0];
  if constexpr (TWR) {
    const cf q = cconj(twA[t * 15 + 1]);                       // e^{+2 pi i t / 240}; conj(twA[t*15 + n2]) = q^n2
    cf qw = q;
#pragma unroll
    for (int n2 = 1; n2 < 15; ++n2) {
      work[t * 15 + n2] = cmul(V[n2], qw);
      if (n2 < 14) qw = cmul(qw, q);
    }
  } else {
#pragma unroll
    for (int n2 = 1; n2 < 15; ++n2) work[t * 15 + n2] = cmul(V[n2], cconj(twA[t * 15 + n2]));
  }
  w.sync();
  if (t < 15) {
    cf a[16];
#pragma unroll
    for (int k1 = 0; k1 < 16; ++k1) a[k1] = work[k1 * 15 + t];
    fft16<+1>(a);
#pragma unroll
    for (int n1 = 0; n1 < 16; ++n1) storez(15 * n1 + t, a[n1]);
  }
  w.sync();
}

// real-input split.  in: Z[k2] = Z[lane + 16 k2];  emit(k2, 2 X[lane + 16 k2]) for the length-480 real sequence
// (TWICE the spectrum: the factor 1/2 is left to the caller, who folds it into its output scale);
// returns 2 X[240] (meaningful on lane 0 only).  Bin 240-k of lane k1 > 0 is register 14-k2 of lane 16-k1;
// lane 0 pairs its own registers (k2 <-> 15-k2, bin 0 with itself).
//   2 X = (Z + conj(Zp)) - i w (Z - conj(Zp)),   w = e^{-2 pi i k / 480}
template <bool TWR, class Warp, class Emit>
FL_HD float rfft480_split(Warp& w, const cf (&Z)[15], Emit emit, const cf* tw480) {
  const int k1 = w.lane16();
  const int src = (16 - k1) & 15;
  const cf base = tw480[k1];
#pragma unroll
  for (int k2 = 0; k2 < 15; ++k2) {
    const cf c = (!TWR || k2 == 0) ? (k2 == 0 ? base : tw480[k1 + 16 * k2]) : cmul(base, w30c(k2));
    const cf zs = mk(w.shfl_idx(Z[14 - k2].x, src), w.shfl_idx(Z[14 - k2].y, src));
    const cf z0 = (k2 == 0) ? Z[0] : Z[15 - k2];
    const cf zp = (k1 == 0) ? z0 : zs;
    const cf zc = cconj(zp);
    const cf p = cmul(csub(Z[k2], zc), c);                    // w (Z - conj(Zp))
    emit(k2, cadd(cadd(Z[k2], zc), fl::mul_i<-1>(p)));        // (Z + conj(Zp)) - i w (Z - conj(Zp))
  }
  return 2.f * (Z[0].x - Z[0].y);
}

// ---------------------------------------------------------------------------------------------
// Row-level composition for nlon = 480 * R (R = 1: 240x480 Gauss grid, R = 3: 721x1440 equiangular).
//   forward, R = 3: x_r[q] = x[3q + r] (three length-480 real FFTs), X[m] = sum_r w1440^(m r) X_r[m], m <= 240
//                   (outputs are TWICE the spectrum, see rfft480_split)
//   inverse, R = 3: x[3q + r] = irfft480( X[m] conj(w1440)^(m r) ), bin 240 doubled (it is not Nyquist there)
// loadz(n, r) returns (x[2 R n + r], x[2 R n + R + r]);  storez(n, r, z) writes that pair back.
template <int R, class Warp, class LoadZ>
FL_HD void rfft_row_fwd(Warp& w, LoadZ loadz, cf* work, cf (&X)[15], cf& x240, const cf* twA, const cf* tw480,
                        const cf* tw1440) {
  const int k1 = w.lane16();
  constexpr bool TWR = tw_regs<R>::value;
  // Horner in the sub-sequence index: X = X_0 + t (X_1 + t X_2),  t = w1440^m  - one complex multiply per bin and r
#pragma unroll
  for (int r = R - 1; r >= 0; --r) {
    cf Z[15];
    fft240p_fwd<TWR>(w, [&](int n) { return loadz(n, r); }, work, twA, Z);
    if (r == R - 1) {
      const float nyq = rfft480_split<TWR>(w, Z, [&](int k2, cf v) { X[k2] = v; }, tw480);
      x240 = mk(nyq, 0.f);
    } else if constexpr (TWR) {
      const cf b1440 = tw1440[k1];
      const float nyq = rfft480_split<TWR>(w, Z, [&](int k2, cf v) {
        const cf t = k2 == 0 ? b1440 : cmul(b1440, w90c(k2));
        X[k2] = cadd(v, cmul(t, X[k2]));
      }, tw480);
      x240 = cadd(mk(nyq, 0.f), cmul(mk(0.5f, -0.86602540378443865f), x240));      // w1440^240 = e^{-i pi/3}
    } else {
      const float nyq = rfft480_split<TWR>(w, Z, [&](int k2, cf v) {
        X[k2] = cadd(v, cmul(tw1440[k1 + 16 * k2], X[k2]));
      }, tw480);
      x240 = cadd(mk(nyq, 0.f), cmul(tw1440[240], x240));
    }
  }
}

// Pre-processing of the inverse real FFT for one bin pair: from h = H[k], hp = H[240-k] (Hermitian half-spectrum of a
// length-480 real sequence) build V[k] = A + i conj(w) B,  A = h + conj(hp), B = h - conj(hp),  so that
// x[2n] + i x[2n+1] = sum_k V[k] e^{+2 pi i k n / 240}.
FL_HD cf irfft480_prep(cf h, cf hp, cf w480_k) {
  const cf hc = cconj(hp);
  const cf p = cmul(csub(h, hc), cconj(w480_k));              // conj(w) B
  return cadd(cadd(h, hc), fl::mul_i<+1>(p));                 // A + i conj(w) B
}

// For nlon = 480 R the sub-sequence r of the output needs the spectrum twisted by conj(w1440^k)^r (bin k) and
// conj(w1440^(240-k))^r (bin 240-k): advance (h, hp) by one step of r.
FL_HD void irfft_twist_step(int k, cf& h, cf& hp, const cf* tw1440) {
  h = cmul(h, cconj(tw1440[k]));
  hp = cmul(hp, cconj(tw1440[240 - k]));
}
// The k = 0 pair is made real (bin 240 is a true Nyquist bin only for R = 1; otherwise an ordinary, doubled bin).
template <int R>
FL_HD void irfft_fix0(int k, cf& h, cf& hp) {
  if (k == 0) {
    h = mk(h.x, 0.f);
    hp = mk((R == 1 ? 1.f : 2.f) * hp.x, 0.f);
  }
}

// H[k2] / HP[k2]: the (already scaled) spectrum at bins k = lane + 16 k2 and 240 - k; destroyed (twisted in place).
template <int R, class Warp, class StoreZ>
FL_HD void irfft_row(Warp& w, cf (&H)[15], cf (&HP)[15], cf* work, StoreZ storez, const cf* twA, const cf* tw480,
                     const cf* tw1440) {
  const int k1 = w.lane16();
  constexpr bool TWR = tw_regs<R>::value;
  const cf b480 = tw480[k1];
#if SFNO_TW_REGS_TWIST
  const cf b1440 = R > 1 ? tw1440[k1] : mk(1.f, 0.f);
#endif
#pragma unroll
  for (int r = 0; r < R; ++r) {
    cf V[15];
#pragma unroll
    for (int k2 = 0; k2 < 15; ++k2) {
      const int k = k1 + 16 * k2;
#if SFNO_TW_REGS_TWIST
      if (r > 0 && TWR) {
        // w1440^k = (lane entry) x (constant of k2);  w1440^(240 - k) = e^{-i pi/3} conj(w1440^k)
        const cf t = k2 == 0 ? b1440 : cmul(b1440, w90c(k2));
        H[k2] = cmul(H[k2], cconj(t));
        HP[k2] = cmul(HP[k2], cmul(mk(0.5f, 0.86602540378443865f), t));       // conj(w^(240-k)) = e^{+i pi/3} w^k
      }
      if (r > 0 && !TWR) irfft_twist_step(k, H[k2], HP[k2], tw1440);
#else
      // the twist keeps its table loads: with H, HP and V live (90 registers) the computed variant spills
      if (r > 0) irfft_twist_step(k, H[k2], HP[k2], tw1440);
#endif
      const cf c480 = (!TWR || k2 == 0) ? (k2 == 0 ? b480 : tw480[k]) : cmul(b480, w30c(k2));
      cf h = H[k2], hp = HP[k2];
      if (k2 == 0) irfft_fix0<R>(k, h, hp);
      V[k2] = irfft480_prep(h, hp, c480);
    }
    fft240p_inv<TWR>(w, V, work, twA, [&](int n, cf z) { storez(n, r, z); });
  }
}

}  // namespace fp
}  // namespace sfno
