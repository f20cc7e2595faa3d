// Per-lane building blocks of the length-480 real FFT used by the longitude DFT kernels.
//
// One row (480 real samples = 240 complex points z[n] = x[2n] + i x[2n+1]) is transformed by 16 lanes.
// With n = 16 a + b (lane b, a in [0,15)) and k = k1 + 15 k2:
//     Z[k1 + 15 k2] = sum_b w16^(b k2) * [ w240^(b k1) * sum_a w15^(a k1) z[16 a + b] ]
//   step A: 15-point DFT over a, in registers (3 x 5 Cooley-Tukey)
//   step B: twiddle by w240^(b k1)
//   step C: 16-point DFT across the 16 lanes - radix-2 butterflies through warp shuffles
//           (DIF: natural in, bit-reversed out, so lane p ends up with k2 = bitrev4(p))
//   step D: split the packed complex FFT into the real-input spectrum X[0..240]
// The inverse runs the transposed flow graph (DIT butterflies take the bit-reversed order back).
//
// The code is written against a "Warp" policy (shfl_xor / shfl_idx over a 16-lane group) so the very
// same functions run on the device (DeviceWarp16, __shfl_*_sync) and in a host emulation with one
// thread per lane (tests/cpp/fft_lane_host_test.cpp) - there is no GPU in the build container.
#pragma once

#if defined(__CUDACC__)
#define FL_HD __host__ __device__ __forceinline__
#else
#define FL_HD inline
#endif

namespace sfno {
namespace fl {

struct cf {
  float x, y;
};
FL_HD cf mk(float x, float y) { cf r; r.x = x; r.y = y; return r; }
// sm_100 has packed fp32 arithmetic (add / sub / mul / fma .f32x2 -> FADD2 / FMUL2 / FFMA2): one issue slot for both
// components of a complex number.  The FP32 pipe needs two cycles for it, so nothing is gained in FLOP/s, but the DFT
// kernels are bound by issue slots (62 % of their instructions are FP32 adds / FMAs), not by the FP32 pipe.
// ptxas folds swaps, sign patterns and scalar broadcasts into operand modifiers (FFMA2 -R.F32x2.LO_HI.NP, R.F32), so a
// complex product is two instructions.  Measured on B200 (profiles/r2t_dft_f32x2_ab.txt; C = 384, bf16): longitude DFT
// 721x1440 forward 494 -> 464 us, inverse 479 -> 424 us; 240x480 70.9 -> 69.6 / 66.6 -> 63.9 us; FP issue slots per row
// 1926 -> 966.  -DSFNO_F32X2=0 keeps the scalar forms (A/B runs; the host emulation always uses them).
#ifndef SFNO_F32X2
#define SFNO_F32X2 1
#endif
#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ >= 1000) && SFNO_F32X2
#define SFNO_PACKED_CF 1
__device__ __forceinline__ unsigned long long cf_pack(cf a) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a.x), "f"(a.y));
  return r;
}
__device__ __forceinline__ cf cf_unpack(unsigned long long v) {
  cf r;
  asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
  return r;
}
__device__ __forceinline__ cf cadd(cf a, cf b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(cf_pack(a)), "l"(cf_pack(b)));
  return cf_unpack(r);
}
__device__ __forceinline__ cf csub(cf a, cf b) {
  unsigned long long r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(cf_pack(a)), "l"(cf_pack(b)));
  return cf_unpack(r);
}
// complex product in two instructions: a * b.x + (i a) * b.y; ptxas folds the swap and the sign pattern of (i a) into
// operand modifiers (FFMA2 -R.F32x2.LO_HI.NP) and the pairs {b.x, b.x}, {b.y, b.y} into broadcast operands
__device__ __forceinline__ unsigned long long cf_pack2(float x, float y) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(x), "f"(y));
  return r;
}
__device__ __forceinline__ cf cmul(cf a, cf b) {
  unsigned long long t, r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(t) : "l"(cf_pack(a)), "l"(cf_pack2(b.x, b.x)));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(cf_pack2(a.y, a.x)), "l"(cf_pack2(-b.y, b.y)), "l"(t));
  return cf_unpack(r);
}
// a * s + c  and  a * s  with a real scalar s (the pair {s, s} becomes a broadcast operand)
__device__ __forceinline__ cf cfma(cf a, float s, cf c) {
  unsigned long long r, ss;
  asm("mov.b64 %0, {%1, %1};" : "=l"(ss) : "f"(s));
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(cf_pack(a)), "l"(ss), "l"(cf_pack(c)));
  return cf_unpack(r);
}
__device__ __forceinline__ cf cscale(cf a, float s) {
  unsigned long long r, ss;
  asm("mov.b64 %0, {%1, %1};" : "=l"(ss) : "f"(s));
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(cf_pack(a)), "l"(ss));
  return cf_unpack(r);
}
#else
#define SFNO_PACKED_CF 0
FL_HD cf cadd(cf a, cf b) { return mk(a.x + b.x, a.y + b.y); }
FL_HD cf csub(cf a, cf b) { return mk(a.x - b.x, a.y - b.y); }
FL_HD cf cmul(cf a, cf b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
FL_HD cf cfma(cf a, float s, cf c) { return mk(a.x * s + c.x, a.y * s + c.y); }
FL_HD cf cscale(cf a, float s) { return mk(a.x * s, a.y * s); }
#endif
FL_HD cf cconj(cf a) { return mk(a.x, -a.y); }
// multiply by  SGN * i
template <int SGN> FL_HD cf mul_i(cf a) { return SGN > 0 ? mk(-a.y, a.x) : mk(a.y, -a.x); }

FL_HD int bitrev4(int v) { return ((v & 1) << 3) | ((v & 2) << 1) | ((v & 4) >> 1) | ((v & 8) >> 3); }

// SGN = -1: forward kernel e^{-i..};  SGN = +1: inverse kernel e^{+i..}
template <int SGN>
FL_HD void dft3(cf& a0, cf& a1, cf& a2) {
  const float S3 = 0.8660254037844386f;
  const cf s = cadd(a1, a2), d = csub(a1, a2);
  const cf ya = cfma(s, -0.5f, a0);
  const cf t = mul_i<SGN>(cscale(d, S3));
  a0 = cadd(a0, s);
  a1 = cadd(ya, t);
  a2 = csub(ya, t);
}

template <int SGN>
FL_HD void dft5(cf& a0, cf& a1, cf& a2, cf& a3, cf& a4) {
  const float C1 = 0.30901699437494745f, C2 = -0.8090169943749473f;
  const float S1 = 0.9510565162951535f, S2 = 0.5877852522924732f;
  const cf s1 = cadd(a1, a4), d1 = csub(a1, a4), s2 = cadd(a2, a3), d2 = csub(a2, a3);
  const cf ya = cfma(s2, C2, cfma(s1, C1, a0));
  const cf yb = cfma(s2, C1, cfma(s1, C2, a0));
  const cf ta = mul_i<SGN>(cfma(d2, S2, cscale(d1, S1)));
  const cf tb = mul_i<SGN>(cfma(d2, -S1, cscale(d1, S2)));
  a0 = cadd(a0, cadd(s1, s2));
  a1 = cadd(ya, ta);
  a4 = csub(ya, ta);
  a2 = cadd(yb, tb);
  a3 = csub(yb, tb);
}

// w15^j = (cos, sin)(2 pi j / 15), j = 0..8
#define SFNO_W15_TABLE                                                                              \
  {1.0f, 0.0f}, {0.9135454576426009f, 0.40673664307580015f}, {0.6691306063588582f, 0.7431448254773941f},    \
  {0.30901699437494745f, 0.9510565162951535f}, {-0.10452846326765333f, 0.9945218953682734f},                \
  {-0.5f, 0.8660254037844387f}, {-0.8090169943749473f, 0.5877852522924732f},                                \
  {-0.9781476007338057f, 0.20791169081775931f}, {-0.9781476007338057f, -0.20791169081775907f}

// 15-point DFT, natural order in and out.  n = 3 n1 + n2,  k = k1 + 5 k2.
template <int SGN>
FL_HD void dft15(cf (&v)[15]) {
  const float w15[9][2] = {SFNO_W15_TABLE};
  cf t[3][5];
#pragma unroll
  for (int n2 = 0; n2 < 3; ++n2) {
#pragma unroll
    for (int n1 = 0; n1 < 5; ++n1) t[n2][n1] = v[3 * n1 + n2];
    dft5<SGN>(t[n2][0], t[n2][1], t[n2][2], t[n2][3], t[n2][4]);      // index is now k1
  }
#pragma unroll
  for (int n2 = 1; n2 < 3; ++n2) {
#pragma unroll
    for (int k1 = 1; k1 < 5; ++k1) {
      const int j = n2 * k1;                                           // <= 8
      t[n2][k1] = cmul(t[n2][k1], mk(w15[j][0], SGN * w15[j][1]));
    }
  }
#pragma unroll
  for (int k1 = 0; k1 < 5; ++k1) {
    dft3<SGN>(t[0][k1], t[1][k1], t[2][k1]);                           // index n2 is now k2
#pragma unroll
    for (int k2 = 0; k2 < 3; ++k2) v[k1 + 5 * k2] = t[k2][k1];
  }
}

// Twiddle tables every lane function takes (built once on the host in double precision):
//   tw240[b*15 + k1] = e^{-2 pi i b k1 / 240}      b in [0,16), k1 in [0,15)
//   tw480[k]         = e^{-2 pi i k / 480}          k in [0,241)
//   bf[s]            = e^{-2 pi i (lane & (d-1)) / (2d)},  d = 1 << s   (per lane, 4 entries)

// Per-lane butterfly constants for the 16-lane radix-2 stages (d = 1 << s):
//   sgn[s] = -1 on "upper" lanes (lane & d), +1 on lower lanes
//   fw[s]  = forward twiddle e^{-2 pi i (lane & (d-1)) / (2d)} on upper lanes, 1 on lower lanes
// so that both butterfly outputs are  (t + sgn*v) * fw  (forward, DIF)  and  t' + sgn*(v * conj(fw))  (inverse, DIT)
// without divergent code.  Stage 0 (d = 1) always has a unit twiddle and skips the multiply.
struct LaneConsts {
  float sgn[4];
  cf fw[4];
};

FL_HD LaneConsts make_lane_consts(int lane16, const cf (&bf)[4]) {
  LaneConsts c;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const bool upper = (lane16 & (1 << s)) != 0;
    c.sgn[s] = upper ? -1.f : 1.f;
    c.fw[s] = upper ? bf[s] : mk(1.f, 0.f);
  }
  return c;
}

// forward complex FFT-240 across 16 lanes.  in: v[a] = z[16 a + lane];  out: v[k1] = Z[k1 + 15 bitrev4(lane)]
template <class Warp>
FL_HD void fft240_fwd(Warp& w, cf (&v)[15], const cf* tw240_lane, const LaneConsts& lc) {
  dft15<-1>(v);
#pragma unroll
  for (int k1 = 1; k1 < 15; ++k1) v[k1] = cmul(v[k1], tw240_lane[k1]);
#pragma unroll
  for (int s = 3; s >= 0; --s) {
    const int d = 1 << s;
    const float sg = lc.sgn[s];
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      const cf t = mk(w.shfl_xor(v[k1].x, d), w.shfl_xor(v[k1].y, d));
      const cf u = mk(t.x + sg * v[k1].x, t.y + sg * v[k1].y);       // lower: v + t ; upper: t - v
      v[k1] = (s == 0) ? u : cmul(u, lc.fw[s]);
    }
  }
}

// inverse (unnormalised, e^{+}) complex FFT-240.  in: v[k1] = V[k1 + 15 bitrev4(lane)];  out: v[a] = z[16 a + lane]
template <class Warp>
FL_HD void fft240_inv(Warp& w, cf (&v)[15], const cf* tw240_lane, const LaneConsts& lc) {
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int d = 1 << s;
    const float sg = lc.sgn[s];
    const cf wc = cconj(lc.fw[s]);
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      const cf val = (s == 0) ? v[k1] : cmul(v[k1], wc);
      const cf t = mk(w.shfl_xor(val.x, d), w.shfl_xor(val.y, d));
      v[k1] = mk(t.x + sg * val.x, t.y + sg * val.y);                // lower: val + t ; upper: t - val
    }
  }
#pragma unroll
  for (int k1 = 1; k1 < 15; ++k1) v[k1] = cmul(v[k1], cconj(tw240_lane[k1]));
  dft15<+1>(v);
}

// real-input split.  in: Z[k1] = Z[k1 + 15 k2] (k2 = bitrev4(lane));  out: X[k1] = X[k1 + 15 k2] for the
// length-480 real sequence, and xnyq = X[240] (meaningful on lane 0 only).
template <class Warp>
FL_HD void rfft480_post(Warp& w, const cf (&Z)[15], cf (&X)[15], float& xnyq, const cf* tw480) {
  const int p = w.lane16();
  const int k2 = bitrev4(p);
#pragma unroll
  for (int k1 = 0; k1 < 15; ++k1) {
    cf zp;
    if (k1 == 0) {
      const int src = bitrev4((16 - k2) & 15);
      zp = mk(w.shfl_idx(Z[0].x, src), w.shfl_idx(Z[0].y, src));
    } else {
      zp = mk(w.shfl_xor(Z[15 - k1].x, 15), w.shfl_xor(Z[15 - k1].y, 15));
    }
    const cf zc = cconj(zp);
    const cf e = cscale(cadd(Z[k1], zc), 0.5f);
    const cf d = cscale(csub(Z[k1], zc), 0.5f);
    const cf wd = cmul(tw480[k1 + 15 * k2], d);
    X[k1] = mk(e.x + wd.y, e.y - wd.x);          // E - i w D
  }
  xnyq = Z[0].x - Z[0].y;
}

// inverse of the split: from the Hermitian half-spectrum H[k], H[240-k] of a length-480 real sequence build
// V[k] = 2 Z[k] such that  x[2n] + i x[2n+1] = sum_k V[k] e^{+2 pi i k n / 240}.
FL_HD cf irfft480_pre(cf h, cf hp, cf tw480_k) {
  const cf hc = cconj(hp);
  const cf a = cadd(h, hc);
  const cf b = csub(h, hc);
  const cf wb = cmul(cconj(tw480_k), b);
  return mk(a.x - wb.y, a.y + wb.x);             // A + i conj(w) B
}


// ---------------------------------------------------------------------------------------------
// Row-level composition for nlon = 480 * R (R = 1: the 240x480 Gauss grid, R = 3: 721x1440).
//   forward, R = 3: x_r[q] = x[3q + r] (three length-480 real FFTs), X[m] = sum_r w1440^(m r) X_r[m], m <= 240
//   inverse, R = 3: x[3q + r] = irfft480( X[m] conj(w1440)^(m r) ), with bin 240 doubled (it is not Nyquist)
//   tw1440[m] = e^{-2 pi i m / 1440}, m in [0, 241)
// `loadz(n, r)` returns the packed complex sample (x[2 R n + r], x[2 R n + R + r]) of the row;
// `storez(n, r, z)` writes it back.
template <int R, class Warp, class LoadZ>
FL_HD void rfft_row_fwd(Warp& w, LoadZ loadz, cf (&X)[15], cf& x240, const cf* tw240, const cf* tw480,
                        const cf* tw1440, const LaneConsts& lc) {
  const int b = w.lane16();
  const int k2 = bitrev4(b);
#pragma unroll
  for (int k1 = 0; k1 < 15; ++k1) X[k1] = mk(0.f, 0.f);
  x240 = mk(0.f, 0.f);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    cf v[15];
#pragma unroll
    for (int a = 0; a < 15; ++a) v[a] = loadz(16 * a + b, r);
    fft240_fwd(w, v, tw240 + b * 15, lc);
    cf xr[15];
    float nyq;
    rfft480_post(w, v, xr, nyq, tw480);
    if (R == 1) {
#pragma unroll
      for (int k1 = 0; k1 < 15; ++k1) X[k1] = xr[k1];
      x240 = mk(nyq, 0.f);
    } else {
#pragma unroll
      for (int k1 = 0; k1 < 15; ++k1) {
        const cf t1 = tw1440[k1 + 15 * k2];
        const cf t = (r == 0) ? mk(1.f, 0.f) : (r == 1 ? t1 : cmul(t1, t1));
        X[k1] = cadd(X[k1], cmul(t, xr[k1]));
      }
      const cf e1 = tw1440[240];
      const cf e = (r == 0) ? mk(1.f, 0.f) : (r == 1 ? e1 : cmul(e1, e1));
      x240 = cadd(x240, cscale(e, nyq));
    }
  }
}

// H[k1] / HP[k1] hold the (already scaled) spectrum at bins k = k1 + 15 k2 and 240 - k (zero for bins not held).
template <int R, class Warp, class StoreZ>
FL_HD void irfft_row(Warp& w, const cf (&H)[15], const cf (&HP)[15], StoreZ storez, const cf* tw240,
                     const cf* tw480, const cf* tw1440, const LaneConsts& lc) {
  const int b = w.lane16();
  const int k2 = bitrev4(b);
#pragma unroll
  for (int r = 0; r < R; ++r) {
    cf v[15];
#pragma unroll
    for (int k1 = 0; k1 < 15; ++k1) {
      const int k = k1 + 15 * k2;
      const int kp = 240 - k;
      cf h = H[k1], hp = HP[k1];
      if (R != 1 && r > 0) {
        const cf t1 = cconj(tw1440[k]), u1 = cconj(tw1440[kp]);
        h = cmul(h, r == 1 ? t1 : cmul(t1, t1));
        hp = cmul(hp, r == 1 ? u1 : cmul(u1, u1));
      }
      if (k == 0) {               // H[0] is real; partner is bin 240: true Nyquist for R = 1, an ordinary (doubled) bin else
        h = mk(h.x, 0.f);
        hp = mk((R == 1 ? 1.f : 2.f) * hp.x, 0.f);
      }
      v[k1] = irfft480_pre(h, hp, tw480[k]);
    }
    fft240_inv(w, v, tw240 + b * 15, lc);
#pragma unroll
    for (int a = 0; a < 15; ++a) storez(16 * a + b, r, v[a]);
  }
}

}  // namespace fl
}  // namespace sfno
