// Small in-register FFTs and the two-pass (R1 x 16) shared-memory FFT used by the STFT / iSTFT kernels.
#pragma once
#include "common.cuh"

namespace ssb {

__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cconj(float2 a) { return make_float2(a.x, -a.y); }
__device__ __forceinline__ float2 mul_mi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)

// forward DFT (e^{-i...}), natural order in / natural order out
__device__ __forceinline__ void fft4(float2& x0, float2& x1, float2& x2, float2& x3) {
  float2 t0 = cadd(x0, x2), t1 = csub(x0, x2), t2 = cadd(x1, x3), t3 = mul_mi(csub(x1, x3));
  x0 = cadd(t0, t2);
  x2 = csub(t0, t2);
  x1 = cadd(t1, t3);
  x3 = csub(t1, t3);
}

__device__ __forceinline__ void fft8(float2* v) {
  const float c = 0.70710678118654752440f;
  float2 e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
  float2 o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
  fft4(e0, e1, e2, e3);
  fft4(o0, o1, o2, o3);
  float2 w1 = make_float2(c * (o1.x + o1.y), c * (o1.y - o1.x));    // o1 * (c, -c)
  float2 w2 = mul_mi(o2);                                           // o2 * (0, -1)
  float2 w3 = make_float2(c * (o3.y - o3.x), -c * (o3.x + o3.y));   // o3 * (-c, -c)
  v[0] = cadd(e0, o0);
  v[4] = csub(e0, o0);
  v[1] = cadd(e1, w1);
  v[5] = csub(e1, w1);
  v[2] = cadd(e2, w2);
  v[6] = csub(e2, w2);
  v[3] = cadd(e3, w3);
  v[7] = csub(e3, w3);
}

__device__ __forceinline__ void fft16(float2* v) {
  const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f, c2 = 0.70710678118654752440f;
  float2 e[8], o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    e[i] = v[2 * i];
    o[i] = v[2 * i + 1];
  }
  fft8(e);
  fft8(o);
  // W16^k = (cos(pi k/8), -sin(pi k/8))
  o[1] = cmul(o[1], make_float2(c1, -s1));
  o[2] = make_float2(c2 * (o[2].x + o[2].y), c2 * (o[2].y - o[2].x));
  o[3] = cmul(o[3], make_float2(s1, -c1));
  o[4] = mul_mi(o[4]);
  o[5] = cmul(o[5], make_float2(-s1, -c1));
  o[6] = make_float2(c2 * (o[6].y - o[6].x), -c2 * (o[6].x + o[6].y));
  o[7] = cmul(o[7], make_float2(-c1, -s1));
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    v[k] = cadd(e[k], o[k]);
    v[k + 8] = csub(e[k], o[k]);
  }
}

// ---- two transforms at once on the packed fp32x2 pipe (FADD2 / FMUL2 / FFMA2 of sm_100a) -------------------------
// A p2 holds the same scalar of two independent transforms (lo = transform A, hi = transform B); every scalar
// operation of the FFT becomes one packed instruction.  ptxas folds neg2() into the operand modifiers.
typedef unsigned long long p2;
struct c2 { p2 re, im; };

__device__ __forceinline__ p2 pk(float lo, float hi) {
  p2 r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void upk(p2 a, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a)); }
__device__ __forceinline__ p2 dup2(float x) { return pk(x, x); }
__device__ __forceinline__ p2 add2(p2 a, p2 b) {
  p2 r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ p2 sub2(p2 a, p2 b) {
  p2 r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ p2 mul2(p2 a, p2 b) {
  p2 r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ p2 fma2(p2 a, p2 b, p2 c) {
  p2 r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ p2 neg2(p2 a) {
  float lo, hi;
  upk(a, lo, hi);
  return pk(-lo, -hi);
}
__device__ __forceinline__ c2 cadd(c2 a, c2 b) { return {add2(a.re, b.re), add2(a.im, b.im)}; }
__device__ __forceinline__ c2 csub(c2 a, c2 b) { return {sub2(a.re, b.re), sub2(a.im, b.im)}; }
// a * (wr + i wi), both components of w duplicated in their p2
__device__ __forceinline__ c2 cmul(c2 a, p2 wr, p2 wi) {
  return {fma2(a.re, wr, neg2(mul2(a.im, wi))), fma2(a.re, wi, mul2(a.im, wr))};
}
// one per-thread scalar factor for both transforms: two scalar ops cost the same two issue slots as a packed op and
// the factor does not have to be duplicated in a table
__device__ __forceinline__ p2 scale2(p2 a, float w) {
  float lo, hi;
  upk(a, lo, hi);
  return pk(lo * w, hi * w);
}
__device__ __forceinline__ c2 cmul_s(c2 a, float wr, float wi) {
  float ra, rb, ia, ib;
  upk(a.re, ra, rb);
  upk(a.im, ia, ib);
  return {pk(fmaf(ra, wr, -ia * wi), fmaf(rb, wr, -ib * wi)), pk(fmaf(ra, wi, ia * wr), fmaf(rb, wi, ib * wr))};
}
// (t + (-i) d, t - (-i) d) without materialising -i d
__device__ __forceinline__ void bfly_mi(c2 t, c2 d, c2& p, c2& m) {
  p = {add2(t.re, d.im), sub2(t.im, d.re)};
  m = {sub2(t.re, d.im), add2(t.im, d.re)};
}

__device__ __forceinline__ void fft4(c2& x0, c2& x1, c2& x2, c2& x3) {
  const c2 t0 = cadd(x0, x2), t1 = csub(x0, x2), t2 = cadd(x1, x3), d = csub(x1, x3);
  x0 = cadd(t0, t2);
  x2 = csub(t0, t2);
  bfly_mi(t1, d, x1, x3);
}

__device__ __forceinline__ void fft8(c2* v) {
  const p2 c = dup2(0.70710678118654752440f);
  c2 e0 = v[0], e1 = v[2], e2 = v[4], e3 = v[6];
  c2 o0 = v[1], o1 = v[3], o2 = v[5], o3 = v[7];
  fft4(e0, e1, e2, e3);
  fft4(o0, o1, o2, o3);
  const c2 w1 = {mul2(c, add2(o1.re, o1.im)), mul2(c, sub2(o1.im, o1.re))};          // o1 * (c, -c)
  const c2 w3 = {mul2(c, sub2(o3.im, o3.re)), neg2(mul2(c, add2(o3.re, o3.im)))};    // o3 * (-c, -c)
  v[0] = cadd(e0, o0);
  v[4] = csub(e0, o0);
  v[1] = cadd(e1, w1);
  v[5] = csub(e1, w1);
  bfly_mi(e2, o2, v[2], v[6]);                                                       // o2 * (0, -1)
  v[3] = cadd(e3, w3);
  v[7] = csub(e3, w3);
}

__device__ __forceinline__ void fft16(c2* v) {
  const p2 c1 = dup2(0.92387953251128675613f), s1 = dup2(0.38268343236508977173f), c2_ = dup2(0.70710678118654752440f);
  c2 e[8], o[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    e[i] = v[2 * i];
    o[i] = v[2 * i + 1];
  }
  fft8(e);
  fft8(o);
  o[1] = cmul(o[1], c1, neg2(s1));
  o[2] = {mul2(c2_, add2(o[2].re, o[2].im)), mul2(c2_, sub2(o[2].im, o[2].re))};
  o[3] = cmul(o[3], s1, neg2(c1));
  o[5] = cmul(o[5], neg2(s1), neg2(c1));
  o[6] = {mul2(c2_, sub2(o[6].im, o[6].re)), neg2(mul2(c2_, add2(o[6].re, o[6].im)))};
  o[7] = cmul(o[7], neg2(c1), neg2(s1));
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k == 4) {
      bfly_mi(e[4], o[4], v[4], v[12]);
    } else {
      v[k] = cadd(e[k], o[k]);
      v[k + 8] = csub(e[k], o[k]);
    }
  }
}

template <int R>
__device__ __forceinline__ void fft_small(float2* v) {
  static_assert(R == 8 || R == 16, "radix");
  if (R == 8) fft8(v); else fft16(v);
}

// Complex N-point forward FFT by 16 cooperating threads (a half warp): N = R1 * 16.
//   tw[j] = exp(-2*pi*i*j/(2N)), j in [0, 2N)  (shared table of the real transform of size 2N)
//   in  : v[n1] = z[16*n1 + lane16]            (R1 values per thread)
//   out : Z[k] natural order in `scr` (N float2; needs N + R1 float2 of scratch for the padded pass)
// All 16 threads of the group must call; uses __syncwarp().
template <int N>
__device__ __forceinline__ void fft_group16(float2* v, float2* scr, const float2* tw, int lane16) {
  constexpr int R1 = N / 16;
  constexpr int PITCH = 17;
  fft_small<R1>(v);
#pragma unroll
  for (int k1 = 0; k1 < R1; ++k1) {
    float2 t = tw[2 * lane16 * k1];          // W_N^(n2*k1)
    scr[k1 * PITCH + lane16] = (k1 == 0) ? v[0] : cmul(v[k1], t);
  }
  __syncwarp();
  float2 u[16];
  if (lane16 < R1) {
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) u[n2] = scr[lane16 * PITCH + n2];
  }
  __syncwarp();
  if (lane16 < R1) {
    fft16(u);
#pragma unroll
    for (int k2 = 0; k2 < 16; ++k2) scr[lane16 + R1 * k2] = u[k2];
  }
  __syncwarp();
}

}  // namespace ssb
