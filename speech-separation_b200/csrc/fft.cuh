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
