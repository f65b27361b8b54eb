// K3/K4 glue: BatchNorm1d statistics over all (t,b) rows -- padding included, as the reference does
// (archs/uPIT.py:119,138) -- folded into the head Linear (archs/uPIT.py:117,141) so that the head is one GEMM,
// and the matching backward reductions.  x is the layer-2 BLSTM output as sb16 [R = T*B][C = 2H].
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int BN_ROWS = 256;

// partial[blk][0][c] = sum_r x, partial[blk][1][c] = sum_r x^2   (if dy == null)
// partial[blk][0][c] = sum_r dy, partial[blk][1][c] = sum_r dy*x (if dy != null)
__global__ void __launch_bounds__(256) bn_partial_kernel(const __nv_bfloat16* __restrict__ x, long long ld,
                                                         long long plane, const float* __restrict__ dy, int R, int C,
                                                         float* __restrict__ partial) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  const int r0 = blockIdx.y * BN_ROWS, r1 = min(R, r0 + BN_ROWS);
  float a = 0.0f, b = 0.0f;
#pragma unroll 8
  for (int r = r0; r < r1; ++r) {                              // 8 rows of loads in flight per thread
    const long long o = (long long)r * ld + c;
    const float v = join_bf16(x[o], x[o + plane]);
    if (dy) {
      const float g = dy[(long long)r * C + c];
      a += g;
      b += g * v;
    } else {
      a += v;
      b += v * v;
    }
  }
  partial[((long long)blockIdx.y * 2 + 0) * C + c] = a;
  partial[((long long)blockIdx.y * 2 + 1) * C + c] = b;
}

// coef layout (5 x C): [0] scale s = gamma*rstd, [1] shift = beta - mean*s, [2] mean, [3] rstd, [4] var (biased)
// nt (optional, data-parallel mode B): the row count is nt[0] * nt[1] (total batch x longest T over all ranks) instead of R
__global__ void __launch_bounds__(256) bn_stats_final_kernel(const float* __restrict__ partial, int parts, int R_, int C,
                                                             const float* __restrict__ nt,
                                                             const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, float eps, float momentum,
                                                             int training, float* __restrict__ running_mean,
                                                             float* __restrict__ running_var, float* __restrict__ coef) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  const double R = nt ? (double)nt[0] * (double)nt[1] : (double)R_;
  double mean, var;
  if (training) {
    double s = 0.0, s2 = 0.0;
    for (int p = 0; p < parts; ++p) {
      s += (double)partial[((long long)p * 2 + 0) * C + c];
      s2 += (double)partial[((long long)p * 2 + 1) * C + c];
    }
    mean = s / R;
    var = s2 / R - mean * mean;
    if (var < 0.0) var = 0.0;
    const double unb = R > 1 ? var * (R / (R - 1.0)) : var;
    running_mean[c] = (float)((1.0 - momentum) * running_mean[c] + momentum * mean);
    running_var[c] = (float)((1.0 - momentum) * running_var[c] + momentum * unb);
  } else {
    mean = running_mean[c];
    var = running_var[c];
  }
  const double rstd = 1.0 / sqrt(var + (double)eps);
  const double s = gamma[c] * rstd;
  coef[0 * C + c] = (float)s;
  coef[1 * C + c] = (float)(beta[c] - mean * s);
  coef[2 * C + c] = (float)mean;
  coef[3 * C + c] = (float)rstd;
  coef[4 * C + c] = (float)var;
}

// b'[o] = b[o] + sum_c W[o][c] * shift[c]      (one warp per output row)
__global__ void __launch_bounds__(256) head_fold_bias_kernel(const float* __restrict__ W, const float* __restrict__ b,
                                                             const float* __restrict__ shift, int O, int C,
                                                             float* __restrict__ bout) {
  const int o = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (o >= O) return;
  float acc = 0.0f;
  for (int c = lane; c < C; c += 32) acc += W[(long long)o * C + c] * shift[c];
  acc = warp_sum(acc);
  if (lane == 0) bout[o] = b[o] + acc;
}

// from partial sums of dy and dy*x: dgamma, dbeta and the coefficients of dx = s*dy + bc*(x - mean) + cc
// bwd layout (2 x C): [0] bc, [1] cc
// glob / nt (optional, data-parallel mode B): dgamma, dbeta come from this rank's sums (the gradient exchange adds the
// ranks up), the coefficients of dx from the sums over ALL ranks (glob[2][C]) and the global row count nt[0] * nt[1]
__global__ void __launch_bounds__(256) bn_bwd_final_kernel(const float* __restrict__ partial, int parts, int R_, int C,
                                                           const float* __restrict__ glob, const float* __restrict__ nt,
                                                           const float* __restrict__ coef, int training,
                                                           float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                           float* __restrict__ bwd) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  double a = 0.0, bx = 0.0;
  for (int p = 0; p < parts; ++p) {
    a += (double)partial[((long long)p * 2 + 0) * C + c];
    bx += (double)partial[((long long)p * 2 + 1) * C + c];
  }
  const double s = coef[0 * C + c], mean = coef[2 * C + c], rstd = coef[3 * C + c];
  double dg = (bx - mean * a) * rstd;            // sum dy * xhat
  dgamma[c] = (float)dg;
  dbeta[c] = (float)a;
  double R = R_;
  if (glob) {
    a = glob[c];
    dg = ((double)glob[C + c] - mean * a) * rstd;
    R = (double)nt[0] * (double)nt[1];
  }
  if (training) {
    bwd[0 * C + c] = (float)(-s * rstd * dg / R);
    bwd[1 * C + c] = (float)(-s * a / R);
  } else {
    bwd[0 * C + c] = 0.0f;
    bwd[1 * C + c] = 0.0f;
  }
}

// in place: dy[r][c] <- s*dy + bc*(x - mean) + cc
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(float* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                           long long ld, long long plane, long long n, int C,
                                                           const float* __restrict__ coef,
                                                           const float* __restrict__ bwd) {
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
    const long long r = e / C;
    const int c = (int)(e - r * C);
    const long long o = r * ld + c;
    const float v = join_bf16(x[o], x[o + plane]);
    dy[e] = coef[c] * dy[e] + bwd[c] * (v - coef[2 * C + c]) + bwd[C + c];
  }
}

// dW[o][c] = s_c * (G[o][c] - mean_c * dsum[o]) + beta_c * dsum[o]   with G = dz^T x
// the same on four consecutive columns per thread (C, ld, plane multiples of 4): 16-byte dy accesses, 8-byte bf16 loads
__global__ void __launch_bounds__(256) bn_bwd_apply4_kernel(float* __restrict__ dy, const __nv_bfloat16* __restrict__ x,
                                                            long long ld, long long plane, unsigned int nq, int C4,
                                                            const float* __restrict__ coef,
                                                            const float* __restrict__ bwd) {
  const int C = 4 * C4;
  for (unsigned int q = blockIdx.x * 256u + threadIdx.x; q < nq; q += gridDim.x * 256u) {
    const unsigned int r = q / (unsigned int)C4;
    const int c = 4 * (int)(q - r * (unsigned int)C4);
    const long long o = (long long)r * ld + c;
    const uint2 h = *reinterpret_cast<const uint2*>(x + o), l = *reinterpret_cast<const uint2*>(x + o + plane);
    float4 g = reinterpret_cast<float4*>(dy)[q];
    const float4 s = *reinterpret_cast<const float4*>(coef + c), m = *reinterpret_cast<const float4*>(coef + 2 * C + c);
    const float4 b0 = *reinterpret_cast<const float4*>(bwd + c), b1 = *reinterpret_cast<const float4*>(bwd + C + c);
    const float v0 = __uint_as_float(h.x << 16) + __uint_as_float(l.x << 16);
    const float v1 = __uint_as_float(h.x & 0xffff0000u) + __uint_as_float(l.x & 0xffff0000u);
    const float v2 = __uint_as_float(h.y << 16) + __uint_as_float(l.y << 16);
    const float v3 = __uint_as_float(h.y & 0xffff0000u) + __uint_as_float(l.y & 0xffff0000u);
    g.x = s.x * g.x + b0.x * (v0 - m.x) + b1.x;
    g.y = s.y * g.y + b0.y * (v1 - m.y) + b1.y;
    g.z = s.z * g.z + b0.z * (v2 - m.z) + b1.z;
    g.w = s.w * g.w + b0.w * (v3 - m.w) + b1.w;
    reinterpret_cast<float4*>(dy)[q] = g;
  }
}

__global__ void __launch_bounds__(256) head_dw_fix_kernel(const float* __restrict__ G, const float* __restrict__ dsum,
                                                          const float* __restrict__ coef,
                                                          const float* __restrict__ beta, int O, int C,
                                                          float* __restrict__ dW) {
  const long long n = (long long)O * C;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
    const int o = (int)(e / C), c = (int)(e - (long long)o * C);
    dW[e] = coef[c] * (G[e] - coef[2 * C + c] * dsum[o]) + beta[c] * dsum[o];
  }
}

// sums[0][c], sums[1][c] = the two column sums of bn_partial_kernel over all row blocks
__global__ void __launch_bounds__(256) bn_reduce_kernel(const float* __restrict__ partial, int parts, int C,
                                                        float* __restrict__ sums) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= C) return;
  double a = 0.0, b = 0.0;
  for (int p = 0; p < parts; ++p) {
    a += (double)partial[((long long)p * 2 + 0) * C + c];
    b += (double)partial[((long long)p * 2 + 1) * C + c];
  }
  sums[c] = (float)a;
  sums[C + c] = (float)b;
}

}  // namespace ssb

extern "C" int ss_bn_parts(int R) { return ssb::cdiv(R, ssb::BN_ROWS); }

extern "C" int ss_bn_stats(const void* x, long long ld, long long plane, int R, int C, const float* gamma,
                           const float* beta, float eps, float momentum, int training, float* running_mean,
                           float* running_var, float* partial, float* coef, void* stream) {
  using namespace ssb;
  cudaStream_t st = (cudaStream_t)stream;
  const int parts = cdiv(R, BN_ROWS);
  if (training)
    bn_partial_kernel<<<dim3(cdiv(C, 256), parts), 256, 0, st>>>((const __nv_bfloat16*)x, ld, plane, nullptr, R, C, partial);
  bn_stats_final_kernel<<<cdiv(C, 256), 256, 0, st>>>(partial, parts, R, C, nullptr, gamma, beta, eps, momentum, training,
                                                      running_mean, running_var, coef);
  return check_launch("ss_bn_stats");
}

// ---- data-parallel mode B: the same statistics over the rows of ALL ranks.  The caller sum-allreduces
// `sums` (and the two counts) between ss_bn_sums and the *_from_sums call.
extern "C" int ss_bn_sums(const void* x, long long ld, long long plane, const float* dy, int R, int C, float* partial,
                          float* sums, void* stream) {
  using namespace ssb;
  cudaStream_t st = (cudaStream_t)stream;
  const int parts = cdiv(R, BN_ROWS);
  bn_partial_kernel<<<dim3(cdiv(C, 256), parts), 256, 0, st>>>((const __nv_bfloat16*)x, ld, plane, dy, R, C, partial);
  bn_reduce_kernel<<<cdiv(C, 256), 256, 0, st>>>(partial, parts, C, sums);
  return check_launch("ss_bn_sums");
}

extern "C" int ss_bn_stats_from_sums(const float* sums, const float* counts, int C, const float* gamma, const float* beta,
                                     float eps, float momentum, float* running_mean, float* running_var, float* coef,
                                     void* stream) {
  using namespace ssb;
  bn_stats_final_kernel<<<cdiv(C, 256), 256, 0, (cudaStream_t)stream>>>(sums, 1, 0, C, counts, gamma, beta, eps, momentum, 1,
                                                                        running_mean, running_var, coef);
  return check_launch("ss_bn_stats_from_sums");
}

extern "C" int ss_bn_bwd_from_sums(float* dy, const void* x, long long ld, long long plane, int R, int C, const float* coef,
                                   const float* sums_local, const float* sums_global, const float* counts, float* dgamma,
                                   float* dbeta, float* bwd, void* stream) {
  using namespace ssb;
  cudaStream_t st = (cudaStream_t)stream;
  bn_bwd_final_kernel<<<cdiv(C, 256), 256, 0, st>>>(sums_local, 1, R, C, sums_global, counts, coef, 1, dgamma, dbeta, bwd);
  const long long n = (long long)R * C;
  bn_bwd_apply_kernel<<<(int)min((long long)ssb::sm_count() * 8, (n + 255) / 256), 256, 0, st>>>(dy, (const __nv_bfloat16*)x, ld,
                                                                                      plane, n, C, coef, bwd);
  return check_launch("ss_bn_bwd_from_sums");
}

extern "C" int ss_head_fold_bias(const float* W, const float* b, const float* coef, int O, int C, float* bout,
                                 void* stream) {
  using namespace ssb;
  head_fold_bias_kernel<<<cdiv(O, 8), 256, 0, (cudaStream_t)stream>>>(W, b, coef + C, O, C, bout);
  return check_launch("ss_head_fold_bias");
}

extern "C" int ss_bn_bwd(float* dy, const void* x, long long ld, long long plane, int R, int C, const float* coef,
                         int training, float* partial, float* dgamma, float* dbeta, float* bwd, void* stream) {
  using namespace ssb;
  cudaStream_t st = (cudaStream_t)stream;
  const int parts = cdiv(R, BN_ROWS);
  bn_partial_kernel<<<dim3(cdiv(C, 256), parts), 256, 0, st>>>((const __nv_bfloat16*)x, ld, plane, dy, R, C, partial);
  bn_bwd_final_kernel<<<cdiv(C, 256), 256, 0, st>>>(partial, parts, R, C, nullptr, nullptr, coef, training, dgamma, dbeta, bwd);
  const long long n = (long long)R * C;
  const bool vec4 = C % 4 == 0 && ld % 4 == 0 && plane % 4 == 0 && n / 4 < 0xffffffffll &&
                    ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(coef) | reinterpret_cast<uintptr_t>(bwd)) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(x) & 7) == 0;
  if (vec4)
    bn_bwd_apply4_kernel<<<(int)min((long long)ssb::sm_count() * 8, (n / 4 + 255) / 256), 256, 0, st>>>(
        dy, (const __nv_bfloat16*)x, ld, plane, (unsigned int)(n / 4), C / 4, coef, bwd);
  else
    bn_bwd_apply_kernel<<<(int)min((long long)ssb::sm_count() * 8, (n + 255) / 256), 256, 0, st>>>(dy, (const __nv_bfloat16*)x, ld,
                                                                                      plane, n, C, coef, bwd);
  return check_launch("ss_bn_bwd");
}

extern "C" int ss_head_dw_fix(const float* G, const float* dsum, const float* coef, const float* beta, int O, int C,
                              float* dW, void* stream) {
  using namespace ssb;
  const long long n = (long long)O * C;
  head_dw_fix_kernel<<<(int)min((long long)ssb::sm_count() * 8, (n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(G, dsum, coef, beta,
                                                                                                   O, C, dW);
  return check_launch("ss_head_dw_fix");
}
