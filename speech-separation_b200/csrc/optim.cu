// f1: global-norm gradient clip + Adam over ONE flat fp32 bucket.
// Replaces torch.nn.utils.clip_grad_norm_(params, 0.25) + torch.optim.Adam.step() (steps/train_qsub.py:95,121-122),
// which launch ~6 kernels per parameter tensor (20 tensors).  HBM-bound: reads g twice, p/m/v once, writes p/m/v.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int OPT_THREADS = 256;
constexpr int OPT_PER_BLOCK = OPT_THREADS * 4 * 8;   // 8 float4 per thread

__global__ void __launch_bounds__(OPT_THREADS) sumsq_partial_kernel(const float* __restrict__ g, long long n,
                                                                     float* __restrict__ partial) {
  const long long base = (long long)blockIdx.x * OPT_PER_BLOCK;
  float acc = 0.0f;
  if (base + OPT_PER_BLOCK <= n && ((uintptr_t)g & 15) == 0) {
    const float4* g4 = reinterpret_cast<const float4*>(g + base);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 v = g4[i * OPT_THREADS + threadIdx.x];
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
  } else {
    for (long long e = base + threadIdx.x; e < min(n, base + OPT_PER_BLOCK); e += OPT_THREADS) acc += g[e] * g[e];
  }
  __shared__ float red[OPT_THREADS / 32];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < OPT_THREADS / 32 ? red[threadIdx.x] : 0.0f;
    v = warp_sum(v);
    if (threadIdx.x == 0) partial[blockIdx.x] = v;
  }
}

// norm_out[0] = gscale * sqrt(sum partial)  (fp64 accumulation of the per-block partials)
__global__ void __launch_bounds__(256) sumsq_final_kernel(const float* __restrict__ partial, int parts, float gscale,
                                                           float* __restrict__ norm_out) {
  double acc = 0.0;
  for (int i = threadIdx.x; i < parts; i += 256) acc += (double)partial[i];
  __shared__ double red[256];
  red[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) {
    if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
    __syncthreads();
  }
  if (threadIdx.x == 0) norm_out[0] = gscale * (float)sqrt(red[0]);
}

struct AdamArgs {
  float gscale, max_norm, lr, beta1, beta2, eps, bc1, bc2_sqrt;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float coef, const AdamArgs& a) {
  g *= coef;
  m = m + (g - m) * (1.0f - a.beta1);                 // exp_avg.lerp_(grad, 1-beta1)
  v = v * a.beta2 + (1.0f - a.beta2) * g * g;         // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1-beta2)
  const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
  p = p - (a.lr / a.bc1) * (m / denom);
}

__global__ void __launch_bounds__(OPT_THREADS) clip_adam_kernel(float* __restrict__ p, const float* __restrict__ g,
                                                                 float* __restrict__ m, float* __restrict__ v,
                                                                 long long n, const float* __restrict__ norm,
                                                                 AdamArgs a) {
  // clip_grad_norm_: coef = max_norm / (total_norm + 1e-6), clamped to 1
  float coef = a.gscale;
  if (a.max_norm > 0.0f) coef *= fminf(1.0f, a.max_norm / (norm[0] + 1e-6f));
  const long long n4 = (((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) & 15) == 0 ? n / 4 : 0;
  const long long stride = (long long)gridDim.x * OPT_THREADS;
  for (long long i = (long long)blockIdx.x * OPT_THREADS + threadIdx.x; i < n4; i += stride) {
    float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    adam_one(pp.x, gg.x, mm.x, vv.x, coef, a);
    adam_one(pp.y, gg.y, mm.y, vv.y, coef, a);
    adam_one(pp.z, gg.z, mm.z, vv.z, coef, a);
    adam_one(pp.w, gg.w, mm.w, vv.w, coef, a);
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  for (long long e = n4 * 4 + (long long)blockIdx.x * OPT_THREADS + threadIdx.x; e < n; e += stride)
    adam_one(p[e], g[e], m[e], v[e], coef, a);
}

// ---- scratch of ss_grad_norm_noalloc: per-block partial sums of squares in a file-scope device array, so that the
// entry point needs no caller workspace and allocates nothing.  Every GPU that loads the library holds these 32 KB
// until exit.  In HOST code the name `g_norm_scratch` is only a stand-in for the symbol, not a device address:
// noalloc_scratch() resolves the address with cudaGetSymbolAddress once per device, with that device current.
constexpr long long NOALLOC_MAX_N = SS_GRAD_NORM_NOALLOC_MAX_N;
constexpr int NOALLOC_MAX_PARTS = (int)((NOALLOC_MAX_N + OPT_PER_BLOCK - 1) / OPT_PER_BLOCK);   // 8192 floats
static __device__ __align__(256) float g_norm_scratch[NOALLOC_MAX_PARTS];

static float* noalloc_scratch() {
  static float* addr[64] = {nullptr};   // per device; a racing first call resolves the same address twice
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    set_error("ss_grad_norm_noalloc: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  float* p = __atomic_load_n(&addr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    void* sym = nullptr;
    e = cudaGetSymbolAddress(&sym, g_norm_scratch);
    if (e != cudaSuccess || !sym) {
      set_error("ss_grad_norm_noalloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
      (void)cudaGetLastError();
      return nullptr;
    }
    p = static_cast<float*>(sym);
    __atomic_store_n(&addr[dev], p, __ATOMIC_RELEASE);
  }
  return p;
}

}  // namespace ssb

extern "C" int ss_optim_parts(long long n) { return ssb::cdiv(n, ssb::OPT_PER_BLOCK); }

extern "C" int ss_grad_norm(const float* g, long long n, float gscale, float* partial, float* norm_out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n >= 0 && norm_out, "ss_grad_norm: bad arguments");
  const int parts = cdiv(n, OPT_PER_BLOCK);
  cudaStream_t st = (cudaStream_t)stream;
  if (parts > 0) sumsq_partial_kernel<<<parts, OPT_THREADS, 0, st>>>(g, n, partial);
  sumsq_final_kernel<<<1, 256, 0, st>>>(partial, parts, gscale, norm_out);
  return check_launch("ss_grad_norm");
}

extern "C" long long ss_grad_norm_noalloc_max_n(void) { return ssb::NOALLOC_MAX_N; }

// ss_grad_norm without a caller workspace and without any allocation: the same two kernels, the partial sums in the
// library's static device array.  All calls on a device share that array: every call overwrites the `parts` entries
// it reads (nothing has to start at zero, so nothing is cleared), and calls that overlap on the GPU (different
// streams, no ordering between them) race.
extern "C" int ss_grad_norm_noalloc(const float* g, long long n, float gscale, float* norm_out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n >= 0 && norm_out && (g || n == 0), "ss_grad_norm_noalloc: bad arguments");
  SS_REQUIRE(n <= NOALLOC_MAX_N,
             "ss_grad_norm_noalloc: n = %lld exceeds the static scratch (max %lld elements); use ss_grad_norm with a "
             "caller workspace of ss_optim_parts(n) floats", n, (long long)NOALLOC_MAX_N);
  float* partial = noalloc_scratch();
  if (!partial) return -3;
  const int parts = cdiv(n, OPT_PER_BLOCK);
  cudaStream_t st = (cudaStream_t)stream;
  if (parts > 0) sumsq_partial_kernel<<<parts, OPT_THREADS, 0, st>>>(g, n, partial);
  sumsq_final_kernel<<<1, 256, 0, st>>>(partial, parts, gscale, norm_out);
  return check_launch("ss_grad_norm_noalloc");
}

extern "C" int ss_clip_adam(float* p, const float* g, float* m, float* v, long long n, const float* norm, float gscale,
                            float max_norm, float lr, float beta1, float beta2, float eps, int step, void* stream) {
  using namespace ssb;
  SS_REQUIRE(step >= 1, "ss_clip_adam: step counts from 1");
  SS_REQUIRE(max_norm <= 0.0f || norm, "ss_clip_adam: clipping needs the device norm");
  if (n == 0) return 0;
  AdamArgs a;
  a.gscale = gscale, a.max_norm = max_norm, a.lr = lr, a.beta1 = beta1, a.beta2 = beta2, a.eps = eps;
  a.bc1 = (float)(1.0 - pow((double)beta1, step));
  a.bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, step));
  const int blocks = (int)min((long long)ssb::sm_count() * 8, (n / 4 + OPT_THREADS - 1) / OPT_THREADS + 1);
  clip_adam_kernel<<<blocks, OPT_THREADS, 0, (cudaStream_t)stream>>>(p, g, m, v, n, norm, a);
  return check_launch("ss_clip_adam");
}
