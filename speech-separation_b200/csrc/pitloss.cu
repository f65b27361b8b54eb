// K5/K6: fused sigmoid-mask * mixture, pairwise squared-error sums, all-permutation (uPIT) or greedy (RSH)
// assignment with per-utterance argmin, loss, and the backward into the head's pre-activations.
// Replaces archs/uPIT.py:178-206 (compute_loss: cat/mul, per-permutation MSELoss sums, torch.min, normaliser)
// and archs/RSH.py:224-241 (per-pass losses, +inf for used sources, torch.min, usage bookkeeping).
#include <math_constants.h>

#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int PL_CHUNK = 8;      // frames per CTA in the partial-sum pass
constexpr int PL_THREADS = 256;

template <int NO, int NS>
__global__ void __launch_bounds__(PL_THREADS)
pit_partial_kernel(const float* __restrict__ z, long long ldz, const float* __restrict__ mix,
                   const float* __restrict__ src, const int* __restrict__ lens, int T, int B, int F,
                   float* __restrict__ mask_out, float* __restrict__ partial) {
  const int b = blockIdx.y, t0 = blockIdx.x * PL_CHUNK;
  const int len = lens[b];
  const int t1 = min(min(T, t0 + PL_CHUNK), len);
  float acc[NO * NS];
#pragma unroll
  for (int i = 0; i < NO * NS; ++i) acc[i] = 0.0f;
  const long long plane = (long long)T * B * F;
  const int nel = (t1 - t0) * F;
  for (int e = threadIdx.x; e < nel; e += PL_THREADS) {
    const int tl = e / F, f = e - tl * F;
    const long long row = (long long)(t0 + tl) * B + b;
    const float x = mix[row * F + f];
    float s[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) s[j] = src[j * plane + row * F + f];
#pragma unroll
    for (int i = 0; i < NO; ++i) {
      const float m = sigmoidf_(z[row * ldz + i * F + f]);
      if (mask_out) mask_out[row * (long long)(NO * F) + i * F + f] = m;
      const float y = m * x;
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        const float d = y - s[j];
        acc[i * NS + j] += d * d;
      }
    }
  }
  __shared__ float red[PL_THREADS / 32][NO * NS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll
  for (int i = 0; i < NO * NS; ++i) {
    const float v = warp_sum(acc[i]);
    if (lane == 0) red[warp][i] = v;
  }
  __syncthreads();
  if (threadIdx.x < NO * NS) {
    float v = 0.0f;
    for (int w = 0; w < PL_THREADS / 32; ++w) v += red[w][threadIdx.x];
    partial[((long long)blockIdx.x * B + b) * (NO * NS) + threadIdx.x] = v;
  }
}

// One CTA.  mode 0 (uPIT): n_out == n_src == S, enumerate permutations in itertools (lexicographic) order.
// mode 1 (RSH pass): n_out == 1, sources already used by utterance b (bitmask used[b]) cost +inf.
__global__ void __launch_bounds__(256)
pit_select_kernel(const float* __restrict__ partial, int chunks, const int* __restrict__ lens, int B, int F, int n_out,
                  int n_src, int mode, int accumulate, float* __restrict__ losses, int* __restrict__ idx,
                  int* __restrict__ assign, int* __restrict__ used, float* __restrict__ stats) {
  __shared__ double s_min[256], s_len[256];
  double my_min = 0.0, my_len = 0.0;
  const int ne = n_out * n_src;
  for (int b = threadIdx.x; b < B; b += 256) {
    double E[16];
    for (int i = 0; i < ne; ++i) {
      double a = 0.0;
      for (int c = 0; c < chunks; ++c) a += (double)partial[((long long)c * B + b) * ne + i];
      E[i] = a;
    }
    float best = CUDART_INF_F;
    int best_i = 0;
    if (mode == 0) {
      int perm[4] = {0, 1, 2, 3}, best_perm[4] = {0, 1, 2, 3};
      int p = 0;
      while (true) {
        double l = 0.0;
        for (int k = 0; k < n_out; ++k) l += E[k * n_src + perm[k]];
        const float lf = (float)l;
        losses[(long long)p * B + b] = lf;
        if (lf < best) {
          best = lf;
          best_i = p;
          for (int k = 0; k < 4; ++k) best_perm[k] = perm[k];
        }
        // next lexicographic permutation of perm[0..n_src)
        int i = n_src - 2;
        while (i >= 0 && perm[i] > perm[i + 1]) --i;
        if (i < 0) break;
        int j = n_src - 1;
        while (perm[j] < perm[i]) --j;
        int tmp = perm[i]; perm[i] = perm[j]; perm[j] = tmp;
        for (int l2 = i + 1, r = n_src - 1; l2 < r; ++l2, --r) { tmp = perm[l2]; perm[l2] = perm[r]; perm[r] = tmp; }
        ++p;
      }
      for (int k = 0; k < n_out; ++k) assign[b * n_out + k] = best_perm[k];
    } else {
      const int u = used[b];
      for (int s = 0; s < n_src; ++s) {
        const float lf = ((u >> s) & 1) ? CUDART_INF_F : (float)E[s];
        losses[(long long)s * B + b] = lf;
        if (lf < best) { best = lf; best_i = s; }
      }
      used[b] = u | (1 << best_i);
      assign[b] = best_i;
    }
    idx[b] = best_i;
    my_min += (double)best;
    my_len += (double)lens[b];
  }
  s_min[threadIdx.x] = my_min;
  s_len[threadIdx.x] = my_len;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0, tl = 0.0;
    for (int i = 0; i < 256; ++i) { tot += s_min[i]; tl += s_len[i]; }
    const double S = (double)n_src;
    double loss_sum = tot / S, norm = tl * (double)F;
    if (accumulate) { loss_sum += (double)stats[0]; norm += (double)stats[1]; }
    stats[0] = (float)loss_sum;          // sum over (passes and) utterances of min / S     (uPIT.py:196, RSH.py:240)
    stats[1] = (float)norm;              // sum_b len_b * F (per pass)                      (uPIT.py:197, RSH.py:241)
    stats[2] = (float)(loss_sum / norm); // the returned loss                               (uPIT.py:206, RSH.py:259)
    stats[3] = (float)(1.0 / (S * norm));
  }
}

// dz = [ 2*(m*x - s_assigned)*x * gout/(S*norm) + dmask_extra ] * m*(1-m), written as sb16 (rows t*B+b): the operand of
// both head backward GEMMs (dY = dz W in the K-major form, dW = dz^T y in the MN-major form).
__global__ void __launch_bounds__(256)
pit_bwd_kernel(const float* __restrict__ z, long long ldz, const float* __restrict__ mix, const float* __restrict__ src,
               const int* __restrict__ lens, const int* __restrict__ assign, int T, int B, int F, int n_out,
               const float* __restrict__ stats, const float* __restrict__ gout, const float* __restrict__ dmask_extra,
               __nv_bfloat16* __restrict__ dz, long long ld_dz, long long dz_plane) {
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int R = T * B, C = n_out * F;
  const float g = 2.0f * stats[3] * (gout ? gout[0] : 1.0f);
  const long long plane = (long long)T * B * F;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + tx;
    float v = 0.0f;
    if (r < R && c < C) {
      const int t = r / B, b = r - t * B;
      if (t < lens[b]) {
        const int k = c / F, f = c - k * F;
        const float m = sigmoidf_(z[(long long)r * ldz + c]);
        const float x = mix[(long long)r * F + f];
        const float s = src[assign[b * n_out + k] * plane + (long long)r * F + f];
        float dm = g * (m * x - s) * x;
        if (dmask_extra) dm += dmask_extra[(long long)r * C + c];
        v = dm * m * (1.0f - m);
      }
      __nv_bfloat16 hi, lo;
      split_bf16(v, hi, lo);
      const long long o = (long long)r * ld_dz + c;
      dz[o] = hi;
      dz[o + dz_plane] = lo;
    }
  }
}

}  // namespace ssb

namespace ssb {
__global__ void pit_stats_finalize_kernel(float* stats, int n_src) {
  stats[2] = stats[0] / stats[1];
  stats[3] = (float)(1.0 / ((double)n_src * (double)stats[1]));
}
}  // namespace ssb

// data-parallel mode B: stats[0] (sum of min losses / S) and stats[1] (norm) have been sum-allreduced over the ranks;
// recompute the loss and the gradient scale from them
extern "C" int ss_pit_stats_finalize(float* stats, int n_src, void* stream) {
  ssb::pit_stats_finalize_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(stats, n_src);
  return ssb::check_launch("ss_pit_stats_finalize");
}

extern "C" int ss_pit_chunks(int T) { return ssb::cdiv(T, ssb::PL_CHUNK); }

extern "C" int ss_pit_loss_fwd(const float* z, long long ldz, const float* mix, const float* src, const int* lens,
                               int T, int B, int F, int n_out, int n_src, float* mask_out, float* partial,
                               void* stream) {
  using namespace ssb;
  SS_REQUIRE(n_out >= 1 && n_src >= 1 && n_src <= 4 && (n_out == n_src || n_out == 1),
             "ss_pit_loss_fwd: unsupported n_out=%d n_src=%d", n_out, n_src);
  if (T == 0 || B == 0) return 0;
  dim3 grid(cdiv(T, PL_CHUNK), B);
  cudaStream_t st = (cudaStream_t)stream;
#define SS_GO(NO, NS) pit_partial_kernel<NO, NS><<<grid, PL_THREADS, 0, st>>>(z, ldz, mix, src, lens, T, B, F, mask_out, partial)
  if (n_out == 1) {
    if (n_src == 1) SS_GO(1, 1); else if (n_src == 2) SS_GO(1, 2); else if (n_src == 3) SS_GO(1, 3); else SS_GO(1, 4);
  } else {
    if (n_src == 2) SS_GO(2, 2); else if (n_src == 3) SS_GO(3, 3); else SS_GO(4, 4);
  }
#undef SS_GO
  return check_launch("ss_pit_loss_fwd");
}

extern "C" int ss_pit_select(const float* partial, int chunks, const int* lens, int B, int F, int n_out, int n_src,
                             int mode, int accumulate, float* losses, int* idx, int* assign, int* used, float* stats,
                             void* stream) {
  using namespace ssb;
  SS_REQUIRE((mode == 0 && n_out == n_src) || (mode == 1 && n_out == 1 && used), "ss_pit_select: bad mode/shape");
  SS_REQUIRE(n_src <= 4, "ss_pit_select: at most 4 sources");
  pit_select_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(partial, chunks, lens, B, F, n_out, n_src, mode, accumulate,
                                                        losses, idx, assign, used, stats);
  return check_launch("ss_pit_select");
}

extern "C" int ss_pit_loss_bwd(const float* z, long long ldz, const float* mix, const float* src, const int* lens,
                               const int* assign, int T, int B, int F, int n_out, const float* stats,
                               const float* gout, const float* dmask_extra, void* dz, long long ld_dz,
                               long long dz_plane, void* stream) {
  using namespace ssb;
  if (T == 0 || B == 0) return 0;
  dim3 grid(cdiv(n_out * F, 32), cdiv((long long)T * B, 32));
  pit_bwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(z, ldz, mix, src, lens, assign, T, B, F, n_out, stats, gout,
                                                        dmask_extra, (__nv_bfloat16*)dz, ld_dz, dz_plane);
  return check_launch("ss_pit_loss_bwd");
}
