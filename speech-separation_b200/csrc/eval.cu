// f4: oracle-mask evaluation on the device.
//   * ss_oracle_masks: the ideal ratio / ideal binary masks of steps/evaluate_oracle.py:133-139,
//       soft: mask_i = |S_i| / |X|            (np.divide; a bin with |X| == 0 exactly gets mask 0, see the kernel)
//       hard: mask_i = [argmax_j |S_j| == i]  (np.argmax: the first maximum wins)
//     written in the per-utterance, per-source block layout ss_istft_masked reads, so the reconstruction
//     (evaluate_oracle.py:140-143) is the SAME kernel as the separation path (K7).
//   * ss_sdr_sums: the five sums (sum r, sum e, sum r^2, sum e^2, sum r*e) of a reference / estimate waveform pair in
//     fp64, from which the host forms SDR-type ratios (scale-invariant SDR; the projections of BSS-eval reduce to such
//     inner products too).  The estimate may be shorter than the reference (the iSTFT yields hop*(T-1) <= N samples;
//     evaluate_oracle.py:142-143 leaves the tail zero).
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

__global__ void __launch_bounds__(256)
oracle_masks_kernel(const float2* __restrict__ spec, const float* __restrict__ src_mag, long long src_stride,
                    const long long* __restrict__ frame_off, int F, int n_src, int hard, float* __restrict__ masks) {
  const int u = blockIdx.y;
  const long long f0 = frame_off[u], n = (frame_off[u + 1] - f0) * F;      // elements of one (F, T_u) block
  const float2* sp = spec + f0 * F;
  float* mk = masks + f0 * F * n_src;
  for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n; e += (long long)gridDim.x * 256) {
    if (hard) {
      int best = 0;
      float bv = src_mag[f0 * F + e];
      for (int s = 1; s < n_src; ++s) {
        const float v = src_mag[s * src_stride + f0 * F + e];
        if (v > bv) { bv = v; best = s; }
      }
      for (int s = 0; s < n_src; ++s) mk[s * n + e] = s == best ? 1.0f : 0.0f;
    } else {
      const float2 x = sp[e];
      const float mag = hypotf(x.x, x.y);                    // np.abs(complex64): no underflow of the squares -> no spurious x/0
      // |X| == 0 exactly: np.divide would give inf / nan and the reference's iSTFT would spread NaN over the whole frame
      // (its score for the recording becomes nan).  Which bins are EXACTLY zero is an accident of the FFT arithmetic
      // (this kernel's fp32 FFT hits a few that numpy's does not), so such a bin contributes nothing here instead.
      for (int s = 0; s < n_src; ++s) mk[s * n + e] = mag > 0.0f ? src_mag[s * src_stride + f0 * F + e] / mag : 0.0f;
    }
  }
}

constexpr int SDR_CHUNK = 8192;     // samples per block

__global__ void __launch_bounds__(256)
sdr_sums_kernel(const float* __restrict__ ref, const long long* __restrict__ ref_off, const float* __restrict__ est,
                const long long* __restrict__ est_off, const int* __restrict__ n_ref, const int* __restrict__ n_est,
                double* __restrict__ sums) {
  const int item = blockIdx.y;
  const int N = n_ref[item], L = n_est[item];
  const int i0 = blockIdx.x * SDR_CHUNK;
  if (i0 >= N) return;
  const int i1 = min(N, i0 + SDR_CHUNK);
  const float* r = ref + ref_off[item];
  const float* e = est + est_off[item];
  double a[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  for (int i = i0 + threadIdx.x; i < i1; i += 256) {
    const double rv = r[i], ev = i < L ? e[i] : 0.0;
    a[0] += rv; a[1] += ev; a[2] += rv * rv; a[3] += ev * ev; a[4] += rv * ev;
  }
  __shared__ double red[8][5];
#pragma unroll
  for (int k = 0; k < 5; ++k) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
  }
  if ((threadIdx.x & 31) == 0)
    for (int k = 0; k < 5; ++k) red[threadIdx.x >> 5][k] = a[k];
  __syncthreads();
  if (threadIdx.x < 5) {
    double t = 0.0;
    for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
    atomicAdd(&sums[(long long)item * 5 + threadIdx.x], t);
  }
}

}  // namespace ssb

extern "C" int ss_oracle_masks(const void* spec, const float* src_mag, long long src_stride, const long long* frame_off,
                               int n_utts, int max_frames, int F, int n_src, int hard, float* masks, void* stream) {
  using namespace ssb;
  if (n_utts <= 0) return 0;
  SS_REQUIRE(n_src >= 1 && F > 0 && max_frames > 0, "ss_oracle_masks: bad shape");
  const long long per = (long long)max_frames * F;
  const int gx = (int)min((long long)sm_count() * 4, (per + 255) / 256);
  oracle_masks_kernel<<<dim3(gx, n_utts), 256, 0, (cudaStream_t)stream>>>((const float2*)spec, src_mag, src_stride,
                                                                         frame_off, F, n_src, hard, masks);
  return check_launch("ss_oracle_masks");
}

extern "C" int ss_sdr_sums(const float* ref, const long long* ref_off, const float* est, const long long* est_off,
                           const int* n_ref, const int* n_est, int n_items, int max_ref, double* sums, void* stream) {
  using namespace ssb;
  if (n_items <= 0) return 0;
  SS_REQUIRE(max_ref > 0, "ss_sdr_sums: max_ref must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(sums, 0, sizeof(double) * 5 * (size_t)n_items, st);
  sdr_sums_kernel<<<dim3(cdiv(max_ref, SDR_CHUNK), n_items), 256, 0, st>>>(ref, ref_off, est, est_off, n_ref, n_est, sums);
  return check_launch("ss_sdr_sums");
}
