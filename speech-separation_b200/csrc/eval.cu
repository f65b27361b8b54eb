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

// ---- BSS Eval inner products (steps/evaluate_oracle.py:144: mir_eval bss_eval_sources; restated in oracle/eval_oracle.py).
// xcorr: out[p][l] = sum_m a_p[m + lag0 + l] * b_p[m] (signals zero outside their lengths), fp64 accumulation.
// These are the entries of the Toeplitz Gram blocks <s_i(.-a), s_j(.-b)> and of the right-hand sides <est, s_i(.-l)>.
constexpr int XC_CHUNK = 2048, XC_LAGS = 256;

__global__ void __launch_bounds__(XC_LAGS)
xcorr_kernel(const float* __restrict__ a, const long long* __restrict__ a_off, const int* __restrict__ a_len,
             const float* __restrict__ b, const long long* __restrict__ b_off, const int* __restrict__ b_len, int lag0,
             int nlag, double* __restrict__ out) {
  __shared__ float sa[XC_CHUNK + XC_LAGS], sb[XC_CHUNK];
  const int p = blockIdx.z, l0 = blockIdx.x * XC_LAGS, m0 = blockIdx.y * XC_CHUNK;
  const int na = a_len[p], nb = b_len[p];
  if (m0 >= nb) return;
  const float* ap = a + a_off[p];
  const float* bp = b + b_off[p];
  for (int i = threadIdx.x; i < XC_CHUNK; i += XC_LAGS) sb[i] = m0 + i < nb ? bp[m0 + i] : 0.0f;
  const int abase = m0 + lag0 + l0;                       // a index of sa[0]
  for (int i = threadIdx.x; i < XC_CHUNK + XC_LAGS; i += XC_LAGS) {
    const int j = abase + i;
    sa[i] = (j >= 0 && j < na) ? ap[j] : 0.0f;
  }
  __syncthreads();
  const int l = l0 + threadIdx.x;
  if (l >= nlag) return;
  double acc = 0.0;
  const float* x = sa + threadIdx.x;
#pragma unroll 8
  for (int i = 0; i < XC_CHUNK; ++i) acc = fma((double)x[i], (double)sb[i], acc);
  atomicAdd(out + (long long)p * nlag + l, acc);
}

// Item = (recording, estimate j).  With the distortion filters solved on the host (c1: flen taps on reference j; c2:
// S x flen taps on all references of the recording), forms for every n in [0, N + flen - 1)
//   p1[n] = sum_l c1[l] s_j[n-l]        (s_target + e_spat)        p2[n] = sum_s sum_l c2[s][l] s_s[n-l]
// and accumulates sums[item] = { sum p1^2, sum (e - p1)^2, sum (p2 - p1)^2, sum p2^2, sum (e - p2)^2 } in fp64 without ever
// storing the projections: SDR = [0]/[1], SIR = [0]/[2], SAR = [3]/[4].
constexpr int BP_OUT = 256, BP_MAXSRC = 4;

__global__ void __launch_bounds__(BP_OUT)
bss_proj_sums_kernel(const float* __restrict__ ref, const long long* __restrict__ ref_off, const float* __restrict__ est,
                     const long long* __restrict__ est_off, const int* __restrict__ n_ref, const int* __restrict__ n_est,
                     const double* __restrict__ c1, const double* __restrict__ c2, int S, int flen,
                     double* __restrict__ sums) {
  extern __shared__ unsigned char bsm[];
  double* f2 = reinterpret_cast<double*>(bsm);            // [S][flen]
  double* f1 = f2 + S * flen;                             // [flen]
  float* sig = reinterpret_cast<float*>(f1 + flen);       // [S][BP_OUT + flen - 1]
  const int item = blockIdx.y, rec = item / S, j = item - rec * S;
  const int N = n_ref[item], L = n_est[item];
  const int n0 = blockIdx.x * BP_OUT;
  if (n0 >= N + flen - 1) return;
  const int span = BP_OUT + flen - 1;
  for (int i = threadIdx.x; i < S * flen; i += BP_OUT) f2[i] = c2[(long long)item * S * flen + i];
  for (int i = threadIdx.x; i < flen; i += BP_OUT) f1[i] = c1[(long long)item * flen + i];
  for (int s = 0; s < S; ++s) {
    const float* rp = ref + ref_off[rec * S + s];
    for (int i = threadIdx.x; i < span; i += BP_OUT) {
      const int k = n0 - (flen - 1) + i;                  // sig[s][i] = s_s[n0 - flen + 1 + i]
      sig[s * span + i] = (k >= 0 && k < N) ? rp[k] : 0.0f;
    }
  }
  __syncthreads();
  const int n = n0 + threadIdx.x;
  double v[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  if (n < N + flen - 1) {
    double p1 = 0.0, p2 = 0.0;
    const float* xj = sig + j * span + threadIdx.x + flen - 1;       // s_j[n]
    for (int l = 0; l < flen; ++l) p1 = fma(f1[l], (double)xj[-l], p1);
    for (int s = 0; s < S; ++s) {
      const float* xs = sig + s * span + threadIdx.x + flen - 1;
      const double* f = f2 + s * flen;
      for (int l = 0; l < flen; ++l) p2 = fma(f[l], (double)xs[-l], p2);
    }
    const double e = n < L ? (double)est[est_off[item] + n] : 0.0;
    v[0] = p1 * p1; v[1] = (e - p1) * (e - p1); v[2] = (p2 - p1) * (p2 - p1); v[3] = p2 * p2; v[4] = (e - p2) * (e - p2);
  }
  __shared__ double red[BP_OUT / 32][5];
#pragma unroll
  for (int k = 0; k < 5; ++k)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
  if ((threadIdx.x & 31) == 0)
    for (int k = 0; k < 5; ++k) red[threadIdx.x >> 5][k] = v[k];
  __syncthreads();
  if (threadIdx.x < 5) {
    double t = 0.0;
    for (int w = 0; w < BP_OUT / 32; ++w) t += red[w][threadIdx.x];
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

extern "C" int ss_xcorr(const float* a, const long long* a_off, const int* a_len, const float* b, const long long* b_off,
                        const int* b_len, int n_pairs, int max_b_len, int lag0, int nlag, double* out, void* stream) {
  using namespace ssb;
  if (n_pairs <= 0 || nlag <= 0) return 0;
  SS_REQUIRE(max_b_len > 0, "ss_xcorr: max_b_len must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, sizeof(double) * (size_t)n_pairs * nlag, st);
  xcorr_kernel<<<dim3(cdiv(nlag, XC_LAGS), cdiv(max_b_len, XC_CHUNK), n_pairs), XC_LAGS, 0, st>>>(a, a_off, a_len, b, b_off,
                                                                                                b_len, lag0, nlag, out);
  return check_launch("ss_xcorr");
}

extern "C" int ss_bss_proj_sums(const float* ref, const long long* ref_off, const float* est, const long long* est_off,
                                const int* n_ref, const int* n_est, const double* c1, const double* c2, int n_items,
                                int n_src, int flen, int max_ref, double* sums, void* stream) {
  using namespace ssb;
  if (n_items <= 0) return 0;
  SS_REQUIRE(n_src >= 1 && n_src <= BP_MAXSRC && flen >= 1 && max_ref > 0, "ss_bss_proj_sums: bad shape");
  const size_t smem = (size_t)(n_src + 1) * flen * 8 + (size_t)n_src * (BP_OUT + flen - 1) * 4;
  SS_REQUIRE(smem <= 200 * 1024, "ss_bss_proj_sums: filter length %d too large", flen);
  cudaStream_t st = (cudaStream_t)stream;
  cudaFuncSetAttribute(bss_proj_sums_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaMemsetAsync(sums, 0, sizeof(double) * 5 * (size_t)n_items, st);
  bss_proj_sums_kernel<<<dim3(cdiv(max_ref + flen - 1, BP_OUT), n_items), BP_OUT, smem, st>>>(
      ref, ref_off, est, est_off, n_ref, n_est, c1, c2, n_src, flen, sums);
  return check_launch("ss_bss_proj_sums");
}
