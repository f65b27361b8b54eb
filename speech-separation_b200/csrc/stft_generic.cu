// K1 / K7 for ANY even n_fft and any hop <= n_fft: direct-summation DFT kernels behind the same two entry points.
// The reference passes --fft-dim / --step-size straight to librosa (steps/extract_feats.py:23-28,
// steps/reconstruct_sources.py:22-27), so every even size is legal there; the packed-FFT kernels (stft.cu, istft.cu)
// cover the sizes the recipes use (256, 512 with an even hop; iSTFT hop >= n_fft/4) and these cover the rest.
// O(n_fft^2) per frame instead of O(n_fft log n_fft): a correctness path, not a roofline path (e.g. n_fft = 400 costs
// ~4x the 512-point FFT kernel per frame).  Same ragged layouts, flags and rounding rules as the fast kernels.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int GEN_FB = 4;          // frames transformed together (shared-memory frame buffer)
constexpr int GEN_TILE = 32;       // frames per tile: ss_stft_tile_frames()

// periodic Hann window and the twiddle table cos / sin(2 pi m / n_fft), built per CTA in shared memory
__device__ __forceinline__ void gen_tables(float* win, float2* tw, int n_fft) {
  for (int m = threadIdx.x; m < n_fft; m += blockDim.x) {
    float s, c;
    sincospif(2.0f * (float)m / (float)n_fft, &s, &c);
    tw[m] = make_float2(c, s);
    win[m] = 0.5f - 0.5f * c;
  }
}

__global__ void __launch_bounds__(256)
stft_generic_kernel(const float* __restrict__ sig, const long long* __restrict__ sig_off,
                    const long long* __restrict__ frame_off, const int* __restrict__ tile_off, int n_utts,
                    int total_tiles, int n_fft, int hop, int flags, float* __restrict__ out) {
  extern __shared__ unsigned char gsm[];
  float2* tw = reinterpret_cast<float2*>(gsm);
  float* win = reinterpret_cast<float*>(tw + n_fft);
  float* fr = win + n_fft;                                    // [GEN_FB][n_fft] windowed frames
  const int F = n_fft / 2 + 1, half = n_fft / 2;
  const bool cplx = flags & SS_STFT_COMPLEX, tf = flags & SS_STFT_LAYOUT_TF, lg = flags & SS_STFT_LOG;
  gen_tables(win, tw, n_fft);
  int u = 0;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    while (u + 1 < n_utts && tile >= tile_off[u + 1]) ++u;    // tiles ascend per CTA
    while (tile < tile_off[u]) --u;
    const long long s0 = sig_off[u];
    const int N = (int)(sig_off[u + 1] - s0);
    const int T = (int)(frame_off[u + 1] - frame_off[u]);
    const int t0 = (tile - tile_off[u]) * GEN_TILE;
    float* ob = out + frame_off[u] * F * (cplx ? 2 : 1);
    for (int tb = t0; tb < min(T, t0 + GEN_TILE); tb += GEN_FB) {
      const int nf = min(GEN_FB, min(T, t0 + GEN_TILE) - tb);
      __syncthreads();
      for (int e = threadIdx.x; e < nf * n_fft; e += blockDim.x) {
        const int f = e / n_fft, n = e - f * n_fft;
        int j = (tb + f) * hop - half + n;                    // np.pad(y, n_fft//2, 'reflect')
        j = j < 0 ? -j : j;
        j = j >= N ? 2 * N - 2 - j : j;
        fr[e] = sig[s0 + j] * win[n];
      }
      __syncthreads();
      for (int e = threadIdx.x; e < nf * F; e += blockDim.x) {
        const int f = e / F, k = e - f * F;
        const float* x = fr + f * n_fft;
        float re = 0.0f, im = 0.0f;
        int m = 0;
        for (int n = 0; n < n_fft; ++n) {
          const float2 w = tw[m];
          re = fmaf(x[n], w.x, re);
          im = fmaf(-x[n], w.y, im);
          m += k;
          m = m >= n_fft ? m - n_fft : m;
        }
        const int t = tb + f;
        const long long o = tf ? (long long)t * F + k : (long long)k * T + t;
        if (cplx) {
          reinterpret_cast<float2*>(ob)[o] = make_float2(re, im);
        } else {
          float v = hypotf(re, im);
          if (lg) v = logf(fmaxf(v, 1e-8f));
          ob[o] = v;
        }
      }
    }
  }
}

// One thread per output sample: every frame that overlaps it is inverse-transformed at that one position,
//   x_t[m] = (1/N) [ S_0 + (-1)^m S_{N/2} + 2 sum_{0<k<N/2} (Re S_k cos(2 pi k m/N) - Im S_k sin(2 pi k m/N)) ]
// (np.fft.irfft: the imaginary parts of the DC and Nyquist bins are ignored), times the synthesis window, summed in
// ascending frame order like librosa's loop, divided by the window sum-square where it exceeds tiny(float32).
__global__ void __launch_bounds__(256)
istft_generic_kernel(const float2* __restrict__ spec, const float* __restrict__ mask,
                     const long long* __restrict__ frame_off, const long long* __restrict__ wav_off,
                     const int* __restrict__ tile_off, int n_utts, int n_src, int total_tiles, int tile_samples,
                     int n_fft, int hop, int flags, void* __restrict__ out) {
  extern __shared__ unsigned char gsm[];
  float2* tw = reinterpret_cast<float2*>(gsm);
  float* win = reinterpret_cast<float*>(tw + n_fft);
  const int F = n_fft / 2 + 1, half = n_fft / 2;
  const bool tf = flags & SS_ISTFT_LAYOUT_TF, f32 = flags & SS_ISTFT_OUT_F32;
  gen_tables(win, tw, n_fft);
  __syncthreads();
  const float inv_n = 1.0f / (float)n_fft;
  int u = 0;
  const long long n_work = (long long)total_tiles * n_src;
  for (long long w = blockIdx.x; w < n_work; w += gridDim.x) {
    const int tile = (int)(w / n_src), s = (int)(w - (long long)tile * n_src);
    while (u + 1 < n_utts && tile >= tile_off[u + 1]) ++u;
    while (tile < tile_off[u]) --u;
    const int T = (int)(frame_off[u + 1] - frame_off[u]);
    const int L = hop * (T - 1);
    const float2* sp = spec + frame_off[u] * F;
    const float* mk = mask + (frame_off[u] * n_src + (long long)s * T) * F;
    const int j0 = (tile - tile_off[u]) * tile_samples;
    for (int j = j0 + threadIdx.x; j < min(L, j0 + tile_samples); j += blockDim.x) {
      const int n = j + half;                                  // position in the untrimmed overlap-add buffer
      const int t_lo = n >= n_fft ? (n - n_fft) / hop + 1 : 0, t_hi = min(T - 1, n / hop);
      float y = 0.0f, wss = 0.0f;
      for (int t = t_lo; t <= t_hi; ++t) {
        const int m = n - t * hop;
        float acc = 0.0f;
        int a = m;                                             // (k m) mod n_fft for k = 1, 2, ...
        for (int k = 1; k < half; ++k) {
          const long long e = tf ? (long long)t * F + k : (long long)k * T + t;
          const float2 z = sp[e];
          const float g = mk[e];
          const float2 c = tw[a];
          acc = fmaf(g * z.x, c.x, acc);
          acc = fmaf(-g * z.y, c.y, acc);
          a += m;
          a = a >= n_fft ? a - n_fft : a;
        }
        const long long e0 = tf ? (long long)t * F : (long long)t, eN = tf ? (long long)t * F + half : (long long)half * T + t;
        const float dc = mk[e0] * sp[e0].x, ny = mk[eN] * sp[eN].x;
        const float x = (dc + ((m & 1) ? -ny : ny) + 2.0f * acc) * inv_n;
        y = fmaf(win[m], x, y);
        wss = fmaf(win[m], win[m], wss);
      }
      if (wss > 1.17549435e-38f) y /= wss;
      const long long o = wav_off[u] + (long long)s * L + j;
      if (f32) reinterpret_cast<float*>(out)[o] = y;
      else reinterpret_cast<short*>(out)[o] = (short)(int)(y * 32767.0f);
    }
  }
}

int launch_stft_generic(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                        int n_utts, int total_tiles, int n_fft, int hop, int flags, float* out, cudaStream_t st) {
  const size_t smem = (size_t)n_fft * (8 + 4 + 4 * GEN_FB);
  SS_REQUIRE(smem <= 200 * 1024, "ss_stft: n_fft %d too large for the generic kernel", n_fft);
  cudaFuncSetAttribute(stft_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int grid = min(total_tiles, sm_count() * 4);
  stft_generic_kernel<<<grid, 256, smem, st>>>(sig, sig_off, frame_off, tile_off, n_utts, total_tiles, n_fft, hop, flags, out);
  return check_launch("ss_stft (generic)");
}

int launch_istft_generic(const void* spec, const float* mask, const long long* frame_off, const long long* wav_off,
                         const int* tile_off, int n_utts, int n_src, int total_tiles, int tile_samples, int n_fft, int hop,
                         int flags, void* out, cudaStream_t st) {
  const size_t smem = (size_t)n_fft * 12;
  SS_REQUIRE(smem <= 200 * 1024, "ss_istft_masked: n_fft %d too large for the generic kernel", n_fft);
  cudaFuncSetAttribute(istft_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const long long n_work = (long long)total_tiles * n_src;
  const int grid = (int)min(n_work, (long long)sm_count() * 4);
  istft_generic_kernel<<<grid, 256, smem, st>>>((const float2*)spec, mask, frame_off, wav_off, tile_off, n_utts, n_src,
                                               total_tiles, tile_samples, n_fft, hop, flags, out);
  return check_launch("ss_istft_masked (generic)");
}

}  // namespace ssb
