// K7: fused mask * mixture-spectrum -> inverse real FFT -> Hann synthesis window -> overlap-add ->
// window-sum-square normalisation -> trim -> *32767 -> int16.
// Replaces steps/reconstruct_sources.py:35-42 (np.multiply + librosa.core.istft + astype('int16')).
#include "../../include/ssb200.h"
#include "fft.cuh"

namespace ssb {

constexpr int IST_TILE_T = 28;     // hops of output per CTA (2 CTAs/SM at n_fft=512)
constexpr int IST_THREADS = 256;
constexpr int IST_MAXF = IST_TILE_T + 5;

template <int NFFT, bool LAYOUT_TF, bool OUT_F32>
__global__ void __launch_bounds__(IST_THREADS) istft_kernel(const float2* __restrict__ spec,
                                                           const float* __restrict__ mask,
                                                           const long long* __restrict__ frame_off,
                                                           const long long* __restrict__ wav_off,
                                                           const int* __restrict__ tile_off, int n_utts, int n_src,
                                                           int hop, void* __restrict__ out) {
  constexpr int N = NFFT / 2, F = N + 1, R1 = N / 16, SCR = R1 * 17, XP = 17;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* tw = reinterpret_cast<float2*>(smem_raw);                 // NFFT
  float2* regA = tw + NFFT;                                         // max(F*XP, 16*SCR) float2: staged X / FFT scratch
  constexpr int REGA = (F * XP > 16 * SCR) ? F * XP : 16 * SCR;
  float* frames = reinterpret_cast<float*>(regA + REGA);            // IST_MAXF * NFFT

  const int tid = threadIdx.x;
  int lo = 0, hi = n_utts;
  const int tile = blockIdx.x;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (tile_off[mid] <= tile) lo = mid; else hi = mid;
  }
  const int u = lo, src = blockIdx.y;
  const int T = (int)(frame_off[u + 1] - frame_off[u]);
  const int L = hop * (T - 1);                                      // trimmed output length
  const int j0 = (tile - tile_off[u]) * IST_TILE_T * hop;
  const int nj = min(IST_TILE_T * hop, L - j0);
  const int q0 = j0 + NFFT / 2, q1 = q0 + nj - 1;
  const int i_lo = max(0, (q0 - NFFT + hop) / hop);                 // ceil((q0-NFFT+1)/hop)
  const int i_hi = min(T - 1, q1 / hop);
  const int nf = i_hi - i_lo + 1;

  for (int j = tid; j < NFFT; j += IST_THREADS) {
    float s, c;
    sincospif(2.0f * (float)j / (float)NFFT, &s, &c);
    tw[j] = make_float2(c, -s);
  }
  const float2* spec_u = spec + frame_off[u] * F;
  const float* mask_u = mask + (frame_off[u] * n_src + (long long)src * T) * F;
  const int grp = tid >> 4, l16 = tid & 15;
  const float invN = 1.0f / (float)N;

  for (int fb = 0; fb < nf; fb += 16) {
    const int nb = min(16, nf - fb);
    __syncthreads();                     // regA free (previous batch's scratch reads done); tw ready
    // stage X = mix_spec * mask for frames [i_lo+fb, +nb) as regA[k*XP + f]
    if (LAYOUT_TF) {
      for (int e = tid; e < nb * F; e += IST_THREADS) {
        const int f = e / F, k = e - f * F;
        const long long g = (long long)(i_lo + fb + f) * F + k;
        const float2 s = spec_u[g];
        const float m = mask_u[g];
        regA[k * XP + f] = make_float2(s.x * m, s.y * m);
      }
    } else {
      for (int e = tid; e < 16 * F; e += IST_THREADS) {
        const int k = e >> 4, f = e & 15;
        if (f < nb) {
          const long long g = (long long)k * T + (i_lo + fb + f);
          const float2 s = spec_u[g];
          const float m = mask_u[g];
          regA[k * XP + f] = make_float2(s.x * m, s.y * m);
        }
      }
    }
    __syncthreads();
    // build conj(Z[k]), Z = E + iO, E = (X[k]+conj(X[N-k]))/2, O = W^-k (X[k]-conj(X[N-k]))/2
    float2 v[R1];
    const int fsel = (grp < nb) ? grp : 0;
#pragma unroll
    for (int n1 = 0; n1 < R1; ++n1) {
      const int k = 16 * n1 + l16;
      float2 a = regA[k * XP + fsel], b = regA[(N - k) * XP + fsel];
      if (k == 0) { a.y = 0.0f; b.y = 0.0f; }                       // irfft ignores Im of DC and Nyquist
      b = cconj(b);
      const float2 e = make_float2(0.5f * (a.x + b.x), 0.5f * (a.y + b.y));
      const float2 d = make_float2(0.5f * (a.x - b.x), 0.5f * (a.y - b.y));
      const float2 o = cmul(cconj(tw[k]), d);                       // W^-k * d
      v[n1] = make_float2(e.x - o.y, -(e.y + o.x));                 // conj(e + i*o)
    }
    __syncthreads();                     // everyone has read regA -> reuse as FFT scratch
    float2* scr = regA + grp * SCR;
    fft_group16<N>(v, scr, tw, l16);
    if (grp < nb) {
      float* fr = frames + (fb + grp) * NFFT;
      for (int n = l16; n < N; n += 16) {
        const float2 r = scr[n];
        const float w0 = 0.5f - 0.5f * tw[2 * n].x, w1 = 0.5f - 0.5f * tw[2 * n + 1].x;
        *reinterpret_cast<float2*>(fr + 2 * n) = make_float2(r.x * invN * w0, -r.y * invN * w1);
      }
    }
  }
  __syncthreads();
  // overlap-add in ascending frame order (as the reference's python loop), normalise, convert
  const long long obase = wav_off[u] + (long long)src * L + j0;
  for (int j = tid; j < nj; j += IST_THREADS) {
    const int q = q0 + j;
    float y = 0.0f, wss = 0.0f;
    int i = max(0, (q - NFFT + hop) / hop);
    const int iend = min(T - 1, q / hop);
    for (; i <= iend; ++i) {
      const int n = q - i * hop;
      y += frames[(i - i_lo) * NFFT + n];
      const float w = 0.5f - 0.5f * tw[n].x;
      wss += w * w;
    }
    if (wss > 1.17549435e-38f) y /= wss;
    if (OUT_F32) reinterpret_cast<float*>(out)[obase + j] = y;
    else reinterpret_cast<short*>(out)[obase + j] = (short)(int)(y * 32767.0f);
  }
}

template <int NFFT, bool TF, bool F32>
static int launch_istft(const float2* spec, const float* mask, const long long* frame_off, const long long* wav_off,
                        const int* tile_off, int n_utts, int n_src, int total_tiles, int hop, void* out,
                        cudaStream_t st) {
  constexpr int N = NFFT / 2, F = N + 1, SCR = (N / 16) * 17;
  constexpr int REGA = (F * 17 > 16 * SCR) ? F * 17 : 16 * SCR;
  const size_t smem = (size_t)NFFT * 8 + (size_t)REGA * 8 + (size_t)IST_MAXF * NFFT * 4;
  auto kern = istft_kernel<NFFT, TF, F32>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kern<<<dim3(total_tiles, n_src), IST_THREADS, smem, st>>>(spec, mask, frame_off, wav_off, tile_off, n_utts, n_src,
                                                            hop, out);
  return check_launch("ss_istft_masked");
}

}  // namespace ssb

extern "C" int ss_istft_masked(const void* spec, const float* mask, const long long* frame_off,
                               const long long* wav_off, const int* tile_off, int n_utts, int n_src, int total_tiles,
                               int n_fft, int hop, int flags, void* out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n_fft == 512 || n_fft == 256, "ss_istft_masked: n_fft must be 256 or 512 (got %d)", n_fft);
  SS_REQUIRE(hop > 0 && hop % 2 == 0 && hop * 4 >= n_fft && hop <= n_fft, "ss_istft_masked: hop %d unsupported for n_fft %d",
             hop, n_fft);
  if (total_tiles == 0 || n_src == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  const float2* sp = (const float2*)spec;
  const bool tf = flags & SS_ISTFT_LAYOUT_TF, f32 = flags & SS_ISTFT_OUT_F32;
#define SS_GO(NF, A, B) return launch_istft<NF, A, B>(sp, mask, frame_off, wav_off, tile_off, n_utts, n_src, total_tiles, hop, out, st)
#define SS_DISPATCH(NF) \
  if (tf) { if (f32) SS_GO(NF, true, true); else SS_GO(NF, true, false); } \
  else { if (f32) SS_GO(NF, false, true); else SS_GO(NF, false, false); }
  if (n_fft == 512) { SS_DISPATCH(512) } else { SS_DISPATCH(256) }
#undef SS_DISPATCH
#undef SS_GO
  return 0;
}
