// K1: batched STFT over a ragged corpus of utterances.
// Replaces steps/extract_feats.py:76-89,98,105 (librosa.core.stft [+ np.abs]) of the reference:
// reflect pad n_fft/2, frames at stride hop, periodic Hann, real FFT -> |X| (train) or complex (test).
#include "../../include/ssb200.h"
#include "fft.cuh"

namespace ssb {

constexpr int STFT_TILE_T = 32;      // frames per CTA
constexpr int STFT_THREADS = 256;    // 16 frames in flight, 16 threads each

template <int NFFT>
struct StftSmem {
  static constexpr int N = NFFT / 2;
  static constexpr int F = N + 1;
  static constexpr int SCR = (N / 16) * 17;   // float2 per frame group
  float2 tw[NFFT];
  float2 scr[16][SCR];
  // followed by: float sig[(TILE_T-1)*hop + NFFT]; then staging out[F][TILE_T+1] (float or float2)
};

template <int NFFT, bool COMPLEX, bool LAYOUT_TF, bool LOGMAG>
__global__ void __launch_bounds__(STFT_THREADS) stft_kernel(const float* __restrict__ sig,
                                                           const long long* __restrict__ sig_off,
                                                           const long long* __restrict__ frame_off,
                                                           const int* __restrict__ tile_off, int n_utts, int hop,
                                                           float* __restrict__ out) {
  using SM = StftSmem<NFFT>;
  constexpr int N = SM::N, F = SM::F, R1 = N / 16, TT = STFT_TILE_T;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SM& sm = *reinterpret_cast<SM*>(smem_raw);
  float* sigs = reinterpret_cast<float*>(smem_raw + sizeof(SM));
  const int span = (TT - 1) * hop + NFFT;
  float* stage = sigs + ((span + 3) & ~3);

  // which utterance does this tile belong to?  (binary search over the tile prefix)
  int lo = 0, hi = n_utts;
  const int tile = blockIdx.x;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (tile_off[mid] <= tile) lo = mid; else hi = mid;
  }
  const int u = lo;
  const int t0 = (tile - tile_off[u]) * TT;
  const long long s0 = sig_off[u];
  const int n_samp = (int)(sig_off[u + 1] - s0);
  const int T = (int)(frame_off[u + 1] - frame_off[u]);
  const int nt = min(TT, T - t0);
  const int tid = threadIdx.x;

  for (int j = tid; j < NFFT; j += STFT_THREADS) {
    float s, c;
    sincospif(2.0f * (float)j / (float)NFFT, &s, &c);
    sm.tw[j] = make_float2(c, -s);
  }
  // stage the signal span with reflect padding (np.pad(y, n_fft//2, 'reflect'))
  const int used = (nt - 1) * hop + NFFT;
  const int p0 = t0 * hop - NFFT / 2;
  for (int j = tid; j < used; j += STFT_THREADS) {
    int p = p0 + j;
    if (p < 0) p = -p;
    if (p >= n_samp) p = 2 * (n_samp - 1) - p;
    sigs[j] = sig[s0 + p];
  }
  __syncthreads();

  const int grp = tid >> 4, l16 = tid & 15;
  float2* scr = sm.scr[grp];
  float* out_u = out + frame_off[u] * (long long)F * (COMPLEX ? 2 : 1);
  for (int fb = 0; fb < TT; fb += 16) {
    const int fl = fb + grp;
    if (fb < nt) {                       // warp-uniform enough: both groups of a warp share fb
      const bool live = fl < nt;
      float2 v[R1];
      const float* x = sigs + (live ? fl : 0) * hop;
#pragma unroll
      for (int n1 = 0; n1 < R1; ++n1) {
        const int n = 16 * n1 + l16;
        const float2 xs = *reinterpret_cast<const float2*>(x + 2 * n);
        const float w0 = 0.5f - 0.5f * sm.tw[2 * n].x, w1 = 0.5f - 0.5f * sm.tw[2 * n + 1].x;
        v[n1] = make_float2(xs.x * w0, xs.y * w1);
      }
      fft_group16<N>(v, scr, sm.tw, l16);
      // split the packed transform: X[k] = (Z[k]+conj(Z[N-k]))/2 - i/2 * e^{-2 pi i k/(2N)} * (Z[k]-conj(Z[N-k]))
      if (live) {
        for (int k = l16; k <= N; k += 16) {
          const float2 a = scr[k & (N - 1)], b = cconj(scr[(N - k) & (N - 1)]);
          const float2 e = make_float2(0.5f * (a.x + b.x), 0.5f * (a.y + b.y));
          const float2 d = make_float2(0.5f * (a.x - b.x), 0.5f * (a.y - b.y));
          const float2 wd = cmul(sm.tw[k], d);
          const float2 X = make_float2(e.x + wd.y, e.y - wd.x);      // e + (-i)*wd
          if (COMPLEX) {
            if (LAYOUT_TF) reinterpret_cast<float2*>(out_u)[(long long)(t0 + fl) * F + k] = X;
            else reinterpret_cast<float2*>(stage)[k * (TT + 1) + fl] = X;
          } else {
            float m = sqrtf(X.x * X.x + X.y * X.y);
            if (LOGMAG) m = logf(fmaxf(m, 1e-8f));
            if (LAYOUT_TF) out_u[(long long)(t0 + fl) * F + k] = m;
            else stage[k * (TT + 1) + fl] = m;
          }
        }
      }
      __syncwarp();
    }
  }
  if (!LAYOUT_TF) {
    __syncthreads();
    const int lane = tid & 31, warp = tid >> 5;
    if (lane < nt) {
      for (int k = warp; k < F; k += STFT_THREADS / 32) {
        if (COMPLEX)
          reinterpret_cast<float2*>(out_u)[(long long)k * T + t0 + lane] =
              reinterpret_cast<float2*>(stage)[k * (TT + 1) + lane];
        else
          out_u[(long long)k * T + t0 + lane] = stage[k * (TT + 1) + lane];
      }
    }
  }
}

template <int NFFT, bool COMPLEX, bool TF, bool LOGMAG>
static int launch_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                       int n_utts, int total_tiles, int hop, float* out, cudaStream_t st) {
  using SM = StftSmem<NFFT>;
  const int span = (STFT_TILE_T - 1) * hop + NFFT;
  size_t smem = sizeof(SM) + (size_t)((span + 3) & ~3) * 4;
  if (!TF) smem += (size_t)SM::F * (STFT_TILE_T + 1) * (COMPLEX ? 8 : 4);
  auto kern = stft_kernel<NFFT, COMPLEX, TF, LOGMAG>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  kern<<<total_tiles, STFT_THREADS, smem, st>>>(sig, sig_off, frame_off, tile_off, n_utts, hop, out);
  return check_launch("ss_stft");
}

}  // namespace ssb

extern "C" int ss_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                       int n_utts, int total_tiles, int n_fft, int hop, int flags, void* out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n_fft == 512 || n_fft == 256, "ss_stft: n_fft must be 256 or 512 (got %d)", n_fft);
  SS_REQUIRE(hop > 0 && hop <= n_fft && hop % 2 == 0, "ss_stft: bad hop %d", hop);
  if (total_tiles == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  float* o = (float*)out;
  const bool cplx = flags & SS_STFT_COMPLEX, tf = flags & SS_STFT_LAYOUT_TF, lg = flags & SS_STFT_LOG;
  SS_REQUIRE(!(cplx && lg), "ss_stft: LOG applies to magnitude output only");
#define SS_GO(NF, C, T, L) return launch_stft<NF, C, T, L>(sig, sig_off, frame_off, tile_off, n_utts, total_tiles, hop, o, st)
#define SS_DISPATCH(NF)                         \
  if (cplx) { if (tf) SS_GO(NF, true, true, false); else SS_GO(NF, true, false, false); } \
  else if (lg) { if (tf) SS_GO(NF, false, true, true); else SS_GO(NF, false, false, true); } \
  else { if (tf) SS_GO(NF, false, true, false); else SS_GO(NF, false, false, false); }
  if (n_fft == 512) { SS_DISPATCH(512) } else { SS_DISPATCH(256) }
#undef SS_DISPATCH
#undef SS_GO
  return 0;
}
