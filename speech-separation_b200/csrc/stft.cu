// K1: batched STFT over a ragged corpus of utterances.
// Replaces steps/extract_feats.py:76-89,98,105 (librosa.core.stft [+ np.abs]) of the reference:
// reflect pad n_fft/2, frames at stride hop, periodic Hann, real FFT -> |X| (train) or complex (test).
#include "../../include/ssb200.h"
#include "fft.cuh"

namespace ssb {

constexpr int STFT_TILE_T = 32;      // frames per tile
constexpr int STFT_THREADS = 256;

// Packed real FFT: z[n] = x[2n] + i x[2n+1] (M = NFFT/2 complex points), M = 16 * P with P = 8 (NFFT 256) or 16 (512).
// P threads own a frame.  Pass 1: thread n2 holds z[P*n1 + n2], n1 < 16, does a radix-16 FFT in registers and
// multiplies by W_M^(n2*k1) (kept in registers).  One shared-memory exchange.  Pass 2: thread j does the radix-P FFT
// over n2 for k1 = j (+ j+8 when P = 8): every lane is busy in both passes.  The split X[k] = E[k] - i w_k D[k] needs
// Z[M-k], which lives in lane (P-j) mod P of the same group: fetched with warp shuffles, no second exchange.
// The CTA is persistent over tiles so the per-thread twiddles / tables are built once.
template <int NFFT>
struct StftCfg {
  static constexpr int M = NFFT / 2, F = M + 1, P = M / 16, NA = 16 / P, G = STFT_THREADS / P;
  static constexpr int SCR = 16 * (P + 1) + (P == 8 ? 8 : 0);     // float2 per frame group (padded rows; the +8
                                                                  // puts the two groups of a half warp on disjoint banks)
  static constexpr int PITCH = P == 8 ? STFT_TILE_T + 4 : STFT_TILE_T + 2;   // staging pitch: conflict-free writes
};

// MUFU.SQRT, ~1 ulp: far inside the 1e-4 feature tolerance and branch-free (sqrtf carries a slow-path call)
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

template <int P>
__device__ __forceinline__ void fft_p(float2* v) {
  if (P == 8) fft8(v); else fft16(v);
}

template <int NFFT, bool COMPLEX, bool LAYOUT_TF, bool LOGMAG>
__global__ void __launch_bounds__(STFT_THREADS, 2) stft_kernel(const float* __restrict__ sig,
                                                              const long long* __restrict__ sig_off,
                                                              const long long* __restrict__ frame_off,
                                                              const int* __restrict__ tile_off, int n_utts,
                                                              int total_tiles, int hop, float* __restrict__ out) {
  using C = StftCfg<NFFT>;
  constexpr int M = C::M, F = C::F, P = C::P, NA = C::NA, G = C::G, TT = STFT_TILE_T, PITCH = C::PITCH;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* wtab = reinterpret_cast<float2*>(smem_raw);          // [M] window pairs (w[2n], w[2n+1])
  float2* ptw = wtab + M;                                      // [M] exp(-i pi k / M)
  float2* scr_all = ptw + M;                                   // [G][SCR]
  float* stage = reinterpret_cast<float*>(scr_all + G * C::SCR);          // [F][PITCH] (float or float2)
  const int span = (TT - 1) * hop + NFFT;
  float* sigs = stage + F * PITCH * (COMPLEX ? 2 : 1);

  const int tid = threadIdx.x, lane = tid & 31;
  const int g = tid / P, j = tid % P;
  for (int n = tid; n < M; n += STFT_THREADS) {
    float s, c;
    sincospif((float)n / (float)M, &s, &c);                    // angle pi*n/M = 2*pi*n/NFFT
    ptw[n] = make_float2(0.5f * c, -0.5f * s);                  // 0.5 * exp(-i pi n / M): the 1/2 of the split folded in
    float c0 = cospif(2.0f * (float)(2 * n) / (float)NFFT), c1 = cospif(2.0f * (float)(2 * n + 1) / (float)NFFT);
    wtab[n] = make_float2(0.5f - 0.5f * c0, 0.5f - 0.5f * c1);
  }
  float2 tw1[16];                                              // W_M^(j*k1)
#pragma unroll
  for (int k1 = 0; k1 < 16; ++k1) {
    float s, c;
    sincospif(2.0f * (float)(j * k1) / (float)M, &s, &c);
    tw1[k1] = make_float2(c, -s);
  }
  float2* scr = scr_all + g * C::SCR;
  const int src_lane = (lane & ~(P - 1)) | ((P - j) & (P - 1));
  (void)span;

  // contiguous tile range per CTA: the utterance index only ever advances, no search per tile
  const int tile_begin = (int)((long long)total_tiles * blockIdx.x / gridDim.x);
  const int tile_end = (int)((long long)total_tiles * (blockIdx.x + 1) / gridDim.x);
  int u = 0;
  {
    int lo = 0, hi = n_utts;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (tile_off[mid] <= tile_begin) lo = mid; else hi = mid;
    }
    u = lo;
  }
  constexpr int PF = NFFT == 256 ? 9 : 18;                                       // samples per thread prefetched in registers
  float pre[PF];
  // loads the span of `tile` (utterance uu) into registers: reflect padding as np.pad(y, n_fft//2, 'reflect')
  auto prefetch = [&](int tile, int uu) {
    const int t0 = (tile - tile_off[uu]) * TT;
    const long long s0 = sig_off[uu];
    const int n_samp = (int)(sig_off[uu + 1] - s0);
    const int T = (int)(frame_off[uu + 1] - frame_off[uu]);
    const int used = (min(TT, T - t0) - 1) * hop + NFFT;
    const int p0 = t0 * hop - NFFT / 2;
#pragma unroll
    for (int i = 0; i < PF; ++i) {
      const int idx = tid + i * STFT_THREADS;
      int q = p0 + idx;
      if (q < 0) q = -q;
      if (q >= n_samp) q = 2 * (n_samp - 1) - q;
      pre[i] = idx < used ? sig[s0 + q] : 0.0f;
    }
  };
  if (tile_begin < tile_end) prefetch(tile_begin, u);

  for (int tile = tile_begin; tile < tile_end; ++tile) {
    const int t0 = (tile - tile_off[u]) * TT;
    const long long s0 = sig_off[u];
    const int n_samp = (int)(sig_off[u + 1] - s0);
    const int T = (int)(frame_off[u + 1] - frame_off[u]);
    const int nt = min(TT, T - t0);
    const int used = (nt - 1) * hop + NFFT;
    __syncthreads();                                           // previous tile fully written out; tables ready
#pragma unroll
    for (int i = 0; i < PF; ++i) {
      const int idx = tid + i * STFT_THREADS;
      if (idx < used) sigs[idx] = pre[i];
    }
    {                                                          // spans longer than PF*256 samples (large hops)
      const int p0 = t0 * hop - NFFT / 2;
      for (int idx = tid + PF * STFT_THREADS; idx < used; idx += STFT_THREADS) {
        int q = p0 + idx;
        if (q < 0) q = -q;
        if (q >= n_samp) q = 2 * (n_samp - 1) - q;
        sigs[idx] = sig[s0 + q];
      }
    }
    __syncthreads();
    const int u_cur = u;
    if (tile + 1 < tile_end) {                                 // next tile's samples fly while this one is transformed
      while (tile_off[u + 1] <= tile + 1) ++u;
      prefetch(tile + 1, u);
    }
    float* out_u = out + frame_off[u_cur] * (long long)F * (COMPLEX ? 2 : 1);
#pragma unroll 1
    for (int fb = 0; fb < TT; fb += G) {
      if (fb >= nt) break;                                     // uniform
      const int fl = fb + g;
      const bool live = fl < nt;
      const float* x = sigs + (live ? fl : 0) * hop;
      float2 v[16];
#pragma unroll
      for (int n1 = 0; n1 < 16; ++n1) {
        const int n = P * n1 + j;
        const float2 xs = *reinterpret_cast<const float2*>(x + 2 * n);
        const float2 w = wtab[n];
        v[n1] = make_float2(xs.x * w.x, xs.y * w.y);
      }
      fft16(v);
      scr[j] = v[0];
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) scr[k1 * (P + 1) + j] = cmul(v[k1], tw1[k1]);
      __syncwarp();
      float2 z[NA][P], r[NA][P];
#pragma unroll
      for (int a = 0; a < NA; ++a) {
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) z[a][n2] = scr[(j + P * a) * (P + 1) + n2];
        fft_p<P>(z[a]);                                        // z[a][k2] = Z[(j + P*a) + 16*k2]
      }
      __syncwarp();                                            // scr may be rewritten by the next frame block
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int i = 0; i < P; ++i) {
          r[a][i].x = __shfl_sync(0xffffffffu, z[a][i].x, src_lane);
          r[a][i].y = __shfl_sync(0xffffffffu, z[a][i].y, src_lane);
        }
      if (!live) continue;
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
          const int k = j + P * a + 16 * k2;
          // partner Z[M-k]: lane (P-j)%P; for j >= 1 array NA-1-a, index P-1-k2; for j == 0 see the derivation above
          float2 b;
          if (NA == 2) {
            const float2 b0 = a == 0 ? r[0][(P - k2) & (P - 1)] : r[1][P - 1 - k2];
            const float2 b1 = r[1 - a][P - 1 - k2];
            b = j == 0 ? b0 : b1;
          } else {
            const float2 b0 = r[0][(P - k2) & (P - 1)], b1 = r[0][P - 1 - k2];
            b = j == 0 ? b0 : b1;
          }
          const float2 za = z[a][k2];
          const float2 e = make_float2(za.x + b.x, za.y - b.y);          // 2E = Z[k] + conj(Z[M-k])
          const float2 d = make_float2(za.x - b.x, za.y + b.y);          // 2D = Z[k] - conj(Z[M-k])
          const float2 wd = cmul(ptw[k], d);                             // (w_k / 2) * 2D
          const float2 X = make_float2(fmaf(0.5f, e.x, wd.y), fmaf(0.5f, e.y, -wd.x));   // E - i w_k D
          if (COMPLEX) {
            if (LAYOUT_TF) reinterpret_cast<float2*>(out_u)[(long long)(t0 + fl) * F + k] = X;
            else reinterpret_cast<float2*>(stage)[k * PITCH + fl] = X;
          } else {
            float m = sqrt_approx(fmaf(X.x, X.x, X.y * X.y));
            if (LOGMAG) m = logf(fmaxf(m, 1e-8f));
            if (LAYOUT_TF) out_u[(long long)(t0 + fl) * F + k] = m;
            else stage[k * PITCH + fl] = m;
          }
          if (k == 0) {                                        // Nyquist bin from Z[0]: X[M] = Re - Im
            const float2 XN = make_float2(za.x - za.y, 0.0f);
            if (COMPLEX) {
              if (LAYOUT_TF) reinterpret_cast<float2*>(out_u)[(long long)(t0 + fl) * F + M] = XN;
              else reinterpret_cast<float2*>(stage)[M * PITCH + fl] = XN;
            } else {
              float m = fabsf(XN.x);
              if (LOGMAG) m = logf(fmaxf(m, 1e-8f));
              if (LAYOUT_TF) out_u[(long long)(t0 + fl) * F + M] = m;
              else stage[M * PITCH + fl] = m;
            }
          }
        }
    }
    if (!LAYOUT_TF) {
      __syncthreads();
      const int warp = tid >> 5;
      if (lane < nt) {
        constexpr int NW = STFT_THREADS / 32;
        if (COMPLEX) {
          float2* o = reinterpret_cast<float2*>(out_u) + (long long)warp * T + t0 + lane;
          const float2* sgp = reinterpret_cast<const float2*>(stage) + warp * PITCH + lane;
          for (int k = warp; k < F; k += NW, o += (long long)NW * T, sgp += NW * PITCH) *o = *sgp;
        } else {
          float* o = out_u + (long long)warp * T + t0 + lane;
          const float* sgp = stage + warp * PITCH + lane;
          for (int k = warp; k < F; k += NW, o += (long long)NW * T, sgp += NW * PITCH) *o = *sgp;
        }
      }
    }
  }
}

template <int NFFT, bool COMPLEX, bool TF, bool LOGMAG>
static int launch_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                       int n_utts, int total_tiles, int hop, float* out, cudaStream_t st) {
  using C = StftCfg<NFFT>;
  const int span = (STFT_TILE_T - 1) * hop + NFFT;
  size_t smem = (size_t)(2 * C::M + C::G * C::SCR) * 8 + (size_t)C::F * C::PITCH * (COMPLEX ? 8 : 4) + (size_t)span * 4;
  auto kern = stft_kernel<NFFT, COMPLEX, TF, LOGMAG>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int grid = total_tiles < 148 * 2 ? total_tiles : 148 * 2;      // persistent: 2 CTAs per SM
  kern<<<grid, STFT_THREADS, smem, st>>>(sig, sig_off, frame_off, tile_off, n_utts, total_tiles, hop, out);
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
