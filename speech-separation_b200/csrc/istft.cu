// K7: fused mask * mixture-spectrum -> inverse real FFT -> Hann synthesis window -> overlap-add ->
// window-sum-square normalisation -> trim -> *32767 -> int16.
// Replaces steps/reconstruct_sources.py:35-42 (np.multiply + librosa.core.istft + astype('int16')).
#include "../../include/ssb200.h"
#include "fft.cuh"

namespace ssb {

constexpr int IST_TILE_T = 28;     // hops of output per tile (ss_istft_tile_hops): at most 32 frames contribute
constexpr int IST_HALF = 16;       // frame g of the tile rides in the lo lanes of the packed ops, frame 16 + g in the hi lanes

// Same machinery as stft.cu run backwards: P = NFFT/32 threads own a PAIR of frames (g, 16 + g) of the (up to) 32
// frames that touch the tile's 28 output hops and run both inverse transforms on the packed fp32x2 pipe.
//   1. X = spec * mask is staged as two planes Re[k][g], Im[k][g] of (frame g, frame 16+g) pairs (coalesced loads
//      along t in the (F,T) file layout);
//   2. thread j rebuilds conj(2 Z[k]) for k = P*n1 + j from X[k] and X[M-k] (2E = X[k] + conj X[M-k],
//      2D = X[k] - conj X[M-k], 2Z = 2E + i conj(w_k) 2D), which is pass-1 input order of the forward FFT;
//   3. radix-16 in registers, twiddle, one shared-memory exchange (re then im), radix-P: r = FFT(conj 2Z), so
//      x[2n] = Re r[n] / 2M, x[2n+1] = -Im r[n] / 2M; the synthesis-window table carries the 1/2M and the sign;
//   4. windowed frames go to shared memory (aliasing the X planes), then overlap-add in ascending frame order like
//      the reference's python loop, divide by the window-sum-square (table for the fully overlapped interior),
//      *32767 and the truncating int16 cast, two samples per thread.
// Persistent CTAs walk (tile, source) work items so the tables are built once and the second source's spectrum
// reads hit L2.
template <int NFFT>
struct IstftCfg {
  static constexpr int M = NFFT / 2, F = M + 1, P = M / 16, NA = 16 / P, G = IST_HALF;
  static constexpr int THREADS = G * P, CTAS = 512 / THREADS;
  static constexpr int SP = 32 + (P == 8 ? 2 : 1);                 // float2 per row of the staged spectrum (32 frames + skew)
  static constexpr int MP = 32 + (P == 8 ? 4 : 2);                 // floats per row of the staged mask
  static constexpr int SCR = 16 * (P + 1) + (P == 8 ? 8 : 0);      // p2 per frame-pair group of the FFT exchange
  static constexpr int FP = NFFT + 16;                             // floats per synthesised frame (bank skew)
  static constexpr int PLANES = F * SP * 8 + F * MP * 4;           // bytes: staged spectrum + mask
  static constexpr int WORK = G * SCR * 8 + 2 * IST_HALF * FP * 4; // bytes: exchange + frames (the planes alias it)
  static constexpr int REGION = PLANES > WORK ? PLANES : WORK;
};

template <int P>
__device__ __forceinline__ void ifft_p(c2* v) {
  if (P == 8) fft8(v); else fft16(v);
}

template <int NFFT, bool LAYOUT_TF, bool OUT_F32>
__global__ void __launch_bounds__(IstftCfg<NFFT>::THREADS, IstftCfg<NFFT>::CTAS)
    istft_kernel(const float2* __restrict__ spec, const float* __restrict__ mask, const long long* __restrict__ frame_off,
                 const long long* __restrict__ wav_off, const int* __restrict__ tile_off, int n_utts, int n_src,
                 int total_tiles, int hop, void* __restrict__ out) {
  using C = IstftCfg<NFFT>;
  constexpr int M = C::M, F = C::F, P = C::P, NA = C::NA, G = C::G, NT = C::THREADS, SP = C::SP, MP = C::MP, FP = C::FP;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* ptw = reinterpret_cast<float2*>(smem_raw);           // [M] exp(+i pi k / M) = conj(w_k)
  float2* tw = ptw + M;                                        // [16][P] W_M^(j*k1)
  float2* wsyn = tw + 16 * P;                                  // [M] (hann[2n], -hann[2n+1]) / 2M
  float* hsq = reinterpret_cast<float*>(wsyn + M);             // [NFFT] hann^2
  float* wfull = hsq + NFFT;                                   // [hop] window-sum-square of the interior, by q mod hop
  float* rfull = wfull + ((hop + 3) & ~3);                     // [hop] its correctly rounded reciprocal
  unsigned char* region = reinterpret_cast<unsigned char*>(rfull + ((hop + 3) & ~3));
  float2* rs = reinterpret_cast<float2*>(region);              // [F][SP] mixture spectrum of the tile's frames
  float* rm = reinterpret_cast<float*>(rs + F * SP);           // [F][MP] mask
  p2* scr_all = reinterpret_cast<p2*>(region);                 // [G][SCR]   (after the planes are consumed)
  float* frames = reinterpret_cast<float*>(scr_all + G * C::SCR);          // [32][FP]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = tid / P, j = tid % P;
  for (int n = tid; n < M; n += NT) {
    float s, c;
    sincospif((float)n / (float)M, &s, &c);
    ptw[n] = make_float2(c, s);
    const float h0 = 0.5f - 0.5f * cospif(2.0f * (float)(2 * n) / (float)NFFT);
    const float h1 = 0.5f - 0.5f * cospif(2.0f * (float)(2 * n + 1) / (float)NFFT);
    wsyn[n] = make_float2(h0 * (0.5f / (float)M), -h1 * (0.5f / (float)M));
    hsq[2 * n] = h0 * h0;
    hsq[2 * n + 1] = h1 * h1;
  }
  for (int e = tid; e < 16 * P; e += NT) {
    float s, c;
    sincospif(2.0f * (float)((e % P) * (e / P)) / (float)M, &s, &c);
    tw[e] = make_float2(c, -s);
  }
  __syncthreads();
  for (int r = tid; r < hop; r += NT) {                        // same ascending-frame order as the general path
    float s = 0.0f;
    for (int n = r + ((NFFT - 1 - r) / hop) * hop; n >= 0; n -= hop) s += hsq[n];
    const bool tiny = !(s > 1.17549435e-38f);                   // the reference leaves such samples undivided
    wfull[r] = tiny ? 1.0f : s;
    rfull[r] = tiny ? 1.0f : 1.0f / s;
  }
  p2* scr = scr_all + g * C::SCR;
  const int n_work = total_tiles * n_src;

  // contiguous range of (tile, source) items per CTA: the utterance cursor only ever advances and its offsets stay
  // in registers; the sources of a tile are consecutive items, so the second one reads the spectrum from L2
  const int w_begin = (int)((long long)n_work * blockIdx.x / gridDim.x);
  const int w_end = (int)((long long)n_work * (blockIdx.x + 1) / gridDim.x);
  int u = 0;
  if (w_begin < w_end) {
    const int t_first = w_begin / n_src;
    int lo = 0, hi = n_utts;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (tile_off[mid] <= t_first) lo = mid; else hi = mid;
    }
    u = lo;
  }
  int u_tlo = tile_off[u], u_thi = tile_off[u + 1];
  long long fo = frame_off[u], wo = wav_off[u];
  int T = (int)(frame_off[u + 1] - fo);
  for (int work = w_begin; work < w_end; ++work) {
    const int tile = work / n_src, src = work - tile * n_src;
    while (tile >= u_thi) {
      ++u;
      u_tlo = u_thi;
      u_thi = tile_off[u + 1];
      fo = frame_off[u];
      wo = wav_off[u];
      T = (int)(frame_off[u + 1] - fo);
    }
    const int L = hop * (T - 1);                               // trimmed output length
    const int j0 = (tile - u_tlo) * IST_TILE_T * hop;
    const int nj = min(IST_TILE_T * hop, L - j0);
    const int q0 = j0 + NFFT / 2, q1 = q0 + nj - 1;
    const int i_lo = max(0, (q0 - NFFT + hop) / hop);           // ceil((q0-NFFT+1)/hop)
    const int i_hi = min(T - 1, q1 / hop);
    const int nf = i_hi - i_lo + 1;                             // <= 32
    const float2* spec_u = spec + fo * F;
    const float* mask_u = mask + (fo * n_src + (long long)src * T) * F;

    __syncthreads();                                           // previous item's frames consumed; tables ready
    // 1. spectrum and mask of the tile's frames -> shared memory by cp.async: every load of the item is in flight at
    //    once and no register is held (the other CTA of the SM computes meanwhile)
    {
      const uint32_t rs_s = smem_u32(rs), rm_s = smem_u32(rm);
      if (LAYOUT_TF) {
        for (int e = tid; e < nf * F; e += NT) {
          const int f = e / F, k = e - f * F;
          const long long gi = (long long)(i_lo + f) * F + k;
          cp_async8(rs_s + (k * SP + f) * 8, spec_u + gi);
          cp_async4(rm_s + (k * MP + f) * 4, mask_u + gi);
        }
      } else if (lane < nf) {
        constexpr int NW = NT / 32;
        const float2* sp = spec_u + (long long)warp * T + i_lo + lane;
        const float* mp = mask_u + (long long)warp * T + i_lo + lane;
        const long long rstride = (long long)NW * T;
        uint32_t ds = rs_s + (warp * SP + lane) * 8, dm = rm_s + (warp * MP + lane) * 4;
#pragma unroll 4
        for (int k = warp; k < F; k += NW, sp += rstride, mp += rstride, ds += NW * SP * 8, dm += NW * MP * 4) {
          cp_async8(ds, sp);
          cp_async4(dm, mp);
        }
      }
      cp_async_commit();
      cp_async_wait_all();
    }
    __syncthreads();
    const int warp_g0 = (tid & ~31) / P;
    const bool warp_live = warp_g0 < nf;
    c2 v[16];
    if (warp_live) {
      // 2. conj(2 Z[k]), k = P*n1 + j
#pragma unroll
      for (int n1 = 0; n1 < 16; ++n1) {
        const int k = P * n1 + j;
        // X = spec * mask of frames g (lo) and 16+g (hi), rows k and M-k
        const float2 sa = rs[k * SP + g], sb = rs[k * SP + IST_HALF + g];
        const float ma = rm[k * MP + g], mb = rm[k * MP + IST_HALF + g];
        const float2 ta = rs[(M - k) * SP + g], tb = rs[(M - k) * SP + IST_HALF + g];
        const float na = rm[(M - k) * MP + g], nb = rm[(M - k) * MP + IST_HALF + g];
        p2 ax = pk(sa.x * ma, sb.x * mb), ay = pk(sa.y * ma, sb.y * mb);
        p2 bx = pk(ta.x * na, tb.x * nb), by = pk(ta.y * na, tb.y * nb);
        if (n1 == 0 && j == 0) {                               // irfft ignores Im of DC and Nyquist
          ay = 0ull;
          by = 0ull;
        }
        const p2 ex = add2(ax, bx), ey = sub2(ay, by), dx = sub2(ax, bx), dy = add2(ay, by);
        const float2 w = ptw[k];
        const c2 o = cmul_s(c2{dx, dy}, w.x, w.y);             // conj(w_k) * 2D
        v[n1] = {sub2(ex, o.im), neg2(add2(ey, o.re))};        // conj(2E + i o)
      }
    }
    __syncthreads();                                           // planes consumed -> exchange buffer + frames
    if (warp_live) {
      // 3. forward FFT of conj(2Z)
      fft16(v);
#pragma unroll
      for (int k1 = 1; k1 < 16; ++k1) {
        const float2 w = tw[k1 * P + j];
        v[k1] = cmul_s(v[k1], w.x, w.y);
      }
      c2 z[NA][P];
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) scr[k1 * (P + 1) + j] = v[k1].re;
      __syncwarp();
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) z[a][n2].re = scr[(j + P * a) * (P + 1) + n2];
      __syncwarp();
#pragma unroll
      for (int k1 = 0; k1 < 16; ++k1) scr[k1 * (P + 1) + j] = v[k1].im;
      __syncwarp();
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int n2 = 0; n2 < P; ++n2) z[a][n2].im = scr[(j + P * a) * (P + 1) + n2];
#pragma unroll
      for (int a = 0; a < NA; ++a) ifft_p<P>(z[a]);            // z[a][k2] = r[(j + P*a) + 16*k2]
      // 4. synthesis window, frames to shared memory
      float* fa = frames + g * FP;
      float* fb = frames + (IST_HALF + g) * FP;
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int k2 = 0; k2 < P; ++k2) {
          const int n = j + P * a + 16 * k2;
          const float2 w = wsyn[n];
          float xa, xb, ya, yb;
          upk(z[a][k2].re, xa, xb);
          upk(z[a][k2].im, ya, yb);
          *reinterpret_cast<float2*>(fa + 2 * n) = make_float2(xa * w.x, ya * w.y);
          *reinterpret_cast<float2*>(fb + 2 * n) = make_float2(xb * w.x, yb * w.y);
        }
    }
    __syncthreads();
    // 5. overlap-add in ascending frame order, normalise, convert; two samples per thread (hop, L, offsets are even)
    const long long obase = wo + (long long)src * L + j0;
    // q = hb*hop + r walks forward by 2*NT samples per iteration: one division per thread and tile
    int hb = (q0 + 2 * tid) / hop, r = (q0 + 2 * tid) - hb * hop;
    const int dstep = (2 * NT) / hop, rstep = (2 * NT) - dstep * hop;
    const int nov = (NFFT - 1) / hop;                           // frames hb-nov .. hb can reach sample q
    const bool hop4 = 4 * hop == NFFT;
    for (int jj = 2 * tid; jj < nj; jj += 2 * NT) {
      int i = hb - nov, n = r + nov * hop;                      // n = q - i*hop
      if (n >= NFFT) {                                          // (hop does not divide NFFT: first frame may miss)
        ++i;
        n -= hop;
      }
      const bool interior = i >= 0 && hb <= T - 1;              // neither end of the frame range is clipped
      const int iend = min(T - 1, hb);
      if (i < 0) {
        n += i * hop;
        i = 0;
      }
      float y0 = 0.0f, y1 = 0.0f;
      const float* fp = frames + (i - i_lo) * FP + n;
      if (interior) {
        if (hop4) {
          const float2 f0 = *reinterpret_cast<const float2*>(fp);
          const float2 f1 = *reinterpret_cast<const float2*>(fp + (FP - hop));
          const float2 f2 = *reinterpret_cast<const float2*>(fp + 2 * (FP - hop));
          const float2 f3 = *reinterpret_cast<const float2*>(fp + 3 * (FP - hop));
          y0 = ((f0.x + f1.x) + f2.x) + f3.x;
          y1 = ((f0.y + f1.y) + f2.y) + f3.y;
        } else {
          for (; i <= iend; ++i, fp += FP - hop) {
            const float2 fv = *reinterpret_cast<const float2*>(fp);
            y0 += fv.x;
            y1 += fv.y;
          }
        }
        // y / s from the tabulated s and 1/s: quotient estimate plus one residual correction (the division's own
        // fast path without its reciprocal refinement)
        const float2 sw = *reinterpret_cast<const float2*>(wfull + r), rc = *reinterpret_cast<const float2*>(rfull + r);
        const float e0 = y0 * rc.x, e1 = y1 * rc.y;
        y0 = fmaf(fmaf(-e0, sw.x, y0), rc.x, e0);
        y1 = fmaf(fmaf(-e1, sw.y, y1), rc.y, e1);
      } else {
        float s0 = 0.0f, s1 = 0.0f;
        for (; i <= iend; ++i, fp += FP - hop, n -= hop) {
          const float2 fv = *reinterpret_cast<const float2*>(fp);
          y0 += fv.x;
          y1 += fv.y;
          s0 += hsq[n];
          s1 += hsq[n + 1];
        }
        if (s0 > 1.17549435e-38f) y0 /= s0;
        if (s1 > 1.17549435e-38f) y1 /= s1;
      }
      hb += dstep;
      r += rstep;
      if (r >= hop) {
        r -= hop;
        ++hb;
      }
      if (OUT_F32) {
        *reinterpret_cast<float2*>(reinterpret_cast<float*>(out) + obase + jj) = make_float2(y0, y1);
      } else {
        short2 o2;
        o2.x = (short)(int)(y0 * 32767.0f);
        o2.y = (short)(int)(y1 * 32767.0f);
        *reinterpret_cast<short2*>(reinterpret_cast<short*>(out) + obase + jj) = o2;
      }
    }
  }
}

template <int NFFT, bool TF, bool F32>
static int launch_istft(const float2* spec, const float* mask, const long long* frame_off, const long long* wav_off,
                        const int* tile_off, int n_utts, int n_src, int total_tiles, int hop, void* out,
                        cudaStream_t st) {
  using C = IstftCfg<NFFT>;
  const size_t smem = (size_t)(2 * C::M + 16 * C::P) * 8 + (size_t)NFFT * 4 + (size_t)((hop + 3) & ~3) * 8 + C::REGION;
  auto kern = istft_kernel<NFFT, TF, F32>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const long long n_work = (long long)total_tiles * n_src;
  const int resident = ssb::sm_count() * C::CTAS;
  const int grid = n_work < resident ? (int)n_work : resident;
  kern<<<grid, C::THREADS, smem, st>>>(spec, mask, frame_off, wav_off, tile_off, n_utts, n_src, total_tiles, hop, out);
  return check_launch("ss_istft_masked");
}

}  // namespace ssb

extern "C" int ss_istft_masked(const void* spec, const float* mask, const long long* frame_off,
                               const long long* wav_off, const int* tile_off, int n_utts, int n_src, int total_tiles,
                               int n_fft, int hop, int flags, void* out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n_fft >= 4 && n_fft % 2 == 0, "ss_istft_masked: n_fft must be even (got %d)", n_fft);
  SS_REQUIRE(hop > 0 && hop <= n_fft, "ss_istft_masked: bad hop %d for n_fft %d", hop, n_fft);
  if (total_tiles == 0 || n_src == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if ((n_fft != 512 && n_fft != 256) || hop % 2 != 0 || hop * 4 < n_fft)   // sizes the packed-FFT kernel does not cover
    return launch_istft_generic(spec, mask, frame_off, wav_off, tile_off, n_utts, n_src, total_tiles,
                                ss_istft_tile_hops() * hop, n_fft, hop, flags, out, st);
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
