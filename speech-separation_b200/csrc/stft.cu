// K1: batched STFT over a ragged corpus of utterances.
// Replaces steps/extract_feats.py:76-89,98,105 (librosa.core.stft [+ np.abs]) of the reference:
// reflect pad n_fft/2, frames at stride hop, periodic Hann, real FFT -> |X| (train) or complex (test).
#include "../../include/ssb200.h"
#include "fft.cuh"

namespace ssb {

constexpr int STFT_TILE_T = 32;      // frames per tile (ss_stft_tile_frames)
constexpr int STFT_HALF = 16;        // frame g of a tile rides in the lo lanes of the packed ops, frame 16 + g in the hi lanes

// Packed real FFT: z[n] = x[2n] + i x[2n+1] (M = NFFT/2 complex points), M = 16 * P with P = 8 (NFFT 256) or 16 (512).
// P threads own a PAIR of frames (g, 16 + g) of the tile and run both transforms at once on the fp32x2 pipe
// (fft.cuh, p2 / c2): the tile's samples sit in shared memory interleaved (A_i, B_i) so one 16-byte load yields
// (re_A, re_B, im_A, im_B) of z[n], and the window / twiddle tables hold every factor duplicated.
// Pass 1: thread n2 holds z[P*n1 + n2], n1 < 16, does a radix-16 FFT in registers and multiplies by W_M^(n2*k1).
// One shared-memory exchange (re then im through the same buffer).  Pass 2: thread j does the radix-P FFT over n2
// for k1 = j (+ j+8 when P = 8): every lane is busy in both passes.  The split X[k] = E[k] - i w_k D[k] needs
// Z[M-k], which another lane of the group holds: second exchange through the same buffer in natural order.
// The CTA is persistent over a contiguous range of tiles so the tables are built once.
template <int NFFT>
struct StftCfg {
  static constexpr int M = NFFT / 2, F = M + 1, P = M / 16, NA = 16 / P, G = STFT_HALF;
  static constexpr int THREADS = G * P, CTAS = 512 / THREADS;
  static constexpr int SCR = 16 * (P + 1) + (P == 8 ? 8 : 0);     // p2 per frame-pair group (padded rows; the +8
                                                                  // puts the two groups of a half warp on disjoint banks)
  static constexpr int PITCH = P == 8 ? STFT_TILE_T + 4 : STFT_TILE_T + 2;   // staging pitch: conflict-free writes
};

// MUFU.SQRT, ~1 ulp: far inside the 1e-4 feature tolerance and branch-free (sqrtf carries a slow-path call)
__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

#ifndef SSB_STFT_SHFL
#define SSB_STFT_SHFL 1   // partner bins by warp shuffle (0: through the shared-memory exchange buffer)
#endif

template <int P>
__device__ __forceinline__ void fft_p(c2* v) {
  if (P == 8) fft8(v); else fft16(v);
}

template <int NFFT, bool COMPLEX, bool LAYOUT_TF, bool LOGMAG>
__global__ void __launch_bounds__(StftCfg<NFFT>::THREADS, StftCfg<NFFT>::CTAS)
    stft_kernel(const float* __restrict__ sig, const long long* __restrict__ sig_off,
                const long long* __restrict__ frame_off, const int* __restrict__ tile_off, int n_utts, int total_tiles,
                int hop, float* __restrict__ out) {
  using C = StftCfg<NFFT>;
  constexpr int M = C::M, F = C::F, P = C::P, NA = C::NA, G = C::G, TT = STFT_TILE_T, PITCH = C::PITCH, NT = C::THREADS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float2* wtab = reinterpret_cast<float2*>(smem_raw);          // [M] window pairs (w[2n], w[2n+1]) / 2
  float2* ptw = wtab + M;                                      // [M] exp(-i pi k / M)
  float2* tw = ptw + M;                                        // [16][P] W_M^(j*k1)
  const int half_span = (STFT_HALF - 1) * hop + NFFT;          // samples one half of the tile reads
  float* sigs = reinterpret_cast<float*>(tw + 16 * P);         // [half_span][2]: (A_i, B_i), B = A shifted by 16 hops
  p2* scr_all = reinterpret_cast<p2*>(sigs + 2 * half_span);   // [G][SCR]
  float* stage = reinterpret_cast<float*>(scr_all + G * C::SCR);          // [F][PITCH] (float or float2)

  const int tid = threadIdx.x, lane = tid & 31;
  const int g = tid / P, j = tid % P;
  for (int n = tid; n < M; n += NT) {
    float s, c;
    sincospif((float)n / (float)M, &s, &c);                    // angle pi*n/M = 2*pi*n/NFFT
    ptw[n] = make_float2(c, -s);
    float c0 = cospif(2.0f * (float)(2 * n) / (float)NFFT), c1 = cospif(2.0f * (float)(2 * n + 1) / (float)NFFT);
    wtab[n] = make_float2(0.5f * (0.5f - 0.5f * c0), 0.5f * (0.5f - 0.5f * c1));   // Hann / 2
  }
  for (int e = tid; e < 16 * P; e += NT) {
    float s, c;
    sincospif(2.0f * (float)((e % P) * (e / P)) / (float)M, &s, &c);
    tw[e] = make_float2(c, -s);
  }
  p2* scr = scr_all + g * C::SCR;
  const int warp_g0 = (tid & ~31) / P;                         // first frame of this warp's groups
  const int boff = STFT_HALF * hop;

  // contiguous tile range per CTA: the utterance cursor only ever advances, and its offsets stay in registers
  // until the range crosses into the next utterance (no dependent global loads per tile)
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
  int u_tlo, u_thi, u_nsamp, u_T;                              // cursor = utterance of the tile being filled
  long long u_s0, u_fo;
  auto load_utt = [&]() {
    u_tlo = tile_off[u];
    u_thi = tile_off[u + 1];
    u_s0 = sig_off[u];
    u_nsamp = (int)(sig_off[u + 1] - u_s0);
    u_fo = frame_off[u];
    u_T = (int)(frame_off[u + 1] - u_fo);
  };
  load_utt();

  // tile samples -> shared memory with 4-byte cp.async: no registers held while they fly, so the next tile's fill is
  // issued before the current tile is written out.  Reflect padding as np.pad(y, n_fft//2, 'reflect').
  const uint32_t sig_s = smem_u32(sigs);
  int f_t0 = 0;                                                // frame offset of the filled tile inside its utterance
  auto fill = [&](int tile_) {
    while (tile_ >= u_thi) {
      ++u;
      load_utt();
    }
    f_t0 = (tile_ - u_tlo) * TT;
    const int used_ = (min(TT, u_T - f_t0) - 1) * hop + NFFT;
    const int p0_ = f_t0 * hop - NFFT / 2;
    const float* sp = sig + u_s0 + p0_ + tid;
    uint32_t d = sig_s + 8 * tid;
    if (p0_ >= 0 && p0_ + used_ <= u_nsamp) {                  // interior tile: straight copy
#pragma unroll 2
      for (int idx = tid; idx < used_; idx += NT, sp += NT, d += 8 * NT) {
        if (idx < half_span) cp_async4(d, sp);
        if (idx >= boff) cp_async4(d - 8 * boff + 4, sp);
      }
    } else {
      for (int idx = tid; idx < used_; idx += NT, d += 8 * NT) {
        int q = p0_ + idx;
        if (q < 0) q = -q;
        if (q >= u_nsamp) q = 2 * (u_nsamp - 1) - q;
        if (idx < half_span) cp_async4(d, sig + u_s0 + q);
        if (idx >= boff) cp_async4(d - 8 * boff + 4, sig + u_s0 + q);
      }
    }
    cp_async_commit();
  };
  if (tile_begin < tile_end) fill(tile_begin);

  for (int tile = tile_begin; tile < tile_end; ++tile) {
    const int t0 = f_t0, T = u_T;                              // the cursor still points at this tile's utterance
    const int nt = min(TT, T - t0);
    float* out_u = out + u_fo * (long long)F * (COMPLEX ? 2 : 1);
    cp_async_wait_all();
    __syncthreads();                                           // samples landed; previous tile fully written out
    if (warp_g0 < nt) {
      c2 v[16];
      {
      const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(sigs) + g * (hop >> 1) + j;
#pragma unroll
      for (int n1 = 0; n1 < 16; ++n1) {
        const ulonglong2 xs = sp[P * n1];
        const float2 w = wtab[P * n1 + j];
        v[n1] = {scale2(xs.x, w.x), scale2(xs.y, w.y)};
      }
    }
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
      __syncwarp();
#pragma unroll
      for (int a = 0; a < NA; ++a) fft_p<P>(z[a]);             // z[a][k2] = Z[(j + P*a) + 16*k2]

      // Bins k and M-k share 2E = Z[k] + conj(Z[M-k]), 2D = Z[k] - conj(Z[M-k]) and the product w_k D:
      //   X[k] = E - i w_k D,  X[M-k] = conj(E) - i conj(w_k D)
      // so each thread finishes the pairs of its own bins below M/2 (k2 < P/2) and only the upper bins travel
      // through the exchange buffer (natural order, re in [0, M/2), im in [M/2, M)).  The window table carries the
      // 1/2 of E and D.
#if !SSB_STFT_SHFL
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int k2 = P / 2; k2 < P; ++k2) {
          scr[j + P * a + 16 * k2 - M / 2] = z[a][k2].re;
          scr[j + P * a + 16 * k2] = z[a][k2].im;
        }
      __syncwarp();
#else
      // The partner bin M - k of k = j + P a + 16 k2 lives in lane P - j of the same group as z[NA-1-a][P-1-k2]
      // (j != 0), or in lane 0 itself as z[a][P - k2 - a] (j == 0): two 64-bit shuffles replace the trip through
      // the exchange buffer (8 of the kernel's 63 shared-memory wavefronts per frame).
      const int src_lane = (lane & ~(P - 1)) | ((P - j) & (P - 1));
#endif
      const bool liveA = g < nt, liveB = STFT_HALF + g < nt;
      const int fa = g, fb = STFT_HALF + g;
      // one output bin of both frames
      auto emit = [&](int k, p2 Xr, p2 Xi) {
        if (COMPLEX) {
          float ra, rb, ia, ib;
          upk(Xr, ra, rb);
          upk(Xi, ia, ib);
          if (LAYOUT_TF) {
            if (liveA) reinterpret_cast<float2*>(out_u)[(long long)(t0 + fa) * F + k] = make_float2(ra, ia);
            if (liveB) reinterpret_cast<float2*>(out_u)[(long long)(t0 + fb) * F + k] = make_float2(rb, ib);
          } else {
            reinterpret_cast<float2*>(stage)[k * PITCH + fa] = make_float2(ra, ia);
            reinterpret_cast<float2*>(stage)[k * PITCH + fb] = make_float2(rb, ib);
          }
        } else {
          float ma, mb;
          upk(fma2(Xr, Xr, mul2(Xi, Xi)), ma, mb);
          ma = sqrt_approx(ma);
          mb = sqrt_approx(mb);
          if (LOGMAG) {
            ma = logf(fmaxf(ma, 1e-8f));
            mb = logf(fmaxf(mb, 1e-8f));
          }
          if (LAYOUT_TF) {
            if (liveA) out_u[(long long)(t0 + fa) * F + k] = ma;
            if (liveB) out_u[(long long)(t0 + fb) * F + k] = mb;
          } else {
            stage[k * PITCH + fa] = ma;
            stage[k * PITCH + fb] = mb;
          }
        }
      };
#pragma unroll
      for (int a = 0; a < NA; ++a)
#pragma unroll
        for (int k2 = 0; k2 < P / 2; ++k2) {
          const int k = j + P * a + 16 * k2;                   // 0 <= k < M/2; the partner bin is M - k
#if !SSB_STFT_SHFL
          const p2 bx = scr[(M / 2 - k) & (M / 2 - 1)], by = scr[M / 2 + ((M / 2 - k) & (M / 2 - 1))];
#else
          const c2 zo = z[NA - 1 - a][P - 1 - k2], zz = z[a][(P - k2 - a) & (P - 1)];   // (k = 0 reads junk, unused)
          const p2 bx = __shfl_sync(0xffffffffu, j == 0 ? zz.re : zo.re, src_lane);
          const p2 by = __shfl_sync(0xffffffffu, j == 0 ? zz.im : zo.im, src_lane);
#endif
          const p2 zx = z[a][k2].re, zy = z[a][k2].im;
          const p2 ex = add2(zx, bx), dx = sub2(zx, bx), ey = sub2(zy, by), dy = add2(zy, by);
          const float2 w = ptw[k];
          const c2 wd = cmul_s(c2{dx, dy}, w.x, w.y);          // w_k D
          const p2 wr = wd.re, wi = wd.im;
          p2 Xr = add2(ex, wi), Xi = sub2(ey, wr);             // X[k]
          p2 Yr = sub2(ex, wi), Yi = neg2(add2(ey, wr));       // X[M-k]
          if (a == 0 && k2 == 0) {                             // k == 0 (lane j == 0): X[0] = Re Z + Im Z, X[M] = Re Z - Im Z
            const p2 s = add2(zx, zy), d = sub2(zx, zy);
            const bool k0 = j == 0;
            Xr = k0 ? add2(s, s) : Xr;
            Xi = k0 ? 0ull : Xi;
            Yr = k0 ? add2(d, d) : Yr;
            Yi = k0 ? 0ull : Yi;
          }
          emit(k, Xr, Xi);
          emit(M - k, Yr, Yi);
        }
      if (j == 0) {                                            // k = M/2 pairs with itself: X = conj(Z)
        const p2 zx = z[0][P / 2].re, zy = z[0][P / 2].im;
        emit(M / 2, add2(zx, zx), neg2(add2(zy, zy)));
      }
    }
    __syncthreads();                                           // every warp is done with the samples; stage complete
    const int T_cur = T;
    if (tile + 1 < tile_end) fill(tile + 1);                    // next tile's samples fly during the write-out
    if (!LAYOUT_TF) {
      const int warp = tid >> 5;
      if (lane < nt) {
        constexpr int NW = NT / 32;
        if (COMPLEX) {
          float2* o = reinterpret_cast<float2*>(out_u) + (long long)warp * T_cur + t0 + lane;
          const float2* sgp = reinterpret_cast<const float2*>(stage) + warp * PITCH + lane;
#pragma unroll 4
          for (int k = warp; k < F; k += NW, o += (long long)NW * T_cur, sgp += NW * PITCH) *o = *sgp;
        } else {
          float* o = out_u + (long long)warp * T_cur + t0 + lane;
          const float* sgp = stage + warp * PITCH + lane;
#pragma unroll 4
          for (int k = warp; k < F; k += NW, o += (long long)NW * T_cur, sgp += NW * PITCH) *o = *sgp;
        }
      }
    }
  }
}

template <int NFFT, bool COMPLEX, bool TF, bool LOGMAG>
static int launch_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                       int n_utts, int total_tiles, int hop, float* out, cudaStream_t st) {
  using C = StftCfg<NFFT>;
  const int half_span = (STFT_HALF - 1) * hop + NFFT;
  size_t smem = (size_t)(2 * C::M + 16 * C::P) * 8 + (size_t)half_span * 8 + (size_t)C::G * C::SCR * 8 +
                (size_t)C::F * C::PITCH * (COMPLEX ? 8 : 4);
  auto kern = stft_kernel<NFFT, COMPLEX, TF, LOGMAG>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int resident = ssb::sm_count() * C::CTAS;                                  // persistent: every SM full
  const int grid = total_tiles < resident ? total_tiles : resident;
  kern<<<grid, C::THREADS, smem, st>>>(sig, sig_off, frame_off, tile_off, n_utts, total_tiles, hop, out);
  return check_launch("ss_stft");
}

}  // namespace ssb

extern "C" int ss_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off,
                       int n_utts, int total_tiles, int n_fft, int hop, int flags, void* out, void* stream) {
  using namespace ssb;
  SS_REQUIRE(n_fft >= 4 && n_fft % 2 == 0, "ss_stft: n_fft must be even (got %d)", n_fft);
  SS_REQUIRE(hop > 0 && hop <= n_fft, "ss_stft: bad hop %d", hop);
  if (total_tiles == 0) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  float* o = (float*)out;
  const bool cplx = flags & SS_STFT_COMPLEX, tf = flags & SS_STFT_LAYOUT_TF, lg = flags & SS_STFT_LOG;
  SS_REQUIRE(!(cplx && lg), "ss_stft: LOG applies to magnitude output only");
  if ((n_fft != 512 && n_fft != 256) || hop % 2 != 0)       // any other even size / odd hop: direct-summation kernel
    return launch_stft_generic(sig, sig_off, frame_off, tile_off, n_utts, total_tiles, n_fft, hop, flags, o, st);
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
