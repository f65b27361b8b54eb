// K1 on the tensor cores: 256-point / hop-64 magnitude STFT as two tcgen05 GEMMs against a resident, windowed
// DFT basis.  Replaces librosa.core.stft + np.abs at steps/extract_feats.py:87,89 (same contract as ss_stft with
// n_fft = 256, hop = 64, flags = 0; the FFT kernel in stft.cu stays the general path).
//
// Algebra.  With the periodic Hann window w[n] = w[256-n], w[0] = 0 and a real frame x[0..255]:
//   Re X[k] =  sum_{n=1..128} (w[n] cos(2 pi k n/256)) e[n],   e[n] = x[n] + x[256-n] (n < 128), e[128] = x[128]
//   Im X[k] = -sum_{n=1..127} (w[n] sin(2 pi k n/256)) o[n],   o[n] = x[n] - x[256-n]
// so a tile of 128 frames is  E[128 x 128] * C^T[128 x 128]  and  O[128 x 128] * S^T[128 x 128]  (K = n-1, bins 0..127;
// the Nyquist bin X[128] = sum_n (-1)^n w[n] e[n] is real and summed outside the GEMMs).  The
// window lives inside the basis (a diagonal scaling along K), the fold halves the flops of a plain 256-point DFT
// matrix, and every operand is split-bf16 (hi + lo) with the three products hi*hi + hi*lo + lo*hi accumulated in
// fp32 in TMEM (error ~2^-16, inside the 1e-4 feature bar).
//
// One persistent CTA per SM, 21 warps:
//   warps 0..15  producers: half a warp per frame and 64-wide K block; two LDG.128 (x[n..n+3] and the mirrored quad)
//                per lane, fold, split into hi/lo bf16, 8-byte stores into a 128B-swizzled K-major A stage.  The odd
//                part o stays packed in registers until the next ring slot is free, so every sample is loaded once
//                per frame.  The same warps sum the Nyquist bin, sum_n (-1)^n w[n] e[n], on the FMA pipe (16-lane
//                shuffle reduction per frame) so that both GEMMs are N = 128 and TMEM holds two full accumulator sets.
//   warp 16      one thread issues the MMAs (M=128, N=128, K=16; 48 per tile) and the commits
//   warps 17..20 epilogue: pipelined tcgen05.ld -> sqrt(re^2 + im^2) -> (F, T_u) rows of the utterance; a warp
//                stores 32 consecutive frames of one bin (128 contiguous bytes)
// Shared memory: the 128 KB basis image (cos hi | cos lo | sin hi | sin lo, bins 0..127, loaded once per CTA by bulk
// copies) and a ring of three 32 KB A stages that the fills E(kb0), O(kb0), E(kb1), O(kb1), ... go round.
// TMEM: 2 x (128 cos + 128 sin) columns; the epilogue drains tile i while the MMAs of tile i+1 run.
// Measured (profiles/r01_summary.md section K): parity <= 3e-5 against the oracle, 16.4 ms on configs[1] against the
// FFT kernel's 13.4 ms -- the MMAs take 5.8 ms alone, the producers 10.3 ms alone -- hence opt-in.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int D_TM = 128;                                 // frames per tile
#ifndef SSB_DFT_EXP
#define SSB_DFT_EXP 0   // timing experiments (wrong results): 1 no MMAs, 2 no producer work, 4 no epilogue stores
#endif
#ifndef SSB_DFT_EPI_WARPS
#define SSB_DFT_EPI_WARPS 4
#endif
constexpr int D_PROD_WARPS = 16, D_EPI_WARPS = SSB_DFT_EPI_WARPS;
constexpr int D_TASKS = D_TM * 16 / (D_PROD_WARPS * 32);  // (frame, 16-lane) tasks per producer thread and K block
constexpr int D_THREADS = (D_PROD_WARPS + 1 + D_EPI_WARPS) * 32;
constexpr int D_KB = 128 * 128;                           // bytes of one 64-wide K block of one basis plane (128 bins)
constexpr int D_PLANE = 2 * D_KB;
constexpr int D_BASIS = 4 * D_PLANE;                      // cos hi | cos lo | sin hi | sin lo = 128 KB
constexpr int D_APLANE = D_TM * 128;                      // one bf16 plane of an A stage: 128 frames x 64 k
constexpr int D_STAGE = 2 * D_APLANE;                     // hi + lo
constexpr int D_NSTAGE = 3;                               // A stages: fills E(g), O(g), E(g+1), ... go round the ring
constexpr int D_SMEM = D_BASIS + D_NSTAGE * D_STAGE + 256 /*barriers*/ + 512 /*Nyquist weights*/ + 1024 /*align*/;
static_assert(D_SMEM <= 232448, "shared memory budget");
static_assert(D_TASKS == 4, "producer mapping assumes 8 frames per warp");

struct DftMeta {
  long long s0, obase;   // first sample of the utterance in sig; first output element of the utterance
  int N, T, t0;          // samples, frames of the utterance; first frame of the tile
};
__device__ __forceinline__ DftMeta dft_meta(const longlong2* __restrict__ tiles, int tile) {
  const longlong2 a = __ldg(tiles + 2 * (long long)tile), b = __ldg(tiles + 2 * (long long)tile + 1);
  DftMeta m;
  m.s0 = a.x;
  m.obase = a.y;
  m.N = (int)(b.x & 0xffffffffll);
  m.T = (int)(b.x >> 32);
  m.t0 = (int)b.y;
  return m;
}

__device__ __forceinline__ float dft_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts64(uint32_t addr, uint2 v) {
  asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr), "r"(v.x), "r"(v.y) : "memory");
}
// four floats -> packed bf16 hi (two words) and the bf16 of the residuals
__device__ __forceinline__ void split4(const float* v, uint2& hi, uint2& lo) {
  const __nv_bfloat162 h01 = __floats2bfloat162_rn(v[0], v[1]), h23 = __floats2bfloat162_rn(v[2], v[3]);
  hi.x = *reinterpret_cast<const uint32_t*>(&h01);
  hi.y = *reinterpret_cast<const uint32_t*>(&h23);
  // bf16 -> f32 is a 16-bit shift: one instruction per element instead of extract + shift
  const __nv_bfloat162 l01 = __floats2bfloat162_rn(v[0] - __uint_as_float(hi.x << 16), v[1] - __uint_as_float(hi.x & 0xffff0000u));
  const __nv_bfloat162 l23 = __floats2bfloat162_rn(v[2] - __uint_as_float(hi.y << 16), v[3] - __uint_as_float(hi.y & 0xffff0000u));
  lo.x = *reinterpret_cast<const uint32_t*>(&l01);
  lo.y = *reinterpret_cast<const uint32_t*>(&l23);
}
// two 16-column TMEM loads, one wait
__device__ __forceinline__ void tmem_ld16x2(uint32_t a0, uint32_t a1, float* v0, float* v1) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%32];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%33];\n\t"
      "tcgen05.wait::ld.sync.aligned;"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(a0), "r"(a1)
      : "memory");
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    v0[i] = __uint_as_float(r[i]);
    v1[i] = __uint_as_float(r[16 + i]);
  }
}

// the same loads without the wait, and a wait that the 32 result registers depend on (software pipelining)
__device__ __forceinline__ void tmem_ld16x2_issue(uint32_t a0, uint32_t a1, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%32];\n\t"
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%33];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(a0), "r"(a1)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait(uint32_t* r) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                 "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                 "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
               :
               : "memory");
}

// Basis image in the exact shared-memory layout the MMA descriptors read (K-major rows of 128 B, 16-byte chunks
// XOR-swizzled with row%8, 8-row atoms 1024 B apart): plane = [K block 0: 128 bins][K block 1: 128 bins].
// Rows are bins 0..127 (the sin row of bin 0 is all zero); the Nyquist bin is summed by the producers.
__global__ void dft256_basis_kernel(unsigned char* __restrict__ img) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= 2 * 2 * 128 * 64) return;
  const bool is_sin = e >= 2 * 128 * 64;
  const int q = e & (2 * 128 * 64 - 1);
  const int kb = q >> 13, r = (q >> 6) & 127, kk = q & 63;
  const int n = kb * 64 + kk + 1;                                       // sample index 1..128
  const double w = 0.5 - 0.5 * cospi((double)n / 128.0);               // periodic Hann, n_fft = 256
  const int m = (r * n) & 255;                                          // angle 2 pi m / 256
  const double v = is_sin ? -w * sinpi((double)m / 128.0) : w * cospi((double)m / 128.0);
  const __nv_bfloat16 hi = __float2bfloat16_rn((float)v);
  const __nv_bfloat16 lo = __float2bfloat16_rn((float)(v - (double)__bfloat162float(hi)));
  const int off = kb * D_KB + r * 128 + ((((kk >> 3) ^ (r & 7))) << 4) + (kk & 7) * 2;
  unsigned char* base = img + (is_sin ? 2 * D_PLANE : 0);
  *reinterpret_cast<__nv_bfloat16*>(base + off) = hi;
  *reinterpret_cast<__nv_bfloat16*>(base + D_PLANE + off) = lo;
}

// waits of the many-warp roles back off so their polling does not take issue slots from the working warps
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
#ifdef SSB_DFT_NOSLEEP
  while (!mbar_try_wait(bar, parity)) {
  }
#else
  while (!mbar_try_wait(bar, parity)) __nanosleep(32);
#endif
}

__global__ void __launch_bounds__(D_THREADS, 1)
stft_dft256_kernel(const float* __restrict__ sig, const longlong2* __restrict__ tiles, int n_tiles,
                   const unsigned char* __restrict__ basis, float* __restrict__ out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* stages = smem + D_BASIS;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + D_BASIS + D_NSTAGE * D_STAGE);
  uint64_t* empty = full + D_NSTAGE;
  uint64_t* tfull = empty + D_NSTAGE;
  uint64_t* tempty = tfull + 2;
  uint64_t* basis_bar = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(basis_bar + 1);
  float* nyq_w = reinterpret_cast<float*>(smem + D_BASIS + D_NSTAGE * D_STAGE + 256);   // (-1)^n w[n], n = 1..128

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < D_NSTAGE; ++s) {
      mbar_init(&full[s], D_PROD_WARPS);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], D_EPI_WARPS);
    }
    mbar_init(basis_bar, 1);
    fence_barrier_init();
  }
  if (warp == D_PROD_WARPS) tmem_alloc<512>(tmem_slot);
  if (threadIdx.x < 128) {
    const int n = threadIdx.x + 1;
    const float w = 0.5f - 0.5f * cospif((float)n * (1.0f / 128.0f));
    nyq_w[threadIdx.x] = (n & 1) ? -w : w;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < D_PROD_WARPS) {
    // ------------------------------------------------------------------ producers
    const int half = lane >> 4, jq = lane & 15;
    const int rbase = warp * (2 * D_TASKS) + half;                      // frame of task i: rbase + 2 i
    float4 LA[D_TASKS], LB[D_TASKS];
    float EX[D_TASKS], NY[D_TASKS];
    uint2 OH[D_TASKS], OL[D_TASKS];
    uint32_t offs[D_TASKS];                                            // byte offset of this lane's 8 bytes in a plane
#pragma unroll
    for (int i = 0; i < D_TASKS; ++i) {
      const int r = rbase + 2 * i;
      offs[i] = (uint32_t)(r * 128 + ((((jq >> 1) ^ (r & 7))) << 4) + ((jq & 1) << 3));
    }
    const uint32_t st_base = smem_u32(stages), nyq_s = smem_u32(nyq_w);
    int fb = 0;                                                        // ring slot of the next fill
    uint32_t fph = 0;                                                  // its use parity
    bool wfast = false;

    // loads of one K block (samples nb+1 .. nb+64 and their mirrors) for this thread's four frames
    auto issue = [&](const DftMeta& m, int kb) {
      if (SSB_DFT_EXP & 2) return;
      const int nb = kb * 64;
      const int tf = m.t0 + warp * (2 * D_TASKS);                       // first of the warp's 8 frames
      const float* ub = sig + m.s0;
      wfast = tf >= 2 && (tf + 2 * D_TASKS - 1) * 64 + 128 <= m.N && (reinterpret_cast<uintptr_t>(ub) & 15) == 0;
      if (wfast) {
        const float* pa = ub + ((m.t0 + rbase) * 64 - 128 + 4 * jq);   // x[4 jq] of task 0's frame
        const float* pb = pa + (252 - 8 * jq);                         // x[252 - 4 jq]
#pragma unroll
        for (int i = 0; i < D_TASKS; ++i) {
          LA[i] = __ldg(reinterpret_cast<const float4*>(pa + nb + 128 * i));
          LB[i] = __ldg(reinterpret_cast<const float4*>(pb - nb + 128 * i));
        }
        if (kb == 0 && lane == 31) {
#pragma unroll
          for (int i = 0; i < D_TASKS; ++i) EX[i] = __ldg(pa + 4 + 128 * i);   // x[64]
        }
      } else {
        // reflect padding (np.pad(y, n_fft//2, 'reflect')), unaligned utterances, frames past the end
#pragma unroll
        for (int i = 0; i < D_TASKS; ++i) {
          const int t = m.t0 + rbase + 2 * i, start = t * 64 - 128;
          const bool valid = t < m.T;
          float xa[4], xb[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int n = nb + 4 * jq + 1 + q;
            int ia = start + n, ib = start + 256 - n;
            ia = ia < 0 ? -ia : ia;
            ib = ib < 0 ? -ib : ib;
            ia = ia >= m.N ? 2 * m.N - 2 - ia : ia;
            ib = ib >= m.N ? 2 * m.N - 2 - ib : ib;
            xa[q] = valid ? __ldg(ub + ia) : 0.0f;
            xb[q] = valid ? __ldg(ub + ib) : 0.0f;
          }
          LA[i] = make_float4(0.0f, xa[0], xa[1], xa[2]);
          EX[i] = xa[3];
          LB[i] = make_float4(xb[3], xb[2], xb[1], xb[0]);
        }
      }
    };

    int tile = blockIdx.x;
    DftMeta cur = {0, 0, 0, 0, 0}, nxt = cur, nn = cur;
    if (tile < n_tiles) {
      cur = dft_meta(tiles, tile);
      issue(cur, 0);
    }
    if (tile + (int)gridDim.x < n_tiles) nxt = dft_meta(tiles, tile + gridDim.x);
    int g = 0;                                                         // steps done: 2 per tile, one per K block
    for (; tile < n_tiles; tile += gridDim.x) {
      const int tile_n = tile + gridDim.x, tile_nn = tile_n + gridDim.x;
      if (tile_nn < n_tiles) nn = dft_meta(tiles, tile_nn);            // consumed one iteration later
      if (tile_n < n_tiles) {
        // the next tile's samples (32 KB + halo) go to L2 now: one 128-byte line per producer thread
        long long a = (long long)nxt.t0 * 64 - 128, b = a + (D_TM - 1) * 64 + 256;
        a = a < 0 ? 0 : a;
        b = b > nxt.N ? nxt.N : b;
        const char* p0 = reinterpret_cast<const char*>(sig + nxt.s0 + a);
        const long long off = (long long)threadIdx.x * 128;
        if (off < (b - a) * 4) asm volatile("prefetch.global.L2 [%0];" ::"l"(p0 + off));
      }
#pragma unroll
      for (int kb = 0; kb < 2; ++kb, ++g) {
        float cn[4];
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(cn[0]), "=f"(cn[1]), "=f"(cn[2]), "=f"(cn[3])
                     : "r"(nyq_s + (uint32_t)(kb * 64 + 4 * jq) * 4u));
        mbar_wait_backoff(&empty[fb], fph ^ 1);
        const uint32_t st0 = st_base + fb * D_STAGE;
#pragma unroll
        for (int i = 0; i < ((SSB_DFT_EXP & 2) ? 0 : D_TASKS); ++i) {
          float nx = __shfl_down_sync(0xffffffffu, LA[i].x, 1);
          if (!wfast) nx = EX[i];
          else if (lane == 31) nx = kb ? LB[i].x : EX[i];
          const float xa[4] = {LA[i].y, LA[i].z, LA[i].w, nx};
          const float xb[4] = {LB[i].w, LB[i].z, LB[i].y, LB[i].x};
          float ev[4], ov[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            ev[q] = xa[q] + xb[q];
            ov[q] = xa[q] - xb[q];
          }
          if (kb == 1 && jq == 15) {                                   // n = 128 is its own mirror
            ev[3] = xa[3];
            ov[3] = 0.0f;
          }
          float ny = kb ? NY[i] : 0.0f;                                // Nyquist bin: sum (-1)^n w[n] e[n]
#pragma unroll
          for (int q = 0; q < 4; ++q) ny = fmaf(cn[q], ev[q], ny);
          NY[i] = ny;
          uint2 eh, el;
          split4(ev, eh, el);
          split4(ov, OH[i], OL[i]);
          sts64(st0 + offs[i], eh);
          sts64(st0 + D_APLANE + offs[i], el);
        }
        fence_proxy_async_smem();                                      // own stores -> async proxy, then one arrival
        __syncwarp();                                                  // per warp (512 arrivals on one word serialise)
        if (lane == 0) mbar_arrive(&full[fb]);
        if (++fb == D_NSTAGE) { fb = 0; fph ^= 1; }
        // loads of the next step fly while the odd part is stored
        if (kb == 0) issue(cur, 1);
        else if (tile_n < n_tiles) issue(nxt, 0);
        mbar_wait_backoff(&empty[fb], fph ^ 1);
        const uint32_t st1 = st_base + fb * D_STAGE;
#pragma unroll
        for (int i = 0; i < ((SSB_DFT_EXP & 2) ? 0 : D_TASKS); ++i) {
          sts64(st1 + offs[i], OH[i]);
          sts64(st1 + D_APLANE + offs[i], OL[i]);
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(&full[fb]);
        if (++fb == D_NSTAGE) { fb = 0; fph ^= 1; }
      }
      // Nyquist bin of the four frames: 16-lane sums, written by the first lane of each half warp
      {
        float* onq = out + cur.obase + 128ll * cur.T;
#pragma unroll
        for (int i = 0; i < D_TASKS; ++i) {
          float v = NY[i];
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          const int t = cur.t0 + rbase + 2 * i;
          if (jq == 0 && t < cur.T) onq[t] = fabsf(v);
        }
      }
      cur = nxt;
      nxt = nn;
    }
  } else if (warp == D_PROD_WARPS) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      mbar_arrive_expect_tx(basis_bar, D_BASIS);
      constexpr int CH = D_BASIS / 8;
      for (int c = 0; c < 8; ++c) bulk_load(smem + c * CH, basis + c * CH, CH, basis_bar);
      mbar_wait(basis_bar, 0);
      constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
      const uint64_t dbase = umma_desc_sw128(smem_u32(smem));
      const uint64_t abase = dbase + (D_BASIS >> 4);
      int fb = 0, lt = 0;
      uint32_t fph = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(&tempty[buf], ((lt >> 1) & 1) ^ 1);                   // epilogue drained this pair of accumulators
        tc_fence_after();
        const uint32_t acc_c = tmem_base + buf * 256, acc_s = acc_c + 128;
#pragma unroll
        for (int kb = 0; kb < 2; ++kb) {
          const uint64_t ch = dbase + (uint32_t)((kb * D_KB) >> 4), cl = ch + (D_PLANE >> 4);
          const uint64_t sh = ch + (2 * D_PLANE >> 4), sl = sh + (D_PLANE >> 4);
          {
            const uint64_t ah = abase + (uint32_t)(fb * (D_STAGE >> 4)), al = ah + (D_APLANE >> 4);
            mbar_wait(&full[fb], fph);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < ((SSB_DFT_EXP & 1) ? 0 : 4); ++k) {     // small terms first
              umma_bf16(acc_c, al + 2 * k, ch + 2 * k, idesc, (kb | k) ? 1u : 0u);
              umma_bf16(acc_c, ah + 2 * k, cl + 2 * k, idesc, 1u);
              umma_bf16(acc_c, ah + 2 * k, ch + 2 * k, idesc, 1u);
            }
            umma_commit(&empty[fb]);
            if (++fb == D_NSTAGE) { fb = 0; fph ^= 1; }
          }
          {
            const uint64_t ah = abase + (uint32_t)(fb * (D_STAGE >> 4)), al = ah + (D_APLANE >> 4);
            mbar_wait(&full[fb], fph);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < ((SSB_DFT_EXP & 1) ? 0 : 4); ++k) {
              umma_bf16(acc_s, al + 2 * k, sh + 2 * k, idesc, (kb | k) ? 1u : 0u);
              umma_bf16(acc_s, ah + 2 * k, sl + 2 * k, idesc, 1u);
              umma_bf16(acc_s, ah + 2 * k, sh + 2 * k, idesc, 1u);
            }
            umma_commit(&empty[fb]);
            if (++fb == D_NSTAGE) { fb = 0; fph ^= 1; }
          }
        }
        umma_commit(&tfull[buf]);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (bins 0..127)
    const int q = warp & 3, h = (warp - (D_PROD_WARPS + 1)) >> 2;       // TMEM lane quarter; which share of the bins
    constexpr int NCH = 8 / (D_EPI_WARPS / 4);                          // 16-bin chunks per warp
    const int r = q * 32 + lane;
    int lt = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
      const DftMeta m = dft_meta(tiles, tile);
      const int buf = lt & 1;
      const int t = m.t0 + r;
      const bool valid = t < m.T;
      const uint32_t acc_c = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256 + h * (NCH * 16), acc_s = acc_c + 128;
      const uint32_t T4 = (uint32_t)m.T * 4u;                             // byte stride between bins
      char* o = reinterpret_cast<char*>(out + m.obase + t) + (unsigned long long)T4 * (uint32_t)(h * NCH * 16);
      mbar_wait_backoff(&tfull[buf], (lt >> 1) & 1);
      tc_fence_after();
      // chunk c+1 is in flight from TMEM while chunk c is squared, rooted and stored
      uint32_t ra[32], rb[32];
      tmem_ld16x2_issue(acc_c, acc_s, ra);
      tmem_ld_wait(ra);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        uint32_t* cur_r = (c & 1) ? rb : ra;
        uint32_t* nxt_r = (c & 1) ? ra : rb;
        if (c + 1 < NCH) tmem_ld16x2_issue(acc_c + (c + 1) * 16, acc_s + (c + 1) * 16, nxt_r);
        if (valid && !((SSB_DFT_EXP & 4) && cur_r[0] != 123u)) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {                                // one IMAD.WIDE.U32 per address
            const float re = __uint_as_float(cur_r[i]), im = __uint_as_float(cur_r[16 + i]);
            *reinterpret_cast<float*>(o + (unsigned long long)T4 * (uint32_t)i) = dft_sqrt(re * re + im * im);
          }
        }
        o += (unsigned long long)T4 * 16u;
        if (c + 1 < NCH) tmem_ld_wait(nxt_r);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[buf]);
    }
  }
  __syncthreads();
  if (warp == D_PROD_WARPS) {
    tc_fence_after();
    tmem_free<512>(tmem_base);
  }
}

}  // namespace ssb

extern "C" int ss_stft_dft256_tile_frames(void) { return ssb::D_TM; }

extern "C" int ss_stft_dft256(const float* sig, const long long* tiles, int n_tiles, float* out, void* stream) {
  using namespace ssb;
  if (n_tiles <= 0) return 0;
  SS_REQUIRE(sig && tiles && out, "ss_stft_dft256: null pointer");
  SS_REQUIRE((reinterpret_cast<uintptr_t>(tiles) & 15) == 0, "ss_stft_dft256: tile table must be 16-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  static unsigned char* basis[64] = {nullptr};                          // per device, built on first use
  static int n_sm[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  SS_REQUIRE(dev >= 0 && dev < 64, "ss_stft_dft256: device index %d", dev);
  if (!basis[dev]) {
    SS_REQUIRE(cudaMalloc(&basis[dev], D_BASIS) == cudaSuccess, "ss_stft_dft256: cudaMalloc of the basis failed");
    dft256_basis_kernel<<<cdiv(2 * 2 * 128 * 64, 256), 256, 0, st>>>(basis[dev]);
    cudaFuncSetAttribute(stft_dft256_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D_SMEM);
    cudaDeviceGetAttribute(&n_sm[dev], cudaDevAttrMultiProcessorCount, dev);
    int rc = check_launch("ss_stft_dft256(basis)");
    if (rc) return rc;
  }
  const int grid = n_tiles < n_sm[dev] ? n_tiles : n_sm[dev];
  stft_dft256_kernel<<<grid, D_THREADS, D_SMEM, st>>>(sig, reinterpret_cast<const longlong2*>(tiles), n_tiles,
                                                      basis[dev], out);
  return check_launch("ss_stft_dft256");
}
