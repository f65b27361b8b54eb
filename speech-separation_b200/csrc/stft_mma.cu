// K1 on the tensor cores: 256-point / hop-64 magnitude STFT as a tcgen05 GEMM against a resident, windowed DFT basis.
// Replaces librosa.core.stft + np.abs at steps/extract_feats.py:87,89 (same contract as ss_stft with n_fft = 256,
// hop = 64, magnitude, (F, T) layout; the FFT kernel in stft.cu stays the general path).
//
// Frames are never materialised.  hop = n_fft / 4, so frame r of an utterance is samples [64 r - 128, 64 r + 128) of
// the reflect-padded signal: with the padded signal stored as ROWS of 64 samples (128 bytes of bf16), frame r is rows
// r .. r+3 side by side.  A K-major SWIZZLE_128B operand whose start address is advanced by kc rows therefore IS the
// K block kc (64 samples) of 128 consecutive frames: every sample is split into bf16 hi / lo ONCE, stored ONCE per
// plane, and read by four different descriptors.  (The previous version folded and re-split every frame: its
// producers, not its MMAs, bounded it -- profiles/r01_summary.md section K.)
//
//   X[k] = sum_{n < 256} w[n] x[n] exp(-2 pi i k n / 256),   w = periodic Hann
//   D[frame][col] = sum_n A[frame][n] * Bas[col][n]   with the window inside the basis (a diagonal scaling along K),
//   three products hi*hi + hi*lo + lo*hi, fp32 accumulation in TMEM (error ~2^-16, inside the 1e-4 feature bar).
//
// The basis (256 outputs x 256 samples x hi/lo = 256 KB) does not fit one SM, so the bins are split between two CTA
// types: type h holds bins 64 h .. 64 h + 63 (columns 0..63 = Re, 64..127 = Im; 128 KB resident) and both types walk the
// same tiles, so a tile's samples come from HBM once and from L2 the second time.  Im X[0] is identically zero, so that
// column of type 0 carries the Nyquist bin instead (real: sum_n (-1)^n w[n] x[n]).
// Per 16-sample k-step two instructions: x_hi . [B_hi ; B_lo] (N = 256: the hi*hi and hi*lo products side by side in
// TMEM) and x_lo . B_hi (N = 128, accumulated onto the first half); the epilogue adds the two halves.  Stacking the planes
// in N fetches 20 KB of operands per k-step from shared memory instead of 24 KB for three N = 128 instructions -- the
// kernel's MMAs are shared-memory-bandwidth bound (both operands come from there: profiles/r02_summary.md).
//
// Work units are 128 consecutive frames of the STREAM of all utterances: utterance u owns T_u + 3 stream rows (its
// reflect-padded signal), so a tile may straddle two utterances; the 3 frames that would mix two utterances are
// computed and dropped (0.4 % instead of the 8.5 % a per-utterance tiling wastes on partial tiles at configs[1]).
//
// One persistent CTA per SM, 21 warps: warps 0..11 producers (global fp32 -> split -> swizzled stage; the loads of tile
// i+1 are in flight while tile i is split and stored), warp 12 MMA issuer (32 MMAs per tile), warps 13..20 epilogue
// (tcgen05.ld -> sqrt(re^2+im^2) -> (F, T_u) rows; a warp stores 32 consecutive frames of one bin = 128 contiguous bytes;
// two warps per TMEM lane quarter, 32 bins each).  Two sample stages and two TMEM accumulators: producers fill tile i+1
// and the epilogue drains tile i-1 while the MMAs of tile i run.  Every role is a latency-bound instruction stream
// (ncu: 0.14 instructions per cycle and scheduler with 8 + 4 warps), hence the warp counts.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int D_TM = 128;                                 // frames per tile
constexpr int D_ROWS = 136;                               // sample rows per stage: 128 + 3, rounded up to 8-row atoms
constexpr int D_PROD_WARPS = 12, D_EPI_WARPS = 8;
constexpr int D_THREADS = (D_PROD_WARPS + 1 + D_EPI_WARPS) * 32;
constexpr int D_KB = 2 * 128 * 128;                       // one 64-sample K block of the basis: [hi 128 columns | lo 128 columns] x 128 B
constexpr int D_BASIS = 4 * D_KB;                         // 128 KB per CTA type
constexpr int D_APLANE = D_ROWS * 128;                    // 17408 B
constexpr int D_STAGE = 2 * D_APLANE;                     // hi | lo = 34 KB
constexpr int D_TASKS = D_ROWS * 8;                       // (row, 16-byte chunk of 8 samples) tasks per stage
constexpr int D_SMEM = D_BASIS + 2 * D_STAGE + 256 /*barriers*/ + 1024 /*align*/;
static_assert(D_SMEM <= 232448, "shared memory budget");
static_assert(D_STAGE % 1024 == 0, "stages must keep the 1024-byte swizzle pattern alignment");

struct DftUtt {            // 32 bytes per utterance
  long long sig_off;       // first sample in sig
  long long out_off;       // first output element (frame_off[u] * 129)
  long long nt;            // N_u | T_u << 32
  long long row_off;       // first stream row (prefix sum of T_u + 3)
};
__device__ __forceinline__ DftUtt dft_utt(const DftUtt* __restrict__ tab, int u) {
  const longlong2 a = __ldg(reinterpret_cast<const longlong2*>(tab + u));
  const longlong2 b = __ldg(reinterpret_cast<const longlong2*>(tab + u) + 1);
  return DftUtt{a.x, a.y, b.x, b.y};
}

__device__ __forceinline__ float dft_sqrt(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// eight floats -> packed bf16 hi (four words) and the bf16 of the residuals
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __nv_bfloat162 hh = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
    h[i] = *reinterpret_cast<const uint32_t*>(&hh);
    // bf16 -> f32 is a 16-bit shift
    const __nv_bfloat162 ll = __floats2bfloat162_rn(v[2 * i] - __uint_as_float(h[i] << 16),
                                                    v[2 * i + 1] - __uint_as_float(h[i] & 0xffff0000u));
    l[i] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}
// the same loads without the wait, and a wait that the result registers depend on (software pipelining)
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
__device__ __forceinline__ void mbar_wait_backoff(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) __nanosleep(32);
}

// Basis image in the exact shared-memory layout the MMA descriptors read: [type h][K block kc][plane][column][64 k],
// rows of 128 B, 16-byte chunks XOR-swizzled with row % 8, 8-row atoms 1024 B apart (the lo plane directly behind the hi
// plane, so that one N = 256 descriptor covers both).
// Column c of type h: bin k = 64 h + (c & 63); c < 64: w[n] cos(2 pi k n / 256), c >= 64: -w[n] sin(2 pi k n / 256),
// except column 64 of type 0 (Im X[0] = 0): the Nyquist bin, (-1)^n w[n].
__global__ void dft256_basis_kernel(unsigned char* __restrict__ img) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;          // (h, kc, c, kk)
  if (e >= 2 * 4 * 128 * 64) return;
  const int kk = e & 63, c = (e >> 6) & 127, kc = (e >> 13) & 3, h = e >> 15;
  const int n = kc * 64 + kk;                                           // sample index within the frame
  const double w = 0.5 - 0.5 * cospi((double)n / 128.0);               // periodic Hann, n_fft = 256
  const int k = 64 * h + (c & 63);
  const int m = (k * n) & 255;                                          // angle 2 pi m / 256
  double v = c >= 64 ? -w * sinpi((double)m / 128.0) : w * cospi((double)m / 128.0);
  if (h == 0 && c == 64) v = (n & 1) ? -w : w;
  const __nv_bfloat16 hi = __float2bfloat16_rn((float)v);
  const __nv_bfloat16 lo = __float2bfloat16_rn((float)(v - (double)__bfloat162float(hi)));
  const int off = kc * D_KB + c * 128 + ((((kk >> 3) ^ (c & 7))) << 4) + (kk & 7) * 2;
  unsigned char* base = img + (size_t)h * D_BASIS;
  *reinterpret_cast<__nv_bfloat16*>(base + off) = hi;
  *reinterpret_cast<__nv_bfloat16*>(base + D_KB / 2 + off) = lo;
}

// Swizzle phase of the row-shifted A descriptors: the tensor core XORs the 16-byte chunk index with bits [7,10) of the
// ABSOLUTE shared-memory address, which is also how the producers place the chunks (the stages are 1024-byte aligned), so
// a start address advanced by kc * 128 B needs nothing else; the descriptor's matrix-base-offset field stays 0 (measured:
// setting it to (start >> 7) & 7 reads the wrong chunks -- errors of order 1, profiles/r02_summary.md).
__global__ void __launch_bounds__(D_THREADS, 1)
stft_dft256_kernel(const float* __restrict__ sig, const DftUtt* __restrict__ utts, int n_utts,
                   const int* __restrict__ tile_u0, int n_tiles, long long total_rows,
                   const unsigned char* __restrict__ basis, float* __restrict__ out, int log_out, int exp) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* stages = smem + D_BASIS;
  uint64_t* full = reinterpret_cast<uint64_t*>(stages + 2 * D_STAGE);
  uint64_t* empty = full + 2;
  uint64_t* tfull = empty + 2;
  uint64_t* tempty = tfull + 2;
  uint64_t* basis_bar = tempty + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(basis_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int h = blockIdx.x & 1, pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&full[s], D_PROD_WARPS);
      mbar_init(&empty[s], 1);
      mbar_init(&tfull[s], 1);
      mbar_init(&tempty[s], D_EPI_WARPS);
    }
    mbar_init(basis_bar, 1);
    fence_barrier_init();
  }
  if (warp == D_PROD_WARPS) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < D_PROD_WARPS) {
    // ------------------------------------------------------------------ producers
    const int tid = threadIdx.x;                                       // 0 .. 32 * D_PROD_WARPS - 1
    const int c = tid & 7;                                             // 16-byte chunk (8 samples) of the row, fixed
    const uint32_t st_base = smem_u32(stages);
    constexpr int NP = D_PROD_WARPS * 32;
    constexpr int NT = (D_TASKS + NP - 1) / NP;                        // tasks per thread (the last one partial)
    static_assert(NP % 8 == 0 && (D_TASKS % 32) == 0, "chunk index fixed per thread, task predicate warp-uniform");

    // all of this thread's loads of one tile, issued together (rows ascend with the task index)
    auto load_tile = [&](int tile, float (&x)[NT][8]) {
      int u = __ldg(tile_u0 + tile);
      DftUtt U = dft_utt(utts, u);
      long long next_off = u + 1 < n_utts ? __ldg(&utts[u + 1].row_off) : total_rows;
      const long long R0 = (long long)tile * D_TM;
#pragma unroll
      for (int it = 0; it < NT; ++it) {
        const int task = tid + it * NP;
        const long long R = R0 + (task >> 3);
        if (task < D_TASKS && R < total_rows && !(exp & 2)) {
          while (R >= next_off) {
            ++u;
            U = dft_utt(utts, u);
            next_off = u + 1 < n_utts ? __ldg(&utts[u + 1].row_off) : total_rows;
          }
          const int N = (int)(U.nt & 0xffffffffll);
          const int i0 = (int)(R - U.row_off) * 64 - 128 + 8 * c;      // first of the 8 samples, utterance-relative
          const float* p = sig + U.sig_off;
          if (i0 >= 0 && i0 + 8 <= N && ((reinterpret_cast<uintptr_t>(p + i0) & 15) == 0)) {
            const float4 a = __ldg(reinterpret_cast<const float4*>(p + i0));
            const float4 b = __ldg(reinterpret_cast<const float4*>(p + i0) + 1);
            x[it][0] = a.x; x[it][1] = a.y; x[it][2] = a.z; x[it][3] = a.w;
            x[it][4] = b.x; x[it][5] = b.y; x[it][6] = b.z; x[it][7] = b.w;
          } else {
            // reflect padding (np.pad(y, n_fft//2, 'reflect')), unaligned utterances
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              int jx = i0 + i;
              jx = jx < 0 ? -jx : jx;
              jx = jx >= N ? 2 * N - 2 - jx : jx;
              jx = jx < 0 ? 0 : jx;                                    // (rows of the 3 dropped frames past a short utterance)
              x[it][i] = __ldg(p + jx);
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) x[it][i] = 0.0f;
        }
      }
    };
    // split + swizzled stores of one tile into stage s, once the MMAs of the tile two back have read it
    auto store_tile = [&](int lt, const float (&x)[NT][8]) {
      const int s = lt & 1;
      mbar_wait_backoff(&empty[s], ((lt >> 1) & 1) ^ 1);
      const uint32_t sa = st_base + s * D_STAGE;
#pragma unroll
      for (int it = 0; it < NT; ++it) {
        const int task = tid + it * NP;
        if (task < D_TASKS && !(exp & 2)) {
          const int m = task >> 3;
          uint4 hi, lo;
          split8(x[it], hi, lo);
          const uint32_t off = (uint32_t)(m * 128 + ((c ^ (m & 7)) << 4));
          sts128(sa + off, hi);
          sts128(sa + D_APLANE + off, lo);
        }
      }
      fence_proxy_async_smem();                                        // own stores -> async proxy, then one arrival
      __syncwarp();
      if (lane == 0) mbar_arrive(&full[s]);
    };
    float xa[NT][8], xb[NT][8];
    int lt = 0, tile = pair;
    if (tile < n_tiles) load_tile(tile, xa);
    while (tile < n_tiles) {                                           // two tiles per trip: the register sets alternate
      int nxt = tile + n_pairs;
      if (nxt < n_tiles) load_tile(nxt, xb);
      store_tile(lt++, xa);
      tile = nxt;
      if (tile >= n_tiles) break;
      nxt = tile + n_pairs;
      if (nxt < n_tiles) load_tile(nxt, xa);
      store_tile(lt++, xb);
      tile = nxt;
    }
  } else if (warp == D_PROD_WARPS) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      mbar_arrive_expect_tx(basis_bar, D_BASIS);
      constexpr int CH = D_BASIS / 8;
      for (int i = 0; i < 8; ++i) bulk_load(smem + i * CH, basis + (size_t)h * D_BASIS + i * CH, CH, basis_bar);
      mbar_wait(basis_bar, 0);
      constexpr uint32_t idesc256 = umma_idesc_bf16(128, 256), idesc128 = umma_idesc_bf16(128, 128);
      const uint64_t bbase = umma_desc_sw128(smem_u32(smem));
      int lt = 0;
      for (int tile = pair; tile < n_tiles; tile += n_pairs, ++lt) {
        const int s = lt & 1;
        const uint32_t ph = (lt >> 1) & 1;
        mbar_wait(&tempty[s], ph ^ 1);                                  // epilogue drained this accumulator
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t acc = tmem_base + s * 256;
        const uint64_t abase = umma_desc_sw128(smem_u32(stages + s * D_STAGE));
#pragma unroll
        for (int kc = 0; kc < 4; ++kc) {
          // K block kc of every frame = the sample rows shifted by kc: start address + kc * 128 B
          const uint64_t ah = abase + (uint32_t)((kc * 128) >> 4), al = ah + (D_APLANE >> 4);
          const uint64_t bb = bbase + (uint32_t)((kc * D_KB) >> 4);      // [B_hi ; B_lo] of this K block, 256 rows
#pragma unroll
          for (int k = 0; k < 4; ++k) {                                  // 16 samples = 32 bytes inside the 128-byte row
            if (exp & 1) break;
            umma_bf16(acc, ah + 2 * k, bb + 2 * k, idesc256, (kc | k) ? 1u : 0u);         // x_hi . [B_hi ; B_lo]
            umma_bf16(acc, al + 2 * k, bb + 2 * k, idesc128, 1u);                        // x_lo . B_hi
          }
        }
        umma_commit(&empty[s]);
        umma_commit(&tfull[s]);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue
    const int q = warp & 3;                                             // TMEM lane quarter this warp may read
    const int eh = (warp - (D_PROD_WARPS + 1)) >> 2;                    // which 32 of the CTA's 64 bins
    const int r = q * 32 + lane;                                        // frame within the tile
    int lt = 0;
    for (int tile = pair; tile < n_tiles; tile += n_pairs, ++lt) {
      const int s = lt & 1;
      const long long R = (long long)tile * D_TM + r;                   // stream frame
      bool valid = false;
      float* o = out;
      uint32_t T4 = 0;
      if (R < total_rows) {
        int u = __ldg(tile_u0 + tile);
        DftUtt U = dft_utt(utts, u);
        while (u + 1 < n_utts && R >= __ldg(&utts[u + 1].row_off)) U = dft_utt(utts, ++u);
        const int T = (int)(U.nt >> 32), t = (int)(R - U.row_off);
        valid = t < T;                                                  // the last 3 stream frames of an utterance are dropped
        o = out + U.out_off + t;
        T4 = (uint32_t)T * 4u;                                          // byte stride between bins
      }
      char* ob = reinterpret_cast<char*>(o) + (unsigned long long)T4 * (uint32_t)(64 * h + 32 * eh);
      mbar_wait_backoff(&tfull[s], (lt >> 1) & 1);
      tc_fence_after();
      // columns: [0,64) Re hi*hi+lo*hi, [64,128) Im ..., [128,192) Re hi*lo, [192,256) Im hi*lo
      const uint32_t acc_re = tmem_base + ((uint32_t)(q * 32) << 16) + s * 256 + 32 * eh, acc_im = acc_re + 64;
      uint32_t ra[32], rb[32];
#pragma unroll
      for (int cch = 0; cch < 2; ++cch) {
        tmem_ld16x2_issue(acc_re + cch * 16, acc_im + cch * 16, ra);
        tmem_ld16x2_issue(acc_re + 128 + cch * 16, acc_im + 128 + cch * 16, rb);
        tmem_ld_wait(ra);
        tmem_ld_wait(rb);
        if (valid && !(exp & 4)) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float re = __uint_as_float(ra[i]) + __uint_as_float(rb[i]);
            float im = __uint_as_float(ra[16 + i]) + __uint_as_float(rb[16 + i]);
            const bool nyq_col = h == 0 && eh == 0 && cch == 0 && i == 0;     // that column holds the Nyquist bin, Im X[0] = 0
            float v = dft_sqrt(re * re + (nyq_col ? 0.0f : im * im));
            if (log_out) v = logf(fmaxf(v, 1e-8f));
            *reinterpret_cast<float*>(ob + (unsigned long long)T4 * (uint32_t)i) = v;
            if (nyq_col) {
              float vn = fabsf(im);
              if (log_out) vn = logf(fmaxf(vn, 1e-8f));
              *reinterpret_cast<float*>(ob + (unsigned long long)T4 * 128u) = vn;          // bin 128
            }
          }
        }
        ob += (unsigned long long)T4 * 16u;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[s]);
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

extern "C" int ss_stft_dft256(const float* sig, const long long* utts, int n_utts, const int* tile_u0, int n_tiles,
                              long long total_rows, float* out, int flags, void* stream) {
  using namespace ssb;
  if (n_tiles <= 0 || n_utts <= 0) return 0;
  SS_REQUIRE(sig && utts && tile_u0 && out, "ss_stft_dft256: null pointer");
  SS_REQUIRE((reinterpret_cast<uintptr_t>(utts) & 15) == 0, "ss_stft_dft256: utterance table must be 16-byte aligned");
  // (flag bits 8..10: timing experiments of tools/diag_dft.py -- no MMAs / no producer work / no epilogue stores; wrong results)
  SS_REQUIRE((flags & ~(SS_STFT_LOG | 0x700)) == 0, "ss_stft_dft256: magnitude (optionally log) in the (F, T) layout only");
  cudaStream_t st = (cudaStream_t)stream;
  static unsigned char* basis[64] = {nullptr};                          // per device, built on first use
  int dev = 0;
  cudaGetDevice(&dev);
  SS_REQUIRE(dev >= 0 && dev < 64, "ss_stft_dft256: device index %d", dev);
  if (!basis[dev]) {
    SS_REQUIRE(cudaMalloc(&basis[dev], 2 * D_BASIS) == cudaSuccess, "ss_stft_dft256: cudaMalloc of the basis failed");
    dft256_basis_kernel<<<cdiv(2 * 4 * 128 * 64, 256), 256, 0, st>>>(basis[dev]);
    cudaFuncSetAttribute(stft_dft256_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, D_SMEM);
    int rc = check_launch("ss_stft_dft256(basis)");
    if (rc) return rc;
  }
  int grid = sm_count() & ~1;                                           // CTA pairs: type 0 / type 1 of the same tiles
  if (grid > 2 * n_tiles) grid = 2 * n_tiles;
  stft_dft256_kernel<<<grid, D_THREADS, D_SMEM, st>>>(sig, reinterpret_cast<const DftUtt*>(utts), n_utts, tile_u0, n_tiles,
                                                      total_rows, basis[dev], out, (flags & SS_STFT_LOG) ? 1 : 0, (flags >> 8) & 7);
  return check_launch("ss_stft_dft256");
}
