// K1 on the tensor cores: 256-point / hop-64 magnitude STFT as two tcgen05 GEMMs against a resident, windowed
// DFT basis.  Replaces librosa.core.stft + np.abs at steps/extract_feats.py:87,89 (same contract as ss_stft with
// n_fft = 256, hop = 64, flags = 0; the FFT kernel in stft.cu stays the general path).
//
// Algebra.  With the periodic Hann window w[n] = w[256-n], w[0] = 0 and a real frame x[0..255]:
//   Re X[k] =  sum_{n=1..128} (w[n] cos(2 pi k n/256)) e[n],   e[n] = x[n] + x[256-n] (n < 128), e[128] = x[128]
//   Im X[k] = -sum_{n=1..127} (w[n] sin(2 pi k n/256)) o[n],   o[n] = x[n] - x[256-n]
// so a tile of 128 frames is  E[128 x 128] * C^T[128 x 129]  and  O[128 x 128] * S^T[128 x 128]  (K = n-1).  The
// window lives inside the basis (a diagonal scaling along K), the fold halves the flops of a plain 256-point DFT
// matrix, and every operand is split-bf16 (hi + lo) with the three products hi*hi + hi*lo + lo*hi accumulated in
// fp32 in TMEM (error ~2^-16, inside the 1e-4 feature bar).
//
// One persistent CTA per SM, 17 warps:
//   warps 0..7   producers: half a warp per frame and K block; two LDG.128 (x[n..n+3] and the mirrored quad) per
//                lane, fold, split into hi/lo bf16 and 8-byte stores into the 128B-swizzled K-major A stage.  The odd
//                part o stays packed in registers until its stage is free, so every sample is loaded once per frame.
//   warp 8       one thread issues the MMAs (M=128, N=144 cos / N=128 sin, K=16; 48 per tile) and the commits
//   warps 9..16  epilogue: TMEM -> sqrt(re^2 + im^2) -> (F, T_u) rows of the utterance; a warp stores 32 consecutive
//                frames of one bin (128 contiguous bytes)
// Shared memory: the 132 KB basis image (loaded once per CTA by bulk copies), two 32 KB A stages cycling through
// (E, kb0), (O, kb0), (E, kb1), (O, kb1).  TMEM: three 144-column accumulator slots; tile i uses slots 2i%3 (cos)
// and (2i+1)%3 (sin), so the cos GEMM of tile i+1 runs while the epilogue drains tile i.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int D_TM = 128;                                 // frames per tile
constexpr int D_PROD_WARPS = 16, D_EPI_WARPS = 4;
constexpr int D_TASKS = D_TM * 16 / (D_PROD_WARPS * 32);   // (frame, 16-lane) tasks per producer thread and K block
constexpr int D_THREADS = (D_PROD_WARPS + 1 + D_EPI_WARPS) * 32;
constexpr int D_COS_ROWS = 136;                           // 129 bins rounded up to whole 8-row swizzle atoms
constexpr int D_COS_KB = D_COS_ROWS * 128;                // bytes of one 64-wide K block of the cos basis plane
constexpr int D_SIN_KB = 128 * 128;
constexpr int D_COS_PLANE = 2 * D_COS_KB, D_SIN_PLANE = 2 * D_SIN_KB;
constexpr int D_BASIS = 2 * D_COS_PLANE + 2 * D_SIN_PLANE;   // cos hi | cos lo | sin hi | sin lo = 135168 B
constexpr int D_APLANE = D_TM * 128;                      // one bf16 plane of an A stage: 128 frames x 64 k
constexpr int D_STAGE = 2 * D_APLANE;                     // hi + lo
constexpr int D_SLOT = 144;                               // TMEM columns per accumulator slot
constexpr int D_SMEM = D_BASIS + 2 * D_STAGE + 256 /*barriers*/ + 1024 /*align*/;
static_assert(D_BASIS % 1024 == 0 && D_COS_KB % 1024 == 0, "swizzle atoms must stay 1024-byte aligned");

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
  const __nv_bfloat162 l01 = __floats2bfloat162_rn(v[0] - __low2float(h01), v[1] - __high2float(h01));
  const __nv_bfloat162 l23 = __floats2bfloat162_rn(v[2] - __low2float(h23), v[3] - __high2float(h23));
  hi.x = *reinterpret_cast<const uint32_t*>(&h01);
  hi.y = *reinterpret_cast<const uint32_t*>(&h23);
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

// Basis image in the exact shared-memory layout the MMA descriptors read (K-major rows of 128 B, 16-byte chunks
// XOR-swizzled with row%8, 8-row atoms 1024 B apart): plane = [K block 0: rows][K block 1: rows].
__global__ void dft256_basis_kernel(unsigned char* __restrict__ img) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  const int n_cos = 2 * D_COS_ROWS * 64, n_sin = 2 * 128 * 64;
  if (e >= n_cos + n_sin) return;
  const bool is_sin = e >= n_cos;
  const int rows = is_sin ? 128 : D_COS_ROWS;
  const int q = is_sin ? e - n_cos : e;
  const int kb = q / (rows * 64), r = (q / 64) % rows, kk = q % 64;
  const int n = kb * 64 + kk + 1;                                       // sample index 1..128
  const double w = 0.5 - 0.5 * cospi((double)n / 128.0);               // periodic Hann, n_fft = 256
  const int m = (r * n) & 255;                                          // angle 2 pi m / 256
  double v;
  if (is_sin) v = -w * sinpi((double)m / 128.0);                        // rows = bins 0..127 (row 0 is all zero)
  else v = r <= 128 ? w * cospi((double)m / 128.0) : 0.0;              // rows = bins 0..128, 129..135 zero
  const __nv_bfloat16 hi = __float2bfloat16_rn((float)v);
  const __nv_bfloat16 lo = __float2bfloat16_rn((float)(v - (double)__bfloat162float(hi)));
  const int off = kb * (rows * 128) + r * 128 + ((((kk >> 3) ^ (r & 7))) << 4) + (kk & 7) * 2;
  unsigned char* base = img + (is_sin ? 2 * D_COS_PLANE : 0);
  *reinterpret_cast<__nv_bfloat16*>(base + off) = hi;
  *reinterpret_cast<__nv_bfloat16*>(base + (is_sin ? D_SIN_PLANE : D_COS_PLANE) + off) = lo;
}

__global__ void __launch_bounds__(D_THREADS, 1)
stft_dft256_kernel(const float* __restrict__ sig, const longlong2* __restrict__ tiles, int n_tiles,
                   const unsigned char* __restrict__ basis, float* __restrict__ out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* stages = smem + D_BASIS;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + D_BASIS + 2 * D_STAGE);
  uint64_t* empty = full + 2;
  uint64_t* tfull = empty + 2;
  uint64_t* drained = tfull + 1;
  uint64_t* basis_bar = drained + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(basis_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int s = 0; s < 2; ++s) {
      mbar_init(&full[s], D_PROD_WARPS * 32);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tfull, 1);
    mbar_init(drained, D_EPI_WARPS);
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
    const int half = lane >> 4, jq = lane & 15;
    float4 LA[D_TASKS], LB[D_TASKS];
    float EX[D_TASKS];
    uint2 OH[D_TASKS], OL[D_TASKS];
    uint32_t fastmask = 0;
    const uint32_t st0 = smem_u32(stages), st1 = st0 + D_STAGE;

    auto issue = [&](const DftMeta& m, int kb) {
      fastmask = 0;
      const int nb = kb * 64;
#pragma unroll
      for (int i = 0; i < D_TASKS; ++i) {
        const int r = warp * (2 * D_TASKS) + 2 * i + half;
        const int t = m.t0 + r, start = t * 64 - 128;
        const float* p = sig + m.s0 + start;                           // x[0] of the frame (may precede the utterance)
        const bool valid = t < m.T;
        const bool fast = valid && start >= 0 && start + 256 <= m.N && (reinterpret_cast<uintptr_t>(p) & 15) == 0;
        if (__all_sync(0xffffffffu, fast)) {
          fastmask |= 1u << i;
          LA[i] = __ldg(reinterpret_cast<const float4*>(p + nb + 4 * jq));
          LB[i] = __ldg(reinterpret_cast<const float4*>(p + 252 - nb - 4 * jq));
          EX[i] = 0.0f;
          if (kb == 0 && lane == 31) EX[i] = __ldg(p + 64);
        } else {
          // reflect padding (np.pad(y, n_fft//2, 'reflect')), unaligned utterances, frames past the end
          float xa[4], xb[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int n = nb + 4 * jq + 1 + q;
            int ia = start + n, ib = start + 256 - n;
            ia = ia < 0 ? -ia : ia;
            ib = ib < 0 ? -ib : ib;
            ia = ia >= m.N ? 2 * m.N - 2 - ia : ia;
            ib = ib >= m.N ? 2 * m.N - 2 - ib : ib;
            xa[q] = valid ? __ldg(sig + m.s0 + ia) : 0.0f;
            xb[q] = valid ? __ldg(sig + m.s0 + ib) : 0.0f;
          }
          LA[i] = make_float4(0.0f, xa[0], xa[1], xa[2]);
          EX[i] = xa[3];
          LB[i] = make_float4(xb[3], xb[2], xb[1], xb[0]);
        }
      }
    };

    int tile = blockIdx.x;
    DftMeta cur = {0, 0, 0, 0, 0}, nxt = cur;
    if (tile < n_tiles) {
      cur = dft_meta(tiles, tile);
      issue(cur, 0);
    }
    int g = 0;                                                         // steps done: 2 per tile, one per K block
    for (; tile < n_tiles; tile += gridDim.x) {
      const int tile_n = tile + gridDim.x;
      if (tile_n < n_tiles) nxt = dft_meta(tiles, tile_n);
#pragma unroll
      for (int kb = 0; kb < 2; ++kb, ++g) {
        mbar_wait(&empty[0], (g & 1) ^ 1);
        const bool kfast_last = kb == 1;
#pragma unroll
        for (int i = 0; i < D_TASKS; ++i) {
          const int r = warp * (2 * D_TASKS) + 2 * i + half;
          float nx = __shfl_down_sync(0xffffffffu, LA[i].x, 1);
          if (!((fastmask >> i) & 1)) nx = EX[i];
          else if (lane == 31) nx = kfast_last ? LB[i].x : EX[i];
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
          uint2 eh, el;
          split4(ev, eh, el);
          split4(ov, OH[i], OL[i]);
          const uint32_t off = (uint32_t)(r * 128 + ((((jq >> 1) ^ (r & 7))) << 4) + ((jq & 1) << 3));
          sts64(st0 + off, eh);
          sts64(st0 + D_APLANE + off, el);
        }
        fence_proxy_async_smem();
        mbar_arrive(&full[0]);
        // loads of the next step fly while the odd part waits for its stage
        if (kb == 0) issue(cur, 1);
        else if (tile_n < n_tiles) issue(nxt, 0);
        mbar_wait(&empty[1], (g & 1) ^ 1);
#pragma unroll
        for (int i = 0; i < D_TASKS; ++i) {
          const int r = warp * (2 * D_TASKS) + 2 * i + half;
          const uint32_t off = (uint32_t)(r * 128 + ((((jq >> 1) ^ (r & 7))) << 4) + ((jq & 1) << 3));
          sts64(st1 + off, OH[i]);
          sts64(st1 + D_APLANE + off, OL[i]);
        }
        fence_proxy_async_smem();
        mbar_arrive(&full[1]);
      }
      cur = nxt;
    }
  } else if (warp == D_PROD_WARPS) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      mbar_arrive_expect_tx(basis_bar, D_BASIS);
      constexpr int CH = D_BASIS / 8;
      for (int c = 0; c < 8; ++c) bulk_load(smem + c * CH, basis + c * CH, CH, basis_bar);
      mbar_wait(basis_bar, 0);
      constexpr uint32_t idesc_cos = umma_idesc_bf16(128, D_SLOT), idesc_sin = umma_idesc_bf16(128, 128);
      const uint64_t dbase = umma_desc_sw128(smem_u32(smem));
      const uint64_t a0h = dbase + (D_BASIS >> 4), a0l = a0h + (D_APLANE >> 4);
      const uint64_t a1h = a0h + (D_STAGE >> 4), a1l = a1h + (D_APLANE >> 4);
      int g = 0, lt = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
        const uint32_t acc_c = tmem_base + ((2 * lt) % 3) * D_SLOT, acc_s = tmem_base + ((2 * lt + 1) % 3) * D_SLOT;
#pragma unroll
        for (int kb = 0; kb < 2; ++kb, ++g) {
          const uint64_t ch = dbase + (uint32_t)((kb * D_COS_KB) >> 4), cl = ch + (D_COS_PLANE >> 4);
          const uint64_t sh = dbase + (uint32_t)((2 * D_COS_PLANE + kb * D_SIN_KB) >> 4), sl = sh + (D_SIN_PLANE >> 4);
          mbar_wait(&full[0], g & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {                                 // small terms first
            umma_bf16(acc_c, a0l + 2 * k, ch + 2 * k, idesc_cos, (kb | k) ? 1u : 0u);
            umma_bf16(acc_c, a0h + 2 * k, cl + 2 * k, idesc_cos, 1u);
            umma_bf16(acc_c, a0h + 2 * k, ch + 2 * k, idesc_cos, 1u);
          }
          umma_commit(&empty[0]);
          if (kb == 0 && lt > 0) {                                      // the sin slot was tile lt-1's cos slot
            mbar_wait(drained, (lt - 1) & 1);
            tc_fence_after();
          }
          mbar_wait(&full[1], g & 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            umma_bf16(acc_s, a1l + 2 * k, sh + 2 * k, idesc_sin, (kb | k) ? 1u : 0u);
            umma_bf16(acc_s, a1h + 2 * k, sl + 2 * k, idesc_sin, 1u);
            umma_bf16(acc_s, a1h + 2 * k, sh + 2 * k, idesc_sin, 1u);
          }
          umma_commit(&empty[1]);
        }
        umma_commit(tfull);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue
    const int q = warp & 3;                                             // TMEM lane quarter of this warp
    const int r = q * 32 + lane;
    int lt = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
      const DftMeta m = dft_meta(tiles, tile);
      const int t = m.t0 + r;
      const bool valid = t < m.T;
      const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
      const uint32_t acc_c = tmem_base + lane_sel + ((2 * lt) % 3) * D_SLOT;
      const uint32_t acc_s = tmem_base + lane_sel + ((2 * lt + 1) % 3) * D_SLOT;
      float* o = out + m.obase + t;
      mbar_wait(tfull, lt & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 8; ++c) {
        float re[16], im[16];
        tmem_ld16x2(acc_c + c * 16, acc_s + c * 16, re, im);
        if (valid) {
#pragma unroll
          for (int i = 0; i < 16; ++i) o[(long long)i * m.T] = dft_sqrt(re[i] * re[i] + im[i] * im[i]);
        }
        o += 16ll * m.T;
      }
      {                                                                 // Nyquist bin: real
        float re[16], im[16];
        tmem_ld16x2(acc_c + 128, acc_c + 128, re, im);
        if (valid) o[0] = fabsf(re[0]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(drained);
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
    cudaMemsetAsync(basis[dev], 0, D_BASIS, st);
    const int n = 2 * D_COS_ROWS * 64 + 2 * 128 * 64;
    dft256_basis_kernel<<<cdiv(n, 256), 256, 0, st>>>(basis[dev]);
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
