// tcgen05 GEMM on split-bf16 operands:  C[m][n] = alpha * sum_k A[m][k] * B[n][k] (+ bias[n]) (+ beta * C[m][n])
// A, B are "sb16" matrices (two bf16 planes hi, lo; x ~= hi + lo) and the product is evaluated as
// hi*hi + hi*lo + lo*hi with fp32 accumulation in TMEM (error ~2^-16, i.e. fp32-class results from the bf16 pipe).
// Persistent CTAs over 128x128 output tiles; warp 0 = TMA producer, warp 1 = MMA issuer, warps 2..5 = epilogue.
// Stands in for every cuBLAS call the reference makes through torch (nn.LSTM input projections,
// nn.Linear: archs/uPIT.py:115-117,132,141) and for their autograd backward GEMMs.
//
// Two operand forms share the kernel:
//   NT (TN = false): A[M][K], B[N][K], both K-major (the reduction index is the contiguous one);
//   TN (TN = true):  C[m][n] = sum_r A[r][m] * B[r + shift][n] -- both operands are stored with the REDUCTION index
//       as their row ("MN-major" for the tensor core).  This is the weight-gradient form dW = dG^T X with r = t*B + b:
//       the activations and gate gradients are read exactly as the forward / backward kernels wrote them, so no
//       transposed copy of anything exists.  TMA brings 64-row x 64-column boxes (128-byte rows, SWIZZLE_128B), the
//       shared-memory descriptors declare MN-major atoms (8 reduction rows x 64 columns, LBO = distance between
//       64-column blocks, SBO = 1024 B between 8-row groups) and the instruction descriptor sets a_major = b_major = 1.
//       `shift` rows (per batch entry) realise h_prev[t] = out[t -/+ 1] for dW_hh; TMA zero-fills rows outside [0, R).
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int GB_M = 128, GB_N = 128, GB_K = 64, G_STAGES = 3, G_EPI_WARPS = 8, G_THREADS = 64 + 32 * G_EPI_WARPS;
constexpr int G_PLANE = GB_M * GB_K * 2;                 // 16 KB: one bf16 plane of one operand tile
constexpr int G_STAGE_BYTES = 4 * G_PLANE;               // A(hi,lo) + B(hi,lo)
constexpr int G_EPI_BYTES = G_EPI_WARPS * 32 * 32 * 4;   // one 32 x 32 fp32 block per epilogue warp (XOR-swizzled, no padding)
constexpr int G_SMEM = G_STAGES * G_STAGE_BYTES + G_EPI_BYTES + 1024 /*align*/ + 256 /*barriers*/;
constexpr int G_TMEM = 2 * GB_N;                         // two accumulators: epilogue of tile i overlaps mainloop of i+1

struct GemmParams {
  float* C[2];               // per batch entry
  long long ldc;
  const float* bias;
  float alpha, beta;
  int M, N, K, rowmap, hidden;
  int batch, a_batch_rows, b_batch_rows;   // batch entry z reads A rows [z*a_batch_rows, +M), B rows [z*b_batch_rows, +N)
  // TN form only: A columns [a_col0 + z*a_batch_rows, +M), B columns [b_col0 + z*b_batch_rows, +N); reduction over the
  // 64-row blocks [kb0[z], kb0[z] + nkb) of both operands, B rows shifted by b_shift[z]
  int a_col0, b_col0, nkb, kb0[2], b_shift[2];
};

// MN-major, SWIZZLE_128B shared-memory matrix descriptor: 64-column blocks `lbo` bytes apart, 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t saddr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Persistent: CTA c works on tiles c, c + gridDim.x, ... (n fastest, then m, then batch).
template <bool TN>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_sb16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  float* epi = reinterpret_cast<float*>(smem + G_STAGES * G_STAGE_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + G_STAGES * G_STAGE_BYTES + G_EPI_BYTES);
  uint64_t* empty = full + G_STAGES;
  uint64_t* tfull = empty + G_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = TN ? p.nkb : (p.K + GB_K - 1) / GB_K;
  const int mt = (p.M + GB_M - 1) / GB_M, nt = (p.N + GB_N - 1) / GB_N;
  const int tiles = p.batch * mt * nt;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < G_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], G_EPI_WARPS);
    }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<G_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      int it = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
        const int z = tile / (mt * nt), rem = tile - z * mt * nt;
        const int m0 = (rem / nt) * GB_M + z * p.a_batch_rows, n0 = (rem % nt) * GB_N + z * p.b_batch_rows;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % G_STAGES, ph = (it / G_STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          unsigned char* st = smem + s * G_STAGE_BYTES;
          mbar_arrive_expect_tx(&full[s], G_STAGE_BYTES);
          if (TN) {
            // four boxes of [plane][64 rows][64 columns]: A columns m0.., m0+64.., B columns n0.., n0+64..
            const int r = (p.kb0[z] + kb) * GB_K;
            tma_load_3d(st, &tmA, &full[s], p.a_col0 + m0, r, 0);
            tma_load_3d(st + G_PLANE, &tmA, &full[s], p.a_col0 + m0 + 64, r, 0);
            tma_load_3d(st + 2 * G_PLANE, &tmB, &full[s], p.b_col0 + n0, r + p.b_shift[z], 0);
            tma_load_3d(st + 3 * G_PLANE, &tmB, &full[s], p.b_col0 + n0 + 64, r + p.b_shift[z], 0);
          } else {
            tma_load_3d(st, &tmA, &full[s], kb * GB_K, m0, 0);               // [plane][128 rows][64 k]
            tma_load_3d(st + 2 * G_PLANE, &tmB, &full[s], kb * GB_K, n0, 0);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(GB_M, GB_N) | (TN ? (1u << 15) | (1u << 16) : 0u);
      const uint64_t base = TN ? umma_desc_sw128_mn(smem_u32(smem), G_PLANE) : umma_desc_sw128(smem_u32(smem));
      // NT: planes 16 KB apart, a k-step is 32 bytes inside the 128-byte row.  TN: a box is [hi 8 KB | lo 8 KB], the two
      // 64-column blocks of an operand are G_PLANE apart (= LBO), a k-step is two 8-row groups = 2048 bytes.
      constexpr uint32_t lo_off = (TN ? G_PLANE / 2 : G_PLANE) >> 4, kstep = (TN ? 2048 : 32) >> 4;
      int it = 0, lt = 0;
      for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++lt) {
        const int acc = lt & 1;
        mbar_wait(&tempty[acc], ((lt >> 1) & 1) ^ 1);                      // epilogue drained this accumulator
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * GB_N;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = it % G_STAGES, ph = (it / G_STAGES) & 1;
          mbar_wait(&full[s], ph);
          tc_fence_after();
          const uint64_t dah = base + (uint32_t)(s * (G_STAGE_BYTES >> 4)), dal = dah + lo_off;
          const uint64_t dbh = dah + (2 * G_PLANE >> 4), dbl = dbh + lo_off;
#pragma unroll
          for (int k = 0; k < GB_K / 16; ++k) {
            umma_bf16(tmem_d, dal + kstep * k, dbh + kstep * k, idesc, (kb | k) ? 1u : 0u);   // small terms first
            umma_bf16(tmem_d, dah + kstep * k, dbl + kstep * k, idesc, 1u);
            umma_bf16(tmem_d, dah + kstep * k, dbh + kstep * k, idesc, 1u);
          }
          umma_commit(&empty[s]);
        }
        umma_commit(&tfull[acc]);
      }
    }
  } else {
    // Epilogue, 8 warps: warp w may touch TMEM lanes [32*(w%4), +32); warps 2..5 drain columns [0, 64) of the tile,
    // warps 6..9 columns [64, 128).  A thread owns an accumulator ROW after tcgen05.ld; one transposing pass through
    // shared memory (4 KB per warp, 16-byte pieces XOR-swizzled with the row so both passes are conflict-free) turns
    // that into the store role -- a fixed column quad, 8 rows -- where a warp instruction covers 4 rows x 128 contiguous
    // bytes and alpha / bias / beta are applied per column quad (one bias load per 32-column chunk, the `beta` reads of a
    // chunk all issued before the first use).  The per-warp instruction stream is what bounds a short-K GEMM:
    // profiles/r02_summary.md (ncu, K = 257: 9.4k warp-instructions per tile with 4 warps -> this form).
    const int q = warp & 3, ch = (warp - 2) >> 2;
    const int sr = lane >> 3, cqi = lane & 7;                  // store role: row within a group of 4, column quad
    float4* stg = reinterpret_cast<float4*>(epi) + (warp - 2) * 256;        // this warp's 32 x 8 float4 staging block
    const bool unit4 = p.rowmap == SS_ROWMAP_UNIT4;
    const long long rstep = unit4 ? p.ldc : 4 * p.ldc;         // rows rl = 4 rr + sr: output pointer advance per rr
    int lt = 0;
    for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++lt) {
      const int z = tile / (mt * nt), rem = tile - z * mt * nt;
      const int m0 = (rem / nt) * GB_M, n0 = (rem % nt) * GB_N;
      const int acc = lt & 1;
      const int mw = m0 + q * 32;                              // first row of this warp (a multiple of 32)
      mbar_wait(&tfull[acc], (lt >> 1) & 1);
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * GB_N + ((uint32_t)(q * 32) << 16);
      float* Cz = p.C[z];
      const bool vec = ((p.ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(Cz) & 15) == 0);
      // row mw + sr of the store role: natural, or gate-interleaved 4u+g -> g*hidden + u (mw % 4 == 0, so g = sr)
      const long long orow0 = unit4 ? (long long)sr * p.hidden + (mw >> 2) : (long long)(mw + sr);
      const int rows_ok = p.M - mw - sr;                       // row rl = 4 rr + sr exists iff 4 rr < rows_ok
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = ch * 64 + cc * 32, nb = n0 + c;
        if (nb >= p.N) break;                                  // warp-uniform
        float v[32];
        tmem_ld32(tmem_d + c, v);
        if (vec && nb + 32 <= p.N) {
#pragma unroll
          for (int i = 0; i < 8; ++i)
            stg[lane * 8 + (i ^ (lane & 7))] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
          __syncwarp();
          float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
          if (p.bias) {
            const float* bp = p.bias + nb + 4 * cqi;
            b4 = make_float4(bp[0], bp[1], bp[2], bp[3]);
          }
          float* dst = Cz + orow0 * p.ldc + nb + 4 * cqi;
          float4 o[8];
#pragma unroll
          for (int rr = 0; rr < 8; ++rr) {
            const int rl = rr * 4 + sr;
            o[rr] = stg[rl * 8 + (cqi ^ (rl & 7))];
            o[rr].x = fmaf(p.alpha, o[rr].x, b4.x); o[rr].y = fmaf(p.alpha, o[rr].y, b4.y);
            o[rr].z = fmaf(p.alpha, o[rr].z, b4.z); o[rr].w = fmaf(p.alpha, o[rr].w, b4.w);
          }
          if (p.beta != 0.0f) {
            float4 old[8];
#pragma unroll
            for (int rr = 0; rr < 8; ++rr)
              old[rr] = 4 * rr < rows_ok ? *reinterpret_cast<const float4*>(dst + rr * rstep) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int rr = 0; rr < 8; ++rr) {
              o[rr].x = fmaf(p.beta, old[rr].x, o[rr].x); o[rr].y = fmaf(p.beta, old[rr].y, o[rr].y);
              o[rr].z = fmaf(p.beta, old[rr].z, o[rr].z); o[rr].w = fmaf(p.beta, old[rr].w, o[rr].w);
            }
          }
#pragma unroll
          for (int rr = 0; rr < 8; ++rr)
            if (4 * rr < rows_ok) *reinterpret_cast<float4*>(dst + rr * rstep) = o[rr];
          __syncwarp();
        } else {
          // edge columns / unaligned output: the thread stores its own row
          const int m = mw + lane;
          if (m < p.M) {
            const long long orow = unit4 ? (long long)(m & 3) * p.hidden + (m >> 2) : (long long)m;
            float* crow = Cz + orow * p.ldc;
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (nb + i < p.N) {
                float r = p.alpha * v[i];
                if (p.bias) r += p.bias[nb + i];
                crow[nb + i] = r + (p.beta != 0.0f ? p.beta * crow[nb + i] : 0.0f);
              }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tempty[acc]);         // this warp's TMEM lanes / columns of the accumulator are free again
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_free<G_TMEM>(tmem_base);
  }
}

// tensor map over an sb16 matrix: dims (K, rows, 2 planes), box (64, 128, 2)
static int sb16_tmap(CUtensorMap* tm, const void* base, long long rows, long long K, long long pitch, long long plane,
                     int box_rows) {
  uint64_t dims[3] = {(uint64_t)K, (uint64_t)rows, 2};
  uint64_t strides[2] = {(uint64_t)pitch * 2, (uint64_t)plane * 2};
  uint32_t box[3] = {64, (uint32_t)box_rows, 2};
  return make_tmap_bf16(tm, base, 3, dims, strides, box);
}

}  // namespace ssb

static int gemm_grid(int tiles, int sm_limit) {
  static int n_sm = 0;
  if (!n_sm) {
    cudaFuncSetAttribute(ssb::gemm_sb16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ssb::G_SMEM);
    cudaFuncSetAttribute(ssb::gemm_sb16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ssb::G_SMEM);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
  }
  int grid = tiles < n_sm ? tiles : n_sm;
  if (sm_limit > 0 && grid > sm_limit) grid = sm_limit;      // run beside a resident recurrence kernel on its idle SMs
  return grid;
}

static int gemm_launch(const void* A, long long lda, long long a_plane, int a_rows, const void* B, long long ldb,
                       long long b_plane, int b_rows, ssb::GemmParams p, int sm_limit, void* stream) {
  using namespace ssb;
  SS_REQUIRE(p.M > 0 && p.N > 0 && p.K > 0, "ss_gemm_sb16: empty problem %dx%dx%d", p.M, p.N, p.K);
  SS_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && a_plane % 8 == 0 && b_plane % 8 == 0,
             "ss_gemm_sb16: pitches/plane strides must be multiples of 8 elements (16 bytes)");
  SS_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "ss_gemm_sb16: operands must be 16-byte aligned");
  SS_REQUIRE(p.rowmap == SS_ROWMAP_NONE || (p.rowmap == SS_ROWMAP_UNIT4 && p.hidden > 0 && p.M == 4 * p.hidden),
             "ss_gemm_sb16: bad rowmap");
  CUtensorMap tmA, tmB;
  int rc = sb16_tmap(&tmA, A, a_rows, p.K, lda, a_plane, GB_M);
  if (rc) return rc;
  rc = sb16_tmap(&tmB, B, b_rows, p.K, ldb, b_plane, GB_N);
  if (rc) return rc;
  const int tiles = p.batch * cdiv(p.N, GB_N) * cdiv(p.M, GB_M);
  gemm_sb16_kernel<false><<<gemm_grid(tiles, sm_limit), G_THREADS, G_SMEM, (cudaStream_t)stream>>>(tmA, tmB, p);
  return check_launch("ss_gemm_sb16");
}

extern "C" int ss_gemm_sb16(const void* A, long long lda, long long a_plane, const void* B, long long ldb,
                            long long b_plane, float* C, long long ldc, int M, int N, int K, const float* bias,
                            float alpha, float beta, int rowmap, int hidden, int sm_limit, void* stream) {
  ssb::GemmParams p{{C, nullptr}, ldc, bias, alpha, beta, M, N, K, rowmap, hidden, 1, 0, 0};
  return gemm_launch(A, lda, a_plane, M, B, ldb, b_plane, N, p, sm_limit, stream);
}

extern "C" int ss_gemm_sb16_x2(const void* A, long long lda, long long a_plane, int a_batch_rows, const void* B,
                               long long ldb, long long b_plane, int b_batch_rows, float* C0, float* C1,
                               long long ldc, int M, int N, int K, const float* bias, float alpha, float beta,
                               int rowmap, int hidden, int sm_limit, void* stream) {
  ssb::GemmParams p{{C0, C1}, ldc, bias, alpha, beta, M, N, K, rowmap, hidden, 2, a_batch_rows, b_batch_rows};
  return gemm_launch(A, lda, a_plane, a_batch_rows + M, B, ldb, b_plane, b_batch_rows + N, p, sm_limit, stream);
}

extern "C" int ss_gemm_sb16_tn(const void* A, long long lda, long long a_plane, int a_cols, int a_col0, int a_col_batch,
                               const void* B, long long ldb, long long b_plane, int b_cols, int b_col0, int b_col_batch,
                               int R, int blk0_z0, int blk0_z1, int n_blk, int b_shift_z0, int b_shift_z1, int batch,
                               float* C0, float* C1, long long ldc, int M, int N, float alpha, float beta, int rowmap,
                               int hidden, int sm_limit, void* stream) {
  using namespace ssb;
  SS_REQUIRE(M > 0 && N > 0 && R > 0 && n_blk > 0 && (batch == 1 || batch == 2), "ss_gemm_sb16_tn: empty problem");
  SS_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && a_plane % 8 == 0 && b_plane % 8 == 0 && a_col0 % 8 == 0 && b_col0 % 8 == 0 &&
                 a_col_batch % 8 == 0 && b_col_batch % 8 == 0,
             "ss_gemm_sb16_tn: pitches, plane strides and column offsets must be multiples of 8 elements (16 bytes)");
  SS_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "ss_gemm_sb16_tn: operands must be 16-byte aligned");
  SS_REQUIRE(rowmap == SS_ROWMAP_NONE || (rowmap == SS_ROWMAP_UNIT4 && hidden > 0 && M == 4 * hidden),
             "ss_gemm_sb16_tn: bad rowmap");
  SS_REQUIRE(blk0_z0 >= 0 && blk0_z1 >= 0 && (long long)(blk0_z0 + n_blk - 1) * GB_K < R &&
                 (long long)(blk0_z1 + n_blk - 1) * GB_K < R,
             "ss_gemm_sb16_tn: row blocks outside the %d reduction rows", R);
  CUtensorMap tmA, tmB;
  {  // (columns, reduction rows, plane), box 64 x 64 x 2; rows outside [0, R) and columns outside the matrix read as zero
    uint64_t dims[3] = {(uint64_t)a_cols, (uint64_t)R, 2};
    uint64_t str[2] = {(uint64_t)lda * 2, (uint64_t)a_plane * 2};
    uint32_t box[3] = {64, 64, 2};
    int rc = make_tmap_bf16(&tmA, A, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)b_cols, (uint64_t)R, 2};
    uint64_t str[2] = {(uint64_t)ldb * 2, (uint64_t)b_plane * 2};
    uint32_t box[3] = {64, 64, 2};
    int rc = make_tmap_bf16(&tmB, B, 3, dims, str, box);
    if (rc) return rc;
  }
  GemmParams p{{C0, batch == 2 ? C1 : nullptr}, ldc, nullptr, alpha, beta, M, N, n_blk * GB_K, rowmap, hidden, batch,
               a_col_batch, b_col_batch};
  p.a_col0 = a_col0;
  p.b_col0 = b_col0;
  p.nkb = n_blk;
  p.kb0[0] = blk0_z0;
  p.kb0[1] = blk0_z1;
  p.b_shift[0] = b_shift_z0;
  p.b_shift[1] = b_shift_z1;
  const int tiles = batch * cdiv(N, GB_N) * cdiv(M, GB_M);
  gemm_sb16_kernel<true><<<gemm_grid(tiles, sm_limit), G_THREADS, G_SMEM, (cudaStream_t)stream>>>(tmA, tmB, p);
  return check_launch("ss_gemm_sb16_tn");
}
