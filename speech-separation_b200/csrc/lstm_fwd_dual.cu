// K2b (dual-context fast path): the forward recurrence of lstm_fwd.cu for H = 600 and 32 < B <= 64, with the batch
// split into two halves ("contexts", rows 0..31 and 32..63) that run as two INDEPENDENT recurrences inside every CTA.
// A step of the recurrence is a latency chain (publish h -> gpu-scope release -> peers observe -> bulk loads ->
// MMA -> gates, profiles/r01_summary.md section C), during most of which the CTA's tensor pipe and copy engine
// idle; the two halves do not depend on each other, so while one waits on its grid barrier the other one loads and
// multiplies.  Same arithmetic per row as the single-context kernel (same k order, same quadrant sums).
//
//   * per direction and context one release/acquire counter (4 per launch) and one ping-pong h buffer
//     hbuf[slot][dir][ctx][kc][plane][32 rows][64], rows pre-swizzled for SWIZZLE_128B;
//   * operand stage = 8 KB = [h_hi 32 rows][h_lo 32 rows]; the MMA still has M = 128 (a tcgen05.mma costs the same
//     for M = 64): context 0 points the descriptor at the stage, so its rows land in TMEM lanes 0..63, context 1
//     points it 8 KB BEFORE the stage, so its rows land in lanes 64..127; the other 64 rows of either tile are
//     whatever lies next to the stage (never read back).  Each epilogue warp therefore serves exactly one context
//     (lane quadrant q: context q >> 1, hi rows for even q, lo rows for odd q);
//   * two TMEM accumulators (columns 0 and 128), one accum_full barrier per context;
//   * producer and MMA threads walk the fixed order ctx0(s), ctx1(s), ctx0(s+1), ... over a 10-stage ring.
//
// STATUS: experimental, opt-in with SSB_FWD_DUAL=1, parity-green but SLOWER than the single-context kernel today
// (8.6 vs 6.7 ms per step of two layers at B=64, T=501).  tools/trace_lstm_dual.py shows why: (1) a bulk copy costs
// the SM ~230-250 cycles to issue whatever its size (tools/microbench/bulk_size.cu), so ten 8 KB copies take as long
// as ten 16 KB ones; (2) the MMA-issuing thread needs ~390 cycles per 4-MMA chunk when alone but 1000-1600 while the
// OTHER context's epilogue warps run (cell math, release, gate stores), which is exactly the overlap this design
// wanted.  Kept as the starting point for the next round (fewer, larger copies; MMA issue off the contended path).
#include <stdlib.h>

#include "../../include/ssb200.h"
#include "lstm_fwd.cuh"

namespace ssb {

constexpr int LD_H = 600, LD_KC = 10, LD_NS = 10;
constexpr int LD_STAGE = 2 * 32 * 128;           // bytes
constexpr int LD_TMEM = 256;
constexpr int LD_CTX_THREADS = 32 * LF_EPI_WARPS / 2;

// DBG: clock64 stamps of CTA 0 into p.dbg[step][32]: slot 16*ctx + {0 poll start, 1 counter observed, 2 loads issued,
// 3 first chunk landed, 4 MMAs committed, 5 accumulator seen by the epilogue, 6 cell math done, 7 context barrier
// passed, 8 release done}
#define LD_STAMP(slot)                                                         \
  do {                                                                         \
    if (DBG && blockIdx.x == 0) p.dbg[s * 64 + 16 * ctx + (slot)] = clock64(); \
  } while (0)

template <bool DBG>
__global__ void __launch_bounds__(LF_THREADS, 1)
lstm_fwd_dual_kernel(const __grid_constant__ CUtensorMap tmW, LstmFwdParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* wsm = smem;                                               // KC chunks of the W slice
  unsigned char* ring = wsm + LD_KC * LF_WCHUNK;                           // NS stages of h_prev chunks
  float* xbuf = reinterpret_cast<float*>(ring + LD_NS * LD_STAGE);         // [ctx][32][LF_XPITCH] lo-row partial sums
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + 64 * LF_XPITCH);
  uint64_t* empty = full + LD_NS;
  uint64_t* accum_full = empty + LD_NS;                                    // [2]
  uint64_t* wfull = accum_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int u0 = cta * LF_U;
  const int T = p.T, B = p.B;
  constexpr int H = LD_H;
  constexpr long long blk = 2LL * 32 * 64;                                 // elements per (slot,dir,ctx,kc) block

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < LD_NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(&accum_full[0], 1);
    mbar_init(&accum_full[1], 1);
    mbar_init(wfull, 1);
    fence_barrier_init();
  }
  // roles: warps 0..11 epilogue, warp 12 producer, warp 13 MMA issuer
  constexpr int W_PROD = LF_EPI_WARPS, W_MMA = LF_EPI_WARPS + 1;
  if (warp == W_MMA) tmem_alloc<LD_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == W_PROD) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, LD_KC * LF_WCHUNK);
      for (int kc = 0; kc < LD_KC; ++kc) tma_load_3d(wsm + kc * LF_WCHUNK, &tmW, wfull, kc * 64, dir * 4 * H + 4 * u0, 0);
      for (int s = 0; s < T; ++s) {
        const unsigned int need = (unsigned int)p.NC * (unsigned int)(s + 1);
#pragma unroll
        for (int ctx = 0; ctx < 2; ++ctx) {
          const unsigned int* counter = p.sync + (dir * 2 + ctx) * 32;
          LD_STAMP(0);
          while (ld_acquire(counter) < need) {
          }
          LD_STAMP(1);
          const __nv_bfloat16* src = p.hbuf + ((((long long)(s & 1)) * 2 + dir) * 2 + ctx) * LD_KC * blk;
#pragma unroll
          for (int i = 0; i < LD_KC; ++i) {
            mbar_wait(&empty[i], ctx ^ 1);                     // use number 2s + ctx of stage i
            mbar_arrive_expect_tx(&full[i], LD_STAGE);
            bulk_load(ring + i * LD_STAGE, src + i * blk, LD_STAGE, &full[i]);
            if (DBG && blockIdx.x == 0 && ctx == 0 && i < 6) p.dbg[s * 64 + 42 + i] = clock64();
          }
          LD_STAMP(2);
        }
      }
    }
  } else if (warp == W_MMA) {
    if (lane == 0) {
      constexpr uint32_t idesc96 = umma_idesc_bf16(128, 2 * LF_N);
      const uint64_t a_base = umma_desc_sw128(smem_u32(ring)), b_base = umma_desc_sw128(smem_u32(wsm));
      mbar_wait(wfull, 0);
      for (int s = 0; s < T; ++s) {
#pragma unroll
        for (int ctx = 0; ctx < 2; ++ctx) {
#pragma unroll
          for (int i = 0; i < LD_KC; ++i) {
            mbar_wait(&full[i], ctx);
            if (i == 0) LD_STAMP(3);
            if (DBG && blockIdx.x == 0) p.dbg[s * 64 + 32 + 16 * ctx + i] = clock64();
            tc_fence_after();
            // context 1: tile starts one stage earlier, its 64 rows are tile rows 64..127
            const uint64_t da = a_base + (uint64_t)((long long)(i - ctx) * (LD_STAGE >> 4));
            const uint64_t db = b_base + (uint32_t)i * (LF_WCHUNK >> 4);
            constexpr int last = (H - (LD_KC - 1) * 64 + 15) / 16;
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (i < LD_KC - 1 || k < last) umma_bf16(tmem_d + 128 * ctx, da + 2 * k, db + 2 * k, idesc96, (i | k) ? 1u : 0u);
            umma_commit(&empty[i]);
          }
          LD_STAMP(14);
          umma_commit(&accum_full[ctx]);
          LD_STAMP(4);
        }
      }
    }
  } else {
    const int q = warp & 3, cg = warp >> 2;
    const int ctx = q >> 1;
    const bool helper = q & 1;                                 // these warps own the h_lo rows of D
    const int b = ctx * 32 + lane;
    const int ug = u0 + cg * 4;                                // this thread's 4 units within the direction
    const bool row_ok = !helper && b < B;
    const bool leader = (warp == 2 * ctx) && lane == 0;
    const int len = row_ok ? p.lens[b] : 0;
    unsigned int* counter = p.sync + (dir * 2 + ctx) * 32;
    const int bar0 = 1 + 3 * ctx;
    float h[4], c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = row_ok ? p.h0[((long long)dir * B + b) * H + ug + j] : 0.0f;
      c[j] = row_ok ? p.c0[((long long)dir * B + b) * H + ug + j] : 0.0f;
    }
    // byte offset of this thread's 4 units inside the pre-swizzled 128-byte row of K-chunk kc_w
    const int kc_w = ug >> 6;
    const int o = (ug & 63) * 2;
    const long long row_off = (long long)lane * 64 + ((((o >> 4) ^ (lane & 7)) << 4) + (o & 15)) / 2;   // elements
    int s = 0;
    auto publish = [&](int slot) {
      if (row_ok) {
        __nv_bfloat16 hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_bf16(h[j], hi[j], lo[j]);
        __nv_bfloat16* dst = p.hbuf + (((((long long)slot) * 2 + dir) * 2 + ctx) * LD_KC + kc_w) * blk + row_off;
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<uint2*>(hi);
        *reinterpret_cast<uint2*>(dst + 32 * 64) = *reinterpret_cast<uint2*>(lo);
      }
      fence_proxy_async();                                     // writer side of the generic -> async proxy hand-over
      // context barrier, then ONE thread releases at gpu scope (cumulative over the other threads' stores)
      asm volatile("bar.sync %0, %1;" ::"r"(bar0), "n"(LD_CTX_THREADS) : "memory");
      if (leader) {
        LD_STAMP(7);
        if (p.EXP & 8) atomicAdd(counter, 1u); else red_release_add(counter, 1u);   // 8: timing experiment, no release
        LD_STAMP(8);
      }
    };
    publish(0);
    const uint32_t trow = tmem_d + 128 * ctx + ((uint32_t)(q * 32) << 16);
    float* xr = xbuf + ((long long)ctx * 32 + lane) * LF_XPITCH + 16 * cg;
    for (s = 0; s < T; ++s) {
      const int t = dir ? (T - 1 - s) : s;
      const bool active = row_ok && t < len;
      const long long row = (long long)t * B + (row_ok ? b : 0);
      float4* gptr = reinterpret_cast<float4*>(p.gx + row * (8LL * H) + (long long)dir * 4 * H + 4 * ug);
      float4 g4[4];
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) g4[j] = gptr[j];
      }
      // one warp per context polls the mbarrier; the others sleep on a hardware barrier
      if (warp == 2 * ctx) mbar_wait(&accum_full[ctx], s & 1);
      asm volatile("bar.sync %0, %1;" ::"r"(bar0 + 1), "n"(LD_CTX_THREADS) : "memory");
      if (leader) LD_STAMP(5);
      tc_fence_after();
      float d[16], e[16];
      tmem_ld16(trow + 16 * cg, d);
      tmem_ld16(trow + LF_N + 16 * cg, e);
      tc_fence_before();
#pragma unroll
      for (int i = 0; i < 16; ++i) d[i] += e[i];
      if (helper) {
#pragma unroll
        for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(xr)[i] = make_float4(d[4 * i], d[4 * i + 1], d[4 * i + 2], d[4 * i + 3]);
      }
      asm volatile("bar.sync %0, %1;" ::"r"(bar0 + 2), "n"(LD_CTX_THREADS) : "memory");
      if (!helper) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 v = reinterpret_cast<const float4*>(xr)[i];
          d[4 * i] += v.x; d[4 * i + 1] += v.y; d[4 * i + 2] += v.z; d[4 * i + 3] += v.w;
        }
      }
      float hout[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gi = fsigmoid(d[4 * j + 0] + g4[j].x), gf = fsigmoid(d[4 * j + 1] + g4[j].y);
          const float gg = ftanh(d[4 * j + 2] + g4[j].z), go = fsigmoid(d[4 * j + 3] + g4[j].w);
          c[j] = gf * c[j] + gi * gg;
          h[j] = go * ftanh(c[j]);
          hout[j] = h[j];
          g4[j] = make_float4(gi, gf, gg, go);
        }
      }
      if (leader) LD_STAMP(6);
      publish((s + 1) & 1);                                    // carried state for the next step (critical path first)
      if (row_ok && !(p.EXP & 4)) {
        if (active) {
#pragma unroll
          for (int j = 0; j < 4; ++j) __stcs(gptr + j, g4[j]);
          if (p.cs) __stcs(reinterpret_cast<float4*>(p.cs + row * (2LL * H) + dir * H + ug), make_float4(c[0], c[1], c[2], c[3]));
        }
        __nv_bfloat16 hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_bf16(hout[j], hi[j], lo[j]);
        __nv_bfloat16* op = p.out + row * (2LL * H) + dir * H + ug;
        __stcs(reinterpret_cast<uint2*>(op), *reinterpret_cast<uint2*>(hi));
        __stcs(reinterpret_cast<uint2*>(op + p.out_plane), *reinterpret_cast<uint2*>(lo));
      }
    }
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        p.hn[((long long)dir * B + b) * H + ug + j] = h[j];
        p.cn[((long long)dir * B + b) * H + ug + j] = c[j];
      }
    }
  }
  __syncthreads();
  if (warp == W_MMA) {
    tc_fence_after();
    tmem_free<LD_TMEM>(tmem_d);
  }
}

// p.hbuf / p.sync point into the zeroed workspace (ss_lstm_ws_bytes): 2*2*2*10 blocks of 8 KB, then 4 counters 128 B apart
int launch_lstm_fwd_dual(const CUtensorMap& tmW, LstmFwdParams p, cudaStream_t st) {
  const int smem = LD_KC * LF_WCHUNK + LD_NS * LD_STAGE + 64 * LF_XPITCH * 4 + 1024 + 256;
  void* kern = p.dbg ? (void*)lstm_fwd_dual_kernel<true> : (void*)lstm_fwd_dual_kernel<false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&tmW, (void*)&p};
  cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(2 * p.NC), dim3(LF_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_fwd: cooperative launch (dual) failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}

}  // namespace ssb
