// K2b: persistent bidirectional LSTM recurrence (forward pass), one cooperative launch per layer.
// Replaces the sequential part of cudnnRNNForward behind `self.blstm(x, self.hidden)` (archs/uPIT.py:115,132;
// archs/RSH.py:155,172) for a length-sorted padded batch with PackedSequence semantics:
//   * the reverse direction starts at each utterance's own last frame, from (h0, c0);
//   * outputs at t >= len_b are zero, (h_n, c_n) are the states after each utterance's last processed frame.
// Each CTA owns LF_U hidden units of one direction with their 4 gates interleaved (column 4*u+gate), keeps its
// slice of W_hh (split-bf16, hi rows then lo rows) resident in shared memory, and per step computes
//   D[b][4u+g] = sum_k h_prev[b][k] * W_hh[g*H+u][k]      (tcgen05, fp32 accumulate in TMEM)
// from h_prev (split-bf16) streamed by bulk async copies out of a small ping-pong buffer that all CTAs of the
// direction publish into; a gpu-scope release/acquire counter per direction is the per-step grid barrier.
//
// tcgen05.mma costs >= ~53 cycles per instruction whatever its shape (tools/microbench/mma_rate.cu), so the
// three split-bf16 products are folded into ONE instruction per 16-wide k-step: for B <= 64 the A operand is the
// 128-row stack [h_hi; h_lo] and the B operand the 96-row stack [W_hi; W_lo]; the four 64x48 quadrants of D
// (hi*hi, hi*lo, lo*hi, lo*lo) are summed in the epilogue.
#include <stdlib.h>

#include "../../include/ssb200.h"
#include "lstm_fwd.cuh"

namespace ssb {

#define LF_STAMP(slot)                                                        \
  do {                                                                        \
    if (!FAST && p.dbg && blockIdx.x == 0) p.dbg[s * 32 + (slot)] = clock64(); \
  } while (0)

// FAST: H = 600, B <= 64 (R = 64), no trace -- the reference's shape; every loop bound is a constant so
// the single MMA-issuing thread runs a fully unrolled instruction stream (its scalar overhead, not the tensor
// pipe, paced the generic loop: tools/microbench/bulk_rate.cu).
template <bool FAST>
__global__ void __launch_bounds__(LF_THREADS, 1)
lstm_fwd_kernel(const __grid_constant__ CUtensorMap tmW, LstmFwdParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int PR = FAST ? 64 : p.R, PKC = FAST ? 10 : p.KC, PNS = FAST ? 5 : p.NS;
  const int stage_bytes = 2 * PR * 128;
  const bool stacked = PR == 64;                               // A operand = [h_hi; h_lo] as one 128-row tile
  unsigned char* ring = smem;                                  // NS stages of h_prev chunks
  unsigned char* wsm = smem + PNS * stage_bytes;               // KC chunks of the W slice
  float* xbuf = reinterpret_cast<float*>(wsm + PKC * LF_WCHUNK);           // [64][LF_XPITCH] lo-row partial sums
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + 64 * LF_XPITCH);
  uint64_t* empty = full + 8;
  uint64_t* accum_full = empty + 8;
  uint64_t* wfull = accum_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int u0 = cta * LF_U;
  const int T = p.T, B = p.B, H = FAST ? 600 : p.H;
  unsigned int* counter = p.sync + dir;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < PNS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(accum_full, 1);
    mbar_init(wfull, 1);
    fence_barrier_init();
  }
  // roles: warps 0..11 epilogue, warp 12 producer, warp 13 MMA issuer
  constexpr int W_PROD = LF_EPI_WARPS, W_MMA = LF_EPI_WARPS + 1;
  if (warp == W_MMA) tmem_alloc<LF_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == W_PROD) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, PKC * LF_WCHUNK);
      for (int kc = 0; kc < PKC; ++kc) tma_load_3d(wsm + kc * LF_WCHUNK, &tmW, wfull, kc * 64, dir * 4 * H + 4 * u0, 0);
      const long long blk = 2LL * PR * 64;                     // elements per (slot,dir,kc) block
      int it = 0;
      for (int s = 0; s < T; ++s) {
        const unsigned int need = (unsigned int)p.NC * (unsigned int)(s + 1);
        LF_STAMP(0);
        while (ld_acquire(counter) < need) {
        }
        LF_STAMP(1);
        // (the generic->async proxy fence sits on the writer side, in publish(): there it overlaps the other
        // threads' stores instead of adding ~1000 cycles between the acquire and the first copy)
        const __nv_bfloat16* src = p.hbuf + ((long long)(s & 1) * 2 + dir) * PKC * blk;
#pragma unroll
        for (int i = 0; i < PKC; ++i, ++it) {
          const int kc = i;
          const int st = FAST ? i % 5 : it % PNS, ph = FAST ? (i / 5) & 1 : (it / PNS) & 1;
          mbar_wait(&empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&full[st], stage_bytes);
          bulk_load(ring + st * stage_bytes, src + kc * blk, stage_bytes, &full[st]);
        }
        // Progress for stream-ordered consumers (ss_stream_wait_value32): the `out` stores of step j precede, in every
        // CTA, the release of step j+1's publish, i.e. counter >= NC*(j+3); the counter observed at the top of step s therefore
        // covers the outputs of steps 0..s-2.  One CTA per direction reports, every 4th step, after its copies are
        // in flight (the fence then overlaps the MMAs instead of delaying them).
        if (p.progress && cta == 0 && s >= 2 && (s & 3) == 0) {
          __threadfence_system();
          *reinterpret_cast<volatile unsigned int*>(p.progress + dir) = p.prog_base + (unsigned int)(s - 1);
        }
        LF_STAMP(2);
      }
    }
  } else if (warp == W_MMA) {
    if (lane == 0) {
      const uint32_t idesc96 = umma_idesc_bf16(128, 2 * LF_N), idesc48 = umma_idesc_bf16(128, LF_N);
      const uint64_t a_base = umma_desc_sw128(smem_u32(ring)), b_base = umma_desc_sw128(smem_u32(wsm));
      const uint32_t a_stage = stage_bytes >> 4, a_lo_off = (PR * 128) >> 4;
      mbar_wait(wfull, 0);
      int it = 0;
      for (int s = 0; s < T; ++s) {
#pragma unroll
        for (int i = 0; i < PKC; ++i, ++it) {
          const int kc = i;
          const int st = FAST ? i % 5 : it % PNS, ph = FAST ? (i / 5) & 1 : (it / PNS) & 1;
          mbar_wait(&full[st], ph);
          LF_STAMP(i == 0 ? 3 : 10 + i);
          tc_fence_after();
          const uint64_t da = a_base + (uint32_t)st * a_stage;
          const uint64_t db = b_base + (uint32_t)kc * (LF_WCHUNK >> 4);
          const int ksteps = min(4, (H - kc * 64 + 15) / 16);
          if (stacked) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (k < ksteps) umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc96, (i | k) ? 1u : 0u);
          } else {
            for (int k = 0; k < ksteps; ++k) {
              umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc96, (i | k) ? 1u : 0u);           // h_hi x [W_hi; W_lo]
              umma_bf16(tmem_d, da + a_lo_off + 2 * k, db + 2 * k, idesc48, 1u);               // h_lo x W_hi
            }
          }
          umma_commit(&empty[st]);
        }
        umma_commit(accum_full);
        LF_STAMP(4);
      }
    }
  } else {
    const int q = warp & 3, cg = warp >> 2;
    const bool helper = stacked && q >= 2;                     // these warps own the h_lo rows of D
    const int b = (stacked ? (q & 1) : q) * 32 + lane;
    const int ug = u0 + cg * 4;                                // this thread's 4 units within the direction
    const bool row_ok = !helper && b < B;
    const int len = row_ok ? p.lens[b] : 0;
    float h[4], c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = row_ok ? p.h0[((long long)dir * B + b) * H + ug + j] : 0.0f;
      c[j] = row_ok ? p.c0[((long long)dir * B + b) * H + ug + j] : 0.0f;
    }
    // byte offset of this thread's 4 units inside the pre-swizzled 128-byte row of K-chunk kc_w
    const int kc_w = ug >> 6;
    const int o = (ug & 63) * 2;
    const long long row_off = (long long)b * 64 + ((((o >> 4) ^ (b & 7)) << 4) + (o & 15)) / 2;   // elements
    const long long blk = 2LL * PR * 64;
    int dbg_s = 0;
    auto publish = [&](int slot) {
      if (row_ok) {
        __nv_bfloat16 hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_bf16(h[j], hi[j], lo[j]);
        __nv_bfloat16* dst = p.hbuf + (((long long)slot * 2 + dir) * PKC + kc_w) * blk + row_off;
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<uint2*>(hi);
        *reinterpret_cast<uint2*>(dst + PR * 64) = *reinterpret_cast<uint2*>(lo);
      }
      // CTA barrier, then ONE thread releases at gpu scope (cumulative over the other threads' stores, the
      // cooperative-groups grid.sync pattern) and orders them against the consumers' async-proxy reads.
      fence_proxy_async();                                     // writer side: this thread's h stores -> async-proxy readers
      if (!FAST && threadIdx.x == 0 && p.dbg && blockIdx.x == 0) p.dbg[dbg_s * 32 + 20] = clock64();
      asm volatile("bar.sync 1, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
      if (threadIdx.x == 0) {
        if (!FAST && p.dbg && blockIdx.x == 0) p.dbg[dbg_s * 32 + 21] = clock64();
        red_release_add(counter, 1u);
        if (!FAST && p.dbg && blockIdx.x == 0) p.dbg[dbg_s * 32 + 23] = clock64();
      }
    };
    publish(0);
    const uint32_t trow = tmem_d + ((uint32_t)(q * 32) << 16);
    for (int s = 0; s < T; ++s) {
      const int t = dir ? (T - 1 - s) : s;
      const bool active = row_ok && t < len;
      const long long row = (long long)t * B + (row_ok ? b : 0);
      float4* gptr = reinterpret_cast<float4*>(p.gx + row * (8LL * H) + (long long)dir * 4 * H + 4 * ug);
      float4 g4[4];
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) g4[j] = gptr[j];
      }
      // one warp polls the mbarrier; the others sleep on a hardware barrier
      if (warp == 0) mbar_wait(accum_full, s & 1);
      asm volatile("bar.sync 2, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
      if (threadIdx.x == 0) LF_STAMP(5);
      tc_fence_after();
      float d[16], e[16];
      tmem_ld16(trow + 16 * cg, d);
      tmem_ld16(trow + LF_N + 16 * cg, e);
      tc_fence_before();
#pragma unroll
      for (int i = 0; i < 16; ++i) d[i] += e[i];
      if (stacked) {
        float* xr = xbuf + (long long)b * LF_XPITCH + 16 * cg;
        if (helper) {
#pragma unroll
          for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(xr)[i] = make_float4(d[4 * i], d[4 * i + 1], d[4 * i + 2], d[4 * i + 3]);
        }
        asm volatile("bar.sync 3, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
        if (!helper) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 v = reinterpret_cast<const float4*>(xr)[i];
            d[4 * i] += v.x; d[4 * i + 1] += v.y; d[4 * i + 2] += v.z; d[4 * i + 3] += v.w;
          }
        }
      }
      if (threadIdx.x == 0) LF_STAMP(8);
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
      if (threadIdx.x == 0) LF_STAMP(6);
      dbg_s = s;
      publish((s + 1) & 1);                                    // carried state for the next step (critical path first)
      if (threadIdx.x == 0) LF_STAMP(7);
      if (row_ok) {
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
    tmem_free<LF_TMEM>(tmem_d);
  }
}

}  // namespace ssb

namespace ssb {
long long* g_lstm_dbg = nullptr;
}
using ssb::g_lstm_dbg;
extern "C" int ss_lstm_debug(long long* buf) {   // development aid: clock64 trace [T][32] of CTA 0 of the next launches
  g_lstm_dbg = buf;
  return 0;
}

extern "C" long long ss_lstm_ws_bytes(int B, int H) {
  const long long R = B <= 64 ? 64 : 128;
  return 2LL * 2 * ssb::cdiv(H, 64) * 2 * R * 64 * 2 + 1024;          // h ping-pong buffer + sync counters
}

extern "C" int ss_lstm_fwd(float* gx, const void* whh_r, long long whh_ld, long long whh_plane, const float* h0,
                           const float* c0, const int* lens, void* out, long long out_plane, float* cs, float* hn,
                           float* cn, void* ws, unsigned int* progress, unsigned int prog_base, int T, int B, int H,
                           void* stream) {
  using namespace ssb;
  SS_REQUIRE(T > 0 && B > 0 && B <= 128, "ss_lstm_fwd: need 1 <= B <= 128 (got %d) and T > 0", B);
  SS_REQUIRE(H % LF_U == 0 && H % 8 == 0 && H <= 64 * LF_MAXKC, "ss_lstm_fwd: unsupported hidden size %d", H);
  cudaStream_t st = (cudaStream_t)stream;
  LstmFwdParams p;
  p.gx = gx; p.h0 = h0; p.c0 = c0; p.lens = lens;
  p.out = (__nv_bfloat16*)out; p.out_plane = out_plane;
  p.cs = cs; p.hn = hn; p.cn = cn;
  p.T = T; p.B = B; p.H = H;
  p.dbg = g_lstm_dbg;
  p.R = B <= 64 ? 64 : 128;
  p.NC = H / LF_U;
  p.KC = cdiv(H, 64);
  p.NS = p.R == 64 ? 5 : 2;
  p.progress = progress;
  p.prog_base = prog_base;
  const long long hbuf_bytes = 2LL * 2 * p.KC * 2 * p.R * 64 * 2;
  p.hbuf = (__nv_bfloat16*)ws;
  p.sync = (unsigned int*)((char*)ws + hbuf_bytes);
  cudaMemsetAsync(ws, 0, hbuf_bytes + 1024, st);

  CUtensorMap tmW;
  {
    uint64_t dims[3] = {(uint64_t)H, (uint64_t)8 * H, 2};
    uint64_t str[2] = {(uint64_t)whh_ld * 2, (uint64_t)whh_plane * 2};
    uint32_t box[3] = {64, LF_N, 2};
    int rc = make_tmap_bf16(&tmW, whh_r, 3, dims, str, box);
    if (rc) return rc;
  }
  const int smem = p.NS * 2 * p.R * 128 + p.KC * LF_WCHUNK + 64 * LF_XPITCH * 4 + 1024 + 256;
  const bool fast = H == 600 && p.R == 64 && p.dbg == nullptr;
  void* kern = fast ? (void*)lstm_fwd_kernel<true> : (void*)lstm_fwd_kernel<false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&tmW, (void*)&p};
  cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(2 * p.NC), dim3(LF_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_fwd: cooperative launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}
