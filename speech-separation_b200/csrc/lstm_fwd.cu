// K2b (generic path: any H <= 640 divisible by 12, B <= 128; the reference's shape H = 600, B <= 64 runs lstm_fwd_tw.cu):
// persistent bidirectional LSTM recurrence (forward pass), one cooperative launch per layer.
// Replaces the sequential part of cudnnRNNForward behind `self.blstm(x, self.hidden)` (archs/uPIT.py:115,132;
// archs/RSH.py:155,172) for a length-sorted padded batch with PackedSequence semantics:
//   * the reverse direction starts at each utterance's own last frame, from (h0, c0);
//   * outputs at t >= len_b are zero, (h_n, c_n) are the states after each utterance's last processed frame.
// Each CTA owns LF_U hidden units of one direction with their 4 gates interleaved (column 4*u+gate), keeps its
// slice of W_hh (split-bf16, hi rows then lo rows) resident in shared memory, and per step computes
//   D[b][4u+g] = sum_k h_prev[b][k] * W_hh[g*H+u][k]      (tcgen05, fp32 accumulate in TMEM)
// from h_prev (split-bf16) streamed by bulk async copies out of a small ping-pong buffer in L2 that all CTAs of the
// direction publish into; a gpu-scope release/acquire counter per direction is the per-step grid barrier.
//
// The exchange is the chain that bounds the kernel.  What it costs is mostly the SM's own memory pipe: a thread holds
// a few units of ONE batch row, so direct stores are 32 scattered 4..16-byte transactions per instruction, the SM
// retires about one transaction per cycle (1.7k per step), and the release and the next step's polls queue behind them.
// So (1) everything is staged in shared memory and stored row-wise (whole sectors), (2) only the carried state is
// stored before the release (ONE per CTA: twelve per-warp releases made 600 adds per step on one address, which queue up
// in L2), no proxy fence on the writer side -- and (3) the activated gates, cell states, `out` rows leave only after the next
// step's gate has opened.  On the reader side one otherwise idle thread polls the counter, crosses to the async proxy
// (fence.proxy.async is cheap in a thread with nothing outstanding) and issues the ten copies.
//
// tcgen05.mma costs >= ~53 cycles per instruction whatever its shape (tools/microbench/mma_rate.cu), so the
// three split-bf16 products are folded into ONE instruction per 16-wide k-step: for B <= 64 the A operand is the
// 128-row stack of h_hi / h_lo rows and the B operand the 96-row stack [W_hi; W_lo]; the four 64x48 quadrants of D
// (hi*hi, hi*lo, lo*hi, lo*lo) are summed in the epilogue.
#include <stdlib.h>

#include "../../include/ssb200.h"
#include "lstm_fwd.cuh"

namespace ssb {

#define LF_STAMP(slot)                                                         \
  do {                                                                         \
    if (TRACE && p.dbg && blockIdx.x == 0) p.dbg[s * 32 + (slot)] = clock64(); \
  } while (0)

__device__ __forceinline__ void lf_st_strong32(void* p, uint32_t v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void lf_st_strong64(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ uint32_t lf_pack(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (uint32_t)__bfloat16_as_ushort(a) | ((uint32_t)__bfloat16_as_ushort(b) << 16);
}

// (The reference's shape, H = 600 and B <= 64, runs lstm_fwd_tw.cu; this kernel serves every other shape.)
template <bool TRACE>
__global__ void __launch_bounds__(LF_THREADS, 1)
lstm_fwd_kernel(const __grid_constant__ CUtensorMap tmW, LstmFwdParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int PR = p.R, PKC = p.KC, PNS = p.NS;
  const int stage_bytes = 2 * PR * 128;
  // B <= 64: the A operand is ONE 128-row tile, rows [h_hi 0..31 | h_lo 0..31 | h_hi 32..63 | h_lo 32..63]: TMEM lane
  // quadrant 2 g + plane of D then belongs to batch rows 32 g .. 32 g + 31, and the two warps of a pair (hi / lo plane of
  // the same rows) swap half of their partial sums so that ALL twelve epilogue warps (all four schedulers' MUFU units)
  // do the gate math, two units per thread.
  const bool stacked = PR == 64;
  unsigned char* ring = smem;                                  // NS stages of h_prev chunks
  unsigned char* wsm = smem + PNS * stage_bytes;               // KC chunks of the W slice
  float* xbuf = reinterpret_cast<float*>(wsm + PKC * LF_WCHUNK);           // [64][LF_XPITCH] partial sums swapped between the planes
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + 64 * LF_XPITCH);     // chunk landed (bulk copy)
  uint64_t* empty = full + 8;                                  // chunk consumed (MMA commit)
  uint64_t* accum_full = empty + 8;
  uint64_t* wfull = accum_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);
  int* step_flag = reinterpret_cast<int*>(tmem_slot + 1);
  // staging tiles of the stacked path (after the 256 bytes of barriers)
  uint32_t* hst = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(full) + 256);   // [128 operand rows][8 words]: published h
  uint32_t* ost = hst + 128 * 8;                               // [plane * 64 + row][8 words]: `out`
  float* cst = reinterpret_cast<float*>(ost + 128 * 8);        // [64][12] cell states
  int* lens_s = reinterpret_cast<int*>(cst + 64 * 12);         // [64]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int u0 = cta * LF_U;
  const int T = p.T, B = p.B, H = p.H;
  const long long slot_elems = 2LL * PKC * stage_bytes / 2;   // bf16 elements of one slot (both directions)
  const long long dir_elems = (long long)PKC * stage_bytes / 2;
  // operand row of (plane, batch row) inside a chunk
  auto op_row = [&](int pl, int bb) { return stacked ? ((bb >> 5) * 64 + pl * 32 + (bb & 31)) : (pl * PR + bb); };
  // element offset inside a slot/direction of (plane, batch row, unit): chunk-major, rows pre-swizzled (SWIZZLE_128B)
  auto elem_off = [&](int pl, int bb, int u) {
    const int r = op_row(pl, bb), k = u & 63;
    return (long long)(u >> 6) * (stage_bytes / 2) + r * 64 + ((((k >> 3) ^ (r & 7)) << 3) | (k & 7));
  };

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    *step_flag = 0;
    for (int s = 0; s < PNS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(accum_full, 1);
    mbar_init(wfull, 1);
    fence_barrier_init();
  }
  // roles: warps 0..11 epilogue, warp 12 gate + copies, warp 13 MMA issuer, warp 14 progress
  constexpr int W_COPY = LF_EPI_WARPS, W_MMA = LF_EPI_WARPS + 1, W_REPORT = LF_EPI_WARPS + 2;
  if (warp == W_MMA) tmem_alloc<LF_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == W_MMA) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, PKC * LF_WCHUNK);
      for (int kc = 0; kc < PKC; ++kc) tma_load_3d(wsm + kc * LF_WCHUNK, &tmW, wfull, kc * 64, dir * 4 * H + 4 * u0, 0);
      const uint32_t idesc96 = umma_idesc_bf16(128, 2 * LF_N), idesc48 = umma_idesc_bf16(128, LF_N);
      const uint64_t a_base = umma_desc_sw128(smem_u32(ring)), b_base = umma_desc_sw128(smem_u32(wsm));
      const uint32_t a_stage = stage_bytes >> 4, a_lo_off = (PR * 128) >> 4;
      mbar_wait(wfull, 0);
      int it = 0;
      for (int s = 0; s < T; ++s) {
#pragma unroll
        for (int i = 0; i < PKC; ++i, ++it) {
          const int kc = i;
          const int st = it % PNS, ph = (it / PNS) & 1;
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
  } else if (warp == W_COPY) {
    // ---------------- gate + copies: one thread waits for the step's releases (one per CTA of the direction), crosses to
    // the async proxy and streams the ten chunks.  (A counter per K chunk, so that the chunks whose producers are done are
    // copied while the slowest CTA is still computing, was tried twice: fence.proxy.async waits for the bulk copies the
    // executing thread has in flight -- 1.3k cycles per chunk -- and one polling lane per chunk made the warp's divergent
    // spin loops delay every copy by 3k cycles: profiles/r02_summary.md section J.)
    unsigned int* counter = p.done + 128 + dir;
    int it = 0;
    for (int s = 0; s < T; ++s) {
      const __nv_bfloat16* src = p.hbuf + (s & 1) * slot_elems + dir * dir_elems;
      if (lane == 0) {
        LF_STAMP(0);
        const unsigned int need = (unsigned int)p.NC * (unsigned int)(s + 1);
        while (ld_acquire(counter) < need) {
        }
        fence_proxy_async();                                   // the producers' generic-proxy stores -> this thread's bulk copies
        LF_STAMP(1);
      }
      __syncwarp();
      // the store path is free again: the epilogue warps may drain the previous step's tiles (stacked path)
      if (stacked && s > 0) asm volatile("bar.arrive 8, %0;" ::"n"(32 + 32 * LF_EPI_WARPS) : "memory");
      if (lane == 0) {
#pragma unroll
        for (int i = 0; i < PKC; ++i, ++it) {
          const int st = it % PNS, ph = (it / PNS) & 1;
          mbar_wait(&empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&full[st], stage_bytes);
          bulk_load(ring + st * stage_bytes, src + (long long)i * (stage_bytes / 2), stage_bytes, &full[st]);
          if (TRACE && i < 10) LF_STAMP(20 + i);
        }
        LF_STAMP(2);
      }
      __syncwarp();
    }
  } else if (warp == W_REPORT) {
    // ---------------- progress for stream-ordered consumers (ss_stream_wait_value32), off the chain ----------------
    // `step_flag` = s says that every epilogue thread of this CTA has issued its `out` stores of steps < s.  Every 4th step
    // the reporter makes them visible at gpu scope and posts the period in its own slot; the reporter of CTA 0 of the
    // direction waits for all NC slots and writes the value the stream waits poll.
    if (lane == 0 && p.progress) {
      unsigned int* mine = p.done + dir * 64 + cta;
      for (int k = 1; 4 * k < T; ++k) {
        while (*reinterpret_cast<volatile int*>(step_flag) < 4 * k) __nanosleep(256);
        __threadfence_block();
        __threadfence();
        *reinterpret_cast<volatile unsigned int*>(mine) = (unsigned int)k;
        if (cta == 0) {
          for (int cc = 0; cc < p.NC; ++cc)
            while (ld_acquire(p.done + dir * 64 + cc) < (unsigned int)k) __nanosleep(256);
          __threadfence_system();
          *reinterpret_cast<volatile unsigned int*>(p.progress + dir) = p.prog_base + (unsigned int)(4 * k - 1);
        }
      }
    }
  } else {
    // ---------------- epilogue role ----------------
    // stacked: quadrant q = 2 g + plane: batch rows 32 g + lane, this thread finishes units (2 plane, 2 plane + 1) of its
    // column group; otherwise quadrant q owns rows 32 q + lane and all 4 units of the group.
    const int q = warp & 3, cg = warp >> 2;
    const int plane = stacked ? (q & 1) : 0;
    const int NU = stacked ? 2 : 4;
    const int b = (stacked ? (q >> 1) : q) * 32 + lane;
    const int ug = u0 + cg * 4 + 2 * plane;                    // first of this thread's NU units within the direction
    const bool row_ok = b < B;
    const int len = row_ok ? p.lens[b] : 0;
    float h[4], c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = (row_ok && j < NU) ? p.h0[((long long)dir * B + b) * H + ug + j] : 0.0f;
      c[j] = (row_ok && j < NU) ? p.c0[((long long)dir * B + b) * H + ug + j] : 0.0f;
    }
    const long long my_hi = dir * dir_elems + elem_off(0, b, ug), my_lo = dir * dir_elems + elem_off(1, b, ug);
    // every row of the R-row operand is published (rows b >= B as zeros): a consumer waits for all of them
    auto publish = [&](int slot, const float* hv) {
      __nv_bfloat16 hi[4], lo[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) split_bf16(hv[j], hi[j], lo[j]);
      __nv_bfloat16* base = p.hbuf + slot * slot_elems;
      if (stacked) {
        lf_st_strong32(base + my_hi, lf_pack(hi[0], hi[1]));
        lf_st_strong32(base + my_lo, lf_pack(lo[0], lo[1]));
      } else {
        lf_st_strong64(base + my_hi, lf_pack(hi[0], hi[1]), lf_pack(hi[2], hi[3]));
        lf_st_strong64(base + my_lo, lf_pack(lo[0], lo[1]), lf_pack(lo[2], lo[3]));
      }
    };
    // staged stores (stacked path): thread -> (operand row, 8-byte part) of the 128 x 24-byte tile this CTA publishes
    const int ul = cg * 4 + 2 * plane;                         // first of this thread's units within the CTA
    const int st_row = (int)threadIdx.x / 3, st_part = (int)threadIdx.x % 3;
    long long st_off;
    {
      const int u = u0 + st_part * 4, k = u & 63;
      st_off = dir * dir_elems + (long long)(u >> 6) * (stage_bytes / 2) + st_row * 64 + ((((k >> 3) ^ (st_row & 7)) << 3) | (k & 7));
    }
    if (threadIdx.x < 64) lens_s[threadIdx.x] = (int)threadIdx.x < B ? p.lens[threadIdx.x] : 0;
    // One release per CTA: twelve (one per warp) made 600 adds per step on one address, which queue up in L2 (6.25 instead
    // of 5.66 us per step).
    unsigned int* counter = p.done + 128 + dir;
    auto release_cta = [&]() {
      asm volatile("bar.sync 9, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
      if (threadIdx.x == 0) red_release_add(counter, 1u);     // cumulative over the CTA's stores ordered before the bar.sync
    };
    publish(0, h);
    release_cta();
    const uint32_t trow = tmem_d + ((uint32_t)(q * 32) << 16);
    for (int s = 0; s < T; ++s) {
      const int slot1 = (s + 1) & 1;                           // slot of the next step's input
      const int t = dir ? (T - 1 - s) : s;
      const bool active = row_ok && t < len;
      const long long row = (long long)t * B + (row_ok ? b : 0);
      float4* gptr = reinterpret_cast<float4*>(p.gx + row * (8LL * H) + (long long)dir * 4 * H + 4 * ug);
      float4 g4[4];
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j < NU) g4[j] = gptr[j];
      }
      // one warp polls the mbarrier; the others sleep on a hardware barrier
      if (warp == 0) mbar_wait(accum_full, s & 1);
      asm volatile("bar.sync 2, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
      if (threadIdx.x == 0) {
        LF_STAMP(5);
        if (p.progress && (s & 3) == 0) {                      // all epilogue threads are past their stores of step s-1
          __threadfence_block();
          *reinterpret_cast<volatile int*>(step_flag) = s;
        }
      }
      tc_fence_after();
      float d[16], e[16];
      tmem_ld16(trow + 16 * cg, d);
      tmem_ld16(trow + LF_N + 16 * cg, e);
      tc_fence_before();
#pragma unroll
      for (int i = 0; i < 16; ++i) d[i] += e[i];
      if (stacked) {
        // swap halves with the thread that holds the other plane of the same batch row: it needs the 8 columns this one
        // does not finish; afterwards d[0..7] are the complete pre-activations of this thread's two units
        float* xr = xbuf + (long long)b * LF_XPITCH + 16 * cg;
        float mine8[8], give8[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          mine8[i] = plane ? d[8 + i] : d[i];
          give8[i] = plane ? d[i] : d[8 + i];
        }
        float4* gp = reinterpret_cast<float4*>(xr + (plane ? 0 : 8));
        gp[0] = make_float4(give8[0], give8[1], give8[2], give8[3]);
        gp[1] = make_float4(give8[4], give8[5], give8[6], give8[7]);
        asm volatile("bar.sync 3, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
        const float4* kp = reinterpret_cast<const float4*>(xr + (plane ? 8 : 0));
        const float4 v0 = kp[0], v1 = kp[1];
        d[0] = mine8[0] + v0.x; d[1] = mine8[1] + v0.y; d[2] = mine8[2] + v0.z; d[3] = mine8[3] + v0.w;
        d[4] = mine8[4] + v1.x; d[5] = mine8[5] + v1.y; d[6] = mine8[6] + v1.z; d[7] = mine8[7] + v1.w;
      }
      if (threadIdx.x == 0) LF_STAMP(8);
      float hout[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j < NU) {
            const float gi = fsigmoid(d[4 * j + 0] + g4[j].x), gf = fsigmoid(d[4 * j + 1] + g4[j].y);
            const float gg = ftanh(d[4 * j + 2] + g4[j].z), go = fsigmoid(d[4 * j + 3] + g4[j].w);
            c[j] = gf * c[j] + gi * gg;
            h[j] = go * ftanh(c[j]);
            hout[j] = h[j];
            g4[j] = make_float4(gi, gf, gg, go);
          }
      }
      if (threadIdx.x == 0) LF_STAMP(6);
      if (stacked) {
        // Every global store of the step leaves the SM as whole sectors: a thread holds 2 units of ONE batch row, so direct
        // stores are 32 scattered 4..16-byte transactions per instruction, and the SM retires about one transaction per
        // cycle (3.5k per step: they queued in front of the next step's polls).  Stage in shared memory, store row-wise.
        {
          __nv_bfloat16 hi[2], lo[2];
          split_bf16(h[0], hi[0], lo[0]);                     // the CARRIED state (rows that have not started: h0)
          split_bf16(h[1], hi[1], lo[1]);
          hst[op_row(0, b) * 8 + ul / 2] = lf_pack(hi[0], hi[1]);
          hst[op_row(1, b) * 8 + ul / 2] = lf_pack(lo[0], lo[1]);
        }
        asm volatile("bar.sync 6, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
        {
          const uint2 v = reinterpret_cast<const uint2*>(hst)[st_row * 4 + st_part];
          lf_st_strong64(p.hbuf + slot1 * slot_elems + st_off, v.x, v.y);   // carried state for the next step: critical path first
        }
        release_cta();
        if (threadIdx.x == 0) LF_STAMP(7);
        {
          float* xr = xbuf + (long long)b * LF_XPITCH + 16 * cg + 8 * plane;   // the part of the swap buffer only this thread read
          reinterpret_cast<float4*>(xr)[0] = g4[0];
          reinterpret_cast<float4*>(xr)[1] = g4[1];
          *reinterpret_cast<float2*>(cst + b * 12 + ul) = make_float2(c[0], c[1]);
          __nv_bfloat16 hi[2], lo[2];
          split_bf16(hout[0], hi[0], lo[0]);
          split_bf16(hout[1], hi[1], lo[1]);
          ost[b * 8 + ul / 2] = lf_pack(hi[0], hi[1]);
          ost[(64 + b) * 8 + ul / 2] = lf_pack(lo[0], lo[1]);
        }
        asm volatile("bar.sync 7, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
        {
          // The tiles are drained only once the next step's gate has opened: until then the SM's store path belongs to the
          // publish above and the release behind it.
          if (s + 1 < T) asm volatile("bar.sync 8, %0;" ::"n"(32 + 32 * LF_EPI_WARPS) : "memory");
          // `out`: 128 (plane, row) x 24 bytes = one 8-byte store per thread
          {
            const int pl = st_row >> 6, bb = st_row & 63;
            if (bb < B) {
              const uint2 v = reinterpret_cast<const uint2*>(ost)[st_row * 4 + st_part];
              __stcs(reinterpret_cast<uint2*>(p.out + pl * p.out_plane + ((long long)t * B + bb) * (2LL * H) + dir * H + u0 + st_part * 4), v);
            }
          }
          // activated gates: 64 rows x 192 bytes; cell states: 64 rows x 48 bytes (rows past their length keep what they had)
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const int f = (int)threadIdx.x + 384 * j, rr = f / 12, c4 = f % 12;
            if (rr < B && t < lens_s[rr])
              __stcs(reinterpret_cast<float4*>(p.gx + ((long long)t * B + rr) * (8LL * H) + (long long)dir * 4 * H + 4 * u0) + c4,
                     reinterpret_cast<const float4*>(xbuf + rr * LF_XPITCH)[c4]);
          }
          if (p.cs && threadIdx.x < 192) {
            const int rr = (int)threadIdx.x / 3, c4 = (int)threadIdx.x % 3;
            if (rr < B && t < lens_s[rr])
              __stcs(reinterpret_cast<float4*>(p.cs + ((long long)t * B + rr) * (2LL * H) + dir * H + u0) + c4,
                     reinterpret_cast<const float4*>(cst + rr * 12)[c4]);
          }
        }
      } else {
        publish(slot1, h);                                     // carried state for the next step (critical path first)
        release_cta();
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
      if (threadIdx.x == 0) LF_STAMP(9);
    }
    if (row_ok) {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < NU) {
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
  // forward: exchange buffer [slot 2][dir 2][kc][2R rows][64] bf16 + counters / report slots; backward (lstm_bwd*.cu): the same size
  const long long fwd = 2LL * 2 * ssb::cdiv(H, 64) * 2 * R * 64 * 2 + 1024;
  const long long bwd = 2LL * 2 * ssb::cdiv(H, 64) * 2 * R * 64 * 2 + 1024;
  return fwd > bwd ? fwd : bwd;
}

extern "C" int ss_lstm_fwd(float* gx, const void* whh_r, long long whh_ld, long long whh_plane, const float* h0,
                           const float* c0, const int* lens, void* out, long long out_plane, float* cs, float* hn,
                           float* cn, void* ws, unsigned int* progress, unsigned int prog_base, int T, int B,
                           int batch_pitch, int H, void* stream) {
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
  p.done = (unsigned int*)((char*)ws + hbuf_bytes);
  cudaMemsetAsync(ws, 0, hbuf_bytes + 1024, st);            // the tail of the last K chunk is never published and must be finite

  p.Bp = batch_pitch > 0 ? batch_pitch : B;
  SS_REQUIRE(p.Bp >= B, "ss_lstm_fwd: batch_pitch %d < B %d", batch_pitch, B);
  if (H == 600 && p.R == 64 && whh_ld % 8 == 0 && whh_plane % 8 == 0)
    return launch_lstm_fwd_tw(p, whh_r, whh_ld, whh_plane, st);
  SS_REQUIRE(p.Bp == B, "ss_lstm_fwd: a sub-batch view (batch_pitch %d != B %d) needs H = 600 and B <= 64", batch_pitch, B);

  CUtensorMap tmW;
  {
    uint64_t dims[3] = {(uint64_t)H, (uint64_t)8 * H, 2};
    uint64_t str[2] = {(uint64_t)whh_ld * 2, (uint64_t)whh_plane * 2};
    uint32_t box[3] = {64, LF_N, 2};
    int rc = make_tmap_bf16(&tmW, whh_r, 3, dims, str, box);
    if (rc) return rc;
  }
  const int smem = p.NS * 2 * p.R * 128 + p.KC * LF_WCHUNK + 64 * LF_XPITCH * 4 + 1024 + 256 + LF_STAGE_TILES;
  void* kern = p.dbg ? (void*)lstm_fwd_kernel<true> : (void*)lstm_fwd_kernel<false>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&tmW, (void*)&p};
  cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(2 * p.NC), dim3(LF_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_fwd: cooperative launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}
