// K2b, reference shape (H = 600, B <= 64): forward LSTM recurrence with W_hh RESIDENT IN TENSOR MEMORY.
// Same contract as lstm_fwd.cu (which stays as the generic path: other hidden sizes, 64 < B <= 128).
//
// lstm_fwd.cu keeps its 115 KB slice of W_hh in shared memory, which leaves room for five 16 KB stages of h_prev: a
// bulk copy lands ~1300 cycles after it is issued, so the five-deep ring -- not the copy engine -- paces the ten
// chunks of a step (370-420 cycles per chunk against a 281-cycle floor), and the tensor core fetches 28 KB of operands
// per chunk through the shared-memory pipe (profiles/r02_summary.md section J).  Here the roles of the operands are
// swapped:
//
//   D^T[c][n] = sum_k W[c][k] * h[n][k],   c = 8 * unit + 2 * gate + plane (96 of 128 TMEM lanes),
//                                          n = 64 * plane_h + batch row (N = 128)
//
// A = the W slice, split-bf16, written ONCE into tensor memory (128 lanes x 304 columns = 38 k-steps of 16), B = the
// h_prev chunk exactly as the exchange buffer holds it.  Shared memory then has room for all ten chunks of a step: the
// ten copies are issued back to back when the gate opens, and the tensor core reads 16 KB per chunk.  The accumulator
// comes back transposed (a thread owns one W row and 32 of the 128 columns), so the hi / lo planes are combined by a
// shuffle between neighbouring lanes and the 4 gates of a unit are gathered by two butterfly exchanges inside groups of
// 8 lanes -- no shared memory, no CTA barrier before the gate math.  The exchange protocol is lstm_fwd.cu's.
#include "../../include/ssb200.h"
#include "lstm_fwd.cuh"

namespace ssb {

constexpr int TW_THREADS = 512;             // 16 warps: warp % 4 < 3 epilogue (TMEM lane quadrants 0..2), warp % 4 == 3 roles
constexpr int TW_EPI = 384;
constexpr int TW_KC = 10, TW_KSTEPS = 38;   // 64-wide chunks / 16-wide k-steps of K = 600
constexpr int TW_STAGE = 16384;             // one chunk at RB = 64: 128 operand rows x 128 B (RB = 16 / 32: a quarter / half)
constexpr int TW_CPC = 2;                   // chunks per bulk copy: a copy costs the engine ~280 cycles whether it moves 16 or 32 KB
constexpr int TW_NCOPY = TW_KC / TW_CPC;
constexpr int TW_TMEM = 512;
constexpr int TW_DCOL = 320;                // accumulator columns [320, 448); W in [0, 304)
constexpr int TW_SMEM = TW_KC * TW_STAGE + 64 * LF_XPITCH * 4 + 256 + LF_STAGE_TILES + 1024;
static_assert(TW_SMEM <= 232448, "shared memory budget");

#define TW_STAMP(slot)                                                         \
  do {                                                                         \
    if (TRACE && p.dbg && blockIdx.x == 0) p.dbg[s * 32 + (slot)] = clock64(); \
  } while (0)

__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, uint4 a, uint4 b) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(a.x), "r"(a.y), "r"(a.z),
               "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}
__device__ __forceinline__ void tw_st_strong64(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.relaxed.gpu.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}

struct LstmFwdTwParams {
  LstmFwdParams p;
  const __nv_bfloat16* whh;   // sb16 [8H][H] rows dir*4H + 4u + gate, hi plane; lo plane at + whh_plane
  long long whh_ld, whh_plane;
};

// RB = padded batch rows of the exchanged operand (16, 32 or 64 >= B): a batch of 8 twenty-second utterances (configs[4])
// moves 40 KB instead of 160 KB of h per step and multiplies N = 32 instead of 128 columns.
template <int RB, bool TRACE>
__global__ void __launch_bounds__(TW_THREADS, 1) lstm_fwd_tw_kernel(LstmFwdTwParams pp) {
  constexpr int STAGE = 2 * RB * 128;                          // one K chunk of the operand: rows [h_hi 0..RB) | h_lo 0..RB)]
  const LstmFwdParams& p = pp.p;
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int H = 600;
  unsigned char* ring = smem;                                  // stage i = K chunk i of the step
  float* xbuf = reinterpret_cast<float*>(ring + TW_KC * STAGE);         // [64][LF_XPITCH] activated gates of the step
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + 64 * LF_XPITCH);     // [10] chunk landed
  uint64_t* empty = full + 10;                                 // [10] chunk consumed (MMA commit)
  uint64_t* accum_full = empty + 10;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_full + 1);
  int* step_flag = reinterpret_cast<int*>(tmem_slot + 1);
  uint16_t* hst = reinterpret_cast<uint16_t*>(reinterpret_cast<unsigned char*>(full) + 256);   // [128 operand rows][16] published h
  uint16_t* ost = hst + 128 * 16;                              // [plane * 64 + row][16] `out`
  float* cst = reinterpret_cast<float*>(ost + 128 * 16);       // [64][12] cell states
  int* lens_s = reinterpret_cast<int*>(cst + 64 * 12);         // [64]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int u0 = cta * LF_U;
  const int T = p.T, B = p.B, Bp = p.Bp;                      // Bp: rows per frame of the padded batch (>= B)
  const long long slot_elems = 2LL * TW_KC * (STAGE / 2), dir_elems = (long long)TW_KC * (STAGE / 2);
  unsigned int* counter = p.done + 128 + dir;

  if (threadIdx.x == 0) {
    *step_flag = 0;
    for (int i = 0; i < TW_KC; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(accum_full, 1);
    fence_barrier_init();
  }
  if (threadIdx.x < 64) lens_s[threadIdx.x] = (int)threadIdx.x < B ? p.lens[threadIdx.x] : 0;
  constexpr int W_COPY = 3, W_MMA = 7, W_REPORT = 11;
  if (warp == W_MMA) tmem_alloc<TW_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  // ---- W_hh slice -> tensor memory (once): TMEM lane m = 8 * unit + 2 * gate + plane, 8 columns per 16-wide k-step ----
  if ((warp & 3) < 3 && (warp >> 2) == 0) {
    const int m = 32 * (warp & 3) + lane;
    const int row = dir * 4 * H + 4 * (u0 + (m >> 3)) + ((m >> 1) & 3);
    const __nv_bfloat16* wr = pp.whh + (long long)(m & 1) * pp.whh_plane + (long long)row * pp.whh_ld;
    const uint32_t tl = tmem + ((uint32_t)(32 * (warp & 3)) << 16);
#pragma unroll 2
    for (int ks = 0; ks < TW_KSTEPS; ++ks) {
      const uint4 a = __ldg(reinterpret_cast<const uint4*>(wr + 16 * ks));
      const uint4 b = 16 * ks + 16 <= H ? __ldg(reinterpret_cast<const uint4*>(wr + 16 * ks + 8)) : make_uint4(0, 0, 0, 0);
      tmem_st8(tl + 8 * ks, a, b);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == W_MMA) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 2 * RB);
      const uint64_t b_base = umma_desc_sw128(smem_u32(ring));
      for (int s = 0; s < T; ++s) {
#pragma unroll
        for (int i = 0; i < TW_KC; ++i) {
          if (i % TW_CPC == 0) {
            mbar_wait(&full[i / TW_CPC], s & 1);
            TW_STAMP(i == 0 ? 3 : 10 + i);
            tc_fence_after();
          }
          const uint64_t db = b_base + (uint32_t)i * (STAGE >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (4 * i + k < TW_KSTEPS) umma_bf16_ts(tmem + TW_DCOL, tmem + 8 * (4 * i + k), db + 2 * k, idesc, (i | k) ? 1u : 0u);
          if (i % TW_CPC == TW_CPC - 1) umma_commit(&empty[i / TW_CPC]);
        }
        umma_commit(accum_full);
        TW_STAMP(4);
      }
    }
  } else if (warp == W_COPY) {
    // ---- gate + copies: wait for the step's releases (one per CTA of the direction), cross to the async proxy, stream
    // the ten chunks back to back (every chunk has a stage of its own; two chunks per copy)
    for (int s = 0; s < T; ++s) {
      const __nv_bfloat16* src = p.hbuf + (s & 1) * slot_elems + dir * dir_elems;
      if (lane == 0) {
        TW_STAMP(0);
        const unsigned int need = (unsigned int)p.NC * (unsigned int)(s + 1);
        while (ld_acquire(counter) < need) {
        }
        fence_proxy_async();                                   // the producers' generic-proxy stores -> this thread's bulk copies
        TW_STAMP(1);
      }
      __syncwarp();
      // the store path is free again: the epilogue warps may drain the previous step's tiles
      if (s > 0) asm volatile("bar.arrive 8, %0;" ::"n"(32 + TW_EPI) : "memory");
      if (lane == 0) {
        // (Multicast to a 2-CTA cluster, each CTA issuing five of the ten copies, was measured: the chunks still land every
        // ~365 cycles -- the SM's ingest, not L2, sets the rate -- and the first one later: 5.80 instead of 5.35 us per step.)
#pragma unroll
        for (int i = 0; i < TW_NCOPY; ++i) {
          mbar_wait(&empty[i], (s & 1) ^ 1);
          mbar_arrive_expect_tx(&full[i], TW_CPC * STAGE);
          bulk_load(ring + i * TW_CPC * STAGE, src + (long long)i * TW_CPC * (STAGE / 2), TW_CPC * STAGE, &full[i]);
          if (TRACE) TW_STAMP(20 + i);
        }
        TW_STAMP(2);
      }
      __syncwarp();
    }
  } else if (warp == W_REPORT) {
    // ---- progress for stream-ordered consumers (ss_stream_wait_value32), off the chain: see lstm_fwd.cu ----
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
  } else if ((warp & 3) < 3) {
    // ---------------- epilogue: quadrant q holds W rows of units 4q..4q+3, column group cgp = batch rows 16 cgp .. +15 ----------------
    const int q = warp & 3, cgp = warp >> 2;
    const int eidx = (cgp * 3 + q) * 32 + lane;                // 0..383
    const int w = lane & 7, g1 = (w >> 2) & 1, g0 = (w >> 1) & 1, pl = w & 1;
    const int ul = 4 * q + (lane >> 3), u = u0 + ul;           // this thread's unit
    const int b0 = 16 * cgp + 8 * pl + 4 * g1 + 2 * g0;        // its two batch rows: b0, b0 + 1
    const bool ok0 = b0 < B, ok1 = b0 + 1 < B;
    const int len0 = lens_s[b0], len1 = lens_s[b0 + 1];
    float h[2], c[2];
    h[0] = ok0 ? p.h0[((long long)dir * Bp + b0) * H + u] : 0.0f;
    h[1] = ok1 ? p.h0[((long long)dir * Bp + b0 + 1) * H + u] : 0.0f;
    c[0] = ok0 ? p.c0[((long long)dir * Bp + b0) * H + u] : 0.0f;
    c[1] = ok1 ? p.c0[((long long)dir * Bp + b0 + 1) * H + u] : 0.0f;
    // coalesced stores: epilogue thread -> (operand row, 8-byte part) of the 128 x 24-byte tile this CTA publishes
    const int st_row = eidx / 3, st_part = eidx % 3;            // (tile row plane * 64 + b, 8-byte part) of the staged tiles
    const int pub_row = (st_row >> 6) * RB + (st_row & 63);    // its operand row; published if b < RB
    const bool pub = (st_row & 63) < RB;
    long long st_off;
    {
      const int uu = u0 + st_part * 4, k = uu & 63;
      st_off = dir * dir_elems + (long long)(uu >> 6) * (STAGE / 2) + pub_row * 64 + ((((k >> 3) ^ (pub_row & 7)) << 3) | (k & 7));
    }
    auto stage_h = [&](uint16_t* tile, const float* hv) {      // rows: hi plane 0..63, lo plane 64..127
      __nv_bfloat16 hi, lo;
      split_bf16(hv[0], hi, lo);
      tile[b0 * 16 + ul] = __bfloat16_as_ushort(hi);
      tile[(64 + b0) * 16 + ul] = __bfloat16_as_ushort(lo);
      split_bf16(hv[1], hi, lo);
      tile[(b0 + 1) * 16 + ul] = __bfloat16_as_ushort(hi);
      tile[(64 + b0 + 1) * 16 + ul] = __bfloat16_as_ushort(lo);
    };
    auto publish = [&](int slot) {                             // hst -> exchange buffer, one release per CTA
      asm volatile("bar.sync 6, %0;" ::"n"(TW_EPI) : "memory");
      if (pub) {
        const uint2 v = reinterpret_cast<const uint2*>(hst)[st_row * 4 + st_part];
        tw_st_strong64(p.hbuf + slot * slot_elems + st_off, v.x, v.y);
      }
      asm volatile("bar.sync 9, %0;" ::"n"(TW_EPI) : "memory");
      if (eidx == 0) red_release_add(counter, 1u);             // cumulative over the CTA's stores ordered before the bar.sync
    };
    stage_h(hst, h);
    publish(0);
    const uint32_t trow = tmem + ((uint32_t)(32 * q) << 16) + TW_DCOL;
    for (int s = 0; s < T; ++s) {
      const int t = dir ? (T - 1 - s) : s;
      const bool act0 = ok0 && t < len0, act1 = ok1 && t < len1;
      float4 gx0 = make_float4(0.f, 0.f, 0.f, 0.f), gx1 = gx0;
      if (act0) gx0 = *reinterpret_cast<const float4*>(p.gx + ((long long)t * Bp + b0) * (8LL * H) + (long long)dir * 4 * H + 4 * u);
      if (act1) gx1 = *reinterpret_cast<const float4*>(p.gx + ((long long)t * Bp + b0 + 1) * (8LL * H) + (long long)dir * 4 * H + 4 * u);
      // one warp polls the mbarrier; the others sleep on a hardware barrier
      if (warp == 0) mbar_wait(accum_full, s & 1);
      asm volatile("bar.sync 2, %0;" ::"n"(TW_EPI) : "memory");
      if (eidx == 0) {
        TW_STAMP(5);
        if (p.progress && (s & 3) == 0) {                      // all epilogue threads are past their stores of step s-1
          __threadfence_block();
          *reinterpret_cast<volatile int*>(step_flag) = s;
        }
      }
      tc_fence_after();
      float d[16], e[16];
      if (16 * cgp < RB) {                                     // (warp-uniform) column groups past the padded batch have no columns
        tmem_ld16(trow + 16 * cgp, d);                         // W row x h_hi of batch rows 16 cgp ..
        tmem_ld16(trow + RB + 16 * cgp, e);                    // W row x h_lo
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) d[i] = e[i] = 0.0f;
      }
      tc_fence_before();
      // hi / lo planes of W are neighbouring lanes: W_hi.(h_hi + h_lo) + W_lo.h_hi
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float v = pl ? d[i] : d[i] + e[i];
        d[i] = v + __shfl_xor_sync(0xffffffffu, v, 1);
      }
      // the 8 lanes of a unit hold gate (g1 g0) for batch index i = (i3 i2 i1 i0); lane (g1 g0 pl) finishes the rows with
      // i3 = pl, i2 = g1, i1 = g0: take half by plane, then two butterfly exchanges move gate bits against row bits
      float t8[8];
#pragma unroll
      for (int m = 0; m < 8; ++m) t8[m] = pl ? d[8 + m] : d[m];
      float x[2][2][2];                                        // [gate bit 0][i2][i0]   (gate bit 1 = g1, i1 = g0)
#pragma unroll
      for (int i2 = 0; i2 < 2; ++i2)
#pragma unroll
        for (int i0 = 0; i0 < 2; ++i0) {
          const float keep = g0 ? t8[4 * i2 + 2 + i0] : t8[4 * i2 + i0];
          const float send = g0 ? t8[4 * i2 + i0] : t8[4 * i2 + 2 + i0];
          const float recv = __shfl_xor_sync(0xffffffffu, send, 2);
          x[0][i2][i0] = g0 ? recv : keep;
          x[1][i2][i0] = g0 ? keep : recv;
        }
      float y[2][2][2];                                        // [gate bit 1][gate bit 0][i0]   (i2 = g1, i1 = g0)
#pragma unroll
      for (int gb0 = 0; gb0 < 2; ++gb0)
#pragma unroll
        for (int i0 = 0; i0 < 2; ++i0) {
          const float keep = g1 ? x[gb0][1][i0] : x[gb0][0][i0];
          const float send = g1 ? x[gb0][0][i0] : x[gb0][1][i0];
          const float recv = __shfl_xor_sync(0xffffffffu, send, 4);
          y[0][gb0][i0] = g1 ? recv : keep;
          y[1][gb0][i0] = g1 ? keep : recv;
        }
      if (eidx == 0) TW_STAMP(8);
      float hout[2] = {0.0f, 0.0f};
      float4 ga[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const float4 gx = j ? gx1 : gx0;
        if (j ? act1 : act0) {
          const float gi = fsigmoid(y[0][0][j] + gx.x), gf = fsigmoid(y[0][1][j] + gx.y);
          const float gg = ftanh(y[1][0][j] + gx.z), go = fsigmoid(y[1][1][j] + gx.w);
          c[j] = gf * c[j] + gi * gg;
          h[j] = go * ftanh(c[j]);
          hout[j] = h[j];
          ga[j] = make_float4(gi, gf, gg, go);
        }
      }
      if (eidx == 0) TW_STAMP(6);
      stage_h(hst, h);                                         // the CARRIED state (rows that have not started: h0)
      publish((s + 1) & 1);
      if (eidx == 0) TW_STAMP(7);
      // the rest of the step's results: staged now, stored row-wise once the next step's gate has opened
      if (p.cs) {                                              // (evaluation keeps neither the gates nor the cell states)
        *reinterpret_cast<float4*>(xbuf + b0 * LF_XPITCH + 4 * ul) = ga[0];
        *reinterpret_cast<float4*>(xbuf + (b0 + 1) * LF_XPITCH + 4 * ul) = ga[1];
        cst[b0 * 12 + ul] = c[0];
        cst[(b0 + 1) * 12 + ul] = c[1];
      }
      stage_h(ost, hout);
      asm volatile("bar.sync 7, %0;" ::"n"(TW_EPI) : "memory");
      if (s + 1 < T) asm volatile("bar.sync 8, %0;" ::"n"(32 + TW_EPI) : "memory");
      {
        const int plo = st_row >> 6, bb = st_row & 63;
        if (bb < B) {
          const uint2 v = reinterpret_cast<const uint2*>(ost)[st_row * 4 + st_part];
          __stcs(reinterpret_cast<uint2*>(p.out + plo * p.out_plane + ((long long)t * Bp + bb) * (2LL * H) + dir * H + u0 + st_part * 4), v);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int f = eidx + TW_EPI * j, rr = f / 12, c4 = f % 12;
          if (p.cs && rr < B && t < lens_s[rr])
            __stcs(reinterpret_cast<float4*>(p.gx + ((long long)t * Bp + rr) * (8LL * H) + (long long)dir * 4 * H + 4 * u0) + c4,
                   reinterpret_cast<const float4*>(xbuf + rr * LF_XPITCH)[c4]);
        }
        if (p.cs && eidx < 192) {
          const int rr = eidx / 3, c4 = eidx % 3;
          if (rr < B && t < lens_s[rr])
            __stcs(reinterpret_cast<float4*>(p.cs + ((long long)t * Bp + rr) * (2LL * H) + dir * H + u0) + c4,
                   reinterpret_cast<const float4*>(cst + rr * 12)[c4]);
        }
      }
      if (eidx == 0) TW_STAMP(9);
    }
    if (ok0) {
      p.hn[((long long)dir * Bp + b0) * H + u] = h[0];
      p.cn[((long long)dir * Bp + b0) * H + u] = c[0];
    }
    if (ok1) {
      p.hn[((long long)dir * Bp + b0 + 1) * H + u] = h[1];
      p.cn[((long long)dir * Bp + b0 + 1) * H + u] = c[1];
    }
  }
  __syncthreads();
  if (warp == W_MMA) {
    tc_fence_after();
    tmem_free<TW_TMEM>(tmem);
  }
}

int tw_rows(int B) { return B <= 16 ? 16 : B <= 32 ? 32 : 64; }

// called by ss_lstm_fwd (lstm_fwd.cu) for H = 600, B <= 64
int launch_lstm_fwd_tw(const LstmFwdParams& p, const void* whh_r, long long whh_ld, long long whh_plane, cudaStream_t st) {
  LstmFwdTwParams pp;
  pp.p = p;
  pp.whh = (const __nv_bfloat16*)whh_r;
  pp.whh_ld = whh_ld;
  pp.whh_plane = whh_plane;
  const int rb = tw_rows(p.B);
  void* kern;
  if (rb == 16) kern = p.dbg ? (void*)lstm_fwd_tw_kernel<16, true> : (void*)lstm_fwd_tw_kernel<16, false>;
  else if (rb == 32) kern = p.dbg ? (void*)lstm_fwd_tw_kernel<32, true> : (void*)lstm_fwd_tw_kernel<32, false>;
  else kern = p.dbg ? (void*)lstm_fwd_tw_kernel<64, true> : (void*)lstm_fwd_tw_kernel<64, false>;
  const int smem = TW_KC * 2 * rb * 128 + 64 * LF_XPITCH * 4 + 256 + LF_STAGE_TILES + 1024;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&pp};
  cudaError_t e = cudaLaunchCooperativeKernel(kern, dim3(2 * p.NC), dim3(TW_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_fwd: cooperative launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}

}  // namespace ssb
