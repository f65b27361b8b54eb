// K2c (fast path): backward LSTM recurrence with the K dimension split over 4-CTA clusters.
// Same contract as lstm_bwd.cu (which stays as the generic path); used for the reference's shape H = 600, B <= 64.
//
//   dh_rec[b][j] = sum_{n < 4H} dG_prev[b][n] * W_hh[n][j]            (n = 4u+gate, interleaved)
//
// has K = 4H = 2400: an output-stationary CTA must stream all of dG_prev (614 KB split-bf16) through shared memory
// every step -> >= 10k cycles per step.  Here a cluster of 4 CTAs owns 64 hidden units; CTA r streams only K-quarter r
// of dG_prev (154 KB) and multiplies it with its [K-quarter x 64 units] slice of W_hh^T, which is RESIDENT IN TENSOR
// MEMORY: A = [W^T rows: lane 2 * unit + plane] written once with tcgen05.st (up to 320 columns), B = the dG chunk as
// the TMA box lands it (rows 2 * b + plane), D^T = 128 x 128 in TMEM columns 320..447.  Shared memory then holds all
// ten chunks of a step (the 160 KB W slice used to leave room for a 3-deep ring: 480 cycles per chunk), and the
// copies are issued back to back.  The four partial products are reduced over distributed shared memory: a thread
// owns one W^T row (unit, plane) and 32 columns, adds the planes (a shuffle between neighbouring lanes, an add between
// neighbouring columns) and sends the peer that finalises its unit eight consecutive batch rows (st.async, completion
// counted on the peer's mbarrier).  Each CTA then runs the cell backward for its 16 units, writes dG[t] and publishes
// it (gpu-scope release counter).
#include <stdlib.h>

#include "../../include/ssb200.h"
#include "common.cuh"
#include "lstm_bwd.cuh"

namespace ssb {

constexpr int KS_CL = 4;                       // CTAs per cluster
constexpr int KS_UC = 64;                      // hidden units per cluster
constexpr int KS_U = KS_UC / KS_CL;            // units finalised per CTA
constexpr int KS_H = 600;
constexpr int KS_NCL = (KS_H + KS_UC - 1) / KS_UC;      // clusters per direction (10)
constexpr int KS_NCTA = KS_NCL * KS_CL;                 // CTAs per direction (40)
constexpr int KS_KCH = (4 * KS_H + 63) / 64;            // 64-wide K chunks in total (38; the last holds 32 columns)
constexpr int KS_MAXCH = 10;                            // chunks per CTA: 10, 10, 9, 9
constexpr int KS_NS = KS_MAXCH;                         // dG stages: one per chunk of the step
constexpr int KS_CHUNK = 16384;                         // one chunk of either operand: 128 rows x 128 B
constexpr int KS_EPI_WARPS = 16;
constexpr int KS_THREADS = 64 + 32 * KS_EPI_WARPS;
constexpr int KS_RP = 68;                               // floats per unit row of the reduction buffer (64 batch rows + bank skew)
constexpr int KS_RBUF = KS_CL * KS_U * KS_RP * 4;       // [src][16 units][KS_RP] floats
constexpr int KS_DGT = 64 * 2 * 128;                    // staging tile of this CTA's dG block: [row 2 b + plane][64 columns] bf16
constexpr int KS_SMEM = KS_NS * KS_CHUNK + KS_DGT + KS_RBUF + 1024 + 256;
constexpr int KS_TMEM = 512;                            // W^T slice in columns [0, 320), accumulator in [320, 448)
constexpr int KS_DCOL = 320;

__device__ __forceinline__ int ks_chunk_begin(int r) { return r < 2 ? 10 * r : 20 + 9 * (r - 2); }

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// 16-byte store into a peer CTA's shared memory; the bytes are counted on that CTA's mbarrier when they land.
__device__ __forceinline__ void st_async_v4(uint32_t raddr, float a, float b, float c, float d, uint32_t rbar) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(raddr),
               "f"(a), "f"(b), "f"(c), "f"(d), "r"(rbar)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait_cluster(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}

__device__ __forceinline__ void ks_umma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accum)
      : "memory");
}
__device__ __forceinline__ void ks_tmem_st8(uint32_t taddr, uint4 a, uint4 b) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(a.x), "r"(a.y), "r"(a.z),
               "r"(a.w), "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w)
               : "memory");
}

// tile in shared memory -> a box of the dG tensor (async proxy), as one bulk group
__device__ __forceinline__ void ks_tma_store_3d(const CUtensorMap* m, const void* src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"((uint64_t)m),
               "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void ks_tma_store_wait() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// all-CTA skew probe (DBG only): globaltimer of 4 events for steps 100..115 behind the per-step rows of CTA 0
#define KS_GSTAMP(ev)                                                                                              \
  do {                                                                                                             \
    if (DBG && s >= 100 && s < 116) {                                                                              \
      unsigned long long gt_;                                                                                      \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt_));                                                      \
      p.dbg[(long long)(p.T + 2) * 32 + ((long long)blockIdx.x * 16 + (s - 100)) * 4 + (ev)] = (long long)gt_;    \
    }                                                                                                              \
  } while (0)

#define KS_STAMP(slot)                                                 \
  do {                                                                 \
    if (DBG && blockIdx.x == 0) p.dbg[s * 32 + (slot)] = clock64();    \
  } while (0)

// RB = batch rows of the dG boxes (16, 32 or 64 >= B): small batches of long utterances (configs[4]) move a quarter of
// the operand per step and multiply N = 32 instead of 128 columns.
template <int RB, bool DBG>
__global__ void __launch_bounds__(KS_THREADS, 1)
lstm_bwd_ks_kernel(const __grid_constant__ CUtensorMap tmG, const __grid_constant__ CUtensorMap tmS64,
                   const __grid_constant__ CUtensorMap tmS32, LstmBwdParams p, const __nv_bfloat16* __restrict__ whhT,
                   long long whhT_ld, long long whhT_plane) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ring = smem;                                  // KS_NS chunks of dG_prev (stage i = chunk i of the step), rows = 2*b + plane
  constexpr int STAGE = 2 * RB * 128;                          // one dG box: rows 2 * b + plane, 64 columns
  unsigned char* dgt = ring + KS_NS * KS_CHUNK;                // this step's dG block, staged for ONE tensor store
  float* rbuf = reinterpret_cast<float*>(dgt + KS_DGT);        // [src][unit 16][KS_RP]
  uint64_t* full = reinterpret_cast<uint64_t*>(reinterpret_cast<unsigned char*>(rbuf) + KS_RBUF);
  uint64_t* empty = full + KS_NS;
  uint64_t* accum_full = empty + KS_NS;
  uint64_t* wfull = accum_full + 1;
  uint64_t* rfull = wfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(rfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = (int)cluster_rank();
  const int cl = blockIdx.x / KS_CL;
  const int dir = cl / KS_NCL, c = cl % KS_NCL;
  const int T = p.T, B = p.B, Bp = p.Bp;                       // Bp: rows per frame of the padded batch (>= B)
  constexpr int H = KS_H;
  const int nsteps = T + (p.want_d0 ? 1 : 0);
  const int ch0 = ks_chunk_begin(rank), nch = rank < 2 ? 10 : 9;
  unsigned int* counter = p.sync + dir;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmG);
    tma_prefetch_desc(&tmS64);
    tma_prefetch_desc(&tmS32);
    for (int s = 0; s < KS_NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(accum_full, 1);
    mbar_init(wfull, 1);
    mbar_init(rfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<KS_TMEM>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();                                          // peers' mbarriers exist before anyone sends
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  // ---- W_hh^T slice -> tensor memory (once): lane m = 2 * unit + plane of the cluster's 64 units, 8 columns per 16-wide
  // k-step of this CTA's K range; units >= H (last cluster) and the k-steps past 4H (last chunk) are never multiplied / zero
  if (warp >= 2 && warp < 6) {
    const int m = 32 * (warp & 3) + lane, unit = c * KS_UC + (m >> 1);
    const __nv_bfloat16* wr = whhT + (long long)(m & 1) * whhT_plane + (long long)(dir * H + unit) * whhT_ld + (long long)ch0 * 64;
    const uint32_t tl = tmem_d + ((uint32_t)(32 * (warp & 3)) << 16);
    const uint4 z4 = make_uint4(0, 0, 0, 0);
    for (int ks = 0; ks < 4 * nch; ++ks) {
      const bool ok = unit < H && ch0 * 64 + 16 * ks + 16 <= 4 * H;
      const uint4 a = ok ? __ldg(reinterpret_cast<const uint4*>(wr + 16 * ks)) : z4;
      const uint4 b = ok ? __ldg(reinterpret_cast<const uint4*>(wr + 16 * ks + 8)) : z4;
      ks_tmem_st8(tl + 8 * ks, a, b);
    }
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  if (warp == 0) {
    if (lane == 0) {
      for (int s = 1; s < nsteps; ++s) {
        const int tp = dir ? (s - 1) : (T - s);                // frame whose dG this step consumes
        const unsigned int need = (unsigned int)KS_NCTA * (unsigned int)s;
        KS_STAMP(0);
        while (ld_acquire(counter) < need) {
        }
        KS_STAMP(1);                                           // (proxy fence: writer side, before the CTA barrier)
        // every chunk has a stage of its own: the copies go out back to back (the stage was consumed a step ago)
#pragma unroll
        for (int i = 0; i < KS_MAXCH; ++i) {
          if (i < nch) {
            mbar_wait(&empty[i], (s & 1));                     // = ((s - 1) & 1) ^ 1
            mbar_arrive_expect_tx(&full[i], STAGE);
            tma_load_3d(ring + i * STAGE, &tmG, &full[i], dir * 4 * H + (ch0 + i) * 64, 0, tp * Bp);
          }
        }
        // progress for stream-ordered consumers (ss_stream_wait_value32): the counter just observed covers the dG rows of
        // steps 0..s-1 of this direction; one CTA per direction reports every 4th step, after its copies are in flight
        if (p.progress && c == 0 && rank == 0 && (s & 3) == 0) {
          __threadfence_system();
          *reinterpret_cast<volatile unsigned int*>(p.progress + dir) = p.prog_base + (unsigned int)s;
        }
        KS_STAMP(2);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, 2 * RB);
      const uint64_t b_base = umma_desc_sw128(smem_u32(ring));
      for (int s = 1; s < nsteps; ++s) {
#pragma unroll
        for (int i = 0; i < KS_MAXCH; ++i) {
          if (i < nch) {
            mbar_wait(&full[i], (s - 1) & 1);
            KS_STAMP(10 + i);
            tc_fence_after();
            const uint64_t db = b_base + (uint32_t)(i * (STAGE >> 4));
            const bool last = ch0 + i == KS_KCH - 1;           // 32 valid columns: 2 k-steps
#pragma unroll
            for (int k = 0; k < 4; ++k)
              if (k < 2 || !last) ks_umma_ts(tmem_d + KS_DCOL, tmem_d + 8 * (4 * i + k), db + 2 * k, idesc, (i | k) ? 1u : 0u);
            umma_commit(&empty[i]);
          }
        }
        umma_commit(accum_full);
        KS_STAMP(3);
      }
    }
  } else {
    const int ew = warp - 2;
    const int q = warp & 3, cgp = ew >> 2;                     // TMEM lane quadrant = peer that finalises its 16 units; batch rows 16 cgp .. +15
    const int pw = lane & 1;                                   // sender role: plane of this lane's W^T row (unit 16 q + lane / 2)
    // final role: lanes run along the 8 unit pairs of a batch row (4 rows per warp), so a warp's loads and stores cover
    // whole sectors: 64 / 128 contiguous bytes per row instead of 32 scattered 8 / 16-byte transactions
    const int fb = 4 * ew + (lane >> 3), u2 = lane & 7;        // batch row, unit pair
    const int ul = 2 * u2;                                     // unit within this CTA's 16
    const int ug = c * KS_UC + rank * KS_U + ul;               // unit within the direction
    const bool row_ok = fb < B && ug < H;
    const int len = row_ok ? p.lens[fb] : 0;
    const uint32_t trow = tmem_d + ((uint32_t)(q * 32) << 16) + KS_DCOL;
    const uint32_t rbar = map_to_rank(smem_u32(rfull), (uint32_t)q);
    const uint32_t rdst = map_to_rank(smem_u32(rbuf), (uint32_t)q) + (uint32_t)(((rank * KS_U + (lane >> 1)) * KS_RP + 16 * cgp + 8 * pw) * 4);
    const int nvalid = max(0, min(KS_U, H - (c * KS_UC + rank * KS_U)));   // units of this CTA that exist: 16, 8 (last cluster, rank 1) or 0
    const uint32_t tile_pitch = nvalid == KS_U ? 128u : 64u;                // bytes per (row, plane) of the staged block
    const uint32_t dgt_u32 = smem_u32(dgt);
    float dc[2] = {0.0f, 0.0f};
    float dh0[2] = {0.0f, 0.0f};
    // running pointers of this thread's (row, unit pair): one add per step instead of 64-bit multiplies
    const long long rstep = dir ? (long long)Bp : -(long long)Bp;         // rows per step in processing order
    const long long row0 = (long long)(dir ? 0 : T - 1) * Bp + (row_ok ? fb : 0);
    const float* gptr = p.gates + row0 * (8LL * H) + (long long)dir * 4 * H + 4 * ug;
    const float* cptr = p.cs + row0 * (2LL * H) + dir * H + ug;
    const float* dptr = p.dout + row0 * (2LL * H) + dir * H + ug;
    const float* c0ptr = p.c0 + ((long long)dir * Bp + (row_ok ? fb : 0)) * H + ug;
    for (int s = 0; s < nsteps; ++s) {
      const int t = dir ? s : (T - 1 - s);                     // t == -1 / T on the extra step
      const bool in_range = s < T;
      const bool active = row_ok && in_range && t < len;
      float4 g4[2];
      float2 ct = make_float2(0.f, 0.f), cp = ct, dov = ct;
      if (s > 0 && ew == 0 && lane == 0) mbar_arrive_expect_tx(rfull, KS_CL * KS_U * RB * 4);
      if (active) {
        const float4* gp = reinterpret_cast<const float4*>(gptr);
        g4[0] = gp[0];
        g4[1] = gp[1];
        ct = *reinterpret_cast<const float2*>(cptr);
        const int tprev = dir ? t + 1 : t - 1;                 // frame the forward pass processed before t
        if (tprev >= 0 && tprev < len)
          cp = *reinterpret_cast<const float2*>(cptr + rstep * (2LL * H));   // tprev = t + (dir ? 1 : -1)
        else
          cp = *reinterpret_cast<const float2*>(c0ptr);
        dov = *reinterpret_cast<const float2*>(dptr);
      }
      float dhr[2] = {0.0f, 0.0f};
      if (s > 0) {
        if (warp == 2) mbar_wait(accum_full, (s - 1) & 1);     // one warp polls, the others sleep on a hardware barrier
        asm volatile("bar.sync 2, %0;" ::"n"(32 * KS_EPI_WARPS) : "memory");
        if (threadIdx.x == 64) KS_STAMP(4);
        if (threadIdx.x == 64) KS_GSTAMP(0);
        tc_fence_after();
        if (16 * cgp < RB) {                                   // (warp-uniform) column groups past the padded batch have no columns
        float v[32];
        tmem_ld32(trow + 32 * cgp, v);                         // this W^T row x dG rows (b = 16 cgp + i, plane): v[2 i + plane]
        tc_fence_before();
        // W_hi . (dG_hi + dG_lo) in the even lane, W_lo . dG_hi in the odd one; their sum is the unit's partial dh
        float sm[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float a = pw ? v[2 * i] : v[2 * i] + v[2 * i + 1];
          sm[i] = a + __shfl_xor_sync(0xffffffffu, a, 1);
        }
        // each lane of the pair sends eight consecutive batch rows to the peer that finalises the unit
        st_async_v4(rdst, pw ? sm[8] : sm[0], pw ? sm[9] : sm[1], pw ? sm[10] : sm[2], pw ? sm[11] : sm[3], rbar);
        st_async_v4(rdst + 16, pw ? sm[12] : sm[4], pw ? sm[13] : sm[5], pw ? sm[14] : sm[6], pw ? sm[15] : sm[7], rbar);
        }
        if (threadIdx.x == 64) KS_STAMP(5);
        while (!mbar_try_wait_cluster(rfull, (s - 1) & 1)) {
        }
        if (threadIdx.x == 64) KS_STAMP(6);
        if (threadIdx.x == 64) KS_GSTAMP(1);
#pragma unroll
        for (int r = 0; r < KS_CL; ++r) {
          dhr[0] += rbuf[(r * KS_U + ul) * KS_RP + fb];
          dhr[1] += rbuf[(r * KS_U + ul + 1) * KS_RP + fb];
        }
      }
      // the row's forward pass started at frame 0 (dir 0) / len-1 (dir 1): one step later dh_rec is d(h0)
      if (row_ok && p.want_d0 && ((dir == 0 && s == T) || (dir == 1 && t == len))) {
        dh0[0] = dhr[0];
        dh0[1] = dhr[1];
      }
      float dgv[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) dgv[i] = 0.0f;
      if (active) {
        const bool first = dir ? (t == 0) : (t == len - 1);    // first frame of this row in backward order
        const float cta_[2] = {ct.x, ct.y}, cpa[2] = {cp.x, cp.y}, doa[2] = {dov.x, dov.y};
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          float dh = doa[j] + dhr[j];
          float dct = dc[j];
          if (first && p.dhn) {
            dh += p.dhn[((long long)dir * Bp + fb) * H + ug + j];
            dct += p.dcn[((long long)dir * Bp + fb) * H + ug + j];
          }
          const float gi = g4[j].x, gf = g4[j].y, gg = g4[j].z, go = g4[j].w;
          const float tc = tanhf_(cta_[j]);
          dct += dh * go * (1.0f - tc * tc);
          dgv[4 * j + 0] = dct * gg * gi * (1.0f - gi);
          dgv[4 * j + 1] = dct * cpa[j] * gf * (1.0f - gf);
          dgv[4 * j + 2] = dct * gi * (1.0f - gg * gg);
          dgv[4 * j + 3] = dh * tc * go * (1.0f - go);
          dc[j] = dct * gf;
        }
      }
      __nv_bfloat16 hi[8], lo[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) split_bf16(dgv[i], hi[i], lo[i]);
      // The CTA's dG block of the step (B rows x 2 planes x 64 columns) is staged in shared memory and leaves as ONE tensor
      // store: written by the async proxy it needs no proxy fence before the consumers' TMA loads, and the 512 threads
      // wait for no store acknowledgement (their stores + fence were 1.3k cycles of the step).  Units >= H (last cluster)
      // have no columns: that CTA stores a 32-column box or nothing.
      if (fb < B && in_range && 2 * u2 < nvalid) {
        const uint32_t a = dgt_u32 + (uint32_t)(2 * fb) * tile_pitch + (uint32_t)u2 * 16u;
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(reinterpret_cast<uint32_t*>(hi)[0]),
                     "r"(reinterpret_cast<uint32_t*>(hi)[1]), "r"(reinterpret_cast<uint32_t*>(hi)[2]), "r"(reinterpret_cast<uint32_t*>(hi)[3])
                     : "memory");
        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a + tile_pitch), "r"(reinterpret_cast<uint32_t*>(lo)[0]),
                     "r"(reinterpret_cast<uint32_t*>(lo)[1]), "r"(reinterpret_cast<uint32_t*>(lo)[2]), "r"(reinterpret_cast<uint32_t*>(lo)[3])
                     : "memory");
      }
      gptr += rstep * (8LL * H);
      cptr += rstep * (2LL * H);
      dptr += rstep * (2LL * H);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // this thread's tile stores -> the tensor store below
      if (threadIdx.x == 64) KS_STAMP(7);
      if (threadIdx.x == 64) KS_GSTAMP(2);
      asm volatile("bar.sync 1, %0;" ::"n"(32 * KS_EPI_WARPS) : "memory");
      if (threadIdx.x == 64) {
        KS_STAMP(8);
        if (in_range && nvalid > 0) {
          ks_tma_store_3d(nvalid == KS_U ? &tmS64 : &tmS32, dgt, dir * 4 * H + 4 * (c * KS_UC + rank * KS_U), 0, t * Bp);
          ks_tma_store_wait();                                 // the block is in global memory (and the tile reusable)
        }
        red_release_add(counter, 1u);
        KS_STAMP(9);
        KS_GSTAMP(3);
      }
    }
    if (row_ok && p.want_d0) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        p.dh0[((long long)dir * Bp + fb) * H + ug + j] = dh0[j];
        p.dc0[((long long)dir * Bp + fb) * H + ug + j] = dc[j];
      }
    }
  }
  __syncthreads();
  cluster_sync_all();                                          // nobody leaves while a peer may still address it
  if (warp == 1) {
    tc_fence_after();
    tmem_free<KS_TMEM>(tmem_d);
  }
}

int launch_lstm_bwd_ks(const LstmBwdParams& p, const void* whhT, long long whhT_ld, long long whhT_plane, cudaStream_t st) {
  const int H = KS_H;
  const int rb = p.B <= 16 ? 16 : p.B <= 32 ? 32 : 64;
  CUtensorMap tmG;
  {  // (columns, plane, row): a box lands in shared memory as rows 2*b + plane
    uint64_t dims[3] = {(uint64_t)8 * H, 2, (uint64_t)p.T * p.Bp - (uint64_t)(p.Bp - p.B)};   // rows of THIS sub-batch's view
    uint64_t str[2] = {(uint64_t)p.dG_plane * 2, (uint64_t)8 * H * 2};
    uint32_t box[3] = {64, 2, (uint32_t)rb};
    int rc = make_tmap_bf16(&tmG, p.dG, 3, dims, str, box);
    if (rc) return rc;
  }
  CUtensorMap tmS64, tmS32;   // the same tensor, boxes of B rows x 2 planes x (64 | 32) columns, no swizzle: the CTAs' stores
  for (int v = 0; v < 2; ++v) {
    uint64_t dims[3] = {(uint64_t)8 * H, 2, (uint64_t)p.T * p.Bp - (uint64_t)(p.Bp - p.B)};
    uint64_t str[2] = {(uint64_t)p.dG_plane * 2, (uint64_t)8 * H * 2};
    uint32_t box[3] = {v ? 32u : 64u, 2, (uint32_t)p.B};
    int rc = make_tmap_bf16(v ? &tmS32 : &tmS64, p.dG, 3, dims, str, box, false);
    if (rc) return rc;
  }
  auto kern = rb == 16 ? (p.dbg ? lstm_bwd_ks_kernel<16, true> : lstm_bwd_ks_kernel<16, false>)
            : rb == 32 ? (p.dbg ? lstm_bwd_ks_kernel<32, true> : lstm_bwd_ks_kernel<32, false>)
                       : (p.dbg ? lstm_bwd_ks_kernel<64, true> : lstm_bwd_ks_kernel<64, false>);
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, KS_SMEM);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * KS_NCTA);
  cfg.blockDim = dim3(KS_THREADS);
  cfg.dynamicSmemBytes = KS_SMEM;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = KS_CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  // The CTAs spin on each other (per-step counter), so all 20 clusters must be able to run at once.  This is checked
  // here instead of asking for a cooperative launch: Nsight Compute cannot replay a launch that is both cooperative
  // and clustered, and work that already occupies SMs (the flag-gated GEMMs of the side stream) always finishes on
  // its own, so late CTAs only wait.
  static int max_clusters[2] = {-1, -1};   // (the three row variants use the same resources)
  if (max_clusters[p.dbg ? 1 : 0] < 0) {
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess) n = 0;
    max_clusters[p.dbg ? 1 : 0] = n;
  }
  if (max_clusters[p.dbg ? 1 : 0] < 2 * KS_NCL) {
    set_error("ss_lstm_bwd (K-split clusters): only %d of the %d clusters can be co-resident on this device",
              max_clusters[p.dbg ? 1 : 0], 2 * KS_NCL);
    return -2;
  }
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmG, tmS64, tmS32, p, (const __nv_bfloat16*)whhT, whhT_ld, whhT_plane);
  if (e != cudaSuccess) {
    set_error("ss_lstm_bwd (K-split clusters): launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}

}  // namespace ssb
