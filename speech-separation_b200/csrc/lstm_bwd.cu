// K2c: persistent bidirectional LSTM recurrence, backward pass (reverse-time), one cooperative launch per layer.
// Stands in for the sequential part of cudnnRNNBackwardData that `loss.backward()` (steps/train_qsub.py:120)
// runs for `self.blstm` (archs/uPIT.py:132).  Each CTA owns LB_U hidden units j of one direction, keeps
// W_hh^T[j][n] (n = interleaved gate index 4u+g, split-bf16) resident in shared memory and per step computes
//   dh_rec[b][j] = sum_n dG_prev[b][n] * W_hh[n][j]            (tcgen05, K = 4H)
// with dG_prev (the gate pre-activation gradients of the previously processed frame) streamed by TMA straight
// from the dG matrix all CTAs of the direction write; then the cell backward for its units, producing dG[t].
// dG is written as sb16 [T*B][8H]: the K-major operand of the dX GEMM and the MN-major operand of the dW GEMMs.
#include <stdlib.h>

#include "../../include/ssb200.h"
#include "common.cuh"
#include "lstm_bwd.cuh"

namespace ssb {

constexpr int LB_U = 16;
constexpr int LB_EPI_WARPS = 16;            // 4 TMEM lane quadrants x 4 column groups of 4 units
constexpr int LB_THREADS = 64 + 32 * LB_EPI_WARPS;
constexpr int LB_WCHUNK = 2 * LB_U * 128;   // 4 KB
constexpr int LB_MAXKC = 38;                // 4H <= 2432
constexpr int LB_XPITCH = 20;               // floats per row of the lo-row exchange buffer


__global__ void __launch_bounds__(LB_THREADS, 1)
lstm_bwd_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmG, LstmBwdParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = 2 * p.R * 128;
  unsigned char* ring = smem;
  unsigned char* wsm = smem + p.NS * stage_bytes;
  float* xbuf = reinterpret_cast<float*>(wsm + p.KC * LB_WCHUNK);         // [64][LB_XPITCH] dG_lo-row partial sums
  uint64_t* full = reinterpret_cast<uint64_t*>(xbuf + 64 * LB_XPITCH);
  uint64_t* empty = full + 4;
  uint64_t* accum_full = empty + 4;
  uint64_t* wfull = accum_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int j0 = cta * LB_U;
  const int T = p.T, B = p.B, H = p.H;
  const int nsteps = T + (p.want_d0 ? 1 : 0);
  unsigned int* counter = p.sync + dir;
  const int kc0 = p.ROT ? (cta * p.ROT) % p.KC : 0;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmG);
    for (int s = 0; s < p.NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(accum_full, 1);
    mbar_init(wfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<32>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, p.KC * LB_WCHUNK);
      for (int kc = 0; kc < p.KC; ++kc) tma_load_3d(wsm + kc * LB_WCHUNK, &tmW, wfull, kc * 64, dir * H + j0, 0);
      int it = 0;
      for (int s = 1; s < nsteps; ++s) {
        // frame processed at step s-1 (whose dG this step consumes)
        const int tp = dir ? (s - 1) : (T - s);
        const unsigned int need = (unsigned int)p.NC * (unsigned int)s;
        while (ld_acquire(counter) < need) {
        }
        fence_proxy_async();
        for (int i = 0; i < p.KC; ++i, ++it) {
          const int kc = (kc0 + i) % p.KC;               // every CTA starts on a different chunk: spreads the L2 slices
          const int st = it % p.NS, ph = (it / p.NS) & 1;
          mbar_wait(&empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&full[st], stage_bytes);
          tma_load_3d(ring + st * stage_bytes, &tmG, &full[st], dir * 4 * H + kc * 64, tp * B, 0);
        }
        // progress for stream-ordered consumers (ss_stream_wait_value32): the counter just observed covers the dG rows of
        // steps 0..s-1 of this direction; one CTA per direction reports every 4th step, after its copies are in flight
        if (p.progress && cta == 0 && (s & 3) == 0) {
          __threadfence_system();
          *reinterpret_cast<volatile unsigned int*>(p.progress + dir) = p.prog_base + (unsigned int)s;
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // B <= 64: ONE instruction per 16-wide k-step on the stacks A = [dG_hi; dG_lo] (128 rows) and
      // B = [W_hi; W_lo] (32 rows); the four 64x16 quadrants of D are summed in the epilogue.  (tcgen05.mma costs
      // >= ~53 cycles per instruction whatever its shape, tools/microbench/mma_rate.cu.)
      constexpr uint32_t idesc32 = umma_idesc_bf16(128, 2 * LB_U), idesc16 = umma_idesc_bf16(128, LB_U);
      const bool stacked = p.R == 64;
      const uint64_t a_base = umma_desc_sw128(smem_u32(ring)), b_base = umma_desc_sw128(smem_u32(wsm));
      const uint32_t a_stage = stage_bytes >> 4, a_lo_off = (p.R * 128) >> 4;
      mbar_wait(wfull, 0);
      int it = 0;
      for (int s = 1; s < nsteps; ++s) {
        for (int i = 0; i < p.KC; ++i, ++it) {
          const int kc = (kc0 + i) % p.KC;
          const int st = it % p.NS, ph = (it / p.NS) & 1;
          mbar_wait(&full[st], ph);
          tc_fence_after();
          const uint64_t da = a_base + (uint32_t)st * a_stage;
          const uint64_t db = b_base + (uint32_t)kc * (LB_WCHUNK >> 4);
          const int ksteps = min(4, (4 * H - kc * 64 + 15) / 16);
          if (stacked) {
            for (int k = 0; k < ksteps; ++k) umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc32, (i | k) ? 1u : 0u);
          } else {
            for (int k = 0; k < ksteps; ++k) {
              umma_bf16(tmem_d, da + 2 * k, db + 2 * k, idesc32, (i | k) ? 1u : 0u);          // dG_hi x [W_hi; W_lo]
              umma_bf16(tmem_d, da + a_lo_off + 2 * k, db + 2 * k, idesc16, 1u);               // dG_lo x W_hi
            }
          }
          umma_commit(&empty[st]);
        }
        umma_commit(accum_full);
      }
    }
  } else {
    const int ew = warp - 2;
    const int q = warp & 3, cg = ew >> 2;
    const bool stacked = p.R == 64;
    const bool helper = stacked && q >= 2;                     // these warps own the dG_lo rows of D
    const int b = (stacked ? (q & 1) : q) * 32 + lane;
    const int ug = j0 + cg * 4;
    const bool row_ok = !helper && b < B && ug < H;
    const int len = row_ok ? p.lens[b] : 0;
    float* xr = xbuf + b * LB_XPITCH + 4 * cg;
    const uint32_t trow = tmem_d + ((uint32_t)(q * 32) << 16) + 4 * cg;
    float dc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    float dh0[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    for (int s = 0; s < nsteps; ++s) {
      const int t = dir ? s : (T - 1 - s);               // t == -1 / T on the extra step
      const bool in_range = s < T;
      const bool active = row_ok && in_range && t < len;
      const long long row = (long long)(in_range ? t : 0) * B + (row_ok ? b : 0);
      float4 g4[4];
      float4 ct = make_float4(0.f, 0.f, 0.f, 0.f), cp = ct, dov = ct;
      if (active) {
        const float4* gp = reinterpret_cast<const float4*>(p.gates + row * (8LL * H) + (long long)dir * 4 * H + 4 * ug);
#pragma unroll
        for (int j = 0; j < 4; ++j) g4[j] = gp[j];
        ct = *reinterpret_cast<const float4*>(p.cs + row * (2LL * H) + dir * H + ug);
        const int tprev = dir ? t + 1 : t - 1;             // frame the forward pass processed before t
        if (tprev >= 0 && tprev < len)
          cp = *reinterpret_cast<const float4*>(p.cs + ((long long)tprev * B + b) * (2LL * H) + dir * H + ug);
        else
          cp = *reinterpret_cast<const float4*>(p.c0 + ((long long)dir * B + b) * H + ug);
        dov = *reinterpret_cast<const float4*>(p.dout + row * (2LL * H) + dir * H + ug);
      }
      float dhr[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      if (s > 0) {
        if (warp == 2) mbar_wait(accum_full, (s - 1) & 1);     // one warp polls, the others sleep on a hardware barrier
        asm volatile("bar.sync 2, %0;" ::"n"(32 * LB_EPI_WARPS) : "memory");
        tc_fence_after();
        float e[4];
        tmem_ld4x2(trow, trow + LB_U, dhr, e);                 // x W_hi and x W_lo columns of this thread's 4 units
        tc_fence_before();
#pragma unroll
        for (int j = 0; j < 4; ++j) dhr[j] += e[j];
        if (stacked) {
          if (helper) *reinterpret_cast<float4*>(xr) = make_float4(dhr[0], dhr[1], dhr[2], dhr[3]);
          asm volatile("bar.sync 3, %0;" ::"n"(32 * LB_EPI_WARPS) : "memory");
          if (!helper) {
            const float4 v = *reinterpret_cast<const float4*>(xr);
            dhr[0] += v.x; dhr[1] += v.y; dhr[2] += v.z; dhr[3] += v.w;
          }
        }
      }
      // the row's forward pass started at frame 0 (dir 0) / len-1 (dir 1): one step later dh_rec is d(h0)
      if (row_ok && p.want_d0 && ((dir == 0 && s == T) || (dir == 1 && t == len))) {
#pragma unroll
        for (int j = 0; j < 4; ++j) dh0[j] = dhr[j];
      }
      float dgv[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) dgv[i] = 0.0f;
      if (active) {
        const bool first = dir ? (t == 0) : (t == len - 1);  // first frame of this row in backward order
        const float cta_[4] = {ct.x, ct.y, ct.z, ct.w}, cpa[4] = {cp.x, cp.y, cp.z, cp.w};
        const float doa[4] = {dov.x, dov.y, dov.z, dov.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float dh = doa[j] + dhr[j];
          float dct = dc[j];
          if (first && p.dhn) {
            dh += p.dhn[((long long)dir * B + b) * H + ug + j];
            dct += p.dcn[((long long)dir * B + b) * H + ug + j];
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
      __nv_bfloat16 hi[16], lo[16];
#pragma unroll
      for (int i = 0; i < 16; ++i) split_bf16(dgv[i], hi[i], lo[i]);
      if (row_ok && in_range) {
        __nv_bfloat16* o = p.dG + row * (8LL * H) + (long long)dir * 4 * H + 4 * ug;
        uint4* o4 = reinterpret_cast<uint4*>(o);
        o4[0] = reinterpret_cast<uint4*>(hi)[0];
        o4[1] = reinterpret_cast<uint4*>(hi)[1];
        uint4* l4 = reinterpret_cast<uint4*>(o + p.dG_plane);
        l4[0] = reinterpret_cast<uint4*>(lo)[0];
        l4[1] = reinterpret_cast<uint4*>(lo)[1];
      }
      asm volatile("bar.sync 1, %0;" ::"n"(32 * LB_EPI_WARPS) : "memory");
      if (threadIdx.x == 64) {
        __threadfence();
        fence_proxy_async();
        red_release_add(counter, 1u);
      }
    }
    if (row_ok && p.want_d0) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        p.dh0[((long long)dir * B + b) * H + ug + j] = dh0[j];
        p.dc0[((long long)dir * B + b) * H + ug + j] = dc[j];
      }
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_free<32>(tmem_d);
  }
}

}  // namespace ssb

extern "C" int ss_lstm_bwd(const float* gates, const float* cs, const float* c0, const float* dout, const float* dhn,
                           const float* dcn, const int* lens, const void* whhT, long long whhT_ld,
                           long long whhT_plane, void* dG, long long dG_plane, float* dh0, float* dc0, void* ws,
                           unsigned int* progress, unsigned int prog_base, int T, int B, int batch_pitch, int H,
                           void* stream) {
  using namespace ssb;
  SS_REQUIRE(T > 0 && B > 0 && B <= 128, "ss_lstm_bwd: need 1 <= B <= 128 (got %d) and T > 0", B);
  if (batch_pitch <= 0) batch_pitch = B;
  SS_REQUIRE(batch_pitch >= B, "ss_lstm_bwd: batch_pitch %d < B %d", batch_pitch, B);
  SS_REQUIRE(batch_pitch == B || (H == 600 && B <= 64), "ss_lstm_bwd: a sub-batch view (batch_pitch != B) needs H = 600, B <= 64");
  SS_REQUIRE(H % 8 == 0 && 4 * H <= 64 * LB_MAXKC, "ss_lstm_bwd: unsupported hidden size %d", H);
  cudaStream_t st = (cudaStream_t)stream;
  LstmBwdParams p;
  p.gates = gates; p.cs = cs; p.c0 = c0; p.dout = dout; p.dhn = dhn; p.dcn = dcn; p.lens = lens;
  p.dG = (__nv_bfloat16*)dG; p.dG_plane = dG_plane;
  p.dh0 = dh0; p.dc0 = dc0;
  p.T = T; p.B = B; p.H = H; p.Bp = batch_pitch;
  p.R = B <= 64 ? 64 : 128;
  p.NC = cdiv(H, LB_U);
  p.KC = cdiv(4 * H, 64);
  p.NS = p.R == 64 ? 4 : 2;
  p.want_d0 = dh0 != nullptr;
  p.dbg = g_lstm_dbg;
  p.ROT = 1;
  p.sync = (unsigned int*)ws;
  p.progress = progress;
  p.prog_base = prog_base;
  cudaMemsetAsync(ws, 0, 256, st);
  if (H == 600 && B <= 64) return launch_lstm_bwd_ks(p, whhT, whhT_ld, whhT_plane, st);

  CUtensorMap tmW, tmG;
  {
    uint64_t dims[3] = {(uint64_t)4 * H, (uint64_t)2 * H, 2};
    uint64_t str[2] = {(uint64_t)whhT_ld * 2, (uint64_t)whhT_plane * 2};
    uint32_t box[3] = {64, LB_U, 2};
    int rc = make_tmap_bf16(&tmW, whhT, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[3] = {(uint64_t)8 * H, (uint64_t)T * B, 2};
    uint64_t str[2] = {(uint64_t)8 * H * 2, (uint64_t)dG_plane * 2};
    uint32_t box[3] = {64, (uint32_t)p.R, 2};
    int rc = make_tmap_bf16(&tmG, dG, 3, dims, str, box);
    if (rc) return rc;
  }
  const int smem = p.NS * 2 * p.R * 128 + p.KC * LB_WCHUNK + 64 * LB_XPITCH * 4 + 1024 + 256;
  cudaFuncSetAttribute(lstm_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&tmW, (void*)&tmG, (void*)&p};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)lstm_bwd_kernel, dim3(2 * p.NC), dim3(LB_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_bwd: cooperative launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}
