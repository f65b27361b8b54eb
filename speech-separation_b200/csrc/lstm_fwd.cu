// K2b: persistent bidirectional LSTM recurrence (forward pass), one cooperative launch per layer.
// Replaces the sequential part of cudnnRNNForward behind `self.blstm(x, self.hidden)` (archs/uPIT.py:115,132;
// archs/RSH.py:155,172) for a length-sorted padded batch with PackedSequence semantics:
//   * the reverse direction starts at each utterance's own last frame, from (h0, c0);
//   * outputs at t >= len_b are zero, (h_n, c_n) are the states after each utterance's last processed frame.
// Each CTA owns LF_U hidden units of one direction with their 4 gates interleaved (column 4*u+gate), keeps its
// slice of W_hh (split-bf16) resident in shared memory, and per step computes
//   D[b][4u+g] = sum_k h_{prev}[b][k] * W_hh[g*H+u][k]      (tcgen05, fp32 accumulate in TMEM)
// from h_prev streamed by TMA out of a small ping-pong buffer that all CTAs of the direction publish into;
// a gpu-scope release/acquire counter per direction is the per-step grid barrier.
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

constexpr int LF_U = 12;                    // hidden units per CTA
constexpr int LF_N = 4 * LF_U;              // gate columns per CTA (MMA N)
constexpr int LF_EPI_WARPS = 12;            // 4 TMEM lane quadrants x 3 column groups of 4 units
constexpr int LF_THREADS = 64 + 32 * LF_EPI_WARPS;
constexpr int LF_WCHUNK = 2 * LF_N * 128;   // one 64-wide K chunk of the W slice, both planes
constexpr int LF_MAXKC = 10;                // H <= 640

struct LstmFwdParams {
  float* gx;                 // [T*B][8H] in: x-projection + biases (interleaved gate columns); out: activated gates
  const float* h0;           // [2][B][H]
  const float* c0;
  const int* lens;           // [B]
  __nv_bfloat16* out;        // sb16 [T*B][2H]
  long long out_plane;
  __nv_bfloat16* outT;       // sb16 [2H][ldT] or null: transposed outputs (X^T of the next layer's dW GEMM)
  __nv_bfloat16* hprevT;     // sb16 [2H][ldT] or null: transposed carried state BEFORE each step (dW_hh GEMM)
  long long ldT, outT_plane;
  float* cs;                 // [T*B][2H] or null
  float* hn;                 // [2][B][H]
  float* cn;
  __nv_bfloat16* hbuf;       // [slot 2][dir 2][plane 2][R][H]
  unsigned int* sync;        // [2]
  int T, B, H, R, NC, KC, NS;
};

__global__ void __launch_bounds__(LF_THREADS, 1)
lstm_fwd_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmH, LstmFwdParams p) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stage_bytes = 2 * p.R * 128;
  unsigned char* ring = smem;                                  // NS stages of h_prev chunks
  unsigned char* wsm = smem + p.NS * stage_bytes;              // KC chunks of the W slice (also absorbs M=128 over-read)
  uint64_t* full = reinterpret_cast<uint64_t*>(wsm + p.KC * LF_WCHUNK);
  uint64_t* empty = full + 4;
  uint64_t* accum_full = empty + 4;
  uint64_t* wfull = accum_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dir = blockIdx.x / p.NC, cta = blockIdx.x % p.NC;
  const int u0 = cta * LF_U;
  const int T = p.T, B = p.B, H = p.H;
  unsigned int* counter = p.sync + dir;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmH);
    for (int s = 0; s < p.NS; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(accum_full, 1);
    mbar_init(wfull, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc<64>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      mbar_arrive_expect_tx(wfull, p.KC * LF_WCHUNK);
      for (int kc = 0; kc < p.KC; ++kc) tma_load_3d(wsm + kc * LF_WCHUNK, &tmW, wfull, kc * 64, dir * 4 * H + 4 * u0, 0);
      int it = 0;
      for (int s = 0; s < T; ++s) {
        const unsigned int need = (unsigned int)p.NC * (unsigned int)(s + 1);
        while (ld_acquire(counter) < need) {
        }
        fence_proxy_async();
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int st = it % p.NS, ph = (it / p.NS) & 1;
          mbar_wait(&empty[st], ph ^ 1);
          mbar_arrive_expect_tx(&full[st], stage_bytes);
          asm volatile(
              "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::
                  "r"(smem_u32(ring + st * stage_bytes)), "l"((uint64_t)&tmH), "r"(smem_u32(&full[st])), "r"(kc * 64), "r"(0),
              "r"(0), "r"((s & 1) * 2 + dir)
              : "memory");
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc_bf16(128, LF_N);
      mbar_wait(wfull, 0);
      int it = 0;
      for (int s = 0; s < T; ++s) {
        for (int kc = 0; kc < p.KC; ++kc, ++it) {
          const int st = it % p.NS, ph = (it / p.NS) & 1;
          mbar_wait(&full[st], ph);
          tc_fence_after();
          const uint32_t a_hi = smem_u32(ring + st * stage_bytes), a_lo = a_hi + p.R * 128;
          const uint32_t b_hi = smem_u32(wsm + kc * LF_WCHUNK), b_lo = b_hi + LF_N * 128;
          const int ksteps = min(4, (H - kc * 64 + 15) / 16);
          for (int k = 0; k < ksteps; ++k) {
            const uint32_t ko = k * 32;
            const uint64_t dah = umma_desc_sw128(a_hi + ko), dal = umma_desc_sw128(a_lo + ko);
            const uint64_t dbh = umma_desc_sw128(b_hi + ko), dbl = umma_desc_sw128(b_lo + ko);
            umma_bf16(tmem_d, dal, dbh, idesc, (kc | k) ? 1u : 0u);
            umma_bf16(tmem_d, dah, dbl, idesc, 1u);
            umma_bf16(tmem_d, dah, dbh, idesc, 1u);
          }
          umma_commit(&empty[st]);
        }
        umma_commit(accum_full);
      }
    }
  } else {
    const int ew = warp - 2;
    const int q = warp & 3, cg = ew >> 2;
    const int b = q * 32 + lane;
    const int ul = cg * 4;                                     // first of this thread's 4 units within the CTA
    const int ug = u0 + ul;                                    // ... within the direction
    const bool row_ok = b < B;
    const int len = row_ok ? p.lens[b] : 0;
    float h[4], c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      h[j] = row_ok ? p.h0[((long long)dir * B + b) * H + ug + j] : 0.0f;
      c[j] = row_ok ? p.c0[((long long)dir * B + b) * H + ug + j] : 0.0f;
    }
    const long long hb_plane = (long long)p.R * H;
    auto publish = [&](int slot) {
      if (row_ok) {
        __nv_bfloat16 hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_bf16(h[j], hi[j], lo[j]);
        __nv_bfloat16* dst = p.hbuf + ((long long)(slot * 2 + dir) * 2) * hb_plane + (long long)b * H + ug;
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<uint2*>(hi);
        *reinterpret_cast<uint2*>(dst + hb_plane) = *reinterpret_cast<uint2*>(lo);
      }
      fence_proxy_async();
      __threadfence();
      asm volatile("bar.sync 1, %0;" ::"n"(32 * LF_EPI_WARPS) : "memory");
      if (threadIdx.x == 64) red_release_add(counter, 1u);
    };
    publish(0);
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
      mbar_wait(accum_full, s & 1);
      tc_fence_after();
      float d[16];
      tmem_ld16(tmem_d + ((uint32_t)(q * 32) << 16) + 16 * cg, d);
      tc_fence_before();
      float hout[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      const float hp[4] = {h[0], h[1], h[2], h[3]};
      if (active) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gi = sigmoidf_(d[4 * j + 0] + g4[j].x), gf = sigmoidf_(d[4 * j + 1] + g4[j].y);
          const float gg = tanhf_(d[4 * j + 2] + g4[j].z), go = sigmoidf_(d[4 * j + 3] + g4[j].w);
          c[j] = gf * c[j] + gi * gg;
          h[j] = go * tanhf_(c[j]);
          hout[j] = h[j];
          g4[j] = make_float4(gi, gf, gg, go);
        }
      }
      publish((s + 1) & 1);                                    // carried state for the next step (critical path first)
      if (row_ok) {
        if (active) {
#pragma unroll
          for (int j = 0; j < 4; ++j) gptr[j] = g4[j];
          if (p.cs) *reinterpret_cast<float4*>(p.cs + row * (2LL * H) + dir * H + ug) = make_float4(c[0], c[1], c[2], c[3]);
        }
        __nv_bfloat16 hi[4], lo[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) split_bf16(hout[j], hi[j], lo[j]);
        __nv_bfloat16* o = p.out + row * (2LL * H) + dir * H + ug;
        *reinterpret_cast<uint2*>(o) = *reinterpret_cast<uint2*>(hi);
        *reinterpret_cast<uint2*>(o + p.out_plane) = *reinterpret_cast<uint2*>(lo);
        if (p.outT) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const long long oT = (long long)(dir * H + ug + j) * p.ldT + row;
            p.outT[oT] = hi[j];
            p.outT[oT + p.outT_plane] = lo[j];
            __nv_bfloat16 phi, plo;
            split_bf16(hp[j], phi, plo);
            p.hprevT[oT] = phi;
            p.hprevT[oT + p.outT_plane] = plo;
          }
        }
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
  if (warp == 1) {
    tc_fence_after();
    tmem_free<64>(tmem_d);
  }
}

}  // namespace ssb

extern "C" long long ss_lstm_ws_bytes(int B, int H) {
  const long long R = B <= 64 ? 64 : 128;
  return 2LL * 2 * 2 * R * H * 2 + 256;       // hbuf + sync counters
}

extern "C" int ss_lstm_fwd(float* gx, const void* whh_r, long long whh_ld, long long whh_plane, const float* h0,
                           const float* c0, const int* lens, void* out, long long out_plane, void* outT, void* hprevT,
                           long long ldT, long long outT_plane, float* cs, float* hn, float* cn, void* ws, int T, int B, int H,
                           void* stream) {
  using namespace ssb;
  SS_REQUIRE(T > 0 && B > 0 && B <= 128, "ss_lstm_fwd: need 1 <= B <= 128 (got %d) and T > 0", B);
  SS_REQUIRE((outT == nullptr) == (hprevT == nullptr), "ss_lstm_fwd: outT and hprevT go together");
  SS_REQUIRE(H % LF_U == 0 && H % 8 == 0 && H <= 64 * LF_MAXKC, "ss_lstm_fwd: unsupported hidden size %d", H);
  cudaStream_t st = (cudaStream_t)stream;
  LstmFwdParams p;
  p.gx = gx; p.h0 = h0; p.c0 = c0; p.lens = lens;
  p.out = (__nv_bfloat16*)out; p.out_plane = out_plane;
  p.outT = (__nv_bfloat16*)outT; p.hprevT = (__nv_bfloat16*)hprevT; p.ldT = ldT; p.outT_plane = outT_plane;
  p.cs = cs; p.hn = hn; p.cn = cn;
  p.T = T; p.B = B; p.H = H;
  p.R = B <= 64 ? 64 : 128;
  p.NC = H / LF_U;
  p.KC = cdiv(H, 64);
  p.NS = p.R == 64 ? 4 : 3;
  const long long hbuf_bytes = 2LL * 2 * 2 * p.R * H * 2;
  p.hbuf = (__nv_bfloat16*)ws;
  p.sync = (unsigned int*)((char*)ws + hbuf_bytes);
  cudaMemsetAsync(ws, 0, hbuf_bytes + 256, st);

  CUtensorMap tmW, tmH;
  {
    uint64_t dims[3] = {(uint64_t)H, (uint64_t)8 * H, 2};
    uint64_t str[2] = {(uint64_t)whh_ld * 2, (uint64_t)whh_plane * 2};
    uint32_t box[3] = {64, LF_N, 2};
    int rc = make_tmap_bf16(&tmW, whh_r, 3, dims, str, box);
    if (rc) return rc;
  }
  {
    uint64_t dims[4] = {(uint64_t)H, (uint64_t)p.R, 2, 4};
    uint64_t str[3] = {(uint64_t)H * 2, (uint64_t)p.R * H * 2, (uint64_t)2 * p.R * H * 2};
    uint32_t box[4] = {64, (uint32_t)p.R, 2, 1};
    int rc = make_tmap_bf16(&tmH, p.hbuf, 4, dims, str, box);
    if (rc) return rc;
  }
  const int smem = p.NS * 2 * p.R * 128 + p.KC * LF_WCHUNK + 1024 + 256;
  cudaFuncSetAttribute(lstm_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  void* args[] = {(void*)&tmW, (void*)&tmH, (void*)&p};
  cudaError_t e = cudaLaunchCooperativeKernel((void*)lstm_fwd_kernel, dim3(2 * p.NC), dim3(LF_THREADS), args, smem, st);
  if (e != cudaSuccess) {
    set_error("ss_lstm_fwd: cooperative launch failed: %s", cudaGetErrorString(e));
    return -2;
  }
  return 0;
}
