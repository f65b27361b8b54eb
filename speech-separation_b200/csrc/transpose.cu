// Transposed sb16 copies of the recurrence outputs, produced AFTER the persistent kernels by one HBM-bound pass
// (the dW GEMMs want K = T*B contiguous).  Writing them from inside the recurrence as scattered 2-byte stores made
// every step's gpu-scope release wait ~7.5k cycles for those stores to drain (tools/trace_lstm_bwd.py).
#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

// in: sb16 [R][C] (pitch ld, planes at +plane)  ->  out: sb16 [C][R] (pitch ldT, planes at +planeT).
// HPREV: the value for row r = t*B+b, column j (dir = j / H) is the state carried INTO step t of that direction:
// in[(t-1)*B+b][j] (dir 0, t >= 1) / in[(t+1)*B+b][j] (dir 1, t+1 < len_b), else h0[dir][b][j % H].
template <bool HPREV>
__global__ void __launch_bounds__(256) transpose_sb16_kernel(const __nv_bfloat16* __restrict__ in, long long ld,
                                                             long long plane, int R, int C,
                                                             __nv_bfloat16* __restrict__ out, long long ldT,
                                                             long long planeT, int B, int H,
                                                             const int* __restrict__ lens, const float* __restrict__ h0) {
  __shared__ __nv_bfloat16 tile[64][66];
  const int c0 = blockIdx.x * 64, r0 = blockIdx.y * 64, pl = blockIdx.z;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const __nv_bfloat16* src = in + (long long)pl * plane;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = r0 + ty + 8 * i, c = c0 + 2 * tx;
    __nv_bfloat162 v = __floats2bfloat162_rn(0.0f, 0.0f);
    if (r < R && c < C) {                                       // C is even (2H, 8H, in_dim padded by the caller)
      if (!HPREV) {
        v = *reinterpret_cast<const __nv_bfloat162*>(src + (long long)r * ld + c);
      } else {
        const int t = r / B, b = r - t * B, dir = c / H;
        const int tp = dir ? t + 1 : t - 1;
        const bool valid = dir ? (tp < lens[b]) : (tp >= 0);
        if (valid) {
          v = *reinterpret_cast<const __nv_bfloat162*>(src + ((long long)tp * B + b) * ld + c);
        } else {
          const float2 f = *reinterpret_cast<const float2*>(h0 + ((long long)dir * B + b) * H + (c - dir * H));
          __nv_bfloat16 hx, lx, hy, ly;
          split_bf16(f.x, hx, lx);
          split_bf16(f.y, hy, ly);
          v.x = pl ? lx : hx;
          v.y = pl ? ly : hy;
        }
      }
    }
    tile[ty + 8 * i][2 * tx] = v.x;
    tile[ty + 8 * i][2 * tx + 1] = v.y;
  }
  __syncthreads();
  __nv_bfloat16* dst = out + (long long)pl * planeT;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = c0 + ty + 8 * i, r = r0 + 2 * tx;
    if (c < C && r < R) {
      __nv_bfloat162 v;
      v.x = tile[2 * tx][ty + 8 * i];
      v.y = tile[2 * tx + 1][ty + 8 * i];
      if (r + 1 < R) *reinterpret_cast<__nv_bfloat162*>(dst + (long long)c * ldT + r) = v;
      else dst[(long long)c * ldT + r] = v.x;
    }
  }
}

int launch_transpose_sb16(const void* in, long long ld, long long plane, int R, int C, void* out, long long ldT,
                          long long planeT, cudaStream_t st) {
  if (R == 0 || C == 0) return 0;
  dim3 grid(cdiv(C, 64), cdiv(R, 64), 2);
  transpose_sb16_kernel<false><<<grid, 256, 0, st>>>((const __nv_bfloat16*)in, ld, plane, R, C, (__nv_bfloat16*)out, ldT,
                                                     planeT, 1, 1, nullptr, nullptr);
  return check_launch("transpose_sb16");
}

int launch_hprev_transpose(const void* out_sb, long long ld, long long plane, int T, int B, int H, const int* lens,
                           const float* h0, void* hprevT, long long ldT, long long planeT, cudaStream_t st) {
  dim3 grid(cdiv(2 * H, 64), cdiv((long long)T * B, 64), 2);
  transpose_sb16_kernel<true><<<grid, 256, 0, st>>>((const __nv_bfloat16*)out_sb, ld, plane, T * B, 2 * H,
                                                    (__nv_bfloat16*)hprevT, ldT, planeT, B, H, lens, h0);
  return check_launch("hprev_transpose");
}

}  // namespace ssb

// The transposed operand copies on their own, so the caller can keep them off the critical stream
// (ss_lstm_fwd with outT = hprevT = NULL, ss_lstm_bwd with dGT = NULL skip them).
extern "C" int ss_transpose_sb16(const void* in, long long ld, long long plane, int R, int C, void* out, long long ldT,
                                 long long planeT, void* stream) {
  return ssb::launch_transpose_sb16(in, ld, plane, R, C, out, ldT, planeT, (cudaStream_t)stream);
}
extern "C" int ss_hprev_transpose(const void* out_sb, long long ld, long long plane, int T, int B, int H, const int* lens,
                                  const float* h0, void* hprevT, long long ldT, long long planeT, void* stream) {
  return ssb::launch_hprev_transpose(out_sb, ld, plane, T, B, H, lens, h0, hprevT, ldT, planeT, (cudaStream_t)stream);
}
