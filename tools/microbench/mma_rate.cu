// Microbenchmark: tcgen05.mma (kind::f16, SS) issue-to-completion rate vs shape, one CTA, data resident in smem.
#include <cstdio>
#include "../../speech-separation_b200/csrc/common.cuh"
using namespace ssb;

__global__ void __launch_bounds__(128, 1) k(int M, int N, int nacc, int iters, long long* out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) ((uint32_t*)smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) tmem_alloc<512>(&slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t td = slot;
  if (threadIdx.x == 0) {
    uint32_t idesc = umma_idesc_bf16(M, N);
    uint32_t a = smem_u32(smem), b = a + 32 * 1024;
    long long t0 = clock64();
    const uint64_t da = umma_desc_sw128(a), db = umma_desc_sw128(b);
    for (int i = 0; i < iters; i += 4) {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        umma_bf16(td + ((i + k) & (nacc - 1)) * N, da + 2 * k, db + 2 * k, idesc, (i + k) >= nacc);
    }
    umma_commit(&bar);
    long long t1 = clock64();
    mbar_wait(&bar, 0);
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_free<512>(td); }
}
int main() {
  long long* d; cudaMalloc(&d, 16);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  int Ms[] = {64, 128};
  int Ns[] = {16, 32, 48, 64, 96, 128, 256};
  for (int M : Ms) for (int N : Ns) for (int nacc : {1, 2}) {
    if (M == 128 && N % 16) continue;
    if (nacc * N > 512) continue;
    long long h[2];
    for (int rep = 0; rep < 2; ++rep) { k<<<1, 128, 80 * 1024>>>(M, N, nacc, 512, d); cudaDeviceSynchronize(); }
    cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
    printf("M=%3d N=%3d nacc=%d issue %.1f cyc/mma, complete %.1f cyc/mma  (%s)\n", M, N, nacc, h[0] / 512.0, h[1] / 512.0,
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
