// Microbenchmark: cp.async.bulk (global -> shared) rate per SM as a function of how many CTAs stream and whether
// they read the SAME region (the recurrence's h exchange pattern) or distinct ones.  Also the same loads followed
// by tcgen05.mma consumption (4 MMAs per 16 KB chunk) to see whether loads or MMA issue pace the pipeline.
#include <cstdio>
#include <cstdlib>
#include "../../speech-separation_b200/csrc/common.cuh"
using namespace ssb;

constexpr int CHUNK = 16384, NS = 5;

__global__ void __launch_bounds__(128, 1) k(const unsigned char* src, long long per_cta_stride, int nchunks, int rounds,
                                            int with_mma, int N, long long* out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  unsigned char* wsm = smem + NS * CHUNK;                      // 32 KB of "weights" for the MMA B operand
  __shared__ uint64_t full[NS], empty[NS], done;
  __shared__ uint32_t slot;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(&done, 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < 32 * 1024 / 4; i += 128) ((uint32_t*)wsm)[i] = 0x3c003c00u;
  if (threadIdx.x < 32) tmem_alloc<256>(&slot);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t td = slot;
  const unsigned char* my = src + (long long)blockIdx.x * per_cta_stride;
  if (threadIdx.x == 0 && !(with_mma & 16)) {                  // producer
    int it = 0;
    for (int r = 0; r < rounds; ++r)
      for (int c = 0; c < nchunks; ++c, ++it) {
        const int st = it % NS, ph = (it / NS) & 1;
        mbar_wait(&empty[st], ph ^ 1);
        if (with_mma & 2) { mbar_arrive(&full[st]); continue; }           // no load: handshake + MMA only
        mbar_arrive_expect_tx(&full[st], CHUNK);
        bulk_load(smem + st * CHUNK, my + (long long)c * CHUNK, CHUNK, &full[st]);
      }
  } else if (threadIdx.x == 32) {                              // consumer
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t a_base = umma_desc_sw128(smem_u32(smem)), b_base = umma_desc_sw128(smem_u32(wsm));
    int it = 0;
    long long t0 = clock64();
    for (int r = 0; r < rounds; ++r)
      for (int c = 0; c < nchunks; ++c, ++it) {
        const int st = it % NS, ph = (it / NS) & 1;
        if (!(with_mma & 16)) mbar_wait(&full[st], ph);
        if (with_mma) {
          if (!(with_mma & 8)) tc_fence_after();
          const uint64_t da = a_base + (uint32_t)st * (CHUNK >> 4);
          for (int kk = 0; kk < 4; ++kk) umma_bf16(td, da + 2 * kk, b_base + 2 * kk, idesc, (c | kk) ? 1u : 0u);
          if (with_mma & 32) {}
          else if (with_mma & 4) mbar_arrive(&empty[st]); else umma_commit(&empty[st]);   // 4: release without commit
        } else {
          mbar_arrive(&empty[st]);
        }
      }
    if (with_mma) { umma_commit(&done); mbar_wait(&done, 0); }
    long long t1 = clock64();
    out[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_free<256>(td); }
}

int main() {
  const int nchunks = 10, rounds = 200;
  unsigned char* d;
  const size_t per = (size_t)nchunks * CHUNK;
  cudaMalloc(&d, per * 160);
  cudaMemset(d, 0, per * 160);
  long long* out; cudaMalloc(&out, 160 * 8);
  const int smem = NS * CHUNK + 32 * 1024 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int with_mma : {1, 3, 5, 7, 15, 31}) for (int N : {96, 128}) for (int same : {1}) for (int ctas : {100}) {
    if (!with_mma && N != 96) continue;
    long long h[160];
    for (int rep = 0; rep < 2; ++rep) {
      k<<<ctas, 128, smem>>>(d, same ? 0 : (long long)per, nchunks, rounds, with_mma, N, out);
      cudaDeviceSynchronize();
    }
    cudaMemcpy(h, out, ctas * 8, cudaMemcpyDeviceToHost);
    double mx = 0, av = 0;
    for (int i = 0; i < ctas; ++i) { av += h[i]; if (h[i] > mx) mx = h[i]; }
    av /= ctas;
    printf("mma=%d N=%3d same=%d ctas=%3d: %.0f cyc/chunk avg (max %.0f) -> %.1f B/clk/SM  (%s)\n", with_mma, N, same, ctas,
           av / (nchunks * rounds), mx / (nchunks * rounds), CHUNK / (av / (nchunks * rounds)),
           cudaGetErrorString(cudaGetLastError()));
  }
  return 0;
}
