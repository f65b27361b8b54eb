// Microbenchmark: cp.async.bulk (global -> shared, L2-resident source shared by all CTAs) as a function of the copy
// size and of the number of copies in flight: is the ~280 cycles per 16 KB copy a bandwidth or a per-copy cost?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o bulk_size bulk_size.cu
#include <cstdio>
#include <cstdlib>
#include "../../speech-separation_b200/csrc/common.cuh"
using namespace ssb;

__global__ void __launch_bounds__(128, 1) k(const unsigned char* src, int chunk, int ns, int nchunks, int rounds, int nprod, long long* out) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = (unsigned char*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[32], empty[32];
  if (threadIdx.x == 0) {
    for (int s = 0; s < ns; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    fence_barrier_init();
  }
  __syncthreads();
  if ((threadIdx.x == 0 || threadIdx.x == 64) && (int)(threadIdx.x >> 6) < nprod) {   // producer(s): alternate copies
    int it = 0;
    const int me = threadIdx.x >> 6;
    for (int r = 0; r < rounds; ++r)
      for (int c = 0; c < nchunks; ++c, ++it) {
        const int st = it % ns, ph = (it / ns) & 1;
        if (nprod == 2 && (it & 1) != me) continue;
        mbar_wait(&empty[st], ph ^ 1);
        mbar_arrive_expect_tx(&full[st], chunk);
        bulk_load(smem + st * chunk, src + (long long)c * chunk, chunk, &full[st]);
      }
  } else if (threadIdx.x == 32) {                              // consumer
    int it = 0;
    long long t0 = clock64();
    for (int r = 0; r < rounds; ++r)
      for (int c = 0; c < nchunks; ++c, ++it) {
        const int st = it % ns, ph = (it / ns) & 1;
        mbar_wait(&full[st], ph);
        mbar_arrive(&empty[st]);
      }
    out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  unsigned char* d;
  const size_t total = 160 * 1024;
  cudaMalloc(&d, total);
  cudaMemset(d, 0, total);
  long long* out; cudaMalloc(&out, 160 * 8);
  const int smem = 200 * 1024 + 1024;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int rounds = 100;
  for (int nprod : {1, 2})
  for (int ctas : {100})
    for (int chunk : {4096, 8192, 16384, 32768})
      for (int ns : {2, 4, 8}) {
        if ((long long)chunk * ns > 192 * 1024) continue;
        const int nchunks = (int)(total / chunk);
        long long h[160];
        for (int rep = 0; rep < 2; ++rep) {
          k<<<ctas, 128, smem>>>(d, chunk, ns, nchunks, rounds, nprod, out);
          cudaDeviceSynchronize();
        }
        cudaMemcpy(h, out, ctas * 8, cudaMemcpyDeviceToHost);
        double av = 0;
        for (int i = 0; i < ctas; ++i) av += h[i];
        av /= ctas;
        const double per = av / ((double)nchunks * rounds);
        printf("producers=%d ctas=%3d chunk=%6d in_flight=%2d: %7.0f cyc/copy  %6.1f B/clk/SM   160 KB in %6.0f cyc (%s)\n", nprod, ctas, chunk, ns, per,
               chunk / per, per * nchunks, cudaGetErrorString(cudaGetLastError()));
      }
  return 0;
}
