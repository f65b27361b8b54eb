// libssb200 C-ABI plumbing: error text, version, TMA tensor-map encoding (driver entry point
// resolved at run time so the library links without libcuda and loads on GPU-less hosts).
#include <cudaTypedefs.h>
#include <stdarg.h>
#include <string.h>

#include "../../include/ssb200.h"
#include "common.cuh"

namespace ssb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int n[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (!n[dev]) {
    if (cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n[dev] <= 0) n[dev] = 148;
  }
  return n[dev];
}

int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("%s: %s", what, cudaGetErrorString(e));
    return -2;
  }
  return 0;
}

int make_tmap_bf16(CUtensorMap* out, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box, bool swizzle128) {
  static PFN_cuTensorMapEncodeTiled_v12000 encode = nullptr;
  if (!encode) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      set_error("cuTensorMapEncodeTiled unavailable (%s)", cudaGetErrorString(e));
      return -3;
    }
    encode = (PFN_cuTensorMapEncodeTiled_v12000)fn;
  }
  cuuint64_t gdim[5], gstr[5];
  cuuint32_t bdim[5], estr[5];
  for (int i = 0; i < rank; ++i) {
    gdim[i] = dims[i];
    bdim[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  CUresult r = encode(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), gdim, gstr, bdim, estr,
                      CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                      CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed: CUresult %d (rank %d dims %llu,%llu box %u,%u)", (int)r, rank,
              (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0), box[0], rank > 1 ? box[1] : 0);
    return -3;
  }
  return 0;
}

}  // namespace ssb

extern "C" {

const char* ss_last_error(void) { return ssb::g_err; }

// Stream-ordered wait until *addr >= value (cuStreamWaitValue32, GEQ): no SM is occupied while waiting.
int ss_stream_wait_value32(void* stream, unsigned int* addr, unsigned int value) {
  typedef CUresult (*wait_fn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
  static wait_fn wait = nullptr;
  if (!wait) {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    cudaError_t e = cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qres);
    if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) {
      ssb::set_error("cuStreamWaitValue32 unavailable (%s)", cudaGetErrorString(e));
      return -3;
    }
    wait = (wait_fn)fn;
  }
  CUresult r = wait((CUstream)stream, (CUdeviceptr)addr, value, CU_STREAM_WAIT_VALUE_GEQ);
  if (r != CUDA_SUCCESS) {
    ssb::set_error("cuStreamWaitValue32 failed: CUresult %d", (int)r);
    return -3;
  }
  return 0;
}

int ss_version(void) { return SSB200_VERSION; }

int ss_device_check(void) {
  int dev = 0;
  cudaDeviceProp p;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaGetDeviceProperties(&p, dev) != cudaSuccess) {
    ssb::set_error("no CUDA device: %s", cudaGetErrorString(cudaGetLastError()));
    return -2;
  }
  if (p.major != 10) {
    ssb::set_error("libssb200 is built for sm_100a only; device is sm_%d%d", p.major, p.minor);
    return -4;
  }
  return p.multiProcessorCount;
}

}  // extern "C"
