// Parameters shared by the two backward-recurrence kernels (lstm_bwd.cu: generic; lstm_bwd_ks.cu: K-split clusters).
#pragma once
#include "common.cuh"

namespace ssb {

struct LstmBwdParams {
  const float* gates;        // [T*B][8H] activated gates saved by the forward kernel
  const float* cs;           // [T*B][2H] cell states
  const float* c0;           // [2][B][H]
  const float* dout;         // [T*B][2H] gradient wrt the layer output
  const float* dhn;          // [2][B][H] or null: gradient wrt (h_n, c_n)
  const float* dcn;
  const int* lens;
  __nv_bfloat16* dG;         // sb16 [T*B][8H]
  long long dG_plane;
  float* dh0;                // [2][B][H] or null
  float* dc0;
  unsigned int* sync;
  unsigned int* progress;    // [2] or null: prog_base + number of steps whose dG rows are globally visible, per direction
  unsigned int prog_base;
  int T, B, H, R, NC, KC, NS, want_d0, ROT;
  int Bp;                    // rows per frame of the padded batch the pointers walk (K-split kernel; >= B)
  long long* dbg;            // optional clock64 trace of CTA 0: [step][32]
};

extern long long* g_lstm_dbg;   // set by ss_lstm_debug (lstm_fwd.cu)

// K-split cluster variant (H = 600, B <= 64); returns 0 / negative error like the C ABI.
int launch_lstm_bwd_ks(const LstmBwdParams& p, const void* whhT, long long whhT_ld, long long whhT_plane, cudaStream_t st);

}  // namespace ssb
