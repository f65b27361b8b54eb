// Declarations of the forward recurrence kernel (lstm_fwd.cu: generic + constant-shape fast path).
#pragma once
#include "common.cuh"

namespace ssb {

constexpr int LF_U = 12;                    // hidden units per CTA
constexpr int LF_N = 4 * LF_U;              // gate columns per CTA
constexpr int LF_EPI_WARPS = 12;            // 4 TMEM lane quadrants x 3 column groups of 4 units
constexpr int LF_LOAD_WARPS = 8;            // epilogue warps 0..7 also check the landed h_prev chunks
constexpr int LF_THREADS = 32 * (LF_EPI_WARPS + 3);   // + warp 12 (gate + copies) + warp 13 (MMA issuer) + warp 14 (progress)
constexpr int LF_WCHUNK = 2 * LF_N * 128;   // one 64-wide K chunk of the W slice: [hi 48 rows][lo 48 rows] x 128 B
constexpr int LF_MAXKC = 10;                // H <= 640
constexpr int LF_XPITCH = 52;               // floats per row of the lo-row exchange buffer
constexpr int LF_TMEM = 128;
constexpr int LF_STAGE_TILES = 2 * 128 * 32 + 64 * 12 * 4 + 64 * 4;   // staging tiles of the coalesced stores

struct LstmFwdParams {
  float* gx;                 // [T*B][8H] in: x-projection + biases (interleaved gate columns); out: activated gates
  const float* h0;           // [2][B][H]
  const float* c0;
  const int* lens;           // [B]
  __nv_bfloat16* out;        // sb16 [T*B][2H]
  long long out_plane;
  float* cs;                 // [T*B][2H] or null
  float* hn;                 // [2][B][H]
  float* cn;
  __nv_bfloat16* hbuf;       // exchange buffer [slot 3][dir 2][kc][2R operand rows][64], rows pre-swizzled (SWIZZLE_128B)
  unsigned int* done;        // [2][64] per direction and CTA: reporting periods (4 steps) whose `out` rows it has made visible
  unsigned int* progress;    // [2] or null: prog_base + number of steps whose `out` rows are globally visible, per direction
  unsigned int prog_base;
  int T, B, H, R, NC, KC, NS;
  int Bp;                    // rows per frame of the padded batch the pointers walk (lstm_fwd_tw.cu; >= B)
  long long* dbg;            // optional clock64 trace of CTA 0: [step][32]
};

// lstm_fwd_tw.cu: H = 600, B <= 64 with W_hh resident in tensor memory
int launch_lstm_fwd_tw(const LstmFwdParams& p, const void* whh_r, long long whh_ld, long long whh_plane, cudaStream_t st);

}  // namespace ssb
