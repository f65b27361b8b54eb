/* libssb200 -- C ABI of the B200-native hot path of mmaciej2/speech-separation.
 *
 * The reference has NO C/FFI boundary: its boundary is the Python arch-module protocol
 * (archs/uPIT.py, archs/RSH.py) plus two CLIs (steps/extract_feats.py, steps/reconstruct_sources.py), and all
 * of its device arithmetic is implicit torch / librosa library calls.  Each entry point below therefore cites
 * the reference call site(s) whose arithmetic it replaces.  The Python host code in
 * speech-separation_b200/ssb200/ binds these with ctypes.
 *
 * Conventions: every pointer is a DEVICE pointer owned by the caller (workspaces included); `stream` is a
 * cudaStream_t passed as void*; functions return 0 on success, <0 on error with text in ss_last_error();
 * nothing synchronises the device.  Activations are time-major padded: row index = t*B + b.
 *
 * "sb16" = split-bf16 matrix: two bf16 planes (hi at base, lo at base + plane elements), x ~= hi + lo, row
 * pitch in elements (multiple of 8).  Tensor-core products use hi*hi + hi*lo + lo*hi (fp32-class accuracy).
 */
#ifndef SSB200_H
#define SSB200_H

#ifdef __cplusplus
extern "C" {
#endif

#define SSB200_VERSION 3

/* ---- library ---- */
const char* ss_last_error(void);
int ss_version(void);
int ss_device_check(void); /* >0: SM count of the current sm_100 device; <0: error */

/* ---- K1: STFT features.  Replaces librosa.core.stft (+np.abs) at steps/extract_feats.py:76-89,98,105 ----
 * Ragged corpus: utterance u owns samples [sig_off[u], sig_off[u+1]) of `sig`, frames
 * [frame_off[u], frame_off[u+1]) with T_u = 1 + N_u/hop, and CTA tiles [tile_off[u], tile_off[u+1]) with
 * ceil(T_u / ss_stft_tile_frames()) tiles.  Output block of utterance u starts at element frame_off[u]*F
 * (F = n_fft/2+1; float for magnitude, float2 for complex) and is (F, T_u) row-major -- the layout of the
 * reference's .npz arrays -- or (T_u, F) with SS_STFT_LAYOUT_TF.  N_u must exceed n_fft/2 (reflect padding).
 * Any even n_fft and any hop <= n_fft (the reference passes --fft-dim / --step-size straight to librosa): n_fft 256 / 512
 * with an even hop run the packed-FFT kernel, every other size a direct-summation kernel (correct, O(n_fft^2)). */
#define SS_STFT_COMPLEX 1   /* keep complex64 (extract_feats.py test branch :98,105) instead of |X| (:87,89) */
#define SS_STFT_LAYOUT_TF 2 /* (T,F) per utterance instead of the file layout (F,T) */
#define SS_STFT_LOG 4       /* log(max(|X|,1e-8)); NOT reference behaviour, off for parity */
int ss_stft_tile_frames(void);
int ss_stft(const float* sig, const long long* sig_off, const long long* frame_off, const int* tile_off, int n_utts,
            int total_tiles, int n_fft, int hop, int flags, void* out, void* stream);

/* Per-utterance, per-bin mean-variance normalisation of the (log-)magnitude blocks ss_stft wrote, in place:
 * x <- (x - mean_t) / sqrt(var_t + eps) over the T_u frames of each (utterance, bin) row (biased variance).  An optional
 * epilogue of K1 (north_star); the reference stores plain magnitudes, so it is off on every parity path. */
int ss_cmvn(float* feat, const long long* frame_off, int n_utts, int F, int layout_tf, float eps, void* stream);

/* K1 on the tensor cores (csrc/stft_mma.cu): the same magnitude STFT as ss_stft(n_fft = 256, hop = 64, flags = 0 or
 * SS_STFT_LOG), replacing librosa.core.stft + np.abs at steps/extract_feats.py:87,89, evaluated as a split-bf16 tcgen05
 * GEMM of the frames against a resident windowed DFT basis.  hop = n_fft/4 makes frame r the four 64-sample rows
 * r..r+3 of the reflect-padded signal, so the operand descriptors read frames straight out of ONE bf16 hi/lo copy of the
 * samples.  Work is the ROW STREAM of all utterances: utterance u owns T_u + 3 rows; utts[u] = four int64
 * { first sample in sig, first output element (frame_off[u]*129), N_u | T_u << 32, first stream row };
 * tile_u0[i] = utterance of stream row i * ss_stft_dft256_tile_frames(); total_rows = sum (T_u + 3).
 * Output as ss_stft: (129, T_u) row-major per utterance.  Utterances whose first sample is 16-byte aligned take the
 * vector load path (same results). */
int ss_stft_dft256_tile_frames(void);
int ss_stft_dft256(const float* sig, const long long* utts, int n_utts, const int* tile_u0, int n_tiles,
                   long long total_rows, float* out, int flags, void* stream);

/* ---- K7: masked iSTFT.  Replaces steps/reconstruct_sources.py:39-42 (np.multiply, librosa.core.istft,
 * *32767, astype('int16')) ----
 * spec: complex64 mixture STFT blocks as written by ss_stft(COMPLEX); mask: for utterance u, n_src blocks of
 * the same (F,T_u) shape starting at element (frame_off[u]*n_src)*F.  Output of (u, s) starts at
 * wav_off[u] + s*L_u, L_u = hop*(T_u-1) samples (int16, or float before the *32767 with SS_ISTFT_OUT_F32).
 * tile_off counts ceil(L_u / (ss_istft_tile_hops()*hop)) tiles per utterance.  Any even n_fft = 2(F-1) and hop <= n_fft;
 * the packed-FFT kernel covers n_fft 256 / 512 with an even hop >= n_fft/4, a direct-summation kernel the rest. */
#define SS_ISTFT_LAYOUT_TF 1
#define SS_ISTFT_OUT_F32 2
int ss_istft_tile_hops(void);
int ss_istft_masked(const void* spec, const float* mask, const long long* frame_off, const long long* wav_off,
                    const int* tile_off, int n_utts, int n_src, int total_tiles, int n_fft, int hop, int flags,
                    void* out, void* stream);

/* ---- tensor-core GEMM on sb16 operands.  Replaces the cuBLAS GEMMs behind nn.LSTM's input projection and
 * nn.Linear (archs/uPIT.py:115-117,132,141; archs/RSH.py:155-157,172,181) and their autograd backward ----
 * C[rowmap(m)][n] = alpha * sum_k A[m][k]*B[n][k] + bias[n] + beta*C[..][n];  A: M x K, B: N x K (both K-major). */
#define SS_ROWMAP_NONE 0
#define SS_ROWMAP_UNIT4 1 /* m = 4*u + gate  ->  output row gate*hidden + u (undo the gate interleave) */
int ss_gemm_sb16(const void* A, long long lda, long long a_plane, const void* B, long long ldb, long long b_plane,
                 float* C, long long ldc, int M, int N, int K, const float* bias, float alpha, float beta, int rowmap,
                 int hidden, int sm_limit, void* stream);

/* Two problems of the same shape in one launch: entry z in {0,1} reads A rows [z*a_batch_rows, +M) and B rows
 * [z*b_batch_rows, +N) of the same sb16 matrices and writes Cz (bias[n] is shared).
 * sm_limit > 0 (all ss_gemm_* entry points) caps the grid to that many CTAs: the GEMM then runs beside a resident
 * recurrence kernel on the SMs that kernel leaves idle; 0 = every SM. */
int ss_gemm_sb16_x2(const void* A, long long lda, long long a_plane, int a_batch_rows, const void* B, long long ldb,
                    long long b_plane, int b_batch_rows, float* C0, float* C1, long long ldc, int M, int N, int K,
                    const float* bias, float alpha, float beta, int rowmap, int hidden, int sm_limit, void* stream);

/* Weight-gradient form ("TN"): both operands are stored with the REDUCTION index r = t*B + b as their row, exactly as
 * the forward / backward kernels write activations and gate gradients -- no transposed copy exists anywhere.
 *   Cz[rowmap(m)][n] = alpha * sum_r A[r][a_col0 + z*a_col_batch + m] * B[r + b_shift_z][b_col0 + z*b_col_batch + n]
 *                      + beta * Cz[..][n]
 * with r running over the n_blk blocks of 64 rows that start at block blk0_z (per entry z; batch = 1 or 2), so a sum
 * over all R rows can be accumulated chunk by chunk (beta = 1) while a recurrence kernel is still producing rows.
 * A: sb16 [R][a_cols] (pitch lda), B: sb16 [R][b_cols] (pitch ldb); rows outside [0, R) read as zero, which is
 * what makes b_shift = -/+B express h_prev[t] = out[t -/+ 1] for dW_hh (the h0 term: ss_dwhh_edge).
 * Replaces the autograd backward GEMMs of nn.LSTM / nn.Linear weights (steps/train_qsub.py:120). */
int ss_gemm_sb16_tn(const void* A, long long lda, long long a_plane, int a_cols, int a_col0, int a_col_batch,
                    const void* B, long long ldb, long long b_plane, int b_cols, int b_col0, int b_col_batch, int R,
                    int blk0_z0, int blk0_z1, int n_blk, int b_shift_z0, int b_shift_z1, int batch, float* C0, float* C1,
                    long long ldc, int M, int N, float alpha, float beta, int rowmap, int hidden, int sm_limit,
                    void* stream);
/* dW_hh[z][gate*H+u][j] += sum_b dG[row_z(b)][z*4H + 4u+gate] * h0[z][b][j], row_0(b) = b, row_1(b) = (len_b-1)*B + b:
 * the first frame each direction processes carries h0, which no row of `out` holds. */
int ss_dwhh_edge(const void* dG, long long ld, long long plane, const int* lens, const float* h0, int B, int H,
                 float* dW0, float* dW1, long long ldw, void* stream);

/* ---- layout / precision conversion kernels (plumbing around the GEMMs) ----
 * ss_split: fp32 in[R][C] (pitch ld_in) -> sb16 out[rp(r)][C] (pitch ld_out, plane out_plane) and/or the
 * transposed sb16 outT[C][rp(r)] (pitch ld_outT).  rowperm = SS_ROWMAP_UNIT4 interleaves LSTM gate rows:
 * natural row gate*hidden+u -> 4*u+gate.  Either output may be NULL.  scale[C] (optional) multiplies column c. */
int ss_split(const float* in, long long ld_in, int R, int C, int rowperm, int hidden, const float* col_scale,
             void* out, long long ld_out, long long out_plane, void* outT, long long ld_outT, long long outT_plane,
             void* stream);
/* column sums of an sb16 matrix: out[c] = sum_r (hi+lo)[r][c]; partial: >= ss_colsum_parts(R)*C floats. */
int ss_colsum_parts(int R);
int ss_colsum_sb16(const void* in, long long ld, long long plane, int R, int C, float* partial, float* out,
                   int rowmap, int hidden, void* stream);

/* ---- K5/K6: fused mask*mix, pairwise squared errors, permutation / greedy assignment, loss; and backward ----
 * Replaces archs/uPIT.py:178-206 and archs/RSH.py:224-241.  z: head pre-activations [T*B][ldz] (n_out*F columns,
 * output slot i in columns [i*F,(i+1)*F)); mix [T*B][F]; src [n_src][T*B][F]; lens[B] frames per utterance.
 * fwd writes partial[ss_pit_chunks(T)][B][n_out*n_src] (+ optionally mask = sigmoid(z) [T*B][n_out*F]).
 * select, mode 0 (uPIT.py:186-194): losses[P][B] in itertools.permutations order, idx[B] = argmin, assign[B][n_out]
 * = chosen source per output slot; mode 1 (RSH.py:229-238): losses[n_src][B] with +inf for sources in the
 * bitmask used[B], idx = argmin, used updated.  stats[4] = {sum_b min/S, sum_b len_b*F, loss = [0]/[1], 1/(S*[1])},
 * accumulated over calls when accumulate != 0 (RSH.py:240-241).
 * bwd (gout NULL = 1): dz = (2*(m*x - src[assign])*x*stats[3]*gout[0] + dmask_extra) * m*(1-m), as sb16 [T*B][ld_dz];
 * rows with t >= len_b are zero. */
int ss_pit_chunks(int T);
int ss_pit_loss_fwd(const float* z, long long ldz, const float* mix, const float* src, const int* lens, int T, int B,
                    int F, int n_out, int n_src, float* mask_out, float* partial, void* stream);
int ss_pit_select(const float* partial, int chunks, const int* lens, int B, int F, int n_out, int n_src, int mode,
                  int accumulate, float* losses, int* idx, int* assign, int* used, float* stats, void* stream);
int ss_pit_loss_bwd(const float* z, long long ldz, const float* mix, const float* src, const int* lens,
                    const int* assign, int T, int B, int F, int n_out, const float* stats, const float* gout,
                    const float* dmask_extra, void* dz, long long ld_dz, long long dz_plane, void* stream);

/* out[R][C] = sigmoid(z[R][ldz]) -- F.sigmoid of the head output (archs/uPIT.py:144) */
int ss_sigmoid(const float* z, long long ldz, int R, int C, float* out, void* stream);

/* y = a*x + b*y (b == 0: y need not be initialised) */
int ss_axpby(const float* x, float* y, long long n, float a, float b, void* stream);

/* out[blk*4H + 4u+gate] = b_ih[blk*4H + gate*H+u] + b_hh[same]: interleaved bias of the x-projection GEMM. */
int ss_bias_prep(const float* b_ih, const float* b_hh, int n, int hidden, float* out, void* stream);

/* PackedSequence data [sum_t n_t][c_src] (columns c_off..c_off+C) -> time-major padded [T][B][C], zeros where
 * absent; toff[T+1] = prefix sums of batch_sizes.  Replaces pad_packed_sequence (archs/uPIT.py:135,165,178). */
int ss_unpack(const float* packed, const long long* toff, int T, int B, int C, int c_src, int c_off, float* out,
              void* stream);

/* On-device hand-offs (no .npz round trip between the stages).
 * ss_ragged_to_padded: the (T_u, F) blocks ss_stft(LAYOUT_TF) writes -> the model's time-major padded input
 * [T][B][F] (zeros past T_u); is_complex takes |X| of complex64 blocks on the way (np.abs, archs/uPIT.py:91).
 * Utterance u of the ragged corpus becomes batch column u, so the corpus must be sorted by descending length
 * like the collator's batches (archs/uPIT.py:40).
 * ss_masks_to_ragged: head pre-activations z [T*B][ldz] -> sigmoid (archs/uPIT.py:144) -> the per-utterance,
 * per-speaker (T_u, F) mask blocks ss_istft_masked(LAYOUT_TF) reads (archs/uPIT.py:220-225 without the files). */
int ss_ragged_to_padded(const void* in, int is_complex, const long long* frame_off, int T, int B, int F, float* out,
                        void* stream);
int ss_masks_to_ragged(const float* z, long long ldz, const long long* frame_off, int T, int B, int F, int n_src,
                       float* out, void* stream);

/* ---- K2b/K2c: persistent bidirectional LSTM recurrence.  Replaces the sequential part of the cuDNN LSTM
 * behind nn.LSTM(..., bidirectional=True) at archs/uPIT.py:115,132 / archs/RSH.py:155,172 (PackedSequence
 * semantics via lens[], descending) and of its backward (steps/train_qsub.py:120) ----
 * Gate columns are interleaved: column dir*4H + 4*u + gate (gate order i,f,g,o as torch).
 * fwd: gx [T*B][8H] = x W_ih^T + b_ih + b_hh (both directions) is overwritten with the activated gates (training,
 *      i.e. cs != NULL; without cs the contents of gx after the call are unspecified: evaluation keeps neither);
 *      whh_r sb16 [8H][H] (rows dir*4H+4u+gate); h0,c0,hn,cn [2][B][H]; out sb16 [T*B][2H] (zero at t>=len_b);
 *      cs [T*B][2H] cell states (training; may be NULL).  ws: ss_lstm_ws_bytes(B,H) bytes.
 * bwd: dout [T*B][2H] gradient wrt out; whhT sb16 [2H][4H] (row dir*H+j, column 4u+gate);
 *      dG sb16 [T*B][8H] gate pre-activation gradients (zero at t>=len_b);
 *      dhn,dcn (optional) gradient wrt (hn,cn); dh0,dc0 (optional) gradient wrt (h0,c0).
 *      batch_pitch (0 = B; fwd with H = 600 and bwd): the pointers may address a SUB-BATCH of a wider padded batch -- rows b0..b0+B of every frame,
 *      every pointer advanced to its row / entry b0 by the caller, batch_pitch = rows per frame of the whole batch (and
 *      entries per direction of c0 / dhn / dcn / dh0 / dc0).  Batch rows are independent chains, so a batch of
 *      64 < B <= 128 is run as two halves through the tensor-memory kernels (H = 600) instead of the slower generic ones.
 * progress (optional, both): two counters, one per direction; while it runs the kernel stores
 *      progress[d] = prog_base + (number of steps of direction d whose out / dG rows are globally visible), every few
 *      steps (never the last ones: kernel completion covers those).  Direction 0 forward and direction 1 backward walk
 *      t = 0..T-1, the other two T-1..0.  Work on another stream gated with ss_stream_wait_value32 (cyclic >=, so
 *      prog_base may grow from launch to launch without ever being reset) can consume finished frames while the
 *      recurrence is still running: engine.py streams the next GEMM through the SMs the recurrence leaves idle. */
long long ss_lstm_ws_bytes(int B, int H);
int ss_lstm_fwd(float* gx, const void* whh_r, long long whh_ld, long long whh_plane, const float* h0, const float* c0,
                const int* lens, void* out, long long out_plane, float* cs, float* hn, float* cn, void* ws,
                unsigned int* progress, unsigned int prog_base, int T, int B, int batch_pitch, int H, void* stream);
int ss_lstm_bwd(const float* gates, const float* cs, const float* c0, const float* dout, const float* dhn,
                const float* dcn, const int* lens, const void* whhT, long long whhT_ld, long long whhT_plane, void* dG,
                long long dG_plane, float* dh0, float* dc0, void* ws, unsigned int* progress, unsigned int prog_base,
                int T, int B, int batch_pitch, int H, void* stream);
/* Stream-ordered wait until (int)(*addr - value) >= 0 (cuStreamWaitValue32, GEQ): no SM is occupied while waiting. */
int ss_stream_wait_value32(void* stream, unsigned int* addr, unsigned int value);
/* ---- K3/K4 glue: BatchNorm1d(2H) over all T*B rows (padding included, archs/uPIT.py:119,138) folded into the
 * head Linear (archs/uPIT.py:117,141), and its backward ----
 * coef[5][C] = {s = gamma*rstd, shift = beta - mean*s, mean, rstd, biased var}; training updates the running
 * stats (momentum 0.1, unbiased var), eval reads them.  partial: 2*ss_bn_parts(R)*C floats.
 * ss_head_fold_bias: bout[o] = b[o] + sum_c W[o][c]*shift[c]  (the folded weight is ss_split(W, col_scale=s)).
 * ss_bn_bwd: dy [R][C] (gradient wrt the BN output) is replaced in place by the gradient wrt x; also dgamma,
 * dbeta; bwd: 2*C floats scratch.  ss_head_dw_fix: dW = s*(G - mean*dsum) + beta*dsum with G = dz^T x. */
int ss_bn_parts(int R);
int ss_bn_stats(const void* x, long long ld, long long plane, int R, int C, const float* gamma, const float* beta,
                float eps, float momentum, int training, float* running_mean, float* running_var, float* partial,
                float* coef, void* stream);
int ss_head_fold_bias(const float* W, const float* b, const float* coef, int O, int C, float* bout, void* stream);
int ss_bn_bwd(float* dy, const void* x, long long ld, long long plane, int R, int C, const float* coef, int training,
              float* partial, float* dgamma, float* dbeta, float* bwd, void* stream);
int ss_head_dw_fix(const float* G, const float* dsum, const float* coef, const float* beta, int O, int C, float* dW,
                   void* stream);
/* Data-parallel mode B (N ranks reproduce the reference's arithmetic on the concatenated batch): the
 * BatchNorm statistics are taken over the rows of ALL ranks.  ss_bn_sums writes sums[2][C] = column sums of {x, x^2}
 * (dy == NULL) or {dy, dy*x}; the caller sum-allreduces them, and counts[2] = {total batch size, longest T of any rank}
 * (their product is the row count of the concatenated, zero-padded batch); then ss_bn_stats_from_sums gives coef and
 * updates the running statistics exactly as ss_bn_stats(training) would on the big batch, and ss_bn_bwd_from_sums
 * applies the backward with the global sums while dgamma / dbeta stay this rank's share (the gradient exchange adds
 * them up).  ss_pit_stats_finalize recomputes stats[2] = loss and stats[3] = 1/(S*norm) after stats[0:2] were
 * sum-allreduced (archs/uPIT.py:204: one normaliser for the whole batch). */
int ss_bn_sums(const void* x, long long ld, long long plane, const float* dy, int R, int C, float* partial, float* sums,
               void* stream);
int ss_bn_stats_from_sums(const float* sums, const float* counts, int C, const float* gamma, const float* beta, float eps,
                          float momentum, float* running_mean, float* running_var, float* coef, void* stream);
int ss_bn_bwd_from_sums(float* dy, const void* x, long long ld, long long plane, int R, int C, const float* coef,
                        const float* sums_local, const float* sums_global, const float* counts, float* dgamma,
                        float* dbeta, float* bwd, void* stream);
int ss_pit_stats_finalize(float* stats, int n_src, void* stream);

/* ---- RSH residual attention.  Replaces archs/RSH.py:254-257 (compute_loss: combos = relu_(combos - [0 || mask]),
 * repack) and :277-281 (compute_masks: no relu) and the autograd backward of that path ----
 * cur,next: pass inputs [T*B][2F] = [mix || attention]; z: head pre-activations of the pass (mask = sigmoid(z)).
 * Rows with t >= len_b of `next` are zero.  bwd: dx = gradient wrt the NEXT pass's input [R][2F], next_in = that
 * input; carry_out[R][F] = (dx[:,F:] + carry_in) * [next_in[:,F:] > 0] is the total gradient reaching this pass's
 * attention half through the residual, dmask = -carry_out is the extra gradient wrt this pass's mask
 * (ss_pit_loss_bwd's dmask_extra).  carry_in may be NULL. */
int ss_rsh_residual(const float* cur, const float* z, long long ldz, const int* lens, int T, int B, int F, int relu,
                    float* next, void* stream);
int ss_rsh_residual_bwd(const float* dx, const float* carry_in, const float* next_in, long long R, int F,
                        float* carry_out, float* dmask, void* stream);

/* ---- f4: oracle-mask evaluation.  Replaces steps/evaluate_oracle.py:133-143 (ideal ratio / ideal binary masks applied
 * to the mixture STFT; the reconstruction itself is ss_istft_masked) and the inner products of the SDR-type scores
 * computed after it (steps/evaluate_oracle.py:144, steps/evaluate_sources.py:57) ----
 * ss_oracle_masks: spec = complex64 mixture STFT blocks (ss_stft COMPLEX, (F,T_u) per utterance), src_mag = n_src
 * magnitude STFTs of the sources in the same ragged layout, source s at src_mag + s*src_stride.  masks: the block layout
 * ss_istft_masked reads (utterance u, source s at element (frame_off[u]*n_src + s*T_u)*F).
 * soft (hard = 0): |S_s| / |X| (np.divide); bins where |X| is EXACTLY zero get mask 0 (np.divide would give inf / nan there
 * and poison the frame -- which bins are exactly zero is an accident of the FFT arithmetic);
 * hard: 1 where s == argmax_j |S_j| (first maximum wins like np.argmax) else 0.
 * ss_sdr_sums: item i compares ref[ref_off[i] .. +n_ref[i]) with est[est_off[i] .. +n_est[i]) (zero beyond n_est, the
 * unfilled tail of evaluate_oracle.py:142-143): sums[i][5] = {sum r, sum e, sum r^2, sum e^2, sum r*e} in fp64. */
int ss_oracle_masks(const void* spec, const float* src_mag, long long src_stride, const long long* frame_off, int n_utts,
                    int max_frames, int F, int n_src, int hard, float* masks, void* stream);
int ss_sdr_sums(const float* ref, const long long* ref_off, const float* est, const long long* est_off, const int* n_ref,
                const int* n_est, int n_items, int max_ref, double* sums, void* stream);
/* BSS Eval (SDR / SIR / SAR with a flen-tap time-invariant distortion filter: what mir_eval.separation.bss_eval_sources
 * computes at steps/evaluate_oracle.py:144 and steps/evaluate_sources.py:57; restated in oracle/eval_oracle.py, parity
 * unpinned -- mir_eval is absent).  The O(N * flen) parts run on the device, the (n_src*flen)^2 solves on the host:
 * ss_xcorr: out[p][l] = sum_m a_p[m + lag0 + l] * b_p[m], l < nlag, fp64 (the Toeplitz Gram blocks of the delayed
 *   references and the right-hand sides <estimate, reference delayed by l>); pair p reads a + a_off[p] (a_len[p] samples)
 *   and b + b_off[p]; signals are zero outside their lengths.
 * ss_bss_proj_sums: item = (recording, estimate j), n_src consecutive items per recording; ref_off[rec*n_src + s] locates
 *   reference s of the recording; c1[item][flen] = filter on reference j, c2[item][n_src][flen] = filters on all
 *   references; with p1, p2 the two projections over n < N + flen - 1,
 *   sums[item] = { sum p1^2, sum (e-p1)^2, sum (p2-p1)^2, sum p2^2, sum (e-p2)^2 }: SDR = [0]/[1], SIR = [0]/[2], SAR = [3]/[4]. */
int ss_xcorr(const float* a, const long long* a_off, const int* a_len, const float* b, const long long* b_off,
             const int* b_len, int n_pairs, int max_b_len, int lag0, int nlag, double* out, void* stream);
int ss_bss_proj_sums(const float* ref, const long long* ref_off, const float* est, const long long* est_off, const int* n_ref,
                     const int* n_est, const double* c1, const double* c2, int n_items, int n_src, int flen, int max_ref,
                     double* sums, void* stream);

/* ---- f1: global-norm clip + Adam over one flat fp32 bucket.  Replaces torch.nn.utils.clip_grad_norm_(params,
 * 0.25) and torch.optim.Adam(lr=1e-3).step() at steps/train_qsub.py:95,121-122 ----
 * ss_grad_norm: norm_out[0] = gscale * ||g||_2 (partial: ss_optim_parts(n) floats).  gscale = 1/world turns an
 * allreduce-summed gradient into the data-parallel mean.
 * ss_clip_adam: g' = g * gscale * min(1, max_norm/(norm[0]+1e-6)) (max_norm <= 0: no clipping, norm unused), then
 * Adam with bias correction (step counts from 1): m += (g'-m)(1-b1); v = b2 v + (1-b2) g'^2;
 * p -= lr/(1-b1^step) * m / (sqrt(v)/sqrt(1-b2^step) + eps). */
int ss_optim_parts(long long n);
int ss_grad_norm(const float* g, long long n, float gscale, float* partial, float* norm_out, void* stream);
int ss_clip_adam(float* p, const float* g, float* m, float* v, long long n, const float* norm, float gscale,
                 float max_norm, float lr, float beta1, float beta2, float eps, int step, void* stream);

/* ---- ss_grad_norm for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same result as ss_grad_norm (norm_out[0] = gscale * ||g||_2), but the per-block partial sums live in a static
 * __device__ array inside the library (32 KB, held by every GPU that loads the library until exit) instead of a
 * caller workspace; the call performs no cudaMalloc / cudaFree / allocating library routine.
 * Limit: n <= ss_grad_norm_noalloc_max_n() = SS_GRAD_NORM_NOALLOC_MAX_N elements; larger n returns -1 with a message
 * in ss_last_error() (use ss_grad_norm).  -3: the scratch symbol could not be resolved on the current device.
 * Every call on a device shares the array (it keeps the previous call's partials; each call overwrites all it reads,
 * nothing needs clearing): calls that OVERLAP on the GPU -- different streams without ordering -- race. */
#define SS_GRAD_NORM_NOALLOC_MAX_N (64LL << 20)
long long ss_grad_norm_noalloc_max_n(void);
int ss_grad_norm_noalloc(const float* g, long long n, float gscale, float* norm_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SSB200_H */
