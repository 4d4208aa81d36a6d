/* ggpm_b200 — C ABI of the B200 (sm_100a) hierarchical message-passing encoder kernels.
 *
 * Drop-in boundary for ggpm's encoder hot path.  The reference has no FFI of its own (it is pure
 * PyTorch); these entry points are what a binding for that path has to call.  Each one cites the
 * reference code it replaces (paths relative to the reference checkout).  The Python host side
 * (ggpm_b200/_lib.py) binds them with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *  - plain pointers and sizes only; every pointer is a DEVICE pointer on the current CUDA device;
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); all work is
 *    enqueued on it, nothing synchronises, nothing is allocated: callers own every buffer;
 *  - float tables are fp32 row-major with a leading dimension (`ld*`, in floats) that is a multiple of
 *    4 and a 16-byte aligned base; the hidden width H is carried in a padded width Hp (multiple of 32,
 *    zero-filled columns H..Hp-1) so rows stay 16-byte aligned for 128-bit loads and tensor-core tiles;
 *  - neighbour lists are CSR (int32 row_ptr[rows+1], col[nnz]) over 1-based row ids with id 0 = padding;
 *  - return value: 0 = ok, non-zero = error; ggpm_last_error() returns the message (thread local).
 *  - no CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef GGPM_B200_H
#define GGPM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGPM_ACT_NONE 0
#define GGPM_ACT_RELU 1
#define GGPM_ACT_TANH 2
#define GGPM_ACT_SIGMOID 3

/* GEMM engines (ggpm_set_gemm_engine): AUTO picks f16x3 where it is wired (graph LSTM gate / state-gradient GEMMs), tcgen05 3xTF32 wherever else the shape allows */
#define GGPM_ENGINE_AUTO 0
#define GGPM_ENGINE_SIMT_FP32 1
#define GGPM_ENGINE_TCGEN05_3XTF32 2
#define GGPM_ENGINE_TCGEN05_F16X3 3   /* fp16 two-term split (kind::f16) where supported, 3xTF32 elsewhere */

const char* ggpm_last_error(void);
int ggpm_abi_version(void);
int ggpm_set_gemm_engine(int engine);
int ggpm_get_gemm_engine(void);
/* number of kernels this library has launched since load (bench.py "gpu_launches") */
long long ggpm_launch_count(void);

/* Optional per-launch timing for the roofline report (bench.py): while enabled, every kernel launch is
 * bracketed by CUDA events on its stream.  Classes: 0 gather-sum, 1 tcgen05 GEMM, 2 fp32 CUDA-core GEMM,
 * 3 weight-gradient GEMM, 4 forward cell, 5 backward cell, 6 misc.  collect() synchronises the device and
 * returns, per class, summed milliseconds, launch count and algorithmic work (FLOPs for GEMMs, bytes for
 * gather-sum excluding the data-dependent neighbour rows, which the caller adds from nnz). */
int ggpm_prof_enable(int on);
int ggpm_prof_collect(double* ms, long long* count, double* work, int n_class);

/* ---- index construction ------------------------------------------------------------------------
 * Padded neighbour tensor (create_pad_tensor, ggpm/nnutils.py:105-110; rows built by
 * MolGraph.tensorize_graph, ggpm/mol_graph.py:238-281) -> CSR of its non-zero entries in slot order.
 * idx: [rows, width] int64 (idx_is_64bit=1, the dtype make_cuda produces, ggpm/nnutils.py:210-214) or
 * int32.  row_ptr: [rows+1]; col: capacity rows*width; *bad_flag is set to 1 if an entry is outside
 * [0, n_src_rows) (bit 0 is OR-ed in; such an entry is stored as 0, the padding id).  nnz stays on the device
 * (row_ptr[rows]). */
int ggpm_csr_from_padded(const void* idx, int idx_is_64bit, int rows, int width, int n_src_rows,
                         int32_t* row_ptr, int32_t* col, int32_t* bad_flag, void* stream);

/* Range check of an id column before it reaches a kernel that indexes with it (the reference's index_select raises
 * on a label >= table rows, a position >= MAX_POS or a source node >= n_nodes: ggpm/encoder.py:103-106,121-125;
 * torch raises IndexError / a device assert).  ids: n entries at `stride` elements apart (a column of fmess),
 * int64 or int32; *bad_flag |= bit if any entry is outside [lo, hi).  The caller zeroes the flag, may OR several
 * checks into it with distinct bits, and reads it back once (no synchronisation in here). */
int ggpm_check_ids(const void* ids, int ids_are_64bit, long long n, long long stride, long long lo, long long hi,
                   int32_t* bad_flag, int bit, void* stream);

/* Reverse CSR: for every source row n the destination rows that list it (t_row, ascending) and the
 * position of that entry in the forward `col` array (t_pos).  This is the deterministic form of the
 * scatter-add that autograd runs for index_select (ggpm/nnutils.py:65-70).  cursor: scratch [n_src_rows]. */
int ggpm_csr_transpose(const int32_t* row_ptr, const int32_t* col, int rows, int n_src_rows,
                       int32_t* t_ptr, int32_t* t_row, int32_t* t_pos, int32_t* cursor, void* stream);

/* ---- neighbour gather-sum ----------------------------------------------------------------------
 * out[r, 0:width] (+)= sum_{j in row r} table[col[j], 0:width]
 * == index_select_ND(table, 0, index).sum(dim=1)   (ggpm/nnutils.py:65-70; ggpm/rnn.py:26,86;
 * ggpm/encoder.py:31-32,100,135-136).  With the reverse CSR it is the backward of the same. */
int ggpm_gather_sum(const float* table, int ld_table, const int32_t* row_ptr, const int32_t* col,
                    float* out, int ld_out, int rows, int width, int accumulate, void* stream);

/* dW [n_table_rows, ldw] = 0; dW[idx[r]] += dOut[r, :width] — backward of the embedding lookups E_c / E_i
 * (nn.Embedding, ggpm/encoder.py:62-66 used at :98,111).  idx: int32 or int64 ids read in place with element stride
 * idx_stride (a column of fnode); ids outside [0, n_table_rows) are skipped. */
int ggpm_embedding_bwd(const float* dOut, int ld, const void* idx, int idx_is_64bit, long long idx_stride, int rows,
                       int width, int n_table_rows, float* dW, int ldw, void* stream);

/* ---- dense layers --------------------------------------------------------------------------------
 * C[M,N] = act( A(M,K) * B(K,N) + beta*C + bias[n] ), row 0 optionally forced to 0.
 * A(m,k) = A[m*sam + k*sak], B(k,n) = B[k*sbk + n*sbn] (one stride of each operand must be 1).
 * Replaces nn.Linear + activation + row-0 mask: W_o (ggpm/encoder.py:15-19,33-38), W_i / W_c
 * (ggpm/encoder.py:62-72,101,112), W_root (ggpm/encoder.py:79-82,138), their autograd backward
 * (input grads: B = W; weight grads: A transposed, split_k > 1 accumulates atomically into C). */
int ggpm_gemm(const float* A, long long sam, long long sak, const float* B, long long sbk, long long sbn,
              float* C, int ldc, int M, int N, int K, float beta, const float* bias, int act,
              int zero_row0, int split_k, void* stream);

/* The same product on the tcgen05 tensor cores (3xTF32 operand split: fp32-level accuracy) with the neighbour
 * gather-sum fused into the operand prologue, so that index_select_ND(h, 0, bgraph).sum(1) feeding a gate
 * Linear (ggpm/rnn.py:26-28, 86-91) is never materialised:
 *   a_row_ptr/a_col (A k-contiguous): A(m, :) = sum_{j in CSR row m} A[a_col[j], :]
 *   b_row_ptr/b_col (B n-contiguous): B(k, :) = sum_{j in CSR row k} B[b_col[j], :]
 * "Second region": indices >= split_at along split_dim (1: output columns n, 2: k, 3: output rows m) take no
 * gather and read A from A2 (leading stride lda2; k resp. m counted from split_at) when A2 is given.  Used to
 * run [S U_g^T | h U_f^T], [gather^T(dA) | dP] [U_g; U_f] and [dA | dP]^T [S | h] as single launches.
 * split_at must be a multiple of the tile (128 rows / 128 columns / 32 k).  tile_n: 0 auto, 128 or 256.
 * B_hi/B_lo (optional, k-contiguous B only): pre-split planes of B, loaded by TMA instead of through registers.
 * Fails (no fallback) unless the engine is tcgen05-capable and operands are 16-byte aligned with strides % 4 == 0. */
int ggpm_gemm_fused(const float* A, long long sam, long long sak, const float* B, long long sbk, long long sbn,
                    float* C, int ldc, int M, int N, int K, float beta, const float* bias, int act, int zero_row0,
                    int split_k, const int32_t* a_row_ptr, const int32_t* a_col, const int32_t* b_row_ptr,
                    const int32_t* b_col, int split_dim, int split_at, const float* A2, long long lda2, int tile_n,
                    const float* B_hi, const float* B_lo, void* stream);

/* hi = tf32(w) (round to nearest), lo = tf32(w - hi), element-wise: the 3xTF32 operand split of a weight matrix, done
 * once so that ggpm_gemm_fused can stream the planes by TMA (B_hi/B_lo: same shape and strides as a k-contiguous B). */
int ggpm_tf32_split(const float* w, long long n, float* hi, float* lo, void* stream);

/* ---- fp16 two-term split engine ("f16x3"): tcgen05.mma.kind::f16 at twice the tf32 rate, same 22-bit operands ----
 * v * s ~ hi + lo with hi = fp16(v s), lo = fp16(v s - hi); s is a power of two that brings the operand's largest
 * magnitude into [2^13, 2^14) (fp16 has a 5-bit exponent), divided out exactly in the epilogue.
 * ggpm_f16_prepare_weights: planes of W [rows, cols] (transpose = 0: planes [rows, cols]; 1: planes [cols, rows]) as
 * fp16 hi / lo of W * scale; scal[0] <- max|W|, scal[1] <- scale (both device floats). */
int ggpm_f16_prepare_weights(const float* W, int rows, int cols, int transpose, void* hi, void* lo, float* scal,
                             void* stream);

/* C[M,N] = epilogue(A(M,K) * B(K,N)) with A fp32 k-contiguous (row stride lda; optional CSR gather-sum a_row_ptr/a_col
 * and second region along n (split_dim 1) or k (split_dim 2, A2/lda2) exactly as in ggpm_gemm_fused) and B given as
 * the k-contiguous fp16 planes B_hi/B_lo [N, K] (row stride ldb halves) made by ggpm_f16_prepare_weights with
 * *b_scale their scale.  The A scale is derived on the device from the bound
 *   max|A(m, k)| <= amax * max(1, *a_deg),   amax = a_amax ? max(a_amax[0 .. a_amax_n)) : a_amax_const
 * (a_deg: largest CSR row length when rows are gather-sums, else NULL).  The graph LSTM uses it for the gate /
 * forget projections (|h| <= 1) and for the state gradient (amax recorded by the cell backward kernel). */
int ggpm_gemm_f16x3(const float* A, long long lda, const void* B_hi, const void* B_lo, const float* b_scale,
                    long long ldb, float* C, int ldc, int M, int N, int K, float beta, const float* bias, int act,
                    int zero_row0, const int32_t* a_row_ptr, const int32_t* a_col, int split_dim, int split_at,
                    const float* A2, long long lda2, const float* a_amax, int a_amax_n, float a_amax_const,
                    const int32_t* a_deg, int tile_n, void* stream);

/* Both operands as fp16 hi / lo planes (ggpm_f16_prepare_weights makes them for any matrix), all loads by TMA:
 *   tn = 0: C[M,N] = beta C + A B^T with A planes [M, K] (row stride lda halves), B planes [N, K] (row stride ldb);
 *   tn = 1: C[M,N] = beta C + A^T B with A planes [K, M], B planes [K, N]  (weight gradients: the reduction index is
 *           the row of both operands; M, N multiples of 64).
 * a_scale / b_scale: device floats holding the planes' scales.  beta is 0 or 1. */
int ggpm_gemm_f16x3_planes(const void* A_hi, const void* A_lo, const float* a_scale, long long lda, const void* B_hi,
                           const void* B_lo, const float* b_scale, long long ldb, float* C, int ldc, int M, int N, int K,
                           float beta, int tn, void* stream);

/* ---- weight packing (one launch each way; the reference layout is nn.Linear(IN + H -> H) per gate on [x | h],
 * ggpm/rnn.py:69-72, with x = [node feature (F) | one-hot columns (IN - F)], ggpm/encoder.py:103-106,114-116,121-125)
 * W4 / b4: host arrays of 4 device pointers, gate order i, o, u (rnn.W), f: weights [H, IN + H] row-major, biases [H].
 * Outputs (zero padded): Wx [4Hp, Fp] node-feature columns; T [ntab, 4Hp] one-hot columns as lookup rows, rows
 * [0, n_bias_rows) also carry the bias (ntab = IN - F, or 1 when IN == F); U4 = [U_i; U_o; U_u; U_f] [4Hp, Hp]. */
int ggpm_lstm_pack(const float* const* W4, const float* const* b4, int H, int Hp, int F, int Fp, int IN, int ntab,
                   int n_bias_rows, float* Wx, float* T, float* U4, void* stream);
/* gradients of (Wx, T, U4) (each may be NULL = zero) back to the four [H, IN + H] weights and [H] biases (overwritten) */
int ggpm_lstm_unpack(float* const* dW4, float* const* db4, int H, int Hp, int F, int Fp, int IN, int n_bias_rows,
                     const float* dWx, const float* dT, const float* dU4, void* stream);
/* W [N, K] (row stride ldw) -> n_pieces (<= 4) column blocks [c0[i], c1[i]) zero padded to out[i] [Np, kp[i]], and
 * bias [N] -> bias_p [Np] (both optional): the operands of the cat-free Linear layers (ggpm/encoder.py:33-34,101,112,138)
 * at tensor-core friendly shapes.  c0 / c1 / kp / out are host arrays.  split_unpad is the gradient (dW contiguous). */
int ggpm_split_pad(const float* W, int ldw, int N, int K, int Np, int n_pieces, const int* c0, const int* c1,
                   const int* kp, float* const* out, const float* bias, float* bias_p, void* stream);
int ggpm_split_unpad(float* dW, int N, int K, int n_pieces, const int* c0, const int* c1, const int* kp,
                     float* const* dpieces, const float* dbias_p, float* dbias, void* stream);

/* dZ = dY * act'(Y) with row 0 optionally zeroed (backward of the fused epilogue above). */
int ggpm_act_bwd(const float* dY, int ld_dy, const float* Y, int ld_y, float* dZ, int ld_dz, int rows,
                 int width, int act, int zero_row0, void* stream);

/* C[rows, width] = act(C + bias) in place, optional zero row 0: the epilogue of a linear layer (nn.Linear + ReLU /
 * Tanh + mask, ggpm/encoder.py:33-38) whose product was accumulated from K-split partial sums. */
int ggpm_bias_act(float* C, int ldc, int rows, int width, const float* bias, int act, int zero_row0, void* stream);

/* out[c] (+)= sum_r X[r, c]  (bias gradients) */
int ggpm_colsum(const float* X, int ld_x, int rows, int width, float* out, int accumulate, void* stream);

/* dT[k, :] += sum_{e: k1[e]==k} dX[e, :] + sum_{e: k2[e]==k} dX[e, :]   (k2 may be NULL)
 * Weight gradient of the one-hot part of the message input (E_b / E_apos / E_pos columns,
 * ggpm/encoder.py:104-106,114-116,121-125): the one-hot GEMM collapses to a row scatter. */
int ggpm_table_wgrad(const float* dX, int ld_dx, int rows, int width, const int32_t* k1, const int32_t* k2,
                     float* dT, int ld_dt, int n_table_rows, void* stream);

/* ---- graph LSTM over a message graph ---------------------------------------------------------------
 * LSTM.forward (ggpm/rnn.py:96-108) with the cell of ggpm/rnn.py:85-94, `depth` rounds, h0 = c0 = 0, row 0
 * masked after every round.  The depth-invariant input projection is hoisted and never materialised:
 *     x-part of gate g for message e:  Yn[src[e], g] + T[k1[e], g] (+ T[k2[e], g])
 * with Yn = (node features) x (x-columns of W_i|W_o|W|W_f)^T, T = the one-hot columns (+bias), both with
 * 4*Hp columns packed [i | o | u | f].  Ug = [U_i;U_o;U_u] ([3Hp, Hp] row-major, h-columns of the gate
 * weights), Uf [Hp, Hp].  The forget gate is factored per SOURCE message: P = h U_f^T, f_en = s(Xf_e + P_n).
 * History buffers (saved for backward): hist_h, hist_c: [slots, E1, Hp]; hist_gp: [slots, E1, 4Hp]
 * ([G_i|G_o|G_u|P]); slots = depth when keep_history, else 2 (ping-pong).  s_scratch: [E1, Hp].
 * w_split: optional workspace [2, 4Hp, Hp] (may be NULL): when Uf directly follows Ug in memory ([U_i;U_o;U_u;U_f] in
 * one buffer) the weights are split once into tf32 hi/lo planes there and streamed by TMA in every round.
 * The final hidden state is slot (depth-1) (keep_history) or slot ((depth-1) & 1). */
int ggpm_lstm_fwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                  const float* Yn, const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2,
                  const float* Ug, const float* Uf, float* hist_h, float* hist_c, float* hist_gp,
                  float* s_scratch, float* w_split, void* hist_h16, int keep_history, int h_rotate, void* stream);

/* Back-propagation through time of the above (what autograd does for ggpm/rnn.py:96-108).
 * dH_final: [E1, Hp] gradient w.r.t. the final hidden state (consumed, used as scratch); dC_final: [E1, Hp]
 * gradient w.r.t. the final cell state, or NULL for zero (LSTM.forward returns (h, c), ggpm/rnn.py:108; the
 * encoders only consume h).
 * Outputs: dX [E1, 4Hp] = gradient of the hoisted x-projection summed over the rounds (written once, after the
 * loop), dUg [3Hp, Hp] and dUf [Hp, Hp] (accumulated into: caller zeroes).  bt_*: reverse CSR of (b_ptr, b_col).
 * hist_gp and hist_c are CONSUMED: every round's gate gradients overwrite that round's [G_i|G_o|G_u] block of hist_gp
 * and its forget-gate input gradient overwrites that round's slot of hist_c (each is dead once its round has been
 * back-propagated), so no [E1, 4Hp] accumulator is read and re-written per round; a second call needs a new forward.
 * bt_pos (position of every reverse entry in b_col; ggpm_csr_transpose's t_pos) and pair_buf [3, pair_rows, Hp] with
 * pair_rows >= nnz (both may be NULL / 0): on the fp16-plane path the scatter-adds of a round (sum_n daf_en into dP,
 * dchat f_en into dC of the previous state) are then stored per pair and gathered over the reverse CSR where they are
 * consumed — no atomics, no zeroing launch; without them they are red.global.add into zeroed dP / dC.
 * Scratch: dA (unused since round 2, may be NULL), dC [2, E1, Hp], dP [E1, Hp], dS [E1, Hp], S [E1, Hp], dHn [E1, Hp];
 * w_split: optional [2, Hp, 4Hp] (may be NULL) for the tf32 planes of [Ug; Uf]^T streamed by TMA. */
int ggpm_lstm_bwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col,
                  const int32_t* bt_ptr, const int32_t* bt_row, const float* Yn, const int32_t* src,
                  const float* T, const int32_t* k1, const int32_t* k2, const float* Ug, const float* Uf,
                  const float* hist_h, float* hist_c, float* hist_gp, float* dH_final,
                  const float* dC_final, float* dX, float* dUg, float* dUf, float* dA, float* dC, float* dP, float* dS,
                  float* S, float* dHn, float* w_split, const void* hist_h16, void* z_planes, int h_rotate,
                  const int32_t* bt_pos, float* pair_buf, long long pair_rows, void* stream);

/* Which optional buffers the two calls above will use under the current engine, so the caller can size them:
 * bit 0: f16x3 gate GEMMs; bit 1: fp16 h planes + Z planes (then hist_h may be 2 rotating slots: h_rotate = 1). */
int ggpm_lstm_plan(int Hp, int depth);

/* f16x3 engine extras of the two calls above (all optional, NULL = not used):
 *   hist_h16 [depth, 2, E1, Hp] fp16: hi / lo planes of h^t * 2^13 written by the forward cell kernel (keep_history)
 *     and read by TMA as the B operand of the weight-gradient GEMM;
 *   z_planes [2, E1, 4Hp] fp16: scratch for Z = [gather^T(dA) | dP] of one backward round (hi / lo planes);
 *   h_rotate: hist_h holds only 2 rotating slots (slot (t - 1) & 1 for h^t) although keep_history is set: the backward
 *     then reads h only through hist_h16 (requires ggpm_lstm_plan bit 1);
 *   w_split doubles as the engine's workspace: fp16 weight planes in its first half, device scalars behind them. */

/* ---- graph GRU over a message graph ----------------------------------------------------------------
 * GRU.forward (ggpm/rnn.py:41-50) with the cell of ggpm/rnn.py:25-39.  x-gates packed [z | h | r] (3*Hp
 * columns of Yn / T); Uz, Ur, Uh: [Hp, Hp]; br: [Hp] (bias of U_r).  History: hist_h [slots, E1, Hp] holds
 * h^{t-1} of every round (slot 0 = zeros) plus the final state in slot `depth`; hist_aux [depth, E1, 5Hp] =
 * [S | Gz | Gh | Pr | Sg] per round (keep_history) or [1, E1, 5Hp]. */
int ggpm_gru_fwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col, const float* Yn,
                 const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2, const float* Uz,
                 const float* Ur, const float* br, const float* Uh, float* hist_h, float* hist_aux,
                 int keep_history, void* stream);

/* BPTT of the above.  Outputs: dX [E1, 3Hp] (zeroed here), dUz, dUr, dUh [Hp, Hp], dbr [Hp] (accumulated
 * into).  Scratch: dA [E1, 2Hp], dS [E1, Hp], dSg [E1, Hp], dPr [E1, Hp], dHn [E1, Hp]. */
int ggpm_gru_bwd(int E1, int H, int Hp, int depth, const int32_t* b_ptr, const int32_t* b_col, const float* Yn,
                 const int32_t* src, const float* T, const int32_t* k1, const int32_t* k2, const float* Uz,
                 const float* Ur, const float* Uh, const float* hist_h, const float* hist_aux, float* dH_final,
                 float* dX, float* dUz, float* dUr, float* dbr, float* dUh, float* dA, float* dS, float* dSg,
                 float* dPr, float* dHn, void* stream);

/* ---- latent head --------------------------------------------------------------------------
 * Element-wise part of PropertyVAE.rsample / HierVAE.rsample (ggpm/property_vae.py:26-34) after the fused
 * [W_mean; W_var] projection.  Y [B, ldy] holds z_mean in columns [0, L) and the variance pre-activation v in columns
 * [var_off, var_off + L):  z_log_var = -|v|,  kl = -0.5 sum(1 + z_log_var - m^2 - exp(z_log_var)) / B (deterministic
 * two-stage sum; row_kl [B] is scratch),  z = m + exp(z_log_var / 2) eps, or z = m when eps is NULL.
 * Backward: dY [B, width] = [dm | .. | dv | 0 ..] from dz [B, L] (may be NULL) and the device scalar dkl (may be NULL). */
int ggpm_rsample_fwd(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L, float* z, int ldz,
                     float* kl, float* row_kl, void* stream);
int ggpm_rsample_bwd(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L, const float* dz,
                     int ld_dz, const float* dkl, float* dY, int ld_dy, int width, void* stream);

/* Allocation-free variant of ggpm_rsample_fwd.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation).  Same arguments and results as ggpm_rsample_fwd minus row_kl: the per-row KL scratch is a
 * file-scope `static __device__` array of 65 536 floats (256 KiB, held by every GPU that loads the library until exit),
 * whose device address the library obtains once per device with cudaGetSymbolAddress.  B > 65 536 is rejected with an
 * error (use ggpm_rsample_fwd with a caller buffer).  The call makes no cudaMalloc / cudaFree / allocating library
 * call.  All calls on a device share the array: calls whose kernels may overlap on the GPU (different streams without
 * ordering) race on it; calls ordered on one stream are safe, and nothing has to be cleared between them (each call
 * writes rows [0, B) before reading them). */
int ggpm_rsample_fwd_noalloc(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L, float* z,
                             int ldz, float* kl, void* stream);

/* ---- property heads behind the latent head ---------------------------------------------------
 * PropertyOptimizer of ggpm/property_optimizer.py:5-51 in its shipped shape: two one-hidden-layer regressors (HOMO,
 * LUMO; PropertyRegressor :54-66) on [B, In] inputs,  a = relu(W1 x + b1),  y = w2 . (a * mult) + b2,  loss = mean
 * (y - t)^2, both heads in one launch.  ptrs: 2 x 17 device pointers, head 0 then head 1, each
 *   {x [B, ldx], W1 [Hd, In], b1 [Hd], w2 [Hd], b2 [1], mult [B, ld_mult] | NULL (dropout multiplier), target [B] | NULL,
 *    hid [B, Hd] (out: saved activations), out [B], sq [B] (scratch), loss [1],
 *    dout [B] | NULL, dloss [1] | NULL, dx [B, ld_dx], dW1 [Hd, In], db1 [Hd], dw2 [Hd]}
 * (the last six are used by the backward call only; db2 [2] is passed separately).  In <= 64, Hd * In <= 8192. */
int ggpm_mlp_heads_fwd(const void* const* ptrs, int ldx0, int ldx1, int ld_mult, int B, int In, int Hd, void* stream);
int ggpm_mlp_heads_bwd(const void* const* ptrs, float* db2, int ldx0, int ldx1, int ld_mult, int ld_dx, int B, int In, int Hd,
                       void* stream);

/* ---- incremental ("sparse") message passing of the decoder's encoder ---------------------
 * Replaces GRU.sparse_forward / LSTM.sparse_forward (ggpm/rnn.py:52-59, 110-121) and index_scatter
 * (ggpm/nnutils.py:124-128): M sub-messages of an R-row state table are recomputed for `depth` rounds, each round
 * reading the K padded neighbour slots of every sub-message from the whole table (slot 0 reads row 0, as
 * index_select_ND does; slots are not compacted because the decoder zeroes them mid-row, ggpm/decoder.py:72-83).
 * Row-major fp32 operands; tables of arbitrary width, idx are int32 row ids of the table.  The cell kernels take one
 * width H for all their operands: callers run them at the padded width (multiple of 4 floats, zero tail) so that the
 * H x H products around them run on the tensor-core GEMM engine; a zero tail stays zero through both cells.
 *   rows_gather:      pairs [M*K, ld_p] = table[idx[p]] (may be NULL), sum [M, ld_s] = sum over the K slots (may be NULL);
 *                     output columns [width, pad_to) are zero-filled (pad_to = width: none)
 *   rows_scatter:     table[idx[m]] = src[m]   (src NULL: zero the rows); idx must not repeat
 *   rows_scatter_add: table[idx[p]] += src[p] + bcast[p / K]   (either may be NULL) — backward of rows_gather
 *   sparse_lstm_cell_fwd: A [M, lda>=4H] pre-activations (i | o | u | f x-part), Pf / Cn [M*K, H] = U_f h_nei / c_nei
 *       -> h_sub, c_sub [M, H]; saves gates [M, 4H] = (i | o | u | tanh c) and fs [M*K, H]   (ggpm/rnn.py:85-94)
 *   sparse_lstm_cell_bwd: -> dA [M, 4H], dPf [M*K, H], dCn [M*K, H]; dXacc [M, 4H] += dA (sum over the rounds)
 *   sparse_gru_gate_fwd:  rs [M*K, H] = sigmoid(Xr[m] + Pr[p]), Sg [M, H] = sum_k rs * Hn        (ggpm/rnn.py:30-35)
 *   sparse_gru_gate_bwd:  -> dPr, dHn [M*K, H] (dHn = direct term dSg * r), dXr [M, ldx] += sum_k dPr
 *   sparse_gru_out_fwd:   A [M, lda>=2H] pre-activations (z | h~), S [M, H] -> h_sub = (1-z) S + z h~; saves zh [M, 2H]
 *   sparse_gru_out_bwd:   -> dA [M, 2H], dS [M, H] = dh (1 - z); dXacc [M, ldx] columns [0, 2H) += dA */
int ggpm_rows_gather(const float* table, int ld_t, const int32_t* idx, int M, int K, int width, int pad_to, float* pairs,
                     int ld_p, float* sum, int ld_s, void* stream);
int ggpm_rows_scatter(float* table, int ld_t, const int32_t* idx, int M, int width, const float* src, int ld_s,
                      void* stream);
int ggpm_rows_scatter_add(float* table, int ld_t, const int32_t* idx, int M, int K, int width, const float* src, int ld_s,
                          const float* bcast, int ld_b, void* stream);
int ggpm_sparse_lstm_cell_fwd(int M, int K, int H, const float* A, int lda, const float* Pf, const float* Cn, float* h_sub,
                              float* c_sub, float* gates, float* fs, void* stream);
int ggpm_sparse_lstm_cell_bwd(int M, int K, int H, const float* dh, const float* dc, const float* gates, const float* fs,
                              const float* Cn, float* dA, float* dPf, float* dCn, float* dXacc, void* stream);
int ggpm_sparse_gru_gate_fwd(int M, int K, int H, const float* Xr, int ldx, const float* Pr, const float* Hn, float* rs,
                             float* Sg, void* stream);
int ggpm_sparse_gru_gate_bwd(int M, int K, int H, const float* dSg, const float* rs, const float* Hn, float* dPr,
                             float* dHn, float* dXr, int ldx, void* stream);
int ggpm_sparse_gru_out_fwd(int M, int H, const float* A, int lda, const float* S, float* h_sub, float* zh, void* stream);
int ggpm_sparse_gru_out_bwd(int M, int H, const float* dh, const float* zh, const float* S, float* dA, float* dS,
                            float* dXacc, int ldx, void* stream);

/* Whole-call drivers of the incremental cells: all `depth` rounds of one sparse_forward (or of its backward) sequenced
 * inside ONE call — per round: rows_gather, two (GRU: three) GEMMs, the cell kernel(s), rows_scatter.  Operands at the
 * padded width Hp (multiple of 4, zero tail); tables hw / cw [R, ld_t] are the caller's copies of h / c (width H) and
 * are updated in place; the backward pass updates the caller's copies of dh / dc in place into the input gradients.
 * Xg [M, 4Hp] (LSTM: i | o | u | f) or [M, 3Hp] (GRU: z | h~ | r) = x-projection + biases; U4 [4Hp, Hp] = (U_i; U_o;
 * U_u; U_f).  Stacks saved for the backward pass, one slice per round: S, Sg [depth, M, Hp]; Hn, Cn, fs, rs
 * [depth, M*K, Hp]; gates [depth, M, 4Hp]; zh [depth, M, 2Hp].  Everything after the stacks is scratch of the sizes
 * the per-kernel entries above take.  Backward outputs: dXg (zeroed here, then summed over the rounds) and the weight
 * gradients dU4 / dUz / dUr / dUh (zeroed here). */
int ggpm_sparse_lstm_fwd(int M, int K, int H, int Hp, int depth, float* hw, float* cw, int ld_t, const int32_t* sub,
                         const int32_t* nbr, const float* Xg, const float* U4, float* S, float* Hn, float* Cn, float* gates,
                         float* fs, float* A, float* h_sub, float* c_sub, void* stream);
int ggpm_sparse_lstm_bwd(int M, int K, int H, int Hp, int depth, float* dHw, float* dCw, int ld_t, const int32_t* sub,
                         const int32_t* nbr, const float* U4, const float* S, const float* Hn, const float* Cn,
                         const float* gates, const float* fs, float* dXg, float* dU4, float* dA, float* dPf, float* dCn,
                         float* dh_sub, float* dc_sub, float* dS, float* dHn, void* stream);
int ggpm_sparse_gru_fwd(int M, int K, int H, int Hp, int depth, float* hw, int ld_t, const int32_t* sub, const int32_t* nbr,
                        const float* Xg, const float* Uz, const float* Ur, const float* Uh, float* S, float* Hn, float* rs,
                        float* Sg, float* zh, float* A, float* h_sub, void* stream);
int ggpm_sparse_gru_bwd(int M, int K, int H, int Hp, int depth, float* dHw, int ld_t, const int32_t* sub, const int32_t* nbr,
                        const float* Uz, const float* Ur, const float* Uh, const float* S, const float* Hn, const float* rs,
                        const float* Sg, const float* zh, float* dXg, float* dUz, float* dUr, float* dUh, float* dA,
                        float* dS, float* dSg, float* dPr, float* dHn, float* dh_sub, void* stream);

/* ---- host-side index builder (no device work, no stream): SURVEY 8(f) rank 3 ----------------------------------
 * Replaces MolGraph.tensorize_graph (ggpm/mol_graph.py:238-281: Python dict / list loops per edge over networkx
 * graphs) and the cgraph assembly of MolGraph.tensorize (ggpm/mol_graph.py:211-222) for callers that hold the batch
 * as flat int32 arrays.  Output layout is the reference's, bit for bit: 1-based ids, row 0 = padding, fnode[0] =
 * fnode[1], edges enumerated by source node then insertion order, rows zero padded to (longest + 1) columns
 * (create_pad_tensor, ggpm/nnutils.py:105-110); cgraph has no extra column.
 *
 * ggpm_tensorize_graph: node_off [n_mols+1], node_label [nodes, label_cols], edge_off [n_mols+1],
 *   edges [n_edges, 4] = (u, v, label, position), u / v local to the molecule, insertion order.  Builds into
 *   thread-local storage; sizes_out[6] = {N+1, E+1, agraph width, bgraph width, nnz(agraph), nnz(bgraph)}.
 * ggpm_tensorize_fetch: copies the last result of this thread into caller arrays (any pointer may be NULL):
 *   fnode [N+1, label_cols], fmess [E+1, 4], agraph [N+1, wa], bgraph [E+1, wb], scope [n_mols, 2] and the CSR of
 *   both neighbour tables a_ptr [N+2], a_col [nnz_a], b_ptr [E+2], b_col [nnz_b].
 * ggpm_tensorize_cgraph: cluster_off [n_motifs+1], cluster_atoms (local atom ids), motif_mol [n_motifs] (motifs
 *   listed molecule by molecule), tree_scope / graph_scope [n_mols, 2]; *width_out = longest cluster; cgraph
 *   [n_motifs+1, width] or NULL to query the width. */
int ggpm_tensorize_graph(int n_mols, const int32_t* node_off, const int32_t* node_label, int label_cols,
                         const int32_t* edge_off, const int32_t* edges, int32_t* sizes_out);
int ggpm_tensorize_fetch(int32_t* fnode, int32_t* fmess, int32_t* agraph, int32_t* bgraph, int32_t* scope,
                         int32_t* a_ptr, int32_t* a_col, int32_t* b_ptr, int32_t* b_col);
int ggpm_tensorize_cgraph(int n_mols, int n_motifs, const int32_t* cluster_off, const int32_t* cluster_atoms,
                          const int32_t* motif_mol, const int32_t* tree_scope, const int32_t* graph_scope,
                          int32_t* width_out, int32_t* cgraph);

#ifdef __cplusplus
}
#endif
#endif /* GGPM_B200_H */
