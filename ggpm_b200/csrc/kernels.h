// Internal launcher prototypes shared between the translation units of libggpm_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ggpm_b200.h"

namespace ggpm {

struct GemmArgs {
    const float* A = nullptr; long long sam = 0, sak = 1;   // A(m,k) = A[m*sam + k*sak]
    const float* B = nullptr; long long sbk = 1, sbn = 0;   // B(k,n) = B[k*sbk + n*sbn]
    float* C = nullptr; int ldc = 0;
    int M = 0, N = 0, K = 0;
    float beta = 0.f;
    const float* bias = nullptr;
    int act = GGPM_ACT_NONE;
    int zero_row0 = 0;
    int split_k = 1;
    // ---- tcgen05 engine only ----
    // A rows are CSR gather-sums of the rows of the table A (K-major A); B rows along k are CSR gather-sums of
    // the rows of the table B (MN-major B).
    const int* a_row_ptr = nullptr; const int* a_col = nullptr;
    const int* b_row_ptr = nullptr; const int* b_col = nullptr;
    // second region: output columns (split_dim 1) / k (2) / output rows (3) at or beyond split_at use no gather
    // and read A from A2 (leading stride lda2, indices relative to split_at along k or m) when A2 is given
    int split_dim = 0, split_at = 0;
    const float* A2 = nullptr; long long lda2 = 0;
    int tile_n = 0;   // 0 = automatic, 128 or 256 forces the N tile
    // k-contiguous B (weights) already split into tf32 hi / lo planes (same shape and strides as B): loaded by TMA
    const float* B_hi = nullptr; const float* B_lo = nullptr;
    // ---- fp16 two-term split engine ("f16x3", launch_gemm_f16) ----
    // k-contiguous B as fp16 hi / lo planes of B * (*b_scale) (same shape, row stride sbn elements): loaded by TMA.
    // A (fp32, K-major) is scaled in the producers by a power of two derived from the bound
    //   max|A row| <= amax * max(1, *a_deg),  amax = a_amax ? max(a_amax[0..a_amax_n)) : a_amax_const
    const void* Bh_hi = nullptr; const void* Bh_lo = nullptr;
    const float* b_scale = nullptr;
    const float* a_amax = nullptr; int a_amax_n = 0; float a_amax_const = 1.f;
    const int* a_deg = nullptr;
    // A given as fp16 hi / lo planes (row stride lda_h halves) of A * (*a_scale), loaded by TMA:
    //   f16_tn = 0: planes [M, K] (k-contiguous);  f16_tn = 1: planes [K, M] and B planes [K, N] (weight gradients:
    //   both operands stored with the reduction index as the row)
    const void* Ah_hi = nullptr; const void* Ah_lo = nullptr; long long lda_h = 0;
    const float* a_scale = nullptr;
    float b_scale_const = 1.f;   // used when b_scale is NULL
    int f16_tn = 0;
    // act = GGPM_ACT_LSTM_GATES (f16x3 engine, fused epilogue): C[e, n] = s(acc + X[e, n]) for n < 2 * gx_hp (input and
    // output gates, sigmoid) and tanh(...) for the candidate block, row 0 forced to zero, with the hoisted x-projection
    // X[e, n] = gx_Yn[gx_src[e], n] + gx_T[gx_k1[e], n] (+ gx_T[gx_k2[e], n]), row stride gx_ld (ggpm/rnn.py:88-89,91,106)
    const float* gx_Yn = nullptr; const int* gx_src = nullptr; const float* gx_T = nullptr;
    const int* gx_k1 = nullptr; const int* gx_k2 = nullptr; int gx_ld = 0, gx_hp = 0;
};
constexpr int GGPM_ACT_LSTM_GATES = 4;   // internal to the library (not a public activation code)

int launch_gemm_simt(const GemmArgs& g, cudaStream_t st);
// generic entry: tcgen05 3xTF32 kernel when the operand layout allows it, else the fp32 CUDA-core kernel
int launch_gemm_nt(const GemmArgs& g, cudaStream_t st);
int launch_gemm_tc(const GemmArgs& g, cudaStream_t st);
bool gemm_tc_eligible(const GemmArgs& g);
int launch_gemm_f16(const GemmArgs& g, cudaStream_t st);
bool gemm_f16_eligible(const GemmArgs& g);
// fp16 planes of w * s (or of its transpose), s = the power of two that brings *amax into [2^13, 2^14); *scale = s.
// launch_absmax accumulates max|w| into *amax (zero it first); n_slots > 1 spreads the atomics.
int launch_absmax(const float* w, long long n, float* amax, cudaStream_t st);
int launch_f16_split(const float* w, long long n, const float* amax, void* hi, void* lo, float* scale, cudaStream_t st);
int launch_f16_split_transpose(const float* w, int rows, int cols, const float* amax, void* hi, void* lo, float* scale,
                               cudaStream_t st);
// Z = [gather^T(dA[:, :3Hp]) | dP] as fp16 hi / lo planes [E1, 4Hp] scaled by the power of two derived from
// max(amax16[0..16)) * max(1, *deg); the scale is stored to *scale_out.  Z[n, :3Hp] = sum_{e in T(n)} dA[e, :3Hp].
int launch_zero2(float* a, float* b, size_t n, cudaStream_t st);   // zero two fp32 arrays of n floats (n % 4 == 0) in one launch
int launch_lstm_zgather(int E1, int Hp, const int* t_ptr, const int* t_row, const float* dA, const float* dP,
                        const int* t_pos, const float* pairA, const float* amax16, const int* deg, void* z_hi, void* z_lo,
                        float* scale_out, cudaStream_t st);
int launch_max_degree(const int* row_ptr, int rows, int* out, cudaStream_t st);   // *out = max(*out, max row length)
// hi = rn_tf32(w), lo = rn_tf32(w - hi), element-wise over n floats (weights are split once per layer call)
int launch_tf32_split(const float* w, long long n, float* hi, float* lo, cudaStream_t st);
int launch_tf32_split_transpose(const float* w, int rows, int cols, float* hi, float* lo, cudaStream_t st);

struct XTabArgs {   // hoisted x-projection: X_g[e] = Yn[src[e], g] + T[k1[e], g] (+ T[k2[e], g])
    const float* Yn; const int* src; const float* T; const int* k1; const int* k2;
};

int launch_lstm_cell_fwd(bool first, int E1, int H, int Hp, const int* rp, const int* col, const XTabArgs& x,
                         const float* gp, const float* c_prev, float* h_out, float* c_out, void* h16_hi, void* h16_lo,
                         cudaStream_t st, bool gates_activated = false);
int launch_lstm_cell_bwd(bool first, int E1, int H, int Hp, const int* rp, const int* col, const XTabArgs& x,
                         float* gp, const float* c_prev, float* c_cur, const float* dH, const float* dC,
                         float* dP, float* dC_prev, float* amax16, cudaStream_t st, const int* t_ptr = nullptr,
                         const int* t_pos = nullptr, const float* pairB_in = nullptr, float* pairA = nullptr,
                         float* pairB = nullptr, bool pairs = false, bool gates_activated = false);
int launch_lstm_sgather16(int E1, int Hp, const int* b_ptr, const int* b_col, const void* h_hi, const void* h_lo, const int* deg,
                          void* s_hi, void* s_lo, float* scale_out, cudaStream_t st);
int launch_lstm_dx_reduce(int E1, int Hp, int depth, const float* hist_gp, const float* hist_c, float* dX, cudaStream_t st);
int launch_gru_gate(int E1, int Hp, const int* rp, const int* col, const XTabArgs& x, const float* h_prev, float* aux,
                    cudaStream_t st);
int launch_gru_out(int E1, int H, int Hp, const XTabArgs& x, const float* aux, float* h_out, cudaStream_t st);
int launch_gru_out_bwd(int E1, int H, int Hp, const XTabArgs& x, const float* aux, const float* dH, float* dA, float* dS,
                       cudaStream_t st);
int launch_gru_pair_bwd(int E1, int Hp, const int* rp, const int* col, const XTabArgs& x, const float* aux,
                        const float* h_prev, const float* dA, const float* dS, const float* dSg, float* dX,
                        bool dx_overwrite, bool with_pairs, float* dPr, float* dHn, cudaStream_t st);

int launch_gather_sum(const float* table, int ld_t, const int* row_ptr, const int* col, float* out, int ld_o,
                      int rows, int width, bool accumulate, cudaStream_t st);

}  // namespace ggpm
