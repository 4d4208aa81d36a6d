// Incremental ("sparse") message passing used by the decoder's encoder:
// GRU / LSTM.sparse_forward (ggpm/rnn.py:52-59, 110-121) and index_scatter (ggpm/nnutils.py:124-128).
//
// A step updates M sub-messages out of a table of R rows.  Every round reads the K padded neighbour slots of each
// sub-message from the WHOLE table (slot value 0 reads table row 0, exactly as index_select_ND does — the decoder
// zeroes slots in the middle of a row with its edge mask, ggpm/decoder.py:72-83, so slots are never compacted here),
// runs the cell on the M rows only and writes them back.  The reference does this with full-table passes
// (zeros_like + scatter_ + ones + scatter_ + mul + add per state tensor per round); here a round touches
// M*K gathered rows + M written rows and nothing else:
//     rows_gather      pairs[p] = table[idx[p]],  sum[m] = sum_k pairs[m*K + k]   (one pass; outputs zero-filled from the
//                      table width up to the padded width the GEMM path wants)
//     cell kernels     element-wise gate math on [M, .] / [M*K, .] operands (the H x H products run on the GEMM path)
//     rows_scatter     table[idx[m]] = src[m]   (or 0)
//     rows_scatter_add table[idx[p]] += src[p] + bcast[p / K]    (backward of the gathers; fp32 atomics)
// Table row widths are arbitrary (H = 250 tables come in unpadded from the caller), so accesses are scalar and coalesced
// along the row; everything between gather and scatter lives at the padded width Hp (multiple of 32, zero tail — the
// cells keep a zero tail zero), which makes every H x H product eligible for the tcgen05 GEMM engine.
#include "common.cuh"

namespace ggpm {

constexpr int kSpThreads = 256;

static inline unsigned sp_blocks(long long n) {
    long long b = ceil_div_ll(n, kSpThreads);
    const long long cap = (long long)sm_count() * 16;
    if (b > cap) b = cap;
    return (unsigned)(b < 1 ? 1 : b);
}

__global__ void __launch_bounds__(kSpThreads)
rows_gather_kernel(const float* __restrict__ table, int ld_t, const int32_t* __restrict__ idx, int M, int K, int width,
                   int pad_to, float* __restrict__ pairs, int ld_p, float* __restrict__ sum, int ld_s) {
    const long long n = (long long)M * pad_to;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / pad_to), j = (int)(i % pad_to);
        float s = 0.f;
        for (int k = 0; k < K; ++k) {
            const long long p = (long long)m * K + k;
            const float v = j < width ? table[(size_t)idx[p] * ld_t + j] : 0.f;     // columns [width, pad_to) are zero fill
            if (pairs != nullptr) pairs[(size_t)p * ld_p + j] = v;
            s += v;
        }
        if (sum != nullptr) sum[(size_t)m * ld_s + j] = s;
    }
}

__global__ void __launch_bounds__(kSpThreads)
rows_scatter_kernel(float* __restrict__ table, int ld_t, const int32_t* __restrict__ idx, int M, int width,
                    const float* __restrict__ src, int ld_s) {
    const long long n = (long long)M * width;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / width), j = (int)(i % width);
        table[(size_t)idx[m] * ld_t + j] = src != nullptr ? src[(size_t)m * ld_s + j] : 0.f;
    }
}

// Most padded slots hold 0 (mean fill of a neighbour row is 1.4 of 4-7 slots), so a plain scatter-add would send the
// majority of its atomics to the `width` addresses of table row 0.  Row-0 contributions are therefore summed per block
// in shared memory first and flushed with one global atomic per column and block.
__global__ void __launch_bounds__(kSpThreads)
rows_scatter_add_kernel(float* __restrict__ table, int ld_t, const int32_t* __restrict__ idx, int M, int K, int width,
                        const float* __restrict__ src, int ld_s, const float* __restrict__ bcast, int ld_b) {
    extern __shared__ float row0[];          // [width]
    for (int j = threadIdx.x; j < width; j += blockDim.x) row0[j] = 0.f;
    __syncthreads();
    const long long n = (long long)M * K * width;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long p = i / width;
        const int j = (int)(i % width);
        float v = src != nullptr ? src[(size_t)p * ld_s + j] : 0.f;
        if (bcast != nullptr) v += bcast[(size_t)(p / K) * ld_b + j];
        const int r = idx[p];
        if (r == 0) atomicAdd(&row0[j], v);
        else atomicAdd(&table[(size_t)r * ld_t + j], v);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < width; j += blockDim.x) {
        const float v = row0[j];
        if (v != 0.f) atomicAdd(&table[j], v);
    }
}

// LSTM cell on M rows (ggpm/rnn.py:85-94).  A [M, 4H] = pre-activations (i | o | u | x-part of f) with the U S products
// already added to the first 3H columns; Pf [M*K, H] = U_f h_nei; Cn [M*K, H] = c_nei.
// Saves gates [M, 4H] = (i | o | u | tanh c) and f [M*K, H] for the backward pass.
__global__ void __launch_bounds__(kSpThreads)
sparse_lstm_cell_fwd_kernel(int M, int K, int H, const float* __restrict__ A, int lda, const float* __restrict__ Pf,
                            const float* __restrict__ Cn, float* __restrict__ h_sub, float* __restrict__ c_sub,
                            float* __restrict__ gates, float* __restrict__ fs) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float* a = A + (size_t)m * lda;
        const float i = sigmoidf_(a[j]), o = sigmoidf_(a[H + j]), u = tanhf(a[2 * H + j]), xf = a[3 * H + j];
        float c = i * u;
        for (int k = 0; k < K; ++k) {
            const size_t p = ((size_t)m * K + k) * H + j;
            const float f = sigmoidf_(xf + Pf[p]);
            fs[p] = f;
            c += f * Cn[p];
        }
        const float tc = tanhf(c);
        h_sub[(size_t)m * H + j] = o * tc;
        c_sub[(size_t)m * H + j] = c;
        float* g = gates + (size_t)m * 4 * H;
        g[j] = i; g[H + j] = o; g[2 * H + j] = u; g[3 * H + j] = tc;
    }
}

// dA [M, 4H] = (da_i | da_o | da_u | sum_k da_f),  dPf [M*K, H] = da_f,  dCn [M*K, H] = dc * f;  dXacc [M, 4H] += dA
__global__ void __launch_bounds__(kSpThreads)
sparse_lstm_cell_bwd_kernel(int M, int K, int H, const float* __restrict__ dh, const float* __restrict__ dc,
                            const float* __restrict__ gates, const float* __restrict__ fs, const float* __restrict__ Cn,
                            float* __restrict__ dA, float* __restrict__ dPf, float* __restrict__ dCn,
                            float* __restrict__ dXacc) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float* g = gates + (size_t)m * 4 * H;
        const float i = g[j], o = g[H + j], u = g[2 * H + j], tc = g[3 * H + j];
        const float gh = dh[(size_t)m * H + j];
        const float dct = dc[(size_t)m * H + j] + gh * o * (1.0f - tc * tc);
        float dxf = 0.f;
        for (int k = 0; k < K; ++k) {
            const size_t p = ((size_t)m * K + k) * H + j;
            const float f = fs[p];
            const float daf = dct * Cn[p] * f * (1.0f - f);
            dPf[p] = daf;
            dCn[p] = dct * f;
            dxf += daf;
        }
        float* d = dA + (size_t)m * 4 * H;
        float* x = dXacc + (size_t)m * 4 * H;
        const float dai = dct * u * i * (1.0f - i), dao = gh * tc * o * (1.0f - o), dau = dct * i * (1.0f - u * u);
        d[j] = dai; d[H + j] = dao; d[2 * H + j] = dau; d[3 * H + j] = dxf;
        x[j] += dai; x[H + j] += dao; x[2 * H + j] += dau; x[3 * H + j] += dxf;
    }
}

// GRU reset gate (ggpm/rnn.py:30-35): r = sigmoid(Xr[m] + Pr[p]),  Sg[m] = sum_k r * h_nei
__global__ void __launch_bounds__(kSpThreads)
sparse_gru_gate_fwd_kernel(int M, int K, int H, const float* __restrict__ Xr, int ldx, const float* __restrict__ Pr,
                           const float* __restrict__ Hn, float* __restrict__ rs, float* __restrict__ Sg) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float xr = Xr[(size_t)m * ldx + j];
        float s = 0.f;
        for (int k = 0; k < K; ++k) {
            const size_t p = ((size_t)m * K + k) * H + j;
            const float r = sigmoidf_(xr + Pr[p]);
            rs[p] = r;
            s += r * Hn[p];
        }
        Sg[(size_t)m * H + j] = s;
    }
}

// dPr [M*K, H] = dSg * h_nei * r (1 - r),  dHn [M*K, H] = dSg * r,  dXr [M, ldx] += sum_k dPr
__global__ void __launch_bounds__(kSpThreads)
sparse_gru_gate_bwd_kernel(int M, int K, int H, const float* __restrict__ dSg, const float* __restrict__ rs,
                           const float* __restrict__ Hn, float* __restrict__ dPr, float* __restrict__ dHn,
                           float* __restrict__ dXr, int ldx) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float g = dSg[(size_t)m * H + j];
        float s = 0.f;
        for (int k = 0; k < K; ++k) {
            const size_t p = ((size_t)m * K + k) * H + j;
            const float r = rs[p];
            const float dar = g * Hn[p] * r * (1.0f - r);
            dPr[p] = dar;
            dHn[p] = g * r;
            s += dar;
        }
        dXr[(size_t)m * ldx + j] += s;
    }
}

// GRU output (ggpm/rnn.py:27-38): A [M, 2H] = pre-activations (z | h~) with the U products added;
// h = (1 - z) S + z h~ ; saves zh [M, 2H] = (z | h~)
__global__ void __launch_bounds__(kSpThreads)
sparse_gru_out_fwd_kernel(int M, int H, const float* __restrict__ A, int lda, const float* __restrict__ S,
                          float* __restrict__ h_sub, float* __restrict__ zh) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float z = sigmoidf_(A[(size_t)m * lda + j]), ht = tanhf(A[(size_t)m * lda + H + j]);
        const float s = S[(size_t)m * H + j];
        h_sub[(size_t)m * H + j] = (1.0f - z) * s + z * ht;
        zh[(size_t)m * 2 * H + j] = z;
        zh[(size_t)m * 2 * H + H + j] = ht;
    }
}

// dA [M, 2H] = (da_z | da_h),  dS [M, H] = dh (1 - z);  dXacc [M, ldx] columns [0, 2H) += dA
__global__ void __launch_bounds__(kSpThreads)
sparse_gru_out_bwd_kernel(int M, int H, const float* __restrict__ dh, const float* __restrict__ zh,
                          const float* __restrict__ S, float* __restrict__ dA, float* __restrict__ dS,
                          float* __restrict__ dXacc, int ldx) {
    const long long n = (long long)M * H;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(t / H), j = (int)(t % H);
        const float z = zh[(size_t)m * 2 * H + j], ht = zh[(size_t)m * 2 * H + H + j];
        const float g = dh[(size_t)m * H + j];
        const float daz = g * (ht - S[(size_t)m * H + j]) * z * (1.0f - z), dah = g * z * (1.0f - ht * ht);
        dA[(size_t)m * 2 * H + j] = daz;
        dA[(size_t)m * 2 * H + H + j] = dah;
        dXacc[(size_t)m * ldx + j] += daz;
        dXacc[(size_t)m * ldx + H + j] += dah;
        dS[(size_t)m * H + j] = g * (1.0f - z);
    }
}

}  // namespace ggpm

using namespace ggpm;

#define SP_SIZES(what) GGPM_REQUIRE(M >= 1 && K >= 1 && H >= 1, what ": bad sizes")

extern "C" int ggpm_rows_gather(const float* table, int ld_t, const int32_t* idx, int M, int K, int width, int pad_to,
                                float* pairs, int ld_p, float* sum, int ld_s, void* stream) {
    GGPM_REQUIRE(table && idx && (pairs || sum), "rows_gather: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && width >= 1 && ld_t >= width, "rows_gather: bad sizes");
    GGPM_REQUIRE(pad_to >= width && (!pairs || ld_p >= pad_to) && (!sum || ld_s >= pad_to), "rows_gather: bad leading dimension");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_GATHER, st, (double)M * (K + 1) * width * 4.0);
    rows_gather_kernel<<<sp_blocks((long long)M * pad_to), kSpThreads, 0, st>>>(table, ld_t, idx, M, K, width, pad_to, pairs,
                                                                               ld_p, sum, ld_s);
    return check_launch("rows_gather");
}

extern "C" int ggpm_rows_scatter(float* table, int ld_t, const int32_t* idx, int M, int width, const float* src, int ld_s,
                                 void* stream) {
    GGPM_REQUIRE(table && idx, "rows_scatter: null pointer");
    GGPM_REQUIRE(M >= 1 && width >= 1 && ld_t >= width && (!src || ld_s >= width), "rows_scatter: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_GATHER, st, (double)M * width * 8.0);
    rows_scatter_kernel<<<sp_blocks((long long)M * width), kSpThreads, 0, st>>>(table, ld_t, idx, M, width, src, ld_s);
    return check_launch("rows_scatter");
}

extern "C" int ggpm_rows_scatter_add(float* table, int ld_t, const int32_t* idx, int M, int K, int width, const float* src,
                                     int ld_s, const float* bcast, int ld_b, void* stream) {
    GGPM_REQUIRE(table && idx && (src || bcast), "rows_scatter_add: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && width >= 1 && ld_t >= width, "rows_scatter_add: bad sizes");
    GGPM_REQUIRE((!src || ld_s >= width) && (!bcast || ld_b >= width), "rows_scatter_add: bad leading dimension");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_GATHER, st, (double)M * K * width * 12.0);
    GGPM_REQUIRE(width <= 12288, "rows_scatter_add: rows wider than 12288 floats are not supported");
    // few, fat blocks: every block ends with one flush of its row-0 partial sums
    long long blocks = ceil_div_ll((long long)M * K * width, (long long)kSpThreads * 8);
    if (blocks > 2LL * sm_count()) blocks = 2LL * sm_count();
    rows_scatter_add_kernel<<<(unsigned)(blocks < 1 ? 1 : blocks), kSpThreads, sizeof(float) * (size_t)width, st>>>(
        table, ld_t, idx, M, K, width, src, ld_s, bcast, ld_b);
    return check_launch("rows_scatter_add");
}

extern "C" int ggpm_sparse_lstm_cell_fwd(int M, int K, int H, const float* A, int lda, const float* Pf, const float* Cn,
                                         float* h_sub, float* c_sub, float* gates, float* fs, void* stream) {
    GGPM_REQUIRE(A && Pf && Cn && h_sub && c_sub && gates && fs, "sparse_lstm_cell_fwd: null pointer");
    SP_SIZES("sparse_lstm_cell_fwd");
    GGPM_REQUIRE(lda >= 4 * H, "sparse_lstm_cell_fwd: lda < 4H");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_FWD, st, (double)M * H * 4.0 * (10 + 3 * K));
    sparse_lstm_cell_fwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, K, H, A, lda, Pf, Cn, h_sub, c_sub,
                                                                                   gates, fs);
    return check_launch("sparse_lstm_cell_fwd");
}

extern "C" int ggpm_sparse_lstm_cell_bwd(int M, int K, int H, const float* dh, const float* dc, const float* gates,
                                         const float* fs, const float* Cn, float* dA, float* dPf, float* dCn, float* dXacc,
                                         void* stream) {
    GGPM_REQUIRE(dh && dc && gates && fs && Cn && dA && dPf && dCn && dXacc, "sparse_lstm_cell_bwd: null pointer");
    SP_SIZES("sparse_lstm_cell_bwd");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_BWD, st, (double)M * H * 4.0 * (10 + 4 * K));
    sparse_lstm_cell_bwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, K, H, dh, dc, gates, fs, Cn, dA, dPf,
                                                                                   dCn, dXacc);
    return check_launch("sparse_lstm_cell_bwd");
}

extern "C" int ggpm_sparse_gru_gate_fwd(int M, int K, int H, const float* Xr, int ldx, const float* Pr, const float* Hn,
                                        float* rs, float* Sg, void* stream) {
    GGPM_REQUIRE(Xr && Pr && Hn && rs && Sg, "sparse_gru_gate_fwd: null pointer");
    SP_SIZES("sparse_gru_gate_fwd");
    GGPM_REQUIRE(ldx >= H, "sparse_gru_gate_fwd: ldx < H");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_FWD, st, (double)M * H * 4.0 * (2 + 3 * K));
    sparse_gru_gate_fwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, K, H, Xr, ldx, Pr, Hn, rs, Sg);
    return check_launch("sparse_gru_gate_fwd");
}

extern "C" int ggpm_sparse_gru_gate_bwd(int M, int K, int H, const float* dSg, const float* rs, const float* Hn, float* dPr,
                                        float* dHn, float* dXr, int ldx, void* stream) {
    GGPM_REQUIRE(dSg && rs && Hn && dPr && dHn && dXr, "sparse_gru_gate_bwd: null pointer");
    SP_SIZES("sparse_gru_gate_bwd");
    GGPM_REQUIRE(ldx >= H, "sparse_gru_gate_bwd: ldx < H");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_BWD, st, (double)M * H * 4.0 * (2 + 4 * K));
    sparse_gru_gate_bwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, K, H, dSg, rs, Hn, dPr, dHn, dXr, ldx);
    return check_launch("sparse_gru_gate_bwd");
}

extern "C" int ggpm_sparse_gru_out_fwd(int M, int H, const float* A, int lda, const float* S, float* h_sub, float* zh,
                                       void* stream) {
    GGPM_REQUIRE(A && S && h_sub && zh, "sparse_gru_out_fwd: null pointer");
    GGPM_REQUIRE(M >= 1 && H >= 1 && lda >= 2 * H, "sparse_gru_out_fwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_FWD, st, (double)M * H * 4.0 * 6);
    sparse_gru_out_fwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, H, A, lda, S, h_sub, zh);
    return check_launch("sparse_gru_out_fwd");
}

extern "C" int ggpm_sparse_gru_out_bwd(int M, int H, const float* dh, const float* zh, const float* S, float* dA, float* dS,
                                       float* dXacc, int ldx, void* stream) {
    GGPM_REQUIRE(dh && zh && S && dA && dS && dXacc, "sparse_gru_out_bwd: null pointer");
    GGPM_REQUIRE(M >= 1 && H >= 1 && ldx >= 2 * H, "sparse_gru_out_bwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_CELL_BWD, st, (double)M * H * 4.0 * 7);
    sparse_gru_out_bwd_kernel<<<sp_blocks((long long)M * H), kSpThreads, 0, st>>>(M, H, dh, zh, S, dA, dS, dXacc, ldx);
    return check_launch("sparse_gru_out_bwd");
}

// ---- whole-call drivers: every round of a sparse_forward (and of its backward) launched from ONE C-ABI call ---------
// The Python host allocates the stacks of saved tensors and the scratch buffers (caching allocator) and makes one
// call per direction; the `depth` rounds are sequenced here, so the per-round cost on the host is a handful of kernel
// launches rather than a handful of interpreter round trips.  Operands live at the padded width Hp (zero tail).
#include "kernels.h"

namespace {

int gemm_rows(const float* A, long long lda, const float* W, long long ldw, float* C, int ldc, int M, int N, int K,
              float beta, cudaStream_t st) {          // C[M, N] (+)= A[M, K] @ W[N, K]^T
    GemmArgs g;
    g.A = A; g.sam = lda; g.sak = 1; g.B = W; g.sbk = 1; g.sbn = ldw; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
    g.beta = beta;
    return launch_gemm_nt(g, st);
}

int gemm_back(const float* dY, long long ldy, const float* W, long long ldw, float* dX, int ldx, int M, int N, int K,
              float beta, cudaStream_t st) {          // dX[M, N] (+)= dY[M, K] @ W[K, N]
    GemmArgs g;
    g.A = dY; g.sam = ldy; g.sak = 1; g.B = W; g.sbk = ldw; g.sbn = 1; g.C = dX; g.ldc = ldx; g.M = M; g.N = N; g.K = K;
    g.beta = beta;
    return launch_gemm_nt(g, st);
}

int gemm_wgrad(const float* dY, long long ldy, const float* X, long long ldx, float* dW, int ldw, int N, int Kw, int rows,
               cudaStream_t st) {                     // dW[N, Kw] += dY[rows, N]^T @ X[rows, Kw], cut along the rows
    GemmArgs g;
    g.A = dY; g.sam = 1; g.sak = ldy; g.B = X; g.sbk = ldx; g.sbn = 1; g.C = dW; g.ldc = ldw; g.M = N; g.N = Kw; g.K = rows;
    g.beta = 1.f;
    const int tiles = ceil_div(N, 128) * ceil_div(Kw, 128);
    int splits = ceil_div(2 * sm_count(), tiles);
    if (splits > ceil_div(rows, 256)) splits = ceil_div(rows, 256);
    g.split_k = splits < 1 ? 1 : splits;
    return launch_gemm_nt(g, st);
}

}  // namespace

extern "C" int ggpm_sparse_lstm_fwd(int M, int K, int H, int Hp, int depth, float* hw, float* cw, int ld_t,
                                    const int32_t* sub, const int32_t* nbr, const float* Xg, const float* U4, float* S,
                                    float* Hn, float* Cn, float* gates, float* fs, float* A, float* h_sub, float* c_sub,
                                    void* stream) {
    GGPM_REQUIRE(hw && cw && sub && nbr && Xg && U4 && S && Hn && Cn && gates && fs && A && h_sub && c_sub,
                 "sparse_lstm_fwd: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && H >= 1 && Hp >= H && Hp % 4 == 0 && depth >= 1 && ld_t >= H, "sparse_lstm_fwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)M * Hp, pair = (size_t)M * K * Hp;
    GGPM_TRY(ggpm_rows_scatter(hw, ld_t, sub, M, H, nullptr, 0, stream));
    GGPM_TRY(ggpm_rows_scatter(cw, ld_t, sub, M, H, nullptr, 0, stream));
    for (int t = 0; t < depth; ++t) {
        float *S_t = S + t * row, *Hn_t = Hn + t * pair, *Cn_t = Cn + t * pair, *fs_t = fs + t * pair;
        GGPM_TRY(ggpm_rows_gather(hw, ld_t, nbr, M, K, H, Hp, Hn_t, Hp, S_t, Hp, stream));
        GGPM_TRY(ggpm_rows_gather(cw, ld_t, nbr, M, K, H, Hp, Cn_t, Hp, nullptr, 0, stream));
        GGPM_CUDA(cudaMemcpyAsync(A, Xg, sizeof(float) * 4 * row, cudaMemcpyDeviceToDevice, st));
        GGPM_TRY(gemm_rows(S_t, Hp, U4, Hp, A, 4 * Hp, M, 3 * Hp, Hp, 1.f, st));
        GGPM_TRY(gemm_rows(Hn_t, Hp, U4 + (size_t)3 * Hp * Hp, Hp, fs_t, Hp, M * K, Hp, Hp, 0.f, st));
        GGPM_TRY(ggpm_sparse_lstm_cell_fwd(M, K, Hp, A, 4 * Hp, fs_t, Cn_t, h_sub, c_sub, gates + t * 4 * row, fs_t, stream));
        GGPM_TRY(ggpm_rows_scatter(hw, ld_t, sub, M, H, h_sub, Hp, stream));
        GGPM_TRY(ggpm_rows_scatter(cw, ld_t, sub, M, H, c_sub, Hp, stream));
    }
    return 0;
}

extern "C" int ggpm_sparse_lstm_bwd(int M, int K, int H, int Hp, int depth, float* dHw, float* dCw, int ld_t,
                                    const int32_t* sub, const int32_t* nbr, const float* U4, const float* S, const float* Hn,
                                    const float* Cn, const float* gates, const float* fs, float* dXg, float* dU4, float* dA,
                                    float* dPf, float* dCn, float* dh_sub, float* dc_sub, float* dS, float* dHn,
                                    void* stream) {
    GGPM_REQUIRE(dHw && dCw && sub && nbr && U4 && S && Hn && Cn && gates && fs && dXg && dU4 && dA && dPf && dCn && dh_sub &&
                     dc_sub && dS && dHn, "sparse_lstm_bwd: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && H >= 1 && Hp >= H && Hp % 4 == 0 && depth >= 1 && ld_t >= H, "sparse_lstm_bwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)M * Hp, pair = (size_t)M * K * Hp;
    const float* Uf = U4 + (size_t)3 * Hp * Hp;
    GGPM_CUDA(cudaMemsetAsync(dXg, 0, sizeof(float) * 4 * row, st));
    GGPM_CUDA(cudaMemsetAsync(dU4, 0, sizeof(float) * 4 * Hp * Hp, st));
    for (int t = depth - 1; t >= 0; --t) {
        const float *S_t = S + t * row, *Hn_t = Hn + t * pair, *Cn_t = Cn + t * pair, *fs_t = fs + t * pair;
        GGPM_TRY(ggpm_rows_gather(dHw, ld_t, sub, M, 1, H, Hp, dh_sub, Hp, nullptr, 0, stream));
        GGPM_TRY(ggpm_rows_gather(dCw, ld_t, sub, M, 1, H, Hp, dc_sub, Hp, nullptr, 0, stream));
        GGPM_TRY(ggpm_rows_scatter(dHw, ld_t, sub, M, H, nullptr, 0, stream));
        GGPM_TRY(ggpm_rows_scatter(dCw, ld_t, sub, M, H, nullptr, 0, stream));
        GGPM_TRY(ggpm_sparse_lstm_cell_bwd(M, K, Hp, dh_sub, dc_sub, gates + t * 4 * row, fs_t, Cn_t, dA, dPf, dCn, dXg, stream));
        GGPM_TRY(gemm_back(dA, 4 * Hp, U4, Hp, dS, Hp, M, Hp, 3 * Hp, 0.f, st));
        GGPM_TRY(gemm_back(dPf, Hp, Uf, Hp, dHn, Hp, M * K, Hp, Hp, 0.f, st));
        GGPM_TRY(gemm_wgrad(dA, 4 * Hp, S_t, Hp, dU4, Hp, 3 * Hp, Hp, M, st));
        GGPM_TRY(gemm_wgrad(dPf, Hp, Hn_t, Hp, dU4 + (size_t)3 * Hp * Hp, Hp, Hp, Hp, M * K, st));
        GGPM_TRY(ggpm_rows_scatter_add(dHw, ld_t, nbr, M, K, H, dHn, Hp, dS, Hp, stream));
        GGPM_TRY(ggpm_rows_scatter_add(dCw, ld_t, nbr, M, K, H, dCn, Hp, nullptr, 0, stream));
    }
    GGPM_TRY(ggpm_rows_scatter(dHw, ld_t, sub, M, H, nullptr, 0, stream));   // the incoming sub rows were discarded by the
    GGPM_TRY(ggpm_rows_scatter(dCw, ld_t, sub, M, H, nullptr, 0, stream));   // forward pass
    return 0;
}

extern "C" int ggpm_sparse_gru_fwd(int M, int K, int H, int Hp, int depth, float* hw, int ld_t, const int32_t* sub,
                                   const int32_t* nbr, const float* Xg, const float* Uz, const float* Ur, const float* Uh,
                                   float* S, float* Hn, float* rs, float* Sg, float* zh, float* A, float* h_sub,
                                   void* stream) {
    GGPM_REQUIRE(hw && sub && nbr && Xg && Uz && Ur && Uh && S && Hn && rs && Sg && zh && A && h_sub, "sparse_gru_fwd: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && H >= 1 && Hp >= H && Hp % 4 == 0 && depth >= 1 && ld_t >= H, "sparse_gru_fwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)M * Hp, pair = (size_t)M * K * Hp;
    GGPM_TRY(ggpm_rows_scatter(hw, ld_t, sub, M, H, nullptr, 0, stream));
    for (int t = 0; t < depth; ++t) {
        float *S_t = S + t * row, *Hn_t = Hn + t * pair, *rs_t = rs + t * pair, *Sg_t = Sg + t * row;
        GGPM_TRY(ggpm_rows_gather(hw, ld_t, nbr, M, K, H, Hp, Hn_t, Hp, S_t, Hp, stream));
        GGPM_TRY(gemm_rows(Hn_t, Hp, Ur, Hp, rs_t, Hp, M * K, Hp, Hp, 0.f, st));
        GGPM_TRY(ggpm_sparse_gru_gate_fwd(M, K, Hp, Xg + 2 * Hp, 3 * Hp, rs_t, Hn_t, rs_t, Sg_t, stream));
        GGPM_CUDA(cudaMemcpy2DAsync(A, sizeof(float) * 2 * Hp, Xg, sizeof(float) * 3 * Hp, sizeof(float) * 2 * Hp, M,
                                    cudaMemcpyDeviceToDevice, st));
        GGPM_TRY(gemm_rows(S_t, Hp, Uz, Hp, A, 2 * Hp, M, Hp, Hp, 1.f, st));
        GGPM_TRY(gemm_rows(Sg_t, Hp, Uh, Hp, A + Hp, 2 * Hp, M, Hp, Hp, 1.f, st));
        GGPM_TRY(ggpm_sparse_gru_out_fwd(M, Hp, A, 2 * Hp, S_t, h_sub, zh + t * 2 * row, stream));
        GGPM_TRY(ggpm_rows_scatter(hw, ld_t, sub, M, H, h_sub, Hp, stream));
    }
    return 0;
}

extern "C" int ggpm_sparse_gru_bwd(int M, int K, int H, int Hp, int depth, float* dHw, int ld_t, const int32_t* sub,
                                   const int32_t* nbr, const float* Uz, const float* Ur, const float* Uh, const float* S,
                                   const float* Hn, const float* rs, const float* Sg, const float* zh, float* dXg, float* dUz,
                                   float* dUr, float* dUh, float* dA, float* dS, float* dSg, float* dPr, float* dHn,
                                   float* dh_sub, void* stream) {
    GGPM_REQUIRE(dHw && sub && nbr && Uz && Ur && Uh && S && Hn && rs && Sg && zh && dXg && dUz && dUr && dUh && dA && dS && dSg &&
                     dPr && dHn && dh_sub, "sparse_gru_bwd: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && H >= 1 && Hp >= H && Hp % 4 == 0 && depth >= 1 && ld_t >= H, "sparse_gru_bwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    const size_t row = (size_t)M * Hp, pair = (size_t)M * K * Hp, sq = sizeof(float) * (size_t)Hp * Hp;
    GGPM_CUDA(cudaMemsetAsync(dXg, 0, sizeof(float) * 3 * row, st));
    GGPM_CUDA(cudaMemsetAsync(dUz, 0, sq, st));
    GGPM_CUDA(cudaMemsetAsync(dUr, 0, sq, st));
    GGPM_CUDA(cudaMemsetAsync(dUh, 0, sq, st));
    for (int t = depth - 1; t >= 0; --t) {
        const float *S_t = S + t * row, *Hn_t = Hn + t * pair, *rs_t = rs + t * pair, *Sg_t = Sg + t * row;
        GGPM_TRY(ggpm_rows_gather(dHw, ld_t, sub, M, 1, H, Hp, dh_sub, Hp, nullptr, 0, stream));
        GGPM_TRY(ggpm_rows_scatter(dHw, ld_t, sub, M, H, nullptr, 0, stream));
        GGPM_TRY(ggpm_sparse_gru_out_bwd(M, Hp, dh_sub, zh + t * 2 * row, S_t, dA, dS, dXg, 3 * Hp, stream));
        GGPM_TRY(gemm_back(dA, 2 * Hp, Uz, Hp, dS, Hp, M, Hp, Hp, 1.f, st));
        GGPM_TRY(gemm_back(dA + Hp, 2 * Hp, Uh, Hp, dSg, Hp, M, Hp, Hp, 0.f, st));
        GGPM_TRY(gemm_wgrad(dA, 2 * Hp, S_t, Hp, dUz, Hp, Hp, Hp, M, st));
        GGPM_TRY(gemm_wgrad(dA + Hp, 2 * Hp, Sg_t, Hp, dUh, Hp, Hp, Hp, M, st));
        GGPM_TRY(ggpm_sparse_gru_gate_bwd(M, K, Hp, dSg, rs_t, Hn_t, dPr, dHn, dXg + 2 * Hp, 3 * Hp, stream));
        GGPM_TRY(gemm_back(dPr, Hp, Ur, Hp, dHn, Hp, M * K, Hp, Hp, 1.f, st));
        GGPM_TRY(gemm_wgrad(dPr, Hp, Hn_t, Hp, dUr, Hp, Hp, Hp, M * K, st));
        GGPM_TRY(ggpm_rows_scatter_add(dHw, ld_t, nbr, M, K, H, dHn, Hp, dS, Hp, stream));
    }
    GGPM_TRY(ggpm_rows_scatter(dHw, ld_t, sub, M, H, nullptr, 0, stream));
    return 0;
}
