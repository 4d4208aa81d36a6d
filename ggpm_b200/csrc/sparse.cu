// Incremental ("sparse") message passing used by the decoder's encoder:
// GRU / LSTM.sparse_forward (ggpm/rnn.py:52-59, 110-121) and index_scatter (ggpm/nnutils.py:124-128).
//
// A step updates M sub-messages out of a table of R rows.  Every round reads the K padded neighbour slots of each
// sub-message from the WHOLE table (slot value 0 reads table row 0, exactly as index_select_ND does — the decoder
// zeroes slots in the middle of a row with its edge mask, ggpm/decoder.py:72-83, so slots are never compacted here),
// runs the cell on the M rows only and writes them back.  The reference does this with full-table passes
// (zeros_like + scatter_ + ones + scatter_ + mul + add per state tensor per round); here a round touches
// M*K gathered rows + M written rows and nothing else:
//     rows_gather      pairs[p] = table[idx[p]],  sum[m] = sum_k pairs[m*K + k]            (one pass)
//     cell kernels     element-wise gate math on [M, .] / [M*K, .] operands (the H x H products run on the GEMM path)
//     rows_scatter     table[idx[m]] = src[m]   (or 0)
//     rows_scatter_add table[idx[p]] += src[p] + bcast[p / K]    (backward of the gathers; fp32 atomics)
// Row widths are arbitrary (H = 250 tables come in unpadded from the caller), so accesses are scalar and coalesced
// along the row.
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
                   float* __restrict__ pairs, int ld_p, float* __restrict__ sum, int ld_s) {
    const long long n = (long long)M * width;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const int m = (int)(i / width), j = (int)(i % width);
        float s = 0.f;
        for (int k = 0; k < K; ++k) {
            const long long p = (long long)m * K + k;
            const float v = table[(size_t)idx[p] * ld_t + j];
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

__global__ void __launch_bounds__(kSpThreads)
rows_scatter_add_kernel(float* __restrict__ table, int ld_t, const int32_t* __restrict__ idx, int M, int K, int width,
                        const float* __restrict__ src, int ld_s, const float* __restrict__ bcast, int ld_b) {
    const long long n = (long long)M * K * width;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long p = i / width;
        const int j = (int)(i % width);
        float v = src != nullptr ? src[(size_t)p * ld_s + j] : 0.f;
        if (bcast != nullptr) v += bcast[(size_t)(p / K) * ld_b + j];
        atomicAdd(&table[(size_t)idx[p] * ld_t + j], v);
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

extern "C" int ggpm_rows_gather(const float* table, int ld_t, const int32_t* idx, int M, int K, int width, float* pairs,
                                int ld_p, float* sum, int ld_s, void* stream) {
    GGPM_REQUIRE(table && idx && (pairs || sum), "rows_gather: null pointer");
    GGPM_REQUIRE(M >= 1 && K >= 1 && width >= 1 && ld_t >= width, "rows_gather: bad sizes");
    GGPM_REQUIRE((!pairs || ld_p >= width) && (!sum || ld_s >= width), "rows_gather: bad leading dimension");
    cudaStream_t st = (cudaStream_t)stream;
    ProfScope prof(PROF_GATHER, st, (double)M * (K + 1) * width * 4.0);
    rows_gather_kernel<<<sp_blocks((long long)M * width), kSpThreads, 0, st>>>(table, ld_t, idx, M, K, width, pairs, ld_p,
                                                                              sum, ld_s);
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
    rows_scatter_add_kernel<<<sp_blocks((long long)M * K * width), kSpThreads, 0, st>>>(table, ld_t, idx, M, K, width, src,
                                                                                       ld_s, bcast, ld_b);
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
