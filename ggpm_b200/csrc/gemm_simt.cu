// FP32 CUDA-core GEMM with a fused epilogue — the exact-fp32 path and the layout-generic fallback
// (transposed operands, ragged K, split-K weight gradients) beside the tcgen05 3xTF32 kernel.
//
//   C[M,N] = act( A(M,K) * B(K,N) + beta*C + bias[n] ) , optional zeroing of row 0
//   A(m,k) = A[m*sam + k*sak]   (one of sam/sak is 1)
//   B(k,n) = B[k*sbk + n*sbn]   (one of sbk/sbn is 1)
//
// Covers every nn.Linear on the path: gate projections (ggpm/rnn.py:13-16,69-72), W_o / W_i / W_c /
// W_root (ggpm/encoder.py:15-19,62-72,79-82) forward, their input gradients and weight gradients.
// 128x128x16 tiles, 256 threads, 8x8 outputs per thread, double-buffered shared memory.
#include "common.cuh"
#include "kernels.h"

namespace ggpm {

constexpr int BM = 128, BN = 128, BK = 16, PAD = 4, NT = 256;

// global -> registers for one BMxBK (or BKxBN) operand tile; `rows` indexes the M/N dimension
template <bool K_CONTIG>
__device__ __forceinline__ void load_tile(const float* __restrict__ base, long long s_row, long long s_k,
                                          int row0, int k0, int n_rows, int k_end, bool vec_ok, float4 (&r)[2]) {
    const int t = threadIdx.x;
    if (K_CONTIG) {
        // thread -> (row = t/4 + 64*i, k-quad = t%4)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int row = row0 + (t >> 2) + 64 * i;
            int k = k0 + (t & 3) * 4;
            float4 v = f4zero();
            if (row < n_rows) {
                const float* p = base + (long long)row * s_row + k;
                if (vec_ok && k + 3 < k_end) {
                    v = ldg4(p);
                } else {
                    if (k + 0 < k_end) v.x = __ldg(p + 0);
                    if (k + 1 < k_end) v.y = __ldg(p + 1);
                    if (k + 2 < k_end) v.z = __ldg(p + 2);
                    if (k + 3 < k_end) v.w = __ldg(p + 3);
                }
            }
            r[i] = v;
        }
    } else {
        // operand stored [K, rows]: thread -> (k = t/32 + 8*i, row-quad = t%32)
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int k = k0 + (t >> 5) + 8 * i;
            int row = row0 + (t & 31) * 4;
            float4 v = f4zero();
            if (k < k_end) {
                const float* p = base + (long long)k * s_k + row;
                if (vec_ok && row + 3 < n_rows) {
                    v = ldg4(p);
                } else {
                    if (row + 0 < n_rows) v.x = __ldg(p + 0);
                    if (row + 1 < n_rows) v.y = __ldg(p + 1);
                    if (row + 2 < n_rows) v.z = __ldg(p + 2);
                    if (row + 3 < n_rows) v.w = __ldg(p + 3);
                }
            }
            r[i] = v;
        }
    }
}

template <bool K_CONTIG>
__device__ __forceinline__ void store_tile(float (*s)[BM + PAD], const float4 (&r)[2]) {
    const int t = threadIdx.x;
    if (K_CONTIG) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int row = (t >> 2) + 64 * i, k = (t & 3) * 4;
            s[k + 0][row] = r[i].x;
            s[k + 1][row] = r[i].y;
            s[k + 2][row] = r[i].z;
            s[k + 3][row] = r[i].w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            int k = (t >> 5) + 8 * i, row = (t & 31) * 4;
            *reinterpret_cast<float4*>(&s[k][row]) = r[i];
        }
    }
}

template <bool A_KC, bool B_KC>
__global__ void __launch_bounds__(NT)
gemm_simt_kernel(const float* __restrict__ A, long long sam, long long sak, const float* __restrict__ B,
                 long long sbk, long long sbn, float* __restrict__ C, int ldc, int M, int N, int K, int k_chunk,
                 float beta, const float* __restrict__ bias, int act, int zero_row0, int a_vec, int b_vec, int c_vec) {
    __shared__ __align__(16) float As[2][BK][BM + PAD];
    __shared__ __align__(16) float Bs[2][BK][BN + PAD];
    const int t = threadIdx.x;
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kb = blockIdx.z * k_chunk;
    const int ke = min(K, kb + k_chunk);
    const int tm = (t >> 4) * 4, tn = (t & 15) * 4;   // thread owns rows {tm..tm+3, tm+64..tm+67}, cols likewise

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    float4 ra[2], rb[2];
    const int ntiles = (ke - kb + BK - 1) / BK;
    if (ntiles > 0) {
        load_tile<A_KC>(A, A_KC ? sam : 1, A_KC ? 1 : sak, m0, kb, M, ke, a_vec, ra);
        load_tile<B_KC>(B, B_KC ? sbn : 1, B_KC ? 1 : sbk, n0, kb, N, ke, b_vec, rb);
        store_tile<A_KC>(As[0], ra);
        store_tile<B_KC>(Bs[0], rb);
    }
    __syncthreads();
    for (int kt = 0; kt < ntiles; ++kt) {
        const int cur = kt & 1;
        if (kt + 1 < ntiles) {
            load_tile<A_KC>(A, A_KC ? sam : 1, A_KC ? 1 : sak, m0, kb + (kt + 1) * BK, M, ke, a_vec, ra);
            load_tile<B_KC>(B, B_KC ? sbn : 1, B_KC ? 1 : sbk, n0, kb + (kt + 1) * BK, N, ke, b_vec, rb);
        }
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float4 a0 = *reinterpret_cast<const float4*>(&As[cur][k][tm]);
            float4 a1 = *reinterpret_cast<const float4*>(&As[cur][k][tm + 64]);
            float4 b0 = *reinterpret_cast<const float4*>(&Bs[cur][k][tn]);
            float4 b1 = *reinterpret_cast<const float4*>(&Bs[cur][k][tn + 64]);
            float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        if (kt + 1 < ntiles) {
            store_tile<A_KC>(As[cur ^ 1], ra);
            store_tile<B_KC>(Bs[cur ^ 1], rb);
        }
        __syncthreads();
    }

    const bool split = gridDim.z > 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        int m = m0 + tm + (i & 3) + (i >> 2) * 64;
        if (m >= M) continue;
#pragma unroll
        for (int jh = 0; jh < 2; ++jh) {
            int n = n0 + tn + jh * 64;
            if (n >= N) continue;
            float v[4] = {acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]};
            float* cp = C + (long long)m * ldc + n;
            if (split) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (n + q < N) atomicAdd(cp + q, v[q]);
                continue;
            }
            const bool full = c_vec && (n + 3 < N);
            if (beta != 0.f) {
                if (full) {
                    float4 o = ld4(cp);
                    v[0] += beta * o.x; v[1] += beta * o.y; v[2] += beta * o.z; v[3] += beta * o.w;
                } else {
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (n + q < N) v[q] += beta * cp[q];
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (n + q >= N) continue;
                float x = v[q];
                if (bias) x += __ldg(bias + n + q);
                if (act == GGPM_ACT_RELU) x = fmaxf(x, 0.f);
                else if (act == GGPM_ACT_TANH) x = tanhf(x);
                else if (act == GGPM_ACT_SIGMOID) x = sigmoidf_(x);
                if (zero_row0 && m == 0) x = 0.f;
                v[q] = x;
            }
            if (full) {
                st4(cp, make_float4(v[0], v[1], v[2], v[3]));
            } else {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (n + q < N) cp[q] = v[q];
            }
        }
    }
}

static bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

int launch_gemm_simt(const GemmArgs& g, cudaStream_t st) {
    GGPM_REQUIRE(g.M >= 0 && g.N >= 0 && g.K >= 0, "gemm: negative size");
    if (g.M == 0 || g.N == 0) return 0;
    GGPM_REQUIRE(g.sam == 1 || g.sak == 1, "gemm: A must be contiguous along m or k");
    GGPM_REQUIRE(g.sbk == 1 || g.sbn == 1, "gemm: B must be contiguous along k or n");
    const bool a_kc = (g.sak == 1), b_kc = (g.sbk == 1);
    int splits = g.split_k > 1 ? g.split_k : 1;
    if (splits > 1) GGPM_REQUIRE(!g.bias && g.act == GGPM_ACT_NONE && !g.zero_row0, "gemm: split-K has no epilogue");
    int k_chunk = g.K;
    if (splits > 1) {
        k_chunk = ceil_div(ceil_div(g.K, splits), BK) * BK;
        if (k_chunk < BK) k_chunk = BK;
        splits = ceil_div(g.K, k_chunk);
        if (splits < 1) splits = 1;
    }
    if (splits > 1 && g.beta == 0.f) {
        // atomically accumulated: start from zero
        GGPM_CUDA(cudaMemset2DAsync(g.C, sizeof(float) * (size_t)g.ldc, 0, sizeof(float) * (size_t)g.N, (size_t)g.M, st));
    }
    const long long a_other = a_kc ? g.sam : g.sak, b_other = b_kc ? g.sbn : g.sbk;
    const int a_vec = aligned16(g.A) && (a_other % 4 == 0);
    const int b_vec = aligned16(g.B) && (b_other % 4 == 0);
    const int c_vec = aligned16(g.C) && (g.ldc % 4 == 0);
    dim3 grid(ceil_div(g.N, BN), ceil_div(g.M, BM), splits);
    ProfScope prof(a_kc ? PROF_GEMM_SIMT : PROF_GEMM_WGRAD, st, 2.0 * g.M * g.N * (double)g.K);
#define LAUNCH(AK, BKC)                                                                                         \
    gemm_simt_kernel<AK, BKC><<<grid, NT, 0, st>>>(g.A, g.sam, g.sak, g.B, g.sbk, g.sbn, g.C, g.ldc, g.M, g.N, g.K, \
                                                  k_chunk, g.beta, g.bias, g.act, g.zero_row0, a_vec, b_vec, c_vec)
    if (a_kc && b_kc) LAUNCH(true, true);
    else if (a_kc && !b_kc) LAUNCH(true, false);
    else if (!a_kc && b_kc) LAUNCH(false, true);
    else LAUNCH(false, false);
#undef LAUNCH
    return check_launch("gemm_simt");
}

}  // namespace ggpm
