// CSR neighbour gather-sum:  out[r, :] (+)= sum_{j in row r} table[col[j], :]
//
// Replaces index_select_ND(...).sum(dim=1) (ggpm/nnutils.py:65-70 used at ggpm/rnn.py:26,86,
// ggpm/encoder.py:31-32,100,135-136) without materialising the padded [rows, K, H] tensor, and, with
// the transposed CSR, its backward (index_select backward = scatter-add) as a deterministic gather.
// One warp per destination row, 128-bit loads, up to 4 neighbour rows in flight per lane.
// HBM-bound: algorithmic bytes = nnz*W*4 read + rows*W*4 written (+ index bytes).
#include "common.cuh"

namespace ggpm {

constexpr int kGatherWarpsPerBlock = 8;
constexpr int kCH = 4;   // float4 chunks per lane per pass -> 512 floats per pass

template <bool ACCUM>
__global__ void __launch_bounds__(kGatherWarpsPerBlock * 32)
gather_sum_kernel(const float* __restrict__ table, int ld_t, const int* __restrict__ row_ptr,
                  const int* __restrict__ col, float* __restrict__ out, int ld_o, int rows, int w4) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int r = warp; r < rows; r += nwarps) {
        const int b = __ldg(row_ptr + r), e = __ldg(row_ptr + r + 1);
        for (int c0 = 0; c0 < w4; c0 += 32 * kCH) {
            float4 acc[kCH];
#pragma unroll
            for (int k = 0; k < kCH; ++k) acc[k] = f4zero();
            int j = b;
            for (; j + 4 <= e; j += 4) {
                const float* p0 = table + (size_t)__ldg(col + j) * ld_t;
                const float* p1 = table + (size_t)__ldg(col + j + 1) * ld_t;
                const float* p2 = table + (size_t)__ldg(col + j + 2) * ld_t;
                const float* p3 = table + (size_t)__ldg(col + j + 3) * ld_t;
#pragma unroll
                for (int k = 0; k < kCH; ++k) {
                    int c = c0 + lane + 32 * k;
                    if (c < w4) {
                        float4 v0 = ld4(p0 + 4 * c), v1 = ld4(p1 + 4 * c), v2 = ld4(p2 + 4 * c), v3 = ld4(p3 + 4 * c);
                        acc[k] += v0; acc[k] += v1; acc[k] += v2; acc[k] += v3;   // slot order, like sum(dim=1)
                    }
                }
            }
            for (; j < e; ++j) {
                const float* p0 = table + (size_t)__ldg(col + j) * ld_t;
#pragma unroll
                for (int k = 0; k < kCH; ++k) {
                    int c = c0 + lane + 32 * k;
                    if (c < w4) acc[k] += ld4(p0 + 4 * c);
                }
            }
#pragma unroll
            for (int k = 0; k < kCH; ++k) {
                int c = c0 + lane + 32 * k;
                if (c < w4) {
                    float* o = out + (size_t)r * ld_o + 4 * c;
                    if (ACCUM) acc[k] += ld4(o);
                    st4(o, acc[k]);
                }
            }
        }
    }
}

// dW[idx[r]] += dOut[r]: gradient of an embedding lookup (nn.Embedding of E_c / E_i, ggpm/encoder.py:62-66,98,111).
// One warp per looked-up row, fp32 reductions into the (pre-zeroed) table; ids are read in place from a strided
// int32 / int64 column (fnode[:, 0] or fnode[:, 1]).
template <typename IdxT>
__global__ void __launch_bounds__(256)
embedding_bwd_kernel(const float* __restrict__ dOut, int ld, const IdxT* __restrict__ idx, long long idx_stride, int rows,
                     int width, int n_table_rows, float* __restrict__ dW, int ldw) {
    const int lane = threadIdx.x & 31;
    const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= rows) return;
    const long long k = (long long)idx[(long long)r * idx_stride];
    if (k < 0 || k >= n_table_rows) return;
    const float* g = dOut + (size_t)r * ld;
    float* w = dW + (size_t)k * ldw;
    for (int c = lane; c < width; c += 32) atomicAdd(w + c, g[c]);
}

int launch_gather_sum(const float* table, int ld_t, const int* row_ptr, const int* col, float* out, int ld_o,
                      int rows, int width, bool accumulate, cudaStream_t st) {
    GGPM_REQUIRE(width % 4 == 0 && ld_t % 4 == 0 && ld_o % 4 == 0, "gather_sum: width/ld must be multiples of 4 floats");
    GGPM_REQUIRE(((uintptr_t)table & 15) == 0 && ((uintptr_t)out & 15) == 0, "gather_sum: pointers must be 16-byte aligned");
    if (rows <= 0) return 0;
    int blocks = ceil_div(rows, kGatherWarpsPerBlock);
    int cap = sm_count() * 16;
    if (blocks > cap) blocks = cap;
    ProfScope prof(PROF_GATHER, st, 4.0 * (double)rows * width);
    if (accumulate)
        gather_sum_kernel<true><<<blocks, kGatherWarpsPerBlock * 32, 0, st>>>(table, ld_t, row_ptr, col, out, ld_o, rows, width / 4);
    else
        gather_sum_kernel<false><<<blocks, kGatherWarpsPerBlock * 32, 0, st>>>(table, ld_t, row_ptr, col, out, ld_o, rows, width / 4);
    return check_launch("gather_sum");
}

}  // namespace ggpm

extern "C" int ggpm_embedding_bwd(const float* dOut, int ld, const void* idx, int idx_is_64bit, long long idx_stride,
                                  int rows, int width, int n_table_rows, float* dW, int ldw, void* stream) {
    GGPM_REQUIRE(dOut && idx && dW && width >= 1 && n_table_rows >= 1, "embedding_bwd: bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    GGPM_CUDA(cudaMemsetAsync(dW, 0, sizeof(float) * ((size_t)(n_table_rows - 1) * ldw + width), st));
    if (rows <= 0) return 0;
    ggpm::ProfScope prof(ggpm::PROF_MISC, st, 0.0);
    if (idx_is_64bit)
        ggpm::embedding_bwd_kernel<long long><<<ggpm::ceil_div(rows, 8), 256, 0, st>>>(dOut, ld, (const long long*)idx, idx_stride, rows,
                                                                                   width, n_table_rows, dW, ldw);
    else
        ggpm::embedding_bwd_kernel<int><<<ggpm::ceil_div(rows, 8), 256, 0, st>>>(dOut, ld, (const int*)idx, idx_stride, rows, width,
                                                                             n_table_rows, dW, ldw);
    return ggpm::check_launch("embedding_bwd");
}

extern "C" int ggpm_gather_sum(const float* table, int ld_table, const int32_t* row_ptr, const int32_t* col,
                               float* out, int ld_out, int rows, int width, int accumulate, void* stream) {
    GGPM_REQUIRE(table && row_ptr && col && out, "gather_sum: null pointer");
    return ggpm::launch_gather_sum(table, ld_table, row_ptr, col, out, ld_out, rows, width, accumulate != 0,
                                   (cudaStream_t)stream);
}
