// Padded neighbour-index tensors -> CSR, and CSR transpose (reverse CSR), on the device.
//
// The reference keeps neighbour lists as zero-padded [rows, width] index tensors built by
// create_pad_tensor (ggpm/nnutils.py:105-110) from MolGraph.tensorize_graph (ggpm/mol_graph.py:238-281)
// and gathers every slot, padding included (index_select_ND, ggpm/nnutils.py:65-70).  Here every
// non-zero entry (not only the leading ones: the decoder zeroes entries mid-row, ggpm/decoder.py:72-83)
// becomes one CSR entry, in slot order, so the encoding is lossless and bit-exact w.r.t. the
// reference's index construction.  No host synchronisation: nnz stays on the device (row_ptr[rows]).
#include "common.cuh"

namespace ggpm {

template <typename IdxT>
__global__ void csr_count_kernel(const IdxT* __restrict__ idx, int rows, int width, int* __restrict__ cnt) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const IdxT* p = idx + (size_t)r * width;
    int c = 0;
    for (int k = 0; k < width; ++k) c += (p[k] != 0);
    cnt[r] = c;
}

template <typename IdxT>
__global__ void csr_fill_kernel(const IdxT* __restrict__ idx, int rows, int width, int n_src,
                                const int* __restrict__ row_ptr, int* __restrict__ col, int* __restrict__ bad) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    const IdxT* p = idx + (size_t)r * width;
    int o = row_ptr[r];
    for (int k = 0; k < width; ++k) {
        IdxT v = p[k];
        if (v != 0) {
            if (v < 0 || v >= (IdxT)n_src) { atomicOr(bad, 1); v = 0; }
            col[o++] = (int)v;
        }
    }
}

// In-place exclusive scan by ONE block: data[0..n-1] counts -> prefix sums, data[n] = total.
__global__ void __launch_bounds__(1024) scan_kernel(int* __restrict__ data, int n) {
    __shared__ int warp_tot[32];
    __shared__ int carry_s;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    constexpr int PER = 4;
    for (int base = 0; base < n; base += 1024 * PER) {
        int v[PER], local = 0;
        int i0 = base + tid * PER;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            v[k] = (i0 + k < n) ? data[i0 + k] : 0;
            local += v[k];
        }
        int incl = local;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        if (lane == 31) warp_tot[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int w = warp_tot[lane], wi = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int t = __shfl_up_sync(0xffffffffu, wi, d);
                if (lane >= d) wi += t;
            }
            warp_tot[lane] = wi - w;   // exclusive prefix of warp totals
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + warp_tot[wid] + incl - local;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            if (i0 + k < n) data[i0 + k] = excl;
            excl += v[k];
        }
        __syncthreads();
        if (tid == 1023) carry_s = excl;   // last thread holds the running total
        __syncthreads();
    }
    if (tid == 0) data[n] = carry_s;
}

__global__ void tr_hist_kernel(const int* __restrict__ row_ptr, const int* __restrict__ col, int rows,
                               int* __restrict__ cnt) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    for (int j = row_ptr[r]; j < row_ptr[r + 1]; ++j) atomicAdd(&cnt[col[j]], 1);
}

__global__ void tr_fill_kernel(const int* __restrict__ row_ptr, const int* __restrict__ col, int rows,
                               const int* __restrict__ t_ptr, int* __restrict__ cursor,
                               int* __restrict__ t_row, int* __restrict__ t_pos) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    for (int j = row_ptr[r]; j < row_ptr[r + 1]; ++j) {
        int n = col[j];
        int slot = t_ptr[n] + atomicAdd(&cursor[n], 1);
        t_row[slot] = r;
        t_pos[slot] = j;
    }
}

// Make the transposed lists deterministic: ascending forward position (== ascending dest row, then slot).
__global__ void tr_sort_kernel(const int* __restrict__ t_ptr, int n_src, int* __restrict__ t_row,
                               int* __restrict__ t_pos) {
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= n_src) return;
    int b = t_ptr[n], e = t_ptr[n + 1];
    for (int i = b + 1; i < e; ++i) {
        int kp = t_pos[i], kr = t_row[i];
        int j = i - 1;
        while (j >= b && t_pos[j] > kp) {
            t_pos[j + 1] = t_pos[j];
            t_row[j + 1] = t_row[j];
            --j;
        }
        t_pos[j + 1] = kp;
        t_row[j + 1] = kr;
    }
}

// ids[i * stride] in [lo, hi) for all i, else *bad |= bit.  One pass, no host synchronisation.
template <typename IdxT>
__global__ void check_ids_kernel(const IdxT* __restrict__ ids, long long n, long long stride, long long lo, long long hi,
                                 int* __restrict__ bad, int bit) {
    int any = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const long long v = (long long)ids[i * stride];
        any |= (v < lo || v >= hi);
    }
    if (__any_sync(0xffffffffu, any) && (threadIdx.x & 31) == 0) atomicOr(bad, bit);
}

template <typename IdxT>
static int csr_from_padded_impl(const IdxT* idx, int rows, int width, int n_src, int* row_ptr, int* col,
                                int* bad_flag, cudaStream_t st) {
    const int T = 256;
    csr_count_kernel<IdxT><<<ceil_div(rows, T), T, 0, st>>>(idx, rows, width, row_ptr);
    scan_kernel<<<1, 1024, 0, st>>>(row_ptr, rows);
    csr_fill_kernel<IdxT><<<ceil_div(rows, T), T, 0, st>>>(idx, rows, width, n_src, row_ptr, col, bad_flag);
    return check_launch("csr_from_padded", 3);
}

}  // namespace ggpm

using namespace ggpm;

extern "C" int ggpm_csr_from_padded(const void* idx, int idx_is_64bit, int rows, int width, int n_src_rows,
                                    int32_t* row_ptr, int32_t* col, int32_t* bad_flag, void* stream) {
    GGPM_REQUIRE(rows > 0 && width >= 0 && n_src_rows > 0, "csr_from_padded: bad sizes rows=%d width=%d n_src=%d",
                 rows, width, n_src_rows);
    GGPM_REQUIRE(idx && row_ptr && col && bad_flag, "csr_from_padded: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (idx_is_64bit)
        return csr_from_padded_impl<long long>((const long long*)idx, rows, width, n_src_rows, row_ptr, col, bad_flag, st);
    return csr_from_padded_impl<int>((const int*)idx, rows, width, n_src_rows, row_ptr, col, bad_flag, st);
}

extern "C" int ggpm_check_ids(const void* ids, int ids_are_64bit, long long n, long long stride, long long lo, long long hi,
                              int32_t* bad_flag, int bit, void* stream) {
    GGPM_REQUIRE(bad_flag && (ids || n == 0) && n >= 0 && stride >= 1 && bit != 0, "check_ids: bad arguments");
    if (n == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 256;
    long long want = ceil_div_ll(n, T);
    const int blocks = (int)(want < 4LL * sm_count() ? want : 4LL * sm_count());
    if (ids_are_64bit)
        check_ids_kernel<long long><<<blocks, T, 0, st>>>((const long long*)ids, n, stride, lo, hi, bad_flag, bit);
    else
        check_ids_kernel<int><<<blocks, T, 0, st>>>((const int*)ids, n, stride, lo, hi, bad_flag, bit);
    return check_launch("check_ids");
}

extern "C" int ggpm_csr_transpose(const int32_t* row_ptr, const int32_t* col, int rows, int n_src_rows,
                                  int32_t* t_ptr, int32_t* t_row, int32_t* t_pos, int32_t* cursor, void* stream) {
    GGPM_REQUIRE(rows > 0 && n_src_rows > 0, "csr_transpose: bad sizes");
    GGPM_REQUIRE(row_ptr && col && t_ptr && t_row && t_pos && cursor, "csr_transpose: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 256;
    GGPM_CUDA(cudaMemsetAsync(t_ptr, 0, sizeof(int) * (size_t)(n_src_rows + 1), st));
    GGPM_CUDA(cudaMemsetAsync(cursor, 0, sizeof(int) * (size_t)n_src_rows, st));
    tr_hist_kernel<<<ceil_div(rows, T), T, 0, st>>>(row_ptr, col, rows, t_ptr);
    scan_kernel<<<1, 1024, 0, st>>>(t_ptr, n_src_rows);
    tr_fill_kernel<<<ceil_div(rows, T), T, 0, st>>>(row_ptr, col, rows, t_ptr, cursor, t_row, t_pos);
    tr_sort_kernel<<<ceil_div(n_src_rows, T), T, 0, st>>>(t_ptr, n_src_rows, t_row, t_pos);
    return check_launch("csr_transpose", 4);
}
