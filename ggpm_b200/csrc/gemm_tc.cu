// tcgen05 (5th-gen tensor core) GEMM with fp32-level accuracy by the 3xTF32 operand split, optional CSR
// gather-sum fused into the operand prologue, fused epilogue (beta / bias / activation / row-0 mask) or split-K
// atomic accumulation.
//
//   C[M,N] (+)= A(M,K) * B(K,N)          A(m,k) = A[m*sam + k*sak],  B(k,n) = B[k*sbk + n*sbn]
//
// Every nn.Linear on the encoder path runs through this: the gate projections of the graph LSTM / GRU
// (ggpm/rnn.py:13-16, 28-37, 69-72, 88-91) forward, their input gradients and their weight gradients.
// With `a_row_ptr/a_col` the A rows are S_e = sum_{n in N(e)} table[n] — index_select_ND(h, 0, bgraph).sum(1)
// of ggpm/rnn.py:26,86 — built in registers by the producer warps, never written to memory.
//
// Structure (one CTA = one 128 x BN output tile, 9 warps):
//   warps 0-7  producers: global -> registers (gather-sum) -> hi/lo tf32 split -> 128B-swizzled smem stage
//                         (K-major or MN-major canonical UMMA layouts, so no operand is ever transposed in memory);
//                         afterwards the same warps run the epilogue: TMEM -> registers -> smem transpose -> global
//   warp  8    allocates TMEM; one lane issues tcgen05.mma.kind::tf32 (3 MMAs per 8-deep K step:
//              lo*hi, hi*lo, hi*hi) into a fp32 accumulator of BN TMEM columns; tcgen05.commit releases stages.
// Operands pass through registers (not TMA) because the A operand is computed (gather-sum + split), and the
// split of B rides the same path.
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

namespace ggpm {

int gemm_engine();

namespace {

constexpr int BM = 128;
constexpr int BK = 32;                 // fp32 elements per stage = one 128-byte swizzle row
constexpr int PRODUCERS = 256;         // warps 0-7
constexpr int MMA_WARP = 8;            // warp 8
constexpr int EPI_WARP0 = 9;           // warps 9-12
constexpr int EPI_THREADS = 128;
constexpr int THREADS = PRODUCERS + 32 + EPI_THREADS;
constexpr int A_BYTES = BM * 128;
constexpr int SCR_LD = 36;              // epilogue transpose scratch: row stride in floats
constexpr int SCRATCH_BYTES = 4 * 32 * SCR_LD * 4;
enum { EPI_STORE = 0, EPI_ATOMIC = 1, EPI_GENERIC = 2, EPI_SCALAR = 3, EPI_SCALAR_ATOMIC = 4 };
constexpr int TMEM_COLS = 512;

// "second region" of a launch (index >= split_at along split_dim): no gather, A replaced by A2 (when given)
enum { SPLIT_NONE = 0, SPLIT_N = 1, SPLIT_K = 2, SPLIT_M = 3 };

struct TcParams {
    const float* A; long long lda;      // leading stride (the non-unit one)
    const float* A2; long long lda2;
    const float* B; long long ldb;
    float* C; int ldc;
    int M, N, K, nk;
    int split_dim, split_at;
    const int* a_row_ptr; const int* a_col;      // CSR gather-sum over A's row index (K-major A only)
    const int* b_row_ptr; const int* b_col;      // CSR gather-sum over B's k index (MN-major B only)
    float beta; const float* bias; int act; int zero_row0;
    int m_tiles, n_tiles, total_tiles, tiles_full;
    int split_s;       // > 0 (fewer tiles than SMs): every tile is cut along K into split_s equal parts, one CTA each
    int epi_mode;
    int seg_cap;       // stream-K segments are at most this many K chunks long (0: unbounded).  A tensor-core accumulator
                       // truncates every partial sum it holds, so its error grows with the number of MMA steps chained into it;
                       // capping the chain and adding the pieces in fp32 (the reduce-add epilogue) bounds it — DESIGN.md 4.2
};

template <int BN> struct Cfg {
    static constexpr int B_BYTES = BN * 128;
    static constexpr int STAGE = 2 * A_BYTES + 2 * B_BYTES;
    static constexpr int NSTAGE = (BN == 256) ? 2 : 3;
    static constexpr int NACC = TMEM_COLS / (2 * BN);          // accumulator sets (main + correction) in TMEM
    static constexpr int SMEM = NSTAGE * STAGE + SCRATCH_BYTES + 1024;
};

// Per-tile cache of the CSR neighbour lists of the (up to 4) rows a producer thread gathers for the A operand:
// keeps the row_ptr -> col -> data dependency chain out of the per-chunk loop.
constexpr int NBR_CACHE = 3;
struct RowNbrs {
    int jb[BM / 32], cnt[BM / 32];
    int id[BM / 32][NBR_CACHE];
};

__device__ __forceinline__ void load_row_nbrs(RowNbrs& rn, const int* __restrict__ rp, const int* __restrict__ col,
                                              int i0, int n_i, int tid) {
    const int rq = tid >> 3;
#pragma unroll
    for (int p = 0; p < BM / 32; ++p) {
        const int gi = i0 + p * 32 + rq;
        int b = 0, e = 0;
        if (gi < n_i) { b = __ldg(rp + gi); e = __ldg(rp + gi + 1); }
        rn.jb[p] = b;
        rn.cnt[p] = e - b;
    }
#pragma unroll
    for (int p = 0; p < BM / 32; ++p)
#pragma unroll
        for (int q = 0; q < NBR_CACHE; ++q) rn.id[p][q] = q < rn.cnt[p] ? __ldg(col + rn.jb[p] + q) : 0;
}

// K-contiguous operand: tile of ROWS rows x 32 k.  Phase 1 (prefetch): global -> registers, with the CSR
// gather-sum when `rn` is given.  Phase 2 (store): tf32 hi/lo split -> K-major SWIZZLE_128B smem (8-row groups
// of 1024 B, 16-byte chunks XOR-ed with the row).
template <int ROWS>
__device__ __forceinline__ void km_prefetch(float4 (&v)[ROWS / 32], const float* __restrict__ base, long long ld, int i0,
                                            int n_i, int k0, int k_end, bool gather, const RowNbrs& rn,
                                            const int* __restrict__ col, int tid) {
    const int c = tid & 7, rq = tid >> 3;
    const int k = k0 + 4 * c;
    constexpr int P = ROWS / 32;
    if (!gather) {
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int gi = i0 + p * 32 + rq;
            v[p] = (gi < n_i && k < k_end) ? ld4(base + (long long)gi * ld + k) : f4zero();
        }
    } else {
        const bool kin = k < k_end;
        // two rows at a time: 6 independent 128-bit loads in flight per thread, bounded register use
#pragma unroll
        for (int p0 = 0; p0 < P; p0 += 2) {
            float4 x[2][NBR_CACHE];
#pragma unroll
            for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                for (int q = 0; q < NBR_CACHE; ++q)
                    x[pp][q] = (kin && q < rn.cnt[p0 + pp]) ? ld4(base + (long long)rn.id[p0 + pp][q] * ld + k) : f4zero();
#pragma unroll
            for (int pp = 0; pp < 2; ++pp) {
                const int p = p0 + pp;
                float4 acc = x[pp][0];
#pragma unroll
                for (int q = 1; q < NBR_CACHE; ++q) acc += x[pp][q];             // slot order, like sum(dim=1)
                if (kin)
                    for (int j = NBR_CACHE; j < rn.cnt[p]; ++j)
                        acc += ld4(base + (long long)__ldg(col + rn.jb[p] + j) * ld + k);
                v[p] = acc;
            }
        }
    }
}

template <int ROWS>
__device__ __forceinline__ void km_store(const float4 (&v)[ROWS / 32], uint32_t s_hi, uint32_t s_lo, int tid) {
    const int c = tid & 7, rq = tid >> 3;
#pragma unroll
    for (int p = 0; p < ROWS / 32; ++p) {
        const int r = p * 32 + rq;
        const uint32_t off = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
        float4 hi, lo;
        tc::split4(v[p], hi, lo);
        tc::sts4(s_hi + off, hi);
        tc::sts4(s_lo + off, lo);
    }
}

// MN-contiguous operand (element (i,k) at base[k*ld + i]): tile of ROWS i x 32 k -> MN-major tf32 layout
// SWIZZLE_128B_BASE32B: atoms of 4 k-rows x 32 i (512 B) whose 32-byte chunks are XOR-ed with the k-row,
// i-blocks contiguous (LBO = 512 B), k-blocks at ROWS/32 * 512 B (SBO).  With (rp, col) every k-row is the CSR
// gather-sum of table rows (S = sum of neighbour states, recomputed for the weight gradient).
template <int ROWS>
__device__ __forceinline__ void mn_prefetch(float4 (&v)[ROWS / 32], const float* __restrict__ base, long long ld, int i0,
                                            int n_i, int k0, int k_end, const int* __restrict__ rp,
                                            const int* __restrict__ col, int tid) {
    constexpr int NB = ROWS / 32;
    const int c = tid & 7, q = tid >> 3;
    if (rp == nullptr) {
#pragma unroll
        for (int p = 0; p < NB; ++p) {
            const int pair = p * 32 + q;
            const int ib = pair % NB, k = pair / NB;
            const int gi = i0 + ib * 32 + 4 * c, gk = k0 + k;
            v[p] = (gi < n_i && gk < k_end) ? ld4(base + (long long)gk * ld + gi) : f4zero();
        }
    } else {
        // four (k-row, i-block) pairs at a time: 8 independent 128-bit loads in flight per thread
#pragma unroll
        for (int p0 = 0; p0 < NB; p0 += 4) {
            int jb[4], je[4], gi[4];
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
                const int pair = (p0 + pp) * 32 + q;
                const int ib = pair % NB, k = pair / NB;
                gi[pp] = i0 + ib * 32 + 4 * c;
                const int gk = k0 + k;
                const bool ok = gi[pp] < n_i && gk < k_end;
                jb[pp] = ok ? __ldg(rp + gk) : 0;
                je[pp] = ok ? __ldg(rp + gk + 1) : 0;
            }
            int n0[4], n1[4];
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
                n0[pp] = jb[pp] < je[pp] ? __ldg(col + jb[pp]) : 0;
                n1[pp] = jb[pp] + 1 < je[pp] ? __ldg(col + jb[pp] + 1) : 0;
            }
            float4 x0[4], x1[4];
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
                x0[pp] = jb[pp] < je[pp] ? ld4(base + (long long)n0[pp] * ld + gi[pp]) : f4zero();
                x1[pp] = jb[pp] + 1 < je[pp] ? ld4(base + (long long)n1[pp] * ld + gi[pp]) : f4zero();
            }
#pragma unroll
            for (int pp = 0; pp < 4; ++pp) {
                float4 acc = x0[pp] + x1[pp];
                for (int j = jb[pp] + 2; j < je[pp]; ++j) acc += ld4(base + (long long)__ldg(col + j) * ld + gi[pp]);
                v[p0 + pp] = acc;
            }
        }
    }
}

template <int ROWS>
__device__ __forceinline__ void mn_store(const float4 (&v)[ROWS / 32], uint32_t s_hi, uint32_t s_lo, int tid) {
    constexpr int NB = ROWS / 32;
    const int c = tid & 7, q = tid >> 3;
#pragma unroll
    for (int p = 0; p < NB; ++p) {
        const int pair = p * 32 + q;
        const int ib = pair % NB, k = pair / NB;
        const uint32_t off = (uint32_t)((k >> 2) * (NB * 512) + ib * 512 + (k & 3) * 128 + (((c >> 1) ^ (k & 3)) << 5) +
                                        ((c & 1) << 4));
        float4 hi, lo;
        tc::split4(v[p], hi, lo);
        tc::sts4(s_hi + off, hi);
        tc::sts4(s_lo + off, lo);
    }
}

template <int ROWS, bool KC>
__device__ __forceinline__ uint64_t operand_desc(uint32_t saddr, int kstep) {
    if (KC) return tc::make_sdesc(saddr + kstep * 32, 16, 1024, 2);
    return tc::make_sdesc(saddr + kstep * 2 * (ROWS / 32) * 512, 512, (ROWS / 32) * 512, 1);
}

// Work distribution of the persistent kernel.  Tiles [0, tiles_full) (a multiple of the grid size) go round-robin,
// whole, with a direct epilogue.  The remaining "tail" tiles — which would leave most SMs idle — are cut along K
// into 32-deep chunks and dealt out evenly (stream-K): every CTA accumulates its chunk range of a tail tile and adds
// it to C with vector reductions (C is pre-zeroed there, or holds the beta = 1 addend).
struct Seg { int m0, n0, kb, ke, nk; bool partial; };

template <int BN, int BKE = BK>
struct SegIter {
    int t_next, u, u_end;
    __device__ __forceinline__ explicit SegIter(const TcParams& p) {
        t_next = blockIdx.x;
        if (p.split_s > 0) {          // one K-slice of one tile per CTA: a single segment, a single epilogue
            const int t = blockIdx.x / p.split_s, part = blockIdx.x % p.split_s;
            u = t * p.nk + (int)((long long)p.nk * part / p.split_s);
            u_end = t * p.nk + (int)((long long)p.nk * (part + 1) / p.split_s);
            return;
        }
        const long long ut = (long long)(p.total_tiles - p.tiles_full) * p.nk;
        u = (int)(ut * blockIdx.x / gridDim.x);
        u_end = (int)(ut * (blockIdx.x + 1) / gridDim.x);
    }
    __device__ __forceinline__ bool next(const TcParams& p, Seg& sg) {
        int t, c0, c1;
        if (t_next < p.tiles_full) {
            t = t_next; c0 = 0; c1 = p.nk; sg.partial = false;
            t_next += gridDim.x;
        } else if (u < u_end) {
            t = p.tiles_full + u / p.nk;
            c0 = u % p.nk;
            c1 = min(p.nk, c0 + (u_end - u));
            if (p.seg_cap > 0) c1 = min(c1, c0 + p.seg_cap);
            u += c1 - c0;
            sg.partial = true;
        } else {
            return false;
        }
        sg.m0 = (t / p.n_tiles) * BM;
        sg.n0 = (t % p.n_tiles) * BN;
        sg.kb = c0 * BKE;
        sg.ke = min(p.K, c1 * BKE);
        sg.nk = c1 - c0;
        return true;
    }
};

// B_MODE: 0 = n-contiguous B through registers (MN-major), 1 = k-contiguous B through registers (K-major),
//         2 = k-contiguous B already split into tf32 hi/lo planes, loaded by TMA (weights)
template <int BN, bool A_KC, int B_MODE>
__global__ void __launch_bounds__(THREADS, 1) gemm_tc_kernel(const TcParams p, const __grid_constant__ CUtensorMap tm_hi,
                                                             const __grid_constant__ CUtensorMap tm_lo) {
    constexpr bool B_KC = B_MODE != 0;
    constexpr bool B_TMA = B_MODE == 2;
    using C = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * C::NSTAGE + 2 * C::NACC];
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t smem0 = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar0 = tc::smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (C::NSTAGE + s); };
    auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * C::NSTAGE + b); };
    auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * C::NSTAGE + C::NACC + b); };

    if (tid == 0) {
        for (int s = 0; s < C::NSTAGE; ++s) {
            tc::mbar_init(full_bar(s), PRODUCERS);
            tc::mbar_init(empty_bar(s), 1);
        }
        for (int b = 0; b < C::NACC; ++b) {
            tc::mbar_init(acc_full_bar(b), 1);
            tc::mbar_init(acc_empty_bar(b), EPI_THREADS);
        }
        tc::mbar_fence_init();
    }
    if (warp == MMA_WARP) tc::tmem_alloc<TMEM_COLS>(tc::smem_u32(&tmem_slot));
    if (B_TMA && tid == 32) { tc::tma_prefetch_desc(&tm_hi); tc::tma_prefetch_desc(&tm_lo); }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);

    if (warp < MMA_WARP) {
        // ---------------- producers: global -> (gather-sum) -> tf32 hi/lo -> swizzled smem ----------------
        // Software pipeline: the global loads of chunk i+1 are issued into registers right after chunk i is
        // handed to the tensor core, so their latency hides behind the wait for a free stage.
        float4 va[BM / 32], vb[BN / 32];
        RowNbrs rn;
        const bool a_gather = A_KC && p.a_row_ptr != nullptr;
        SegIter<BN> pit(p);
        Seg ptl;                                             // prefetch cursor: (segment, chunk)
        int pc = 0;
        bool more = pit.next(p, ptl);
        bool rn_valid = false;
        int st_k0 = 0, st_n0 = 0;                            // coordinates of the chunk held in registers
        auto prefetch = [&]() {
            const bool n_reg2 = p.split_dim == SPLIT_N && ptl.n0 >= p.split_at;
            const bool m_reg2 = p.split_dim == SPLIT_M && ptl.m0 >= p.split_at;
            const int k0 = ptl.kb + pc * BK;
            const bool k_reg2 = p.split_dim == SPLIT_K && k0 >= p.split_at;
            const bool reg2 = n_reg2 || m_reg2 || k_reg2;
            const float* Ab = (reg2 && p.A2 != nullptr) ? p.A2 : p.A;
            const long long lda = (reg2 && p.A2 != nullptr) ? p.lda2 : p.lda;
            int a_i0 = ptl.m0, a_ni = p.M, a_k0 = k0, a_ke = ptl.ke;
            if (p.split_dim == SPLIT_M) {
                if (m_reg2) { a_i0 = ptl.m0 - p.split_at; a_ni = p.M - p.split_at; }
                else a_ni = min(p.M, p.split_at);
            }
            if (p.split_dim == SPLIT_K) {
                if (k_reg2) { a_k0 = k0 - p.split_at; a_ke = ptl.ke - p.split_at; }
                else a_ke = min(ptl.ke, p.split_at);
            }
            st_k0 = k0; st_n0 = ptl.n0;
            if (B_TMA) {
            } else if (B_KC)
                km_prefetch<BN>(vb, p.B, p.ldb, ptl.n0, p.N, k0, ptl.ke, false, rn, nullptr, tid);
            else
                mn_prefetch<BN>(vb, p.B, p.ldb, ptl.n0, p.N, k0, ptl.ke, reg2 ? nullptr : p.b_row_ptr, p.b_col, tid);
            if (A_KC) {
                const bool g = a_gather && !reg2;
                if (g && !rn_valid) { load_row_nbrs(rn, p.a_row_ptr, p.a_col, a_i0, a_ni, tid); rn_valid = true; }
                km_prefetch<BM>(va, Ab, lda, a_i0, a_ni, a_k0, a_ke, g, rn, p.a_col, tid);
            } else {
                mn_prefetch<BM>(va, Ab, lda, a_i0, a_ni, a_k0, a_ke, nullptr, nullptr, tid);
            }
        };
        if (more) prefetch();
        int chunk = 0;
        while (more) {
            const int s = chunk % C::NSTAGE;
            const uint32_t ph = (uint32_t)(chunk / C::NSTAGE) & 1u;
            tc::mbar_wait(empty_bar(s), ph ^ 1u);
            const uint32_t sa = smem0 + s * C::STAGE;
            const uint32_t sb = sa + 2 * A_BYTES;
            if (B_TMA) {
                if (tid == 0) {                                  // weights: two bulk tensor copies straight into the stage
                    tc::mbar_expect_tx(full_bar(s), 2 * C::B_BYTES);
                    tc::tma_load_2d(sb, &tm_hi, full_bar(s), st_k0, st_n0);
                    tc::tma_load_2d(sb + C::B_BYTES, &tm_lo, full_bar(s), st_k0, st_n0);
                }
            }
            if (A_KC) km_store<BM>(va, sa, sa + A_BYTES, tid); else mn_store<BM>(va, sa, sa + A_BYTES, tid);
            if (!B_TMA) {
                if (B_KC) km_store<BN>(vb, sb, sb + C::B_BYTES, tid); else mn_store<BN>(vb, sb, sb + C::B_BYTES, tid);
            }
            tc::fence_async_smem();
            tc::mbar_arrive(full_bar(s));
            ++chunk;
            if (++pc == ptl.nk) {                                // next segment of this CTA
                pc = 0;
                rn_valid = false;
                more = pit.next(p, ptl);
            }
            if (more) prefetch();
        }
    } else if (warp == MMA_WARP) {
        // ---------------- MMA issuer: one lane, tcgen05.mma.kind::tf32 into TMEM ----------------
        if (lane == 0) {
            constexpr uint32_t idesc = tc::make_idesc_tf32(BM, BN, A_KC ? 0 : 1, B_KC ? 0 : 1);
            int chunk = 0, it = 0;
            SegIter<BN> sit(p);
            Seg tl;
            for (; sit.next(p, tl); ++it) {
                const int b = it % C::NACC;
                const uint32_t aph = (uint32_t)(it / C::NACC) & 1u;
                tc::mbar_wait(acc_empty_bar(b), aph ^ 1u);          // epilogue has drained this accumulator set
                tc::fence_after_sync();
                const uint32_t d_main = tmem + (uint32_t)(b * 2 * BN), d_corr = d_main + BN;
                for (int c = 0; c < tl.nk; ++c, ++chunk) {
                    const int s = chunk % C::NSTAGE;
                    const uint32_t ph = (uint32_t)(chunk / C::NSTAGE) & 1u;
                    tc::mbar_wait(full_bar(s), ph);
                    tc::fence_after_sync();
                    const uint32_t sa = smem0 + s * C::STAGE;
                    const uint32_t sb = sa + 2 * A_BYTES;
#pragma unroll
                    for (int j = 0; j < BK / 8; ++j) {
                        const uint64_t a_hi = operand_desc<BM, A_KC>(sa, j), a_lo = operand_desc<BM, A_KC>(sa + A_BYTES, j);
                        const uint64_t b_hi = operand_desc<BN, B_KC>(sb, j), b_lo = operand_desc<BN, B_KC>(sb + C::B_BYTES, j);
                        // the small cross terms go to their own accumulator: the main one then sees K/8 roundings only
                        tc::mma_tf32(d_corr, a_lo, b_hi, idesc, (c | j) != 0);
                        tc::mma_tf32(d_corr, a_hi, b_lo, idesc, 1u);
                        tc::mma_tf32(d_main, a_hi, b_hi, idesc, (c | j) != 0);
                    }
                    tc::mma_commit(empty_bar(s));
                }
                tc::mma_commit(acc_full_bar(b));
            }
        }
        __syncwarp();
    } else {
        // ---------------- epilogue warps: TMEM -> registers -> smem transpose -> coalesced global ----------------
        const int q = warp & 3;                                  // TMEM lane quarter this warp may read
        const int ew = warp - EPI_WARP0;
        float* scratch = reinterpret_cast<float*>(smem_raw + (smem0 - tc::smem_u32(smem_raw)) + C::NSTAGE * C::STAGE) +
                         ew * (32 * SCR_LD);
        int it = 0;
        SegIter<BN> sit(p);
        Seg tl;
        for (; sit.next(p, tl); ++it) {
            const int epi_mode = tl.partial ? (p.epi_mode == EPI_SCALAR ? EPI_SCALAR_ATOMIC : EPI_ATOMIC) : p.epi_mode;
            const int b = it % C::NACC;
            const uint32_t aph = (uint32_t)(it / C::NACC) & 1u;
            tc::mbar_wait(acc_full_bar(b), aph);
            tc::fence_after_sync();
            const uint32_t d_main = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * 2 * BN);
            const int row_base = tl.m0 + q * 32;
            const int rows_valid = min(32, p.M - row_base);                 // rows of this warp inside the matrix
            for (int g = 0; g < BN / 32; ++g) {
                const int col0 = g * 32;
                if (tl.n0 + col0 >= p.N) break;
                uint32_t v[32], w[32];
                tc::tmem_ld32(d_main + (uint32_t)col0, v);
                tc::tmem_ld32(d_main + (uint32_t)(BN + col0), w);
                tc::tmem_ld_wait();
                // transpose through smem (row stride 36 floats: conflict-free for 128-bit accesses both ways)
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    st4(scratch + lane * SCR_LD + 4 * j,
                        make_float4(__uint_as_float(v[4 * j]) + __uint_as_float(w[4 * j]),
                                    __uint_as_float(v[4 * j + 1]) + __uint_as_float(w[4 * j + 1]),
                                    __uint_as_float(v[4 * j + 2]) + __uint_as_float(w[4 * j + 2]),
                                    __uint_as_float(v[4 * j + 3]) + __uint_as_float(w[4 * j + 3])));
                __syncwarp();
                const int n = tl.n0 + col0 + 4 * (lane & 7);
                if (epi_mode != EPI_SCALAR && epi_mode != EPI_SCALAR_ATOMIC) {
                    // lanes 0-7 cover the 128 bytes of one row: 4 rows per instruction, fully coalesced
                    if (n < p.N) {
                        float4 bias4 = f4zero();
                        if (epi_mode == EPI_GENERIC && p.bias != nullptr) bias4 = ldg4(p.bias + n);
#pragma unroll 2
                        for (int r4 = 0; r4 < 8; ++r4) {
                            const int rr = r4 * 4 + (lane >> 3);
                            if (rr >= rows_valid) break;
                            float4 x = ld4(scratch + rr * SCR_LD + 4 * (lane & 7));
                            float* cp = p.C + (long long)(row_base + rr) * p.ldc + n;
                            if (epi_mode == EPI_STORE) {
                                st4(cp, x);
                            } else if (epi_mode == EPI_ATOMIC) {
                                red_add4(cp, x);
                            } else {
                                if (p.beta != 0.f) {
                                    const float4 o = ld4(cp);
                                    x.x += p.beta * o.x; x.y += p.beta * o.y; x.z += p.beta * o.z; x.w += p.beta * o.w;
                                }
                                x += bias4;
                                if (p.act == GGPM_ACT_RELU) x = make_float4(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f), fmaxf(x.z, 0.f), fmaxf(x.w, 0.f));
                                else if (p.act == GGPM_ACT_TANH) x = tanh4(x);
                                else if (p.act == GGPM_ACT_SIGMOID) x = sigmoid4(x);
                                if (p.zero_row0 && row_base + rr == 0) x = f4zero();
                                st4(cp, x);
                            }
                        }
                    }
                } else {
                    // ragged N or unaligned C: one column per lane, one row per iteration
                    const int ns = tl.n0 + col0 + lane;
                    const float bias = (p.bias != nullptr && ns < p.N) ? __ldg(p.bias + ns) : 0.f;
                    for (int rr = 0; rr < rows_valid; ++rr) {
                        if (ns >= p.N) break;
                        float x = scratch[rr * SCR_LD + lane];
                        float* cp = p.C + (long long)(row_base + rr) * p.ldc + ns;
                        if (epi_mode == EPI_SCALAR_ATOMIC) {
                            atomicAdd(cp, x);
                        } else {
                            if (p.beta != 0.f) x += p.beta * (*cp);
                            x += bias;
                            if (p.act == GGPM_ACT_RELU) x = fmaxf(x, 0.f);
                            else if (p.act == GGPM_ACT_TANH) x = tanhf(x);
                            else if (p.act == GGPM_ACT_SIGMOID) x = sigmoidf_(x);
                            if (p.zero_row0 && row_base + rr == 0) x = 0.f;
                            *cp = x;
                        }
                    }
                }
                __syncwarp();
            }
            tc::fence_before_sync();
            tc::mbar_arrive(acc_empty_bar(b));
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == MMA_WARP) tc::tmem_dealloc<TMEM_COLS>(tmem);
}

// ====================================================================================================
// fp16 two-term split engine ("f16x3"): same structure as gemm_tc_kernel<BN, true, 2>, but operands are
// fp16 hi / lo pairs of power-of-two scaled values and the MMAs are kind::f16 (K = 16 per instruction, twice
// the tf32 rate, half the shared-memory bytes per FLOP).  A: fp32 K-major rows through registers (optional CSR
// gather-sum, optional second region), B: pre-split fp16 weight planes by TMA.  One stage = 64 k.
// ====================================================================================================
constexpr int BK16 = 64;

struct F16Params {
    TcParams t;
    const float* a_amax; int a_amax_n; float a_amax_const;
    const int* a_deg;
    const float* a_scale;                  // when given: the A scale itself (A already is fp16 planes)
    const float* b_scale; float b_scale_const;
    // GGPM_ACT_LSTM_GATES epilogue: the hoisted x-projection tables (see GemmArgs)
    const float* gx_Yn; const int* gx_src; const float* gx_T; const int* gx_k1; const int* gx_k2; int gx_ld, gx_hp;
#ifdef GGPM_TRACE
    unsigned long long* dbg;               // [gridDim.x][8] globaltimer stamps; only in -DGGPM_TRACE builds (tools/f16_trace.py)
#endif
};

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
#ifdef GGPM_TRACE
#define F16_STAMP(i) do { if (fp.dbg) fp.dbg[blockIdx.x * 8 + (i)] = gtimer(); } while (0)
#else
#define F16_STAMP(i) do { } while (0)       // the stamp hook is compiled out of production builds
#endif

// modes of gemm_f16_kernel
enum { F16_A_REGS = 0,      // A: fp32 K-major rows through registers (gather-sum, scaling, split); B: K-major planes by TMA
       F16_A_TMA = 1,       // A and B: K-major fp16 hi / lo planes by TMA (no producer arithmetic at all)
       F16_TN_TMA = 2 };    // A(m,k) and B(k,n) both stored k-row-major (MN-major operands), planes by TMA: weight gradients

__device__ __forceinline__ float f16_b_scale(const F16Params& p) {
    return p.b_scale != nullptr ? __ldg(p.b_scale) : p.b_scale_const;
}

__device__ __forceinline__ float f16_a_scale(const F16Params& p) {
    if (p.a_scale != nullptr) return __ldg(p.a_scale);
    float amax = p.a_amax_const;
    if (p.a_amax != nullptr) {
        amax = 0.f;
        for (int i = 0; i < p.a_amax_n; ++i) amax = fmaxf(amax, __ldg(p.a_amax + i));
    }
    const int deg = p.a_deg != nullptr ? max(1, __ldg(p.a_deg)) : 1;
    return tc::pow2_scale(amax * (float)deg);
}

// K-contiguous fp32 operand, tile of ROWS rows x 64 k: thread (rq = tid >> 3, c = tid & 7) holds k = k0 + 8c .. +7
// of rows p*32 + rq.
template <int ROWS>
__device__ __forceinline__ void km_prefetch16(float4 (&v)[ROWS / 32][2], const float* __restrict__ base, long long ld,
                                              int i0, int n_i, int k0, int k_end, bool gather, const RowNbrs& rn,
                                              const int* __restrict__ col, int tid) {
    const int c = tid & 7, rq = tid >> 3;
    const int k = k0 + 8 * c;
    const bool kin0 = k < k_end, kin1 = k + 4 < k_end;
    constexpr int P = ROWS / 32;
    if (!gather) {
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int gi = i0 + p * 32 + rq;
            const float* r = base + (long long)gi * ld + k;
            v[p][0] = (gi < n_i && kin0) ? ld4(r) : f4zero();
            v[p][1] = (gi < n_i && kin1) ? ld4(r + 4) : f4zero();
        }
    } else {
        // two rows at a time: 12 independent 128-bit loads in flight per thread
#pragma unroll
        for (int p0 = 0; p0 < P; p0 += 2) {
            float4 x[2][NBR_CACHE][2];
#pragma unroll
            for (int pp = 0; pp < 2; ++pp)
#pragma unroll
                for (int q = 0; q < NBR_CACHE; ++q) {
                    const bool ok = q < rn.cnt[p0 + pp];
                    const float* r = base + (long long)rn.id[p0 + pp][q] * ld + k;
                    x[pp][q][0] = (ok && kin0) ? ld4(r) : f4zero();
                    x[pp][q][1] = (ok && kin1) ? ld4(r + 4) : f4zero();
                }
#pragma unroll
            for (int pp = 0; pp < 2; ++pp) {
                const int p = p0 + pp;
                float4 a0 = x[pp][0][0], a1 = x[pp][0][1];
#pragma unroll
                for (int q = 1; q < NBR_CACHE; ++q) { a0 += x[pp][q][0]; a1 += x[pp][q][1]; }   // slot order
                if (kin0)
                    for (int j = NBR_CACHE; j < rn.cnt[p]; ++j) {
                        const float* r = base + (long long)__ldg(col + rn.jb[p] + j) * ld + k;
                        a0 += ld4(r);
                        if (kin1) a1 += ld4(r + 4);
                    }
                v[p][0] = a0;
                v[p][1] = a1;
            }
        }
    }
}

// scale, split into fp16 hi / lo, store to the K-major SWIZZLE_128B stage (rows of 128 B = 64 fp16)
template <int ROWS>
__device__ __forceinline__ void km_store16(const float4 (&v)[ROWS / 32][2], uint32_t s_hi, uint32_t s_lo, float scale,
                                           int tid) {
    const int c = tid & 7, rq = tid >> 3;
#pragma unroll
    for (int p = 0; p < ROWS / 32; ++p) {
        const int r = p * 32 + rq;
        const uint32_t off = (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c ^ (r & 7)) << 4));
        uint4 hi, lo;
        tc::split8_f16(v[p][0], v[p][1], scale, hi, lo);
        tc::sts4u(s_hi + off, hi);
        tc::sts4u(s_lo + off, lo);
    }
}

// shared-memory descriptor of the 16-deep K step `j` of one fp16 plane of a stage.
//   K-major (rows of 64 k = 128 B, 8-row groups of 1024 B): the step advances 32 B inside the swizzled row.
//   MN-major (TN): swizzle atoms of 64 mn x 8 k (1024 B); a TMA box lays the 8 k-atoms of one 64-wide mn block
//   contiguously (8192 B), mn blocks one after the other: LBO (mn-atom stride) = 8192, SBO (k-atom stride) = 1024,
//   and a step (16 k = 2 atoms) advances 2048 B.
template <bool TN>
__device__ __forceinline__ uint64_t f16_desc(uint32_t saddr, int kstep) {
    if (TN) return tc::make_sdesc(saddr + kstep * 2048, 8192, 1024, 2);
    return tc::make_sdesc(saddr + kstep * 32, 16, 1024, 2);
}

template <int BN, int MODE>
__global__ void __launch_bounds__(THREADS, 1) gemm_f16_kernel(const F16Params fp, const __grid_constant__ CUtensorMap tm_hi,
                                                              const __grid_constant__ CUtensorMap tm_lo,
                                                              const __grid_constant__ CUtensorMap ta_hi,
                                                              const __grid_constant__ CUtensorMap ta_lo,
                                                              const __grid_constant__ CUtensorMap tc_out) {
    using C = Cfg<BN>;
    constexpr bool TN = MODE == F16_TN_TMA;
    if (threadIdx.x == 0) F16_STAMP(0);
    using Iter = SegIter<BN, BK16>;
    const TcParams& p = fp.t;
    extern __shared__ uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t bars[2 * C::NSTAGE + 2 * C::NACC];
    __shared__ uint32_t tmem_slot;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t smem0 = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar0 = tc::smem_u32(bars);
    auto full_bar = [&](int s) { return bar0 + 8u * s; };
    auto empty_bar = [&](int s) { return bar0 + 8u * (C::NSTAGE + s); };
    auto acc_full_bar = [&](int b) { return bar0 + 8u * (2 * C::NSTAGE + b); };
    auto acc_empty_bar = [&](int b) { return bar0 + 8u * (2 * C::NSTAGE + C::NACC + b); };

    if (tid == 0) {
        for (int s = 0; s < C::NSTAGE; ++s) {
            tc::mbar_init(full_bar(s), MODE == F16_A_REGS ? PRODUCERS : 1);
            tc::mbar_init(empty_bar(s), 1);
        }
        for (int b = 0; b < C::NACC; ++b) {
            tc::mbar_init(acc_full_bar(b), 1);
            tc::mbar_init(acc_empty_bar(b), EPI_THREADS);
        }
        tc::mbar_fence_init();
    }
    if (warp == MMA_WARP) tc::tmem_alloc<TMEM_COLS>(tc::smem_u32(&tmem_slot));
    if (tid == 32) {
        tc::tma_prefetch_desc(&tm_hi); tc::tma_prefetch_desc(&tm_lo);
        if (MODE != F16_A_REGS) { tc::tma_prefetch_desc(&ta_hi); tc::tma_prefetch_desc(&ta_lo); }
    }
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(&tmem_slot);
    pdl_wait();                       // setup above overlapped the previous grid's tail; its results are visible from here
    pdl_launch_dependents();
    if (threadIdx.x == 0) F16_STAMP(1);

    if (MODE != F16_A_REGS && warp < MMA_WARP) {
        // ---------------- TMA-only producer: one thread streams both operands' planes into the stages ----------------
        if (tid == 0) {
            Iter pit(p);
            Seg tl;
            int chunk = 0;
            while (pit.next(p, tl)) {
                for (int c = 0; c < tl.nk; ++c, ++chunk) {
                    const int s = chunk % C::NSTAGE;
                    const uint32_t ph = (uint32_t)(chunk / C::NSTAGE) & 1u;
                    tc::mbar_wait(empty_bar(s), ph ^ 1u);
                    const uint32_t sa = smem0 + s * C::STAGE;
                    const uint32_t sb = sa + 2 * A_BYTES;
                    const int k0 = tl.kb + c * BK16;
                    tc::mbar_expect_tx(full_bar(s), 2 * A_BYTES + 2 * C::B_BYTES);
                    if (TN) {
                        // 3-D maps (64 mn, k rows, mn / 64): one box = all swizzle atoms of the tile, atom-major in smem
                        tc::tma_load_3d(sa, &ta_hi, full_bar(s), 0, k0, tl.m0 / 64);
                        tc::tma_load_3d(sa + A_BYTES, &ta_lo, full_bar(s), 0, k0, tl.m0 / 64);
                        tc::tma_load_3d(sb, &tm_hi, full_bar(s), 0, k0, tl.n0 / 64);
                        tc::tma_load_3d(sb + C::B_BYTES, &tm_lo, full_bar(s), 0, k0, tl.n0 / 64);
                    } else {
                        tc::tma_load_2d(sa, &ta_hi, full_bar(s), k0, tl.m0);
                        tc::tma_load_2d(sa + A_BYTES, &ta_lo, full_bar(s), k0, tl.m0);
                        tc::tma_load_2d(sb, &tm_hi, full_bar(s), k0, tl.n0);
                        tc::tma_load_2d(sb + C::B_BYTES, &tm_lo, full_bar(s), k0, tl.n0);
                    }
                    tc::mbar_arrive(full_bar(s));
                }
            }
        }
    } else if (warp < MMA_WARP) {
        // ---------------- producers: global -> (gather-sum) -> scaled fp16 hi/lo -> swizzled smem ----------------
        float4 va[BM / 32][2];
        RowNbrs rn;
        const float a_scale = f16_a_scale(fp);
        const bool a_gather = p.a_row_ptr != nullptr;
        Iter pit(p);
        Seg ptl;
        int pc = 0;
        bool more = pit.next(p, ptl);
        bool rn_valid = false;
        int st_k0 = 0, st_n0 = 0;
        auto prefetch = [&]() {
            const bool n_reg2 = p.split_dim == SPLIT_N && ptl.n0 >= p.split_at;
            const int k0 = ptl.kb + pc * BK16;
            const bool k_reg2 = p.split_dim == SPLIT_K && k0 >= p.split_at;
            const bool reg2 = n_reg2 || k_reg2;
            const float* Ab = (reg2 && p.A2 != nullptr) ? p.A2 : p.A;
            const long long lda = (reg2 && p.A2 != nullptr) ? p.lda2 : p.lda;
            int a_k0 = k0, a_ke = ptl.ke;
            if (p.split_dim == SPLIT_K) {
                if (k_reg2) { a_k0 = k0 - p.split_at; a_ke = ptl.ke - p.split_at; }
                else a_ke = min(ptl.ke, p.split_at);
            }
            st_k0 = k0; st_n0 = ptl.n0;
            const bool g = a_gather && !reg2;
            if (g && !rn_valid) { load_row_nbrs(rn, p.a_row_ptr, p.a_col, ptl.m0, p.M, tid); rn_valid = true; }
            km_prefetch16<BM>(va, Ab, lda, ptl.m0, p.M, a_k0, a_ke, g, rn, p.a_col, tid);
        };
        if (more) prefetch();
        int chunk = 0;
        while (more) {
            const int s = chunk % C::NSTAGE;
            const uint32_t ph = (uint32_t)(chunk / C::NSTAGE) & 1u;
            tc::mbar_wait(empty_bar(s), ph ^ 1u);
            const uint32_t sa = smem0 + s * C::STAGE;
            const uint32_t sb = sa + 2 * A_BYTES;
            if (tid == 0) {                                      // weights: two bulk tensor copies straight into the stage
                tc::mbar_expect_tx(full_bar(s), 2 * C::B_BYTES);
                tc::tma_load_2d(sb, &tm_hi, full_bar(s), st_k0, st_n0);
                tc::tma_load_2d(sb + C::B_BYTES, &tm_lo, full_bar(s), st_k0, st_n0);
            }
            km_store16<BM>(va, sa, sa + A_BYTES, a_scale, tid);
            tc::fence_async_smem();
            tc::mbar_arrive(full_bar(s));
            ++chunk;
            if (++pc == ptl.nk) {
                pc = 0;
                rn_valid = false;
                more = pit.next(p, ptl);
            }
            if (more) prefetch();
        }
    } else if (warp == MMA_WARP) {
        // ---------------- MMA issuer: one lane, tcgen05.mma.kind::f16 into TMEM ----------------
        if (lane == 0) {
            constexpr uint32_t idesc = tc::make_idesc_f16(BM, BN, TN ? 1 : 0, TN ? 1 : 0);
            int chunk = 0, it = 0;
            Iter sit(p);
            Seg tl;
            for (; sit.next(p, tl); ++it) {
                const int b = it % C::NACC;
                const uint32_t aph = (uint32_t)(it / C::NACC) & 1u;
                tc::mbar_wait(acc_empty_bar(b), aph ^ 1u);
                tc::fence_after_sync();
                const uint32_t d_main = tmem + (uint32_t)(b * 2 * BN), d_corr = d_main + BN;
                for (int c = 0; c < tl.nk; ++c, ++chunk) {
                    const int s = chunk % C::NSTAGE;
                    const uint32_t ph = (uint32_t)(chunk / C::NSTAGE) & 1u;
                    tc::mbar_wait(full_bar(s), ph);
                    tc::fence_after_sync();
                    if (chunk == 0) F16_STAMP(2);
                    const uint32_t sa = smem0 + s * C::STAGE;
                    const uint32_t sb = sa + 2 * A_BYTES;
#pragma unroll
                    for (int j = 0; j < BK16 / 16; ++j) {          // 16 fp16 = 32 bytes along the swizzled row
                        const uint64_t a_hi = f16_desc<TN>(sa, j), a_lo = f16_desc<TN>(sa + A_BYTES, j);
                        const uint64_t b_hi = f16_desc<TN>(sb, j), b_lo = f16_desc<TN>(sb + C::B_BYTES, j);
                        tc::mma_f16(d_corr, a_lo, b_hi, idesc, (c | j) != 0);
                        tc::mma_f16(d_corr, a_hi, b_lo, idesc, 1u);
                        tc::mma_f16(d_main, a_hi, b_hi, idesc, (c | j) != 0);
                    }
                    tc::mma_commit(empty_bar(s));
                }
                tc::mma_commit(acc_full_bar(b));
            }
            F16_STAMP(3);
        }
        __syncwarp();
    } else {
        // ---------------- epilogue warps: TMEM -> registers (x 1/scales) -> smem transpose -> coalesced global ----
        const int q = warp & 3;
        const int ew = warp - EPI_WARP0;
        const float inv = 1.0f / (f16_a_scale(fp) * f16_b_scale(fp));
        float* scratch = reinterpret_cast<float*>(smem_raw + (smem0 - tc::smem_u32(smem_raw)) + C::NSTAGE * C::STAGE) +
                         ew * (32 * SCR_LD);
        int it = 0;
        int boxes_issued = 0;
        Iter sit(p);
        Seg tl;
        for (; sit.next(p, tl); ++it) {
            const int epi_mode = tl.partial ? EPI_ATOMIC : p.epi_mode;
            const int b = it % C::NACC;
            const uint32_t aph = (uint32_t)(it / C::NACC) & 1u;
            tc::mbar_wait(acc_full_bar(b), aph);
            tc::fence_after_sync();
            if (warp == EPI_WARP0 && lane == 0) F16_STAMP(4);
            const uint32_t d_main = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * 2 * BN);
            const int row_base = tl.m0 + q * 32;
            const int rows_valid = min(32, p.M - row_base);
            if (epi_mode != EPI_GENERIC) {
                // Plain store / accumulate: TMEM -> registers (x 1/scales) -> 128B-swizzled smem box of 32 rows x 32
                // columns -> ONE bulk tensor store (or fp32 reduce-add: stream-K partial sums and beta = 1) per box.
                // Rows / columns beyond the matrix are clipped by the TMA unit.  Two boxes per warp alternate; the
                // TMEM loads of the next column group are issued before the current one is written out.
                const bool reduce = epi_mode == EPI_ATOMIC || p.beta != 0.f;
                const uint32_t box0 = smem0 + C::NSTAGE * C::STAGE + (uint32_t)ew * 8192u;
                constexpr int NU = BN / 16;                      // units of 16 columns; two units fill one box
                uint32_t v[2][16], w[2][16];
                int nu = 0;
                while (nu < NU && tl.n0 + nu * 16 < p.N) ++nu;   // N % 4 == 0; columns beyond N are clipped by the store
                nu = (nu + 1) & ~1;
                if (rows_valid > 0 && nu > 0) {
                    tc::tmem_ld16(d_main, v[0]);
                    tc::tmem_ld16(d_main + (uint32_t)BN, w[0]);
#pragma unroll
                    for (int u = 0; u < NU; ++u) {
                        if (u < nu) {
                            tc::tmem_ld_wait();
                            if (u + 1 < nu) {
                                tc::tmem_ld16(d_main + (uint32_t)((u + 1) * 16), v[(u + 1) & 1]);
                                tc::tmem_ld16(d_main + (uint32_t)(BN + (u + 1) * 16), w[(u + 1) & 1]);
                            }
                            const uint32_t box = box0 + (uint32_t)(boxes_issued & 1) * 4096u;
                            if ((u & 1) == 0 && boxes_issued >= 2) {     // the box used two stores ago must have been read out
                                if (lane == 0) tc::bulk_wait_read<1>();
                                __syncwarp();
                            }
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                const int jj = (u & 1) * 4 + j;          // 16-byte chunk of the 128-byte box row
                                const uint32_t off = (uint32_t)(lane * 128 + ((jj ^ (lane & 7)) << 4));
                                tc::sts4(box + off,
                                         make_float4((__uint_as_float(v[u & 1][4 * j]) + __uint_as_float(w[u & 1][4 * j])) * inv,
                                                     (__uint_as_float(v[u & 1][4 * j + 1]) + __uint_as_float(w[u & 1][4 * j + 1])) * inv,
                                                     (__uint_as_float(v[u & 1][4 * j + 2]) + __uint_as_float(w[u & 1][4 * j + 2])) * inv,
                                                     (__uint_as_float(v[u & 1][4 * j + 3]) + __uint_as_float(w[u & 1][4 * j + 3])) * inv));
                            }
                            if (u & 1) {
                                tc::fence_async_smem();
                                __syncwarp();
                                if (lane == 0) {
                                    if (reduce) tc::tma_reduce_add_2d(&tc_out, box, tl.n0 + (u >> 1) * 32, row_base);
                                    else tc::tma_store_2d(&tc_out, box, tl.n0 + (u >> 1) * 32, row_base);
                                    tc::bulk_commit();
                                }
                                ++boxes_issued;
                            }
                        }
                    }
                }
                tc::fence_before_sync();
                tc::mbar_arrive(acc_empty_bar(b));
                continue;
            }
            for (int g = 0; g < BN / 32; ++g) {
                const int col0 = g * 32;
                if (tl.n0 + col0 >= p.N) break;
                uint32_t v[32], w[32];
                tc::tmem_ld32(d_main + (uint32_t)col0, v);
                tc::tmem_ld32(d_main + (uint32_t)(BN + col0), w);
                tc::tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    st4(scratch + lane * SCR_LD + 4 * j,
                        make_float4((__uint_as_float(v[4 * j]) + __uint_as_float(w[4 * j])) * inv,
                                    (__uint_as_float(v[4 * j + 1]) + __uint_as_float(w[4 * j + 1])) * inv,
                                    (__uint_as_float(v[4 * j + 2]) + __uint_as_float(w[4 * j + 2])) * inv,
                                    (__uint_as_float(v[4 * j + 3]) + __uint_as_float(w[4 * j + 3])) * inv));
                __syncwarp();
                const int n = tl.n0 + col0 + 4 * (lane & 7);
                if (n < p.N) {
                    float4 bias4 = f4zero();
                    if (epi_mode == EPI_GENERIC && p.bias != nullptr) bias4 = ldg4(p.bias + n);
#pragma unroll 2
                    for (int r4 = 0; r4 < 8; ++r4) {
                        const int rr = r4 * 4 + (lane >> 3);
                        if (rr >= rows_valid) break;
                        float4 x = ld4(scratch + rr * SCR_LD + 4 * (lane & 7));
                        float* cp = p.C + (long long)(row_base + rr) * p.ldc + n;
                        if (epi_mode == EPI_STORE) {
                            st4(cp, x);
                        } else if (epi_mode == EPI_ATOMIC) {
                            red_add4(cp, x);
                        } else {
                            if (p.beta != 0.f) {
                                const float4 o = ld4(cp);
                                x.x += p.beta * o.x; x.y += p.beta * o.y; x.z += p.beta * o.z; x.w += p.beta * o.w;
                            }
                            x += bias4;
                            if (p.act == GGPM_ACT_LSTM_GATES) {
                                // gate non-linearities and padding mask of the graph LSTM in the GEMM epilogue
                                // (ggpm/rnn.py:88-89, 91, 106): add the hoisted x-projection of this message, sigmoid
                                // for the input / output gates, tanh for the candidate
                                const int row = row_base + rr;
                                float4 xv = ldg4(fp.gx_Yn + (size_t)__ldg(fp.gx_src + row) * fp.gx_ld + n) +
                                            ldg4(fp.gx_T + (size_t)__ldg(fp.gx_k1 + row) * fp.gx_ld + n);
                                if (fp.gx_k2 != nullptr) xv += ldg4(fp.gx_T + (size_t)__ldg(fp.gx_k2 + row) * fp.gx_ld + n);
                                x = xv + x;
                                x = n < 2 * fp.gx_hp ? sigmoid4(x) : tanh4(x);
                                if (row == 0) x = f4zero();
                            }
                            else if (p.act == GGPM_ACT_RELU) x = make_float4(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f), fmaxf(x.z, 0.f), fmaxf(x.w, 0.f));
                            else if (p.act == GGPM_ACT_TANH) x = tanh4(x);
                            else if (p.act == GGPM_ACT_SIGMOID) x = sigmoid4(x);
                            if (p.zero_row0 && row_base + rr == 0) x = f4zero();
                            st4(cp, x);
                        }
                    }
                }
                __syncwarp();
            }
            tc::fence_before_sync();
            tc::mbar_arrive(acc_empty_bar(b));
        }
        // every bulk store / reduce of this warp has been performed before the CTA retires: the next grid may already be
        // resident (programmatic dependent launch) and reads C right after its griddepcontrol.wait
        if (lane == 0) tc::bulk_wait_all();
        if (warp == EPI_WARP0 && lane == 0) F16_STAMP(5);
    }
    tc::fence_before_sync();
    __syncthreads();
    if (threadIdx.x == 0) F16_STAMP(6);
    if (warp == MMA_WARP) tc::tmem_dealloc<TMEM_COLS>(tmem);
    if (warp == MMA_WARP && lane == 0) F16_STAMP(7);
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

typedef CUresult (*TmapEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

TmapEncodeFn tmap_encoder() {
    static TmapEncodeFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (TmapEncodeFn)ptr;
    }
    return fn;
}

// fp32 [rows, cols] row-major plane (leading stride ld floats), box = 32 columns (one 128-byte swizzle row) x box_rows
int make_plane_tmap(CUtensorMap* tm, const float* base, int rows, int cols, long long ld, int box_rows) {
    TmapEncodeFn enc = tmap_encoder();
    GGPM_REQUIRE(enc != nullptr, "gemm_tcgen05: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * sizeof(float)};
    cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    GGPM_REQUIRE(r == CUDA_SUCCESS, "gemm_tcgen05: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}

// cuTensorMapEncodeTiled costs about a microsecond and the same maps recur in every round of a layer call (weight
// planes, Z planes, gradient outputs): a small per-thread cache keyed by everything that goes into the encoding
struct TmapKey {
    const void* base; long long a, b, c, ld; int kind;
    bool operator==(const TmapKey& o) const {
        return base == o.base && a == o.a && b == o.b && c == o.c && ld == o.ld && kind == o.kind;
    }
};
constexpr int kTmapCache = 32;
struct TmapCache {
    TmapKey key[kTmapCache];
    CUtensorMap map[kTmapCache];
    int n = 0, next = 0;
    const CUtensorMap* find(const TmapKey& k) const {
        for (int i = 0; i < n; ++i)
            if (key[i] == k) return &map[i];
        return nullptr;
    }
    void put(const TmapKey& k, const CUtensorMap& m) {
        const int i = n < kTmapCache ? n++ : (next = (next + 1) % kTmapCache);
        key[i] = k;
        map[i] = m;
    }
};
static thread_local TmapCache g_tmaps;

// fp16 [rows, cols] row-major plane (leading stride ld halves), box = 64 columns (one 128-byte swizzle row) x box_rows
int make_plane_tmap16(CUtensorMap* tm, const void* base, int rows, int cols, long long ld, int box_rows) {
    const TmapKey key{base, rows, cols, box_rows, ld, 1};
    if (const CUtensorMap* hit = g_tmaps.find(key)) { *tm = *hit; return 0; }
    TmapEncodeFn enc = tmap_encoder();
    GGPM_REQUIRE(enc != nullptr, "gemm_f16: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)BK16, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    GGPM_REQUIRE(r == CUDA_SUCCESS, "gemm_f16: cuTensorMapEncodeTiled failed (%d)", (int)r);
    g_tmaps.put(key, *tm);
    return 0;
}

// 3-D view of a k-row-major fp16 plane [k_rows, mn] (row stride ld halves) for MN-major operand tiles:
// dims (64 mn, k rows, mn / 64), box (64, 64, mn_blocks): one box = mn_blocks swizzle-atom columns of 64 k rows each
int make_plane_tmap16_tn(CUtensorMap* tm, const void* base, int k_rows, int mn, long long ld, int mn_blocks) {
    const TmapKey key{base, k_rows, mn, mn_blocks, ld, 2};
    if (const CUtensorMap* hit = g_tmaps.find(key)) { *tm = *hit; return 0; }
    TmapEncodeFn enc = tmap_encoder();
    GGPM_REQUIRE(enc != nullptr, "gemm_f16: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[3] = {64, (cuuint64_t)k_rows, (cuuint64_t)(mn / 64)};
    cuuint64_t strides[2] = {(cuuint64_t)ld * 2, 128};
    cuuint32_t box[3] = {64, (cuuint32_t)BK16, (cuuint32_t)mn_blocks};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    GGPM_REQUIRE(r == CUDA_SUCCESS, "gemm_f16: cuTensorMapEncodeTiled (3-D) failed (%d)", (int)r);
    g_tmaps.put(key, *tm);
    return 0;
}

// fp32 C [M, N] (row stride ldc floats) for the epilogue's bulk stores: boxes of 32 columns (128 B, swizzled) x 32 rows
int make_out_tmap(CUtensorMap* tm, const float* base, int M, int N, long long ldc) {
    const TmapKey key{base, M, N, 0, ldc, 3};
    if (const CUtensorMap* hit = g_tmaps.find(key)) { *tm = *hit; return 0; }
    TmapEncodeFn enc = tmap_encoder();
    GGPM_REQUIRE(enc != nullptr, "gemm_f16: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)M};
    cuuint64_t strides[1] = {(cuuint64_t)ldc * sizeof(float)};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    GGPM_REQUIRE(r == CUDA_SUCCESS, "gemm_f16: cuTensorMapEncodeTiled (output) failed (%d)", (int)r);
    g_tmaps.put(key, *tm);
    return 0;
}

constexpr int F16_EPI_BYTES = 4 * 2 * 4096;      // per epilogue warp: two 32 x 32 fp32 boxes
template <int BN> constexpr int f16_smem() { return Cfg<BN>::NSTAGE * Cfg<BN>::STAGE + F16_EPI_BYTES + 1024; }

template <int BN, int MODE>
int launch_f16_variant(const F16Params& p, dim3 grid, const CUtensorMap& tm_hi, const CUtensorMap& tm_lo,
                       const CUtensorMap& ta_hi, const CUtensorMap& ta_lo, const CUtensorMap& tc_out, cudaStream_t st) {
    static bool configured = false;
    auto kern = gemm_f16_kernel<BN, MODE>;
    if (!configured) {
        GGPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, f16_smem<BN>()));
        configured = true;
    }
    launch_pdl(kern, grid, dim3(THREADS), (size_t)f16_smem<BN>(), st, p, tm_hi, tm_lo, ta_hi, ta_lo, tc_out);
    return check_launch("gemm_f16x3");
}

template <int BN, bool A_KC, int B_MODE>
int launch_variant(const TcParams& p, dim3 grid, const CUtensorMap& tm_hi, const CUtensorMap& tm_lo, cudaStream_t st) {
    static bool configured = false;
    auto kern = gemm_tc_kernel<BN, A_KC, B_MODE>;
    if (!configured) {
        GGPM_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::SMEM));
        configured = true;
    }
    kern<<<grid, THREADS, Cfg<BN>::SMEM, st>>>(p, tm_hi, tm_lo);
    return check_launch("gemm_tcgen05");
}

}  // namespace

bool gemm_tc_eligible(const GemmArgs& g) {
    if (gemm_engine() == GGPM_ENGINE_SIMT_FP32) return false;
    if (g.M <= 0 || g.N <= 0 || g.K <= 0) return false;
    if (!aligned16(g.A) || !aligned16(g.B)) return false;
    const bool a_kc = (g.sak == 1), b_kc = (g.sbk == 1);
    if (!a_kc && g.sam != 1) return false;
    if (!b_kc && g.sbn != 1) return false;
    if (a_kc ? (g.sam % 4 != 0 || g.K % 4 != 0) : (g.sak % 4 != 0 || g.M % 4 != 0)) return false;
    if (b_kc ? (g.sbn % 4 != 0 || g.K % 4 != 0) : (g.sbk % 4 != 0 || g.N % 4 != 0)) return false;
    if (g.a_row_ptr && !a_kc) return false;
    if (g.b_row_ptr && b_kc) return false;
    return true;
}

// GGPM_SEG_CAP: longest accumulation chain, in K chunks, a stream-K segment may hold (default 16; 0 = unbounded)
static int seg_cap_chunks() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("GGPM_SEG_CAP");
        int c = e != nullptr ? atoi(e) : 16;
        v = c < 0 ? 0 : c;
    }
    return v;
}

int launch_gemm_tc(const GemmArgs& g, cudaStream_t st) {
    GGPM_REQUIRE(gemm_tc_eligible(g), "gemm_tcgen05: operand layout not supported (alignment / strides)");
    const bool a_kc = (g.sak == 1), b_kc = (g.sbk == 1);
    TcParams p;
    p.A = g.A; p.lda = a_kc ? g.sam : g.sak;
    p.A2 = g.A2; p.lda2 = g.lda2;
    p.B = g.B; p.ldb = b_kc ? g.sbn : g.sbk;
    p.C = g.C; p.ldc = g.ldc; p.M = g.M; p.N = g.N; p.K = g.K;
    p.a_row_ptr = g.a_row_ptr; p.a_col = g.a_col;
    p.b_row_ptr = g.b_row_ptr; p.b_col = g.b_col;
    p.split_dim = g.split_dim; p.split_at = g.split_at;
    p.beta = g.beta; p.bias = g.bias; p.act = g.act; p.zero_row0 = g.zero_row0;
    // split_k > 1 is a hint from the fp32 path ("this is a long-K accumulation"): here K is cut automatically
    const bool plain_epilogue = !g.bias && g.act == GGPM_ACT_NONE && !g.zero_row0 && (g.beta == 0.f || g.beta == 1.f);
    if (g.split_k > 1) GGPM_REQUIRE(!g.bias && g.act == GGPM_ACT_NONE && !g.zero_row0, "gemm: split-K has no epilogue");
    bool wide = g.N > 128;
    if (g.tile_n == 128) wide = false;
    if (g.tile_n == 256) wide = true;
    if (g.split_dim != SPLIT_NONE) {
        GGPM_REQUIRE(g.split_at > 0, "gemm_tcgen05: split_at must be positive");
        if (g.split_dim == SPLIT_N && g.split_at % 256 != 0) wide = false;
        GGPM_REQUIRE(g.split_dim != SPLIT_N || g.split_at % (wide ? 256 : 128) == 0, "gemm_tcgen05: N split must be a multiple of 128");
        GGPM_REQUIRE(g.split_dim != SPLIT_M || g.split_at % BM == 0, "gemm_tcgen05: M split must be a multiple of 128");
        GGPM_REQUIRE(g.split_dim != SPLIT_K || g.split_at % BK == 0, "gemm_tcgen05: K split must be a multiple of 32");
        GGPM_REQUIRE(g.A2 == nullptr || (aligned16(g.A2) && g.lda2 % 4 == 0), "gemm_tcgen05: A2 alignment");
    }
    const int bnf = wide ? 256 : 128;
    p.split_s = 0;
    p.nk = ceil_div(g.K, BK);
    p.seg_cap = seg_cap_chunks();
    p.m_tiles = ceil_div(g.M, BM);
    p.n_tiles = ceil_div(g.N, bnf);
    p.total_tiles = p.m_tiles * p.n_tiles;
    const int sms = sm_count();
    int ctas = sms;
    // tail tiles are cut along K (stream-K) when the epilogue is a plain store / accumulate and there is enough K
    // to share; otherwise every tile is processed whole
    p.tiles_full = p.total_tiles;
    if (plain_epilogue && p.total_tiles % sms != 0 && p.nk >= 2) {
        p.tiles_full = (p.total_tiles / sms) * sms;
        const long long tail_units = (long long)(p.total_tiles - p.tiles_full) * p.nk;
        if (p.tiles_full == 0 && tail_units < sms) ctas = (int)tail_units;
    } else if (p.total_tiles < sms) {
        ctas = p.total_tiles;
    }
    const bool c_vec = aligned16(g.C) && g.ldc % 4 == 0 && g.N % 4 == 0 && (g.bias == nullptr || aligned16(g.bias));
    if (!c_vec) p.epi_mode = EPI_SCALAR;
    else if (g.beta != 0.f || g.bias != nullptr || g.act != GGPM_ACT_NONE || g.zero_row0) p.epi_mode = EPI_GENERIC;
    else p.epi_mode = EPI_STORE;
    if (p.tiles_full < p.total_tiles && g.beta == 0.f) {
        // tail tiles are accumulated with reductions: zero their rows (whole-tile stores of the same rows come later)
        const int m_first = (p.tiles_full / p.n_tiles) * BM;
        GGPM_CUDA(cudaMemset2DAsync(g.C + (size_t)m_first * g.ldc, sizeof(float) * (size_t)g.ldc, 0,
                                    sizeof(float) * (size_t)g.N, (size_t)(g.M - m_first), st));
    }
    dim3 grid(ctas, 1, 1);
    ProfScope prof(a_kc ? PROF_GEMM_TC : PROF_GEMM_WGRAD, st, 2.0 * g.M * g.N * (double)g.K);
    CUtensorMap tm_hi, tm_lo;
    memset(&tm_hi, 0, sizeof(tm_hi));
    memset(&tm_lo, 0, sizeof(tm_lo));
    const bool b_tma = g.B_hi != nullptr && g.B_lo != nullptr && b_kc && a_kc && g.b_row_ptr == nullptr;
    if (b_tma) {
        GGPM_REQUIRE(aligned16(g.B_hi) && aligned16(g.B_lo), "gemm_tcgen05: split weight planes must be 16-byte aligned");
        GGPM_TRY(make_plane_tmap(&tm_hi, g.B_hi, g.N, g.K, g.sbn, bnf));
        GGPM_TRY(make_plane_tmap(&tm_lo, g.B_lo, g.N, g.K, g.sbn, bnf));
    }
#define GGPM_TC(BNV, AK, BM_) return launch_variant<BNV, AK, BM_>(p, grid, tm_hi, tm_lo, st)
    if (wide) {
        if (b_tma) GGPM_TC(256, true, 2);
        if (a_kc && b_kc) GGPM_TC(256, true, 1);
        if (a_kc && !b_kc) GGPM_TC(256, true, 0);
        if (!a_kc && b_kc) GGPM_TC(256, false, 1);
        GGPM_TC(256, false, 0);
    } else {
        if (b_tma) GGPM_TC(128, true, 2);
        if (a_kc && b_kc) GGPM_TC(128, true, 1);
        if (a_kc && !b_kc) GGPM_TC(128, true, 0);
        if (!a_kc && b_kc) GGPM_TC(128, false, 1);
        GGPM_TC(128, false, 0);
    }
#undef GGPM_TC
}

static int f16_mode(const GemmArgs& g) { return g.f16_tn ? F16_TN_TMA : (g.Ah_hi != nullptr ? F16_A_TMA : F16_A_REGS); }

bool gemm_f16_eligible(const GemmArgs& g) {
    const int e = gemm_engine();
    if (e == GGPM_ENGINE_SIMT_FP32 || e == GGPM_ENGINE_TCGEN05_3XTF32) return false;
    if (g.M <= 0 || g.N <= 0 || g.K <= 0) return false;
    if (!g.Bh_hi || !g.Bh_lo || g.b_row_ptr) return false;
    if (!aligned16(g.Bh_hi) || !aligned16(g.Bh_lo) || g.sbn % 8 != 0) return false;
    if (!(aligned16(g.C) && g.ldc % 4 == 0 && g.N % 4 == 0 && (g.bias == nullptr || aligned16(g.bias)))) return false;
    const int mode = f16_mode(g);
    if (mode == F16_A_REGS) {
        if (!g.b_scale || !g.A || g.sak != 1 || g.sbk != 1) return false;
        if (!aligned16(g.A) || g.sam % 4 != 0 || g.K % 4 != 0) return false;
        if (g.split_dim == SPLIT_M) return false;
        if (g.split_dim == SPLIT_K && g.split_at % BK16 != 0) return false;
        if (g.split_dim == SPLIT_N && g.split_at % 128 != 0) return false;
        if (g.A2 && (!aligned16(g.A2) || g.lda2 % 4 != 0)) return false;
        return true;
    }
    // plane modes: A is fp16 hi / lo planes with its scale on the device; no gathers, no second region
    if (!g.Ah_hi || !g.Ah_lo || !g.a_scale || g.a_row_ptr || g.split_dim != SPLIT_NONE) return false;
    if (!aligned16(g.Ah_hi) || !aligned16(g.Ah_lo) || g.lda_h % 8 != 0) return false;
    if (mode == F16_TN_TMA) return g.M % 64 == 0 && g.N % 64 == 0;
    return g.K % 8 == 0;
}

int launch_gemm_f16(const GemmArgs& g, cudaStream_t st) {
    GGPM_REQUIRE(gemm_f16_eligible(g), "gemm_f16x3: operand layout not supported");
    const int mode = f16_mode(g);
    F16Params fp;
    TcParams& p = fp.t;
    p.A = g.A; p.lda = g.sam; p.A2 = g.A2; p.lda2 = g.lda2;
    p.B = nullptr; p.ldb = g.sbn;
    p.C = g.C; p.ldc = g.ldc; p.M = g.M; p.N = g.N; p.K = g.K;
    p.a_row_ptr = g.a_row_ptr; p.a_col = g.a_col; p.b_row_ptr = nullptr; p.b_col = nullptr;
    p.split_dim = g.split_dim; p.split_at = g.split_at;
    p.beta = g.beta; p.bias = g.bias; p.act = g.act; p.zero_row0 = g.zero_row0;
    fp.a_amax = g.a_amax; fp.a_amax_n = g.a_amax_n; fp.a_amax_const = g.a_amax_const; fp.a_deg = g.a_deg;
    fp.a_scale = g.a_scale; fp.b_scale = g.b_scale; fp.b_scale_const = g.b_scale_const;
    fp.gx_Yn = g.gx_Yn; fp.gx_src = g.gx_src; fp.gx_T = g.gx_T; fp.gx_k1 = g.gx_k1; fp.gx_k2 = g.gx_k2;
    fp.gx_ld = g.gx_ld; fp.gx_hp = g.gx_hp;
    GGPM_REQUIRE(g.act != GGPM_ACT_LSTM_GATES || (g.gx_Yn && g.gx_src && g.gx_T && g.gx_k1 && g.gx_hp > 0 && g.gx_hp % 4 == 0 &&
                                                 g.gx_ld % 4 == 0 && g.beta == 0.f && g.N <= 3 * g.gx_hp),
                 "gemm_f16x3: the LSTM-gate epilogue needs the x-projection tables");
#ifdef GGPM_TRACE
    fp.dbg = nullptr;
    if (const char* tr = getenv("GGPM_F16_TRACE")) fp.dbg = (unsigned long long*)strtoull(tr, nullptr, 0);
#endif
    const bool plain_epilogue = !g.bias && g.act == GGPM_ACT_NONE && !g.zero_row0 && (g.beta == 0.f || g.beta == 1.f);
    bool wide = g.N > 128;
    // state gradient of a small layer (N = Hp = 256, fewer than two waves of 128 x 256 tiles): 128-wide tiles double the
    // tile count and leave TMEM room for a second accumulator set (measured -0.2 ms per cfg2 step, neutral at cfg3)
    if (mode == F16_A_TMA && g.tile_n == 0 && g.N == 256 && ceil_div(g.M, BM) < 2 * sm_count()) wide = false;
    // plane-fed GEMMs of large layers (forward gate / forget projections, state gradient): 128-wide tiles leave TMEM room
    // for a second accumulator set, so the epilogue of tile i overlaps the MMAs of tile i + 1 — measured on cfg4 (A/B in
    // one gpurun call, profiles/r2_s4_*): GEMM class 35.7 -> 32.2 ms per step, step 138.2 -> 136.2 ms
    if (mode == F16_A_TMA && g.tile_n == 0 && g.K <= 1024) wide = false;
    // (weight gradients keep 256-wide tiles: 128-wide ones measured -4 % (cfg2) / -12 % (cfg1) on that class, -0.5 % on the
    // step, but the doubled number of row-range partial sums moved the K = 19001 kernel test from 1.9e-6 to 2.7e-6)
    if (g.tile_n == 128) wide = false;
    if (g.tile_n == 256) wide = true;
    {   // experiment knob: GGPM_F16_BN_FWD / _DGRAD / _WGRAD = 128 | 256
        static int forced[3] = {-1, -1, -1};
        if (forced[0] < 0) {
            const char* names[3] = {"GGPM_F16_BN_FWD", "GGPM_F16_BN_DGRAD", "GGPM_F16_BN_WGRAD"};
            for (int i = 0; i < 3; ++i) { const char* e = getenv(names[i]); forced[i] = e ? atoi(e) : 0; }
        }
        if (forced[mode] == 128) wide = false;
        if (forced[mode] == 256 && g.N > 128) wide = true;
    }
    if (g.split_dim == SPLIT_N && g.split_at % 256 != 0) wide = false;
    const int bnf = wide ? 256 : 128;
    p.nk = ceil_div(g.K, BK16);
    p.seg_cap = seg_cap_chunks();
    p.m_tiles = ceil_div(g.M, BM);
    p.n_tiles = ceil_div(g.N, bnf);
    p.total_tiles = p.m_tiles * p.n_tiles;
    const int sms = sm_count();
    int ctas = sms;
    p.tiles_full = p.total_tiles;
    p.split_s = 0;
    if (p.total_tiles < sms) {
        // fewer tiles than SMs: cut every tile along K into equal slices so that each CTA runs one segment and one
        // (reducing) epilogue; slices of a tile are summed by the TMA reduce-add of the epilogue
        int s_ = plain_epilogue ? sms / p.total_tiles : 1;
        if (s_ > p.nk) s_ = p.nk;
        if (const char* cap = getenv("GGPM_F16_SPLIT")) {     // experiment knob
            const int c = atoi(cap);
            if (c > 0 && c < s_) s_ = c;
        }
        if (s_ >= 2) { p.tiles_full = 0; p.split_s = s_; ctas = p.total_tiles * s_; }
        else ctas = p.total_tiles;
    } else if (plain_epilogue && p.total_tiles % sms != 0 && p.nk >= 2) {
        p.tiles_full = (p.total_tiles / sms) * sms;       // tail tiles after the full waves: stream-K over all CTAs
    }
    if (g.beta != 0.f || g.bias != nullptr || g.act != GGPM_ACT_NONE || g.zero_row0) p.epi_mode = EPI_GENERIC;
    else p.epi_mode = EPI_STORE;
    if (p.tiles_full < p.total_tiles && g.beta == 0.f) {
        const int m_first = (p.tiles_full / p.n_tiles) * BM;
        GGPM_CUDA(cudaMemset2DAsync(g.C + (size_t)m_first * g.ldc, sizeof(float) * (size_t)g.ldc, 0,
                                    sizeof(float) * (size_t)g.N, (size_t)(g.M - m_first), st));
    }
    ProfScope prof(mode == F16_TN_TMA ? PROF_GEMM_WGRAD : PROF_GEMM_TC, st, 2.0 * g.M * g.N * (double)g.K);
    CUtensorMap tm_hi, tm_lo, ta_hi, ta_lo, tc_out;
    memset(&ta_hi, 0, sizeof(ta_hi));
    memset(&ta_lo, 0, sizeof(ta_lo));
    GGPM_TRY(make_out_tmap(&tc_out, g.C, g.M, g.N, g.ldc));
    dim3 grid(ctas, 1, 1);
    if (mode == F16_TN_TMA) {
        GGPM_TRY(make_plane_tmap16_tn(&tm_hi, g.Bh_hi, g.K, g.N, g.sbn, bnf / 64));
        GGPM_TRY(make_plane_tmap16_tn(&tm_lo, g.Bh_lo, g.K, g.N, g.sbn, bnf / 64));
        GGPM_TRY(make_plane_tmap16_tn(&ta_hi, g.Ah_hi, g.K, g.M, g.lda_h, BM / 64));
        GGPM_TRY(make_plane_tmap16_tn(&ta_lo, g.Ah_lo, g.K, g.M, g.lda_h, BM / 64));
        if (wide) return launch_f16_variant<256, F16_TN_TMA>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
        return launch_f16_variant<128, F16_TN_TMA>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
    }
    GGPM_TRY(make_plane_tmap16(&tm_hi, g.Bh_hi, g.N, g.K, g.sbn, bnf));
    GGPM_TRY(make_plane_tmap16(&tm_lo, g.Bh_lo, g.N, g.K, g.sbn, bnf));
    if (mode == F16_A_TMA) {
        GGPM_TRY(make_plane_tmap16(&ta_hi, g.Ah_hi, g.M, g.K, g.lda_h, BM));
        GGPM_TRY(make_plane_tmap16(&ta_lo, g.Ah_lo, g.M, g.K, g.lda_h, BM));
        if (wide) return launch_f16_variant<256, F16_A_TMA>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
        return launch_f16_variant<128, F16_A_TMA>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
    }
    if (wide) return launch_f16_variant<256, F16_A_REGS>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
    return launch_f16_variant<128, F16_A_REGS>(fp, grid, tm_hi, tm_lo, ta_hi, ta_lo, tc_out, st);
}

// ---- helpers of the f16x3 engine: max|w|, scaled fp16 hi / lo planes, max CSR row length ------------------
__global__ void absmax_kernel(const float* __restrict__ w, long long n, float* __restrict__ amax) {
    float m = 0.f;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        m = fmaxf(m, fabsf(w[i]));
    unsigned u = __reduce_max_sync(0xffffffffu, __float_as_uint(m));     // non-negative floats order like their bits
    if ((threadIdx.x & 31) == 0 && u != 0) atomicMax(reinterpret_cast<unsigned*>(amax), u);
}

int launch_absmax(const float* w, long long n, float* amax, cudaStream_t st) {
    if (n <= 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    const int blocks = (int)(ceil_div_ll(n, 256 * 8) < 592 ? ceil_div_ll(n, 256 * 8) : 592);
    absmax_kernel<<<blocks, 256, 0, st>>>(w, n, amax);
    return check_launch("absmax");
}

__global__ void f16_split_kernel(const float* __restrict__ w, long long n2, const float* __restrict__ amax,
                                 uint32_t* __restrict__ hi, uint32_t* __restrict__ lo, float* __restrict__ scale) {
    const float s = tc::pow2_scale(__ldg(amax));
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) *scale = s;
    if (i >= n2) return;
    const float2 v = *reinterpret_cast<const float2*>(w + 2 * i);
    uint32_t h, l;
    tc::split2_f16(v.x * s, v.y * s, h, l);
    hi[i] = h;
    lo[i] = l;
}

int launch_f16_split(const float* w, long long n, const float* amax, void* hi, void* lo, float* scale, cudaStream_t st) {
    GGPM_REQUIRE(n % 2 == 0 && (((uintptr_t)w | (uintptr_t)hi | (uintptr_t)lo) & 7) == 0, "f16_split: needs 8-byte alignment");
    if (n == 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    f16_split_kernel<<<(unsigned)ceil_div_ll(n / 2, 256), 256, 0, st>>>(w, n / 2, amax, (uint32_t*)hi, (uint32_t*)lo, scale);
    return check_launch("f16_split");
}

// planes of the TRANSPOSE of w [rows, cols]: [cols, rows] fp16
__global__ void f16_split_transpose_kernel(const float* __restrict__ w, int rows, int cols, const float* __restrict__ amax,
                                           __half* __restrict__ hi, __half* __restrict__ lo, float* __restrict__ scale) {
    __shared__ float tile[32][33];
    const float s = tc::pow2_scale(__ldg(amax));
    if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0 && threadIdx.y == 0) *scale = s;
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? w[(size_t)r * cols + c] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < cols && r < rows) {
            const float v = tile[threadIdx.x][i] * s;
            const __half h = __float2half_rn(v);
            hi[(size_t)c * rows + r] = h;
            lo[(size_t)c * rows + r] = __float2half_rn(v - __half2float(h));
        }
    }
}

int launch_f16_split_transpose(const float* w, int rows, int cols, const float* amax, void* hi, void* lo, float* scale,
                               cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    dim3 grid(ceil_div(cols, 32), ceil_div(rows, 32)), block(32, 8);
    f16_split_transpose_kernel<<<grid, block, 0, st>>>(w, rows, cols, amax, (__half*)hi, (__half*)lo, scale);
    return check_launch("f16_split_transpose");
}

__global__ void max_degree_kernel(const int* __restrict__ rp, int rows, int* __restrict__ out) {
    int m = 0;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < rows; i += gridDim.x * blockDim.x)
        m = max(m, __ldg(rp + i + 1) - __ldg(rp + i));
    m = __reduce_max_sync(0xffffffffu, m);
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

int launch_max_degree(const int* row_ptr, int rows, int* out, cudaStream_t st) {
    if (rows <= 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    const int blocks = ceil_div(rows, 256) < 296 ? ceil_div(rows, 256) : 296;
    max_degree_kernel<<<blocks, 256, 0, st>>>(row_ptr, rows, out);
    return check_launch("max_degree");
}

__global__ void tf32_split_kernel(const float* __restrict__ w, long long n4, float* __restrict__ hi, float* __restrict__ lo) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 h, l;
    tc::split4(ld4(w + 4 * i), h, l);
    st4(hi + 4 * i, h);
    st4(lo + 4 * i, l);
}

int launch_tf32_split(const float* w, long long n, float* hi, float* lo, cudaStream_t st) {
    GGPM_REQUIRE(n % 4 == 0 && (((uintptr_t)w | (uintptr_t)hi | (uintptr_t)lo) & 15) == 0, "tf32_split: needs 16-byte alignment");
    if (n == 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    tf32_split_kernel<<<(unsigned)ceil_div_ll(n / 4, 256), 256, 0, st>>>(w, n / 4, hi, lo);
    return check_launch("tf32_split");
}

// hi/lo planes of the TRANSPOSE of w [rows, cols] (row-major): planes are [cols, rows]
__global__ void tf32_split_transpose_kernel(const float* __restrict__ w, int rows, int cols, float* __restrict__ hi,
                                            float* __restrict__ lo) {
    __shared__ float tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int r = r0 + i, c = c0 + threadIdx.x;
        tile[i][threadIdx.x] = (r < rows && c < cols) ? w[(size_t)r * cols + c] : 0.f;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        const int c = c0 + i, r = r0 + threadIdx.x;
        if (c < cols && r < rows) {
            const float v = tile[threadIdx.x][i];
            const float h = tc::tf32_hi(v);
            hi[(size_t)c * rows + r] = h;
            lo[(size_t)c * rows + r] = tc::tf32_hi(v - h);
        }
    }
}

int launch_tf32_split_transpose(const float* w, int rows, int cols, float* hi, float* lo, cudaStream_t st) {
    if (rows <= 0 || cols <= 0) return 0;
    ProfScope prof(PROF_MISC, st, 0.0);
    dim3 grid(ceil_div(cols, 32), ceil_div(rows, 32)), block(32, 8);
    tf32_split_transpose_kernel<<<grid, block, 0, st>>>(w, rows, cols, hi, lo);
    return check_launch("tf32_split_transpose");
}

// generic entry: tensor cores when the layout allows, else the fp32 CUDA-core kernel
// GGPM_GLUE_KSLICE: K-slice length of the generic tcgen05 GEMMs (default 256; 0 = unsliced);
// GGPM_GLUE_KSLICE_MIN_ROWS: only GEMMs with at least this many rows are sliced (default 4096)
static int env_int(const char* name, int dflt) {
    const char* e = getenv(name);
    const int c = e != nullptr ? atoi(e) : dflt;
    return c < 0 ? 0 : c;
}
static int glue_kslice() {
    static int v = -1;
    if (v < 0) v = (env_int("GGPM_GLUE_KSLICE", 256) / 32) * 32;
    return v;
}
static int glue_kslice_min_rows() {
    static int v = -1;
    if (v < 0) v = env_int("GGPM_GLUE_KSLICE_MIN_ROWS", 4096);
    return v;
}

int launch_gemm_nt(const GemmArgs& g, cudaStream_t st) {
    static int glue_simt = -1;          // diagnosis knob: generic GEMMs (glue linears, x-projections) on the fp32 CUDA-core engine
    if (glue_simt < 0) glue_simt = env_int("GGPM_GLUE_SIMT", 0) != 0 ? 1 : 0;
    if (glue_simt && !g.a_row_ptr && !g.b_row_ptr) return launch_gemm_simt(g, st);
    if (gemm_tc_eligible(g)) {
        // Long-K products of the generic path (glue linears: K up to 2H, non-negative post-ReLU inputs) run as K-slices that
        // the epilogue accumulates in fp32: a 3xTF32 accumulator takes K / 8 MMA steps per tile, each rounding toward zero,
        // and the loss grows with that chain and is coherent across rows (measured at K = 1000: rms error
        // 2.3e-6 unsliced, 5.6e-7 with 256-deep slices, cuBLAS fp32 4.0e-7).  Slices are multiples of 32; bias, activation
        // and the row-0 mask run in the last slice only.
        const int ks = glue_kslice();
        if (ks > 0 && g.K > ks + ks / 2 && g.split_k <= 1 && !g.a_row_ptr && !g.b_row_ptr && g.split_dim == SPLIT_NONE &&
            g.M >= glue_kslice_min_rows()) {
            for (int k0 = 0; k0 < g.K;) {
                const int len = g.K - k0 > ks + ks / 2 ? ks : g.K - k0;
                GemmArgs q = g;
                q.A = g.A + (long long)k0 * g.sak;
                q.B = g.B + (long long)k0 * g.sbk;
                q.K = len;
                const bool first = k0 == 0, last = k0 + len >= g.K;
                if (!first) q.beta = 1.f;
                if (!last) { q.bias = nullptr; q.act = GGPM_ACT_NONE; q.zero_row0 = 0; }
                if (!gemm_tc_eligible(q)) {
                    GGPM_REQUIRE(first, "gemm: K-slice %d of a sliced product is not eligible for the tcgen05 engine", k0);
                    return launch_gemm_tc(g, st);
                }
                GGPM_TRY(launch_gemm_tc(q, st));
                k0 += len;
            }
            return 0;
        }
        return launch_gemm_tc(g, st);
    }
    GGPM_REQUIRE(!g.a_row_ptr && !g.b_row_ptr, "gemm: fused gather needs the tcgen05 engine");
    return launch_gemm_simt(g, st);
}

}  // namespace ggpm
