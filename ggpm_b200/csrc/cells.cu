// Per-round cell kernels of the graph LSTM / GRU (forward and backward) and small element-wise helpers.
//
// One warp per message row, lanes across the hidden dimension in 128-bit chunks.  Neighbour rows are
// reached through the CSR of bgraph; the padded [E, K, H] tensors of the reference (ggpm/rnn.py:26-38,
// 86-93) are never built.  The depth-invariant x-projection is read as a sum of table rows:
//     X_g[e] = Yn[src[e], g] + T[k1[e], g] (+ T[k2[e], g])
// Math: SURVEY.md Appendix A (restated and checked against autograd in tests/_proto.py).
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "tc_ptx.cuh"

namespace ggpm {

constexpr int kCellWarps = 8;
constexpr int kCellGridBlocks = 8;
constexpr int kCellMinBlocks = 4;       // 64 registers per thread: 32 resident warps per SM for the row kernels
constexpr float kH16Scale = 8192.f;      // 2^13: fp16 planes of h (|h| <= 1), GGPM_H16_SCALE

struct XTab {
    const float* Yn; const int* src; const float* T; const int* k1; const int* k2; int ld;   // ld = G*Hp
};

struct XRow {   // resolved row pointers of one message
    const float* y; const float* t1; const float* t2;
};

__device__ __forceinline__ XRow xrow(const XTab& x, int e) {
    XRow r;
    r.y = x.Yn + (size_t)__ldg(x.src + e) * x.ld;
    r.t1 = x.T + (size_t)__ldg(x.k1 + e) * x.ld;
    r.t2 = x.k2 ? x.T + (size_t)__ldg(x.k2 + e) * x.ld : nullptr;
    return r;
}
__device__ __forceinline__ float4 xval(const XRow& r, int off) {
    float4 v = ldg4(r.y + off) + ldg4(r.t1 + off);
    if (r.t2) v += ldg4(r.t2 + off);
    return v;
}

// ------------------------------------------------------------------------------------------------
// LSTM forward cell:  reads gp = [G_i|G_o|G_u|P] of this round (rows e: G, rows n: P), c_prev; writes h, c
// ------------------------------------------------------------------------------------------------
// NC = column chunks (128 columns each) a lane works on together: with NC = 2 (Hp a multiple of 256, the shipped width) the
// neighbour ids are loaded once for both chunks and twice as many gathered rows are in flight per warp.
// ACT: the round's gate block of gp already holds the ACTIVATED gates [i|o|u] (the gate GEMM applied the x-projection, the
// non-linearities and the row-0 mask in its epilogue): only the per-neighbour forget gate is evaluated here.
template <bool FIRST, int NC, bool ACT>
__global__ void __launch_bounds__(kCellWarps * 32, kCellMinBlocks)
lstm_cell_fwd_kernel(int E1, int H, int Hp, const int* __restrict__ rp, const int* __restrict__ col, XTab xt,
                     const float* __restrict__ gp, const float* __restrict__ c_prev, float* __restrict__ h_out,
                     float* __restrict__ c_out, uint2* __restrict__ h16_hi, uint2* __restrict__ h16_lo) {
    pdl_wait();
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31;
    const int ld4h = 4 * Hp;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    // grid-stride over message rows: neighbouring warps take neighbouring rows, no partial last wave
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < E1; e += n_warps) {
    if (e == 0) {   // padding row: forced to zero after every round (ggpm/rnn.py:106-107)
        for (int c = lane * 4; c < Hp; c += 128) {
            if (h_out) st4(h_out + c, f4zero());
            st4(c_out + c, f4zero());
            if (h16_hi) { h16_hi[c >> 2] = make_uint2(0u, 0u); h16_lo[c >> 2] = make_uint2(0u, 0u); }
        }
        continue;
    }
    const XRow xr = xrow(xt, e);
    const int b = FIRST ? 0 : __ldg(rp + e), en = FIRST ? 0 : __ldg(rp + e + 1);
    const float* ge = gp + (size_t)e * ld4h;
    for (int c0 = lane * 4; c0 < Hp; c0 += 128 * NC) {
        float4 ai[NC], ao[NC], au[NC], chat[NC];
        if (!(ACT && !FIRST)) {
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const int c = c0 + 128 * k;
                ai[k] = xval(xr, c); ao[k] = xval(xr, Hp + c); au[k] = xval(xr, 2 * Hp + c);
            }
        }
        if (FIRST) {
#pragma unroll
            for (int k = 0; k < NC; ++k) { chat[k] = sigmoid4(ai[k]) * tanh4(au[k]); ao[k] = sigmoid4(ao[k]); }
        } else {
            float4 xf[NC];
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const int c = c0 + 128 * k;
                if (ACT) {
                    chat[k] = ld4(ge + c) * ld4(ge + 2 * Hp + c);        // i * u, both activated by the GEMM epilogue
                    ao[k] = ld4(ge + Hp + c);                             // o
                } else {
                    ai[k] += ld4(ge + c);
                    ao[k] += ld4(ge + Hp + c);
                    au[k] += ld4(ge + 2 * Hp + c);
                }
                xf[k] = xval(xr, 3 * Hp + c);
            }
            if (!ACT) {
#pragma unroll
                for (int k = 0; k < NC; ++k) { chat[k] = sigmoid4(ai[k]) * tanh4(au[k]); ao[k] = sigmoid4(ao[k]); }
            }
            int j = b;
            if (NC == 1) {
                for (; j + 2 <= en; j += 2) {
                    const int n0 = __ldg(col + j), n1 = __ldg(col + j + 1);
                    float4 p0 = ld4(gp + (size_t)n0 * ld4h + 3 * Hp + c0), p1 = ld4(gp + (size_t)n1 * ld4h + 3 * Hp + c0);
                    float4 q0 = ld4(c_prev + (size_t)n0 * Hp + c0), q1 = ld4(c_prev + (size_t)n1 * Hp + c0);
                    chat[0] = fma4(sigmoid4(xf[0] + p0), q0, chat[0]);
                    chat[0] = fma4(sigmoid4(xf[0] + p1), q1, chat[0]);
                }
            }
            for (; j < en; ++j) {
                const int n0 = __ldg(col + j);
                float4 p[NC], q[NC];
#pragma unroll
                for (int k = 0; k < NC; ++k) {
                    p[k] = ld4(gp + (size_t)n0 * ld4h + 3 * Hp + c0 + 128 * k);
                    q[k] = ld4(c_prev + (size_t)n0 * Hp + c0 + 128 * k);
                }
#pragma unroll
                for (int k = 0; k < NC; ++k) chat[k] = fma4(sigmoid4(xf[k] + p[k]), q[k], chat[k]);
            }
        }
#pragma unroll
        for (int k = 0; k < NC; ++k) {
            const int c = c0 + 128 * k;
            const float4 ch = clip_cols(chat[k], c, H);
            const float4 h = clip_cols(ao[k] * tanh4(ch), c, H);       // ao holds the activated output gate here
            st4(c_out + (size_t)e * Hp + c, ch);
            if (h_out) st4(h_out + (size_t)e * Hp + c, h);      // large layers read h through the fp16 planes below only
            if (h16_hi) {   // fp16 hi / lo planes of h * 2^13 (|h| <= 1): the B operand of the f16x3 weight-gradient GEMM
                uint2 hi, lo;
                tc::split2_f16(h.x * kH16Scale, h.y * kH16Scale, hi.x, lo.x);
                tc::split2_f16(h.z * kH16Scale, h.w * kH16Scale, hi.y, lo.y);
                h16_hi[((size_t)e * Hp + c) >> 2] = hi;
                h16_lo[((size_t)e * Hp + c) >> 2] = lo;
            }
        }
    }
    }
}

// ------------------------------------------------------------------------------------------------
// Z = [gather^T(dA[:, :3Hp]) | dP]  ->  fp16 hi / lo planes [E1, 4Hp], scaled.  Both backward GEMMs of a round
// consume it: dH^{t-1} = Z [Ug; Uf] and [dUg; dUf] += Z^T h^{t-1} (the neighbour sum moved onto the gradient side
// by linearity), so neither needs a gather in its operand path and both are fed by TMA alone.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCellWarps * 32, kCellMinBlocks)
lstm_zgather_kernel(int E1, int Hp, const int* __restrict__ tp, const int* __restrict__ trow,
                    const float* __restrict__ dA, const float* __restrict__ dP, const int* __restrict__ tpos,
                    const float* __restrict__ pairA, const float* __restrict__ amax16,
                    const int* __restrict__ deg, uint4* __restrict__ z_hi, uint4* __restrict__ z_lo,
                    float* __restrict__ scale_out) {
    pdl_wait();
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31;
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) amax = fmaxf(amax, __ldg(amax16 + i));
    const float s = tc::pow2_scale(amax * (float)max(1, __ldg(deg)));
    if (blockIdx.x == 0 && threadIdx.x == 0) *scale_out = s;
    const int ld4h = 4 * Hp;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    for (int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < E1; n += n_warps) {
    const int b = __ldg(tp + n), en = __ldg(tp + n + 1);
    for (int c = lane * 8; c < ld4h; c += 256) {
        float4 a0 = f4zero(), a1 = f4zero();
        if (c < 3 * Hp) {
            int j = b;
            for (; j + 2 <= en; j += 2) {
                const float* r0 = dA + (size_t)__ldg(trow + j) * ld4h + c;
                const float* r1 = dA + (size_t)__ldg(trow + j + 1) * ld4h + c;
                const float4 x0 = ld4(r0), x1 = ld4(r0 + 4), y0 = ld4(r1), y1 = ld4(r1 + 4);
                a0 += x0; a1 += x1; a0 += y0; a1 += y1;
            }
            if (j < en) {
                const float* r0 = dA + (size_t)__ldg(trow + j) * ld4h + c;
                a0 += ld4(r0); a1 += ld4(r0 + 4);
            }
        } else if (pairA != nullptr) {
            // dP[n] = sum over the pairs (e, n) of daf_en, which the cell kernel stored in pair order: the scatter-add of
            // autograd as a gather over the reverse CSR (no atomics, nothing to zero)
            const int cc = c - 3 * Hp;
            for (int j = b; j < en; ++j) {
                const float* r = pairA + (size_t)__ldg(tpos + j) * Hp + cc;
                a0 += ld4(r); a1 += ld4(r + 4);
            }
        } else {
            const float* r = dP + (size_t)n * Hp + (c - 3 * Hp);
            a0 = ld4(r); a1 = ld4(r + 4);
        }
        uint4 hi, lo;
        tc::split8_f16(a0, a1, s, hi, lo);
        z_hi[((size_t)n * ld4h + c) >> 3] = hi;
        z_lo[((size_t)n * ld4h + c) >> 3] = lo;
    }
    }
}

// ------------------------------------------------------------------------------------------------
// S = gather-sum of the neighbours' h^{t-1} (index_select_ND(h, 0, bgraph).sum(1), ggpm/rnn.py:86) as scaled fp16
// hi / lo planes [E1, Hp]: for large layers the forward gate GEMM then takes both operands by TMA (no producer
// arithmetic, the sum is formed once per row instead of once per N tile).  |h| <= 1, so the largest neighbour count
// bounds every row.  The source is the pair of fp16 hi / lo planes of h^{t-1} * 2^13 that the forward cell kernel saves
// anyway (the operands of the forget-projection and weight-gradient GEMMs): large layers never write or read an fp32
// copy of h between rounds.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void add_h16(uint4 hi, uint4 lo, float4& a0, float4& a1) {
    const __half2* ph = reinterpret_cast<const __half2*>(&hi);
    const __half2* pl = reinterpret_cast<const __half2*>(&lo);
    const float2 h0 = __half22float2(ph[0]), h1 = __half22float2(ph[1]), h2 = __half22float2(ph[2]), h3 = __half22float2(ph[3]);
    const float2 l0 = __half22float2(pl[0]), l1 = __half22float2(pl[1]), l2 = __half22float2(pl[2]), l3 = __half22float2(pl[3]);
    a0.x += h0.x + l0.x; a0.y += h0.y + l0.y; a0.z += h1.x + l1.x; a0.w += h1.y + l1.y;
    a1.x += h2.x + l2.x; a1.y += h2.y + l2.y; a1.z += h3.x + l3.x; a1.w += h3.y + l3.y;
}

__global__ void __launch_bounds__(kCellWarps * 32, kCellMinBlocks)
lstm_sgather16_kernel(int E1, int Hp, const int* __restrict__ rp, const int* __restrict__ col,
                      const uint4* __restrict__ h_hi, const uint4* __restrict__ h_lo, const int* __restrict__ deg,
                      uint4* __restrict__ s_hi, uint4* __restrict__ s_lo, float* __restrict__ scale_out) {
    pdl_wait();
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31;
    // the planes hold h * 2^13: the sum of <= deg rows is scaled so that deg * 2^13 lands in [2^13, 2^14)
    const float s = tc::pow2_scale((float)max(1, __ldg(deg)));
    const float s_in = s / kH16Scale;                 // planes -> S * s
    if (blockIdx.x == 0 && threadIdx.x == 0) *scale_out = s;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < E1; e += n_warps) {
        const int b = __ldg(rp + e), en = __ldg(rp + e + 1);
        for (int c = lane * 8; c < Hp; c += 256) {
            float4 a0 = f4zero(), a1 = f4zero();
            int j = b;
            for (; j + 2 <= en; j += 2) {
                const size_t o0 = ((size_t)__ldg(col + j) * Hp + c) >> 3, o1 = ((size_t)__ldg(col + j + 1) * Hp + c) >> 3;
                const uint4 x0 = h_hi[o0], x1 = h_lo[o0], y0 = h_hi[o1], y1 = h_lo[o1];
                add_h16(x0, x1, a0, a1);
                add_h16(y0, y1, a0, a1);
            }
            if (j < en) {
                const size_t o0 = ((size_t)__ldg(col + j) * Hp + c) >> 3;
                add_h16(h_hi[o0], h_lo[o0], a0, a1);
            }
            uint4 hi, lo;
            tc::split8_f16(a0, a1, s_in, hi, lo);
            s_hi[((size_t)e * Hp + c) >> 3] = hi;
            s_lo[((size_t)e * Hp + c) >> 3] = lo;
        }
    }
}

// ------------------------------------------------------------------------------------------------
// LSTM backward cell (one round).  dP / dC_prev are accumulated with vector reductions (zeroed by caller).
// The gate gradients of the round overwrite the round's saved pre-activations IN PLACE (gp: [G_i|G_o|G_u] of row e are
// read, then [dai|dao|dau] written to the same words by the same thread; the P block, which other rows gather, is left
// alone) and the forget-gate input gradient sum_n daf_en overwrites the round's saved cell state c_cur (read by this
// thread only; neighbours gather the PREVIOUS slot).  Nothing is accumulated across rounds here: the sum over rounds
// that the x-projection gradient needs is one streaming pass after the loop (lstm_dx_reduce_kernel), instead of a
// read-modify-write of a [E, 4Hp] accumulator in every round (8 row-units of HBM traffic per message per round).
// ------------------------------------------------------------------------------------------------
// PAIRS: instead of reducing daf_en into dP[n] and dchat_e f_en into dC_prev[n] with red.global.add (targets zeroed by a
// kernel of their own right before), the two per-pair terms are STORED in pair order (pairA / pairB, [nnz, Hp], coalesced:
// the pairs of a row are consecutive) and summed where they are consumed, through the reverse CSR: dP by the zgather kernel
// of this round, dC of the previous round's rows by THIS kernel one round later (tp / tpos / pairB_in).  Same bytes as the
// zero + reduce + re-read they replace, but plain stores and gathers: no atomics, one launch less per round.
// ACT: the gate block of gp holds the activated gates (see the forward kernel): nothing to recompute for i, o, u.
template <bool FIRST, bool PAIRS, bool ACT>
__global__ void __launch_bounds__(kCellWarps * 32, kCellMinBlocks)
lstm_cell_bwd_kernel(int E1, int H, int Hp, const int* __restrict__ rp, const int* __restrict__ col, XTab xt,
                     float* gp, const float* __restrict__ c_prev, float* c_cur,
                     const float* __restrict__ dH, const float* __restrict__ dC, float* __restrict__ dP,
                     float* __restrict__ dC_prev, float* __restrict__ amax16, const int* __restrict__ tp,
                     const int* __restrict__ tpos, const float* __restrict__ pairB_in, float* __restrict__ pairA,
                     float* __restrict__ pairB) {
    pdl_wait();
    pdl_launch_dependents();
    const int lane = threadIdx.x & 31;
    const int ld4h = 4 * Hp;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    float tmax = 0.f;      // max |dai|, |dao|, |dau|, |daf_en| seen by this thread (scale of the f16x3 state-gradient GEMM)
    for (int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; e < E1; e += n_warps) {
    float* ge = gp + (size_t)e * ld4h;
    float* ce = c_cur + (size_t)e * Hp;
    if (e == 0) {
        for (int c = lane * 4; c < 3 * Hp; c += 128) st4(ge + c, f4zero());
        for (int c = lane * 4; c < Hp; c += 128) st4(ce + c, f4zero());
        continue;
    }
    const XRow xr = xrow(xt, e);
    const int b = FIRST ? 0 : __ldg(rp + e), en = FIRST ? 0 : __ldg(rp + e + 1);
    const int tb = (PAIRS && pairB_in != nullptr) ? __ldg(tp + e) : 0, te = (PAIRS && pairB_in != nullptr) ? __ldg(tp + e + 1) : 0;
    const float4 one = make_float4(1.f, 1.f, 1.f, 1.f);
    for (int c = lane * 4; c < Hp; c += 128) {
        float4 gi, go, gu;
        if (ACT && !FIRST) {
            gi = ld4(ge + c); go = ld4(ge + Hp + c); gu = ld4(ge + 2 * Hp + c);
        } else {
            float4 ai = xval(xr, c), ao = xval(xr, Hp + c), au = xval(xr, 2 * Hp + c);
            if (!FIRST) {
                ai += ld4(ge + c);
                ao += ld4(ge + Hp + c);
                au += ld4(ge + 2 * Hp + c);
            }
            gi = sigmoid4(ai); go = sigmoid4(ao); gu = tanh4(au);
        }
        const float4 th = tanh4(ld4(ce + c));
        const float4 dh = clip_cols(ld4(dH + (size_t)e * Hp + c), c, H);
        float4 dchat = dh * go * (one - th * th);
        if (dC) dchat += clip_cols(ld4(dC + (size_t)e * Hp + c), c, H);
        if (PAIRS) {       // dL/dc of this row from the round processed before: the pairs (e', this row) in reverse-CSR order
            float4 acc = f4zero();
            for (int j = tb; j < te; ++j) acc += ld4(pairB_in + (size_t)__ldg(tpos + j) * Hp + c);
            dchat += clip_cols(acc, c, H);
        }
        const float4 dai = dchat * gu * gi * (one - gi);
        const float4 dao = dh * th * go * (one - go);
        const float4 dau = dchat * gi * (one - gu * gu);
        float4 dafs = f4zero();
        if (!FIRST) {
            const float4 xf = xval(xr, 3 * Hp + c);
            for (int j = b; j < en; ++j) {
                const int n = __ldg(col + j);
                const float4 f = sigmoid4(xf + ld4(gp + (size_t)n * ld4h + 3 * Hp + c));
                const float4 cn = ld4(c_prev + (size_t)n * Hp + c);
                const float4 daf = dchat * cn * f * (one - f);
                tmax = fmaxf(tmax, amax4(daf));
                dafs += daf;
                if (PAIRS) {
                    st4(pairA + (size_t)j * Hp + c, daf);
                    st4(pairB + (size_t)j * Hp + c, dchat * f);
                } else {
                    red_add4(dP + (size_t)n * Hp + c, daf);
                    red_add4(dC_prev + (size_t)n * Hp + c, dchat * f);
                }
            }
        }
        tmax = fmaxf(tmax, fmaxf(amax4(dai), fmaxf(amax4(dao), amax4(dau))));
        st4(ge + c, dai);
        st4(ge + Hp + c, dao);
        st4(ge + 2 * Hp + c, dau);
        st4(ce + c, dafs);
    }
    }
    if (amax16 != nullptr) {
        // non-negative floats order like their bit patterns
        const unsigned u = __reduce_max_sync(0xffffffffu, __float_as_uint(tmax));
        if (lane == 0 && u != 0) atomicMax(reinterpret_cast<unsigned*>(amax16) + (blockIdx.x & 15), u);
    }
}

// dX[e] = [ sum_t dai_t | sum_t dao_t | sum_t dau_t | sum_t sum_n daf_en,t ](e): the gradient of the hoisted x-projection,
// summed over the rounds in the order the backward loop produced them (t = depth first), one streaming pass.
// hist_gp: [depth, E1, 4Hp] (gate-gradient blocks written by lstm_cell_bwd_kernel), hist_c: [depth, E1, Hp] (slots 1..
// depth-1 hold the forget-gate sums; round 1 has no neighbours and contributes none).
__global__ void __launch_bounds__(256)
lstm_dx_reduce_kernel(int E1, int Hp, int depth, const float* __restrict__ hist_gp, const float* __restrict__ hist_c,
                      float* __restrict__ dX) {
    const int ld4h = 4 * Hp, q = ld4h >> 2;
    const size_t n4 = (size_t)E1 * q;
    const size_t gp_slot = (size_t)E1 * ld4h, c_slot = (size_t)E1 * Hp;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const size_t e = i / q;
        const int c = (int)(i % q) * 4;
        float4 acc = f4zero();
        if (c < 3 * Hp) {
            const float* p = hist_gp + e * ld4h + c;
            acc = ld4(p + (size_t)(depth - 1) * gp_slot);
            for (int t = depth - 2; t >= 0; --t) acc += ld4(p + (size_t)t * gp_slot);
        } else if (depth > 1) {
            const float* p = hist_c + e * Hp + (c - 3 * Hp);
            acc = ld4(p + (size_t)(depth - 1) * c_slot);
            for (int t = depth - 2; t >= 1; --t) acc += ld4(p + (size_t)t * c_slot);
        }
        st4(dX + e * ld4h + c, acc);
    }
}

// ------------------------------------------------------------------------------------------------
// GRU kernels.  aux row layout (5*Hp): [S | Gz | Gh | Pr | Sg];  x-gates (3*Hp): [z | h | r]
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kCellWarps * 32)
gru_gate_kernel(int E1, int Hp, const int* __restrict__ rp, const int* __restrict__ col, XTab xt,
                const float* __restrict__ h_prev, float* __restrict__ aux) {
    const int lane = threadIdx.x & 31;
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= E1) return;
    const int ld = 5 * Hp;
    const XRow xr = xrow(xt, e);
    const int b = __ldg(rp + e), en = __ldg(rp + e + 1);
    for (int c = lane * 4; c < Hp; c += 128) {
        const float4 xrr = xval(xr, 2 * Hp + c);
        float4 sg = f4zero();
        for (int j = b; j < en; ++j) {
            const int n = __ldg(col + j);
            const float4 r = sigmoid4(xrr + ld4(aux + (size_t)n * ld + 3 * Hp + c));
            sg = fma4(r, ld4(h_prev + (size_t)n * Hp + c), sg);
        }
        st4(aux + (size_t)e * ld + 4 * Hp + c, sg);
    }
}

__global__ void __launch_bounds__(kCellWarps * 32)
gru_out_kernel(int E1, int H, int Hp, XTab xt, const float* __restrict__ aux, float* __restrict__ h_out) {
    const int lane = threadIdx.x & 31;
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= E1) return;
    const int ld = 5 * Hp;
    const float4 one = make_float4(1.f, 1.f, 1.f, 1.f);
    if (e == 0) {
        for (int c = lane * 4; c < Hp; c += 128) st4(h_out + c, f4zero());
        return;
    }
    const XRow xr = xrow(xt, e);
    const float* a = aux + (size_t)e * ld;
    for (int c = lane * 4; c < Hp; c += 128) {
        const float4 z = sigmoid4(xval(xr, c) + ld4(a + Hp + c));
        const float4 cand = tanh4(xval(xr, Hp + c) + ld4(a + 2 * Hp + c));
        const float4 s = ld4(a + c);
        st4(h_out + (size_t)e * Hp + c, clip_cols((one - z) * s + z * cand, c, H));
    }
}

// dA[e] = [daz | dah], dS[e] = dH (1 - z)
__global__ void __launch_bounds__(kCellWarps * 32)
gru_out_bwd_kernel(int E1, int H, int Hp, XTab xt, const float* __restrict__ aux, const float* __restrict__ dH,
                   float* __restrict__ dA, float* __restrict__ dS) {
    const int lane = threadIdx.x & 31;
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= E1) return;
    const int ld = 5 * Hp;
    const float4 one = make_float4(1.f, 1.f, 1.f, 1.f);
    if (e == 0) {
        for (int c = lane * 4; c < Hp; c += 128) {
            st4(dA + c, f4zero());
            st4(dA + Hp + c, f4zero());
            st4(dS + c, f4zero());
        }
        return;
    }
    const XRow xr = xrow(xt, e);
    const float* a = aux + (size_t)e * ld;
    for (int c = lane * 4; c < Hp; c += 128) {
        const float4 z = sigmoid4(xval(xr, c) + ld4(a + Hp + c));
        const float4 cand = tanh4(xval(xr, Hp + c) + ld4(a + 2 * Hp + c));
        const float4 s = ld4(a + c);
        const float4 dh = clip_cols(ld4(dH + (size_t)e * Hp + c), c, H);
        st4(dA + (size_t)e * 2 * Hp + c, dh * (cand - s) * z * (one - z));
        st4(dA + (size_t)e * 2 * Hp + Hp + c, dh * z * (one - cand * cand));
        st4(dS + (size_t)e * Hp + c, dh * (one - z));
    }
}

// per pair (e, n): reset-gate backward; dX[e] += [daz | dah | sum_n dar_en]; atomics into dPr[n], dHn[n]
__global__ void __launch_bounds__(kCellWarps * 32)
gru_pair_bwd_kernel(int E1, int Hp, const int* __restrict__ rp, const int* __restrict__ col, XTab xt,
                    const float* __restrict__ aux, const float* __restrict__ h_prev, const float* __restrict__ dA,
                    const float* __restrict__ dS, const float* __restrict__ dSg, float* __restrict__ dX,
                    int dx_overwrite, int with_pairs, float* __restrict__ dPr, float* __restrict__ dHn) {
    const int lane = threadIdx.x & 31;
    const int e = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (e >= E1) return;
    const int ld = 5 * Hp;
    const float4 one = make_float4(1.f, 1.f, 1.f, 1.f);
    const XRow xr = xrow(xt, e);
    const int b = with_pairs ? __ldg(rp + e) : 0, en = with_pairs ? __ldg(rp + e + 1) : 0;
    float* dXe = dX + (size_t)e * 3 * Hp;
    for (int c = lane * 4; c < Hp; c += 128) {
        float4 dxr = f4zero();
        if (en > b) {
            const float4 xrr = xval(xr, 2 * Hp + c);
            const float4 dsg = ld4(dSg + (size_t)e * Hp + c);
            const float4 ds = ld4(dS + (size_t)e * Hp + c);
            for (int j = b; j < en; ++j) {
                const int n = __ldg(col + j);
                const float4 r = sigmoid4(xrr + ld4(aux + (size_t)n * ld + 3 * Hp + c));
                const float4 dar = dsg * ld4(h_prev + (size_t)n * Hp + c) * r * (one - r);
                dxr += dar;
                red_add4(dPr + (size_t)n * Hp + c, dar);
                red_add4(dHn + (size_t)n * Hp + c, ds + dsg * r);
            }
        }
        const float4 daz = ld4(dA + (size_t)e * 2 * Hp + c), dah = ld4(dA + (size_t)e * 2 * Hp + Hp + c);
        if (dx_overwrite) {
            st4(dXe + c, daz);
            st4(dXe + Hp + c, dah);
            st4(dXe + 2 * Hp + c, dxr);
        } else {
            st4(dXe + c, ld4(dXe + c) + daz);
            st4(dXe + Hp + c, ld4(dXe + Hp + c) + dah);
            st4(dXe + 2 * Hp + c, ld4(dXe + 2 * Hp + c) + dxr);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// element-wise helpers
// ------------------------------------------------------------------------------------------------
__global__ void act_bwd_kernel(const float* __restrict__ dY, int ld_dy, const float* __restrict__ Y, int ld_y,
                               float* __restrict__ dZ, int ld_dz, int rows, int width, int act, int zero_row0) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)rows * width) return;
    int r = (int)(i / width), c = (int)(i % width);
    float g = dY[(size_t)r * ld_dy + c];
    float y = Y[(size_t)r * ld_y + c];
    if (act == GGPM_ACT_RELU) g = y > 0.f ? g : 0.f;
    else if (act == GGPM_ACT_TANH) g = g * (1.f - y * y);
    else if (act == GGPM_ACT_SIGMOID) g = g * y * (1.f - y);
    if (zero_row0 && r == 0) g = 0.f;
    dZ[(size_t)r * ld_dz + c] = g;
}

// in-place epilogue of a linear layer whose product was accumulated by K-split partial sums:
// C = act(C + bias), optional zero row 0 (the fused GEMM epilogue can only do this on whole-K tiles)
__global__ void bias_act_kernel(float* __restrict__ C, int ldc, int rows, int width, const float* __restrict__ bias, int act,
                                int zero_row0) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int w4 = width >> 2;
    if (i >= (long long)rows * w4) return;
    const int r = (int)(i / w4), c = (int)(i % w4) * 4;
    float4 x = ld4(C + (size_t)r * ldc + c);
    if (bias != nullptr) x += ldg4(bias + c);
    if (act == GGPM_ACT_RELU) x = make_float4(fmaxf(x.x, 0.f), fmaxf(x.y, 0.f), fmaxf(x.z, 0.f), fmaxf(x.w, 0.f));
    else if (act == GGPM_ACT_TANH) x = tanh4(x);
    else if (act == GGPM_ACT_SIGMOID) x = sigmoid4(x);
    if (zero_row0 && r == 0) x = f4zero();
    st4(C + (size_t)r * ldc + c, x);
}

constexpr int kColsumRows = 256;
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, int ld_x, int rows, int width,
                                                     float* __restrict__ out) {
    // block: 32 columns x 8 row-lanes; grid.x over column groups, grid.y over row chunks
    __shared__ float part[8][33];
    const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cx;
    const int r0 = blockIdx.y * kColsumRows;
    const int r1 = min(rows, r0 + kColsumRows);
    float s = 0.f;
    if (c < width)
        for (int r = r0 + ry; r < r1; r += 8) s += X[(size_t)r * ld_x + c];
    part[ry][cx] = s;
    __syncthreads();
    if (ry == 0 && c < width) {
        float t = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) t += part[k][cx];
        atomicAdd(out + c, t);
    }
}

constexpr int kTabRows = 32;      // max table rows held in shared memory
constexpr int kTabChunk = 128;    // message rows per block
__global__ void __launch_bounds__(256)
table_wgrad_kernel(const float* __restrict__ dX, int ld_dx, int rows, int width, const int* __restrict__ k1,
                   const int* __restrict__ k2, float* __restrict__ dT, int ld_dt, int ntab) {
    // block: 128 columns (32 lanes x float4) x 8 warps over a chunk of rows; accumulate per table row in smem.
    // Column lane*4 + i of a table row lives at word i*32 + lane, so the 32 lanes of one atomic hit 32 banks.
    __shared__ float acc[kTabRows][128];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < ntab * 128; i += blockDim.x) (&acc[0][0])[i] = 0.f;
    __syncthreads();
    const int c = blockIdx.x * 128 + lane * 4;
    const int r0 = blockIdx.y * kTabChunk, r1 = min(rows, r0 + kTabChunk);
    if (c < width) {
        for (int r = r0 + w; r < r1; r += 8) {
            const float4 v = ld4(dX + (size_t)r * ld_dx + c);
            const int a = min(max(__ldg(k1 + r), 0), ntab - 1);      // ids are validated by the callers; clamp so a bad one cannot write outside acc
            atomicAdd(&acc[a][lane], v.x);
            atomicAdd(&acc[a][32 + lane], v.y);
            atomicAdd(&acc[a][64 + lane], v.z);
            atomicAdd(&acc[a][96 + lane], v.w);
            if (k2) {
                const int b = min(max(__ldg(k2 + r), 0), ntab - 1);
                atomicAdd(&acc[b][lane], v.x);
                atomicAdd(&acc[b][32 + lane], v.y);
                atomicAdd(&acc[b][64 + lane], v.z);
                atomicAdd(&acc[b][96 + lane], v.w);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < ntab * 128; i += blockDim.x) {
        const int k = i >> 7, word = i & 127;
        const int cc = blockIdx.x * 128 + (word & 31) * 4 + (word >> 5);
        const float v = acc[k][word];
        if (cc < width && v != 0.f) atomicAdd(dT + (size_t)k * ld_dt + cc, v);
    }
}

// Zero the two reduction targets of a backward round in one launch, right before the cell kernel that reduces into
// them: the lines are then L2-resident when the red.global.add traffic arrives (zeroing them a kernel or two earlier
// lets them be evicted first, which was measured to slow the cell kernel by 60 %).
__global__ void __launch_bounds__(256) zero2_kernel(float4* __restrict__ a, float4* __restrict__ b, size_t n4) {
    pdl_wait();
    pdl_launch_dependents();
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        a[i] = z;
        b[i] = z;
    }
}

// ---- launchers ----------------------------------------------------------------------------------
int launch_zero2(float* a, float* b, size_t n, cudaStream_t st) {     // n % 4 == 0, 16-byte aligned
    const size_t n4 = n / 4;
    const int blocks = (int)((n4 + 255) / 256 < (size_t)(8 * sm_count()) ? (n4 + 255) / 256 : (size_t)(8 * sm_count()));
    launch_pdl(zero2_kernel, dim3(blocks > 0 ? blocks : 1), dim3(256), 0, st, (float4*)a, (float4*)b, n4);
    return check_launch("zero2");
}

// grid of the LSTM row kernels: one warp per row up to kCellGridBlocks blocks per SM (two residency rounds), grid-stride
// beyond that (GGPM_CELL_BPS=n overrides; 0: one block per 8 rows, no striding — A/B timing)
static inline int cell_blocks(int E1) { return ceil_div(E1, kCellWarps); }
static inline int cell_blocks_persistent(int E1) {
    static int bps = -1;
    if (bps < 0) {
        const char* e = getenv("GGPM_CELL_BPS");
        bps = e != nullptr ? atoi(e) : kCellGridBlocks;
    }
    const int need = ceil_div(E1, kCellWarps);
    if (bps <= 0) return need;
    const int cap = bps * sm_count();
    return need < cap ? need : cap;
}

// GGPM_CELL_FWD_NC=2 selects the two-chunk variant.  Measured (same box, A/B/A/B): cfg4 137.6 vs 137.0 ms, cfg2 12.43 vs 12.29 ms
// per step in favour of ONE chunk per pass (the kernel is not short of loads in flight; the second chunk costs registers),
// so one chunk is the default.
static int cell_fwd_chunks(int Hp) {
    static int nc = -1;
    if (nc < 0) {
        const char* e = getenv("GGPM_CELL_FWD_NC");
        nc = e != nullptr ? atoi(e) : 1;
    }
    return (nc == 2 && Hp % 256 == 0) ? 2 : 1;
}

int launch_lstm_cell_fwd(bool first, int E1, int H, int Hp, const int* rp, const int* col, const XTabArgs& x,
                         const float* gp, const float* c_prev, float* h_out, float* c_out, void* h16_hi, void* h16_lo,
                         cudaStream_t st, bool gates_activated) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 4 * Hp};
    ProfScope prof(PROF_CELL_FWD, st, 0.0);
    const bool two = cell_fwd_chunks(Hp) == 2 && !gates_activated;
    auto kern = first ? (two ? lstm_cell_fwd_kernel<true, 2, false> : lstm_cell_fwd_kernel<true, 1, false>)
                      : (gates_activated ? lstm_cell_fwd_kernel<false, 1, true>
                                         : (two ? lstm_cell_fwd_kernel<false, 2, false> : lstm_cell_fwd_kernel<false, 1, false>));
    launch_pdl(kern, dim3(cell_blocks_persistent(E1)), dim3(kCellWarps * 32), 0,
               st, E1, H, Hp, rp, col, xt, gp, c_prev, h_out, c_out, (uint2*)h16_hi, (uint2*)h16_lo);
    return check_launch("lstm_cell_fwd");
}

int launch_lstm_zgather(int E1, int Hp, const int* t_ptr, const int* t_row, const float* dA, const float* dP,
                        const int* t_pos, const float* pairA, const float* amax16, const int* deg, void* z_hi, void* z_lo,
                        float* scale_out, cudaStream_t st) {
    ProfScope prof(PROF_GATHER, st, 0.0);
    launch_pdl(lstm_zgather_kernel, dim3(cell_blocks_persistent(E1)), dim3(kCellWarps * 32), 0, st, E1, Hp, t_ptr, t_row, dA, dP,
               t_pos, pairA, amax16, deg, (uint4*)z_hi, (uint4*)z_lo, scale_out);
    return check_launch("lstm_zgather");
}

int launch_lstm_sgather16(int E1, int Hp, const int* b_ptr, const int* b_col, const void* h_hi, const void* h_lo, const int* deg,
                          void* s_hi, void* s_lo, float* scale_out, cudaStream_t st) {
    ProfScope prof(PROF_GATHER, st, 0.0);
    launch_pdl(lstm_sgather16_kernel, dim3(cell_blocks_persistent(E1)), dim3(kCellWarps * 32), 0, st, E1, Hp, b_ptr, b_col,
               (const uint4*)h_hi, (const uint4*)h_lo, deg, (uint4*)s_hi, (uint4*)s_lo, scale_out);
    return check_launch("lstm_sgather16");
}

int launch_lstm_cell_bwd(bool first, int E1, int H, int Hp, const int* rp, const int* col, const XTabArgs& x,
                         float* gp, const float* c_prev, float* c_cur, const float* dH, const float* dC,
                         float* dP, float* dC_prev, float* amax16, cudaStream_t st, const int* t_ptr, const int* t_pos,
                         const float* pairB_in, float* pairA, float* pairB, bool pairs, bool gates_activated) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 4 * Hp};
    ProfScope prof(PROF_CELL_BWD, st, 0.0);
    auto kern = lstm_cell_bwd_kernel<false, false, false>;
    if (first) kern = pairs ? lstm_cell_bwd_kernel<true, true, false> : lstm_cell_bwd_kernel<true, false, false>;
    else if (gates_activated) kern = pairs ? lstm_cell_bwd_kernel<false, true, true> : lstm_cell_bwd_kernel<false, false, true>;
    else if (pairs) kern = lstm_cell_bwd_kernel<false, true, false>;
    launch_pdl(kern, dim3(cell_blocks_persistent(E1)), dim3(kCellWarps * 32), 0, st, E1, H, Hp, rp, col, xt, gp, c_prev, c_cur, dH,
               dC, dP, dC_prev, amax16, t_ptr, t_pos, pairB_in, pairA, pairB);
    return check_launch("lstm_cell_bwd");
}

int launch_lstm_dx_reduce(int E1, int Hp, int depth, const float* hist_gp, const float* hist_c, float* dX, cudaStream_t st) {
    const size_t n4 = (size_t)E1 * Hp;
    const size_t want = (n4 + 255) / 256;
    const int blocks = (int)(want < (size_t)(16 * sm_count()) ? want : (size_t)(16 * sm_count()));
    ProfScope prof(PROF_CELL_BWD, st, 0.0);
    lstm_dx_reduce_kernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(E1, Hp, depth, hist_gp, hist_c, dX);
    return check_launch("lstm_dx_reduce");
}

int launch_gru_gate(int E1, int Hp, const int* rp, const int* col, const XTabArgs& x, const float* h_prev, float* aux,
                    cudaStream_t st) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 3 * Hp};
    ProfScope prof(PROF_CELL_FWD, st, 0.0);
    gru_gate_kernel<<<cell_blocks(E1), kCellWarps * 32, 0, st>>>(E1, Hp, rp, col, xt, h_prev, aux);
    return check_launch("gru_gate");
}

int launch_gru_out(int E1, int H, int Hp, const XTabArgs& x, const float* aux, float* h_out, cudaStream_t st) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 3 * Hp};
    ProfScope prof(PROF_CELL_FWD, st, 0.0);
    gru_out_kernel<<<cell_blocks(E1), kCellWarps * 32, 0, st>>>(E1, H, Hp, xt, aux, h_out);
    return check_launch("gru_out");
}

int launch_gru_out_bwd(int E1, int H, int Hp, const XTabArgs& x, const float* aux, const float* dH, float* dA, float* dS,
                       cudaStream_t st) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 3 * Hp};
    ProfScope prof(PROF_CELL_BWD, st, 0.0);
    gru_out_bwd_kernel<<<cell_blocks(E1), kCellWarps * 32, 0, st>>>(E1, H, Hp, xt, aux, dH, dA, dS);
    return check_launch("gru_out_bwd");
}

int launch_gru_pair_bwd(int E1, int Hp, const int* rp, const int* col, const XTabArgs& x, const float* aux,
                        const float* h_prev, const float* dA, const float* dS, const float* dSg, float* dX,
                        bool dx_overwrite, bool with_pairs, float* dPr, float* dHn, cudaStream_t st) {
    XTab xt{x.Yn, x.src, x.T, x.k1, x.k2, 3 * Hp};
    ProfScope prof(PROF_CELL_BWD, st, 0.0);
    gru_pair_bwd_kernel<<<cell_blocks(E1), kCellWarps * 32, 0, st>>>(E1, Hp, rp, col, xt, aux, h_prev, dA, dS, dSg, dX,
                                                                    dx_overwrite, with_pairs, dPr, dHn);
    return check_launch("gru_pair_bwd");
}

}  // namespace ggpm

using namespace ggpm;

extern "C" int ggpm_act_bwd(const float* dY, int ld_dy, const float* Y, int ld_y, float* dZ, int ld_dz, int rows,
                            int width, int act, int zero_row0, void* stream) {
    GGPM_REQUIRE(dY && Y && dZ, "act_bwd: null pointer");
    if (rows <= 0 || width <= 0) return 0;
    long long n = (long long)rows * width;
    act_bwd_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(dY, ld_dy, Y, ld_y, dZ, ld_dz, rows,
                                                                                 width, act, zero_row0);
    return check_launch("act_bwd");
}

extern "C" int ggpm_bias_act(float* C, int ldc, int rows, int width, const float* bias, int act, int zero_row0, void* stream) {
    GGPM_REQUIRE(C, "bias_act: null pointer");
    GGPM_REQUIRE(width % 4 == 0 && ldc % 4 == 0 && ((uintptr_t)C & 15) == 0 && (bias == nullptr || ((uintptr_t)bias & 15) == 0),
                 "bias_act: width / ld must be multiples of 4 and the pointers 16-byte aligned");
    if (rows <= 0 || width <= 0) return 0;
    const long long n = (long long)rows * (width / 4);
    bias_act_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(C, ldc, rows, width, bias, act, zero_row0);
    return check_launch("bias_act");
}

extern "C" int ggpm_colsum(const float* X, int ld_x, int rows, int width, float* out, int accumulate, void* stream) {
    GGPM_REQUIRE(X && out, "colsum: null pointer");
    cudaStream_t st = (cudaStream_t)stream;
    if (width <= 0) return 0;
    if (!accumulate) GGPM_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * (size_t)width, st));
    if (rows <= 0) return 0;
    dim3 grid(ceil_div(width, 32), ceil_div(rows, kColsumRows));
    colsum_kernel<<<grid, 256, 0, st>>>(X, ld_x, rows, width, out);
    return check_launch("colsum");
}

extern "C" int ggpm_table_wgrad(const float* dX, int ld_dx, int rows, int width, const int32_t* k1, const int32_t* k2,
                                float* dT, int ld_dt, int n_table_rows, void* stream) {
    GGPM_REQUIRE(dX && k1 && dT, "table_wgrad: null pointer");
    GGPM_REQUIRE(n_table_rows > 0 && n_table_rows <= kTabRows, "table_wgrad: at most %d table rows", kTabRows);
    GGPM_REQUIRE(width % 4 == 0 && ld_dx % 4 == 0, "table_wgrad: width/ld must be multiples of 4");
    if (rows <= 0) return 0;
    dim3 grid(ceil_div(width, 128), ceil_div(rows, kTabChunk));
    table_wgrad_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(dX, ld_dx, rows, width, k1, k2, dT, ld_dt, n_table_rows);
    return check_launch("table_wgrad");
}
