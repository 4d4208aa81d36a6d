// Weight packing for the message-passing kernels as single launches (forward: reference parameter layout ->
// padded kernel layout; backward: gradients of the packed buffers -> gradients at the reference shapes).
//
// The reference stores every LSTM gate as nn.Linear(IN + H -> H) applied to [x | h] (ggpm/rnn.py:69-72) with
// x = [node feature (F columns) | one-hot columns (IN - F)] (ggpm/encoder.py:103-106, 114-116, 121-125).  The kernels
// want Wx [4Hp, Fp], T [ntab, 4Hp] and U4 = [U_i; U_o; U_u; U_f] [4Hp, Hp] (see ggpm_b200/packing.py); doing this with
// torch slicing costs ~100 tiny launches and autograd nodes per step, which made the step host-bound.
#include "common.cuh"
#include "kernels.h"

namespace ggpm {

struct GateW {
    const float* w[4];      // [H, IN + H] row-major, gate order i, o, u, f
    const float* b[4];      // [H]
};
struct GateG {
    float* w[4];
    float* b[4];
};

__global__ void lstm_pack_kernel(GateW g, int H, int Hp, int F, int Fp, int IN, int ntab, int has_tab, int n_bias_rows,
                                 float* __restrict__ Wx, float* __restrict__ T, float* __restrict__ U4) {
    const int ldw = IN + H;
    const long long n_wx = (long long)4 * Hp * Fp, n_t = (long long)ntab * 4 * Hp, n_u = (long long)4 * Hp * Hp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n_wx + n_t + n_u;
         i += (long long)gridDim.x * blockDim.x) {
        if (i < n_wx) {
            const int r = (int)(i / Fp), c = (int)(i % Fp);
            const int gi = r / Hp, j = r % Hp;
            Wx[i] = (j < H && c < F) ? g.w[gi][(size_t)j * ldw + c] : 0.f;
        } else if (i < n_wx + n_t) {
            const long long q = i - n_wx;
            const int k = (int)(q / (4 * Hp)), cc = (int)(q % (4 * Hp));
            const int gi = cc / Hp, j = cc % Hp;
            float v = 0.f;
            if (j < H) {
                if (has_tab) v = g.w[gi][(size_t)j * ldw + F + k];
                if (k < n_bias_rows) v += g.b[gi][j];
            }
            T[q] = v;
        } else {
            const long long q = i - n_wx - n_t;
            const int r = (int)(q / Hp), c = (int)(q % Hp);
            const int gi = r / Hp, j = r % Hp;
            U4[q] = (j < H && c < H) ? g.w[gi][(size_t)j * ldw + IN + c] : 0.f;
        }
    }
}

__global__ void lstm_unpack_kernel(GateG g, int H, int Hp, int F, int Fp, int IN, int has_tab, int n_bias_rows,
                                   const float* __restrict__ dWx, const float* __restrict__ dT,
                                   const float* __restrict__ dU4) {
    const int ldw = IN + H;
    const long long per_gate = (long long)H * ldw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < 4 * per_gate + 4 * H;
         i += (long long)gridDim.x * blockDim.x) {
        if (i < 4 * per_gate) {
            const int gi = (int)(i / per_gate);
            const long long q = i % per_gate;
            const int j = (int)(q / ldw), c = (int)(q % ldw);
            float v;
            if (c < F) v = dWx ? dWx[(size_t)(gi * Hp + j) * Fp + c] : 0.f;
            else if (c < IN) v = dT ? dT[(size_t)(c - F) * 4 * Hp + gi * Hp + j] : 0.f;
            else v = dU4 ? dU4[(size_t)(gi * Hp + j) * Hp + (c - IN)] : 0.f;
            g.w[gi][q] = v;
        } else {
            const int q = (int)(i - 4 * per_gate);
            const int gi = q / H, j = q % H;
            float v = 0.f;
            if (dT)
                for (int k = 0; k < n_bias_rows; ++k) v += dT[(size_t)k * 4 * Hp + gi * Hp + j];
            g.b[gi][j] = v;
        }
    }
}

constexpr int kMaxPieces = 4;
struct Pieces {
    float* out[kMaxPieces];     // [Np, kp[i]]
    int c0[kMaxPieces], c1[kMaxPieces], kp[kMaxPieces];
    int n;
};

// W [N, K] (row stride ldw) -> column pieces [c0, c1) zero-padded to [Np, kp]; bias [N] -> bias_p [Np]
__global__ void split_pad_kernel(const float* __restrict__ W, int ldw, int N, int Np, Pieces pc,
                                 const float* __restrict__ bias, float* __restrict__ bias_p) {
    long long total = 0;
    for (int p = 0; p < pc.n; ++p) total += (long long)Np * pc.kp[p];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total + Np; i += (long long)gridDim.x * blockDim.x) {
        if (i >= total) {
            const int j = (int)(i - total);
            if (bias_p) bias_p[j] = (bias && j < N) ? bias[j] : 0.f;
            continue;
        }
        long long q = i;
        int p = 0;
        while (q >= (long long)Np * pc.kp[p]) { q -= (long long)Np * pc.kp[p]; ++p; }
        const int r = (int)(q / pc.kp[p]), c = (int)(q % pc.kp[p]);
        const int sc = pc.c0[p] + c;
        pc.out[p][q] = (r < N && sc < pc.c1[p]) ? W[(size_t)r * ldw + sc] : 0.f;
    }
}

// inverse for gradients: dW [N, K] contiguous gets every column from its piece (columns outside all pieces: 0)
__global__ void split_unpad_kernel(float* __restrict__ dW, int K, int N, Pieces pc, const float* __restrict__ dbias_p,
                                   float* __restrict__ dbias) {
    const long long total = (long long)N * K;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total + N; i += (long long)gridDim.x * blockDim.x) {
        if (i >= total) {
            const int j = (int)(i - total);
            if (dbias) dbias[j] = dbias_p ? dbias_p[j] : 0.f;
            continue;
        }
        const int r = (int)(i / K), c = (int)(i % K);
        float v = 0.f;
        for (int p = 0; p < pc.n; ++p)
            if (c >= pc.c0[p] && c < pc.c1[p] && pc.out[p] != nullptr) v = pc.out[p][(size_t)r * pc.kp[p] + (c - pc.c0[p])];
        dW[i] = v;
    }
}

static int blocks_for(long long n) {
    long long b = ceil_div_ll(n, 256);
    const long long cap = (long long)sm_count() * 8;
    return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

}  // namespace ggpm

using namespace ggpm;

extern "C" int ggpm_lstm_pack(const float* const* W4, const float* const* b4, int H, int Hp, int F, int Fp, int IN,
                              int ntab, int n_bias_rows, float* Wx, float* T, float* U4, void* stream) {
    GGPM_REQUIRE(W4 && b4 && Wx && T && U4, "lstm_pack: null pointer");
    GGPM_REQUIRE(H >= 1 && Hp >= H && F >= 0 && Fp >= F && IN >= F && ntab >= 1, "lstm_pack: bad sizes");
    const int has_tab = IN > F;
    GGPM_REQUIRE(has_tab ? ntab == IN - F : ntab == 1, "lstm_pack: ntab must be IN - F (or 1 without one-hot columns)");
    GGPM_REQUIRE(n_bias_rows >= 1 && n_bias_rows <= ntab, "lstm_pack: n_bias_rows out of range");
    GateW g;
    for (int i = 0; i < 4; ++i) {
        GGPM_REQUIRE(W4[i] && b4[i], "lstm_pack: null gate pointer");
        g.w[i] = W4[i];
        g.b[i] = b4[i];
    }
    const long long n = (long long)4 * Hp * Fp + (long long)ntab * 4 * Hp + (long long)4 * Hp * Hp;
    lstm_pack_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(g, H, Hp, F, Fp, IN, ntab, has_tab, n_bias_rows, Wx, T, U4);
    return check_launch("lstm_pack");
}

extern "C" int ggpm_lstm_unpack(float* const* dW4, float* const* db4, int H, int Hp, int F, int Fp, int IN, int n_bias_rows,
                                const float* dWx, const float* dT, const float* dU4, void* stream) {
    GGPM_REQUIRE(dW4 && db4, "lstm_unpack: null pointer");
    GateG g;
    for (int i = 0; i < 4; ++i) {
        GGPM_REQUIRE(dW4[i] && db4[i], "lstm_unpack: null gate pointer");
        g.w[i] = dW4[i];
        g.b[i] = db4[i];
    }
    const long long n = (long long)4 * H * (IN + H) + 4 * H;
    lstm_unpack_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(g, H, Hp, F, Fp, IN, IN > F, n_bias_rows, dWx, dT, dU4);
    return check_launch("lstm_unpack");
}

static int fill_pieces(Pieces& pc, int n_pieces, const int* c0, const int* c1, const int* kp, float* const* out, int K) {
    GGPM_REQUIRE(n_pieces >= 1 && n_pieces <= kMaxPieces && c0 && c1 && kp && out, "split_pad: 1..%d pieces", kMaxPieces);
    pc.n = n_pieces;
    for (int p = 0; p < n_pieces; ++p) {
        GGPM_REQUIRE(c0[p] >= 0 && c1[p] >= c0[p] && c1[p] <= K && kp[p] >= c1[p] - c0[p], "split_pad: bad piece %d", p);
        pc.c0[p] = c0[p]; pc.c1[p] = c1[p]; pc.kp[p] = kp[p]; pc.out[p] = out[p];
    }
    return 0;
}

extern "C" int ggpm_split_pad(const float* W, int ldw, int N, int K, int Np, int n_pieces, const int* c0, const int* c1,
                              const int* kp, float* const* out, const float* bias, float* bias_p, void* stream) {
    GGPM_REQUIRE(W && N >= 1 && Np >= N && ldw >= K, "split_pad: bad arguments");
    Pieces pc;
    GGPM_TRY(fill_pieces(pc, n_pieces, c0, c1, kp, out, K));
    long long n = Np;
    for (int p = 0; p < n_pieces; ++p) {
        GGPM_REQUIRE(out[p] != nullptr, "split_pad: null output");
        n += (long long)Np * kp[p];
    }
    split_pad_kernel<<<blocks_for(n), 256, 0, (cudaStream_t)stream>>>(W, ldw, N, Np, pc, bias, bias_p);
    return check_launch("split_pad");
}

extern "C" int ggpm_split_unpad(float* dW, int N, int K, int n_pieces, const int* c0, const int* c1, const int* kp,
                                float* const* dpieces, const float* dbias_p, float* dbias, void* stream) {
    GGPM_REQUIRE(dW && N >= 1 && K >= 1, "split_unpad: bad arguments");
    Pieces pc;
    GGPM_TRY(fill_pieces(pc, n_pieces, c0, c1, kp, dpieces, K));
    split_unpad_kernel<<<blocks_for((long long)N * K + N), 256, 0, (cudaStream_t)stream>>>(dW, K, N, pc, dbias_p, dbias);
    return check_launch("split_unpad");
}
