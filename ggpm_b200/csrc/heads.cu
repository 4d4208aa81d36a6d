// Property heads behind the latent head: PropertyOptimizer / PropertyRegressor of
// ggpm/property_optimizer.py:5-66 — two independent MLPs (HOMO, LUMO) on the two halves of the latent vector,
// Linear(In, Hd) + ReLU + Dropout + Linear(Hd, 1), each followed by a mean-squared-error loss (:21-23).
// Both heads run in ONE launch per direction (grid.y = head).  Dropout stays a torch op for RNG parity: the caller
// passes the multiplier the reference's nn.Dropout would have applied (mask / (1 - p)) or NULL.
//     a   = relu(W1 x + b1)                      [B, Hd]   (saved: the backward needs it)
//     y   = w2 . (a * mult) + b2                 [B]
//     loss = mean_b (y_b - t_b)^2                (deterministic two-stage sum; skipped when t is NULL)
#include "common.cuh"

namespace ggpm {

constexpr int kHeadRows = 8;          // rows per block
constexpr int kHeadThreads = 128;
constexpr int kHeadMaxIn = 64;
constexpr int kHeadMaxW1 = 8192;      // Hd * In floats of W1 staged in shared memory

struct HeadArgs {        // one MLP head
    const float* x; int ldx;                     // [B, In]
    const float* W1; const float* b1;            // [Hd, In], [Hd]
    const float* w2; const float* b2;            // [Hd], [1]
    const float* mult; int ld_mult;              // [B, Hd] dropout multiplier or NULL
    const float* target;                         // [B] or NULL
    float* hid;                                  // [B, Hd] post-ReLU activations (saved)
    float* out;                                  // [B]
    float* sq;                                   // [B] squared errors (scratch)
    float* loss;                                 // [1]
    // backward
    const float* dout; const float* dloss;       // [B] or NULL, device scalar or NULL
    float* dx; int ld_dx;                        // [B, In]
    float* dW1; float* db1; float* dw2; float* db2;   // accumulated into (caller zeroes)
};
struct HeadPair { HeadArgs h[2]; };

__global__ void __launch_bounds__(kHeadThreads)
mlp_head_fwd_kernel(HeadPair hp, int B, int In, int Hd) {
    extern __shared__ float smem[];
    float* sW = smem;                              // [Hd][In]
    float* sx = sW + Hd * In;                      // [kHeadRows][In]
    float* red = sx + kHeadRows * In;              // [kHeadRows][kHeadThreads / 32]
    const HeadArgs& a = hp.h[blockIdx.y];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int r0 = blockIdx.x * kHeadRows;
    for (int i = tid; i < Hd * In; i += kHeadThreads) sW[i] = __ldg(a.W1 + i);
    for (int i = tid; i < kHeadRows * In; i += kHeadThreads) {
        const int r = r0 + i / In;
        sx[i] = r < B ? a.x[(size_t)r * a.ldx + i % In] : 0.f;
    }
    __syncthreads();
    float part[kHeadRows];
#pragma unroll
    for (int r = 0; r < kHeadRows; ++r) part[r] = 0.f;
    for (int j = tid; j < Hd; j += kHeadThreads) {
        const float bj = __ldg(a.b1 + j), wj = __ldg(a.w2 + j);
        const float* w = sW + j * In;
#pragma unroll
        for (int r = 0; r < kHeadRows; ++r) {
            if (r0 + r >= B) break;
            float acc = bj;
            for (int k = 0; k < In; ++k) acc = fmaf(w[k], sx[r * In + k], acc);
            acc = fmaxf(acc, 0.f);
            a.hid[(size_t)(r0 + r) * Hd + j] = acc;
            const float m = a.mult != nullptr ? a.mult[(size_t)(r0 + r) * a.ld_mult + j] : 1.f;
            part[r] = fmaf(wj, acc * m, part[r]);
        }
    }
#pragma unroll
    for (int r = 0; r < kHeadRows; ++r) {
        float v = part[r];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) red[r * (kHeadThreads / 32) + wid] = v;
    }
    __syncthreads();
    if (tid < kHeadRows && r0 + tid < B) {
        float y = __ldg(a.b2);
        for (int w = 0; w < kHeadThreads / 32; ++w) y += red[tid * (kHeadThreads / 32) + w];
        a.out[r0 + tid] = y;
        if (a.target != nullptr) {
            const float d = y - a.target[r0 + tid];
            a.sq[r0 + tid] = d * d;
        }
    }
}

__global__ void __launch_bounds__(256) mlp_head_loss_kernel(HeadPair hp, int B) {
    __shared__ double part[256];
    const HeadArgs& a = hp.h[blockIdx.x];
    if (a.target == nullptr) return;
    double s = 0.0;
    for (int b = threadIdx.x; b < B; b += 256) s += (double)a.sq[b];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *a.loss = (float)(part[0] / (double)B);
}

// dy_b = dout_b + dloss * 2 (y_b - t_b) / B ;  da = dy w2 mult [a > 0] ;  dW1 += da^T x, db1 += sum da, dw2 += sum dy a mult,
// db2 += sum dy, dx = da W1.  Thread j owns hidden unit j: its weight-gradient row is accumulated in registers over the
// block's rows and leaves the block as ONE atomicAdd per element.
constexpr int kHeadBwdRows = 32;
__global__ void __launch_bounds__(kHeadThreads)
mlp_head_bwd_kernel(HeadPair hp, int B, int In, int Hd) {
    extern __shared__ float smem[];
    float* sW = smem;                              // [Hd][In]
    float* sx = sW + Hd * In;                      // [kHeadBwdRows][In]
    float* sdy = sx + kHeadBwdRows * In;           // [kHeadBwdRows]
    float* sdx = sdy + kHeadBwdRows;               // [kHeadBwdRows][In] accumulated with shared atomics
    const HeadArgs& a = hp.h[blockIdx.y];
    const int tid = threadIdx.x;
    const int r0 = blockIdx.x * kHeadBwdRows;
    const int nr = min(kHeadBwdRows, B - r0);
    for (int i = tid; i < Hd * In; i += kHeadThreads) sW[i] = __ldg(a.W1 + i);
    for (int i = tid; i < kHeadBwdRows * In; i += kHeadThreads) {
        const int r = r0 + i / In;
        sx[i] = r < B ? a.x[(size_t)r * a.ldx + i % In] : 0.f;
        sdx[i] = 0.f;
    }
    if (tid < kHeadBwdRows) {
        float dy = 0.f;
        const int r = r0 + tid;
        if (r < B) {
            if (a.dout != nullptr) dy += a.dout[r];
            if (a.dloss != nullptr && a.target != nullptr) dy += __ldg(a.dloss) * 2.f * (a.out[r] - a.target[r]) / (float)B;
        }
        sdy[tid] = dy;
    }
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
        for (int r = 0; r < nr; ++r) s += sdy[r];
        if (s != 0.f) atomicAdd(a.db2, s);
    }
    for (int j = tid; j < Hd; j += kHeadThreads) {
        const float wj = __ldg(a.w2 + j);
        const float* w = sW + j * In;
        float gw[kHeadMaxIn];
#pragma unroll
        for (int k = 0; k < kHeadMaxIn; ++k) gw[k] = 0.f;
        float g2 = 0.f, g1 = 0.f;
        for (int r = 0; r < nr; ++r) {
            const float act = a.hid[(size_t)(r0 + r) * Hd + j];
            const float m = a.mult != nullptr ? a.mult[(size_t)(r0 + r) * a.ld_mult + j] : 1.f;
            const float dy = sdy[r];
            g2 = fmaf(dy, act * m, g2);
            const float da = act > 0.f ? dy * wj * m : 0.f;
            if (da != 0.f) {
                g1 += da;
#pragma unroll
                for (int k = 0; k < kHeadMaxIn; ++k)
                    if (k < In) {
                        gw[k] = fmaf(da, sx[r * In + k], gw[k]);
                        atomicAdd(sdx + r * In + k, da * w[k]);
                    }
            }
        }
#pragma unroll
        for (int k = 0; k < kHeadMaxIn; ++k)
            if (k < In && gw[k] != 0.f) atomicAdd(a.dW1 + (size_t)j * In + k, gw[k]);
        if (g2 != 0.f) atomicAdd(a.dw2 + j, g2);
        if (g1 != 0.f) atomicAdd(a.db1 + j, g1);
    }
    __syncthreads();
    if (a.dx != nullptr)
        for (int i = tid; i < kHeadBwdRows * In; i += kHeadThreads) {
            const int r = r0 + i / In;
            if (r < B) a.dx[(size_t)r * a.ld_dx + i % In] = sdx[i];
        }
}

}  // namespace ggpm

using namespace ggpm;

static int head_pair_from(const void* const* p, HeadPair& hp) {
    for (int h = 0; h < 2; ++h) {
        const void* const* q = p + h * 17;
        HeadArgs& a = hp.h[h];
        a.x = (const float*)q[0]; a.W1 = (const float*)q[1]; a.b1 = (const float*)q[2]; a.w2 = (const float*)q[3];
        a.b2 = (const float*)q[4]; a.mult = (const float*)q[5]; a.target = (const float*)q[6]; a.hid = (float*)q[7];
        a.out = (float*)q[8]; a.sq = (float*)q[9]; a.loss = (float*)q[10]; a.dout = (const float*)q[11];
        a.dloss = (const float*)q[12]; a.dx = (float*)q[13]; a.dW1 = (float*)q[14]; a.db1 = (float*)q[15];
        a.dw2 = (float*)q[16];
    }
    return 0;
}

// ptrs: 2 x 17 device pointers (head 0 = HOMO, head 1 = LUMO) in the order of HeadArgs' pointer members
// {x, W1, b1, w2, b2, mult, target, hid, out, sq, loss, dout, dloss, dx, dW1, db1, dw2}; db2: [2] (one per head).
extern "C" int ggpm_mlp_heads_fwd(const void* const* ptrs, int ldx0, int ldx1, int ld_mult, int B, int In, int Hd, void* stream) {
    GGPM_REQUIRE(ptrs && B >= 1 && In >= 1 && Hd >= 1, "mlp_heads_fwd: bad arguments");
    GGPM_REQUIRE(In <= kHeadMaxIn && Hd * In <= kHeadMaxW1, "mlp_heads_fwd: head too large for the fused kernel (In <= %d, Hd * In <= %d)",
                 kHeadMaxIn, kHeadMaxW1);
    HeadPair hp;
    head_pair_from(ptrs, hp);
    hp.h[0].ldx = ldx0; hp.h[1].ldx = ldx1;
    for (int h = 0; h < 2; ++h) {
        hp.h[h].ld_mult = ld_mult;
        GGPM_REQUIRE(hp.h[h].x && hp.h[h].W1 && hp.h[h].b1 && hp.h[h].w2 && hp.h[h].b2 && hp.h[h].hid && hp.h[h].out,
                     "mlp_heads_fwd: null pointer");
        GGPM_REQUIRE(hp.h[h].target == nullptr || (hp.h[h].sq && hp.h[h].loss), "mlp_heads_fwd: targets need sq / loss buffers");
    }
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = sizeof(float) * ((size_t)Hd * In + kHeadRows * In + kHeadRows * (kHeadThreads / 32));
    mlp_head_fwd_kernel<<<dim3(ceil_div(B, kHeadRows), 2), kHeadThreads, smem, st>>>(hp, B, In, Hd);
    mlp_head_loss_kernel<<<2, 256, 0, st>>>(hp, B);
    return check_launch("mlp_heads_fwd", 2);
}

extern "C" int ggpm_mlp_heads_bwd(const void* const* ptrs, float* db2, int ldx0, int ldx1, int ld_mult, int ld_dx, int B, int In,
                                  int Hd, void* stream) {
    GGPM_REQUIRE(ptrs && db2 && B >= 1 && In >= 1 && Hd >= 1, "mlp_heads_bwd: bad arguments");
    GGPM_REQUIRE(In <= kHeadMaxIn && Hd * In <= kHeadMaxW1, "mlp_heads_bwd: head too large for the fused kernel");
    HeadPair hp;
    head_pair_from(ptrs, hp);
    hp.h[0].ldx = ldx0; hp.h[1].ldx = ldx1;
    cudaStream_t st = (cudaStream_t)stream;
    for (int h = 0; h < 2; ++h) {
        HeadArgs& a = hp.h[h];
        a.ld_mult = ld_mult; a.ld_dx = ld_dx; a.db2 = db2 + h;
        GGPM_REQUIRE(a.x && a.W1 && a.w2 && a.hid && a.out && a.dW1 && a.db1 && a.dw2, "mlp_heads_bwd: null pointer");
        GGPM_CUDA(cudaMemsetAsync(a.dW1, 0, sizeof(float) * (size_t)Hd * In, st));
        GGPM_CUDA(cudaMemsetAsync(a.db1, 0, sizeof(float) * (size_t)Hd, st));
        GGPM_CUDA(cudaMemsetAsync(a.dw2, 0, sizeof(float) * (size_t)Hd, st));
    }
    GGPM_CUDA(cudaMemsetAsync(db2, 0, sizeof(float) * 2, st));
    const size_t smem = sizeof(float) * ((size_t)Hd * In + 2 * kHeadBwdRows * In + kHeadBwdRows);
    mlp_head_bwd_kernel<<<dim3(ceil_div(B, kHeadBwdRows), 2), kHeadThreads, smem, st>>>(hp, B, In, Hd);
    return check_launch("mlp_heads_bwd");
}
