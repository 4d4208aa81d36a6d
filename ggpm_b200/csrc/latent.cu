// Latent head on the encoder output: the element-wise part of
// PropertyVAE.rsample / HierVAE.rsample (ggpm/property_vae.py:26-34, ggpm/hpgnn.py) after the two [B,H]->[B,L]
// projections (which run as ONE fused linear over [W_mean; W_var] on the library's GEMM path):
//     z_log_var = -|v|          kl = -0.5 * sum(1 + z_log_var - m^2 - exp(z_log_var)) / B
//     z = m + exp(z_log_var / 2) * eps        (perturb)      |      z = m      (no perturbation)
// and its hand-written backward.  Y holds [m | v] per row (v starts at column var_off).  The KL sum is deterministic:
// one warp per row writes the row sum, one block adds the rows in a fixed order.
#include "common.cuh"

namespace ggpm {

__global__ void __launch_bounds__(256)
rsample_fwd_kernel(const float* __restrict__ Y, int ldy, int var_off, const float* __restrict__ eps, int ld_eps, int B,
                   int L, float* __restrict__ z, int ldz, float* __restrict__ row_kl) {
    const int lane = threadIdx.x & 31;
    const int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (b >= B) return;
    const float* y = Y + (size_t)b * ldy;
    float acc = 0.f;
    for (int j = lane; j < L; j += 32) {
        const float m = y[j], lv = -fabsf(y[var_off + j]);
        acc += 1.0f + lv - m * m - expf(lv);
        z[(size_t)b * ldz + j] = eps != nullptr ? fmaf(expf(0.5f * lv), eps[(size_t)b * ld_eps + j], m) : m;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) row_kl[b] = acc;
}

__global__ void __launch_bounds__(256) rsample_kl_kernel(const float* __restrict__ row_kl, int B, float* __restrict__ kl) {
    __shared__ double part[256];
    double s = 0.0;
    for (int b = threadIdx.x; b < B; b += 256) s += (double)row_kl[b];
    part[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) part[threadIdx.x] += part[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) *kl = (float)(-0.5 * part[0] / (double)B);
}

// dY = [dm | dv | 0...]:  dm = dz + dkl * m / B ;  dlv = dz * eps * 0.5 * exp(lv / 2) - 0.5 * dkl / B * (1 - exp(lv)) ;
// dv = -sign(v) * dlv   (d|v|/dv = 0 at v = 0, as torch.abs)
__global__ void __launch_bounds__(256)
rsample_bwd_kernel(const float* __restrict__ Y, int ldy, int var_off, const float* __restrict__ eps, int ld_eps, int B,
                   int L, const float* __restrict__ dz, int ld_dz, const float* __restrict__ dkl, float* __restrict__ dY,
                   int ld_dy, int width) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)B * width) return;
    const int b = (int)(i / width), c = (int)(i % width);
    const float gk = dkl != nullptr ? __ldg(dkl) / (float)B : 0.f;
    float out = 0.f;
    if (c < L) {
        const float g = dz != nullptr ? dz[(size_t)b * ld_dz + c] : 0.f;
        out = g + gk * Y[(size_t)b * ldy + c];
    } else if (c >= var_off && c < var_off + L) {
        const int j = c - var_off;
        const float v = Y[(size_t)b * ldy + c];
        const float lv = -fabsf(v);
        float dlv = -0.5f * gk * (1.0f - expf(lv));
        if (eps != nullptr && dz != nullptr) dlv += dz[(size_t)b * ld_dz + j] * eps[(size_t)b * ld_eps + j] * 0.5f * expf(0.5f * lv);
        out = v > 0.f ? -dlv : (v < 0.f ? dlv : 0.f);
    }
    dY[(size_t)b * ld_dy + c] = out;
}

}  // namespace ggpm

using namespace ggpm;

extern "C" int ggpm_rsample_fwd(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L, float* z,
                                int ldz, float* kl, float* row_kl, void* stream) {
    GGPM_REQUIRE(Y && z && kl && row_kl, "rsample_fwd: null pointer");
    GGPM_REQUIRE(B >= 1 && L >= 1 && var_off >= L && ldy >= var_off + L && ldz >= L, "rsample_fwd: bad sizes");
    cudaStream_t st = (cudaStream_t)stream;
    rsample_fwd_kernel<<<ceil_div(B, 8), 256, 0, st>>>(Y, ldy, var_off, eps, ld_eps, B, L, z, ldz, row_kl);
    rsample_kl_kernel<<<1, 256, 0, st>>>(row_kl, B, kl);
    return check_launch("rsample_fwd", 2);
}

// ---- allocation-free variant ----------------------------------------------------------------------------------------
// Same two kernels as ggpm_rsample_fwd, with the per-row KL scratch in a file-scope device array instead of a caller
// buffer, for callers that cannot allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// The array is part of the module image: every GPU that loads the library holds kRowKlMax floats (256 KiB) until exit.
// All calls on a device share it and it keeps the last call's contents; every element a call reads (rows [0, B)) is
// written by that call's first kernel before its second kernel reads it, so nothing needs clearing — but two calls
// whose kernels overlap on the GPU (different streams, no ordering) race on it.
constexpr int kRowKlMax = 65536;
static __device__ __align__(256) float g_row_kl_scratch[kRowKlMax];

constexpr int kMaxDevices = 64;
static float* g_row_kl_addr[kMaxDevices];          // device address of the array, per device (0: not looked up yet)

// In host code `g_row_kl_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, asked once per
// device with that device current.
static int row_kl_scratch(float** out) {
    int dev = -1;
    GGPM_CUDA(cudaGetDevice(&dev));
    GGPM_REQUIRE(dev >= 0 && dev < kMaxDevices, "rsample_fwd_noalloc: device ordinal %d out of range", dev);
    float* p = __atomic_load_n(&g_row_kl_addr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        GGPM_CUDA(cudaGetSymbolAddress((void**)&p, g_row_kl_scratch));
        GGPM_REQUIRE(p != nullptr, "rsample_fwd_noalloc: no device address for the scratch array");
        __atomic_store_n(&g_row_kl_addr[dev], p, __ATOMIC_RELEASE);      // same value from every thread
    }
    *out = p;
    return 0;
}

extern "C" int ggpm_rsample_fwd_noalloc(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L,
                                        float* z, int ldz, float* kl, void* stream) {
    GGPM_REQUIRE(Y && z && kl, "rsample_fwd_noalloc: null pointer");
    GGPM_REQUIRE(B >= 1 && L >= 1 && var_off >= L && ldy >= var_off + L && ldz >= L, "rsample_fwd_noalloc: bad sizes");
    GGPM_REQUIRE(B <= kRowKlMax, "rsample_fwd_noalloc: %d rows exceed the built-in scratch (at most %d); use ggpm_rsample_fwd "
                 "with a caller-provided row_kl buffer", B, kRowKlMax);
    float* row_kl = nullptr;
    GGPM_TRY(row_kl_scratch(&row_kl));
    cudaStream_t st = (cudaStream_t)stream;
    rsample_fwd_kernel<<<ceil_div(B, 8), 256, 0, st>>>(Y, ldy, var_off, eps, ld_eps, B, L, z, ldz, row_kl);
    rsample_kl_kernel<<<1, 256, 0, st>>>(row_kl, B, kl);
    return check_launch("rsample_fwd_noalloc", 2);
}

extern "C" int ggpm_rsample_bwd(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L,
                                const float* dz, int ld_dz, const float* dkl, float* dY, int ld_dy, int width, void* stream) {
    GGPM_REQUIRE(Y && dY, "rsample_bwd: null pointer");
    GGPM_REQUIRE(B >= 1 && L >= 1 && var_off >= L && width >= var_off + L && ldy >= width && ld_dy >= width, "rsample_bwd: bad sizes");
    const long long n = (long long)B * width;
    rsample_bwd_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(Y, ldy, var_off, eps, ld_eps, B, L, dz, ld_dz,
                                                                                     dkl, dY, ld_dy, width);
    return check_launch("rsample_bwd");
}
