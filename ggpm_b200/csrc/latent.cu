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

extern "C" int ggpm_rsample_bwd(const float* Y, int ldy, int var_off, const float* eps, int ld_eps, int B, int L,
                                const float* dz, int ld_dz, const float* dkl, float* dY, int ld_dy, int width, void* stream) {
    GGPM_REQUIRE(Y && dY, "rsample_bwd: null pointer");
    GGPM_REQUIRE(B >= 1 && L >= 1 && var_off >= L && width >= var_off + L && ldy >= width && ld_dy >= width, "rsample_bwd: bad sizes");
    const long long n = (long long)B * width;
    rsample_bwd_kernel<<<(unsigned)ceil_div_ll(n, 256), 256, 0, (cudaStream_t)stream>>>(Y, ldy, var_off, eps, ld_eps, B, L, dz, ld_dz,
                                                                                     dkl, dY, ld_dy, width);
    return check_launch("rsample_bwd");
}
