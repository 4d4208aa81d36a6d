// Shared helpers for the ggpm_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace ggpm {

constexpr int kWarp = 32;

// ---- error plumbing (C-ABI functions return int; ggpm_last_error() gives the text) -------------
void set_error(const char* fmt, ...);
int check_launch(const char* what, int n_kernels = 1);   // cudaGetLastError() -> status; counts launches

#define GGPM_REQUIRE(cond, ...)                         \
    do {                                                \
        if (!(cond)) {                                  \
            ::ggpm::set_error(__VA_ARGS__);             \
            return 1;                                   \
        }                                               \
    } while (0)

#define GGPM_TRY(expr)                                  \
    do {                                                \
        int _st = (expr);                               \
        if (_st != 0) return _st;                       \
    } while (0)

#define GGPM_CUDA(expr)                                                               \
    do {                                                                              \
        cudaError_t _e = (expr);                                                      \
        if (_e != cudaSuccess) {                                                      \
            ::ggpm::set_error("%s failed: %s", #expr, cudaGetErrorString(_e));        \
            return 2;                                                                 \
        }                                                                             \
    } while (0)

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

int sm_count();   // cached cudaDevAttrMultiProcessorCount of the current device

// ---- optional per-launch timing (bench.py roofline pass): CUDA events around launches of a class ----
enum ProfClass { PROF_GATHER = 0, PROF_GEMM_TC = 1, PROF_GEMM_SIMT = 2, PROF_GEMM_WGRAD = 3, PROF_CELL_FWD = 4,
                 PROF_CELL_BWD = 5, PROF_MISC = 6, PROF_NCLASS = 7 };
struct ProfScope {   // no-op unless ggpm_prof_enable(1) was called
    int slot;
    cudaStream_t st;
    ProfScope(int cls, cudaStream_t stream, double work);
    ~ProfScope();
};

// ---- programmatic dependent launch: the kernels of a depth loop are launched with stream serialisation relaxed, so
// the next grid is scheduled (and runs its setup) while the current one drains; every such kernel calls pdl_wait()
// before its first global access, which blocks until the preceding grid has completed and flushed ----
bool pdl_enabled();   // GGPM_PDL=0 turns it off (A/B timing)

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = at; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// ---- device helpers ---------------------------------------------------------------------------
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ float4 ldg4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }
__device__ __forceinline__ float4 operator+(float4 a, float4 b) {
    return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
}
__device__ __forceinline__ float4 operator-(float4 a, float4 b) {
    return make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
}
__device__ __forceinline__ float4 operator*(float4 a, float4 b) {
    return make_float4(a.x * b.x, a.y * b.y, a.z * b.z, a.w * b.w);
}
__device__ __forceinline__ float4& operator+=(float4& a, float4 b) {
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    return a;
}
__device__ __forceinline__ float4 fma4(float4 a, float4 b, float4 c) {
    return make_float4(fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y), fmaf(a.z, b.z, c.z), fmaf(a.w, b.w, c.w));
}
__device__ __forceinline__ float amax4(float4 a) {
    return fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(a.w)));
}
// accurate fp32 transcendental forms (parity bar is 1e-5 against an fp64 oracle: no fast-math)
__device__ __forceinline__ float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float4 sigmoid4(float4 a) {
    return make_float4(sigmoidf_(a.x), sigmoidf_(a.y), sigmoidf_(a.z), sigmoidf_(a.w));
}
__device__ __forceinline__ float4 tanh4(float4 a) {
    return make_float4(tanhf(a.x), tanhf(a.y), tanhf(a.z), tanhf(a.w));
}
// vector reduction into global memory (sm_90+): one 16-byte RED instead of four
__device__ __forceinline__ void red_add4(float* p, float4 v) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}
// zero the lanes of a 4-chunk starting at column c that fall at or beyond `H`
__device__ __forceinline__ float4 clip_cols(float4 v, int c, int H) {
    if (c + 0 >= H) v.x = 0.f;
    if (c + 1 >= H) v.y = 0.f;
    if (c + 2 >= H) v.z = 0.f;
    if (c + 3 >= H) v.w = 0.f;
    return v;
}

}  // namespace ggpm
