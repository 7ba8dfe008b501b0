// Host launchers for the DMMA GEMM engine (see gemm_dmma.cuh).
#include "gemm_dmma.cuh"
#include <cstdio>

namespace dpb {

int gemm_init_attributes() {
    cudaError_t e;
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CfgBig>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)CfgBig::SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CfgSmall>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)CfgSmall::SMEM);
    return (int)e;
}

int gemm_launch(const GemmDesc& d, int nItems, bool smallCfg, cudaStream_t stream) {
    if (nItems <= 0) return 0;
    if (smallCfg)
        gemm_dmma_kernel<CfgSmall><<<nItems, CfgSmall::THREADS, CfgSmall::SMEM, stream>>>(d);
    else
        gemm_dmma_kernel<CfgBig><<<nItems, CfgBig::THREADS, CfgBig::SMEM, stream>>>(d);
    return (int)cudaGetLastError();
}

}  // namespace dpb
