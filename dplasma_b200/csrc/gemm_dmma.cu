// Host launchers for the DMMA GEMM engine (see gemm_dmma.cuh).
#include "gemm_dmma.cuh"
#include <cstdio>
#include <cstdlib>

namespace dpb {

int gemm_init_attributes() {
    cudaError_t e;
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CfgBig>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgBig::SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CfgWide>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgWide::SMEM);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CfgSmall>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CfgSmall::SMEM);
    return (int)e;
}

void gemm_cfg_dims(int cfg, int* bm, int* bn) {
    if (cfg == CFG_BIG) { *bm = CfgBig::BM; *bn = CfgBig::BN; }
    else if (cfg == CFG_WIDE) { *bm = CfgWide::BM; *bn = CfgWide::BN; }
    else { *bm = CfgSmall::BM; *bn = CfgSmall::BN; }
}

// Tile configuration for a problem that would be `bigItems` CTA tiles of 128x128.
// DPB_GEMM_CFG=big|wide|small overrides (profiling).
int gemm_pick_cfg(long long bigItems) {
    static int forced = -2;
    if (forced == -2) {
        const char* e = getenv("DPB_GEMM_CFG");
        forced = -1;
        if (e) {
            if (e[0] == 'b') forced = CFG_BIG;
            else if (e[0] == 'w') forced = CFG_WIDE;
            else if (e[0] == 's') forced = CFG_SMALL;
        }
    }
    if (forced >= 0) return forced;
    // Measured on B200 (profiles/): the 64x64 tile at 3 CTAs/SM beats the larger tiles at every
    // size (34.9 vs 32.6 / 32.3 TFLOP/s on 16384^2 x 512): co-resident CTAs hide each other's
    // prologue and epilogue, which a single resident CTA cannot.
    (void)bigItems;
    return CFG_SMALL;
}

int gemm_launch(const GemmDesc& d, int nItems, int cfg, cudaStream_t stream) {
    if (nItems <= 0) return 0;
    if (cfg == CFG_SMALL)
        gemm_dmma_kernel<CfgSmall><<<nItems, CfgSmall::THREADS, CfgSmall::SMEM, stream>>>(d);
    else if (cfg == CFG_WIDE)
        gemm_dmma_kernel<CfgWide><<<nItems, CfgWide::THREADS, CfgWide::SMEM, stream>>>(d);
    else
        gemm_dmma_kernel<CfgBig><<<nItems, CfgBig::THREADS, CfgBig::SMEM, stream>>>(d);
    return (int)cudaGetLastError();
}

}  // namespace dpb
