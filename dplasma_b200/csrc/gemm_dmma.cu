// Host launchers for the DMMA GEMM engine (see gemm_dmma.cuh).
#include "gemm_dmma.cuh"
#include <cstdio>
#include <cstdlib>

namespace dpb {

int gemm_init_attributes() {
    cudaError_t e;
#define DPB_ATTR(CFG, AK, BK)                                                                     \
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CFG, AK, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                             (int)CFG::smem(AK, BK));                                                    \
    if (e != cudaSuccess) return (int)e;
    DPB_ATTR(CfgBig, false, false)
    DPB_ATTR(CfgWide, false, false)
    DPB_ATTR(CfgSmall, false, false)
    DPB_ATTR(CfgSmall, true, false)
    DPB_ATTR(CfgSmall, false, true)
    DPB_ATTR(CfgSmall, true, true)
#undef DPB_ATTR
    return 0;
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
    // operands stored K-major (transposed forms) exist for the 64x64 tile only
    if (d.a_kmajor || d.b_kmajor) {
        if (d.a_kmajor && d.b_kmajor)
            gemm_dmma_kernel<CfgSmall, true, true><<<nItems, CfgSmall::THREADS, CfgSmall::smem(true, true), stream>>>(d);
        else if (d.a_kmajor)
            gemm_dmma_kernel<CfgSmall, true, false><<<nItems, CfgSmall::THREADS, CfgSmall::smem(true, false), stream>>>(d);
        else
            gemm_dmma_kernel<CfgSmall, false, true><<<nItems, CfgSmall::THREADS, CfgSmall::smem(false, true), stream>>>(d);
    } else if (cfg == CFG_SMALL)
        gemm_dmma_kernel<CfgSmall, false, false><<<nItems, CfgSmall::THREADS, CfgSmall::SMEM, stream>>>(d);
    else if (cfg == CFG_WIDE)
        gemm_dmma_kernel<CfgWide, false, false><<<nItems, CfgWide::THREADS, CfgWide::SMEM, stream>>>(d);
    else
        gemm_dmma_kernel<CfgBig, false, false><<<nItems, CfgBig::THREADS, CfgBig::SMEM, stream>>>(d);
    return (int)cudaGetLastError();
}

}  // namespace dpb
