// Host launchers for the DMMA GEMM engine (see gemm_dmma.cuh).
#include <algorithm>
#include <cstdlib>
#include "gemm_dmma.cuh"
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace dpb {

int gemm_init_attributes() {
    cudaError_t e;
#define DPB_ATTR(CFG, AK, BK)                                                                     \
    e = cudaFuncSetAttribute(gemm_dmma_kernel<CFG, AK, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                             (int)CFG::smem(AK, BK));                                                    \
    if (e != cudaSuccess) return (int)e;
    DPB_ATTR(CfgSmall, false, false)
    DPB_ATTR(CfgSmall, true, false)
    DPB_ATTR(CfgSmall, false, true)
    DPB_ATTR(CfgSmall, true, true)
#undef DPB_ATTR
    return 0;
}

void gemm_cfg_dims(int cfg, int* bm, int* bn) {
    (void)cfg;
    *bm = CfgSmall::BM;
    *bn = CfgSmall::BN;
}

// One tile configuration: 64x64 CTA tiles at 3 CTAs/SM (measured in round 1 against 128x128 and 128x64:
// co-resident CTAs hide each other's prologue and epilogue, which one resident CTA cannot).
int gemm_pick_cfg(long long bigItems) {
    (void)bigItems;
    return CFG_SMALL;
}

// ---- tensor maps of the M-major operands ----
// cuTensorMapEncodeTiled is host arithmetic inside the driver library; it is resolved through the runtime so
// that the library keeps linking against nothing but cudart.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encoder() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = (EncodeTiledFn)p;
        else
            fprintf(stderr, "dplasma_b200: cuTensorMapEncodeTiled is not available from this driver\n");
    }
    return fn;
}

// A 4-D view {rows, k, tile, k-tile} of an operand stored M-major: element (r, kk) of tile t, k-tile q at
// base + r + kk*ld + t*tileStride + q*kTileStride (doubles).  Box = {16 rows, BK}, 128-byte swizzle.
static bool encode_operand(CUtensorMap* tm, const double* base, long long rows, long long kdim, long long ld,
                           long long tiles, long long tileStride, long long ktiles, long long kTileStride) {
    EncodeTiledFn enc = encoder();
    if (!enc) return false;
    if (rows <= 0 || kdim <= 0) return false;
    if (tiles < 1) tiles = 1;
    if (ktiles < 1) ktiles = 1;
    // strides of dimensions that have extent 1 are never used, but must still be legal (non-zero multiples of 16 B)
    if (tiles == 1 || tileStride <= 0) tileStride = ld * kdim + (ld * kdim) % 2;
    if (ktiles == 1 || kTileStride <= 0) kTileStride = tileStride;
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld & 1) || (tileStride & 1) || (kTileStride & 1)) return false;
    const cuuint64_t dims[4] = {(cuuint64_t)rows, (cuuint64_t)kdim, (cuuint64_t)tiles, (cuuint64_t)ktiles};
    const cuuint64_t strides[3] = {(cuuint64_t)ld * 8, (cuuint64_t)tileStride * 8, (cuuint64_t)kTileStride * 8};
    const cuuint32_t box[4] = {(cuuint32_t)CfgSmall::MBLK, (cuuint32_t)CfgSmall::BK, 1, 1};
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, const_cast<double*>(base), dims, strides, box, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS;
}

// fills d.tmA / d.tmB for the operands stored M-major; false: this launch has to use the element-wise producer
static bool prepare_tensor_maps(GemmDesc& d) {
    const bool multiK = d.kTile < (1 << 30);
    const long long kdim = multiK ? d.kTile : d.K;
    const long long ktiles = multiK ? (d.K + d.kTile - 1) / d.kTile : 1;
    if (d.mode == GEMM_SINGLE) {
        if (!d.a_kmajor && !encode_operand(&d.tmA, d.A, d.M, kdim, d.lda, d.batchStrideA != 0 ? d.batch : 1, d.batchStrideA,
                                           ktiles, d.aKTileStride))
            return false;
        if (!d.b_kmajor && !encode_operand(&d.tmB, d.B, d.N, kdim, d.ldb, d.batchStrideB != 0 ? d.batch : 1,
                                           d.batchStrideB, ktiles, d.bKTileStride))
            return false;
        return true;
    }
    if (!encode_operand(&d.tmA, d.rowPanel, d.nb, kdim, d.nb, d.rowEnd, d.rowPanelStride, ktiles, d.aKTileStride)) return false;
    return encode_operand(&d.tmB, d.colPanel, d.nb, kdim, d.nb, d.colFirst + d.ncols, d.colPanelStride, ktiles,
                          d.bKTileStride);
}

int gemm_launch(const GemmDesc& din, int nItems, int cfg, cudaStream_t stream) {
    (void)cfg;
    if (nItems <= 0) return 0;
    GemmDesc d = din;
    if (d.fast && !(d.a_kmajor && d.b_kmajor) && !prepare_tensor_maps(d)) d.fast = 0;
    // work order (decode_work): DPB_GEMM_ORDER=0 keeps the strip-by-strip order of round 1 (A/B measurements)
    static const int order = [] { const char* e = getenv("DPB_GEMM_ORDER"); return e ? atoi(e) : -1; }();
    d.colGroup = d.blkGroup = 0;
    if (d.mode == GEMM_SINGLE) {
        d.blkGroup = order >= 0 ? order : 16;
    } else {
        int g = order >= 0 ? order : std::max(1, std::min(8, 4096 / std::max(d.nb, 1)));
        for (int c = 1; c < d.ncols && g > 0; c++)
            if (d.rowFirst[c] < d.rowFirst[c - 1]) g = 0;  // not a trapezoid: the level walk of decode_work does not apply
        d.colGroup = std::min(g, 8);
    }
    if (d.a_kmajor && d.b_kmajor)
        gemm_dmma_kernel<CfgSmall, true, true><<<nItems, CfgSmall::THREADS, CfgSmall::smem(true, true), stream>>>(d);
    else if (d.a_kmajor)
        gemm_dmma_kernel<CfgSmall, true, false><<<nItems, CfgSmall::THREADS, CfgSmall::smem(true, false), stream>>>(d);
    else if (d.b_kmajor)
        gemm_dmma_kernel<CfgSmall, false, true><<<nItems, CfgSmall::THREADS, CfgSmall::smem(false, true), stream>>>(d);
    else
        gemm_dmma_kernel<CfgSmall, false, false><<<nItems, CfgSmall::THREADS, CfgSmall::SMEM, stream>>>(d);
    return (int)cudaGetLastError();
}

}  // namespace dpb
