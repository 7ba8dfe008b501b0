// Tile-level operations built on the DMMA engine: the recursive in-tile Cholesky, the panel
// triangular solve through inverted diagonal blocks, and the kernel-seam entry points that
// take the place of the reference's per-tile cuSOLVER / cuBLAS calls
// (src/zpotrf_L.jdf:143,236-240,342-346,463-468).
#include <cstdio>
#include <cstdlib>
#include "internal.h"

namespace dpb {

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// C = beta*C + alpha * op(A) * op(B)^T over a batch of equally strided problems.
//   aK / bK : operand stored K-major (op(A)(i,k) at A[k + i*lda]) instead of M-major (A[i + k*lda])
//   lower   : touch only i >= j (SYRK)
int gemm_single(parsec_context_s* ctx, cudaStream_t s, int M, int N, int K, double alpha, const double* A,
                int lda, bool aK, const double* B, int ldb, bool bK, double beta, double* C, int ldc,
                int lower, int batch, long long sA, long long sB, long long sC, int lastBatchM) {
    if (M <= 0 || N <= 0 || batch <= 0) return 0;
    if (alpha == 0.0) return -7;  // not used on this path
    if (ctx->cplx) {
        // complex double: interleaved tiles, C = beta C + alpha A B^H only (the forms zpotrf needs)
        if (aK || bK || lower == 2) return -8;
        GemmDesc z;
        z.mode = GEMM_SINGLE;
        z.alpha = alpha; z.beta = beta; z.K = K;
        z.a_kmajor = z.b_kmajor = 0; z.preload = 0;
        z.kTile = 1 << 30; z.aKTileStride = z.bKTileStride = 0;
        z.A = A; z.B = B; z.C = C;
        z.M = M; z.N = N; z.lda = lda; z.ldb = ldb; z.ldc = ldc;
        z.lower = lower; z.rowOff = z.colOff = 0;
        z.batch = batch; z.batchStrideA = sA; z.batchStrideB = sB; z.batchStrideC = sC;
        z.lastBatchM = lastBatchM < 0 ? M : lastBatchM;
        z.fast = ((K % 8 == 0) && aligned16(A) && aligned16(B) && (sA % 2 == 0) && (sB % 2 == 0)) ? 1 : 0;
        z.gridM = (M + 63) / 64;
        z.gridN = (N + 63) / 64;
        ctx->launches++;
        return zgemm_launch(z, z.gridM * z.gridN * batch, s);
    }
    if (lastBatchM < 0) lastBatchM = M;
    GemmDesc d;
    d.mode = GEMM_SINGLE;
    d.alpha = alpha; d.beta = beta; d.K = K;
    d.a_kmajor = aK; d.b_kmajor = bK;
    d.kTile = 1 << 30; d.aKTileStride = d.bKTileStride = 0;
    d.A = A; d.B = B; d.C = C;
    d.M = M; d.N = N; d.lda = lda; d.ldb = ldb; d.ldc = ldc;
    d.lower = lower; d.rowOff = d.colOff = 0;
    d.preload = ctx->gemmPreload;
    d.batch = batch; d.batchStrideA = sA; d.batchStrideB = sB; d.batchStrideC = sC;
    d.lastBatchM = lastBatchM;
    // TMA producer: tensor-map copies for M-major operands (any K: the part of a box beyond K arrives as zeros),
    // 128-byte bulk rows for K-major ones (whole 16-deep chunks only)
    d.fast = ((!(aK || bK) || K % 16 == 0) && (lda % 2 == 0) && (ldb % 2 == 0) && aligned16(A) && aligned16(B) &&
              (sA % 2 == 0) && (sB % 2 == 0))
                 ? 1 : 0;
    // tile configuration: the large CTA tiles only when they still fill the machine
    const long long bigItems = (long long)((M + 127) / 128) * ((N + 127) / 128) * batch;
    const int cfg = (aK || bK) ? CFG_SMALL : gemm_pick_cfg(bigItems);
    int bm, bn;
    gemm_cfg_dims(cfg, &bm, &bn);
    d.gridM = (M + bm - 1) / bm;
    d.gridN = (N + bn - 1) / bn;
    const long long items = (long long)d.gridM * d.gridN * batch;
    ctx->launches++;
    return gemm_launch(d, (int)items, cfg, s);
}

int ensure_winv(parsec_context_s* ctx, int nb) {
    // [inverses | factored diagonal blocks], sized for complex elements
    const size_t slot = 2 * 2 * (size_t)((nb + IB - 1) / IB) * IB * IB;
    if (ctx->wInv && ctx->wInvSlot >= slot) return 0;
    if (ctx->wInv) cudaFree(ctx->wInv);
    ctx->wInv = nullptr;
    DPB_CUDA(cudaMalloc(&ctx->wInv, 2 * slot * sizeof(double)));
    ctx->wInvSlot = slot;
    return 0;
}
int ensure_wnorm(parsec_context_s* ctx, size_t n) {
    if (ctx->wNorm && ctx->wNormLen >= n) return 0;
    if (ctx->wNorm) cudaFree(ctx->wNorm);
    ctx->wNorm = nullptr;
    DPB_CUDA(cudaMalloc(&ctx->wNorm, (n + 8) * sizeof(double)));
    ctx->wNormLen = n;
    return 0;
}

// split point of the recursion: a multiple of IB, about half
static inline int split_of(int n) {
    int blocks = (n + IB - 1) / IB;
    return (blocks / 2) * IB;
}

// B(:, j0:j0+n) <- B(:, j0:j0+n) * L(j0:j0+n, j0:j0+n)^-T, L lower triangular with the inverses
// of its IB x IB diagonal blocks in winv (block b at winv + b*IB*IB, ld IB).
static int trsm_rec(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, const double* winv, int j0,
                    int n, double* B, int ldb, int M, int batch, long long strideB, int lastBatchM) {
    const size_t ez = ctx->cplx ? 2 : 1;  // doubles per element
    if (n <= IB) {
        // in place: every CTA owns whole rows and has consumed all of A before its epilogue
        return gemm_single(ctx, s, M, n, n, 1.0, B + ez * (size_t)j0 * ldb, ldb, false,
                           winv + ez * (size_t)(j0 / IB) * IB * IB, IB, false, 0.0, B + ez * (size_t)j0 * ldb, ldb, 0,
                           batch, strideB, 0, strideB, lastBatchM);
    }
    const int n1 = split_of(n), n2 = n - n1;
    int rc = trsm_rec(ctx, s, L, ldl, winv, j0, n1, B, ldb, M, batch, strideB, lastBatchM);
    if (rc) return rc;
    rc = gemm_single(ctx, s, M, n2, n1, -1.0, B + ez * (size_t)j0 * ldb, ldb, false,
                     L + ez * ((j0 + n1) + (size_t)j0 * ldl), ldl, false, 1.0, B + ez * (size_t)(j0 + n1) * ldb, ldb, 0,
                     batch, strideB, 0, strideB, lastBatchM);
    if (rc) return rc;
    return trsm_rec(ctx, s, L, ldl, winv, j0 + n1, n2, B, ldb, M, batch, strideB, lastBatchM);
}

// B <- B L^-H through the recursion (both precisions), for callers that bring their own inverses
int trsm_rec_entry(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, const double* winv, int n, double* B,
                   int ldb, int M, int batch, long long strideB, int lastBatchM) {
    return trsm_rec(ctx, s, L, ldl, winv, 0, n, B, ldb, M, batch, strideB, lastBatchM);
}

int trsm_panel_lower(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, int n, const double* winv,
                     double* B, int ldb, int M, int batch, long long strideB, int lastBatchM) {
    if (batch <= 0 || M <= 0) return 0;
    static int mode = -1;  // DPB_TRSM=rec selects the recursive multi-launch variant
    if (mode < 0) {
        const char* e = getenv("DPB_TRSM");
        mode = (e && e[0] == 'r') ? 1 : 0;
    }
    // Many tiles below the diagonal: the step is bound by the bulk update and the panel has slack, so
    // the recursive form wins (its products run on the 3-CTA/SM GEMM engine at full rate).  Few tiles:
    // the panel IS the critical path and the single-launch fused kernel wins on latency.
    if (mode == 0 && !ctx->cplx && (long long)batch * M <= 20LL * 512)
        return launch_trsm_panel_fused(ctx, s, L, ldl, n, winv, B, ldb, strideB, M, lastBatchM, batch);
    return trsm_rec(ctx, s, L, ldl, winv, 0, n, B, ldb, M, batch, strideB, lastBatchM);
}

static int potrf_rec(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, double* winv, int j0, int n,
                     int* dInfo, int infoBase) {
    const size_t ez = ctx->cplx ? 2 : 1;
    double* Ajj = A + ez * (j0 + (size_t)j0 * lda);
    if (n <= IB) {
        double* w = winv + ez * (size_t)(j0 / IB) * IB * IB;
        return ctx->cplx ? launch_zpotrf_leaf(ctx, s, Ajj, lda, n, w, dInfo, infoBase + j0)
                         : launch_potrf_leaf(ctx, s, Ajj, lda, n, w, dInfo, infoBase + j0);
    }
    const int n1 = split_of(n), n2 = n - n1;
    int rc = potrf_rec(ctx, s, A, lda, winv, j0, n1, dInfo, infoBase);
    if (rc) return rc;
    // A21 <- A21 * L11^-T
    double* A21 = A + ez * ((j0 + n1) + (size_t)j0 * lda);
    rc = trsm_rec(ctx, s, Ajj, lda, winv + ez * (size_t)(j0 / IB) * IB * IB, 0, n1, A21, lda, n2, 1, 0, n2);
    if (rc) return rc;
    // A22 <- A22 - A21 A21^T (lower)
    double* A22 = A + ez * ((j0 + n1) + (size_t)(j0 + n1) * lda);
    rc = gemm_single(ctx, s, n2, n2, n1, -1.0, A21, lda, false, A21, lda, false, 1.0, A22, lda, 1);
    if (rc) return rc;
    return potrf_rec(ctx, s, A, lda, winv, j0 + n1, n2, dInfo, infoBase);
}

// In-place lower Cholesky of one n x n tile; leaves the inverses of the diagonal IB-blocks
// in winv for the panel solve that follows.
int potrf_tile_lower(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                     int infoBase) {
    if (n <= 0) return 0;
    return potrf_rec(ctx, s, A, lda, winv, 0, n, dInfo, infoBase);
}

// Same result through the fused left-looking kernel: one launch per 64-wide block column plus
// one write-back launch.  winv points at a slot of ensure_winv(): [inverses | factored blocks].
int potrf_tile_fused(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                     int infoBase) {
    if (n <= 0) return 0;
    static int mode = -1;  // DPB_POTRF_TILE=rec: recursive multi-launch variant; =col: one launch per block column
    if (mode < 0) {
        const char* e = getenv("DPB_POTRF_TILE");
        mode = (e && e[0] == 'r') ? 1 : ((e && e[0] == 'c') ? 2 : 0);
    }
    // the fused kernels are real-only; complex tiles take the recursive variant
    if (mode == 1 || ctx->cplx) return potrf_rec(ctx, s, A, lda, winv, 0, n, dInfo, infoBase);
    if (mode == 0) return launch_potrf_tile(ctx, s, A, lda, n, winv, dInfo, infoBase);  // one launch for the whole tile
    double* wdiag = winv + ctx->wInvSlot / 2;
    for (int j0 = 0; j0 < n; j0 += IB) {
        int rc = launch_potrf_colblock(ctx, s, A, lda, n, j0, winv + (size_t)(j0 / IB) * IB * IB,
                                       wdiag + (size_t)(j0 / IB) * IB * IB, dInfo, infoBase);
        if (rc) return rc;
    }
    return launch_potrf_diag_writeback(ctx, s, A, lda, n, wdiag);
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_b200_dpotrf_tile(parsec_context_t* ctx, void* stream, dplasma_enum_t uplo, int n, double* A, int lda,
                             int* d_info, int info_offset) {
    if (!ctx) return -1;
    if (uplo != dplasmaLower && uplo != dplasmaUpper) return -3;
    if (n < 0) return -4;
    if (lda < (n > 1 ? n : 1)) return -6;
    cudaStream_t s = (cudaStream_t)stream;
    int rc = ensure_winv(ctx, n);
    if (rc) return rc;
    int* info = d_info ? d_info : ctx->dInfo + 2;
    if (uplo == dplasmaLower) return potrf_tile_fused(ctx, s, A, lda, n, ctx->wInv, info, info_offset);
    // Upper: factor the transposed tile (U = L^T) in a scratch tile
    double* tmp = nullptr;
    DPB_CUDA(cudaMallocAsync(&tmp, (size_t)n * n * sizeof(double), s));
    rc = launch_transpose(ctx, s, A, lda, tmp, n, n, n);
    if (!rc) rc = potrf_tile_fused(ctx, s, tmp, n, n, ctx->wInv, info, info_offset);
    // only the upper triangle of A is ours to write: transpose the lower factor back into it
    if (!rc) rc = launch_transpose_upper_into(ctx, s, tmp, n, A, lda, n);
    cudaFreeAsync(tmp, s);
    return rc;
}

int dplasma_b200_dsyrk_tile(parsec_context_t* ctx, void* stream, dplasma_enum_t uplo, dplasma_enum_t trans, int n,
                            int k, double alpha, const double* A, int lda, double beta, double* C, int ldc) {
    if (!ctx) return -1;
    if (uplo != dplasmaLower && uplo != dplasmaUpper) return -3;
    cudaStream_t s = (cudaStream_t)stream;
    const bool aK = trans != dplasmaNoTrans;  // C = alpha A^T A + beta C : op(A)(i,kk) = A[kk + i*lda]
    if (alpha == 0.0 || k == 0)  // BLAS: A is not referenced, C = beta C on the uplo triangle
        return launch_scale_tile(ctx, s, C, ldc, n, n, beta, uplo == dplasmaLower ? 1 : 2);
    ctx->gemmPreload = 0;  // library-call semantics: product first, then beta*C
    int rc = gemm_single(ctx, s, n, n, k, alpha, A, lda, aK, A, lda, aK, beta, C, ldc, uplo == dplasmaLower ? 1 : 2);
    ctx->gemmPreload = 1;
    return rc;
}

int dplasma_b200_dgemm_tile(parsec_context_t* ctx, void* stream, dplasma_enum_t transA, dplasma_enum_t transB,
                            int m, int n, int k, double alpha, const double* A, int lda, const double* B, int ldb,
                            double beta, double* C, int ldc) {
    if (!ctx) return -1;
    cudaStream_t s = (cudaStream_t)stream;
    // engine form: C[i][j] = sum_kk A'[i][kk] B'[j][kk].  A' = op(A): NoTrans -> M-major, Trans -> K-major.
    // B'[j][kk] = op(B)[kk][j]: NoTrans -> B[kk + j*ldb] (K-major), Trans -> B[j + kk*ldb] (M-major).
    const bool aK = transA != dplasmaNoTrans;
    const bool bK = transB == dplasmaNoTrans;
    if (alpha == 0.0 || k == 0) return launch_scale_tile(ctx, s, C, ldc, m, n, beta, 0);  // BLAS: C = beta C
    ctx->gemmPreload = getenv("DPB_GEMM_PRELOAD") ? 1 : 0;  // library-call semantics unless overridden (profiling)
    int rc = gemm_single(ctx, s, m, n, k, alpha, A, lda, aK, B, ldb, bK, beta, C, ldc, 0);
    ctx->gemmPreload = 1;
    return rc;
}

}  // extern "C"
