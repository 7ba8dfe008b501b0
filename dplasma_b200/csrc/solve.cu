// Triangular solves and the posv / potrs drivers: src/ztrsm_wrapper.c:263-310 (Left-side
// DAGs ztrsm_LLN.jdf / ztrsm_LLT.jdf / ztrsm_LUN.jdf / ztrsm_LUT.jdf), src/zpotrs_wrapper.c:58-98,
// src/zposv_wrapper.c:64-114.
//
// The reference solves each RHS tile with CORE_ztrsm on the CPU only (ztrsm_LLN.jdf:40-57 has
// no CUDA body) and updates with legacy cublasZgemm (:111-116).  Here both run on the DMMA
// engine: the diagonal tile's IB x IB diagonal blocks are inverted once (trtri leaf) and the
// tile solve becomes a recursion of products.
#include <cstdio>
#include <algorithm>
#include "internal.h"

namespace dpb {

static inline int split_of(int n) {
    int blocks = (n + IB - 1) / IB;
    return (blocks / 2) * IB;
}

// invert the diagonal IB-blocks of the n x n triangular tile T into winv (always stored as the
// inverse of a lower block, see trtri_leaf_kernel)
static int invert_diag_blocks(parsec_context_s* ctx, cudaStream_t s, int uplo, int diag, int n, const double* T,
                              int ldt, double* winv) {
    return launch_trtri_blocks(ctx, s, T, ldt, n, uplo, diag == dplasmaUnit, winv);
}

int trsm_left_rec_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* winv, double* B, int ldb);

// B(r0:r0+n, :) <- alpha * op(T(r0:r0+n, r0:r0+n))^-1 * B(r0:r0+n, :)      (B is m x ncols, Left side)
// lowerEff: op(T) is lower triangular  (Lower/NoTrans or Upper/Trans)
static int trsm_left_rec(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int r0, int n, int ncols,
                         double alpha, const double* T, int ldt, const double* winv, double* B, int ldb) {
    const bool lowerEff = (uplo == dplasmaLower) != trans;
    if (n <= IB) {
        // X = alpha * inv(op(Tjj)) * B.  winv holds W = inv(lower form of the block):
        //   Lower/NoTrans: inv = W     Lower/Trans: inv = W^T
        //   Upper/NoTrans: inv = W^T   Upper/Trans: inv = W      (W = inv(U^T))
        const bool useWT = (uplo == dplasmaLower) == trans;
        const double* W = winv + (size_t)(r0 / IB) * IB * IB;
        double* Br = B + r0;
        // in place: M = n <= 64 rows -> a single CTA row block per column block, which loads all
        // of B(:, cols) before it writes
        return gemm_single(ctx, s, n, ncols, n, alpha, W, IB, useWT, Br, ldb, true, 0.0, Br, ldb, 0);
    }
    const int n1 = split_of(n), n2 = n - n1;
    int rc;
    if (lowerEff) {
        // forward: X1 = inv(T11) B1 ; B2 = alpha B2 - opT21 X1 ; X2 = inv(T22) B2
        rc = trsm_left_rec(ctx, s, uplo, trans, r0, n1, ncols, alpha, T, ldt, winv, B, ldb);
        if (rc) return rc;
        // op(T)21 (n2 x n1): Lower/NoTrans -> T(r0+n1.., r0..) M-major ; Upper/Trans -> T(r0.., r0+n1..)^T K-major
        const double* T21 = trans ? T + r0 + (size_t)(r0 + n1) * ldt : T + (r0 + n1) + (size_t)r0 * ldt;
        rc = gemm_single(ctx, s, n2, ncols, n1, -1.0, T21, ldt, trans, B + r0, ldb, true, alpha, B + r0 + n1, ldb, 0);
        if (rc) return rc;
        return trsm_left_rec(ctx, s, uplo, trans, r0 + n1, n2, ncols, 1.0, T, ldt, winv, B, ldb);
    }
    // backward: X2 = inv(T22) B2 ; B1 = alpha B1 - opT12 X2 ; X1 = inv(T11) B1
    rc = trsm_left_rec(ctx, s, uplo, trans, r0 + n1, n2, ncols, alpha, T, ldt, winv, B, ldb);
    if (rc) return rc;
    // op(T)12 (n1 x n2): Upper/NoTrans -> T(r0.., r0+n1..) M-major ; Lower/Trans -> T(r0+n1.., r0..)^T K-major
    const double* T12 = trans ? T + (r0 + n1) + (size_t)r0 * ldt : T + r0 + (size_t)(r0 + n1) * ldt;
    rc = gemm_single(ctx, s, n1, ncols, n2, -1.0, T12, ldt, trans, B + r0 + n1, ldb, true, alpha, B + r0, ldb, 0);
    if (rc) return rc;
    return trsm_left_rec(ctx, s, uplo, trans, r0, n1, ncols, 1.0, T, ldt, winv, B, ldb);
}

// One tile solve, Left side: B <- alpha op(T)^-1 B, T m x m triangular, B m x n.
int trsm_left(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, int m, int n, double alpha,
              const double* T, int ldt, double* B, int ldb, double* winv) {
    if (m <= 0 || n <= 0) return 0;
    int rc = invert_diag_blocks(ctx, s, uplo, diag, m, T, ldt, winv);
    if (rc) return rc;
    return trsm_left_rec(ctx, s, uplo, trans != dplasmaNoTrans, 0, m, n, alpha, T, ldt, winv, B, ldb);
}

int trsm_left_rec_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* winv, double* B, int ldb) {
    return trsm_left_rec(ctx, s, uplo, trans, 0, n, ncols, alpha, T, ldt, winv, B, ldb);
}

// dplasma_ztrsm, Left side, on staged matrices (single process).
static int trsm_left_tiled(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, double alpha,
                           const DevMatrix& A, const DevMatrix& B) {
    int rc = ensure_winv(ctx, A.nb);
    if (rc) return rc;
    const int mt = B.mt, ntr = B.nt;
    const bool tr = trans != dplasmaNoTrans;
    const bool forward = (uplo == dplasmaLower) != tr;
    for (int kk = 0; kk < mt; kk++) {
        const int k = forward ? kk : mt - 1 - kk;
        const int tk = B.rowsOfTile(k);
        const double lalpha = (kk == 0) ? alpha : 1.0;  // src/ztrsm_LLN.jdf:44
        double* winv = ctx->wInv + (size_t)(kk & 1) * ctx->wInvSlot;
        const double* Akk = A.tile(k, k);
        rc = invert_diag_blocks(ctx, s, uplo, diag, tk, Akk, A.mb, winv);
        if (rc) return rc;
        for (int n = 0; n < ntr; n++) {
            rc = trsm_left_rec(ctx, s, uplo, tr, 0, tk, B.colsOfTile(n), lalpha, Akk, A.mb, winv, B.tile(k, n), B.mb);
            if (rc) return rc;
        }
        // B(m,:) = lalpha B(m,:) - op(A)(m,k) B(k,:) for all remaining m (src/ztrsm_LLN.jdf:131-135,
        // ztrsm_LLT.jdf:132-137): one launch per RHS tile column, batched over the tile rows m
        const int lo = forward ? k + 1 : 0, hi = forward ? mt : k;
        if (hi > lo) {
            const bool stored_mk = (uplo == dplasmaLower) ? forward : !forward;  // is A(m,k) in the stored triangle?
            const double* A0 = stored_mk ? A.tile(lo, k) : A.tile(k, lo);
            const long long sA = stored_mk ? A.tileElems : (long long)A.lmt * A.tileElems;
            for (int n = 0; n < ntr; n++) {
                rc = gemm_single(ctx, s, B.mb, B.colsOfTile(n), tk, -1.0, A0, A.mb, !stored_mk, B.tile(k, n), B.mb, true,
                                 lalpha, B.tile(lo, n), B.mb, 0, hi - lo, sA, 0, B.tileElems, B.rowsOfTile(hi - 1));
                if (rc) return rc;
            }
        }
    }
    return 0;
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_dtrsm(parsec_context_t* parsec, dplasma_enum_t side, dplasma_enum_t uplo, dplasma_enum_t trans,
                  dplasma_enum_t diag, double alpha, const parsec_tiled_matrix_t* A, parsec_tiled_matrix_t* B) {
    // argument checks and codes: src/ztrsm_wrapper.c:274-296
    if (side != dplasmaLeft && side != dplasmaRight) { fprintf(stderr, "dplasma_dtrsm: illegal value of side\n"); return -1; }
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dtrsm: illegal value of uplo\n"); return -2; }
    if (trans != dplasmaNoTrans && trans != dplasmaTrans && trans != dplasmaConjTrans) { fprintf(stderr, "dplasma_dtrsm: illegal value of trans\n"); return -3; }
    if (diag != dplasmaUnit && diag != dplasmaNonUnit) { fprintf(stderr, "dplasma_dtrsm: illegal value of diag\n"); return -4; }
    if (!parsec) return -100;
    if (side == dplasmaRight) {
        fprintf(stderr, "dplasma_dtrsm: side Right is outside the Cholesky path and not built here\n");
        return -1;
    }
    cudaSetDevice(parsec->device);
    DevMatrix dA, dB;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, B, &dB, true, dplasmaUpperLower);
    if (!rc && alpha == 0.0) {
        // BLAS semantics: alpha == 0 -> B = 0, A is not referenced
        rc = launch_fill(parsec, parsec->sPanel, dB.base, (long long)dB.lmt * dB.lnt * dB.tileElems, 0.0);
    } else
    if (!rc) rc = (A->super.nodes > 1) ? trsm_left_multi(parsec, parsec->sPanel, uplo, trans, diag, alpha, dA, dB)
                                       : trsm_left_tiled(parsec, parsec->sPanel, uplo, trans, diag, alpha, dA, dB);
    if (!rc) rc = stage_out(parsec, B, &dB, dplasmaUpperLower);
    cudaError_t e = cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    stage_free(&dB);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    return e == cudaSuccess ? 0 : -1000 - (int)e;
}

int dplasma_dpotrs(parsec_context_t* parsec, dplasma_enum_t uplo, const parsec_tiled_matrix_t* A,
                   parsec_tiled_matrix_t* B) {
    if (uplo != dplasmaUpper && uplo != dplasmaLower) {
        fprintf(stderr, "dplasma_dpotrs: illegal value of uplo\n");  // src/zpotrs_wrapper.c:64-67
        return -1;
    }
    if (!parsec) return -100;
    cudaSetDevice(parsec->device);
    DevMatrix dA, dB;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, B, &dB, true, dplasmaUpperLower);
    cudaStream_t s = parsec->sPanel;
    const bool multi = A->super.nodes > 1;
    auto solve = [&](int tr) {
        return multi ? trsm_left_multi(parsec, s, uplo, tr, dplasmaNonUnit, 1.0, dA, dB)
                     : trsm_left_tiled(parsec, s, uplo, tr, dplasmaNonUnit, 1.0, dA, dB);
    };
    if (!rc) {
        // src/zpotrs_wrapper.c:78-98: Upper: U^T then U ; Lower: L then L^T
        rc = solve(uplo == dplasmaUpper ? dplasmaTrans : dplasmaNoTrans);
        if (!rc) rc = solve(uplo == dplasmaUpper ? dplasmaNoTrans : dplasmaTrans);
    }
    if (!rc) rc = stage_out(parsec, B, &dB, dplasmaUpperLower);
    cudaError_t e = cudaStreamSynchronize(s);
    stage_free(&dA);
    stage_free(&dB);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    return e == cudaSuccess ? 0 : -1000 - (int)e;
}

int dplasma_dposv(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A, parsec_tiled_matrix_t* B) {
    if (uplo != dplasmaUpper && uplo != dplasmaLower) {
        fprintf(stderr, "dplasma_dposv: illegal value of uplo\n");  // src/zposv_wrapper.c:73-76
        return -1;
    }
    int info = dplasma_dpotrf(parsec, uplo, A);                // :102
    if (info == 0) info = dplasma_dpotrs(parsec, uplo, A, B);  // :103-110
    return info;
}

int dplasma_b200_dtrsm_tile(parsec_context_t* ctx, void* stream, dplasma_enum_t side, dplasma_enum_t uplo,
                            dplasma_enum_t trans, dplasma_enum_t diag, int m, int n, double alpha, const double* T,
                            int ldt, double* C, int ldc) {
    if (!ctx) return -1;
    cudaStream_t s = (cudaStream_t)stream;
    const int na = (side == dplasmaLeft) ? m : n;
    int rc = ensure_winv(ctx, na);
    if (rc) return rc;
    if (side == dplasmaLeft) return trsm_left(ctx, s, uplo, trans, diag, m, n, alpha, T, ldt, C, ldc, ctx->wInv);
    // Right side, the form the factorization uses: C <- alpha C op(T)^-1  (src/zpotrf_L.jdf:236-240)
    if (uplo == dplasmaLower && trans != dplasmaNoTrans && alpha == 1.0) {
        rc = 0;
        for (int j0 = 0; j0 < n && !rc; j0 += IB)
            rc = launch_trtri_leaf(ctx, s, T + j0 + (size_t)j0 * ldt, ldt, std::min(IB, n - j0), uplo,
                                   diag == dplasmaUnit, ctx->wInv + (size_t)(j0 / IB) * IB * IB);
        if (rc) return rc;
        return trsm_panel_lower(ctx, s, T, ldt, n, ctx->wInv, C, ldc, m, 1, 0, m);
    }
    fprintf(stderr, "dplasma_b200_dtrsm_tile: this side/uplo/trans combination is not on the Cholesky path\n");
    return -2;
}

}  // extern "C"
