// Complex solves: dplasma_zpotrs / dplasma_zposv (src/zpotrs_wrapper.c:58-98, src/zposv_wrapper.c:64-114;
// DAGs src/ztrsm_LLN.jdf and ztrsm_LLT.jdf with ConjTrans) and check_zaxmb (src/dplasma_zcheck.c:180-225).
//
// The complex DMMA engine offers one product form, C -= A B^H on column-major operands, which is
// exactly what the factorization needs.  The two triangular solves are brought to that form by
// working on transposed right-hand sides:
//   L Y' = B      <=>  Y L^H = B^H          with Y = Y'^H   : Right / Lower / ConjTrans, the
//                                                             panel solve of the factorization;
//   L^H X = Y'    <=>  P conj(L) = Y'^T     with P = X^T    : products against plainly transposed
//                                                             tiles of L and of its inverted blocks.
// Between the passes the work array is conjugated in place ((Y'^H) -> (Y'^T)).
// Single process, dplasmaLower.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include "internal.h"

namespace dpb {

static inline int zsplit_of(int n) {
    int blocks = (n + IB - 1) / IB;
    return (blocks / 2) * IB;
}

// P(:, c0:c0+n) <- Q(:, c0:c0+n) conj(L(c0.., c0..))^-1 for the batch of row blocks in Q.
// Lt = plain transpose of the triangular tile, wT = plain transposes of its inverted 64x64 blocks.
static int ztrsm_conj_rec(parsec_context_s* ctx, cudaStream_t s, const double* Lt, int ldl, const double* wT, int c0,
                          int n, double* Q, int ldq, int M, int batch, long long strideQ, int lastBatchM) {
    if (n <= IB) {
        double* Qc = Q + 2 * (size_t)c0 * ldq;
        return gemm_single(ctx, s, M, n, n, 1.0, Qc, ldq, false, wT + 2 * (size_t)(c0 / IB) * IB * IB, IB, false, 0.0, Qc,
                           ldq, 0, batch, strideQ, 0, strideQ, lastBatchM);
    }
    const int n1 = zsplit_of(n), n2 = n - n1;
    int rc = ztrsm_conj_rec(ctx, s, Lt, ldl, wT, c0 + n1, n2, Q, ldq, M, batch, strideQ, lastBatchM);
    if (rc) return rc;
    // Q1 -= P2 conj(L21):  B[c][t] = L21[t][c] = Lt[c0+c][c0+n1+t]
    rc = gemm_single(ctx, s, M, n1, n2, -1.0, Q + 2 * (size_t)(c0 + n1) * ldq, ldq, false,
                     Lt + 2 * (c0 + (size_t)(c0 + n1) * ldl), ldl, false, 1.0, Q + 2 * (size_t)c0 * ldq, ldq, 0, batch,
                     strideQ, 0, strideQ, lastBatchM);
    if (rc) return rc;
    return ztrsm_conj_rec(ctx, s, Lt, ldl, wT, c0, n1, Q, ldq, M, batch, strideQ, lastBatchM);
}

// X <- L^-H L^-1 B on staged matrices
static int zpotrs_lower(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, const DevMatrix& B) {
    if (A.mb != A.nb || B.mb != A.mb || A.P * A.Q != 1) return -102;
    const int mt = B.mt, ntr = B.nt, mb = A.mb, nbr = B.nb;
    const size_t tileY = (size_t)nbr * mb * 2;  // doubles: a nbr x mb complex tile, ld nbr
    const size_t nblk = (size_t)((mb + IB - 1) / IB);
    double *Y = nullptr, *W = nullptr, *Lt = nullptr;
    const size_t yElems = std::max<size_t>(tileY * ntr * mt, 2), wElems = 2 * nblk * IB * IB * 2;
    const size_t ltElems = (size_t)A.tileElems * (size_t)(mt + 1);
    if (cudaMalloc(&Y, yElems * 8) != cudaSuccess || cudaMalloc(&W, wElems * 8) != cudaSuccess ||
        cudaMalloc(&Lt, ltElems * 8) != cudaSuccess) {
        if (Y) cudaFree(Y);
        if (W) cudaFree(W);
        return -200;
    }
    double* WT = W + nblk * IB * IB * 2;
    auto Yt = [&](int c, int m) { return Y + ((size_t)c + (size_t)m * ntr) * tileY; };
    const int lastC = B.colsOfTile(ntr - 1);
    ctx->cplx = 1;
    int rc = launch_fill(ctx, s, Y, (long long)yElems, 0.0);
    // Y(c,m) = B(m,c)^H
    for (int m = 0; m < mt && !rc; m++)
        for (int c = 0; c < ntr && !rc; c++)
            rc = launch_ztranspose(ctx, s, B.tile(m, c), mb, Yt(c, m), nbr, B.rowsOfTile(m), B.colsOfTile(c), 1);
    // ---- forward: Y L^H = B^H, tile columns left to right (ztrsm_LLN.jdf in transposed form) ----
    for (int k = 0; k < mt && !rc; k++) {
        const int tk = B.rowsOfTile(k);
        rc = launch_ztrtri_blocks(ctx, s, A.tile(k, k), mb, tk, W, WT);
        if (!rc) rc = trsm_rec_entry(ctx, s, A.tile(k, k), mb, W, tk, Yt(0, k), nbr, nbr, ntr, (long long)tileY, lastC);
        for (int m = k + 1; m < mt && !rc; m++)
            rc = gemm_single(ctx, s, nbr, B.rowsOfTile(m), tk, -1.0, Yt(0, k), nbr, false, A.tile(m, k), mb, false, 1.0,
                             Yt(0, m), nbr, 0, ntr, (long long)tileY, 0, (long long)tileY, lastC);
    }
    // (L^-1 B)^H -> (L^-1 B)^T
    if (!rc) rc = launch_zconj(ctx, s, Y, (long long)(yElems / 2));
    // ---- backward: P conj(L) = Q, tile columns right to left (ztrsm_LLT.jdf, ConjTrans, transposed form) ----
    double* LtDiag = Lt + (size_t)mt * A.tileElems;
    for (int k = mt - 1; k >= 0 && !rc; k--) {
        const int tk = B.rowsOfTile(k);
        rc = launch_ztrtri_blocks(ctx, s, A.tile(k, k), mb, tk, W, WT);
        if (!rc) rc = launch_fill(ctx, s, LtDiag, A.tileElems, 0.0);
        if (!rc) rc = launch_ztranspose(ctx, s, A.tile(k, k), mb, LtDiag, mb, tk, tk, 0);
        if (!rc) rc = ztrsm_conj_rec(ctx, s, LtDiag, mb, WT, 0, tk, Yt(0, k), nbr, nbr, ntr, (long long)tileY, lastC);
        for (int m = 0; m < k && !rc; m++) {
            // Q(:,m) -= P(:,k) conj(L(k,m)):  B[c][t] = L(k,m)[t][c] -> plain transpose of the tile
            double* Ltm = Lt + (size_t)m * A.tileElems;
            rc = launch_ztranspose(ctx, s, A.tile(k, m), mb, Ltm, mb, tk, mb, 0);
            if (!rc)
                rc = gemm_single(ctx, s, nbr, mb, tk, -1.0, Yt(0, k), nbr, false, Ltm, mb, false, 1.0, Yt(0, m), nbr, 0, ntr,
                                 (long long)tileY, 0, (long long)tileY, lastC);
        }
    }
    // X(m,c) = P(c,m)^T
    for (int m = 0; m < mt && !rc; m++)
        for (int c = 0; c < ntr && !rc; c++)
            rc = launch_ztranspose(ctx, s, Yt(c, m), nbr, B.tile(m, c), mb, B.colsOfTile(c), B.rowsOfTile(m), 0);
    ctx->cplx = 0;
    cudaStreamSynchronize(s);
    cudaFree(Y);
    cudaFree(W);
    cudaFree(Lt);
    return rc;
}

// b <- b - A x, A hermitian stored in its lower tiles (dplasma_zhemm(Left, Lower, -1, A, x, 1, b))
static int zhemm_residual_lower(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, const DevMatrix& X,
                                const DevMatrix& B) {
    const int mt = A.mt, ntr = B.nt, mb = A.mb, nbr = B.nb;
    const size_t tileX = (size_t)nbr * mb * 2;
    double *Xh = nullptr, *tmp = nullptr;
    const size_t xElems = std::max<size_t>(tileX * ntr * mt, 2);
    if (cudaMalloc(&Xh, xElems * 8) != cudaSuccess || cudaMalloc(&tmp, (size_t)A.tileElems * (mt + 1) * 8) != cudaSuccess) {
        if (Xh) cudaFree(Xh);
        return -200;
    }
    auto Xt = [&](int c, int n) { return Xh + ((size_t)c + (size_t)n * ntr) * tileX; };
    ctx->cplx = 1;
    int rc = launch_fill(ctx, s, Xh, (long long)xElems, 0.0);
    for (int n = 0; n < mt && !rc; n++)
        for (int c = 0; c < ntr && !rc; c++)
            rc = launch_ztranspose(ctx, s, X.tile(n, c), mb, Xt(c, n), nbr, X.rowsOfTile(n), X.colsOfTile(c), 1);
    double* diagCopy = tmp + (size_t)mt * A.tileElems;
    for (int n = 0; n < mt && !rc; n++) {
        const int tn = A.colsOfTile(n);
        rc = launch_copy_tril(ctx, s, A.tile(n, n), mb, diagCopy, mb, tn, dplasmaLower, 1);  // hermitian completion
        // tiles above the diagonal tile: A(m,n) = A(n,m)^H
        for (int m = 0; m < n && !rc; m++)
            rc = launch_ztranspose(ctx, s, A.tile(n, m), mb, tmp + (size_t)m * A.tileElems, mb, tn, mb, 1);
        for (int c = 0; c < ntr && !rc; c++) {
            const int nc = B.colsOfTile(c);
            rc = gemm_single(ctx, s, tn, nc, tn, -1.0, diagCopy, mb, false, Xt(c, n), nbr, false, 1.0, B.tile(n, c), mb, 0);
            const int below = mt - n - 1;
            if (!rc && below > 0)
                rc = gemm_single(ctx, s, mb, nc, tn, -1.0, A.tile(n + 1, n), mb, false, Xt(c, n), nbr, false, 1.0,
                                 B.tile(n + 1, c), mb, 0, below, A.tileElems, 0, B.tileElems, A.rowsOfTile(mt - 1));
            if (!rc && n > 0)
                rc = gemm_single(ctx, s, mb, nc, tn, -1.0, tmp, mb, false, Xt(c, n), nbr, false, 1.0, B.tile(0, c), mb, 0, n,
                                 A.tileElems, 0, B.tileElems, mb);
        }
    }
    ctx->cplx = 0;
    cudaStreamSynchronize(s);
    cudaFree(Xh);
    cudaFree(tmp);
    return rc;
}

int inf_norm_public(parsec_context_s* ctx, const DevMatrix& A, bool sym, int uplo, double* out);  // check.cu

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_zpotrs(parsec_context_t* parsec, dplasma_enum_t uplo, const parsec_tiled_matrix_t* A,
                   parsec_tiled_matrix_t* B) {
    if (uplo != dplasmaUpper && uplo != dplasmaLower) {
        fprintf(stderr, "dplasma_zpotrs: illegal value of uplo\n");  // src/zpotrs_wrapper.c:64-67
        return -1;
    }
    if (!parsec) return -100;
    if (A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE || B->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return -3;
    if (uplo != dplasmaLower || A->super.nodes > 1) {
        fprintf(stderr, "dplasma_zpotrs: only dplasmaLower on one process is built\n");
        return -2;
    }
    cudaSetDevice(parsec->device);
    DevMatrix dA, dB;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, B, &dB, true, dplasmaUpperLower);
    if (!rc) rc = zpotrs_lower(parsec, parsec->sPanel, dA, dB);
    if (!rc) rc = stage_out(parsec, B, &dB, dplasmaUpperLower);
    cudaError_t e = cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    stage_free(&dB);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    return e == cudaSuccess ? 0 : -1000 - (int)e;
}

int dplasma_zpotrf(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A);

int dplasma_zposv(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A, parsec_tiled_matrix_t* B) {
    if (uplo != dplasmaUpper && uplo != dplasmaLower) {
        fprintf(stderr, "dplasma_zposv: illegal value of uplo\n");  // src/zposv_wrapper.c:73-76
        return -1;
    }
    int info = dplasma_zpotrf(parsec, uplo, A);
    if (info == 0) info = dplasma_zpotrs(parsec, uplo, A, B);
    return info;
}

double dplasma_zlange(parsec_context_t* parsec, dplasma_enum_t ntype, const parsec_tiled_matrix_t* A) {
    return dplasma_dlange(parsec, ntype, A);  // element type comes from the descriptor
}

int check_zaxmb(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                parsec_tiled_matrix_t* b, parsec_tiled_matrix_t* x) {
    if (!parsec) return 1;
    if (uplo != dplasmaLower || A->super.nodes > 1 || A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) {
        fprintf(stderr, "check_zaxmb: only dplasmaLower on one process is built\n");
        return 1;
    }
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    DevMatrix dA, dB, dX;
    double Anorm = 0, Bnorm = 0, Xnorm = 0, Rnorm = 0;
    const int N = b->m;
    const double eps = 0x1.0p-53;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, b, &dB, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, x, &dX, true, dplasmaUpperLower);
    if (!rc) rc = inf_norm_public(parsec, dA, true, uplo, &Anorm);
    if (!rc) rc = inf_norm_public(parsec, dB, false, 0, &Bnorm);
    if (!rc) rc = inf_norm_public(parsec, dX, false, 0, &Xnorm);
    if (!rc) rc = zhemm_residual_lower(parsec, s, dA, dX, dB);
    if (!rc) rc = inf_norm_public(parsec, dB, false, 0, &Rnorm);
    if (!rc) rc = stage_out(parsec, b, &dB, dplasmaUpperLower);
    cudaStreamSynchronize(s);
    stage_free(&dA);
    stage_free(&dB);
    stage_free(&dX);
    if (rc) {
        fprintf(stderr, "check_zaxmb: internal error %d\n", rc);
        return 1;
    }
    const double result = Rnorm / ((Anorm * Xnorm + Bnorm) * N * eps);
    parsec->lastCheck = result;
    if (loud > 2) {
        printf("============\n");
        printf("Checking the Residual of the solution \n");
        if (loud > 3)
            printf("-- ||A||_oo = %e, ||X||_oo = %e, ||B||_oo= %e, ||A X - B||_oo = %e\n", Anorm, Xnorm, Bnorm, Rnorm);
        printf("-- ||Ax-B||_oo/((||A||_oo||x||_oo+||B||_oo).N.eps) = %e \n", result);
    }
    int bad = (std::isnan(Xnorm) || std::isinf(Xnorm) || std::isnan(result) || std::isinf(result) || result > 60.0);
    if (loud) printf(bad ? "-- Solution is suspicious ! \n" : "-- Solution is CORRECT ! \n");
    return bad;
}

}  // extern "C"
