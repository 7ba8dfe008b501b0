// Generators, norms and the reference's two residual checks, computed on the GPU:
//   dplasma_dplgsy / dplasma_dplrnt   src/zplgsy_wrapper.c, src/zplrnt_wrapper.c
//   dplasma_dlansy / dplasma_dlange   (InfNorm)  as used by src/dplasma_zcheck.c:101-103,194-196
//   check_dpotrf                      src/dplasma_zcheck.c:64-134
//   check_daxmb                       src/dplasma_zcheck.c:180-225
#include <cmath>
#include <cstdio>
#include <algorithm>
#include "internal.h"

namespace dpb {

// max row sum; sym: symmetric matrix stored in `uplo`
static int inf_norm(parsec_context_s* ctx, const DevMatrix& A, bool sym, int uplo, double* out) {
    cudaStream_t s = ctx->sPanel;
    const int savedCplx = ctx->cplx;
    ctx->cplx = (A.esz == 2);
    struct Restore { parsec_context_s* c; int v; ~Restore() { c->cplx = v; } } restore{ctx, savedCplx};
    int rc = ensure_wnorm(ctx, (size_t)A.M + 8);
    if (rc) return rc;
    DPB_CUDA(cudaMemsetAsync(ctx->wNorm, 0, ((size_t)A.M + 8) * sizeof(double), s));
    rc = sym ? launch_rowsum_sym(ctx, s, A, uplo, ctx->wNorm) : launch_rowsum_ge(ctx, s, A, ctx->wNorm);
    if (rc) return rc;
    if (A.P * A.Q > 1) {
        rc = sum_allreduce(ctx, ctx->wNorm, (size_t)A.M, s);
        if (rc) return rc;
    }
    rc = launch_absmax(ctx, s, ctx->wNorm, A.M, ctx->wNorm + A.M);
    if (rc) return rc;
    DPB_CUDA(cudaMemcpyAsync(out, ctx->wNorm + A.M, sizeof(double), cudaMemcpyDeviceToHost, s));
    DPB_CUDA(cudaStreamSynchronize(s));
    return 0;
}

// B <- B - sym(A) X on one GPU, A symmetric and stored in `uplo` (dplasma_zhemm(Left, uplo, -1, A, X, 1, B),
// src/dplasma_zcheck.c:199,289): per tile column n of A, one batched launch for the stored tiles (m,n),
// one for the mirrored tiles (n,m)^T and one for the symmetrised diagonal tile, per tile column of X.
// diagCopy: one tile of scratch.
int symm_residual_1gpu(parsec_context_s* parsec, cudaStream_t s, int uplo, const DevMatrix& dA, const DevMatrix& dX,
                       const DevMatrix& dB, double* diagCopy) {
    int rc = 0;
    for (int n = 0; !rc && n < dA.nt; n++) {
        const int tn = dA.colsOfTile(n);
        rc = launch_copy_tril(parsec, s, dA.tile(n, n), dA.mb, diagCopy, dA.mb, tn, uplo, 1);
        for (int c = 0; !rc && c < dB.nt; c++) {
            const int nc = dB.colsOfTile(c);
            const double* Xn = dX.tile(n, c);
            // diagonal tile
            rc = gemm_single(parsec, s, tn, nc, tn, -1.0, diagCopy, dA.mb, false, Xn, dX.mb, true, 1.0, dB.tile(n, c), dB.mb, 0);
            // rows below n: A(m,n) stored (Lower) or A(n,m)^T (Upper)
            const int below = dA.mt - n - 1;
            if (!rc && below > 0) {
                const bool st = (uplo == dplasmaLower);
                rc = gemm_single(parsec, s, dA.mb, nc, tn, -1.0, st ? dA.tile(n + 1, n) : dA.tile(n, n + 1), dA.mb, !st, Xn,
                                 dX.mb, true, 1.0, dB.tile(n + 1, c), dB.mb, 0, below,
                                 st ? dA.tileElems : (long long)dA.lmt * dA.tileElems, 0, dB.tileElems,
                                 dA.rowsOfTile(dA.mt - 1));
            }
            // rows above n: A(m,n) stored (Upper) or A(n,m)^T (Lower)
            if (!rc && n > 0) {
                const bool st = (uplo == dplasmaUpper);
                rc = gemm_single(parsec, s, dA.mb, nc, tn, -1.0, st ? dA.tile(0, n) : dA.tile(n, 0), dA.mb, !st, Xn, dX.mb,
                                 true, 1.0, dB.tile(0, c), dB.mb, 0, n, st ? dA.tileElems : (long long)dA.lmt * dA.tileElems,
                                 0, dB.tileElems, dA.mb);
            }
        }
    }
    return rc;
}

int inf_norm_public(parsec_context_s* ctx, const DevMatrix& A, bool sym, int uplo, double* out) {
    return inf_norm(ctx, A, sym, uplo, out);
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_dplgsy(parsec_context_t* parsec, double bump, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                   unsigned long long int seed) {
    if (!parsec) return -1;
    if (uplo != dplasmaLower && uplo != dplasmaUpper && uplo != dplasmaUpperLower) return -3;
    cudaSetDevice(parsec->device);
    DevMatrix dA;
    int rc = stage_in(parsec, A, &dA, false, uplo);
    if (!rc) rc = launch_generate(parsec, parsec->sPanel, dA, 0, bump, uplo, seed);
    if (!rc) rc = stage_out(parsec, A, &dA, uplo);
    cudaError_t e = cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    return e == cudaSuccess ? 0 : -1000 - (int)e;
}

int dplasma_dplrnt(parsec_context_t* parsec, int diagdom, parsec_tiled_matrix_t* A, unsigned long long int seed) {
    if (!parsec) return -1;
    if (diagdom != 0) {
        fprintf(stderr, "dplasma_dplrnt: diagdom != 0 is not used on the Cholesky path\n");
        return -2;
    }
    cudaSetDevice(parsec->device);
    DevMatrix dA;
    int rc = stage_in(parsec, A, &dA, false, dplasmaUpperLower);
    if (!rc) rc = launch_generate(parsec, parsec->sPanel, dA, 1, 0.0, dplasmaUpperLower, seed);
    if (!rc) rc = stage_out(parsec, A, &dA, dplasmaUpperLower);
    cudaError_t e = cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    return e == cudaSuccess ? 0 : -1000 - (int)e;
}

// complex generators: src/zplghe_wrapper.c -> core_zplghe.c, src/zplrnt_wrapper.c -> core_zplrnt.c.
// The descriptor's mtype (PARSEC_MATRIX_COMPLEX_DOUBLE) selects the element type; the kernels are shared.
int dplasma_zplghe(parsec_context_t* parsec, double bump, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                   unsigned long long int seed) {
    if (A && A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return -4;
    return dplasma_dplgsy(parsec, bump, uplo, A, seed);
}
int dplasma_zplrnt(parsec_context_t* parsec, int diagdom, parsec_tiled_matrix_t* A, unsigned long long int seed) {
    if (A && A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return -3;
    return dplasma_dplrnt(parsec, diagdom, A, seed);
}
double dplasma_zlanhe(parsec_context_t* parsec, dplasma_enum_t ntype, dplasma_enum_t uplo,
                      const parsec_tiled_matrix_t* A);
int check_zpotrf(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                 parsec_tiled_matrix_t* A0);

double dplasma_dlansy(parsec_context_t* parsec, dplasma_enum_t ntype, dplasma_enum_t uplo,
                      const parsec_tiled_matrix_t* A) {
    if (!parsec || ntype != dplasmaInfNorm) return -1.0;
    cudaSetDevice(parsec->device);
    DevMatrix dA;
    double r = -1.0;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = inf_norm(parsec, dA, true, uplo, &r);
    cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    return rc ? -1.0 : r;
}

double dplasma_dlange(parsec_context_t* parsec, dplasma_enum_t ntype, const parsec_tiled_matrix_t* A) {
    if (!parsec || ntype != dplasmaInfNorm) return -1.0;
    cudaSetDevice(parsec->device);
    DevMatrix dA;
    double r = -1.0;
    int rc = stage_in(parsec, A, &dA, true, dplasmaUpperLower);
    if (!rc) rc = inf_norm(parsec, dA, false, dplasmaUpperLower, &r);
    cudaStreamSynchronize(parsec->sPanel);
    stage_free(&dA);
    return rc ? -1.0 : r;
}

// R = A0 - L L^T on the stored triangle, by "un-factoring": the same panel-by-panel rank-nb
// updates as the factorization itself, run through the same grouped DMMA kernel.
int check_dpotrf(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                 parsec_tiled_matrix_t* A0) {
    if (!parsec) return 1;
    const bool multi = A->super.nodes > 1;
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    DevMatrix dA, dA0, R;
    double* diagCopy = nullptr;
    double Anorm = 0.0, Rnorm = 0.0, result = NAN;
    bool swapped = false;
    const int N = A->n;
    const double eps = 0x1.0p-53;  // LAPACKE_dlamch_work('e'), src/dplasma_zcheck.c:77
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, A0, &dA0, true, uplo);
    if (!rc) rc = inf_norm(parsec, dA0, true, uplo, &Anorm);
    parsec->cplx = (!rc && dA.esz == 2);
    if (!rc) {
        R = dA0;
        R.owned = true;
        R.base = nullptr;
        const size_t elems = (size_t)R.lmt * R.lnt * R.tileElems;
        if (cudaMalloc(&R.base, elems * sizeof(double)) != cudaSuccess) rc = -200;
        if (!rc && cudaMalloc(&diagCopy, (size_t)R.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
        if (!rc && cudaMemcpyAsync(R.base, dA0.base, elems * sizeof(double), cudaMemcpyDeviceToDevice, s) != cudaSuccess) rc = -201;
    }
    if (!rc && uplo == dplasmaUpper && multi) {
        // grid: both the factor and A0 in lower form on the swapped grid (local transposes, potrf_dist.cu)
        DevMatrix Lt, Rt;
        Lt.base = Rt.base = nullptr;
        rc = ensure_grid_comms(parsec, dA.P, dA.Q, dA.ip, dA.jq);
        if (!rc) rc = transposed_copy(parsec, s, dA, dplasmaUpper, &Lt);
        if (!rc) rc = transposed_copy(parsec, s, R, dplasmaUpper, &Rt);
        if (!rc) {
            stage_free(&dA);
            stage_free(&R);
            dA = Lt;
            R = Rt;
            swapped = true;
        } else {
            if (Lt.base) cudaFree(Lt.base);
            if (Rt.base) cudaFree(Rt.base);
        }
    }
    if (!rc && uplo == dplasmaUpper && !multi) {
        // run the lower algorithm on the transposed grids
        rc = launch_transpose_grid(parsec, s, R);
        if (!rc && !dA.owned) {
            // A is used in place by the caller: work on a private copy
            DevMatrix c = dA;
            const size_t elems = (size_t)c.lmt * c.lnt * c.tileElems;
            c.base = nullptr;
            if (cudaMalloc(&c.base, elems * sizeof(double)) != cudaSuccess) rc = -200;
            if (!rc) cudaMemcpyAsync(c.base, dA.base, elems * sizeof(double), cudaMemcpyDeviceToDevice, s);
            c.owned = true;
            dA = c;
        }
        if (!rc) rc = launch_transpose_grid(parsec, s, dA);
    }
    if (!rc && multi) {
        if (swapped) swap_grid_comms(parsec);
        rc = check_update_multi(parsec, dA, R);
        if (swapped) swap_grid_comms(parsec);
    }
    for (int k = 0; !rc && !multi && k < dA.nt; k++) {
        const int tk = dA.rowsOfTile(k), below = dA.nt - k - 1;
        // D = tril(L(k,k)) with the strict upper part zeroed (dplasma_zlacpy(uplo), src/dplasma_zcheck.c:90)
        rc = launch_copy_tril(parsec, s, dA.tile(k, k), dA.mb, diagCopy, dA.mb, tk, dplasmaLower, 0);
        // R(k,k) -= D D^T (lower)
        if (!rc) rc = gemm_single(parsec, s, tk, tk, tk, -1.0, diagCopy, dA.mb, false, diagCopy, dA.mb, false, 1.0,
                                  R.tile(k, k), R.mb, 1);
        // R(m,k) -= L(m,k) D^T, m > k
        if (!rc && below > 0)
            rc = gemm_single(parsec, s, dA.mb, tk, tk, -1.0, dA.tile(k + 1, k), dA.mb, false, diagCopy, dA.mb, false,
                             1.0, R.tile(k + 1, k), R.mb, 0, below, dA.tileElems, 0, R.tileElems,
                             dA.rowsOfTile(dA.nt - 1));
        // R(m,n) -= L(m,k) L(n,k)^T, m >= n > k
        if (!rc && below > 0)
            rc = trailing_update(parsec, s, R, dA.tile(0, k), dA.tileElems, dA.tile(0, k), dA.tileElems, k + 1,
                                 dA.nt, -1.0, tk);
    }
    parsec->cplx = 0;
    if (!rc) rc = inf_norm(parsec, R, true, dplasmaLower == uplo ? dplasmaLower : dplasmaLower, &Rnorm);
    cudaStreamSynchronize(s);
    if (diagCopy) cudaFree(diagCopy);
    stage_free(&R);
    stage_free(&dA);
    stage_free(&dA0);
    if (rc) {
        fprintf(stderr, "check_dpotrf: internal error %d\n", rc);
        return 1;
    }
    result = Rnorm / (Anorm * N * eps);
    parsec->lastCheck = result;
    if (loud > 2) {
        printf("============\n");
        printf("Checking the Cholesky factorization \n");
        if (loud > 3) printf("-- ||A||_oo = %e, ||L'L-A||_oo = %e\n", Anorm, Rnorm);
        printf("-- ||L'L-A||_oo/(||A||_oo.N.eps) = %e \n", result);
    }
    int bad = (std::isnan(Rnorm) || std::isinf(Rnorm) || std::isnan(result) || std::isinf(result) || result > 60.0);
    if (loud) printf(bad ? "-- Factorization is suspicious ! \n" : "-- Factorization is CORRECT ! \n");
    return bad;
}

int check_zpotrf(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                 parsec_tiled_matrix_t* A0) {
    if (A && A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return 1;
    return check_dpotrf(parsec, loud, uplo, A, A0);
}
double dplasma_zlanhe(parsec_context_t* parsec, dplasma_enum_t ntype, dplasma_enum_t uplo,
                      const parsec_tiled_matrix_t* A) {
    return dplasma_dlansy(parsec, ntype, uplo, A);
}

int check_daxmb(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                parsec_tiled_matrix_t* b, parsec_tiled_matrix_t* x) {
    if (!parsec) return 1;
    const bool multi = A->super.nodes > 1;
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    DevMatrix dA, dB, dX;
    double* diagCopy = nullptr;
    double Anorm = 0, Bnorm = 0, Xnorm = 0, Rnorm = 0;
    const int N = b->m;
    const double eps = 0x1.0p-53;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, b, &dB, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, x, &dX, true, dplasmaUpperLower);
    if (!rc) rc = inf_norm(parsec, dA, true, uplo, &Anorm);
    if (!rc) rc = inf_norm(parsec, dB, false, 0, &Bnorm);
    if (!rc) rc = inf_norm(parsec, dX, false, 0, &Xnorm);
    if (!rc && cudaMalloc(&diagCopy, (size_t)dA.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
    // b <- b - A x with A symmetric, stored in uplo (dplasma_zhemm(Left, uplo, -1, A, x, 1, b), :199)
    if (!rc && multi) rc = symm_residual_multi(parsec, s, uplo, dA, dX, dB);
    if (!rc && !multi) rc = symm_residual_1gpu(parsec, s, uplo, dA, dX, dB, diagCopy);
    if (!rc) rc = inf_norm(parsec, dB, false, 0, &Rnorm);
    if (!rc) rc = stage_out(parsec, b, &dB, dplasmaUpperLower);
    cudaStreamSynchronize(s);
    if (diagCopy) cudaFree(diagCopy);
    stage_free(&dA);
    stage_free(&dB);
    stage_free(&dX);
    if (rc) {
        fprintf(stderr, "check_daxmb: internal error %d\n", rc);
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
