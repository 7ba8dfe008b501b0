/*
 * TEST INFRASTRUCTURE ONLY (never linked into, or called from, the product path).
 *
 * ref_driver.c -- drives the REFERENCE's own CPU tile kernels (CORE_d* / CORE_z*,
 * compiled by oracle/build_ref.sh from /root/reference/src/cores through the reference's
 * PrecisionGenerator and linked to OpenBLAS) in the order the reference's task graph
 * prescribes.  The reference's runtime (PaRSEC) is not in its tree, so the DAG of
 * src/zpotrf_L.jdf / zpotrf_U.jdf is restated here with OpenMP task dependencies:
 *
 *   potrf(k)      src/zpotrf_L.jdf:174-188  (U: zpotrf_U.jdf:174-190)
 *   trsm(m,k)     src/zpotrf_L.jdf:281-300  (U: zpotrf_U.jdf:283-300)
 *   herk(k,m)     src/zpotrf_L.jdf:383-401  (U: zpotrf_U.jdf:388-405)
 *   gemm(m,n,k)   src/zpotrf_L.jdf:514-535  (U: zpotrf_U.jdf:528-548)
 *   flows         src/zpotrf_L.jdf:109-114,209-217,321-325,423-428
 *   solves        src/ztrsm_LLN.jdf:40-51,122-135, ztrsm_LLT.jdf:40-51,123-136,
 *                 ztrsm_LUN.jdf / ztrsm_LUT.jdf, sequenced by src/zposv_wrapper.c:102-110
 *   generators    src/zplgsy_wrapper.c:36-42 (tile operator) -> CORE_dplgsy / CORE_zplghe,
 *                 src/zplrnt_wrapper.c -> CORE_dplrnt / CORE_zplrnt
 *
 * Matrices here are plain column-major (LAPACK) arrays with leading dimension lda; a
 * tile (m,n) is the nb x nb block at A + m*nb + n*nb*lda (the reference runs the same
 * bodies on LAPACK storage too: src/dplasmajdf_lapack_dtt.h:20-35).
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <complex.h>
#include <omp.h>
#include "common.h" /* reference: src/cores/common.h -> CORE_* prototypes, PLASMA enums */

#define TILE(A, lda, nb, m, n) ((A) + (size_t)(m) * (nb) + (size_t)(n) * (nb) * (size_t)(lda))
static inline int tdim(int N, int nb, int t) {
    int nt = (N + nb - 1) / nb;
    return (t == nt - 1) ? N - t * nb : nb;
}

int ref_num_threads(void) { return omp_get_max_threads(); }
void ref_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }
extern void openblas_set_num_threads(int);
extern char *openblas_get_config(void);
extern char *openblas_get_corename(void);
const char *ref_blas_config(void) { return openblas_get_config(); }
const char *ref_blas_corename(void) { return openblas_get_corename(); }
void ref_blas_set_threads(int n) { openblas_set_num_threads(n); }

/* ---- generators: one CORE call per tile, exactly like the tile operator ---- */
void ref_dplgsy(double bump, int uplo, int M, int N, int mb, int nb, double *A, int lda,
                unsigned long long seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            if (uplo == PlasmaLower && m < n) continue;
            if (uplo == PlasmaUpper && m > n) continue;
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            CORE_dplgsy(bump, tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb,
                        n * nb, seed);
        }
}
void ref_zplghe(double bump, int uplo, int M, int N, int mb, int nb, double _Complex *A, int lda,
                unsigned long long seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            if (uplo == PlasmaLower && m < n) continue;
            if (uplo == PlasmaUpper && m > n) continue;
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            CORE_zplghe(bump, tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb,
                        n * nb, seed);
        }
}
void ref_dplrnt(int M, int N, int mb, int nb, double *A, int lda, unsigned long long seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            CORE_dplrnt(tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb, n * nb,
                        seed);
        }
}
void ref_zplrnt(int M, int N, int mb, int nb, double _Complex *A, int lda, unsigned long long seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            CORE_zplrnt(tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb, n * nb,
                        seed);
        }
}

/* ---- tiled Cholesky, real double ---- */
int ref_dpotrf(int uplo, int N, int nb, double *A, int lda)
{
    int nt = (N + nb - 1) / nb;
    int INFO = 0;
    openblas_set_num_threads(1); /* serial BLAS under the task runtime: INSTALL.rst:642-650 */
    char *dep = (char *)calloc((size_t)nt * nt, 1); /* one dependency token per tile */
#define D(m, n) dep[(size_t)(m) + (size_t)(n) * nt]
#pragma omp parallel
#pragma omp single
    {
        for (int k = 0; k < nt; k++) {
            int tk = tdim(N, nb, k);
            if (uplo == PlasmaLower) {
#pragma omp task depend(inout : D(k, k)) firstprivate(k, tk) shared(INFO) priority(3)
                {
                    int iinfo = 0;
                    CORE_dpotrf(PlasmaLower, tk, TILE(A, lda, nb, k, k), lda, &iinfo);
                    if (iinfo != 0) {
#pragma omp critical
                        if (INFO == 0 || k * nb + iinfo < INFO) INFO = k * nb + iinfo;
                    }
                }
                for (int m = k + 1; m < nt; m++) {
                    int tm = tdim(N, nb, m);
#pragma omp task depend(in : D(k, k)) depend(inout : D(m, k)) firstprivate(k, m, tm) priority(2)
                    CORE_dtrsm(PlasmaRight, PlasmaLower, PlasmaTrans, PlasmaNonUnit, tm, nb, 1.0,
                               TILE(A, lda, nb, k, k), lda, TILE(A, lda, nb, m, k), lda);
                }
                for (int m = k + 1; m < nt; m++) {
                    int tm = tdim(N, nb, m);
#pragma omp task depend(in : D(m, k)) depend(inout : D(m, m)) firstprivate(k, m, tm) priority(1)
                    CORE_dsyrk(PlasmaLower, PlasmaNoTrans, tm, nb, -1.0, TILE(A, lda, nb, m, k),
                               lda, 1.0, TILE(A, lda, nb, m, m), lda);
                    for (int n = k + 1; n < m; n++) {
#pragma omp task depend(in : D(m, k), D(n, k)) depend(inout : D(m, n)) firstprivate(k, m, n, tm)
                        CORE_dgemm(PlasmaNoTrans, PlasmaTrans, tm, nb, nb, -1.0,
                                   TILE(A, lda, nb, m, k), lda, TILE(A, lda, nb, n, k), lda, 1.0,
                                   TILE(A, lda, nb, m, n), lda);
                    }
                }
            } else {
#pragma omp task depend(inout : D(k, k)) firstprivate(k, tk) shared(INFO) priority(3)
                {
                    int iinfo = 0;
                    CORE_dpotrf(PlasmaUpper, tk, TILE(A, lda, nb, k, k), lda, &iinfo);
                    if (iinfo != 0) {
#pragma omp critical
                        if (INFO == 0 || k * nb + iinfo < INFO) INFO = k * nb + iinfo;
                    }
                }
                for (int n = k + 1; n < nt; n++) {
                    int tn = tdim(N, nb, n);
#pragma omp task depend(in : D(k, k)) depend(inout : D(k, n)) firstprivate(k, n, tn) priority(2)
                    CORE_dtrsm(PlasmaLeft, PlasmaUpper, PlasmaTrans, PlasmaNonUnit, nb, tn, 1.0,
                               TILE(A, lda, nb, k, k), lda, TILE(A, lda, nb, k, n), lda);
                }
                for (int n = k + 1; n < nt; n++) {
                    int tn = tdim(N, nb, n);
#pragma omp task depend(in : D(k, n)) depend(inout : D(n, n)) firstprivate(k, n, tn) priority(1)
                    CORE_dsyrk(PlasmaUpper, PlasmaTrans, tn, nb, -1.0, TILE(A, lda, nb, k, n), lda,
                               1.0, TILE(A, lda, nb, n, n), lda);
                    for (int m = k + 1; m < n; m++) {
#pragma omp task depend(in : D(k, m), D(k, n)) depend(inout : D(m, n)) firstprivate(k, m, n, tn)
                        CORE_dgemm(PlasmaTrans, PlasmaNoTrans, nb, tn, nb, -1.0,
                                   TILE(A, lda, nb, k, m), lda, TILE(A, lda, nb, k, n), lda, 1.0,
                                   TILE(A, lda, nb, m, n), lda);
                    }
                }
            }
        }
    }
#undef D
    free(dep);
    return INFO;
}

/* ---- tiled Cholesky, complex double ---- */
int ref_zpotrf(int uplo, int N, int nb, double _Complex *A, int lda)
{
    int nt = (N + nb - 1) / nb;
    int INFO = 0;
    openblas_set_num_threads(1);
    char *dep = (char *)calloc((size_t)nt * nt, 1);
    const double _Complex one = 1.0, mone = -1.0;
#define D(m, n) dep[(size_t)(m) + (size_t)(n) * nt]
#pragma omp parallel
#pragma omp single
    {
        for (int k = 0; k < nt; k++) {
            int tk = tdim(N, nb, k);
            if (uplo == PlasmaLower) {
#pragma omp task depend(inout : D(k, k)) firstprivate(k, tk) shared(INFO) priority(3)
                {
                    int iinfo = 0;
                    CORE_zpotrf(PlasmaLower, tk, TILE(A, lda, nb, k, k), lda, &iinfo);
                    if (iinfo != 0) {
#pragma omp critical
                        if (INFO == 0 || k * nb + iinfo < INFO) INFO = k * nb + iinfo;
                    }
                }
                for (int m = k + 1; m < nt; m++) {
                    int tm = tdim(N, nb, m);
#pragma omp task depend(in : D(k, k)) depend(inout : D(m, k)) firstprivate(k, m, tm) priority(2)
                    CORE_ztrsm(PlasmaRight, PlasmaLower, PlasmaConjTrans, PlasmaNonUnit, tm, nb,
                               one, TILE(A, lda, nb, k, k), lda, TILE(A, lda, nb, m, k), lda);
                }
                for (int m = k + 1; m < nt; m++) {
                    int tm = tdim(N, nb, m);
#pragma omp task depend(in : D(m, k)) depend(inout : D(m, m)) firstprivate(k, m, tm) priority(1)
                    CORE_zherk(PlasmaLower, PlasmaNoTrans, tm, nb, -1.0, TILE(A, lda, nb, m, k),
                               lda, 1.0, TILE(A, lda, nb, m, m), lda);
                    for (int n = k + 1; n < m; n++) {
#pragma omp task depend(in : D(m, k), D(n, k)) depend(inout : D(m, n)) firstprivate(k, m, n, tm)
                        CORE_zgemm(PlasmaNoTrans, PlasmaConjTrans, tm, nb, nb, mone,
                                   TILE(A, lda, nb, m, k), lda, TILE(A, lda, nb, n, k), lda, one,
                                   TILE(A, lda, nb, m, n), lda);
                    }
                }
            } else {
#pragma omp task depend(inout : D(k, k)) firstprivate(k, tk) shared(INFO) priority(3)
                {
                    int iinfo = 0;
                    CORE_zpotrf(PlasmaUpper, tk, TILE(A, lda, nb, k, k), lda, &iinfo);
                    if (iinfo != 0) {
#pragma omp critical
                        if (INFO == 0 || k * nb + iinfo < INFO) INFO = k * nb + iinfo;
                    }
                }
                for (int n = k + 1; n < nt; n++) {
                    int tn = tdim(N, nb, n);
#pragma omp task depend(in : D(k, k)) depend(inout : D(k, n)) firstprivate(k, n, tn) priority(2)
                    CORE_ztrsm(PlasmaLeft, PlasmaUpper, PlasmaConjTrans, PlasmaNonUnit, nb, tn,
                               one, TILE(A, lda, nb, k, k), lda, TILE(A, lda, nb, k, n), lda);
                }
                for (int n = k + 1; n < nt; n++) {
                    int tn = tdim(N, nb, n);
#pragma omp task depend(in : D(k, n)) depend(inout : D(n, n)) firstprivate(k, n, tn) priority(1)
                    CORE_zherk(PlasmaUpper, PlasmaConjTrans, tn, nb, -1.0, TILE(A, lda, nb, k, n),
                               lda, 1.0, TILE(A, lda, nb, n, n), lda);
                    for (int m = k + 1; m < n; m++) {
#pragma omp task depend(in : D(k, m), D(k, n)) depend(inout : D(m, n)) firstprivate(k, m, n, tn)
                        CORE_zgemm(PlasmaConjTrans, PlasmaNoTrans, nb, tn, nb, mone,
                                   TILE(A, lda, nb, k, m), lda, TILE(A, lda, nb, k, n), lda, one,
                                   TILE(A, lda, nb, m, n), lda);
                    }
                }
            }
        }
    }
#undef D
    free(dep);
    return INFO;
}

/* ---- distributed-TRSM restatement (Left side only; the variants posv uses) ----
 * trans == NoTrans with Lower (ztrsm_LLN.jdf) or ConjTrans with Upper (ztrsm_LUT.jdf)
 * walk k upward; the other two walk k downward.  alpha applied at the first k only
 * (ztrsm_LLN.jdf:44). */
static void ref_dtrsm_left_seq(int uplo, int trans, int diag, int M, int NRHS, int mb, int nbr,
                               double alpha, const double *A, int lda, double *B, int ldb)
{
    int mt = (M + mb - 1) / mb, nt = (NRHS + nbr - 1) / nbr;
    int forward = (uplo == PlasmaLower) == (trans == PlasmaNoTrans);
    openblas_set_num_threads(1);
    for (int kk = 0; kk < mt; kk++) {
        int k = forward ? kk : mt - 1 - kk;
        int tk = tdim(M, mb, k);
        double lalpha = (kk == 0) ? alpha : 1.0;
#pragma omp parallel for schedule(dynamic, 1)
        for (int n = 0; n < nt; n++) {
            int tn = (n == nt - 1) ? NRHS - n * nbr : nbr;
            double *Bk = B + (size_t)k * mb + (size_t)n * nbr * ldb;
            CORE_dtrsm(PlasmaLeft, uplo, trans, diag, tk, tn, lalpha,
                       A + (size_t)k * mb + (size_t)k * mb * lda, lda, Bk, ldb);
        }
        int lo = forward ? k + 1 : 0, hi = forward ? mt : k;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
        for (int m = lo; m < hi; m++)
            for (int n = 0; n < nt; n++) {
                int tm = tdim(M, mb, m);
                int tn = (n == nt - 1) ? NRHS - n * nbr : nbr;
                const double *Bk = B + (size_t)k * mb + (size_t)n * nbr * ldb;
                double *Bm = B + (size_t)m * mb + (size_t)n * nbr * ldb;
                if (trans == PlasmaNoTrans) /* B(m) = lalpha B(m) - A(m,k) B(k) */
                    CORE_dgemm(PlasmaNoTrans, PlasmaNoTrans, tm, tn, tk, -1.0,
                               A + (size_t)m * mb + (size_t)k * mb * lda, lda, Bk, ldb, lalpha, Bm,
                               ldb);
                else /* B(m) = lalpha B(m) - A(k,m)^T B(k) */
                    CORE_dgemm(trans, PlasmaNoTrans, tm, tn, tk, -1.0,
                               A + (size_t)k * mb + (size_t)m * mb * lda, lda, Bk, ldb, lalpha, Bm,
                               ldb);
            }
    }
}
int ref_dtrsm(int side, int uplo, int trans, int diag, int M, int NRHS, int mb, int nbr,
              double alpha, const double *A, int lda, double *B, int ldb)
{
    if (side != PlasmaLeft) return -1;
    ref_dtrsm_left_seq(uplo, trans, diag, M, NRHS, mb, nbr, alpha, A, lda, B, ldb);
    return 0;
}
/* src/zpotrs_wrapper.c:58-98 / src/zposv_wrapper.c:102-110 */
int ref_dpotrs(int uplo, int N, int NRHS, int nb, int nbr, const double *A, int lda, double *B,
               int ldb)
{
    if (uplo == PlasmaUpper) {
        ref_dtrsm_left_seq(uplo, PlasmaTrans, PlasmaNonUnit, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
        ref_dtrsm_left_seq(uplo, PlasmaNoTrans, PlasmaNonUnit, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
    } else {
        ref_dtrsm_left_seq(uplo, PlasmaNoTrans, PlasmaNonUnit, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
        ref_dtrsm_left_seq(uplo, PlasmaTrans, PlasmaNonUnit, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
    }
    return 0;
}
int ref_dposv(int uplo, int N, int NRHS, int nb, int nbr, double *A, int lda, double *B, int ldb)
{
    int info = ref_dpotrf(uplo, N, nb, A, lda);
    if (info == 0) ref_dpotrs(uplo, N, NRHS, nb, nbr, A, lda, B, ldb);
    return info;
}

/* ---- inversion after the factorization -------------------------------
 * trtri: src/ztrtri_L.jdf:21-58 (trsm Right, alpha=-1), :60-106 (gemm), :108-142 (trsm Left),
 *        :160-187 (CORE_ztrtri, INFO = k*mb + info); upper mirror src/ztrtri_U.jdf:24-58,
 *        62-110,114-149,170-194.
 * lauum: src/zlauum_L.jdf:38-65 (herk ConjTrans), :66-98 (gemm C,N), :99-130 (trmm Left,
 *        ConjTrans), :131-155 (CORE_zlauum); upper mirror src/zlauum_U.jdf:42-70,75-108,
 *        112-143,147-170.
 * Step k of each DAG is run here with the tasks of the step in the order the flows force
 * (the steps are sequential along the RW chain of every tile).                          */
int ref_dtrtri(int uplo, int diag, int N, int nb, double *A, int lda)
{
    int nt = (N + nb - 1) / nb, INFO = 0;
    openblas_set_num_threads(1);
    for (int k = 0; k < nt; k++) {
        int tk = tdim(N, nb, k);
        double *Akk = TILE(A, lda, nb, k, k);
        if (uplo == PlasmaLower) {
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++)
                CORE_dtrsm(PlasmaRight, uplo, PlasmaNoTrans, diag, tdim(N, nb, m), tk, -1.0, Akk, lda,
                           TILE(A, lda, nb, m, k), lda);
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++)
                for (int n = 0; n < k; n++)
                    CORE_dgemm(PlasmaNoTrans, PlasmaNoTrans, tdim(N, nb, m), nb, tk, 1.0,
                               TILE(A, lda, nb, m, k), lda, TILE(A, lda, nb, k, n), lda, 1.0,
                               TILE(A, lda, nb, m, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++)
                CORE_dtrsm(PlasmaLeft, uplo, PlasmaNoTrans, diag, tk, nb, 1.0, Akk, lda,
                           TILE(A, lda, nb, k, n), lda);
        } else {
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = k + 1; n < nt; n++)
                CORE_dtrsm(PlasmaLeft, uplo, PlasmaNoTrans, diag, tk, tdim(N, nb, n), -1.0, Akk, lda,
                           TILE(A, lda, nb, k, n), lda);
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
            for (int m = 0; m < k; m++)
                for (int n = k + 1; n < nt; n++)
                    CORE_dgemm(PlasmaNoTrans, PlasmaNoTrans, nb, tdim(N, nb, n), tk, 1.0,
                               TILE(A, lda, nb, m, k), lda, TILE(A, lda, nb, k, n), lda, 1.0,
                               TILE(A, lda, nb, m, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++)
                CORE_dtrsm(PlasmaRight, uplo, PlasmaNoTrans, diag, nb, tk, 1.0, Akk, lda,
                           TILE(A, lda, nb, m, k), lda);
        }
        int info = 0;
        CORE_dtrtri(uplo, diag, tk, Akk, lda, &info);
        if (info != 0 && INFO == 0) INFO = k * nb + info;
    }
    return INFO;
}
int ref_dlauum(int uplo, int N, int nb, double *A, int lda)
{
    int nt = (N + nb - 1) / nb;
    openblas_set_num_threads(1);
    for (int k = 0; k < nt; k++) {
        int tk = tdim(N, nb, k);
        double *Akk = TILE(A, lda, nb, k, k);
        if (uplo == PlasmaLower) {
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++) {
                CORE_dsyrk(uplo, PlasmaTrans, nb, tk, 1.0, TILE(A, lda, nb, k, n), lda, 1.0,
                           TILE(A, lda, nb, n, n), lda);
                for (int m = n + 1; m < k; m++)
                    CORE_dgemm(PlasmaTrans, PlasmaNoTrans, nb, nb, tk, 1.0, TILE(A, lda, nb, k, m), lda,
                               TILE(A, lda, nb, k, n), lda, 1.0, TILE(A, lda, nb, m, n), lda);
            }
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++)
                CORE_dtrmm(PlasmaLeft, uplo, PlasmaTrans, PlasmaNonUnit, tk, nb, 1.0, Akk, lda,
                           TILE(A, lda, nb, k, n), lda);
        } else {
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++) {
                CORE_dsyrk(uplo, PlasmaNoTrans, nb, tk, 1.0, TILE(A, lda, nb, m, k), lda, 1.0,
                           TILE(A, lda, nb, m, m), lda);
                for (int n = m + 1; n < k; n++)
                    CORE_dgemm(PlasmaNoTrans, PlasmaTrans, nb, nb, tk, 1.0, TILE(A, lda, nb, m, k), lda,
                               TILE(A, lda, nb, n, k), lda, 1.0, TILE(A, lda, nb, m, n), lda);
            }
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++)
                CORE_dtrmm(PlasmaRight, uplo, PlasmaTrans, PlasmaNonUnit, nb, tk, 1.0, Akk, lda,
                           TILE(A, lda, nb, m, k), lda);
        }
        CORE_dlauum(uplo, tk, Akk, lda);
    }
    return 0;
}
/* src/zpotri_wrapper.c:55-86 ; src/zpoinv_wrapper.c:255-272 (the _sync form: three DAGs in a row) */
int ref_dpotri(int uplo, int N, int nb, double *A, int lda)
{
    int info = ref_dtrtri(uplo, PlasmaNonUnit, N, nb, A, lda);
    ref_dlauum(uplo, N, nb, A, lda);
    return info;
}
int ref_dpoinv(int uplo, int N, int nb, double *A, int lda)
{
    int info = ref_dpotrf(uplo, N, nb, A, lda);
    info = ref_dtrtri(uplo, PlasmaNonUnit, N, nb, A, lda);
    ref_dlauum(uplo, N, nb, A, lda);
    return info;
}

/* ---- dplasma_dgemm / dplasma_dsyrk as public ops: src/zgemm_NN.jdf:197-211 (and the NT/TN/TT
 * twins), src/zsyrk_LN.jdf:24-55,66-112 (and LT/UN/UT): per C tile, k in sequence, beta at k == 0 only. */
int ref_dgemm(int ta, int tb, int M, int N, int K, int mb, int nb, int kb, double alpha, const double *A,
              int lda, const double *B, int ldb, double beta, double *C, int ldc)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb, kt = (K + kb - 1) / kb;
    openblas_set_num_threads(1);
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            double *Cmn = C + (size_t)m * mb + (size_t)n * nb * ldc;
            for (int k = 0; k < kt; k++) {
                int tk = (k == kt - 1) ? K - k * kb : kb;
                const double *Amk = (ta == PlasmaNoTrans) ? A + (size_t)m * mb + (size_t)k * kb * lda
                                                          : A + (size_t)k * kb + (size_t)m * mb * lda;
                const double *Bkn = (tb == PlasmaNoTrans) ? B + (size_t)k * kb + (size_t)n * nb * ldb
                                                          : B + (size_t)n * nb + (size_t)k * kb * ldb;
                CORE_dgemm(ta, tb, tm, tn, tk, alpha, Amk, lda, Bkn, ldb, (k == 0) ? beta : 1.0, Cmn, ldc);
            }
        }
    return 0;
}
int ref_dsyrk(int uplo, int trans, int N, int K, int nb, int kb, double alpha, const double *A, int lda,
              double beta, double *C, int ldc)
{
    int nt = (N + nb - 1) / nb, kt = (K + kb - 1) / kb;
    int tB = (trans == PlasmaNoTrans) ? PlasmaTrans : PlasmaNoTrans;
    openblas_set_num_threads(1);
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < nt; m++) {
            if ((uplo == PlasmaLower) ? (m < n) : (m > n)) continue;
            int tm = (m == nt - 1) ? N - m * nb : nb, tn = (n == nt - 1) ? N - n * nb : nb;
            double *Cmn = C + (size_t)m * nb + (size_t)n * nb * ldc;
            for (int k = 0; k < kt; k++) {
                int tk = (k == kt - 1) ? K - k * kb : kb;
                double zb = (k == 0) ? beta : 1.0;
                const double *Amk = (trans == PlasmaNoTrans) ? A + (size_t)m * nb + (size_t)k * kb * lda
                                                             : A + (size_t)k * kb + (size_t)m * nb * lda;
                const double *Ank = (trans == PlasmaNoTrans) ? A + (size_t)n * nb + (size_t)k * kb * lda
                                                             : A + (size_t)k * kb + (size_t)n * nb * lda;
                if (m == n) CORE_dsyrk(uplo, trans, tn, tk, alpha, Ank, lda, zb, Cmn, ldc);
                else CORE_dgemm(trans, tB, tm, tn, tk, alpha, Amk, lda, Ank, lda, zb, Cmn, ldc);
            }
        }
    return 0;
}

/* Timed run used by bench.py (--impl reference and cpu_baseline): generate with the
 * reference generator, factor with the reference tile kernels under the OpenMP DAG with
 * single-threaded BLAS inside tasks (INSTALL.rst:642-650).  Returns seconds of the
 * factorization only (tests/common.h:258-263 times only start..wait). */
double ref_time_dpotrf(int uplo, int N, int nb, unsigned long long seed, int *info_out)
{
    size_t lda = (size_t)N;
    double *A = (double *)malloc(lda * N * sizeof(double));
    if (!A) return -1.0;
    openblas_set_num_threads(1);
    ref_dplgsy((double)N, uplo, N, N, nb, nb, A, (int)lda, seed);
    double t0 = omp_get_wtime();
    int info = ref_dpotrf(uplo, N, nb, A, (int)lda);
    double t1 = omp_get_wtime();
    if (info_out) *info_out = info;
    free(A);
    return t1 - t0;
}

/* Second CPU line of BASELINE.md section 5: the same matrix factored by ONE monolithic, internally
 * multithreaded LAPACKE_dpotrf call (what a user without a tile runtime would run).  Returns seconds. */
extern int LAPACKE_dpotrf_work(int layout, char uplo, int n, double *a, int lda);
double ref_time_lapacke_dpotrf(int uplo, int N, int nb, unsigned long long seed, int threads, int *info_out)
{
    size_t lda = (size_t)N;
    double *A = (double *)malloc(lda * N * sizeof(double));
    if (!A) return -1.0;
    openblas_set_num_threads(1);
    ref_dplgsy((double)N, uplo, N, N, nb, nb, A, (int)lda, seed);
    openblas_set_num_threads(threads > 0 ? threads : omp_get_max_threads());
    double t0 = omp_get_wtime();
    int info = LAPACKE_dpotrf_work(102 /* LAPACK_COL_MAJOR */, uplo == PlasmaLower ? 'L' : 'U', N, A, (int)lda);
    double t1 = omp_get_wtime();
    openblas_set_num_threads(1);
    if (info_out) *info_out = info;
    free(A);
    return t1 - t0;
}
