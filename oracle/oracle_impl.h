/*
 * TEST INFRASTRUCTURE ONLY.  Precision-generic body of the CPU oracle; included twice by
 * oracle.c (once real: T=double, P(x)=orc_d##x ; once complex: T=double _Complex,
 * P(x)=orc_z##x, ORC_COMPLEX defined).  Plain C loops, no BLAS: every routine restates
 * the arithmetic the reference obtains from CBLAS/LAPACKE at the cited call site.
 * (The reference generates its s/d/c precisions from the z source the same way:
 * tools/PrecisionGenerator/subs.py:22,41-43.)
 */

/* ---------------- generators (src/cores/core_zplgsy.c:72-150, core_zplghe.c:72-154,
 * core_zplrnt.c:68-91): element k of the LCG stream is the matrix entry with
 * column-major linear index k = i + j*gM of the lower triangle; NBELEM states per entry. */
static inline T P(rnd_elem)(uint64_t *ran, int conj_im)
{
    /* float arithmetic on purpose: `0.5f - ran * RndF_Mul` (core_zplgsy.c:89) */
    float re = 0.5f - (float)(*ran) * ORC_RNDF_MUL;
    *ran = ORC_RND64_A * (*ran) + ORC_RND64_C;
#ifdef ORC_COMPLEX
    float im = 0.5f - (float)(*ran) * ORC_RNDF_MUL;
    *ran = ORC_RND64_A * (*ran) + ORC_RND64_C;
    return conj_im ? (double)re - I * (double)im : (double)re + I * (double)im;
#else
    (void)conj_im;
    return (double)re;
#endif
}

/* d: CORE_dplgsy ; z: CORE_zplghe (hermitian: real diagonal, conj mirror) */
void P(plgsy_tile)(double bump, int m, int n, T *A, int lda, int gM, int m0, int n0, uint64_t seed)
{
    uint64_t ran, jump = (uint64_t)m0 + (uint64_t)n0 * (uint64_t)gM;
    if (m0 == n0) {
        for (int j = 0; j < n; j++) {
            ran = orc_rnd64_jump(ORC_NBELEM * jump, seed);
            for (int i = j; i < m; i++) A[i + (size_t)j * lda] = P(rnd_elem)(&ran, 0);
            jump += (uint64_t)gM + 1;
        }
        for (int j = 0; j < n; j++) {
#ifdef ORC_COMPLEX
            A[j + (size_t)j * lda] = creal(A[j + (size_t)j * lda]) + bump; /* core_zplghe.c:102-104 */
#else
            A[j + (size_t)j * lda] += bump;
#endif
            for (int i = 0; i < j; i++) A[i + (size_t)j * lda] = ORC_CONJ(A[j + (size_t)i * lda]);
        }
    } else if (m0 > n0) {
        for (int j = 0; j < n; j++) {
            ran = orc_rnd64_jump(ORC_NBELEM * jump, seed);
            for (int i = 0; i < m; i++) A[i + (size_t)j * lda] = P(rnd_elem)(&ran, 0);
            jump += (uint64_t)gM;
        }
    } else {
        jump = (uint64_t)n0 + (uint64_t)m0 * (uint64_t)gM;
        for (int i = 0; i < m; i++) {
            ran = orc_rnd64_jump(ORC_NBELEM * jump, seed);
            for (int j = 0; j < n; j++) A[i + (size_t)j * lda] = P(rnd_elem)(&ran, 1);
            jump += (uint64_t)gM;
        }
    }
}

void P(plrnt_tile)(int m, int n, T *A, int lda, int gM, int m0, int n0, uint64_t seed)
{
    uint64_t ran, jump = (uint64_t)m0 + (uint64_t)n0 * (uint64_t)gM;
    for (int j = 0; j < n; j++) {
        ran = orc_rnd64_jump(ORC_NBELEM * jump, seed);
        for (int i = 0; i < m; i++) A[i + (size_t)j * lda] = P(rnd_elem)(&ran, 0);
        jump += (uint64_t)gM;
    }
}

/* tile operator applied over the uplo part (src/zplgsy_wrapper.c:36-42 + parsec_apply) */
void P(plgsy)(double bump, int uplo, int M, int N, int mb, int nb, T *A, int lda, uint64_t seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            if (uplo == ORC_LOWER && m < n) continue;
            if (uplo == ORC_UPPER && m > n) continue;
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            P(plgsy_tile)(bump, tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb,
                          n * nb, seed);
        }
}
void P(plrnt)(int M, int N, int mb, int nb, T *A, int lda, uint64_t seed)
{
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb;
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            P(plrnt_tile)(tm, tn, A + (size_t)m * mb + (size_t)n * nb * lda, lda, M, m * mb, n * nb,
                          seed);
        }
}

/* ---------------- tile kernels ---------------- */
/* CORE_zpotrf -> LAPACKE_zpotrf_work (src/cores/core_zpotrf.c:68-74).  Unblocked
 * left-looking column Cholesky (the definition LAPACK's xPOTF2 implements); info = j+1 at
 * the first non-positive (or NaN) pivot, factorization stops there. */
void P(potrf_tile)(int uplo, int n, T *A, int lda, int *info)
{
    *info = 0;
    for (int j = 0; j < n; j++) {
        double d = creal(A[j + (size_t)j * lda]);
        for (int k = 0; k < j; k++) {
            T l = (uplo == ORC_LOWER) ? A[j + (size_t)k * lda] : A[k + (size_t)j * lda];
            d -= creal(l * ORC_CONJ(l));
        }
        if (!(d > 0.0)) {
            A[j + (size_t)j * lda] = d;
            *info = j + 1;
            return;
        }
        d = sqrt(d);
        A[j + (size_t)j * lda] = d;
        if (uplo == ORC_LOWER) {
            for (int i = j + 1; i < n; i++) {
                T s = A[i + (size_t)j * lda];
                for (int k = 0; k < j; k++)
                    s -= A[i + (size_t)k * lda] * ORC_CONJ(A[j + (size_t)k * lda]);
                A[i + (size_t)j * lda] = s / d;
            }
        } else {
            for (int i = j + 1; i < n; i++) { /* row j of U, columns i>j */
                T s = A[j + (size_t)i * lda];
                for (int k = 0; k < j; k++)
                    s -= ORC_CONJ(A[k + (size_t)j * lda]) * A[k + (size_t)i * lda];
                A[j + (size_t)i * lda] = s / d;
            }
        }
    }
}

static inline T P(opA)(const T *A, int lda, int trans, int i, int j)
{ /* element (i,j) of op(A) */
    if (trans == ORC_NOTRANS) return A[i + (size_t)j * lda];
    T v = A[j + (size_t)i * lda];
    return (trans == ORC_CONJTRANS) ? ORC_CONJ(v) : v;
}

/* CORE_zgemm -> cblas_zgemm (src/cores/core_zgemm.c:90-103): C = alpha op(A) op(B) + beta C */
void P(gemm_tile)(int ta, int tb, int m, int n, int k, T alpha, const T *A, int lda, const T *B,
                  int ldb, T beta, T *C, int ldc)
{
    T *tmp = (T *)malloc((size_t)(m > 0 ? m : 1) * sizeof(T));
    for (int j = 0; j < n; j++) {
        for (int i = 0; i < m; i++) tmp[i] = 0;
        if (ta == ORC_NOTRANS) { /* axpy form: unit stride down the column of A */
            for (int l = 0; l < k; l++) {
                T b = P(opA)(B, ldb, tb, l, j);
                const T *a = A + (size_t)l * lda;
                for (int i = 0; i < m; i++) tmp[i] += a[i] * b;
            }
        } else { /* dot form: unit stride along l in A */
            for (int i = 0; i < m; i++) {
                T s = 0;
                for (int l = 0; l < k; l++) s += P(opA)(A, lda, ta, i, l) * P(opA)(B, ldb, tb, l, j);
                tmp[i] = s;
            }
        }
        for (int i = 0; i < m; i++)
            C[i + (size_t)j * ldc] = alpha * tmp[i] + beta * C[i + (size_t)j * ldc];
    }
    free(tmp);
}

/* CORE_zherk -> cblas_zherk (src/cores/core_zherk.c:84-95); d: CORE_dsyrk.
 * trans==NoTrans: C = alpha A A^H + beta C (A n x k); else C = alpha A^H A + beta C (A k x n).
 * Only the uplo triangle is referenced/written; complex: imag of the diagonal forced to 0. */
void P(herk_tile)(int uplo, int trans, int n, int k, double alpha, const T *A, int lda,
                  double beta, T *C, int ldc)
{
    T *tmp = (T *)malloc((size_t)(n > 0 ? n : 1) * sizeof(T));
    for (int j = 0; j < n; j++) {
        int i0 = (uplo == ORC_LOWER) ? j : 0, i1 = (uplo == ORC_LOWER) ? n : j + 1;
        for (int i = i0; i < i1; i++) tmp[i] = 0;
        if (trans == ORC_NOTRANS) {
            for (int l = 0; l < k; l++) {
                T b = ORC_CONJ(A[j + (size_t)l * lda]);
                const T *a = A + (size_t)l * lda;
                for (int i = i0; i < i1; i++) tmp[i] += a[i] * b;
            }
        } else {
            for (int i = i0; i < i1; i++) {
                T s = 0;
                for (int l = 0; l < k; l++)
                    s += ORC_CONJ(A[l + (size_t)i * lda]) * A[l + (size_t)j * lda];
                tmp[i] = s;
            }
        }
        for (int i = i0; i < i1; i++) {
            T v = alpha * tmp[i] + beta * C[i + (size_t)j * ldc];
            C[i + (size_t)j * ldc] = (i == j) ? (T)creal(v) : v;
        }
    }
    free(tmp);
}

/* CORE_ztrsm -> cblas_ztrsm (src/cores/core_ztrsm.c:80-93): solve op(A) X = alpha B
 * (side Left) or X op(A) = alpha B (side Right), X overwrites B (m x n). */
void P(trsm_tile)(int side, int uplo, int trans, int diag, int m, int n, T alpha, const T *A,
                  int lda, T *B, int ldb)
{
    int na = (side == ORC_LEFT) ? m : n;
    /* Is op(A) lower triangular? */
    int lower = (uplo == ORC_LOWER) == (trans == ORC_NOTRANS);
    for (int j = 0; j < n; j++)
        for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] *= alpha;
    if (side == ORC_LEFT) {
        for (int j = 0; j < n; j++) {
            T *b = B + (size_t)j * ldb;
            if (lower) {
                for (int i = 0; i < na; i++) {
                    T s = b[i];
                    for (int l = 0; l < i; l++) s -= P(opA)(A, lda, trans, i, l) * b[l];
                    b[i] = (diag == ORC_UNIT) ? s : s / P(opA)(A, lda, trans, i, i);
                }
            } else {
                for (int i = na - 1; i >= 0; i--) {
                    T s = b[i];
                    for (int l = i + 1; l < na; l++) s -= P(opA)(A, lda, trans, i, l) * b[l];
                    b[i] = (diag == ORC_UNIT) ? s : s / P(opA)(A, lda, trans, i, i);
                }
            }
        }
    } else { /* X op(A) = B: column j of X depends on columns l with op(A)(l,j) != 0 */
        if (!lower) { /* op(A) upper: X(:,j) = (B(:,j) - sum_{l<j} X(:,l) opA(l,j)) / opA(j,j) */
            for (int j = 0; j < na; j++) {
                for (int l = 0; l < j; l++) {
                    T a = P(opA)(A, lda, trans, l, j);
                    for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] -= B[i + (size_t)l * ldb] * a;
                }
                if (diag != ORC_UNIT) {
                    T d = P(opA)(A, lda, trans, j, j);
                    for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] /= d;
                }
            }
        } else {
            for (int j = na - 1; j >= 0; j--) {
                for (int l = j + 1; l < na; l++) {
                    T a = P(opA)(A, lda, trans, l, j);
                    for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] -= B[i + (size_t)l * ldb] * a;
                }
                if (diag != ORC_UNIT) {
                    T d = P(opA)(A, lda, trans, j, j);
                    for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] /= d;
                }
            }
        }
    }
}

/* ---------------- tiled Cholesky: the bodies of src/zpotrf_L.jdf:174-188,281-300,383-401,
 * 514-535 (U: zpotrf_U.jdf) executed in the sequential insertion order of
 * tests/testing_zpotrf_dtd.c:130-205, which is a topological order of the DAG. */
int P(potrf)(int uplo, int N, int nb, T *A, int lda)
{
    int nt = (N + nb - 1) / nb, INFO = 0;
#define TL(m, n) (A + (size_t)(m) * nb + (size_t)(n) * nb * (size_t)lda)
#define TD(t) (((t) == nt - 1) ? N - (t) * nb : nb)
    for (int k = 0; k < nt; k++) {
        int iinfo = 0, tk = TD(k);
        P(potrf_tile)(uplo, tk, TL(k, k), lda, &iinfo);
        if (iinfo != 0 && INFO == 0) INFO = k * nb + iinfo; /* zpotrf_L.jdf:180-182 */
        if (uplo == ORC_LOWER) {
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++)
                P(trsm_tile)(ORC_RIGHT, ORC_LOWER, ORC_CONJTRANS, ORC_NONUNIT, TD(m), nb, 1.0,
                             TL(k, k), lda, TL(m, k), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++) {
                P(herk_tile)(ORC_LOWER, ORC_NOTRANS, TD(m), nb, -1.0, TL(m, k), lda, 1.0, TL(m, m),
                             lda);
                for (int n = k + 1; n < m; n++)
                    P(gemm_tile)(ORC_NOTRANS, ORC_CONJTRANS, TD(m), nb, nb, -1.0, TL(m, k), lda,
                                 TL(n, k), lda, 1.0, TL(m, n), lda);
            }
        } else {
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = k + 1; n < nt; n++)
                P(trsm_tile)(ORC_LEFT, ORC_UPPER, ORC_CONJTRANS, ORC_NONUNIT, nb, TD(n), 1.0,
                             TL(k, k), lda, TL(k, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = k + 1; n < nt; n++) {
                P(herk_tile)(ORC_UPPER, ORC_CONJTRANS, TD(n), nb, -1.0, TL(k, n), lda, 1.0,
                             TL(n, n), lda);
                for (int m = k + 1; m < n; m++)
                    P(gemm_tile)(ORC_CONJTRANS, ORC_NOTRANS, nb, TD(n), nb, -1.0, TL(k, m), lda,
                                 TL(k, n), lda, 1.0, TL(m, n), lda);
            }
        }
    }
    return INFO;
}

/* ---------------- dplasma_ztrsm, Left side (src/ztrsm_wrapper.c:94-163 picks the DAG):
 * LLN ztrsm_LLN.jdf:40-51,122-135 ; LLT ztrsm_LLT.jdf:40-51,123-136 ; LUN / LUT mirror.
 * and the four Right-side DAGs (ztrsm_R{L,U}{N,T}.jdf).
 * Returns the reference's argument-check codes (ztrsm_wrapper.c:274-296). */
int P(trsm)(int side, int uplo, int trans, int diag, int M, int NRHS, int mb, int nbr, T alpha,
            const T *A, int lda, T *B, int ldb)
{
    if (side != ORC_LEFT && side != ORC_RIGHT) return -1;
    if (uplo != ORC_UPPER && uplo != ORC_LOWER) return -2;
    if (trans != ORC_NOTRANS && trans != ORC_TRANS && trans != ORC_CONJTRANS) return -3;
    if (diag != ORC_UNIT && diag != ORC_NONUNIT) return -4;
    int mt = (M + mb - 1) / mb, nt = (NRHS + nbr - 1) / nbr;
    if (side == ORC_RIGHT) {
        /* X op(A) = alpha B, A is NRHS x NRHS with nbr x nbr tiles, lda given.
         * ztrsm_RUN.jdf:40-57,120-140 / ztrsm_RLT.jdf:40-56,120-138 walk the tile columns upward,
         * ztrsm_RLN.jdf / ztrsm_RUT.jdf downward.  NoTrans DAGs scale by alpha at their first step and
         * update with beta = lalpha; Trans DAGs solve with alpha at every step and update with -1/alpha. */
        int upperEff = (uplo == ORC_UPPER) == (trans == ORC_NOTRANS);
        for (int kk = 0; kk < nt; kk++) {
            int k = upperEff ? kk : nt - 1 - kk;
            int tk = (k == nt - 1) ? NRHS - k * nbr : nbr;
            int tr = (trans != ORC_NOTRANS);
            T salpha = tr ? alpha : ((kk == 0) ? alpha : (T)1.0);
            T galpha = tr ? (T)(-1.0) / alpha : (T)(-1.0);
            T gbeta = tr ? (T)1.0 : ((kk == 0) ? alpha : (T)1.0);
            for (int m = 0; m < mt; m++) {
                int tm = (m == mt - 1) ? M - m * mb : mb;
                P(trsm_tile)(ORC_RIGHT, uplo, trans, diag, tm, tk, salpha, A + (size_t)k * nbr + (size_t)k * nbr * lda, lda,
                             B + (size_t)m * mb + (size_t)k * nbr * ldb, ldb);
            }
            int lo = upperEff ? k + 1 : 0, hi = upperEff ? nt : k;
            for (int n = lo; n < hi; n++) {
                int tn = (n == nt - 1) ? NRHS - n * nbr : nbr;
                for (int m = 0; m < mt; m++) {
                    int tm = (m == mt - 1) ? M - m * mb : mb;
                    const T *Bk = B + (size_t)m * mb + (size_t)k * nbr * ldb;
                    T *Bn = B + (size_t)m * mb + (size_t)n * nbr * ldb;
                    if (!tr)
                        P(gemm_tile)(ORC_NOTRANS, ORC_NOTRANS, tm, tn, tk, galpha, Bk, ldb,
                                     A + (size_t)k * nbr + (size_t)n * nbr * lda, lda, gbeta, Bn, ldb);
                    else
                        P(gemm_tile)(ORC_NOTRANS, trans, tm, tn, tk, galpha, Bk, ldb,
                                     A + (size_t)n * nbr + (size_t)k * nbr * lda, lda, gbeta, Bn, ldb);
                }
            }
        }
        return 0;
    }
    int forward = (uplo == ORC_LOWER) == (trans == ORC_NOTRANS);
    for (int kk = 0; kk < mt; kk++) {
        int k = forward ? kk : mt - 1 - kk;
        int tk = (k == mt - 1) ? M - k * mb : mb;
        T lalpha = (kk == 0) ? alpha : (T)1.0;
        for (int n = 0; n < nt; n++) {
            int tn = (n == nt - 1) ? NRHS - n * nbr : nbr;
            P(trsm_tile)(ORC_LEFT, uplo, trans, diag, tk, tn, lalpha,
                         A + (size_t)k * mb + (size_t)k * mb * lda, lda,
                         B + (size_t)k * mb + (size_t)n * nbr * ldb, ldb);
        }
        int lo = forward ? k + 1 : 0, hi = forward ? mt : k;
#pragma omp parallel for schedule(dynamic, 1)
        for (int m = lo; m < hi; m++)
            for (int n = 0; n < nt; n++) {
                int tm = (m == mt - 1) ? M - m * mb : mb;
                int tn = (n == nt - 1) ? NRHS - n * nbr : nbr;
                const T *Bk = B + (size_t)k * mb + (size_t)n * nbr * ldb;
                T *Bm = B + (size_t)m * mb + (size_t)n * nbr * ldb;
                if (trans == ORC_NOTRANS)
                    P(gemm_tile)(ORC_NOTRANS, ORC_NOTRANS, tm, tn, tk, -1.0,
                                 A + (size_t)m * mb + (size_t)k * mb * lda, lda, Bk, ldb, lalpha, Bm,
                                 ldb);
                else
                    P(gemm_tile)(trans, ORC_NOTRANS, tm, tn, tk, -1.0,
                                 A + (size_t)k * mb + (size_t)m * mb * lda, lda, Bk, ldb, lalpha, Bm,
                                 ldb);
            }
    }
    return 0;
}

/* src/zpotrs_wrapper.c:58-98 */
int P(potrs)(int uplo, int N, int NRHS, int nb, int nbr, const T *A, int lda, T *B, int ldb)
{
    if (uplo != ORC_UPPER && uplo != ORC_LOWER) return -1;
    if (uplo == ORC_UPPER) {
        P(trsm)(ORC_LEFT, uplo, ORC_CONJTRANS, ORC_NONUNIT, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
        P(trsm)(ORC_LEFT, uplo, ORC_NOTRANS, ORC_NONUNIT, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
    } else {
        P(trsm)(ORC_LEFT, uplo, ORC_NOTRANS, ORC_NONUNIT, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
        P(trsm)(ORC_LEFT, uplo, ORC_CONJTRANS, ORC_NONUNIT, N, NRHS, nb, nbr, 1.0, A, lda, B, ldb);
    }
    return 0;
}
/* src/zposv_wrapper.c:64-114 */
int P(posv)(int uplo, int N, int NRHS, int nb, int nbr, T *A, int lda, T *B, int ldb)
{
    if (uplo != ORC_UPPER && uplo != ORC_LOWER) return -1;
    int info = P(potrf)(uplo, N, nb, A, lda);
    if (info == 0) P(potrs)(uplo, N, NRHS, nb, nbr, A, lda, B, ldb);
    return info;
}

/* ---------------- norms and checks (src/dplasma_zcheck.c) ---------------- */
/* dplasma_zlanhe(InfNorm, uplo): max row sum of the hermitian matrix stored in uplo */
double P(lanhe_inf)(int uplo, int N, const T *A, int lda)
{
    double *rs = (double *)calloc((size_t)N, sizeof(double)), r = 0.0;
    for (int j = 0; j < N; j++) {
        int i0 = (uplo == ORC_LOWER) ? j : 0, i1 = (uplo == ORC_LOWER) ? N : j + 1;
        for (int i = i0; i < i1; i++) {
            double a = (i == j) ? fabs(creal(A[i + (size_t)j * lda])) : ORC_ABS(A[i + (size_t)j * lda]);
            rs[i] += a;
            if (i != j) rs[j] += a;
        }
    }
    for (int i = 0; i < N; i++) {
        if (rs[i] > r || isnan(rs[i])) r = rs[i];
    }
    free(rs);
    return r;
}
double P(lange_inf)(int M, int N, const T *A, int lda)
{
    double *rs = (double *)calloc((size_t)M, sizeof(double)), r = 0.0;
    for (int j = 0; j < N; j++)
        for (int i = 0; i < M; i++) rs[i] += ORC_ABS(A[i + (size_t)j * lda]);
    for (int i = 0; i < M; i++)
        if (rs[i] > r || isnan(rs[i])) r = rs[i];
    free(rs);
    return r;
}

/* check_zpotrf (src/dplasma_zcheck.c:64-134): R = L L^H - A0 on the uplo triangle (U^H U
 * for Upper), result = ||R||_inf(hermitian) / (||A0||_inf N eps), eps = dlamch('e') = 2^-53;
 * returns 1 ("suspicious") iff result is not finite or > 60. */
int P(check_potrf)(int uplo, int N, const T *A, int lda, const T *A0, int lda0, double *result,
                   double *Rnorm_out, double *Anorm_out)
{
    T *R = (T *)calloc((size_t)N * N, sizeof(T));
#pragma omp parallel for schedule(dynamic, 8)
    for (int j = 0; j < N; j++) {
        int i0 = (uplo == ORC_LOWER) ? j : 0, i1 = (uplo == ORC_LOWER) ? N : j + 1;
        for (int i = i0; i < i1; i++) {
            T s = 0;
            int kmax = (i < j ? i : j) + 1;
            if (uplo == ORC_LOWER)
                for (int k = 0; k < kmax; k++)
                    s += A[i + (size_t)k * lda] * ORC_CONJ(A[j + (size_t)k * lda]);
            else
                for (int k = 0; k < kmax; k++)
                    s += ORC_CONJ(A[k + (size_t)i * lda]) * A[k + (size_t)j * lda];
            R[i + (size_t)j * N] = s - A0[i + (size_t)j * lda0];
        }
    }
    double Anorm = P(lanhe_inf)(uplo, N, A0, lda0);
    double Rnorm = P(lanhe_inf)(uplo, N, R, N);
    double eps = 0x1.0p-53;
    double res = Rnorm / (Anorm * N * eps);
    free(R);
    if (result) *result = res;
    if (Rnorm_out) *Rnorm_out = Rnorm;
    if (Anorm_out) *Anorm_out = Anorm;
    return (isnan(Rnorm) || isinf(Rnorm) || isnan(res) || isinf(res) || res > 60.0) ? 1 : 0;
}

/* check_zaxmb (src/dplasma_zcheck.c:180-225): b <- b - A x (A hermitian, stored in uplo);
 * result = ||b||_inf / ((||A|| ||x|| + ||b0||) N eps); b is overwritten like the reference. */
int P(check_axmb)(int uplo, int N, int NRHS, const T *A, int lda, T *b, int ldb, const T *x,
                  int ldx, double *result)
{
    double Anorm = P(lanhe_inf)(uplo, N, A, lda);
    double Bnorm = P(lange_inf)(N, NRHS, b, ldb);
    double Xnorm = P(lange_inf)(N, NRHS, x, ldx);
#pragma omp parallel for schedule(dynamic, 1)
    for (int c = 0; c < NRHS; c++)
        for (int i = 0; i < N; i++) {
            T s = 0;
            for (int j = 0; j < N; j++) {
                T a;
                if (i == j) a = (T)creal(A[i + (size_t)i * lda]);
                else if ((uplo == ORC_LOWER) == (i > j)) a = A[i + (size_t)j * lda];
                else a = ORC_CONJ(A[j + (size_t)i * lda]);
                s += a * x[j + (size_t)c * ldx];
            }
            b[i + (size_t)c * ldb] -= s;
        }
    double Rnorm = P(lange_inf)(N, NRHS, b, ldb);
    double eps = 0x1.0p-53;
    double res = Rnorm / ((Anorm * Xnorm + Bnorm) * N * eps);
    if (result) *result = res;
    return (isnan(Xnorm) || isinf(Xnorm) || isnan(res) || isinf(res) || res > 60.0) ? 1 : 0;
}
/* ---------------- inversion after the factorization ---------------- */
/* CORE_ztrtri -> LAPACKE_ztrtri_work (src/cores/core_ztrtri.c:66-72): in-place inverse of the
 * uplo triangle; info = i (1-based) if A(i,i) == 0 with NonUnit (nothing is modified then). */
void P(trtri_tile)(int uplo, int diag, int n, T *A, int lda, int *info)
{
    *info = 0;
    if (diag != ORC_UNIT)
        for (int i = 0; i < n; i++)
            if (A[i + (size_t)i * lda] == 0) { *info = i + 1; return; }
    if (uplo == ORC_LOWER) {
        /* column j of inv(L): x_jj = 1/l_jj ; x_ij = -(sum_{l=j}^{i-1} l_il x_lj) / l_ii, against a
         * private copy of the triangle (the inverse overwrites A column by column) */
        T *L = (T *)malloc((size_t)n * n * sizeof(T));
        for (int j = 0; j < n; j++)
            for (int i = j; i < n; i++) L[i + (size_t)j * n] = A[i + (size_t)j * lda];
        for (int j = 0; j < n; j++) {
            T xjj = (diag == ORC_UNIT) ? (T)1 : 1 / L[j + (size_t)j * n];
            if (diag != ORC_UNIT) A[j + (size_t)j * lda] = xjj;
            for (int i = j + 1; i < n; i++) {
                T s = L[i + (size_t)j * n] * xjj;
                for (int l = j + 1; l < i; l++) s += L[i + (size_t)l * n] * A[l + (size_t)j * lda];
                A[i + (size_t)j * lda] = (diag == ORC_UNIT) ? -s : -s / L[i + (size_t)i * n];
            }
        }
        free(L);
    } else {
        T *U = (T *)malloc((size_t)n * n * sizeof(T));
        for (int j = 0; j < n; j++)
            for (int i = 0; i <= j; i++) U[i + (size_t)j * n] = A[i + (size_t)j * lda];
        /* column j of inv(U), rows j down to 0 */
        for (int j = 0; j < n; j++) {
            T xjj = (diag == ORC_UNIT) ? (T)1 : 1 / U[j + (size_t)j * n];
            if (diag != ORC_UNIT) A[j + (size_t)j * lda] = xjj;
            for (int i = j - 1; i >= 0; i--) {
                T s = U[i + (size_t)j * n] * xjj;
                for (int l = i + 1; l < j; l++) s += U[i + (size_t)l * n] * A[l + (size_t)j * lda];
                A[i + (size_t)j * lda] = (diag == ORC_UNIT) ? -s : -s / U[i + (size_t)i * n];
            }
        }
        free(U);
    }
}

/* CORE_zlauum -> LAPACKE_zlauum_work (src/cores/core_zlauum.c): Lower: A <- L^H L ; Upper: A <- U U^H,
 * the uplo triangle only. */
void P(lauum_tile)(int uplo, int n, T *A, int lda)
{
    T *W = (T *)malloc((size_t)(n > 0 ? n : 1) * n * sizeof(T));
    for (int j = 0; j < n; j++)
        for (int i = 0; i < n; i++) {
            int stored = (uplo == ORC_LOWER) ? (i >= j) : (i <= j);
            W[i + (size_t)j * n] = stored ? A[i + (size_t)j * lda] : 0;
        }
    for (int j = 0; j < n; j++) {
        int i0 = (uplo == ORC_LOWER) ? j : 0, i1 = (uplo == ORC_LOWER) ? n : j + 1;
        for (int i = i0; i < i1; i++) {
            T s = 0;
            if (uplo == ORC_LOWER)
                for (int l = i; l < n; l++) s += ORC_CONJ(W[l + (size_t)i * n]) * W[l + (size_t)j * n];
            else
                for (int l = j; l < n; l++) s += W[i + (size_t)l * n] * ORC_CONJ(W[j + (size_t)l * n]);
            A[i + (size_t)j * lda] = (i == j) ? (T)creal(s) : s;
        }
    }
    free(W);
}

/* CORE_ztrmm -> cblas_ztrmm (src/cores/core_ztrmm.c): B <- alpha op(A) B (Left) or alpha B op(A) (Right) */
void P(trmm_tile)(int side, int uplo, int trans, int diag, int m, int n, T alpha, const T *A, int lda,
                  T *B, int ldb)
{
    int na = (side == ORC_LEFT) ? m : n;
    T *W = (T *)malloc((size_t)na * na * sizeof(T)); /* op(A) as a dense matrix */
    for (int j = 0; j < na; j++)
        for (int i = 0; i < na; i++) {
            int r = (trans == ORC_NOTRANS) ? i : j, c = (trans == ORC_NOTRANS) ? j : i;
            int stored = (uplo == ORC_LOWER) ? (r >= c) : (r <= c);
            T v = 0;
            if (r == c) v = (diag == ORC_UNIT) ? (T)1 : A[r + (size_t)c * lda];
            else if (stored) v = A[r + (size_t)c * lda];
            if (trans == ORC_CONJTRANS) v = ORC_CONJ(v);
            W[i + (size_t)j * na] = v;
        }
    T *R = (T *)malloc((size_t)m * n * sizeof(T));
    for (int j = 0; j < n; j++)
        for (int i = 0; i < m; i++) {
            T s = 0;
            if (side == ORC_LEFT)
                for (int l = 0; l < m; l++) s += W[i + (size_t)l * na] * B[l + (size_t)j * ldb];
            else
                for (int l = 0; l < n; l++) s += B[i + (size_t)l * ldb] * W[l + (size_t)j * na];
            R[i + (size_t)j * m] = alpha * s;
        }
    for (int j = 0; j < n; j++)
        for (int i = 0; i < m; i++) B[i + (size_t)j * ldb] = R[i + (size_t)j * m];
    free(R);
    free(W);
}

/* dplasma_ztrtri: src/ztrtri_L.jdf:21-58 (trsm Right, -1), :60-106 (gemm), :108-142 (trsm Left),
 * :160-187 (trtri, INFO = k*mb + info); upper mirror src/ztrtri_U.jdf.  Step k of the DAG. */
int P(trtri)(int uplo, int diag, int N, int nb, T *A, int lda)
{
    int nt = (N + nb - 1) / nb, INFO = 0;
    for (int k = 0; k < nt; k++) {
        int tk = TD(k), info = 0;
        if (uplo == ORC_LOWER) {
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++)
                P(trsm_tile)(ORC_RIGHT, uplo, ORC_NOTRANS, diag, TD(m), tk, -1.0, TL(k, k), lda, TL(m, k), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = k + 1; m < nt; m++)
                for (int n = 0; n < k; n++)
                    P(gemm_tile)(ORC_NOTRANS, ORC_NOTRANS, TD(m), nb, tk, 1.0, TL(m, k), lda, TL(k, n), lda, 1.0,
                                 TL(m, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++)
                P(trsm_tile)(ORC_LEFT, uplo, ORC_NOTRANS, diag, tk, nb, 1.0, TL(k, k), lda, TL(k, n), lda);
        } else {
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = k + 1; n < nt; n++)
                P(trsm_tile)(ORC_LEFT, uplo, ORC_NOTRANS, diag, tk, TD(n), -1.0, TL(k, k), lda, TL(k, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = k + 1; n < nt; n++)
                for (int m = 0; m < k; m++)
                    P(gemm_tile)(ORC_NOTRANS, ORC_NOTRANS, nb, TD(n), tk, 1.0, TL(m, k), lda, TL(k, n), lda, 1.0,
                                 TL(m, n), lda);
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++)
                P(trsm_tile)(ORC_RIGHT, uplo, ORC_NOTRANS, diag, nb, tk, 1.0, TL(k, k), lda, TL(m, k), lda);
        }
        P(trtri_tile)(uplo, diag, tk, TL(k, k), lda, &info);
        if (info != 0 && INFO == 0) INFO = k * nb + info;
    }
    return INFO;
}

/* dplasma_zlauum: src/zlauum_L.jdf:38-65 (herk C), :66-98 (gemm C,N), :99-130 (trmm Left C),
 * :131-155 (lauum tile); upper mirror src/zlauum_U.jdf. */
int P(lauum)(int uplo, int N, int nb, T *A, int lda)
{
    int nt = (N + nb - 1) / nb;
    for (int k = 0; k < nt; k++) {
        int tk = TD(k);
        if (uplo == ORC_LOWER) {
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++) {
                P(herk_tile)(uplo, ORC_CONJTRANS, nb, tk, 1.0, TL(k, n), lda, 1.0, TL(n, n), lda);
                for (int m = n + 1; m < k; m++)
                    P(gemm_tile)(ORC_CONJTRANS, ORC_NOTRANS, nb, nb, tk, 1.0, TL(k, m), lda, TL(k, n), lda, 1.0,
                                 TL(m, n), lda);
            }
#pragma omp parallel for schedule(dynamic, 1)
            for (int n = 0; n < k; n++)
                P(trmm_tile)(ORC_LEFT, uplo, ORC_CONJTRANS, ORC_NONUNIT, tk, nb, 1.0, TL(k, k), lda, TL(k, n), lda);
        } else {
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++) {
                P(herk_tile)(uplo, ORC_NOTRANS, nb, tk, 1.0, TL(m, k), lda, 1.0, TL(m, m), lda);
                for (int n = m + 1; n < k; n++)
                    P(gemm_tile)(ORC_NOTRANS, ORC_CONJTRANS, nb, nb, tk, 1.0, TL(m, k), lda, TL(n, k), lda, 1.0,
                                 TL(m, n), lda);
            }
#pragma omp parallel for schedule(dynamic, 1)
            for (int m = 0; m < k; m++)
                P(trmm_tile)(ORC_RIGHT, uplo, ORC_CONJTRANS, ORC_NONUNIT, nb, tk, 1.0, TL(k, k), lda, TL(m, k), lda);
        }
        P(lauum_tile)(uplo, tk, TL(k, k), lda);
    }
    return 0;
}

/* src/zpotri_wrapper.c:55-86 */
int P(potri)(int uplo, int N, int nb, T *A, int lda)
{
    int info = P(trtri)(uplo, ORC_NONUNIT, N, nb, A, lda);
    P(lauum)(uplo, N, nb, A, lda);
    return info;
}
/* src/zpoinv_wrapper.c:184-207 (one DAG, INFO of the factorization) */
int P(poinv)(int uplo, int N, int nb, T *A, int lda)
{
    int info = P(potrf)(uplo, N, nb, A, lda);
    if (info) return info;
    return P(potri)(uplo, N, nb, A, lda);
}

/* check_zpoinv (src/dplasma_zcheck.c:265-315): R = Id - Ainv A (hemm Left, Ainv stored in uplo, A read
 * as a general matrix); result = ||R||_1 / (||A||_1 ||Ainv||_1 N eps); suspicious iff not finite or > 10. */
int P(check_poinv)(int uplo, int N, const T *A, int lda, const T *Ainv, int ldi, double *result)
{
    double Anorm = P(lanhe_inf)(uplo, N, A, lda);      /* one norm == inf norm for hermitian */
    double Inorm = P(lanhe_inf)(uplo, N, Ainv, ldi);
    double *cs = (double *)calloc((size_t)N, sizeof(double)), Rnorm = 0.0;
#pragma omp parallel for schedule(dynamic, 1)
    for (int c = 0; c < N; c++) {
        double colsum = 0.0;
        for (int i = 0; i < N; i++) {
            T s = (i == c) ? (T)1 : (T)0;
            for (int j = 0; j < N; j++) {
                T a;
                if (i == j) a = (T)creal(Ainv[i + (size_t)i * ldi]);
                else if ((uplo == ORC_LOWER) == (i > j)) a = Ainv[i + (size_t)j * ldi];
                else a = ORC_CONJ(Ainv[j + (size_t)i * ldi]);
                s -= a * A[j + (size_t)c * lda];
            }
            colsum += ORC_ABS(s);
        }
        cs[c] = colsum;
    }
    for (int c = 0; c < N; c++)
        if (cs[c] > Rnorm || isnan(cs[c])) Rnorm = cs[c];
    free(cs);
    double eps = 0x1.0p-53;
    double res = Rnorm / ((Anorm * Inorm) * N * eps);
    if (result) *result = res;
    return (isinf(Inorm) || isnan(res) || isinf(res) || res > 10.0) ? 1 : 0;
}
/* ---------------- dplasma_zgemm / dplasma_zherk (d: dsyrk) as public ops ----------------
 * zgemm_{NN,NT,TN,TT}.jdf: C(m,n) = lbeta C(m,n) + alpha op(A)(m,k) op(B)(k,n), k = 0..KT-1 in sequence,
 * lbeta = beta at k == 0 and 1 afterwards (src/zgemm_NN.jdf:197-211).  A is M x K (NoTrans) or K x M,
 * B is K x N (NoTrans) or N x K; tiles mb x kb, kb x nb, mb x nb.  Argument codes src/zgemm_wrapper.c:657-710. */
int P(gemm)(int ta, int tb, int M, int N, int K, int mb, int nb, int kb, T alpha, const T *A, int lda,
            const T *B, int ldb, T beta, T *C, int ldc)
{
    if (ta != ORC_NOTRANS && ta != ORC_TRANS && ta != ORC_CONJTRANS) return -1;
    if (tb != ORC_NOTRANS && tb != ORC_TRANS && tb != ORC_CONJTRANS) return -2;
    int mt = (M + mb - 1) / mb, nt = (N + nb - 1) / nb, kt = (K + kb - 1) / kb;
    if (M == 0 || N == 0 || ((alpha == 0 || K == 0) && beta == 1)) return 0; /* :716-719 */
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < mt; m++) {
            int tm = (m == mt - 1) ? M - m * mb : mb, tn = (n == nt - 1) ? N - n * nb : nb;
            T *Cmn = C + (size_t)m * mb + (size_t)n * nb * ldc;
            if (kt == 0) P(gemm_tile)(ta, tb, tm, tn, 0, alpha, A, lda, B, ldb, beta, Cmn, ldc);
            for (int k = 0; k < kt; k++) {
                int tk = (k == kt - 1) ? K - k * kb : kb;
                const T *Amk = (ta == ORC_NOTRANS) ? A + (size_t)m * mb + (size_t)k * kb * lda
                                                   : A + (size_t)k * kb + (size_t)m * mb * lda;
                const T *Bkn = (tb == ORC_NOTRANS) ? B + (size_t)k * kb + (size_t)n * nb * ldb
                                                   : B + (size_t)n * nb + (size_t)k * kb * ldb;
                P(gemm_tile)(ta, tb, tm, tn, tk, alpha, Amk, lda, Bkn, ldb, (k == 0) ? beta : (T)1, Cmn, ldc);
            }
        }
    return 0;
}

/* zherk_{L,U}{N,C}.jdf / zsyrk_{L,U}{N,T}.jdf (d: both are dsyrk): diagonal tiles by herk (src/zsyrk_LN.jdf:24-55),
 * the others by gemm with the transposed partner (:66-112), k in sequence with zbeta = beta at k == 0.
 * trans == NoTrans: C = alpha A A^H + beta C, A is N x K ; else C = alpha A^H A + beta C, A is K x N.
 * Argument codes src/zsyrk_wrapper.c:229-245. */
int P(herk)(int uplo, int trans, int N, int K, int nb, int kb, double alpha, const T *A, int lda, double beta,
            T *C, int ldc)
{
    if (uplo != ORC_LOWER && uplo != ORC_UPPER) return -1;
    if (trans != ORC_NOTRANS && trans != ORC_TRANS && trans != ORC_CONJTRANS) return -2;
    int nt = (N + nb - 1) / nb, kt = (K + kb - 1) / kb;
    int tA = (trans == ORC_NOTRANS) ? ORC_NOTRANS : ORC_CONJTRANS, tB = (trans == ORC_NOTRANS) ? ORC_CONJTRANS : ORC_NOTRANS;
#pragma omp parallel for collapse(2) schedule(dynamic, 1)
    for (int n = 0; n < nt; n++)
        for (int m = 0; m < nt; m++) {
            if ((uplo == ORC_LOWER) ? (m < n) : (m > n)) continue;
            int tm = (m == nt - 1) ? N - m * nb : nb, tn = (n == nt - 1) ? N - n * nb : nb;
            T *Cmn = C + (size_t)m * nb + (size_t)n * nb * ldc;
            for (int k = 0; k < kt; k++) {
                int tk = (k == kt - 1) ? K - k * kb : kb;
                double zb = (k == 0) ? beta : 1.0;
                const T *Amk = (trans == ORC_NOTRANS) ? A + (size_t)m * nb + (size_t)k * kb * lda
                                                      : A + (size_t)k * kb + (size_t)m * nb * lda;
                const T *Ank = (trans == ORC_NOTRANS) ? A + (size_t)n * nb + (size_t)k * kb * lda
                                                      : A + (size_t)k * kb + (size_t)n * nb * lda;
                if (m == n) P(herk_tile)(uplo, tA, tn, tk, alpha, Ank, lda, zb, Cmn, ldc);
                else P(gemm_tile)(tA, tB, tm, tn, tk, (T)alpha, Amk, lda, Ank, lda, (T)zb, Cmn, ldc);
            }
        }
    return 0;
}
#undef TL
#undef TD
