/* A driver that follows the four checks of the reference's tests/testing_zposv.c (d precision), compiled by a
 * plain C compiler against include/dplasma_b200.h and linked with libdplasma_b200.so.  For both triangles:
 *   1. dplasma_dposv                                   -> check_dpotrf + check_daxmb   (testing_zposv.c:75-115)
 *   2. dplasma_dpotrf + dplasma_dpotrs                 -> check_dpotrf + check_daxmb   (:117-165)
 *   3. dplasma_dpotrf + dplasma_dtrsm + dplasma_dtrsm  -> check_dpotrf + check_daxmb   (:167-215)
 *   4. dplasma_dpotrf + dplasma_dpotri                 -> check_dpoinv                 (:217-254)
 * A0 is a full block-cyclic matrix, A a sym block-cyclic one holding only the uplo triangle (:39-42,67-70).  Where the
 * reference copies A0 into A with dplasma_zlacpy, the generator is simply run again (same seed, same values).
 * usage: testing_dposv N NB NRHS ; exit code 0 iff every check passed. */
#include <stdio.h>
#include <stdlib.h>
#include "dplasma_b200.h"

#define TM(x) ((parsec_tiled_matrix_t *)&(x))

int main(int argc, char **argv)
{
    int N = argc > 1 ? atoi(argv[1]) : 378, NB = argc > 2 ? atoi(argv[2]) : 93, NRHS = argc > 3 ? atoi(argv[3]) : 367;
    int rank = 0, P = 1, Q = 1, ret = 0, loud = 1, u;
    dplasma_enum_t uplo[2] = {dplasmaUpper, dplasmaLower};
    const char *uplostr[2] = {"Upper", "Lower"};
    parsec_context_t *parsec = parsec_init(-1, &argc, &argv);
    if (!parsec) return 2;

    parsec_matrix_block_cyclic_t dcA0, dcB, dcX;
    parsec_matrix_block_cyclic_init(&dcA0, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, N, 0, 0, N, N, P, Q, 1, 1, 0, 0);
    parsec_matrix_block_cyclic_init(&dcB, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, NRHS, 0, 0, N, NRHS, P, Q, 1, 1, 0, 0);
    parsec_matrix_block_cyclic_init(&dcX, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, NRHS, 0, 0, N, NRHS, P, Q, 1, 1, 0, 0);
    dcA0.mat = parsec_data_allocate((size_t)dcA0.super.nb_local_tiles * (size_t)dcA0.super.bsiz * sizeof(double));
    dcB.mat = parsec_data_allocate((size_t)dcB.super.nb_local_tiles * (size_t)dcB.super.bsiz * sizeof(double));
    dcX.mat = parsec_data_allocate((size_t)dcX.super.nb_local_tiles * (size_t)dcX.super.bsiz * sizeof(double));
    dplasma_dplgsy(parsec, (double)N, dplasmaUpperLower, TM(dcA0), 3872); /* testing_zposv.c:56-57 */

    for (u = 0; u < 2; u++) {
        dplasma_enum_t t1 = (uplo[u] == dplasmaUpper) ? dplasmaTrans : dplasmaNoTrans;
        dplasma_enum_t t2 = (uplo[u] == dplasmaUpper) ? dplasmaNoTrans : dplasmaTrans;
        parsec_matrix_sym_block_cyclic_t dcA;
        int check, info, info_facto, info_solve;
        parsec_matrix_sym_block_cyclic_init(&dcA, PARSEC_MATRIX_DOUBLE, rank, NB, NB, N, N, 0, 0, N, N, P, Q, uplo[u]);
        dcA.mat = parsec_data_allocate((size_t)dcA.super.nb_local_tiles * (size_t)dcA.super.bsiz * sizeof(double));
        for (check = 1; check <= 4; check++) {
            static const char *what[5] = {"", "DPOSV", "DPOTRF + DPOTRS", "DPOTRF + DTRSM + DTRSM", "DPOTRF + DPOTRI"};
            info_facto = info_solve = 0;
            printf("***************************************************\n");
            /* create A and X */
            dplasma_dplgsy(parsec, (double)N, uplo[u], TM(dcA), 3872);
            dplasma_dplrnt(parsec, 0, TM(dcB), 2354);
            dplasma_dplrnt(parsec, 0, TM(dcX), 2354);
            /* compute */
            if (check == 1) {
                info = dplasma_dposv(parsec, uplo[u], TM(dcA), TM(dcX));
            } else {
                info = dplasma_dpotrf(parsec, uplo[u], TM(dcA));
                if (info == 0 && check == 2) info = dplasma_dpotrs(parsec, uplo[u], TM(dcA), TM(dcX));
                if (info == 0 && check == 3) {
                    dplasma_dtrsm(parsec, dplasmaLeft, uplo[u], t1, dplasmaNonUnit, 1.0, TM(dcA), TM(dcX));
                    dplasma_dtrsm(parsec, dplasmaLeft, uplo[u], t2, dplasmaNonUnit, 1.0, TM(dcA), TM(dcX));
                }
                if (info == 0 && check == 4) info = dplasma_dpotri(parsec, uplo[u], TM(dcA));
            }
            if (info != 0) printf("%d: Info = %d\n", rank, info);
            /* check the result */
            if (info == 0 && check < 4) {
                info_facto = check_dpotrf(parsec, loud, uplo[u], TM(dcA), TM(dcA0));
                info_solve = check_daxmb(parsec, loud, uplo[u], TM(dcA0), TM(dcB), TM(dcX));
            }
            if (info == 0 && check == 4) info_solve = check_dpoinv(parsec, loud, uplo[u], TM(dcA0), TM(dcA));
            if (info_solve || info_facto || info) {
                printf(" ----- TESTING %s (%s) ... FAILED !\n", what[check], uplostr[u]);
                ret |= 1;
            } else {
                printf(" ----- TESTING %s (%s) ....... PASSED !\n", what[check], uplostr[u]);
            }
            printf("***************************************************\n");
        }
        parsec_data_free(dcA.mat);
        parsec_tiled_matrix_destroy(TM(dcA));
    }
    parsec_data_free(dcA0.mat); parsec_tiled_matrix_destroy(TM(dcA0));
    parsec_data_free(dcB.mat); parsec_tiled_matrix_destroy(TM(dcB));
    parsec_data_free(dcX.mat); parsec_tiled_matrix_destroy(TM(dcX));
    parsec_fini(&parsec);
    return ret;
}
