/* A driver written the way the reference's tests/testing_zpotrf.c is (d precision), compiled by a
 * plain C compiler against include/dplasma_b200.h and linked with libdplasma_b200.so:
 *   parsec_init -> allocate a sym block-cyclic matrix -> dplasma_dplgsy -> _New / add / start /
 *   wait / _Destruct (the timed region of tests/common.h:253-267) -> check_dpotrf -> dplrnt ->
 *   dplasma_dpotrs -> check_daxmb, printing the reference's "[****] TIME(s)" line.
 * usage: testing_dpotrf N NB [L|U] */
#define _POSIX_C_SOURCE 200809L
#include <stdio.h>
#include <stdlib.h>
#include <time.h>
#include "dplasma_b200.h"

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

int main(int argc, char **argv)
{
    int N = argc > 1 ? atoi(argv[1]) : 4096, NB = argc > 2 ? atoi(argv[2]) : 256, NRHS = 3;
    dplasma_enum_t uplo = (argc > 3 && argv[3][0] == 'U') ? dplasmaUpper : dplasmaLower;
    int rank = 0, P = 1, Q = 1, info = 0, ret = 0;
    unsigned long long random_seed = 3872; /* tests/common.c:274 */
    parsec_context_t *parsec = parsec_init(-1, &argc, &argv);
    if (!parsec) return 2;

    parsec_matrix_sym_block_cyclic_t dcA;
    parsec_matrix_sym_block_cyclic_init(&dcA, PARSEC_MATRIX_DOUBLE, rank, NB, NB, N, N, 0, 0, N, N, P, Q, uplo);
    dcA.mat = parsec_data_allocate((size_t)dcA.super.nb_local_tiles * (size_t)dcA.super.bsiz * sizeof(double));
    dplasma_dplgsy(parsec, (double)N, uplo, (parsec_tiled_matrix_t *)&dcA, random_seed); /* testing_zpotrf.c:50 */

    double t0 = now();
    parsec_taskpool_t *tp = dplasma_dpotrf_New(uplo, (parsec_tiled_matrix_t *)&dcA, &info);
    parsec_context_add_taskpool(parsec, tp);
    parsec_context_start(parsec);
    parsec_context_wait(parsec);
    double t1 = now();
    dplasma_dpotrf_Destruct(tp);
    printf("[****] TIME(s) %12.5f : dpotrf\tPxQxg= %3d %-3d %d NB= %4d N= %7d : %14f gflops (device %.5f s)\n", t1 - t0, P, Q,
           1, NB, N, dplasma_b200_flops_dpotrf((double)N) / 1e9 / (t1 - t0), dplasma_b200_last_elapsed(parsec));
    if (info != 0) { printf("-- Factorization is suspicious (info = %d) ! \n", info); ret |= 1; }

    /* check: regenerate A0, residual of the factorization, then a solve (testing_zpotrf.c:88-117) */
    parsec_matrix_sym_block_cyclic_t dcA0;
    parsec_matrix_sym_block_cyclic_init(&dcA0, PARSEC_MATRIX_DOUBLE, rank, NB, NB, N, N, 0, 0, N, N, P, Q, uplo);
    dcA0.mat = parsec_data_allocate((size_t)dcA0.super.nb_local_tiles * (size_t)dcA0.super.bsiz * sizeof(double));
    dplasma_dplgsy(parsec, (double)N, uplo, (parsec_tiled_matrix_t *)&dcA0, random_seed);
    ret |= check_dpotrf(parsec, 4, uplo, (parsec_tiled_matrix_t *)&dcA, (parsec_tiled_matrix_t *)&dcA0);

    parsec_matrix_block_cyclic_t dcB, dcX;
    parsec_matrix_block_cyclic_init(&dcB, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, NRHS, 0, 0, N, NRHS, P, Q, 1, 1, 0, 0);
    parsec_matrix_block_cyclic_init(&dcX, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_TILE, rank, NB, NB, N, NRHS, 0, 0, N, NRHS, P, Q, 1, 1, 0, 0);
    dcB.mat = parsec_data_allocate((size_t)dcB.super.nb_local_tiles * (size_t)dcB.super.bsiz * sizeof(double));
    dcX.mat = parsec_data_allocate((size_t)dcX.super.nb_local_tiles * (size_t)dcX.super.bsiz * sizeof(double));
    dplasma_dplrnt(parsec, 0, (parsec_tiled_matrix_t *)&dcB, random_seed + 1);
    dplasma_dplrnt(parsec, 0, (parsec_tiled_matrix_t *)&dcX, random_seed + 1);
    dplasma_dpotrs(parsec, uplo, (parsec_tiled_matrix_t *)&dcA, (parsec_tiled_matrix_t *)&dcX);
    ret |= check_daxmb(parsec, 4, uplo, (parsec_tiled_matrix_t *)&dcA0, (parsec_tiled_matrix_t *)&dcB, (parsec_tiled_matrix_t *)&dcX);

    parsec_data_free(dcA.mat); parsec_tiled_matrix_destroy((parsec_tiled_matrix_t *)&dcA);
    parsec_data_free(dcA0.mat); parsec_tiled_matrix_destroy((parsec_tiled_matrix_t *)&dcA0);
    parsec_data_free(dcB.mat); parsec_tiled_matrix_destroy((parsec_tiled_matrix_t *)&dcB);
    parsec_data_free(dcX.mat); parsec_tiled_matrix_destroy((parsec_tiled_matrix_t *)&dcX);
    parsec_fini(&parsec);
    return ret;
}
