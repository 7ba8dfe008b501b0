/*
 * TEST INFRASTRUCTURE ONLY -- CPU oracle for the tiled-Cholesky hot path.
 *
 * A plain-C restatement (no BLAS, no CUDA) of the arithmetic of therault/dplasma's
 * dpotrf / zpotrf / dposv path: matrix generators, the four tile kernels, the tiled
 * loop nest, the two solves and the reference's residual checks.  Every function cites
 * the reference file:line it follows (see oracle_impl.h).  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library; the product (libdplasma_b200.so) never does.
 *
 * Parity pinning: the reference has no golden vectors (every reference test is a
 * residual threshold, src/dplasma_zcheck.c:119,215).  This oracle is pinned instead
 * against outputs of the reference's own compiled tile kernels and generators
 * (oracle/_ref/libdplasma_ref.so, built by oracle/build_ref.sh from /root/reference),
 * frozen as fixtures in tests/golden/ by tests/golden/make_golden.py:
 * generators bit-exact, factor/solve values to 1e-12 relative (BLAS summation order).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <complex.h>

/* enum values: src/include/dplasma/constants.h:78-96 */
#define ORC_NOTRANS 111
#define ORC_TRANS 112
#define ORC_CONJTRANS 113
#define ORC_UPPER 121
#define ORC_LOWER 122
#define ORC_UPPERLOWER 123
#define ORC_NONUNIT 131
#define ORC_UNIT 132
#define ORC_LEFT 141
#define ORC_RIGHT 142

/* src/cores/random.h:19-22 */
#define ORC_RND64_A 6364136223846793005ULL
#define ORC_RND64_C 1ULL
#define ORC_RNDF_MUL 5.4210108624275222e-20f

/* src/cores/random.h:24-41: n-fold application of x -> a x + c by repeated squaring */
uint64_t orc_rnd64_jump(uint64_t n, uint64_t seed)
{
    uint64_t a_k = ORC_RND64_A, c_k = ORC_RND64_C, ran = seed;
    for (; n; n >>= 1) {
        if (n & 1) ran = a_k * ran + c_k;
        c_k *= (a_k + 1);
        a_k *= a_k;
    }
    return ran;
}

/* src/flops.h:86-87,92-93,270-273,282 */
double orc_flops_dpotrf(double n) { return n * n * n / 3.0 + n * n / 2.0 + n / 6.0; }
double orc_flops_zpotrf(double n)
{
    double fmuls = n * (((1. / 6.) * n + 0.5) * n + (1. / 3.));
    double fadds = n * (((1. / 6.) * n) * n - (1. / 6.));
    return 6.0 * fmuls + 2.0 * fadds;
}
double orc_flops_dpotrs(double n, double nrhs) { return 2.0 * nrhs * n * n; }

/* real double */
#define T double
#define P(x) orc_d##x
#define ORC_CONJ(x) (x)
#define ORC_ABS(x) fabs(x)
#define ORC_NBELEM 1ULL
#include "oracle_impl.h"
#undef T
#undef P
#undef ORC_CONJ
#undef ORC_ABS
#undef ORC_NBELEM

/* complex double */
#define ORC_COMPLEX 1
#define T double _Complex
#define P(x) orc_z##x
#define ORC_CONJ(x) conj(x)
#define ORC_ABS(x) cabs(x)
#define ORC_NBELEM 2ULL
#include "oracle_impl.h"
