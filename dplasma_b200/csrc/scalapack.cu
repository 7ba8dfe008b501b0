// ScaLAPACK-compatible entry point pdpotrf_ (src/scalapack_wrappers/dplasma_wrapper_pdpotrf.c:133-297):
// Fortran calling convention, ScaLAPACK array descriptor, block-cyclic LAPACK-storage local
// array owned by the caller.  The reference wraps the caller's memory zero-copy in a
// parsec_matrix_block_cyclic (LAPACK storage) descriptor (:177-188); here the same descriptor
// is built and the blocking dplasma_dpotrf stages it through the GPU tile grid.
#include <cstdio>
#include <cstdlib>
#include <dlfcn.h>
#include <map>
#include "internal.h"

namespace {
struct GridInfo { int nprow, npcol, myrow, mycol; };
std::map<int, GridInfo> g_grids;
parsec_context_t* g_wrapper_ctx = nullptr;
bool g_wrapper_owned = false;

// The process grid of a BLACS context.  The reference binds Cblacs_gridinfo / Cblacs_get / numroc_ at link time
// (src/scalapack_wrappers/common.h:235-263,298-301).  This library must also load into processes that have no
// BLACS at all (the Python tests, a C driver without ScaLAPACK), so the same symbols are looked up in the
// running process: an application linked with ScaLAPACK/BLACS (tools/cscalapack/pdpotrf.c) is served by its own
// BLACS exactly as with the reference; otherwise the grid registered with dplasma_b200_blacs_gridinit is used.
typedef void (*blacs_gridinfo_fn)(int, int*, int*, int*, int*);
typedef int (*numroc_fn)(int*, int*, int*, int*, int*);
bool grid_of_context(int ictxt, GridInfo* g) {
    blacs_gridinfo_fn gridinfo = (blacs_gridinfo_fn)dlsym(RTLD_DEFAULT, "Cblacs_gridinfo");
    if (gridinfo) {
        int P = -1, Q = -1, r = -1, c = -1;
        gridinfo(ictxt, &P, &Q, &r, &c);
        if (P > 0 && Q > 0 && r >= 0 && c >= 0) {  // BLACS answers -1 for a process outside the grid
            *g = GridInfo{P, Q, r, c};
            return true;
        }
    }
    auto it = g_grids.find(ictxt);
    if (it != g_grids.end()) { *g = it->second; return true; }
    *g = GridInfo{1, 1, 0, 0};
    return false;
}
// NUMROC: rows / columns of a block-cyclic dimension that land on process iproc (ScaLAPACK TOOLS/numroc.f);
// the application's own numroc_ when it has one
int local_count(int n, int nb, int iproc, int isrc, int nprocs) {
    numroc_fn ext = (numroc_fn)dlsym(RTLD_DEFAULT, "numroc_");
    if (ext) return ext(&n, &nb, &iproc, &isrc, &nprocs);
    const int mydist = (nprocs + iproc - isrc) % nprocs, nblocks = n / nb;
    int cnt = (nblocks / nprocs) * nb;
    const int extra = nblocks % nprocs;
    if (mydist < extra) cnt += nb;
    else if (mydist == extra) cnt += n % nb;
    return cnt;
}
}  // namespace

extern "C" {

void dplasma_b200_blacs_gridinit(int ictxt, int nprow, int npcol, int myrow, int mycol) {
    g_grids[ictxt] = GridInfo{nprow, npcol, myrow, mycol};
}

// parsec_init_wrapper_ / parsec_fini_wrapper_: dplasma_wrapper_parsec_init.c:98-124
void parsec_init_wrapper_(void) {
    if (!g_wrapper_ctx) {
        g_wrapper_ctx = parsec_init(-1, nullptr, nullptr);
        g_wrapper_owned = true;
    }
}
void parsec_fini_wrapper_(void) {
    if (g_wrapper_ctx && g_wrapper_owned) parsec_fini(&g_wrapper_ctx);
    g_wrapper_ctx = nullptr;
}
// lets an application that already owns a context (and its NCCL communicators) lend it to the wrapper
void dplasma_b200_wrapper_set_context(parsec_context_t* ctx) {
    g_wrapper_ctx = ctx;
    g_wrapper_owned = false;
}

void pdpotrf_(char* UPLO, int* N, double* A, int* IA, int* JA, int* DESCA, int* info) {
    *info = 0;                      // :146
    if (*N == 0) return;            // :148-150
    const int ictxt = DESCA[1], gM = DESCA[2], gN = DESCA[3], MB = DESCA[4], NB = DESCA[5];
    const int RSRC = DESCA[6], CSRC = DESCA[7], LLD = DESCA[8];  // common.h:283-292
    GridInfo g{1, 1, 0, 0};
    grid_of_context(ictxt, &g);  // PASTE_SETUP: Cblacs_gridinfo (common.h:250)
    if (MB != NB) {
        fprintf(stderr, "pdpotrf_: square blocks required (MB=%d NB=%d)\n", MB, NB);  // :83
        *info = -606;
        return;
    }
    const int cIA = *IA - 1, cJA = *JA - 1;
    if (cIA < 0 || cJA < 0 || cIA % MB != 0 || cJA % NB != 0) {
        // the reference asserts tile-aligned offsets (:209-211)
        fprintf(stderr, "pdpotrf_: IA and JA must address the first row / column of a block (IA=%d JA=%d MB=%d)\n", *IA, *JA, MB);
        *info = -4;
        return;
    }
    if (cIA + *N > gM || cJA + *N > gN || (cIA + *N < gM && *N % MB != 0) || (cJA + *N < gN && *N % NB != 0)) {
        fprintf(stderr, "pdpotrf_: sub( A ) must lie inside A and, unless it reaches the last row / column, end on a block boundary\n");
        *info = -2;
        return;
    }
    const char u = *UPLO;
    const int uplo = (u == 'U' || u == 'u') ? dplasmaUpper : dplasmaLower;
    if (!g_wrapper_ctx) parsec_init_wrapper_();
    if (!g_wrapper_ctx) { *info = -1000; return; }
    const int rank = g.myrow * g.npcol + g.mycol;
    parsec_matrix_block_cyclic_t dcA;
    // :177-188: the caller's local array, zero-copy, as a LAPACK-storage block-cyclic descriptor of sub( A )
    const int nloc = local_count(gN, NB, g.mycol, CSRC, g.npcol);  // numroc_ (common.h:252)
    parsec_matrix_block_cyclic_lapack_init(&dcA, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_LAPACK, rank, MB, NB, gM, gN, cIA, cJA,
                                           *N, *N, g.nprow, g.npcol, 1, 1, RSRC, CSRC, LLD, nloc);
    dcA.mat = A;
    // NB: like the reference wrapper (:219-222) the local info is returned without a MAX-reduce
    int rc = dplasma_dpotrf(g_wrapper_ctx, uplo, (parsec_tiled_matrix_t*)&dcA);
    *info = rc;
}

// pdtrsm_: src/scalapack_wrappers/dplasma_wrapper_pdtrsm.c:203-327.  op(A) X = alpha B (Left) or
// X op(A) = alpha B (Right); A is M x M (Left) or N x N (Right) (:263-268), B is M x N.
void pdtrsm_(char* SIDE, char* UPLO, char* TRANS, char* DIAG, int* M, int* N, double* ALPHA, double* A, int* IA,
             int* JA, int* DESCA, double* B, int* IB, int* JB, int* DESCB) {
    if (*M == 0 || *N == 0) return;
    auto up = [](char c) { return (c >= 'a' && c <= 'z') ? (char)(c - 32) : c; };
    // OP_SIDE / OP_UPLO / OP_TRANS / OP_DIAG, common.h:146-181
    const int side = up(*SIDE) == 'L' ? dplasmaLeft : dplasmaRight;
    const int uplo = up(*UPLO) == 'U' ? dplasmaUpper : dplasmaLower;
    const int trans = up(*TRANS) == 'N' ? dplasmaNoTrans : (up(*TRANS) == 'T' ? dplasmaTrans : dplasmaConjTrans);
    const int diag = up(*DIAG) == 'U' ? dplasmaUnit : dplasmaNonUnit;
    const int Am = (side == dplasmaLeft) ? *M : *N;
    const int ictxt = DESCA[1];
    if (DESCA[4] != DESCA[5] || DESCB[4] != DESCB[5] || DESCA[4] != DESCB[4]) {
        fprintf(stderr, "pdtrsm_: square blocks of one size required (A %dx%d, B %dx%d)\n", DESCA[4], DESCA[5], DESCB[4],
                DESCB[5]);
        return;
    }
    if (*IA != 1 || *JA != 1 || *IB != 1 || *JB != 1 || DESCA[2] != Am || DESCA[3] != Am || DESCB[2] != *M ||
        DESCB[3] != *N || DESCB[1] != ictxt) {
        fprintf(stderr, "pdtrsm_: only IA=JA=IB=JB=1 and whole matrices on one context are supported\n");
        return;
    }
    GridInfo g{1, 1, 0, 0};
    grid_of_context(ictxt, &g);
    if (!g_wrapper_ctx) parsec_init_wrapper_();
    if (!g_wrapper_ctx) return;
    const int rank = g.myrow * g.npcol + g.mycol, nb = DESCA[4];
    parsec_matrix_block_cyclic_t dcA, dcB;
    parsec_matrix_block_cyclic_lapack_init(&dcA, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_LAPACK, rank, nb, nb, Am, Am, 0, 0, Am,
                                           Am, g.nprow, g.npcol, 1, 1, DESCA[6], DESCA[7], DESCA[8], 0);
    dcA.super.lln = dcA.super.sln;
    dcA.mat = A;
    parsec_matrix_block_cyclic_lapack_init(&dcB, PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_LAPACK, rank, nb, nb, *M, *N, 0, 0, *M,
                                           *N, g.nprow, g.npcol, 1, 1, DESCB[6], DESCB[7], DESCB[8], 0);
    dcB.super.lln = dcB.super.sln;
    dcB.mat = B;
    const int rc = dplasma_dtrsm(g_wrapper_ctx, side, uplo, trans, diag, *ALPHA, (parsec_tiled_matrix_t*)&dcA,
                                 (parsec_tiled_matrix_t*)&dcB);
    if (rc) fprintf(stderr, "pdtrsm_: dplasma_dtrsm returned %d\n", rc);
}

// pdgemm_: src/scalapack_wrappers/dplasma_wrapper_pdgemm.c:215-221.  C (M x N) = alpha op(A) op(B) + beta C.
void pdgemm_(char* TRANSA, char* TRANSB, int* M, int* N, int* K, double* ALPHA, double* A, int* IA, int* JA, int* DESCA,
             double* B, int* IB, int* JB, int* DESCB, double* BETA, double* C, int* IC, int* JC, int* DESCC) {
    if (*M == 0 || *N == 0) return;
    auto up = [](char c) { return (c >= 'a' && c <= 'z') ? (char)(c - 32) : c; };
    const int ta = up(*TRANSA) == 'N' ? dplasmaNoTrans : (up(*TRANSA) == 'T' ? dplasmaTrans : dplasmaConjTrans);
    const int tb = up(*TRANSB) == 'N' ? dplasmaNoTrans : (up(*TRANSB) == 'T' ? dplasmaTrans : dplasmaConjTrans);
    const int Am = ta == dplasmaNoTrans ? *M : *K, An = ta == dplasmaNoTrans ? *K : *M;
    const int Bm = tb == dplasmaNoTrans ? *K : *N, Bn = tb == dplasmaNoTrans ? *N : *K;
    const int ictxt = DESCC[1], nb = DESCC[4];
    for (const int* d : {DESCA, DESCB, DESCC})
        if (d[4] != nb || d[5] != nb || d[1] != ictxt) {
            fprintf(stderr, "pdgemm_: square blocks of one size on one context required\n");
            return;
        }
    if (*IA != 1 || *JA != 1 || *IB != 1 || *JB != 1 || *IC != 1 || *JC != 1 || DESCA[2] != Am || DESCA[3] != An ||
        DESCB[2] != Bm || DESCB[3] != Bn || DESCC[2] != *M || DESCC[3] != *N) {
        fprintf(stderr, "pdgemm_: only IA=JA=IB=JB=IC=JC=1 and whole matrices are supported\n");
        return;
    }
    GridInfo g{1, 1, 0, 0};
    grid_of_context(ictxt, &g);
    if (!g_wrapper_ctx) parsec_init_wrapper_();
    if (!g_wrapper_ctx) return;
    const int rank = g.myrow * g.npcol + g.mycol;
    parsec_matrix_block_cyclic_t dc[3];
    const int rows[3] = {Am, Bm, *M}, cols[3] = {An, Bn, *N};
    const int* desc[3] = {DESCA, DESCB, DESCC};
    double* mem[3] = {A, B, C};
    for (int i = 0; i < 3; i++) {
        parsec_matrix_block_cyclic_lapack_init(&dc[i], PARSEC_MATRIX_DOUBLE, PARSEC_MATRIX_LAPACK, rank, nb, nb, rows[i],
                                               cols[i], 0, 0, rows[i], cols[i], g.nprow, g.npcol, 1, 1, desc[i][6],
                                               desc[i][7], desc[i][8], 0);
        dc[i].super.lln = dc[i].super.sln;
        dc[i].mat = mem[i];
    }
    const int rc = dplasma_dgemm(g_wrapper_ctx, ta, tb, *ALPHA, (parsec_tiled_matrix_t*)&dc[0],
                                 (parsec_tiled_matrix_t*)&dc[1], *BETA, (parsec_tiled_matrix_t*)&dc[2]);
    if (rc) fprintf(stderr, "pdgemm_: dplasma_dgemm returned %d\n", rc);
}

}  // extern "C"
