// Taskpool forms (_New / _Destruct) of the operations beside the factorization, for callers that compose
// several of them in one context the way the reference does (src/zpotri_wrapper.c:68-80 under
// PARSEC_COMPOSITION: trtri_New + lauum_New, one wait):
//   dplasma_dtrsm_New    src/ztrsm_wrapper.c:94-163,185-192
//   dplasma_dtrtri_New   src/ztrtri_wrapper.c:76-130      dplasma_dlauum_New   src/zlauum_wrapper.c:70-123
//   dplasma_dpoinv_New   src/zpoinv_wrapper.c:70-132      dplasma_dgemm_New    src/zgemm_wrapper.c:436-470
//   dplasma_dsyrk_New    src/zsyrk_wrapper.c:100-160
// _New only records the operation (it "builds but does not run", src/zpotrf_wrapper.c:123); a queue holding any
// of them is executed by parsec_context_start operation by operation, in the order it was filled -- the order
// the data dependencies of a composition need -- each one as the blocking call does.
#include <cstdio>
#include "internal.h"

namespace dpb {

int run_queue_in_order(parsec_context_s* ctx) {
    double elapsed = 0.0;
    for (parsec_taskpool_s* tp : ctx->queue) {
        int info = 0, rc = 0;
        ctx->lastElapsed = 0.0;
        switch (tp->kind) {
            case TP_POTRF: rc = run_inverse_chain(ctx, "dplasma_dpotrf", OP_POTRF, tp->uplo, dplasmaNonUnit, tp->A, &info); break;
            case TP_TRTRI: rc = run_inverse_chain(ctx, "dplasma_dtrtri", OP_TRTRI, tp->uplo, tp->diag, tp->A, &info); break;
            case TP_LAUUM: rc = run_inverse_chain(ctx, "dplasma_dlauum", OP_LAUUM, tp->uplo, dplasmaNonUnit, tp->A, &info); break;
            case TP_POINV:
                rc = run_inverse_chain(ctx, "dplasma_dpoinv", OP_POTRF | OP_TRTRI | OP_LAUUM, tp->uplo, dplasmaNonUnit, tp->A, &info);
                break;
            case TP_TRSM: rc = dplasma_dtrsm(ctx, tp->side, tp->uplo, tp->trans, tp->diag, tp->alpha, tp->A, tp->C); break;
            case TP_GEMM: rc = dplasma_dgemm(ctx, tp->trans, tp->trans2, tp->alpha, tp->A, tp->B, tp->beta, tp->C); break;
            case TP_SYRK: rc = dplasma_dsyrk(ctx, tp->uplo, tp->trans, tp->alpha, tp->A, tp->beta, tp->C); break;
            default: rc = -1;
        }
        elapsed += ctx->lastElapsed;
        tp->done = true;
        if (tp->info && *tp->info == 0) *tp->info = info;
        if (rc) return rc;
    }
    ctx->seqElapsed = elapsed;
    return 0;
}

static parsec_taskpool_s* make_tp(int kind) {
    parsec_taskpool_s* tp = new parsec_taskpool_s();
    tp->kind = kind;
    tp->uplo = dplasmaLower;
    tp->A = nullptr;
    tp->info = nullptr;
    tp->hmb = 0;
    tp->staged = false;
    return tp;
}
static bool bad_uplo(int u) { return u != dplasmaUpper && u != dplasmaLower; }
static bool bad_trans(int t) { return t != dplasmaNoTrans && t != dplasmaTrans && t != dplasmaConjTrans; }

}  // namespace dpb

using namespace dpb;

extern "C" {

parsec_taskpool_t* dplasma_dtrsm_New(dplasma_enum_t side, dplasma_enum_t uplo, dplasma_enum_t trans, dplasma_enum_t diag,
                                     double alpha, const parsec_tiled_matrix_t* A, parsec_tiled_matrix_t* B) {
    // src/ztrsm_wrapper.c:104-119
    if (side != dplasmaLeft && side != dplasmaRight) { fprintf(stderr, "dplasma_dtrsm_New: illegal value of side\n"); return nullptr; }
    if (bad_uplo(uplo)) { fprintf(stderr, "dplasma_dtrsm_New: illegal value of uplo\n"); return nullptr; }
    if (bad_trans(trans)) { fprintf(stderr, "dplasma_dtrsm_New: illegal value of trans\n"); return nullptr; }
    if (diag != dplasmaUnit && diag != dplasmaNonUnit) { fprintf(stderr, "dplasma_dtrsm_New: illegal value of diag\n"); return nullptr; }
    parsec_taskpool_s* tp = make_tp(TP_TRSM);
    tp->side = side; tp->uplo = uplo; tp->trans = trans; tp->diag = diag; tp->alpha = alpha;
    tp->A = const_cast<parsec_tiled_matrix_t*>(A);
    tp->C = B;
    return tp;
}

parsec_taskpool_t* dplasma_dtrtri_New(dplasma_enum_t uplo, dplasma_enum_t diag, parsec_tiled_matrix_t* A, int* info) {
    // src/ztrtri_wrapper.c:84-92
    if (bad_uplo(uplo)) { fprintf(stderr, "dplasma_dtrtri_New: illegal value of uplo\n"); return nullptr; }
    if (diag != dplasmaUnit && diag != dplasmaNonUnit) { fprintf(stderr, "dplasma_dtrtri_New: illegal value of diag\n"); return nullptr; }
    if (info) *info = 0;
    parsec_taskpool_s* tp = make_tp(TP_TRTRI);
    tp->uplo = uplo; tp->diag = diag; tp->A = A; tp->info = info;
    return tp;
}

parsec_taskpool_t* dplasma_dlauum_New(dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    if (bad_uplo(uplo)) { fprintf(stderr, "dplasma_dlauum_New: illegal value of uplo\n"); return nullptr; }  // src/zlauum_wrapper.c:76-79
    parsec_taskpool_s* tp = make_tp(TP_LAUUM);
    tp->uplo = uplo; tp->A = A;
    return tp;
}

parsec_taskpool_t* dplasma_dpoinv_New(dplasma_enum_t uplo, parsec_tiled_matrix_t* A, int* info) {
    if (bad_uplo(uplo)) { fprintf(stderr, "dplasma_dpoinv_New: illegal value of uplo\n"); return nullptr; }  // src/zpoinv_wrapper.c:78-81
    if (info) *info = 0;
    parsec_taskpool_s* tp = make_tp(TP_POINV);
    tp->uplo = uplo; tp->A = A; tp->info = info;
    return tp;
}

parsec_taskpool_t* dplasma_dgemm_New(dplasma_enum_t transA, dplasma_enum_t transB, double alpha, const parsec_tiled_matrix_t* A,
                                     const parsec_tiled_matrix_t* B, double beta, parsec_tiled_matrix_t* C) {
    // src/zgemm_wrapper.c:445-452
    if (bad_trans(transA)) { fprintf(stderr, "dplasma_dgemm_New: illegal value of transA\n"); return nullptr; }
    if (bad_trans(transB)) { fprintf(stderr, "dplasma_dgemm_New: illegal value of transB\n"); return nullptr; }
    parsec_taskpool_s* tp = make_tp(TP_GEMM);
    tp->trans = transA; tp->trans2 = transB; tp->alpha = alpha; tp->beta = beta;
    tp->A = const_cast<parsec_tiled_matrix_t*>(A);
    tp->B = B;
    tp->C = C;
    return tp;
}

parsec_taskpool_t* dplasma_dsyrk_New(dplasma_enum_t uplo, dplasma_enum_t trans, double alpha, const parsec_tiled_matrix_t* A,
                                     double beta, parsec_tiled_matrix_t* C) {
    if (bad_uplo(uplo)) { fprintf(stderr, "dplasma_dsyrk_New: illegal value of uplo\n"); return nullptr; }
    if (bad_trans(trans)) { fprintf(stderr, "dplasma_dsyrk_New: illegal value of trans\n"); return nullptr; }
    parsec_taskpool_s* tp = make_tp(TP_SYRK);
    tp->uplo = uplo; tp->trans = trans; tp->alpha = alpha; tp->beta = beta;
    tp->A = const_cast<parsec_tiled_matrix_t*>(A);
    tp->C = C;
    return tp;
}

// every taskpool of this library is the same object: one destructor behind the reference's names
void dplasma_dtrsm_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_dtrtri_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_dlauum_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_dpoinv_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_dgemm_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_dsyrk_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }

}  // extern "C"
