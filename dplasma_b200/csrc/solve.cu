// Triangular solves and the posv / potrs drivers: src/ztrsm_wrapper.c:263-310 (Left-side
// DAGs ztrsm_LLN.jdf / ztrsm_LLT.jdf / ztrsm_LUN.jdf / ztrsm_LUT.jdf), src/zpotrs_wrapper.c:58-98,
// src/zposv_wrapper.c:64-114.
//
// The reference solves each RHS tile with CORE_ztrsm on the CPU only (ztrsm_LLN.jdf:40-57 has
// no CUDA body) and updates with legacy cublasZgemm (:111-116).  Here both run on the DMMA
// engine: the diagonal tile's IB x IB diagonal blocks are inverted once (trtri leaf) and the
// tile solve becomes a recursion of products.
#include <cstdio>
#include <cstdlib>
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

// ---------------------------------------------------------------------------------------
// Left-side tile solve with LARGE inverted diagonal blocks.
//
// The recursion over 64-blocks costs 2 n/64 - 1 dependent launches per RHS tile (31 at n = 1024), each with a handful of
// CTAs: measured 0.76 ms for a 1024 x 256 RHS tile, the whole critical path of a step of the grid solves
// (profiles/solve_trace_r02.md).  Here the diagonal blocks of size `ibs` (256 by default, DPB_SOLVE_IBS) of the lower
// form of T are inverted explicitly -- built bottom-up from the 64 x 64 inverses,
//     inv [T11 0; T21 T22] = [W1 0; -W2 T21 W1, W2],
// two batched products per level, none of which depends on the right-hand side (the grid solve does them one step
// ahead on the communication stream) -- and the solve recurses only down to ibs: 2 n/ibs - 1 launches plus one copy.
// cuBLAS / MAGMA dtrsm invert diagonal blocks the same way; the error constant involves the conditioning of the
// ibs x ibs diagonal blocks of the factor instead of the 64 x 64 ones.
// Storage: wbig is an n x n array (ld = ldw >= n) that holds the inverse blocks on its block diagonal.
// ---------------------------------------------------------------------------------------
__global__ void place_inverse_blocks_kernel(const double* __restrict__ winv64, double* __restrict__ wbig, int ldw, int n,
                                            int ibs) {
    // element (i, j) of big block b: the 64 x 64 inverse when i and j fall into the same 64-block, else zero
    const long long total = (long long)((n + ibs - 1) / ibs) * ibs * ibs;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
        const int b = (int)(e / ((long long)ibs * ibs)), r = (int)(e % ((long long)ibs * ibs));
        const int i = b * ibs + r % ibs, j = b * ibs + r / ibs;
        if (i >= n || j >= n) continue;
        double v = 0.0;
        if (i / IB == j / IB) v = winv64[(size_t)(i / IB) * IB * IB + (i % IB) + (size_t)(j % IB) * IB];
        wbig[i + (size_t)j * ldw] = v;
    }
}

int solve_ibs_for(int n) {
    static const int want = [] {
        const char* e = getenv("DPB_SOLVE_IBS");
        int v = e ? atoi(e) : 256;
        int p = IB;
        while (p * 2 <= v && p < 1024) p *= 2;
        return p;
    }();
    int ibs = want;
    while (ibs > IB && (n % ibs) != 0) ibs /= 2;  // ragged tiles fall back to the block size that divides them
    if (n % IB != 0) ibs = IB;
    return ibs;
}

int ensure_solve_ws(parsec_context_s* ctx, int nb, int ncols) {
    // two sets of [wbig (nb x nb) | y (nb x nb / 2)] and one solved-tile buffer (nb x ncols)
    const size_t set = (size_t)nb * nb + (size_t)nb * nb / 2, need = 2 * set + (size_t)nb * std::max(ncols, 1);
    if (ctx->wSolveInv && ctx->wSolveInvLen >= need && ctx->wSolveInvNb == nb) return 0;
    if (ctx->wSolveInv) { cudaDeviceSynchronize(); cudaFree(ctx->wSolveInv); }
    ctx->wSolveInv = nullptr;
    ctx->wSolveInvLen = 0;
    DPB_CUDA(cudaMalloc(&ctx->wSolveInv, need * sizeof(double)));
    ctx->wSolveInvLen = need;
    ctx->wSolveInvNb = nb;
    return 0;
}
double* solve_ws_big(parsec_context_s* ctx, int set) {
    const size_t nb = (size_t)ctx->wSolveInvNb;
    return ctx->wSolveInv + (size_t)set * (nb * nb + nb * nb / 2);
}
double* solve_ws_xt(parsec_context_s* ctx) {
    const size_t nb = (size_t)ctx->wSolveInvNb;
    return ctx->wSolveInv + 2 * (nb * nb + nb * nb / 2);
}

// inverse blocks of size ibs of the lower form of T (n x n, n % ibs == 0) into wbig (ld ldw); winv64 and y are scratch
int build_inverse_blocks(parsec_context_s* ctx, cudaStream_t s, int uplo, int diag, int n, const double* T, int ldt,
                         double* winv64, double* wbig, int ldw, double* y, int ibs) {
    int rc = launch_trtri_blocks(ctx, s, T, ldt, n, uplo, diag == dplasmaUnit, winv64);
    if (rc || ibs <= IB) return rc;
    const long long total = (long long)(n / ibs) * ibs * ibs;
    place_inverse_blocks_kernel<<<(int)std::min<long long>((total + 255) / 256, 1184), 256, 0, s>>>(winv64, wbig, ldw, n, ibs);
    ctx->launches++;
    if (cudaGetLastError() != cudaSuccess) return -9;
    const bool upper = (uplo != dplasmaLower);
    for (int m = IB; m < ibs && !rc; m *= 2) {
        const int pairs = n / (2 * m);
        // lower-form block (rows of the second block, columns of the first): T(o1.., o0..) or T(o0.., o1..)^T
        const double* L21 = upper ? T + (size_t)m * ldt : T + m;
        // Y = L21 W1
        rc = gemm_single(ctx, s, m, m, m, 1.0, L21, ldt, upper, wbig, ldw, true, 0.0, y, m, 0, pairs,
                         2LL * m * (ldt + 1), 2LL * m * (ldw + 1), (long long)m * m, -1);
        // Z = -W2 Y, into the lower-left quarter of the merged block
        if (!rc)
            rc = gemm_single(ctx, s, m, m, m, -1.0, wbig + (size_t)m * (ldw + 1), ldw, false, y, m, true, 0.0, wbig + m, ldw, 0,
                             pairs, 2LL * m * (ldw + 1), (long long)m * m, 2LL * m * (ldw + 1), -1);
    }
    return rc;
}

struct LeftWs {
    const double* winv;  // ibs == IB: 64 x 64 inverse blocks (block b at winv + b IB IB, ld IB); else wbig
    int ldw, ibs;
    double* xt;          // ibs > IB: solved rows are written here (ld ldx), B keeps the updated right-hand side
    int ldx;
};

// B(r0:r0+n, :) <- alpha * op(T(r0:r0+n, r0:r0+n))^-1 * B(r0:r0+n, :)      (B is m x ncols, Left side)
// lowerEff: op(T) is lower triangular  (Lower/NoTrans or Upper/Trans)
static int trsm_left_rec(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int r0, int n, int ncols,
                         double alpha, const double* T, int ldt, const LeftWs& w, double* B, int ldb) {
    const bool lowerEff = (uplo == dplasmaLower) != trans;
    const bool big = w.ibs > IB;
    if (n <= w.ibs) {
        // X = alpha * inv(op(Tjj)) * B.  W = inv(lower form of the block):
        //   Lower/NoTrans: inv = W     Lower/Trans: inv = W^T
        //   Upper/NoTrans: inv = W^T   Upper/Trans: inv = W      (W = inv(U^T))
        const bool useWT = (uplo == dplasmaLower) == trans;
        double* Br = B + r0;
        if (big)  // out of place: several CTA row blocks read all of B(r0:r0+n, cols)
            return gemm_single(ctx, s, n, ncols, n, alpha, w.winv + (size_t)r0 * (w.ldw + 1), w.ldw, useWT, Br, ldb, true, 0.0,
                               w.xt + r0, w.ldx, 0);
        // in place: M = n <= 64 rows -> a single CTA row block per column block, which loads all
        // of B(:, cols) before it writes
        return gemm_single(ctx, s, n, ncols, n, alpha, w.winv + (size_t)(r0 / IB) * IB * IB, IB, useWT, Br, ldb, true, 0.0, Br, ldb, 0);
    }
    const int n1 = big ? (n / w.ibs / 2) * w.ibs : split_of(n), n2 = n - n1;
    const double* X = big ? w.xt : B;  // where solved rows are read from
    const int ldx = big ? w.ldx : ldb;
    int rc;
    if (lowerEff) {
        // forward: X1 = inv(T11) B1 ; B2 = alpha B2 - opT21 X1 ; X2 = inv(T22) B2
        rc = trsm_left_rec(ctx, s, uplo, trans, r0, n1, ncols, alpha, T, ldt, w, B, ldb);
        if (rc) return rc;
        // op(T)21 (n2 x n1): Lower/NoTrans -> T(r0+n1.., r0..) M-major ; Upper/Trans -> T(r0.., r0+n1..)^T K-major
        const double* T21 = trans ? T + r0 + (size_t)(r0 + n1) * ldt : T + (r0 + n1) + (size_t)r0 * ldt;
        rc = gemm_single(ctx, s, n2, ncols, n1, -1.0, T21, ldt, trans, X + r0, ldx, true, alpha, B + r0 + n1, ldb, 0);
        if (rc) return rc;
        return trsm_left_rec(ctx, s, uplo, trans, r0 + n1, n2, ncols, 1.0, T, ldt, w, B, ldb);
    }
    // backward: X2 = inv(T22) B2 ; B1 = alpha B1 - opT12 X2 ; X1 = inv(T11) B1
    rc = trsm_left_rec(ctx, s, uplo, trans, r0 + n1, n2, ncols, alpha, T, ldt, w, B, ldb);
    if (rc) return rc;
    // op(T)12 (n1 x n2): Upper/NoTrans -> T(r0.., r0+n1..) M-major ; Lower/Trans -> T(r0+n1.., r0..)^T K-major
    const double* T12 = trans ? T + (r0 + n1) + (size_t)r0 * ldt : T + r0 + (size_t)(r0 + n1) * ldt;
    rc = gemm_single(ctx, s, n1, ncols, n2, -1.0, T12, ldt, trans, X + r0 + n1, ldx, true, alpha, B + r0, ldb, 0);
    if (rc) return rc;
    return trsm_left_rec(ctx, s, uplo, trans, r0, n1, ncols, 1.0, T, ldt, w, B, ldb);
}

// the 64-block recursion, in place (winv: 64 x 64 inverse blocks)
int trsm_left_rec_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* winv, double* B, int ldb) {
    LeftWs w{winv, IB, IB, nullptr, 0};
    return trsm_left_rec(ctx, s, uplo, trans, 0, n, ncols, alpha, T, ldt, w, B, ldb);
}
// with inverse blocks of size ibs in wbig (build_inverse_blocks); xt: n x ncols scratch (ld n)
int trsm_left_big_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* wbig, int ldw, int ibs, double* xt, double* B, int ldb) {
    LeftWs w{wbig, ldw, ibs, xt, n};
    int rc = trsm_left_rec(ctx, s, uplo, trans, 0, n, ncols, alpha, T, ldt, w, B, ldb);
    if (!rc && cudaMemcpy2DAsync(B, (size_t)ldb * sizeof(double), xt, (size_t)n * sizeof(double), (size_t)n * sizeof(double), ncols,
                                 cudaMemcpyDeviceToDevice, s) != cudaSuccess)
        rc = -10;
    return rc;
}

// One tile solve, Left side: B <- alpha op(T)^-1 B, T m x m triangular, B m x n.
int trsm_left(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, int m, int n, double alpha,
              const double* T, int ldt, double* B, int ldb, double* winv) {
    if (m <= 0 || n <= 0) return 0;
    const int ibs = ctx->cplx ? IB : solve_ibs_for(m);
    if (ibs > IB) {
        int rc = ensure_solve_ws(ctx, m, n);
        double* wbig = solve_ws_big(ctx, 0);
        if (!rc) rc = build_inverse_blocks(ctx, s, uplo, diag, m, T, ldt, winv, wbig, m, wbig + (size_t)m * m, ibs);
        if (!rc) rc = trsm_left_big_entry(ctx, s, uplo, trans != dplasmaNoTrans, m, n, alpha, T, ldt, wbig, m, ibs, solve_ws_xt(ctx), B, ldb);
        return rc;
    }
    int rc = invert_diag_blocks(ctx, s, uplo, diag, m, T, ldt, winv);
    if (rc) return rc;
    return trsm_left_rec_entry(ctx, s, uplo, trans != dplasmaNoTrans, m, n, alpha, T, ldt, winv, B, ldb);
}

// dplasma_ztrsm, Left side, on staged matrices (single process).
static int trsm_left_tiled(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, double alpha,
                           const DevMatrix& A, const DevMatrix& B) {
    int rc = ensure_winv(ctx, A.nb);
    if (rc) return rc;
    const int mt = B.mt, ntr = B.nt;
    const bool tr = trans != dplasmaNoTrans;
    const bool forward = (uplo == dplasmaLower) != tr;
    // one transposed solved tile (nb_rhs x mb), context-owned
    double* xT = nullptr;
    if (!ctx->cplx) {
        const size_t need = (size_t)B.tileElems;
        if (ctx->wSolveLen < need) {
            if (ctx->wSolve) { cudaStreamSynchronize(s); cudaFree(ctx->wSolve); }
            ctx->wSolve = nullptr;
            ctx->wSolveLen = 0;
            DPB_CUDA(cudaMalloc(&ctx->wSolve, need * sizeof(double)));
            ctx->wSolveLen = need;
        }
        xT = ctx->wSolve;
    }
    for (int kk = 0; kk < mt; kk++) {
        const int k = forward ? kk : mt - 1 - kk;
        const int tk = B.rowsOfTile(k);
        const double lalpha = (kk == 0) ? alpha : 1.0;  // src/ztrsm_LLN.jdf:44
        double* winv = ctx->wInv + (size_t)(kk & 1) * ctx->wInvSlot;
        const double* Akk = A.tile(k, k);
        const int ibs = ctx->cplx ? IB : solve_ibs_for(tk);
        if (ibs > IB) {
            rc = ensure_solve_ws(ctx, A.mb, B.nb);
            if (rc) return rc;
            double* wbig = solve_ws_big(ctx, kk & 1);
            rc = build_inverse_blocks(ctx, s, uplo, diag, tk, Akk, A.mb, winv, wbig, A.mb, wbig + (size_t)A.mb * A.mb, ibs);
            for (int n = 0; n < ntr && !rc; n++)
                rc = trsm_left_big_entry(ctx, s, uplo, tr, tk, B.colsOfTile(n), lalpha, Akk, A.mb, wbig, A.mb, ibs, solve_ws_xt(ctx),
                                         B.tile(k, n), B.mb);
            if (rc) return rc;
        } else {
            rc = invert_diag_blocks(ctx, s, uplo, diag, tk, Akk, A.mb, winv);
            if (rc) return rc;
            for (int n = 0; n < ntr; n++) {
                rc = trsm_left_rec_entry(ctx, s, uplo, tr, tk, B.colsOfTile(n), lalpha, Akk, A.mb, winv, B.tile(k, n), B.mb);
                if (rc) return rc;
            }
        }
        // B(m,:) = lalpha B(m,:) - op(A)(m,k) B(k,:) for all remaining m (src/ztrsm_LLN.jdf:131-135,
        // ztrsm_LLT.jdf:132-137): one launch per RHS tile column, batched over the tile rows m
        const int lo = forward ? k + 1 : 0, hi = forward ? mt : k;
        if (hi > lo) {
            const bool stored_mk = (uplo == dplasmaLower) ? forward : !forward;  // is A(m,k) in the stored triangle?
            const double* A0 = stored_mk ? A.tile(lo, k) : A.tile(k, lo);
            const long long sA = stored_mk ? A.tileElems : (long long)A.lmt * A.tileElems;
            for (int n = 0; n < ntr; n++) {
                // the solved tile as an M-major operand: its transpose (nc x tk, ld = B.nb); see solve_panels_multi
                const double* Xn = B.tile(k, n);
                int ldx = B.mb;
                bool xK = true;
                if (!ctx->cplx && xT) {
                    rc = launch_transpose(ctx, s, B.tile(k, n), B.mb, xT, B.nb, tk, B.colsOfTile(n));
                    if (rc) return rc;
                    Xn = xT; ldx = B.nb; xK = false;
                }
                rc = gemm_single(ctx, s, B.mb, B.colsOfTile(n), tk, -1.0, A0, A.mb, !stored_mk, Xn, ldx, xK,
                                 lalpha, B.tile(lo, n), B.mb, 0, hi - lo, sA, 0, B.tileElems, B.rowsOfTile(hi - 1));
                if (rc) return rc;
            }
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// Right side:  B <- alpha * B * op(T)^-1  (src/ztrsm_RLN.jdf, ztrsm_RLT.jdf, ztrsm_RUN.jdf,
// ztrsm_RUT.jdf; CORE_ztrsm(Right, ...) at :48 of each).  B is batched over tiles (all tile
// rows of one tile column of B share the triangular tile).  Same inverse-block recursion as
// the left side; which operand form (M-major / K-major) a block takes follows from where
// op(T)(k,c) sits in memory.
// ---------------------------------------------------------------------------------------
static int trsm_right_rec(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int c0, int n, double alpha,
                          const double* T, int ldt, const double* winv, double* B, int ldb, int M, int batch,
                          long long strideB, int lastBatchM) {
    const bool upperEff = (uplo == dplasmaUpper) != trans;  // op(T) upper triangular -> forward over the columns
    if (n <= IB) {
        // X = alpha * B inv(op(Tjj)); winv = inverse of the block's lower form
        const bool leafK = (uplo == dplasmaLower) != trans;   // RLN, RUT: inv(op)[k][c] = W[k + c*IB]
        double* Bc = B + (size_t)c0 * ldb;
        return gemm_single(ctx, s, M, n, n, alpha, Bc, ldb, false, winv + (size_t)(c0 / IB) * IB * IB, IB, leafK, 0.0, Bc,
                           ldb, 0, batch, strideB, 0, strideB, lastBatchM);
    }
    const int n1 = split_of(n), n2 = n - n1;
    int rc;
    if (upperEff) {
        // X1 = a B1 inv(T11) ; B2 = a B2 - X1 opT(1,2) ; X2 = B2 inv(T22)
        rc = trsm_right_rec(ctx, s, uplo, trans, c0, n1, alpha, T, ldt, winv, B, ldb, M, batch, strideB, lastBatchM);
        if (rc) return rc;
        // opT(1,2)[k][c]: Upper/NoTrans -> T[(c0+k) + (c0+n1+c) ldt] (K-major) ; Lower/Trans -> T[(c0+n1+c) + (c0+k) ldt] (M-major)
        const double* T12 = trans ? T + (c0 + n1) + (size_t)c0 * ldt : T + c0 + (size_t)(c0 + n1) * ldt;
        rc = gemm_single(ctx, s, M, n2, n1, -1.0, B + (size_t)c0 * ldb, ldb, false, T12, ldt, !trans, alpha,
                         B + (size_t)(c0 + n1) * ldb, ldb, 0, batch, strideB, 0, strideB, lastBatchM);
        if (rc) return rc;
        return trsm_right_rec(ctx, s, uplo, trans, c0 + n1, n2, 1.0, T, ldt, winv, B, ldb, M, batch, strideB, lastBatchM);
    }
    // op(T) lower: X2 = a B2 inv(T22) ; B1 = a B1 - X2 opT(2,1) ; X1 = B1 inv(T11)
    rc = trsm_right_rec(ctx, s, uplo, trans, c0 + n1, n2, alpha, T, ldt, winv, B, ldb, M, batch, strideB, lastBatchM);
    if (rc) return rc;
    // opT(2,1)[k][c]: Lower/NoTrans -> T[(c0+n1+k) + (c0+c) ldt] (K-major) ; Upper/Trans -> T[(c0+c) + (c0+n1+k) ldt] (M-major)
    const double* T21 = trans ? T + c0 + (size_t)(c0 + n1) * ldt : T + (c0 + n1) + (size_t)c0 * ldt;
    rc = gemm_single(ctx, s, M, n1, n2, -1.0, B + (size_t)(c0 + n1) * ldb, ldb, false, T21, ldt, !trans, alpha,
                     B + (size_t)c0 * ldb, ldb, 0, batch, strideB, 0, strideB, lastBatchM);
    if (rc) return rc;
    return trsm_right_rec(ctx, s, uplo, trans, c0, n1, 1.0, T, ldt, winv, B, ldb, M, batch, strideB, lastBatchM);
}

// one tile (or a batch of tiles): B <- alpha B op(T)^-1
int trsm_right(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, int m, int n, double alpha,
               const double* T, int ldt, double* B, int ldb, double* winv, int batch, long long strideB, int lastBatchM) {
    if (m <= 0 || n <= 0 || batch <= 0) return 0;
    int rc = invert_diag_blocks(ctx, s, uplo, diag, n, T, ldt, winv);
    if (rc) return rc;
    return trsm_right_rec(ctx, s, uplo, trans != dplasmaNoTrans, 0, n, alpha, T, ldt, winv, B, ldb, m, batch, strideB,
                          lastBatchM);
}

// dplasma_ztrsm, Right side, on staged matrices (single process): X op(A) = alpha B.
static int trsm_right_tiled(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, double alpha,
                            const DevMatrix& A, const DevMatrix& B) {
    int rc = ensure_winv(ctx, A.nb);
    if (rc) return rc;
    const int mt = B.mt, nt = B.nt;
    const bool tr = trans != dplasmaNoTrans;
    const bool forward = (uplo == dplasmaUpper) != tr;  // op(A) upper: tile columns left to right
    for (int kk = 0; kk < nt; kk++) {
        const int k = forward ? kk : nt - 1 - kk;
        const int tk = B.colsOfTile(k);
        // alpha: the NoTrans DAGs scale at their first step (ztrsm_RUN.jdf:46), the Trans DAGs solve with alpha
        // at every step and update with -1/alpha (ztrsm_RLT.jdf:48,128-132)
        const double salpha = tr ? alpha : ((kk == 0) ? alpha : 1.0);
        const double galpha = tr ? -1.0 / alpha : -1.0;
        const double gbeta = tr ? 1.0 : ((kk == 0) ? alpha : 1.0);
        double* winv = ctx->wInv + (size_t)(kk & 1) * ctx->wInvSlot;
        rc = trsm_right(ctx, s, uplo, trans, diag, B.mb, tk, salpha, A.tile(k, k), A.mb, B.tile(0, k), B.mb, winv, mt,
                        B.tileElems, B.rowsOfTile(mt - 1));
        if (rc) return rc;
        const int lo = forward ? k + 1 : 0, hi = forward ? nt : k;
        for (int n = lo; n < hi; n++) {
            // B(:,n) = gbeta B(:,n) + galpha B(:,k) op(A)(k,n), batched over the tile rows of B
            const bool stored_kn = (uplo == dplasmaUpper) ? (k < n) : (k > n);   // is A(k,n) in the stored triangle?
            // NoTrans: op(A)(k,n) = A(k,n) (always stored in the direction we walk) -> B'[c][kk] = A(k,n)[kk + c ld]: K-major
            // Trans:   op(A)(k,n) = A(n,k)^T                                       -> B'[c][kk] = A(n,k)[c + kk ld]: M-major
            const double* Akn = tr ? A.tile(n, k) : A.tile(k, n);
            (void)stored_kn;
            rc = gemm_single(ctx, s, B.mb, B.colsOfTile(n), tk, galpha, B.tile(0, k), B.mb, false, Akn, A.mb, !tr, gbeta,
                             B.tile(0, n), B.mb, 0, mt, B.tileElems, 0, B.tileElems, B.rowsOfTile(mt - 1));
            if (rc) return rc;
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
    if (side == dplasmaRight && A->super.nodes > 1) {
        fprintf(stderr, "dplasma_dtrsm: side Right on a process grid is not built here\n");
        return -1;
    }
    cudaSetDevice(parsec->device);
    DevMatrix dA, dB;
    int rc = stage_in(parsec, A, &dA, true, uplo);
    if (!rc) rc = stage_in(parsec, B, &dB, true, dplasmaUpperLower);
    if (!rc && alpha == 0.0) {
        // BLAS semantics: alpha == 0 -> B = 0, A is not referenced
        rc = launch_fill(parsec, parsec->sPanel, dB.base, (long long)dB.lmt * dB.lnt * dB.tileElems, 0.0);
    } else if (!rc && side == dplasmaRight) {
        rc = trsm_right_tiled(parsec, parsec->sPanel, uplo, trans, diag, alpha, dA, dB);
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
    if (!(uplo == dplasmaLower && trans != dplasmaNoTrans && alpha == 1.0))
        return trsm_right(ctx, s, uplo, trans, diag, m, n, alpha, T, ldt, C, ldc, ctx->wInv, 1, 0, m);
    // Right side, the form the factorization uses: C <- C L^-T  (src/zpotrf_L.jdf:236-240)
    if (uplo == dplasmaLower && trans != dplasmaNoTrans && alpha == 1.0) {
        // inverses of all 64 x 64 diagonal blocks in one launch
        rc = launch_trtri_blocks(ctx, s, T, ldt, n, uplo, diag == dplasmaUnit, ctx->wInv);
        if (rc) return rc;
        return trsm_panel_lower(ctx, s, T, ldt, n, ctx->wInv, C, ldc, m, 1, 0, m);
    }
    fprintf(stderr, "dplasma_b200_dtrsm_tile: this side/uplo/trans combination is not on the Cholesky path\n");
    return -2;
}

}  // extern "C"
