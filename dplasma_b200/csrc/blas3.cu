// Public level-3 operations on the engine of the factorization:
//   dplasma_dgemm   src/zgemm_wrapper.c:645-733   (DAGs src/zgemm_{NN,NT,TN,TT}.jdf; GPU body cublasZgemm per tile,
//                                                  zgemm_NN.jdf:165-195)
//   dplasma_dsyrk   src/zsyrk_wrapper.c:219-258   (DAGs src/zsyrk_{L,U}{N,T}.jdf; CPU bodies only)
// One process.  The contraction runs tile column by tile column of op(A) (k in sequence, beta applied at
// k == 0 only, zgemm_NN.jdf:197-211); operands that are not already in the engine's NT M-major form are
// transposed once per k into scratch (an HBM-bound pass over one tile row), so every product is
// C += A B^T with both operands M-major.  dgemm: one launch per tile column of C, batched over its tile
// rows.  dsyrk: the factorization's grouped trailing-update launch (diagonal tiles masked to their lower
// part), one per k; Upper runs as Lower on the transposed tile grid.
#include <cstdio>
#include <algorithm>
#include "internal.h"

namespace dpb {

// C <- beta C on the stored part of the local tiles of a (distributed) tile grid: uplo = Lower (global i >= j),
// Upper (i <= j) or UpperLower (everything); beta == 0 stores zeros (no 0 * NaN)
__global__ void scale_grid_kernel(DevMatrix C, int uplo, double beta) {
    const long long total = (long long)C.lmt * C.lnt * C.tileElems;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long t = e / C.tileElems, r = e % C.tileElems;
        const int il = (int)(t % C.lmt), jl = (int)(t / C.lmt);
        const long long gi = ((long long)il * C.P + C.prow) * C.mb + (int)(r % C.mb);  // global element indices
        const long long gj = ((long long)jl * C.Q + C.pcol) * C.nb + (int)(r / C.mb);
        if (gi >= C.M || gj >= C.N) continue;
        if (uplo == dplasmaLower && gi < gj) continue;
        if (uplo == dplasmaUpper && gi > gj) continue;
        C.base[e] = (beta == 0.0) ? 0.0 : beta * C.base[e];
    }
}
static int launch_scale_grid(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& C, int uplo, double beta) {
    if (beta == 1.0 || C.lmt * C.lnt == 0) return 0;
    scale_grid_kernel<<<1184, 256, 0, s>>>(C, uplo, beta);
    ctx->launches++;
    return (int)cudaGetLastError();
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_dgemm(parsec_context_t* parsec, dplasma_enum_t transA, dplasma_enum_t transB, double alpha,
                  const parsec_tiled_matrix_t* A, const parsec_tiled_matrix_t* B, double beta, parsec_tiled_matrix_t* C) {
    // argument checks and codes: src/zgemm_wrapper.c:657-710
    if (transA != dplasmaNoTrans && transA != dplasmaTrans && transA != dplasmaConjTrans) { fprintf(stderr, "dplasma_dgemm: illegal value of transA\n"); return -1; }
    if (transB != dplasmaNoTrans && transB != dplasmaTrans && transB != dplasmaConjTrans) { fprintf(stderr, "dplasma_dgemm: illegal value of transB\n"); return -2; }
    const bool tA = transA != dplasmaNoTrans, tB = transB != dplasmaNoTrans;
    const int Am = tA ? A->n : A->m, An = tA ? A->m : A->n, Amb = tA ? A->nb : A->mb, Anb = tA ? A->mb : A->nb;
    const int Bm = tB ? B->n : B->m, Bn = tB ? B->m : B->n, Bmb = tB ? B->nb : B->mb, Bnb = tB ? B->mb : B->nb;
    if (Amb != C->mb || Anb != Bmb || Bnb != C->nb) { fprintf(stderr, "dplasma_dgemm: tile sizes have to match\n"); return -101; }
    if (Am != C->m || An != Bm || Bn != C->n) { fprintf(stderr, "dplasma_dgemm: sizes of matrices have to match\n"); return -101; }
    const int M = C->m, N = C->n, K = An;
    if (M == 0 || N == 0 || ((alpha == 0.0 || K == 0) && beta == 1.0)) return 0;  // :716-719
    if (!parsec) return -100;
    if (A->mtype != PARSEC_MATRIX_DOUBLE || B->mtype != PARSEC_MATRIX_DOUBLE || C->mtype != PARSEC_MATRIX_DOUBLE) {
        fprintf(stderr, "dplasma_dgemm: only real double precision is built\n");
        return -101;
    }
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    if (C->super.nodes > 1) {
        // process grid: C = alpha A B^T + beta C on square C with one square tile size (src/zgemm_NT.jdf); the
        // panels of A and B travel like the factorization's (potrf_dist.cu: gemm_nt_multi)
        if (tA || !tB || C->m != C->n || C->mb != C->nb || Amb != C->mb || Anb != C->mb || Bmb != C->mb) {
            fprintf(stderr, "dplasma_dgemm: on a process grid only C = alpha A B^T + beta C with square C and one square tile size is built\n");
            return -101;
        }
        DevMatrix gA, gB, gC;
        int grc = stage_in(parsec, A, &gA, true, dplasmaUpperLower);
        if (!grc) grc = stage_in(parsec, B, &gB, true, dplasmaUpperLower);
        if (!grc) grc = stage_in(parsec, C, &gC, beta != 0.0, dplasmaUpperLower);
        if (!grc) grc = (int)cudaEventRecord(parsec->evStart, s);
        if (!grc) grc = launch_scale_grid(parsec, s, gC, dplasmaUpperLower, beta);
        if (!grc && alpha != 0.0 && K > 0) grc = gemm_nt_multi(parsec, s, alpha, gA, gB, gC);
        if (!grc) grc = (int)cudaEventRecord(parsec->evStop, s);
        if (!grc) grc = stage_out(parsec, C, &gC, dplasmaUpperLower);
        cudaError_t ge = cudaStreamSynchronize(s);
        float gms = 0.f;
        if (!grc && cudaEventElapsedTime(&gms, parsec->evStart, parsec->evStop) == cudaSuccess) parsec->lastElapsed = gms * 1e-3;
        stage_free(&gA);
        stage_free(&gB);
        stage_free(&gC);
        if (grc) { fprintf(stderr, "dplasma_dgemm: internal error %d\n", grc); return -101; }
        return ge == cudaSuccess ? 0 : -101;
    }
    DevMatrix dA, dB, dC;
    double *At = nullptr, *Bt = nullptr;
    int rc = stage_in(parsec, A, &dA, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, B, &dB, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, C, &dC, beta != 0.0, dplasmaUpperLower);
    const int mb = C->mb, nb = C->nb, kb = Anb;
    const int mt = dC.mt, nt = dC.nt, kt = (K + kb - 1) / kb;
    const long long teAt = (long long)mb * kb, teBt = (long long)nb * kb;
    if (!rc && tA && cudaMalloc(&At, (size_t)std::max(1, mt) * teAt * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc && !tB && cudaMalloc(&Bt, (size_t)std::max(1, nt) * teBt * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc) rc = (int)cudaEventRecord(parsec->evStart, s);
    if (!rc && (alpha == 0.0 || K == 0)) {
        rc = launch_scale_grid(parsec, s, dC, dplasmaUpperLower, beta);
    } else {
        for (int k = 0; !rc && k < kt; k++) {
            const int tk = (k == kt - 1) ? K - k * kb : kb;
            // op(A)(m,k) = A(k,m)^T: transpose tile row k of A (tiles kb x mb) into At(m), mb x kb
            if (tA) rc = launch_transpose_tiles(parsec, s, dA.tile(k, 0), (long long)dA.lmt * dA.tileElems, dA.mb, At, teAt, mb, tk, mb, mt);
            // op(B)(k,n) = B(k,n): the engine takes B^T, so transpose tile row k of B (tiles kb x nb) into Bt(n), nb x kb
            if (!rc && !tB) rc = launch_transpose_tiles(parsec, s, dB.tile(k, 0), (long long)dB.lmt * dB.tileElems, dB.mb, Bt, teBt, nb, tk, nb, nt);
            const double* Aop = tA ? At : dA.tile(0, k);
            const long long sA = tA ? teAt : dA.tileElems;
            for (int n = 0; !rc && n < nt; n++) {
                const double* Bop = tB ? dB.tile(n, k) : Bt + (size_t)n * teBt;
                // C(:,n) = lbeta C(:,n) + alpha op(A)(:,k) op(B)(k,n), batched over the tile rows of C
                rc = gemm_single(parsec, s, mb, dC.colsOfTile(n), tk, alpha, Aop, mb, false, Bop, nb, false,
                                 (k == 0) ? beta : 1.0, dC.tile(0, n), mb, 0, mt, sA, 0, dC.tileElems, dC.rowsOfTile(mt - 1));
            }
        }
    }
    if (!rc) rc = (int)cudaEventRecord(parsec->evStop, s);
    if (!rc) rc = stage_out(parsec, C, &dC, dplasmaUpperLower);
    cudaError_t e = cudaStreamSynchronize(s);
    float ms = 0.f;
    if (!rc && cudaEventElapsedTime(&ms, parsec->evStart, parsec->evStop) == cudaSuccess) parsec->lastElapsed = ms * 1e-3;
    if (At) cudaFree(At);
    if (Bt) cudaFree(Bt);
    stage_free(&dA);
    stage_free(&dB);
    stage_free(&dC);
    if (rc) { fprintf(stderr, "dplasma_dgemm: internal error %d\n", rc); return -101; }
    return e == cudaSuccess ? 0 : -101;
}

int dplasma_dsyrk(parsec_context_t* parsec, dplasma_enum_t uplo, dplasma_enum_t trans, double alpha,
                  const parsec_tiled_matrix_t* A, double beta, parsec_tiled_matrix_t* C) {
    // argument checks and codes: src/zsyrk_wrapper.c:229-245
    if (uplo != dplasmaLower && uplo != dplasmaUpper) { fprintf(stderr, "dplasma_dsyrk: illegal value of uplo\n"); return -1; }
    if (trans != dplasmaTrans && trans != dplasmaNoTrans && trans != dplasmaConjTrans) { fprintf(stderr, "dplasma_dsyrk: illegal value of trans\n"); return -2; }
    if (C->m != C->n) { fprintf(stderr, "dplasma_dsyrk: illegal size of matrix C which should be square\n"); return -6; }
    const bool tr = trans != dplasmaNoTrans;
    if ((!tr && A->m != C->m) || (tr && A->n != C->m)) { fprintf(stderr, "dplasma_dsyrk: illegal size of matrix A\n"); return -4; }
    if (!parsec) return -100;
    if (C->super.nodes > 1 && (uplo != dplasmaLower || tr || A->nb != C->nb)) {
        fprintf(stderr, "dplasma_dsyrk: on a process grid only dplasmaLower / dplasmaNoTrans with one square tile size is built\n");
        return -101;
    }
    if (A->mtype != PARSEC_MATRIX_DOUBLE || C->mtype != PARSEC_MATRIX_DOUBLE || C->mb != C->nb || (tr ? A->nb : A->mb) != C->nb) {
        fprintf(stderr, "dplasma_dsyrk: real double and square tiles of C matching A's are required\n");
        return -101;
    }
    const int N = C->n, K = tr ? A->m : A->n, nb = C->nb, kb = tr ? A->mb : A->nb;
    if (N == 0 || ((alpha == 0.0 || K == 0) && beta == 1.0)) return 0;
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    DevMatrix dA, dC;
    double* At = nullptr;
    int rc = stage_in(parsec, A, &dA, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, C, &dC, true, uplo);
    const int nt = dC.nt, kt = (K + kb - 1) / kb;
    const long long teAt = (long long)nb * kb;
    if (!rc && tr && cudaMalloc(&At, (size_t)std::max(1, nt) * teAt * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc) rc = (int)cudaEventRecord(parsec->evStart, s);
    // Upper: C(m,n)^T = alpha op(A)(n,:) op(A)(m,:)^T + beta C(m,n)^T is the lower problem on the transposed grid
    if (!rc && uplo == dplasmaUpper) rc = launch_transpose_grid(parsec, s, dC);
    if (!rc) rc = launch_scale_grid(parsec, s, dC, dplasmaLower, beta);
    for (int k = 0; !rc && alpha != 0.0 && k < kt && C->super.nodes == 1; k++) {
        const int tk = (k == kt - 1) ? K - k * kb : kb;
        const double* panel = dA.tile(0, k);
        long long stride = dA.tileElems;
        if (tr) {  // op(A)(m,k) = A(k,m)^T: transpose tile row k of A (tiles kb x nb) into At(m), nb x kb
            rc = launch_transpose_tiles(parsec, s, dA.tile(k, 0), (long long)dA.lmt * dA.tileElems, dA.mb, At, teAt, nb, tk, nb, nt);
            panel = At;
            stride = teAt;
        }
        // C(m,n) += alpha op(A)(m,k) op(A)(n,k)^T, m >= n (zsyrk_LN.jdf:24-55 diagonal tiles, :66-112 the others)
        if (!rc) rc = trailing_update(parsec, s, dC, panel, stride, panel, stride, 0, nt, alpha, tk);
    }
    // process grid: the same per-k update, the panels of A exchanged like the factorization's (syrk_lower_multi)
    if (!rc && alpha != 0.0 && K > 0 && C->super.nodes > 1) rc = syrk_lower_multi(parsec, s, alpha, dA, dC);
    if (!rc && uplo == dplasmaUpper) rc = launch_transpose_grid(parsec, s, dC);
    if (!rc) rc = (int)cudaEventRecord(parsec->evStop, s);
    if (!rc) rc = stage_out(parsec, C, &dC, uplo);
    cudaError_t e = cudaStreamSynchronize(s);
    float ms = 0.f;
    if (!rc && cudaEventElapsedTime(&ms, parsec->evStart, parsec->evStop) == cudaSuccess) parsec->lastElapsed = ms * 1e-3;
    if (At) cudaFree(At);
    stage_free(&dA);
    stage_free(&dC);
    if (rc) { fprintf(stderr, "dplasma_dsyrk: internal error %d\n", rc); return -101; }
    return e == cudaSuccess ? 0 : -101;
}

}  // extern "C"
