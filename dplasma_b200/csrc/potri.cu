// Inversion after the factorization:
//   dplasma_dtrtri   src/ztrtri_wrapper.c:194-230   (DAG src/ztrtri_L.jdf, ztrtri_U.jdf)
//   dplasma_dlauum   src/zlauum_wrapper.c:184-214   (DAG src/zlauum_L.jdf, zlauum_U.jdf)
//   dplasma_dpotri   src/zpotri_wrapper.c:55-86     (trtri NonUnit, then lauum)
//   dplasma_dpoinv / dplasma_dpoinv_sync   src/zpoinv_wrapper.c:184-207,255-272
//   check_dpoinv     src/dplasma_zcheck.c:265-315
//
// The reference runs every tile body of these DAGs on the CPU (CORE_ztrsm / CORE_zgemm /
// CORE_ztrmm / CORE_zherk / CORE_ztrtri / CORE_zlauum; none of the four JDFs has a CUDA body).
// Here step k of each DAG is a handful of launches of the DMMA engine, every product in the engine's
// fastest form (C += A B^T, both operands M-major) after one transposition of the step's tile row:
//   trtri, step k:  W = inv(A(k,k))                            (tile solve against the identity)
//                   A(m,k) = -A(m,k) W            m > k        (one launch, batched over m)
//                   A(m,n) += A(m,k) A(k,n)       m > k > n    (ONE grouped launch over the rectangle)
//                   A(k,n) = W A(k,n)             n < k        (one launch, batched over n)
//   lauum, step k:  A(m,n) += A(k,m)^T A(k,n)     n <= m < k   (ONE grouped launch: the factorization's
//                                                               trailing-update kernel on the leading tiles)
//                   A(k,n) = tril(A(k,k))^T A(k,n)             (one launch, batched over n)
//                   A(k,k) = tril(A(k,k))^T tril(A(k,k))
// Upper runs as Lower on the transposed tile grid, like the factorization (potrf.cu).  Single process only.
#include <cmath>
#include <cstdio>
#include <algorithm>
#include "internal.h"

extern "C" int info_allreduce_max(parsec_context_s* ctx, int info);  // comm.cu

namespace dpb {

int potrf_enqueue(parsec_context_s* ctx, parsec_taskpool_s* tp);

// dst(triangle) <- src(triangle); strict: leave the diagonal of dst alone (Unit diagonal: LAPACK does not
// reference it).  The other triangle of dst is never written.
__global__ void copy_tri_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd, int n,
                                int strict) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)n * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % n), j = (int)(e / n);
        if (i < j || (strict && i == j)) continue;
        dst[i + (size_t)j * ldd] = src[i + (size_t)j * lds];
    }
}
static int launch_copy_tri(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd, int n,
                           int strict) {
    if (n <= 0) return 0;
    int blocks = (int)std::min<long long>(((long long)n * n + 255) / 256, 1184);
    copy_tri_kernel<<<blocks, 256, 0, s>>>(src, lds, dst, ldd, n, strict);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// W (ld x n) <- identity
__global__ void identity_kernel(double* __restrict__ W, int ld, int n) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)n * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % n), j = (int)(e / n);
        W[i + (size_t)j * ld] = (i == j) ? 1.0 : 0.0;
    }
}
static int launch_identity(parsec_context_s* ctx, cudaStream_t s, double* W, int ld, int n) {
    if (n <= 0) return 0;
    int blocks = (int)std::min<long long>(((long long)n * n + 255) / 256, 1184);
    identity_kernel<<<blocks, 256, 0, s>>>(W, ld, n);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// whole single-process tile grid <- identity (dplasma_zlaset(UpperLower, 0, 1), src/dplasma_zcheck.c:286)
__global__ void identity_grid_kernel(DevMatrix A) {
    const long long total = (long long)A.lmt * A.lnt * A.tileElems;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long t = e / A.tileElems, r = e % A.tileElems;
        const int il = (int)(t % A.lmt), jl = (int)(t / A.lmt);
        const int i = (int)(r % A.mb), j = (int)(r / A.mb);
        A.base[e] = (il == jl && i == j && il * A.mb + i < A.M) ? 1.0 : 0.0;
    }
}

// info <- 1-based index of the first exactly-zero diagonal entry (CORE_ztrtri's singularity test,
// src/ztrtri_L.jdf:180-186: INFO = k*mb + info), if *info is still 0.  One CTA.
__global__ void __launch_bounds__(1024) diag_zero_scan_kernel(DevMatrix A, int* info) {
    __shared__ int best[1024];
    int mine = 0x7fffffff;
    for (int g = threadIdx.x; g < A.N; g += 1024) {
        const int t = g / A.nb, r = g % A.nb;
        if (A.tile(t, t)[r + (size_t)r * A.mb] == 0.0) { mine = g; break; }
    }
    best[threadIdx.x] = mine;
    __syncthreads();
    for (int w = 512; w > 0; w >>= 1) {
        if (threadIdx.x < w) best[threadIdx.x] = min(best[threadIdx.x], best[threadIdx.x + w]);
        __syncthreads();
    }
    if (threadIdx.x == 0 && best[0] != 0x7fffffff && *info == 0) *info = best[0] + 1;
}

// dst_t = (src_t)^T for t = 0..cnt-1: src tile t (rows x cols valid, ld lds) at src + t*srcStride, dst tile t
// (cols x rows, ld ldd) at dst + t*dstStride.  One CTA per 32x32 block; blockIdx.z walks the tiles.
__global__ void __launch_bounds__(256) transpose_tiles_kernel(const double* __restrict__ src, long long srcStride, int lds,
                                                              double* __restrict__ dst, long long dstStride, int ldd,
                                                              int rows, int cols) {
    __shared__ double t[32][33];
    const double* S = src + (size_t)blockIdx.z * srcStride;
    double* D = dst + (size_t)blockIdx.z * dstStride;
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8)
        if (bi + tx < rows && bj + j < cols) t[j][tx] = S[(bi + tx) + (size_t)(bj + j) * lds];
    __syncthreads();
    for (int j = ty; j < 32; j += 8)
        if (bj + tx < cols && bi + j < rows) D[(bj + tx) + (size_t)(bi + j) * ldd] = t[tx][j];
}
int launch_transpose_tiles(parsec_context_s* ctx, cudaStream_t s, const double* src, long long srcStride, int lds,
                           double* dst, long long dstStride, int ldd, int rows, int cols, int cnt) {
    if (cnt <= 0 || rows <= 0 || cols <= 0) return 0;
    for (int t0 = 0; t0 < cnt; t0 += 65535) {
        const int c = std::min(65535, cnt - t0);
        dim3 grid((rows + 31) / 32, (cols + 31) / 32, c);
        transpose_tiles_kernel<<<grid, 256, 0, s>>>(src + (size_t)t0 * srcStride, srcStride, lds,
                                                    dst + (size_t)t0 * dstStride, dstStride, ldd, rows, cols);
        ctx->launches++;
    }
    return (int)cudaGetLastError();
}

// Every product below is brought to the engine's fastest form, C += A B^T with both operands M-major
// (the form of the factorization's trailing update), by transposing the tile row of the step once
// (an HBM-bound pass over k tiles); the transposed copy doubles as the private copy a product needs
// when it overwrites its own operand.

// A <- inv(A), A lower triangular on a single-process tile grid.
// Scratch: W, W2 one tile each, Wp nt tiles.
static int trtri_lower_1gpu(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int diag, double* W, double* W2,
                            double* Wp) {
    const int nt = A.nt, nb = A.nb, ld = A.mb;
    const long long te = A.tileElems, cs = (long long)A.lmt * te;
    int rc = ensure_winv(ctx, nb);
    if (rc) return rc;
    for (int k = 0; k < nt; k++) {
        const int tk = A.rowsOfTile(k), below = nt - k - 1;
        double* Akk = A.tile(k, k);
        double* winv = ctx->wInv + (size_t)(k & 1) * ctx->wInvSlot;
        // W = inv(tril(A(k,k))): the tile solve A(k,k) W = I (trtri_ztrtri, src/ztrtri_L.jdf:160-187)
        rc = launch_identity(ctx, s, W, ld, tk);
        if (!rc) rc = trsm_left(ctx, s, dplasmaLower, dplasmaNoTrans, diag, tk, tk, 1.0, Akk, ld, W, ld, winv);
        if (rc) return rc;
        if (below > 0) {
            // A(m,k) = -A(m,k) inv(A(k,k)) = -A(m,k) (W^T)^T  (trtri_ztrsmR, :21-58), one launch batched over m
            rc = launch_transpose(ctx, s, W, ld, W2, ld, tk, tk);
            if (rc) return rc;
            DPB_CUDA(cudaMemcpyAsync(Wp, A.tile(k + 1, k), (size_t)below * te * sizeof(double), cudaMemcpyDeviceToDevice, s));
            rc = gemm_single(ctx, s, ld, tk, tk, -1.0, Wp, ld, false, W2, ld, false, 0.0, A.tile(k + 1, k), ld, 0, below, te,
                             0, te, A.rowsOfTile(nt - 1));
            if (rc) return rc;
        }
        if (k > 0) {
            // Wp(n) = A(k,n)^T, n < k
            rc = launch_transpose_tiles(ctx, s, A.tile(k, 0), cs, ld, Wp, te, ld, tk, nb, k);
            // A(m,n) += A(m,k) A(k,n) = A(m,k) Wp(n)^T for the whole rectangle m > k > n  (trtri_zgemm, :60-106):
            // one grouped launch
            if (!rc && below > 0)
                rc = trailing_update(ctx, s, A, A.tile(0, k), te, Wp, te, 0, k, 1.0, tk, 0, 0, k + 1);
            // A(k,n) = inv(A(k,k)) A(k,n) = W Wp(n)^T  (trtri_ztrsmL, :108-142), one launch batched over n
            if (!rc) rc = gemm_single(ctx, s, tk, nb, tk, 1.0, W, ld, false, Wp, ld, false, 0.0, A.tile(k, 0), ld, 0, k, 0, te,
                                      cs, tk);
            if (rc) return rc;
        }
        rc = launch_copy_tri(ctx, s, W, ld, Akk, ld, tk, diag == dplasmaUnit);
        if (rc) return rc;
    }
    return 0;
}

// A <- L^T L on the lower triangle of a single-process tile grid
static int lauum_lower_1gpu(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, double* W, double* W2, double* Wp) {
    const int nt = A.nt, nb = A.nb, ld = A.mb;
    const long long te = A.tileElems, cs = (long long)A.lmt * te;
    int rc = 0;
    for (int k = 0; k < nt; k++) {
        const int tk = A.rowsOfTile(k);
        double* Akk = A.tile(k, k);
        // W2 = tril(A(k,k))^T (zeros below the diagonal)
        rc = launch_copy_tril(ctx, s, Akk, ld, W, ld, tk, dplasmaLower, 0);
        if (!rc) rc = launch_transpose(ctx, s, W, ld, W2, ld, tk, tk);
        if (rc) return rc;
        if (k > 0) {
            // Wp(n) = A(k,n)^T, n < k
            rc = launch_transpose_tiles(ctx, s, A.tile(k, 0), cs, ld, Wp, te, ld, tk, nb, k);
            // A(m,n) += A(k,m)^T A(k,n) = Wp(m) Wp(n)^T, n <= m < k, diagonal tiles in their lower part only
            // (lauum_zherk, src/zlauum_L.jdf:38-65, and lauum_zgemm, :66-98): the factorization's grouped
            // trailing-update launch on the leading k x k tiles
            if (!rc) {
                DevMatrix V = A;
                V.mt = V.nt = k;
                V.M = V.N = k * nb;
                rc = trailing_update(ctx, s, V, Wp, te, Wp, te, 0, k, 1.0, tk);
            }
            // A(k,n) = A(k,k)^T A(k,n) = W2 Wp(n)^T  (lauum_ztrmm, :99-130), one launch batched over n
            if (!rc) rc = gemm_single(ctx, s, tk, nb, tk, 1.0, W2, ld, false, Wp, ld, false, 0.0, A.tile(k, 0), ld, 0, k, 0,
                                      te, cs, tk);
            if (rc) return rc;
        }
        // A(k,k) = A(k,k)^T A(k,k) = W2 W2^T, lower  (lauum_zlauum, :131-155)
        rc = gemm_single(ctx, s, tk, tk, tk, 1.0, W2, ld, false, W2, ld, false, 0.0, Akk, ld, 1);
        if (rc) return rc;
    }
    return 0;
}

// local tiles of a matrix distributed over a P x Q grid <- identity
__global__ void identity_dist_kernel(DevMatrix A) {
    const long long total = (long long)A.lmt * A.lnt * A.tileElems;
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
         e += (long long)gridDim.x * blockDim.x) {
        const long long t = e / A.tileElems, r = e % A.tileElems;
        const int il = (int)(t % A.lmt), jl = (int)(t / A.lmt);
        const long long gi = (long long)(il * A.P + A.prow) * A.mb + (int)(r % A.mb);
        const long long gj = (long long)(jl * A.Q + A.pcol) * A.nb + (int)(r / A.mb);
        A.base[e] = (gi == gj && gi < A.M) ? 1.0 : 0.0;
    }
}

// Process grid.  The inverse comes from the distributed triangular solves that dpotrs uses, applied to the
// identity: op(T) X = I for trtri, and A X = I (two solves with the factor, src/zpotrs_wrapper.c:78-98) for
// potri; the uplo triangle of X then replaces the factor.  Three times the flops of trtri + lauum, but built
// only from the grid primitives the parity tests already pin; the panel-broadcast version of the two DAGs
// replaces it next.  NonUnit only; lauum alone is not built on a grid.
static int inverse_chain_grid(parsec_context_s* ctx, const char* name, int ops, int uplo, int diag,
                              parsec_tiled_matrix_t* A, int* info_out) {
    if (!(ops & OP_TRTRI) || diag != dplasmaNonUnit || A->mtype != PARSEC_MATRIX_DOUBLE) {
        fprintf(stderr, "%s: on a process grid only the real NonUnit inversion (dtrtri / dpotri / dpoinv) is built\n", name);
        return -2;
    }
    cudaSetDevice(ctx->device);
    cudaStream_t s = ctx->sPanel;
    DevMatrix dA, X;
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    ctx->evIn.clear();
    int rc = stage_in(ctx, A, &dA, true, uplo);
    if (!rc && (dA.mb != dA.nb || dA.mt != dA.nt)) rc = -3;
    if (!rc && cudaMemsetAsync(ctx->dInfo, 0, sizeof(int), s) != cudaSuccess) rc = -201;
    if (!rc) *ctx->hInfo = 0;
    if (!rc) rc = (int)cudaEventRecord(ctx->evStart, s);
    int info = 0;
    if (!rc && (ops & OP_POTRF)) {
        parsec_taskpool_s tp{};
        tp.kind = TP_POTRF;
        tp.uplo = uplo;
        tp.A = A;
        tp.dA = dA;
        tp.dA.owned = false;
        ctx->curTp = nullptr;
        ctx->earlyOut = -1;
        rc = potrf_enqueue(ctx, &tp);
        if (!rc && cudaMemcpyAsync(ctx->hInfo, ctx->dInfo, sizeof(int), cudaMemcpyDeviceToHost, s) != cudaSuccess) rc = -201;
        if (!rc && cudaStreamSynchronize(s) != cudaSuccess) rc = -202;
        if (!rc) info = info_allreduce_max(ctx, *ctx->hInfo);  // every rank takes the same branch below
    }
    if (!rc && info == 0) {
        rc = alloc_like(dA, &X);
        if (!rc) {
            identity_dist_kernel<<<1184, 256, 0, s>>>(X);
            ctx->launches++;
            rc = (int)cudaGetLastError();
        }
        if (!rc && (ops & OP_LAUUM)) {
            // Lower: L Y = I, L^T X = Y ; Upper: U^T Y = I, U X = Y
            rc = trsm_left_multi(ctx, s, uplo, uplo == dplasmaLower ? dplasmaNoTrans : dplasmaTrans, dplasmaNonUnit, 1.0, dA, X);
            if (!rc) rc = trsm_left_multi(ctx, s, uplo, uplo == dplasmaLower ? dplasmaTrans : dplasmaNoTrans, dplasmaNonUnit, 1.0, dA, X);
        } else if (!rc) {
            rc = trsm_left_multi(ctx, s, uplo, dplasmaNoTrans, dplasmaNonUnit, 1.0, dA, X);
        }
        // the uplo triangle of X replaces the factor; the other triangle of A is not touched
        const int rowEnd = std::min(dA.firstLocalRowGE(dA.mt), dA.lmt), colEnd = std::min(dA.firstLocalColGE(dA.nt), dA.lnt);
        for (int jl = 0; !rc && jl < colEnd; jl++)
            for (int il = 0; !rc && il < rowEnd; il++) {
                const int gm = il * dA.P + dA.prow, gn = jl * dA.Q + dA.pcol;
                if (gm == gn) rc = launch_copy_tril(ctx, s, X.tile(il, jl), X.mb, dA.tile(il, jl), dA.mb, dA.rowsOfTile(gm), uplo, 2);
                else if ((uplo == dplasmaLower) ? (gm > gn) : (gm < gn)) {
                    if (cudaMemcpyAsync(dA.tile(il, jl), X.tile(il, jl), (size_t)dA.tileElems * sizeof(double),
                                        cudaMemcpyDeviceToDevice, s) != cudaSuccess) rc = -201;
                }
            }
    }
    if (!rc) rc = (int)cudaEventRecord(ctx->evStop, s);
    if (!rc) rc = stage_out(ctx, A, &dA, uplo);
    cudaError_t e = cudaStreamSynchronize(s);
    float ms = 0.f;
    if (!rc && cudaEventElapsedTime(&ms, ctx->evStart, ctx->evStop) == cudaSuccess) ctx->lastElapsed = ms * 1e-3;
    stage_free(&X);
    stage_free(&dA);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    if (e != cudaSuccess) return -1000 - (int)e;
    *info_out = info;
    return 0;
}

// stage A, run the requested chain of operations on the GPU, stage the triangle back
int run_inverse_chain(parsec_context_s* ctx, const char* name, int ops, int uplo, int diag, parsec_tiled_matrix_t* A,
                      int* info_out) {
    *info_out = 0;
    if (!ctx) return -100;
    if (A->m == 0 || A->n == 0) return 0;  // LAPACK quick return
    if (A->super.nodes > 1) return inverse_chain_grid(ctx, name, ops, uplo, diag, A, info_out);
    if (A->mtype != PARSEC_MATRIX_DOUBLE && (ops & (OP_TRTRI | OP_LAUUM))) {
        fprintf(stderr, "%s: only real double precision is built\n", name);
        return -3;
    }
    cudaSetDevice(ctx->device);
    cudaStream_t s = ctx->sPanel;
    DevMatrix dA;
    double *W = nullptr, *W2 = nullptr, *Wp = nullptr;
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    ctx->evIn.clear();
    int rc = stage_in(ctx, A, &dA, true, uplo);
    if (!rc && (dA.mb != dA.nb || dA.mt != dA.nt)) rc = -3;
    const bool inv = (ops & (OP_TRTRI | OP_LAUUM)) != 0;
    if (!rc && inv && cudaMalloc(&W, (size_t)dA.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc && inv && cudaMalloc(&W2, (size_t)dA.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc && inv && cudaMalloc(&Wp, (size_t)dA.nt * dA.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc && cudaMemsetAsync(ctx->dInfo, 0, sizeof(int), s) != cudaSuccess) rc = -201;
    if (!rc) *ctx->hInfo = 0;
    if (!rc) rc = (int)cudaEventRecord(ctx->evStart, s);
    bool skip = false;
    if (!rc && (ops & OP_POTRF)) {
        parsec_taskpool_s tp{};
        tp.kind = TP_POTRF;
        tp.uplo = uplo;
        tp.A = A;
        tp.dA = dA;
        tp.dA.owned = false;
        ctx->curTp = nullptr;
        ctx->earlyOut = -1;
        rc = potrf_enqueue(ctx, &tp);
        // a failed factorization leaves the matrix as dpotrf would: nothing is inverted
        if (!rc && cudaMemcpyAsync(ctx->hInfo, ctx->dInfo, sizeof(int), cudaMemcpyDeviceToHost, s) != cudaSuccess) rc = -201;
        if (!rc && cudaStreamSynchronize(s) != cudaSuccess) rc = -202;
        if (!rc && *ctx->hInfo != 0) skip = true;
    }
    if (!rc && !skip && (ops & (OP_TRTRI | OP_LAUUM))) {
        if (uplo == dplasmaUpper) rc = launch_transpose_grid(ctx, s, dA);
        if (!rc && (ops & OP_TRTRI)) {
            if (diag == dplasmaNonUnit && !(ops & OP_POTRF)) {
                diag_zero_scan_kernel<<<1, 1024, 0, s>>>(dA, ctx->dInfo);
                ctx->launches++;
                rc = (int)cudaGetLastError();
            }
            if (!rc) rc = trtri_lower_1gpu(ctx, s, dA, diag, W, W2, Wp);
        }
        if (!rc && (ops & OP_LAUUM)) rc = lauum_lower_1gpu(ctx, s, dA, W, W2, Wp);
        if (!rc && uplo == dplasmaUpper) rc = launch_transpose_grid(ctx, s, dA);
    }
    if (!rc) rc = (int)cudaEventRecord(ctx->evStop, s);
    if (!rc) rc = stage_out(ctx, A, &dA, uplo);
    if (!rc && !skip && cudaMemcpyAsync(ctx->hInfo, ctx->dInfo, sizeof(int), cudaMemcpyDeviceToHost, s) != cudaSuccess) rc = -201;
    cudaError_t e = cudaStreamSynchronize(s);
    float ms = 0.f;
    if (!rc && cudaEventElapsedTime(&ms, ctx->evStart, ctx->evStop) == cudaSuccess) ctx->lastElapsed = ms * 1e-3;
    if (W) cudaFree(W);
    if (W2) cudaFree(W2);
    if (Wp) cudaFree(Wp);
    stage_free(&dA);
    if (rc) return rc < 0 ? rc : -1000 - rc;
    if (e != cudaSuccess) return -1000 - (int)e;
    *info_out = *ctx->hInfo;
    return 0;
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int dplasma_dtrtri(parsec_context_t* parsec, dplasma_enum_t uplo, dplasma_enum_t diag, parsec_tiled_matrix_t* A) {
    // argument checks and codes: src/ztrtri_wrapper.c:203-216
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dtrtri: illegal value of uplo\n"); return -1; }
    if (diag != dplasmaUnit && diag != dplasmaNonUnit) { fprintf(stderr, "dplasma_dtrtri: illegal value of diag\n"); return -2; }
    if (A->m != A->n) { fprintf(stderr, "dplasma_dtrtri: illegal matrix A\n"); return -6; }
    int info = 0;
    const int rc = run_inverse_chain(parsec, "dplasma_dtrtri", OP_TRTRI, uplo, diag, A, &info);
    return rc ? (rc == -100 ? -100 : -101) : info;
}

int dplasma_dlauum(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    // src/zlauum_wrapper.c:192-200
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dlauum: illegal value of uplo\n"); return -1; }
    if (A->m != A->n) { fprintf(stderr, "dplasma_dlauum: illegal matrix A\n"); return -6; }
    int info = 0;
    const int rc = run_inverse_chain(parsec, "dplasma_dlauum", OP_LAUUM, uplo, dplasmaNonUnit, A, &info);
    return rc ? (rc == -100 ? -100 : -101) : 0;
}

int dplasma_dpotri(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    // src/zpotri_wrapper.c:62-65; the two DAGs run back to back on the resident matrix
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dpotri: illegal value of uplo\n"); return -1; }
    if (A->m != A->n) { fprintf(stderr, "dplasma_dpotri: illegal matrix A\n"); return -6; }
    int info = 0;
    const int rc = run_inverse_chain(parsec, "dplasma_dpotri", OP_TRTRI | OP_LAUUM, uplo, dplasmaNonUnit, A, &info);
    return rc ? (rc == -100 ? -100 : -101) : info;
}

int dplasma_dpoinv(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    // src/zpoinv_wrapper.c:78-81 (the check lives in _New there), :184-207
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dpoinv: illegal value of uplo\n"); return -1; }
    if (A->m != A->n) { fprintf(stderr, "dplasma_dpoinv: illegal matrix A\n"); return -6; }
    int info = 0;
    const int rc = run_inverse_chain(parsec, "dplasma_dpoinv", OP_POTRF | OP_TRTRI | OP_LAUUM, uplo, dplasmaNonUnit, A, &info);
    return rc ? (rc == -100 ? -100 : -101) : info;
}

int dplasma_dpoinv_sync(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    // src/zpoinv_wrapper.c:255-272: three blocking calls in a row (the reference overwrites the factorization's
    // info with the inversion's, :267-268; here a failed factorization stops the chain and is reported)
    if (uplo != dplasmaUpper && uplo != dplasmaLower) { fprintf(stderr, "dplasma_dpoinv_sync: illegal value of uplo\n"); return -1; }
    // a process grid has no stand-alone lauum yet: the three steps run as the one chain of dplasma_dpoinv
    if (A->super.nodes > 1) return dplasma_dpoinv(parsec, uplo, A);
    int info = dplasma_dpotrf(parsec, uplo, A);
    if (info != 0) return info;
    info = dplasma_dtrtri(parsec, uplo, dplasmaNonUnit, A);
    if (info != 0) return info;
    return dplasma_dlauum(parsec, uplo, A);
}

// ||Id - Ainv A||_1 / (||A||_1 ||Ainv||_1 N eps), Ainv stored in `uplo`, A read as a general matrix
int check_dpoinv(parsec_context_t* parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t* A,
                 parsec_tiled_matrix_t* Ainv) {
    if (!parsec) return 1;
    if (A->super.nodes > 1) {
        fprintf(stderr, "check_dpoinv: not built for a process grid yet (single process only)\n");
        return 1;
    }
    if (((const parsec_matrix_block_cyclic_t*)A)->uplo != dplasmaUpperLower) {
        // dplasma_zhemm reads every tile of its B operand (src/dplasma_zcheck.c:289-291)
        fprintf(stderr, "check_dpoinv: A must be stored in full (it is the general operand of the product)\n");
        return 1;
    }
    cudaSetDevice(parsec->device);
    cudaStream_t s = parsec->sPanel;
    DevMatrix dA, dI, R;
    double* diagCopy = nullptr;
    double Anorm = 0, Inorm = 0, Rnorm = 0;
    const int N = A->n;
    const double eps = 0x1.0p-53;
    int rc = stage_in(parsec, A, &dA, true, dplasmaUpperLower);
    if (!rc) rc = stage_in(parsec, Ainv, &dI, true, uplo);
    if (!rc && (dA.mb != dA.nb || dA.mt != dA.nt || dI.mb != dA.mb || dI.mt != dA.mt)) rc = -3;
    // one norm == infinity norm for a symmetric matrix
    if (!rc) rc = inf_norm_public(parsec, dA, true, uplo, &Anorm);
    if (!rc) rc = inf_norm_public(parsec, dI, true, uplo, &Inorm);
    if (!rc) rc = alloc_like(dA, &R);
    if (!rc && cudaMalloc(&diagCopy, (size_t)dA.tileElems * sizeof(double)) != cudaSuccess) rc = -200;
    if (!rc) {
        identity_grid_kernel<<<1184, 256, 0, s>>>(R);
        parsec->launches++;
        rc = (int)cudaGetLastError();
    }
    // Id - Ainv A  (dplasma_zhemm(Left, uplo, -1, Ainv, A, 1, Id), :289-291)
    if (!rc) rc = symm_residual_1gpu(parsec, s, uplo, dI, dA, R, diagCopy);
    // ||R||_1 = ||R^T||_oo
    if (!rc) rc = launch_transpose_grid(parsec, s, R);
    if (!rc) rc = inf_norm_public(parsec, R, false, 0, &Rnorm);
    cudaStreamSynchronize(s);
    if (diagCopy) cudaFree(diagCopy);
    stage_free(&R);
    stage_free(&dA);
    stage_free(&dI);
    if (rc) {
        fprintf(stderr, "check_dpoinv: internal error %d\n", rc);
        return 1;
    }
    const double result = Rnorm / ((Anorm * Inorm) * N * eps);
    parsec->lastCheck = result;
    if (loud > 2)
        printf("  ||A||_one = %e, ||A^(-1)||_one = %e, ||I - A * A^(-1)||_one = %e, result = %e\n", Anorm, Inorm, Rnorm,
               result);
    return (std::isinf(Inorm) || std::isnan(result) || std::isinf(result) || result > 10.0) ? 1 : 0;
}

}  // extern "C"
