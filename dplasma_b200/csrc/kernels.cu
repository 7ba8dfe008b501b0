// Small kernels around the DMMA engine: the leaf Cholesky (+ inverse of the factor), the
// bit-exact matrix generators, norms and tile copies.
#include <cstdio>
#include "internal.h"
#include "leaf.cuh"

namespace dpb {

#ifndef DPB_DEBUG_CLOCKS
__device__ long long g_leaf_clk[8];
#endif

// ---------------------------------------------------------------------------------------
// Leaf of the in-tile recursion: factor an n x n (n <= 64) diagonal block A = L L^T in
// shared memory and invert L.  Replaces cusolverDnDpotrf on the innermost block
// (src/zpotrf_L.jdf:143); the inverse turns every triangular solve against this block
// (cublasDtrsm, src/zpotrf_L.jdf:236-240) into a DMMA product.
//   winv : 64 x 64 column-major (ld 64), lower triangular inverse, zero padded / identity
//          beyond n.
//   info : first non-positive pivot -> *dInfo = infoBase + j + 1 (if *dInfo was 0), the
//          reference's global index rule (src/zpotrf_L.jdf:180-182).
// ---------------------------------------------------------------------------------------
constexpr int LEAF_LD = IB + 1;
__global__ void __launch_bounds__(256) potrf_leaf_kernel(double* __restrict__ A, int lda, int n,
                                                         double* __restrict__ winv, int* dInfo, int infoBase) {
    extern __shared__ double leaf_smem[];
    double(*S)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem);
    double(*X)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + IB * LEAF_LD);
    __shared__ int failed;
    const int tid = threadIdx.x;
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < n && j < n && i >= j) v = A[i + (size_t)j * lda];
        S[i][j] = v;
        X[i][j] = 0.0;
    }
    if (tid == 0) failed = 0;
    __syncthreads();
    leaf_cholesky_64(S, n, tid, &failed, dInfo, infoBase, true);
    // write the factor back (lower part only: the strict upper part of the tile is not ours)
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        if (i < n && j < n && i >= j) A[i + (size_t)j * lda] = S[i][j];
    }
    invert_lower_64(S, X, tid);
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        winv[e] = (i >= j) ? X[i][j] : 0.0;
    }
}

// Inverse of the IB x IB diagonal block of a triangular tile (the trtri step that lets the
// solves of src/ztrsm_LLN.jdf:48 / ztrsm_LLT.jdf:48 run as DMMA products).  Upper blocks are
// loaded transposed, so winv always holds the inverse of a LOWER block: inv(U) = winv^T.
__global__ void __launch_bounds__(256) trtri_leaf_kernel(const double* __restrict__ A, int lda, int n, int uplo,
                                                         int unit, double* __restrict__ winv) {
    extern __shared__ double leaf_smem[];
    double(*S)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem);
    double(*X)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + IB * LEAF_LD);
    const int tid = threadIdx.x;
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < n && j < n && i >= j) {
            if (i == j && unit) v = 1.0;
            else v = (uplo == dplasmaLower) ? A[i + (size_t)j * lda] : A[j + (size_t)i * lda];
        }
        S[i][j] = v;
        X[i][j] = 0.0;
    }
    __syncthreads();
    invert_lower_64(S, X, tid);
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        winv[e] = (i >= j) ? X[i][j] : 0.0;
    }
}

// all IB x IB diagonal blocks of an n x n triangular tile in one launch (one CTA per block)
__global__ void __launch_bounds__(256) trtri_blocks_kernel(const double* __restrict__ T, int ldt, int n, int uplo,
                                                           int unit, double* __restrict__ winv) {
    extern __shared__ double leaf_smem[];
    double(*S)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem);
    double(*X)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + IB * LEAF_LD);
    const int j0 = blockIdx.x * IB, nn = min(IB, n - j0), tid = threadIdx.x;
    const double* A = T + j0 + (size_t)j0 * ldt;
    double* W = winv + (size_t)blockIdx.x * IB * IB;
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        double v = (i == j) ? 1.0 : 0.0;
        if (i < nn && j < nn && i >= j) {
            if (i == j && unit) v = 1.0;
            else v = (uplo == dplasmaLower) ? A[i + (size_t)j * ldt] : A[j + (size_t)i * ldt];
        }
        S[i][j] = v;
        X[i][j] = 0.0;
    }
    __syncthreads();
    invert_lower_64(S, X, tid);
    for (int e = tid; e < IB * IB; e += 256) {
        int i = e % IB, j = e / IB;
        W[e] = (i >= j) ? X[i][j] : 0.0;
    }
}
int launch_trtri_blocks(parsec_context_s* ctx, cudaStream_t s, const double* T, int ldt, int n, int uplo, int unit,
                        double* winv) {
    if (n <= 0) return 0;
    const int smem = 2 * IB * LEAF_LD * (int)sizeof(double);
    DPB_SMEM_OPT_IN(trtri_blocks_kernel, smem);
    trtri_blocks_kernel<<<(n + IB - 1) / IB, 256, smem, s>>>(T, ldt, n, uplo, unit, winv);
    ctx->launches++;
    return (int)cudaGetLastError();
}

int launch_trtri_leaf(parsec_context_s* ctx, cudaStream_t s, const double* A, int lda, int n, int uplo, int unit,
                      double* winv) {
    const int smem = 2 * IB * LEAF_LD * (int)sizeof(double);
    DPB_SMEM_OPT_IN(trtri_leaf_kernel, smem);
    trtri_leaf_kernel<<<1, 256, smem, s>>>(A, lda, n, uplo, unit, winv);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Fused left-looking column-block step of the diagonal-tile factorization (replaces
// cusolverDnDpotrf on the tile, src/zpotrf_L.jdf:143, with nb/64 launches instead of a deep
// recursion of tiny kernels).  For the 64-wide block column starting at j0, CTA b owns the
// 64-row block i0 = j0 + 64 b and does, without any inter-CTA synchronisation:
//   1. D = A(j0.., j0..) - A(j0.., 0:j0) A(j0.., 0:j0)^T      (every CTA, redundantly)
//      R = A(i0.., j0..) - A(i0.., 0:j0) A(j0.., 0:j0)^T      (its own block; DMMA, K = j0)
//   2. D = L L^T in shared memory, X = L^-1                    (every CTA, redundantly)
//   3. A(i0.., j0..) = R X^T                                   (the panel solve of its block)
// CTA 0 writes the inverse (for the panel TRSM), info, and the factored diagonal block -- the
// latter to workspace (wdiag), not in place: the other CTAs of the launch still read the
// original block.  Nothing reads a factored diagonal block before the tile is finished, so
// one copy launch at the end puts them all in place.
// ---------------------------------------------------------------------------------------
constexpr int PC_KC = 32;            // contraction chunk
constexpr int PC_LD = IB + 4;        // staged operand rows: conflict-free for the 2+2 fragment layout
constexpr int PC_SMEM = (2 * IB * (IB + 1) + IB * PC_LD + 2 * 2 * PC_KC * PC_LD) * (int)sizeof(double);

// phase clocks of the last CTA of each launch (build with -DDPB_DEBUG_CLOCKS; tools/tile_phases.py reads them)
__device__ long long g_dbg_clk[16 * 8];
#ifdef DPB_DEBUG_CLOCKS
#define DBG_CLK(ph)                                                                      \
    do {                                                                                 \
        if (tid == 0 && blockIdx.x == gridDim.x - 1 && (j0 / IB) < 16) g_dbg_clk[(j0 / IB) * 8 + (ph)] = clock64(); \
    } while (0)
#else
#define DBG_CLK(ph) do { } while (0)
#endif

__global__ void __launch_bounds__(256) potrf_colblock_kernel(double* __restrict__ A, int lda, int n, int j0,
                                                             double* __restrict__ winv, double* __restrict__ wdiag,
                                                             int* dInfo, int infoBase) {
    extern __shared__ double pc_smem[];
    double(*D)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(pc_smem);
    double* Rk = pc_smem + IB * LEAF_LD;  // R transposed: [col][row], stride PC_LD (A operand of R X^T)
    double(*X)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(pc_smem + IB * LEAF_LD + IB * PC_LD);
    double* stage = pc_smem + 2 * IB * LEAF_LD + IB * PC_LD;  // [buf][op][PC_KC][PC_LD], op 0: rows of block j, 1: block i
    __shared__ int failed;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int i0 = j0 + blockIdx.x * IB;
    const int nj = min(IB, n - j0), ni = min(IB, n - i0);
    const bool isDiag = blockIdx.x == 0;
    if (tid == 0) failed = 0;
    DBG_CLK(0);
    // source blocks, coalesced: D = lower triangle of A(j0.., j0..) with identity padding, Rk = A(i0.., j0..)^T
    for (int e = tid; e < IB * IB; e += 256) {
        const int r = e % IB, c = e / IB;
        double dv = (r == c) ? 1.0 : 0.0;
        if (r < nj && c < nj) dv = (r >= c) ? A[j0 + r + (size_t)(j0 + c) * lda] : 0.0;
        D[r][c] = dv;
        double rv = 0.0;
        if (!isDiag && r < ni && c < nj) rv = A[i0 + r + (size_t)(j0 + c) * lda];
        Rk[c * PC_LD + r] = rv;
        X[r][c] = 0.0;
    }
    __syncthreads();  // the fragment-wise "-= acc" below touches elements other threads loaded

    // ---- 1. DMMA accumulation of the two 64x64 products over K = j0 ----
    // warps 0-3: D quadrants, warps 4-7: R quadrants; 32x32 per warp
    const int which = warp >> 2, q = warp & 3;
    const int wm0 = (q & 1) * 32, wn0 = (q >> 1) * 32;
    const int g = lane >> 2, t = lane & 3;
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
    const int nChunks = (j0 + PC_KC - 1) / PC_KC;
    // 16-byte asynchronous copies when the tile allows it (even lda, 16 B aligned base, full row blocks)
    const bool async16 = ((lda & 1) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0) && (nj == IB) &&
                         (isDiag || ni == IB);
    auto load_chunk = [&](int c, int buf) {
        double* dst = stage + (size_t)buf * 2 * PC_KC * PC_LD;
        if (async16) {
            // 2 operands x PC_KC k-columns x 32 pairs of rows
            for (int e = tid; e < 2 * PC_KC * (IB / 2); e += 256) {
                const int op = e / (PC_KC * (IB / 2)), r2 = e % (IB / 2), kk = (e / (IB / 2)) % PC_KC;
                if (op && isDiag) continue;
                const int k = c * PC_KC + kk;  // j0 is a multiple of 64: every chunk is full
                const double* src = A + ((op ? i0 : j0) + 2 * r2) + (size_t)k * lda;
                const uint32_t d32 = smem_u32(dst + (op * PC_KC + kk) * PC_LD + 2 * r2);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d32), "l"(src) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        } else {
            // predicated element loads (ragged edges read as zero)
            for (int e = tid; e < 2 * PC_KC * IB; e += 256) {
                const int op = e / (PC_KC * IB), r = e % IB, kk = (e / IB) % PC_KC;
                const int k = c * PC_KC + kk, row = (op ? i0 : j0) + r;
                double v = 0.0;
                if (k < j0 && row < n && (op == 0 || !isDiag)) v = A[row + (size_t)k * lda];
                dst[(op * PC_KC + kk) * PC_LD + r] = v;
            }
        }
    };
    if (nChunks > 0) load_chunk(0, 0);
    for (int c = 0; c < nChunks; c++) {
        const int buf = c & 1;
        if (c + 1 < nChunks) load_chunk(c + 1, buf ^ 1);
        if (async16) {
            if (c + 1 < nChunks) asm volatile("cp.async.wait_group 1;" ::: "memory");
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();  // chunk c has landed for every thread
        const double* sJ = stage + (size_t)buf * 2 * PC_KC * PC_LD;
        const double* sI = sJ + PC_KC * PC_LD;
        const double* sAop = which ? sI : sJ;
        if (!(which && isDiag)) {
#pragma unroll
            for (int ks = 0; ks < PC_KC / 4; ks++) {
                double a[4], b[4];
                const double* ra = sAop + (ks * 4 + t) * PC_LD + wm0 + 2 * g;
                const double* rb = sJ + (ks * 4 + t) * PC_LD + wn0 + 2 * g;
#pragma unroll
                for (int v = 0; v < 2; v++) {
                    const double2 x = *reinterpret_cast<const double2*>(ra + 16 * v);
                    a[2 * v] = x.x; a[2 * v + 1] = x.y;
                    const double2 y = *reinterpret_cast<const double2*>(rb + 16 * v);
                    b[2 * v] = y.x; b[2 * v + 1] = y.y;
                }
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 4; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
            }
        }
        __syncthreads();  // everyone is done with buf before it is refilled (two iterations ahead)
    }
    DBG_CLK(1);
    // ---- D / R -= acc, in place in shared memory (the source blocks were loaded coalesced up front) ----
    if (!(which && isDiag)) {
#pragma unroll
        for (int j = 0; j < 4; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int col = wn0 + frag_col(2 * t + h, j);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int row = wm0 + frag_row<32>(g, i);
                    if (which) Rk[col * PC_LD + row] -= acc[i][j][h];
                    else if (row >= col && row < nj) D[row][col] -= acc[i][j][h];
                }
            }
    }
    __syncthreads();
    DBG_CLK(2);
    leaf64(D, X, nj, &failed, dInfo, infoBase + j0, isDiag);
    DBG_CLK(3);
    // ---- 3. write back ----
    if (isDiag) {
        for (int e = tid; e < IB * IB; e += 256) {
            const int i = e % IB, j = e / IB;
            wdiag[e] = D[i][j];
            winv[e] = (i >= j) ? X[i][j] : 0.0;
        }
    } else {
        // A(i0.., j0..) = R X^T : out(r,c) = sum_k Rk[k][r] Xk[k][c], Xk = X^T staged in the (free) operand buffers
        double* Xk = stage;
        for (int e = tid; e < IB * IB; e += 256) {
            const int c = e % IB, k = e / IB;
            Xk[k * PC_LD + c] = (c >= k) ? X[c][k] : 0.0;
        }
        __syncthreads();
        const int pm0 = (warp & 1) * 32, pn0 = (warp >> 1) * 16;  // 8 warps: 2 x 4, 32 x 16 each
        double o[4][2][2];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 2; j++) o[i][j][0] = o[i][j][1] = 0.0;
#pragma unroll
        for (int ks = 0; ks < IB / 4; ks++) {
            double a[4], bb[2];
            const double* ra = Rk + (ks * 4 + t) * PC_LD + pm0 + 2 * g;
            const double2 y = *reinterpret_cast<const double2*>(Xk + (ks * 4 + t) * PC_LD + pn0 + 2 * g);
            bb[0] = y.x; bb[1] = y.y;
#pragma unroll
            for (int v = 0; v < 2; v++) {
                const double2 x = *reinterpret_cast<const double2*>(ra + 16 * v);
                a[2 * v] = x.x; a[2 * v + 1] = x.y;
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 2; j++) dmma884(o[i][j][0], o[i][j][1], a[i], bb[j]);
        }
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int col = pn0 + frag_col(2 * t + h, j);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int row = pm0 + frag_row<32>(g, i);
                    if (row < ni && col < nj) A[i0 + row + (size_t)(j0 + col) * lda] = o[i][j][h];
                }
            }
    }
    DBG_CLK(4);
}

extern "C" int dplasma_b200_debug_read(long long* out, int n) {
    return (int)cudaMemcpyFromSymbol(out, g_dbg_clk, sizeof(long long) * (n < 128 ? n : 128));
}
extern "C" int dplasma_b200_debug_read_leaf(long long* out) {
    return (int)cudaMemcpyFromSymbol(out, g_leaf_clk, sizeof(long long) * 8);
}

// put the factored diagonal blocks (lower triangles) of an n x n tile in place
__global__ void __launch_bounds__(256) potrf_diag_writeback_kernel(double* __restrict__ A, int lda, int n,
                                                                   const double* __restrict__ wdiag) {
    const int j0 = blockIdx.x * IB, nj = min(IB, n - j0);
    const double* W = wdiag + (size_t)blockIdx.x * IB * IB;
    for (int e = threadIdx.x; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        if (i < nj && j < nj && i >= j) A[j0 + i + (size_t)(j0 + j) * lda] = W[e];
    }
}
int launch_potrf_diag_writeback(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, const double* wdiag) {
    potrf_diag_writeback_kernel<<<(n + IB - 1) / IB, 256, 0, s>>>(A, lda, n, wdiag);
    ctx->launches++;
    return (int)cudaGetLastError();
}

int launch_potrf_colblock(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, int j0, double* winv,
                          double* wdiag, int* dInfo, int infoBase) {
    DPB_SMEM_OPT_IN(potrf_colblock_kernel, PC_SMEM);
    const int nblk = (n - j0 + IB - 1) / IB;
    potrf_colblock_kernel<<<nblk, 256, PC_SMEM, s>>>(A, lda, n, j0, winv, wdiag, dInfo, infoBase);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Complex leaf (zpotrf): A = L L^H on an n x n (n <= 64) interleaved-complex block in shared
// memory, then X = L^-1.  Straightforward block-synchronous version (the real path has the
// fast warp-register leaf; this one favours brevity).  winv: 64 x 64 complex, ld 64.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) zpotrf_leaf_kernel(double* __restrict__ A, int lda, int n,
                                                          double* __restrict__ winv, int* dInfo, int infoBase) {
    extern __shared__ double leaf_smem[];
    double(*Sr)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem);
    double(*Si)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + IB * LEAF_LD);
    double(*Xr)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + 2 * IB * LEAF_LD);
    double(*Xi)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + 3 * IB * LEAF_LD);
    __shared__ int failed;
    const int tid = threadIdx.x;
    for (int e = tid; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        double re = (i == j) ? 1.0 : 0.0, im = 0.0;
        if (i < n && j < n && i >= j) {
            re = A[2 * (i + (size_t)j * lda)];
            im = (i == j) ? 0.0 : A[2 * (i + (size_t)j * lda) + 1];
        }
        Sr[i][j] = re; Si[i][j] = im;
        Xr[i][j] = 0.0; Xi[i][j] = 0.0;
    }
    if (tid == 0) failed = 0;
    __syncthreads();
    for (int j = 0; j < n; j++) {
        const double d = Sr[j][j];
        if (tid == 0 && !(d > 0.0) && !failed) {
            failed = 1;
            if (*dInfo == 0) *dInfo = infoBase + j + 1;
        }
        const double s = sqrt(d), inv = 1.0 / s;
        __syncthreads();
        if (tid < IB) {
            if (tid == j) { Sr[j][j] = s; Si[j][j] = 0.0; }
            else if (tid > j && tid < n) { Sr[tid][j] *= inv; Si[tid][j] *= inv; }
        }
        __syncthreads();
        const int i = tid & (IB - 1);
        const double lr = Sr[i][j], li = Si[i][j];
        for (int c = j + 1 + (tid >> 6); c < n; c += 4)
            if (i >= c && i < n) {
                const double cr = Sr[c][j], ci = Si[c][j];  // a_ic -= l_ij conj(l_cj)
                Sr[i][c] -= lr * cr + li * ci;
                Si[i][c] -= li * cr - lr * ci;
            }
        __syncthreads();
    }
    for (int e = tid; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        if (i < n && j < n && i >= j) {
            A[2 * (i + (size_t)j * lda)] = Sr[i][j];
            A[2 * (i + (size_t)j * lda) + 1] = Si[i][j];
        }
    }
    // inverse by forward substitution, one column per thread
    if (tid < IB) {
        const int c = tid;
        Xr[c][c] = 1.0 / Sr[c][c];
        for (int i = c + 1; i < IB; i++) {
            double sr = 0.0, si = 0.0;
            for (int l = c; l < i; l++) {
                sr += Sr[i][l] * Xr[l][c] - Si[i][l] * Xi[l][c];
                si += Sr[i][l] * Xi[l][c] + Si[i][l] * Xr[l][c];
            }
            const double dinv = 1.0 / Sr[i][i];
            Xr[i][c] = -sr * dinv;
            Xi[i][c] = -si * dinv;
        }
    }
    __syncthreads();
    for (int e = tid; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        winv[2 * e] = (i >= j) ? Xr[i][j] : 0.0;
        winv[2 * e + 1] = (i >= j) ? Xi[i][j] : 0.0;
    }
}
// ---------------------------------------------------------------------------------------
// Helpers of the complex solves (zpotrs): rectangular (conjugate) transposition of interleaved
// complex tiles, in-place conjugation, and the inverses of the 64x64 diagonal blocks of a
// lower-triangular complex tile (W and its plain transpose W^T).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) ztranspose_rect_kernel(const double2* __restrict__ src, int lds,
                                                              double2* __restrict__ dst, int ldd, int m, int n, int conj) {
    __shared__ double2 t[32][33];
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8)
        if (bi + tx < m && bj + j < n) t[j][tx] = src[(bi + tx) + (size_t)(bj + j) * lds];
    __syncthreads();
    for (int j = ty; j < 32; j += 8)
        if (bj + tx < n && bi + j < m) {
            double2 v = t[tx][j];
            if (conj) v.y = -v.y;
            dst[(bj + tx) + (size_t)(bi + j) * ldd] = v;
        }
}
// dst (n x m, ld ldd) = src (m x n, ld lds)^T or ^H ; pointers in doubles, dimensions in complex elements
int launch_ztranspose(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd, int m,
                      int n, int conj) {
    if (m <= 0 || n <= 0) return 0;
    dim3 grid((m + 31) / 32, (n + 31) / 32);
    ztranspose_rect_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const double2*>(src), lds,
                                                reinterpret_cast<double2*>(dst), ldd, m, n, conj);
    ctx->launches++;
    return (int)cudaGetLastError();
}
__global__ void zconj_kernel(double* p, long long ncomplex) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < ncomplex; e += (long long)gridDim.x * blockDim.x)
        p[2 * e + 1] = -p[2 * e + 1];
}
int launch_zconj(parsec_context_s* ctx, cudaStream_t s, double* p, long long ncomplex) {
    if (ncomplex <= 0) return 0;
    long long blocks = (ncomplex + 255) / 256;
    if (blocks > 1184) blocks = 1184;
    zconj_kernel<<<(int)blocks, 256, 0, s>>>(p, ncomplex);
    ctx->launches++;
    return (int)cudaGetLastError();
}
// one CTA per 64x64 diagonal block of the lower-triangular n x n complex tile T:
// winv[b] = inverse (64x64 complex, ld 64), winvT[b] = its plain transpose
__global__ void __launch_bounds__(256) ztrtri_blocks_kernel(const double* __restrict__ T, int ldt, int n,
                                                            double* __restrict__ winv, double* __restrict__ winvT) {
    extern __shared__ double leaf_smem[];
    double(*Sr)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem);
    double(*Si)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + IB * LEAF_LD);
    double(*Xr)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + 2 * IB * LEAF_LD);
    double(*Xi)[LEAF_LD] = reinterpret_cast<double(*)[LEAF_LD]>(leaf_smem + 3 * IB * LEAF_LD);
    const int j0 = blockIdx.x * IB, nn = min(IB, n - j0), tid = threadIdx.x;
    const double* A = T + 2 * (j0 + (size_t)j0 * ldt);
    for (int e = tid; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        double re = (i == j) ? 1.0 : 0.0, im = 0.0;
        if (i < nn && j < nn && i >= j) { re = A[2 * (i + (size_t)j * ldt)]; im = A[2 * (i + (size_t)j * ldt) + 1]; }
        Sr[i][j] = re; Si[i][j] = im; Xr[i][j] = 0.0; Xi[i][j] = 0.0;
    }
    __syncthreads();
    if (tid < IB) {
        const int c = tid;
        // 1 / L_cc (complex in general; real for a Cholesky factor)
        {
            const double dr = Sr[c][c], di = Si[c][c], den = dr * dr + di * di;
            Xr[c][c] = dr / den; Xi[c][c] = -di / den;
        }
        for (int i = c + 1; i < IB; i++) {
            double sr = 0.0, si = 0.0;
            for (int l = c; l < i; l++) {
                sr += Sr[i][l] * Xr[l][c] - Si[i][l] * Xi[l][c];
                si += Sr[i][l] * Xi[l][c] + Si[i][l] * Xr[l][c];
            }
            const double dr = Sr[i][i], di = Si[i][i], den = dr * dr + di * di;
            // x = -(s) / L_ii
            Xr[i][c] = -(sr * dr + si * di) / den;
            Xi[i][c] = -(si * dr - sr * di) / den;
        }
    }
    __syncthreads();
    double* W = winv + (size_t)blockIdx.x * IB * IB * 2;
    double* Wt = winvT + (size_t)blockIdx.x * IB * IB * 2;
    for (int e = tid; e < IB * IB; e += 256) {
        const int i = e % IB, j = e / IB;
        const double re = (i >= j) ? Xr[i][j] : 0.0, im = (i >= j) ? Xi[i][j] : 0.0;
        W[2 * e] = re; W[2 * e + 1] = im;
        Wt[2 * (j + (size_t)i * IB)] = re; Wt[2 * (j + (size_t)i * IB) + 1] = im;
    }
}
int launch_ztrtri_blocks(parsec_context_s* ctx, cudaStream_t s, const double* T, int ldt, int n, double* winv,
                         double* winvT) {
    if (n <= 0) return 0;
    const int smem = 4 * IB * LEAF_LD * (int)sizeof(double);
    DPB_SMEM_OPT_IN(ztrtri_blocks_kernel, smem);
    ztrtri_blocks_kernel<<<(n + IB - 1) / IB, 256, smem, s>>>(T, ldt, n, winv, winvT);
    ctx->launches++;
    return (int)cudaGetLastError();
}

int launch_zpotrf_leaf(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                       int infoBase) {
    const int smem = 4 * IB * LEAF_LD * (int)sizeof(double);
    DPB_SMEM_OPT_IN(zpotrf_leaf_kernel, smem);
    zpotrf_leaf_kernel<<<1, 256, smem, s>>>(A, lda, n, winv, dInfo, infoBase);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Fused panel solve  B <- B L^-T  for all tiles below the diagonal tile (cublasDtrsm Right /
// Lower / Trans at src/zpotrf_L.jdf:236-240), ONE launch per step.  Rows are independent, so a
// CTA owns a 64-row strip of one tile and walks the nb/64 block columns by itself:
//     T   = B(strip, j) - sum_{l<j} X(strip, l) L(j, l)^T        (DMMA, K = 64 j)
//     X_j = T inv(L_jj)^T                                         (DMMA, K = 64; inverse from the leaf)
// X_j is written in place and read back (L2) as an operand of the later block columns.
// ---------------------------------------------------------------------------------------
constexpr int TP_SMEM = (2 * IB * PC_LD + 2 * 2 * PC_KC * PC_LD) * (int)sizeof(double);

__global__ void __launch_bounds__(256) trsm_panel_fused_kernel(const double* __restrict__ L, int ldl, int n,
                                                               const double* __restrict__ winv, double* B, int ldb,
                                                               long long tileStride, int rowsPerTile, int lastTileRows,
                                                               int ntiles) {
    extern __shared__ double tp_smem[];
    double* Tk = tp_smem;                 // [k][row], stride PC_LD : T transposed = A operand of the inverse product
    double* Wk = Tk + IB * PC_LD;         // [k][col], stride PC_LD : inv(L_jj)^T as B operand
    double* stage = Wk + IB * PC_LD;      // [buf][op][PC_KC][PC_LD]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int spt = (rowsPerTile + IB - 1) / IB;
    const int b = blockIdx.x / spt, r0 = (blockIdx.x - b * spt) * IB;
    const int rows = (b == ntiles - 1) ? lastTileRows : rowsPerTile;
    const int nr = min(IB, rows - r0);
    if (nr <= 0) return;
    double* Bs = B + (size_t)b * tileStride + r0;
    // 8 warps: 2 (rows) x 4 (cols), 32 x 16 each
    const int wm0 = (warp & 1) * 32, wn0 = (warp >> 1) * 16;
    const int g = lane >> 2, t = lane & 3;
    const bool aligned = ((ldb & 1) == 0) && ((ldl & 1) == 0) && ((reinterpret_cast<uintptr_t>(B) & 15) == 0) &&
                         ((reinterpret_cast<uintptr_t>(L) & 15) == 0) && ((tileStride & 1) == 0) && (nr == IB);
    const int nblk = (n + IB - 1) / IB;
    for (int jb = 0; jb < nblk; jb++) {
        const int j0 = jb * IB, nj = min(IB, n - j0);
        double acc[4][2][2];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 2; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
        const int nChunks = j0 / PC_KC;  // j0 is a multiple of 64
        const bool async16 = aligned && (nj == IB);
        auto load_chunk = [&](int c, int buf) {
            double* dst = stage + (size_t)buf * 2 * PC_KC * PC_LD;
            if (async16) {
                for (int e = tid; e < 2 * PC_KC * (IB / 2); e += 256) {
                    const int op = e / (PC_KC * (IB / 2)), r2 = e % (IB / 2), kk = (e / (IB / 2)) % PC_KC;
                    const int k = c * PC_KC + kk;
                    const double* src = op ? L + (j0 + 2 * r2) + (size_t)k * ldl : Bs + 2 * r2 + (size_t)k * ldb;
                    const uint32_t d32 = smem_u32(dst + (op * PC_KC + kk) * PC_LD + 2 * r2);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d32), "l"(src) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            } else {
                for (int e = tid; e < 2 * PC_KC * IB; e += 256) {
                    const int op = e / (PC_KC * IB), r = e % IB, kk = (e / IB) % PC_KC;
                    const int k = c * PC_KC + kk;
                    double v = 0.0;
                    if (op == 0) { if (r < nr) v = __ldcg(Bs + r + (size_t)k * ldb); }
                    else if (r < nj) v = L[(j0 + r) + (size_t)k * ldl];
                    dst[(op * PC_KC + kk) * PC_LD + r] = v;
                }
            }
        };
        if (nChunks > 0) load_chunk(0, 0);
        for (int c = 0; c < nChunks; c++) {
            const int buf = c & 1;
            if (c + 1 < nChunks) load_chunk(c + 1, buf ^ 1);
            if (async16) {
                if (c + 1 < nChunks) asm volatile("cp.async.wait_group 1;" ::: "memory");
                else asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncthreads();
            const double* sX = stage + (size_t)buf * 2 * PC_KC * PC_LD;
            const double* sL = sX + PC_KC * PC_LD;
#pragma unroll
            for (int ks = 0; ks < PC_KC / 4; ks++) {
                double a[4], bb[2];
                const double* ra = sX + (ks * 4 + t) * PC_LD + wm0 + 2 * g;
                const double2 y = *reinterpret_cast<const double2*>(sL + (ks * 4 + t) * PC_LD + wn0 + 2 * g);
                bb[0] = y.x; bb[1] = y.y;
#pragma unroll
                for (int v = 0; v < 2; v++) {
                    const double2 x = *reinterpret_cast<const double2*>(ra + 16 * v);
                    a[2 * v] = x.x; a[2 * v + 1] = x.y;
                }
#pragma unroll
                for (int i = 0; i < 4; i++)
#pragma unroll
                    for (int j = 0; j < 2; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], bb[j]);
            }
            __syncthreads();
        }
        // T = B(strip, j) - acc, stored transposed; W = inv(L_jj) as [k][col]
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int col = wn0 + frag_col(2 * t + h, j);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int row = wm0 + frag_row<32>(g, i);
                    double v = 0.0;
                    if (row < nr && col < nj) v = __ldcg(Bs + row + (size_t)(j0 + col) * ldb) - acc[i][j][h];
                    Tk[col * PC_LD + row] = v;
                }
            }
        const double* Wj = winv + (size_t)jb * IB * IB;
        for (int e = tid; e < IB * IB; e += 256) Wk[(e / IB) * PC_LD + (e % IB)] = Wj[e];  // Wk[k][c] = W(c,k)
        __syncthreads();
        // X_j = T W^T : out(row, c) = sum_k Tk[k][row] Wk[k][c]
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 2; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
#pragma unroll
        for (int ks = 0; ks < IB / 4; ks++) {
            double a[4], bb[2];
            const double* ra = Tk + (ks * 4 + t) * PC_LD + wm0 + 2 * g;
            const double2 y = *reinterpret_cast<const double2*>(Wk + (ks * 4 + t) * PC_LD + wn0 + 2 * g);
            bb[0] = y.x; bb[1] = y.y;
#pragma unroll
            for (int v = 0; v < 2; v++) {
                const double2 x = *reinterpret_cast<const double2*>(ra + 16 * v);
                a[2 * v] = x.x; a[2 * v + 1] = x.y;
            }
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 2; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], bb[j]);
        }
#pragma unroll
        for (int j = 0; j < 2; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int col = wn0 + frag_col(2 * t + h, j);
#pragma unroll
                for (int i = 0; i < 4; i++) {
                    const int row = wm0 + frag_row<32>(g, i);
                    if (row < nr && col < nj) Bs[row + (size_t)(j0 + col) * ldb] = acc[i][j][h];
                }
            }
        __threadfence_block();
        __syncthreads();  // X_j visible to the whole CTA before it is read back as an operand; Tk/Wk free
    }
}

int launch_trsm_panel_fused(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, int n, const double* winv,
                            double* B, int ldb, long long tileStride, int rowsPerTile, int lastTileRows, int ntiles) {
    if (ntiles <= 0 || n <= 0) return 0;
    DPB_SMEM_OPT_IN(trsm_panel_fused_kernel, TP_SMEM);
    const int spt = (rowsPerTile + IB - 1) / IB;
    trsm_panel_fused_kernel<<<ntiles * spt, 256, TP_SMEM, s>>>(L, ldl, n, winv, B, ldb, tileStride, rowsPerTile,
                                                                lastTileRows, ntiles);
    ctx->launches++;
    return (int)cudaGetLastError();
}

int launch_potrf_leaf(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv,
                      int* dInfo, int infoBase) {
    const int smem = 2 * IB * LEAF_LD * (int)sizeof(double);
    DPB_SMEM_OPT_IN(potrf_leaf_kernel, smem);
    potrf_leaf_kernel<<<1, 256, smem, s>>>(A, lda, n, winv, dInfo, infoBase);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Matrix generators, bit-exact with CORE_dplgsy / CORE_dplrnt
// (src/cores/core_zplgsy.c:72-150, core_zplrnt.c:68-91, random.h:19-41).
// Entry (I,J) of the symmetric matrix is produced by LCG state number
// max(I,J) + min(I,J)*gM; plrnt uses I + J*gM.  The jump-ahead is evaluated per element from
// a table of the 64 squarings (a_k, c_k) -- popcount(index) affine steps.
// ---------------------------------------------------------------------------------------
__constant__ unsigned long long c_jumpA[64];
__constant__ unsigned long long c_jumpC[64];
static bool g_jumpInit[64] = {false};

static int init_jump_tables(int device) {
    if (device < 64 && g_jumpInit[device]) return 0;
    unsigned long long a[64], c[64];
    unsigned long long a_k = 6364136223846793005ULL, c_k = 1ULL;
    for (int i = 0; i < 64; i++) {
        a[i] = a_k;
        c[i] = c_k;
        c_k *= (a_k + 1);
        a_k *= a_k;
    }
    DPB_CUDA(cudaMemcpyToSymbol(c_jumpA, a, sizeof(a)));
    DPB_CUDA(cudaMemcpyToSymbol(c_jumpC, c, sizeof(c)));
    if (device < 64) g_jumpInit[device] = true;
    return 0;
}

__device__ __forceinline__ unsigned long long rnd64_jump(unsigned long long n, unsigned long long seed) {
    unsigned long long ran = seed;
    while (n) {
        const int k = __ffsll((long long)n) - 1;
        ran = c_jumpA[k] * ran + c_jumpC[k];
        n &= n - 1;
    }
    return ran;
}
__device__ __forceinline__ double rnd_value(unsigned long long ran) {
    // `0.5f - ran * RndF_Mul` evaluated in float without contraction (core_zplgsy.c:89)
    return (double)__fsub_rn(0.5f, __fmul_rn(__ull2float_rn(ran), 5.4210108624275222e-20f));
}

__global__ void __launch_bounds__(256) generate_kernel(DevMatrix A, int kind, double bump, int uplo,
                                                       unsigned long long seed) {
    const int t = blockIdx.y;  // local tile index
    const int il = t % A.lmt, jl = t / A.lmt;
    const int gm = il * A.P + A.prow, gn = jl * A.Q + A.pcol;
    if (gm >= A.mt || gn >= A.nt) return;
    if (kind == 0) {
        if (uplo == dplasmaLower && gm < gn) return;
        if (uplo == dplasmaUpper && gm > gn) return;
    }
    const int tm = A.rowsOfTile(gm), tn = A.colsOfTile(gn);
    double* T = A.tile(il, jl);
    for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < tm * tn; e += gridDim.x * blockDim.x) {
        const int i = e % tm, j = e / tm;
        const unsigned long long I = (unsigned long long)gm * A.mb + i, J = (unsigned long long)gn * A.nb + j;
        unsigned long long idx;
        if (kind == 0) idx = (I >= J) ? I + J * (unsigned long long)A.M : J + I * (unsigned long long)A.M;
        else idx = I + J * (unsigned long long)A.M;
        if (A.esz == 1) {
            double v = rnd_value(rnd64_jump(idx, seed));
            if (kind == 0 && I == J) v += bump;
            T[i + (size_t)j * A.mb] = v;
        } else {
            // two consecutive states per entry (NBELEM = 2); plghe: real diagonal, upper = conj of the
            // mirrored entry (src/cores/core_zplghe.c:95-110,140-150)
            unsigned long long ran = rnd64_jump(2ULL * idx, seed);
            double re = rnd_value(ran);
            ran = 6364136223846793005ULL * ran + 1ULL;
            double im = rnd_value(ran);
            if (kind == 0) {
                if (I == J) { re += bump; im = 0.0; }
                else if (I < J) im = -im;
            }
            T[2 * (i + (size_t)j * A.mb)] = re;
            T[2 * (i + (size_t)j * A.mb) + 1] = im;
        }
    }
}

int launch_generate(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int kind, double bump, int uplo,
                    unsigned long long seed) {
    int rc = init_jump_tables(ctx->device);
    if (rc) return rc;
    const int tiles = A.lmt * A.lnt;
    if (tiles == 0) return 0;
    int bx = (int)((A.tileElems + 256 * 8 - 1) / (256 * 8));
    if (bx < 1) bx = 1;
    dim3 grid(bx, tiles);
    generate_kernel<<<grid, 256, 0, s>>>(A, kind, bump, uplo, seed);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Norm helpers (dplasma_zlanhe / dplasma_zlange, InfNorm: src/dplasma_zcheck.c:101-103,194-196)
// rowsum[I] accumulates sum_J |A(I,J)| of the symmetric matrix whose uplo triangle is stored.
// One warp per tile column; lanes stride down the rows.
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) rowsum_kernel(DevMatrix A, int sym, int uplo, double* __restrict__ rowsum) {
    const int t = blockIdx.y;
    const int il = t % A.lmt, jl = t / A.lmt;
    const int gm = il * A.P + A.prow, gn = jl * A.Q + A.pcol;
    if (gm >= A.mt || gn >= A.nt) return;
    if (sym && uplo == dplasmaLower && gm < gn) return;
    if (sym && uplo == dplasmaUpper && gm > gn) return;
    const int tm = A.rowsOfTile(gm), tn = A.colsOfTile(gn);
    const double* T = A.tile(il, jl);
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int j = warp; j < tn; j += nwarps) {
        const long long J = (long long)gn * A.nb + j;
        double colsum = 0.0;
        for (int i = lane; i < tm; i += 32) {
            const long long I = (long long)gm * A.mb + i;
            if (sym) {
                if (uplo == dplasmaLower ? (I < J) : (I > J)) continue;  // other triangle of a diagonal tile
            }
            double v;
            if (A.esz == 1) v = fabs(T[i + (size_t)j * A.mb]);
            else {
                const double re = T[2 * (i + (size_t)j * A.mb)], im = T[2 * (i + (size_t)j * A.mb) + 1];
                v = (sym && I == J) ? fabs(re) : hypot(re, im);  // zlanhe takes |Re| on the diagonal
            }
            atomicAdd(&rowsum[I], v);
            if (sym && I != J) colsum += v;
        }
        if (sym) {
            for (int o = 16; o; o >>= 1) colsum += __shfl_xor_sync(0xffffffffu, colsum, o);
            if (lane == 0 && colsum != 0.0) atomicAdd(&rowsum[J], colsum);
            // NaN must propagate even though NaN != 0.0 is true -> handled: NaN != 0.0 is true
        }
    }
}

int launch_rowsum_sym(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int uplo, double* rowsum) {
    const int tiles = A.lmt * A.lnt;
    if (tiles == 0) return 0;
    int bx = (A.nb + 7) / 8;
    dim3 grid(bx, tiles);
    rowsum_kernel<<<grid, 256, 0, s>>>(A, 1, uplo, rowsum);
    ctx->launches++;
    return (int)cudaGetLastError();
}
int launch_rowsum_ge(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, double* rowsum) {
    const int tiles = A.lmt * A.lnt;
    if (tiles == 0) return 0;
    int bx = (A.nb + 7) / 8;
    dim3 grid(bx, tiles);
    rowsum_kernel<<<grid, 256, 0, s>>>(A, 0, dplasmaUpperLower, rowsum);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// max |v[i]| with NaN propagation; single block
__global__ void __launch_bounds__(1024) absmax_kernel(const double* __restrict__ v, long long n, double* out) {
    __shared__ double sm[32];
    __shared__ int snan[32];
    double m = 0.0;
    int isn = 0;
    for (long long i = threadIdx.x; i < n; i += blockDim.x) {
        double x = fabs(v[i]);
        if (x != x) isn = 1;
        else if (x > m) m = x;
    }
    for (int o = 16; o; o >>= 1) {
        m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        isn |= __shfl_xor_sync(0xffffffffu, isn, o);
    }
    if ((threadIdx.x & 31) == 0) { sm[threadIdx.x >> 5] = m; snan[threadIdx.x >> 5] = isn; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 0; w < (blockDim.x >> 5); w++) { m = fmax(m, sm[w]); isn |= snan[w]; }
        *out = isn ? nan("") : m;
    }
}
int launch_absmax(parsec_context_s* ctx, cudaStream_t s, const double* v, long long n, double* out) {
    absmax_kernel<<<1, 1024, 0, s>>>(v, n, out);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// dst <- triangle of src (n x n); the other triangle zeroed (symmetrize=0) or mirrored (1)
__global__ void copy_tril_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd,
                                 int n, int uplo, int symmetrize) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)n * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % n), j = (int)(e / n);
        const bool stored = (uplo == dplasmaLower) ? (i >= j) : (i <= j);
        double v;
        if (stored) v = src[i + (size_t)j * lds];
        else if (symmetrize == 2) continue;  // leave the other triangle of dst alone
        else v = symmetrize ? src[j + (size_t)i * lds] : 0.0;
        dst[i + (size_t)j * ldd] = v;
    }
}
// complex variant: triangle copy with the other triangle zeroed (symmetrize 0) or conj-mirrored (1)
__global__ void zcopy_tril_kernel(const double* __restrict__ src, int lds, double* __restrict__ dst, int ldd, int n,
                                  int uplo, int symmetrize) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)n * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % n), j = (int)(e / n);
        const bool stored = (uplo == dplasmaLower) ? (i >= j) : (i <= j);
        double re = 0.0, im = 0.0;
        if (stored) { re = src[2 * (i + (size_t)j * lds)]; im = (i == j) ? 0.0 : src[2 * (i + (size_t)j * lds) + 1]; }
        else if (symmetrize == 2) continue;
        else if (symmetrize) { re = src[2 * (j + (size_t)i * lds)]; im = -src[2 * (j + (size_t)i * lds) + 1]; }
        dst[2 * (i + (size_t)j * ldd)] = re;
        dst[2 * (i + (size_t)j * ldd) + 1] = im;
    }
}
int launch_copy_tril(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd,
                     int n, int uplo, int symmetrize) {
    if (n <= 0) return 0;
    if (ctx->cplx) {
        int zb = (int)(((long long)n * n + 255) / 256);
        if (zb > 1184) zb = 1184;
        zcopy_tril_kernel<<<zb, 256, 0, s>>>(src, lds, dst, ldd, n, uplo, symmetrize);
        ctx->launches++;
        return (int)cudaGetLastError();
    }
    int blocks = (int)(((long long)n * n + 255) / 256);
    if (blocks > 1184) blocks = 1184;
    copy_tril_kernel<<<blocks, 256, 0, s>>>(src, lds, dst, ldd, n, uplo, symmetrize);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// ---------------------------------------------------------------------------------------
// Transposition (the Upper variants run as the Lower algorithm on the transposed tiles:
// U(k,n) = L(n,k)^T, src/zpotrf_U.jdf computes exactly the transposed operations).
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) transpose_kernel(const double* __restrict__ src, int lds,
                                                        double* __restrict__ dst, int ldd, int m, int n) {
    __shared__ double t[32][33];
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8)
        if (bi + tx < m && bj + j < n) t[j][tx] = src[(bi + tx) + (size_t)(bj + j) * lds];
    __syncthreads();
    for (int j = ty; j < 32; j += 8)
        if (bj + tx < n && bi + j < m) dst[(bj + tx) + (size_t)(bi + j) * ldd] = t[tx][j];
}
// dst(upper triangle incl. diagonal) <- src(lower triangle)^T ; the strict lower part of dst is not written
__global__ void __launch_bounds__(256) transpose_upper_into_kernel(const double* __restrict__ src, int lds,
                                                                   double* __restrict__ dst, int ldd, int n) {
    __shared__ double t[32][33];
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;  // block of src
    if (bi < bj) return;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8)
        if (bi + tx < n && bj + j < n) t[j][tx] = src[(bi + tx) + (size_t)(bj + j) * lds];
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        const int r = bj + tx, c = bi + j;  // dst(r,c) = src(c,r)
        if (r < n && c < n && r <= c) dst[r + (size_t)c * ldd] = t[tx][j];
    }
}
int launch_transpose_upper_into(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst,
                                int ldd, int n) {
    if (n <= 0) return 0;
    dim3 grid((n + 31) / 32, (n + 31) / 32);
    transpose_upper_into_kernel<<<grid, 256, 0, s>>>(src, lds, dst, ldd, n);
    ctx->launches++;
    return (int)cudaGetLastError();
}
int launch_transpose(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd,
                     int m, int n) {
    if (m <= 0 || n <= 0) return 0;
    dim3 grid((m + 31) / 32, (n + 31) / 32);
    transpose_kernel<<<grid, 256, 0, s>>>(src, lds, dst, ldd, m, n);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// In-place transposition of a whole single-process tile grid (square tiles, P = Q = 1):
// tile (m,n) <-> tile (n,m)^T.  One CTA per 32x32 block pair.
__global__ void __launch_bounds__(256) transpose_grid_kernel(DevMatrix A) {
    __shared__ double ta[32][33], tb[32][33];
    // blockIdx.z walks the tile pairs m >= n (grid-stride: gridDim.z is capped at 65535); x,y the 32x32 blocks of tile (m,n)
    const long long pairs = (long long)A.mt * (A.mt + 1) / 2;
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    for (long long pz = blockIdx.z; pz < pairs; pz += gridDim.z) {
    int m = (int)((sqrt(8.0 * (double)pz + 1.0) - 1.0) * 0.5);
    while ((long long)(m + 1) * (m + 2) / 2 <= pz) m++;
    while ((long long)m * (m + 1) / 2 > pz) m--;
    const int n = (int)(pz - (long long)m * (m + 1) / 2);
    if (m == n && bi < bj) continue;
    __syncthreads();  // the shared tiles of the previous pair have been consumed
    double* Ta = A.tile(m, n);
    double* Tb = A.tile(n, m);
    const int ld = A.mb, nb = A.mb;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        if (bi + tx < nb && bj + j < nb) ta[j][tx] = Ta[(bi + tx) + (size_t)(bj + j) * ld];
        if (bj + tx < nb && bi + j < nb) tb[j][tx] = Tb[(bj + tx) + (size_t)(bi + j) * ld];
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        // Tb block (bj.., bi..) <- (Ta block)^T ; Ta block (bi.., bj..) <- (Tb block)^T
        if (bj + tx < nb && bi + j < nb) Tb[(bj + tx) + (size_t)(bi + j) * ld] = ta[tx][j];
        if (bi + tx < nb && bj + j < nb) Ta[(bi + tx) + (size_t)(bj + j) * ld] = tb[tx][j];
    }
    }
}
// complex variant: tile (m,n) <-> tile (n,m)^H (conjugate transpose), interleaved (re,im)
__global__ void __launch_bounds__(256) ztranspose_grid_kernel(DevMatrix A) {
    __shared__ double2 ta[32][33], tb[32][33];
    const long long pairs = (long long)A.mt * (A.mt + 1) / 2;
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    for (long long pz = blockIdx.z; pz < pairs; pz += gridDim.z) {
    int m = (int)((sqrt(8.0 * (double)pz + 1.0) - 1.0) * 0.5);
    while ((long long)(m + 1) * (m + 2) / 2 <= pz) m++;
    while ((long long)m * (m + 1) / 2 > pz) m--;
    const int n = (int)(pz - (long long)m * (m + 1) / 2);
    if (m == n && bi < bj) continue;
    __syncthreads();
    double2* Ta = reinterpret_cast<double2*>(A.tile(m, n));
    double2* Tb = reinterpret_cast<double2*>(A.tile(n, m));
    const int ld = A.mb, nb = A.mb;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    for (int j = ty; j < 32; j += 8) {
        if (bi + tx < nb && bj + j < nb) ta[j][tx] = Ta[(bi + tx) + (size_t)(bj + j) * ld];
        if (bj + tx < nb && bi + j < nb) tb[j][tx] = Tb[(bj + tx) + (size_t)(bi + j) * ld];
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
        if (bj + tx < nb && bi + j < nb) { double2 v = ta[tx][j]; v.y = -v.y; Tb[(bj + tx) + (size_t)(bi + j) * ld] = v; }
        if (bi + tx < nb && bj + j < nb) { double2 v = tb[tx][j]; v.y = -v.y; Ta[(bi + tx) + (size_t)(bj + j) * ld] = v; }
    }
    }
}
int launch_transpose_grid(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A) {
    if (A.P * A.Q != 1 || A.mb != A.nb || A.mt != A.nt) return -3;
    const int nblk = (A.mb + 31) / 32;
    const long long pairs = (long long)A.mt * (A.mt + 1) / 2;
    dim3 grid(nblk, nblk, (unsigned)(pairs > 65535 ? 65535 : pairs));
    if (A.esz == 2) ztranspose_grid_kernel<<<grid, 256, 0, s>>>(A);
    else transpose_grid_kernel<<<grid, 256, 0, s>>>(A);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// Transposition of the LOCAL part of a block-cyclic tile grid into a second buffer: local tile (il, jl) of S goes,
// (conjugate-)transposed, to local tile (jl, il) of D.  On the process that holds tile (m, n) of a P x Q distribution the
// result is tile (n, m) of the transposed matrix distributed over the Q x P grid whose coordinates are the swapped
// ones, so no tile leaves its GPU.  Only the tiles of the `uplo` triangle of S are moved; full mb x nb buffers.
template <typename T>
__global__ void __launch_bounds__(256) transpose_local_kernel(DevMatrix S, DevMatrix D, int upper) {
    __shared__ T t[32][33];
    const int bi = blockIdx.x * 32, bj = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int nb = S.mb;
    const long long tiles = (long long)S.lmt * S.lnt;
    for (long long z = blockIdx.z; z < tiles; z += gridDim.z) {
        const int il = (int)(z % S.lmt), jl = (int)(z / S.lmt);
        const int gm = S.prow + il * S.P, gn = S.pcol + jl * S.Q;
        if (gm >= S.mt || gn >= S.nt || (upper ? gm > gn : gm < gn)) continue;
        const T* src = reinterpret_cast<const T*>(S.tile(il, jl));
        T* dst = reinterpret_cast<T*>(D.tile(jl, il));
        __syncthreads();
        for (int j = ty; j < 32; j += 8)
            if (bi + tx < nb && bj + j < nb) t[j][tx] = src[(bi + tx) + (size_t)(bj + j) * nb];
        __syncthreads();
        for (int j = ty; j < 32; j += 8)
            if (bj + tx < nb && bi + j < nb) {
                T v = t[tx][j];
                if constexpr (sizeof(T) == 16) v.y = -v.y;
                dst[(bj + tx) + (size_t)(bi + j) * nb] = v;
            }
    }
}
int launch_transpose_local(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& S, const DevMatrix& D, int uplo) {
    if (S.mb != S.nb || S.lmt != D.lnt || S.lnt != D.lmt || S.tileElems != D.tileElems) return -3;
    const long long tiles = (long long)S.lmt * S.lnt;
    if (tiles == 0) return 0;
    const int nblk = (S.mb + 31) / 32;
    dim3 grid(nblk, nblk, (unsigned)(tiles > 65535 ? 65535 : tiles));
    if (S.esz == 2) transpose_local_kernel<double2><<<grid, 256, 0, s>>>(S, D, uplo == dplasmaUpper);
    else transpose_local_kernel<double><<<grid, 256, 0, s>>>(S, D, uplo == dplasmaUpper);
    ctx->launches++;
    return (int)cudaGetLastError();
}

// C <- beta C on an m x n column-major tile (part: 0 all, 1 lower, 2 upper); beta == 0 stores zeros
__global__ void scale_tile_kernel(double* __restrict__ C, int ldc, int m, int n, double beta, int part) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < (long long)m * n;
         e += (long long)gridDim.x * blockDim.x) {
        const int i = (int)(e % m), j = (int)(e / m);
        if ((part == 1 && i < j) || (part == 2 && i > j)) continue;
        double* c = C + i + (size_t)j * ldc;
        *c = (beta == 0.0) ? 0.0 : beta * *c;
    }
}
int launch_scale_tile(parsec_context_s* ctx, cudaStream_t s, double* C, int ldc, int m, int n, double beta, int part) {
    if (m <= 0 || n <= 0 || beta == 1.0) return 0;
    long long blocks = ((long long)m * n + 255) / 256;
    if (blocks > 1184) blocks = 1184;
    scale_tile_kernel<<<(int)blocks, 256, 0, s>>>(C, ldc, m, n, beta, part);
    ctx->launches++;
    return (int)cudaGetLastError();
}

__global__ void fill_kernel(double* p, long long n, double v) {
    for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < n;
         e += (long long)gridDim.x * blockDim.x)
        p[e] = v;
}
int launch_fill(parsec_context_s* ctx, cudaStream_t s, double* p, long long n, double v) {
    if (n <= 0) return 0;
    long long blocks = (n + 255) / 256;
    if (blocks > 1184) blocks = 1184;
    fill_kernel<<<(int)blocks, 256, 0, s>>>(p, n, v);
    ctx->launches++;
    return (int)cudaGetLastError();
}

}  // namespace dpb
