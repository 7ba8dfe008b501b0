// Diagonal-tile Cholesky in ONE launch (replaces cusolverDnDpotrf on the tile, src/zpotrf_L.jdf:143).
//
// The tile is cut into 64 x 64 blocks.  A small persistent grid (G = 2 x block-rows CTAs, all resident at once)
// walks the block columns j = 0 .. nblk-1, right-looking, with two grid barriers per step:
//
//     leaf(j)      L(j,j) = chol(A(j,j)),  X_j = inv(L(j,j))             one CTA (the "leaf CTA" of step j)
//   -- barrier --
//     solve(j)     A(b,j) <- A(b,j) X_j^T            for b > j            64^3 DMMA products, spread over the grid
//   -- barrier --
//     update(j)    A(b,c) -= A(b,j) A(c,j)^T         for j < c <= b       spread over the grid; the block (j+1,j+1)
//                                                                          goes to the leaf CTA of step j+1, which
//                                                                          runs leaf(j+1) straight from shared memory
//                                                                          while the others finish their updates
//
// so the sequential chain of a step is  leaf -> one solve product -> one update product  instead of the
// launch + full-depth left-looking update + redundant leaf of the one-launch-per-block-column version
// (potrf_colblock_kernel, kept for the allocation-free entry point and as DPB_POTRF_TILE=col).
// Blocks travel between CTAs through global memory / L2: every inter-CTA read bypasses L1 (cp.async.cg, ld.cg).
// X_j goes to the workspace the panel TRSM of the step reads (winv, 64 x 64 column-major per block, zeros above
// the diagonal).  The strict upper triangle of the tile is never written.
#include <cstdio>
#include <cstdlib>
#include "internal.h"
#include "leaf.cuh"

namespace dpb {

namespace {

constexpr int PT_LD = IB + 4;                                 // operand rows [k][64 + pad]: conflict-free 2+2 fragments
constexpr int PT_OPER = IB * PT_LD;                           // doubles per staged operand
constexpr int PT_SMEM = (2 * IB * (IB + 1) + 2 * PT_OPER) * (int)sizeof(double);

__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Grid barrier on a monotonic counter (never reset inside a launch), split in two: a CTA that does not depend on
// the others at this point arrives and goes on.  `target` = arrivals expected so far.
__device__ __forceinline__ void grid_arrive(int* counter) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1);
    }
}
__device__ __forceinline__ void grid_wait(int* counter, int target) {
    if (threadIdx.x == 0) {
        while (ld_acquire(counter) < target) __nanosleep(32);
        __threadfence();
    }
    __syncthreads();
}
__device__ __forceinline__ void grid_barrier(int* counter, int target) {
    grid_arrive(counter);
    grid_wait(counter, target);
}

// stage a 64 x 64 column-major block (rows r0.., cols c0.. of A) as dst[k][row] (k = column): column k of the block
// is row k of the staged operand.  Bypasses L1.  Ragged blocks (nr x nc valid) are zero-filled.
__device__ __forceinline__ void stage_block(double* dst, const double* A, int lda, int r0, int c0, int nr, int nc,
                                            bool fast, int tid) {
    if (fast && nr == IB && nc == IB) {
        for (int e = tid; e < IB * (IB / 2); e += 256) {
            const int k = e / (IB / 2), r2 = e % (IB / 2);
            const double* src = A + (r0 + 2 * r2) + (size_t)(c0 + k) * lda;
            const uint32_t d32 = smem_u32(dst + k * PT_LD + 2 * r2);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d32), "l"(src) : "memory");
        }
    } else {
        for (int e = tid; e < IB * IB; e += 256) {
            const int k = e / IB, r = e % IB;
            double v = 0.0;
            if (r < nr && k < nc) v = __ldcg(A + (r0 + r) + (size_t)(c0 + k) * lda);
            dst[k * PT_LD + r] = v;
        }
    }
}
__device__ __forceinline__ void stage_commit_wait() {
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
}

// acc(r,c) = sum_k opA[k][r] * opB[k][c], 8 warps as 2 x 4, 32 x 16 outputs per warp
__device__ __forceinline__ void block_product(const double* opA, const double* opB, double (&o)[4][2][2], int warp, int lane) {
    const int pm0 = (warp & 1) * 32, pn0 = (warp >> 1) * 16;
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) o[i][j][0] = o[i][j][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < IB / 4; ks++) {
        double a[4], bb[2];
        const double* ra = opA + (ks * 4 + t) * PT_LD + pm0 + 2 * g;
        const double2 y = *reinterpret_cast<const double2*>(opB + (ks * 4 + t) * PT_LD + pn0 + 2 * g);
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
}

#ifdef DPB_DEBUG_CLOCKS
__device__ long long g_tile_clk[32 * 8];
#define TILE_CLK(step, ph) do { if (tid == 0 && (step) < 32) g_tile_clk[(step) * 8 + (ph)] = clock64(); } while (0)
#else
#define TILE_CLK(step, ph) do { } while (0)
#endif

__global__ void __launch_bounds__(256) potrf_tile_kernel(double* __restrict__ A, int lda, int n, double* __restrict__ winv,
                                                         int* counter, int counterBase, int* dInfo, int infoBase) {
    extern __shared__ double pt_smem[];
    double(*S)[IB + 1] = reinterpret_cast<double(*)[IB + 1]>(pt_smem);
    double(*X)[IB + 1] = reinterpret_cast<double(*)[IB + 1]>(pt_smem + IB * (IB + 1));
    double* opA = pt_smem + 2 * IB * (IB + 1);
    double* opB = opA + PT_OPER;
    __shared__ int failed;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int G = gridDim.x, me = blockIdx.x;
    const int nblk = (n + IB - 1) / IB;
    const bool fast = ((lda & 1) == 0) && ((reinterpret_cast<uintptr_t>(A) & 15) == 0);
    const int pm0 = (warp & 1) * 32, pn0 = (warp >> 1) * 16, g = lane >> 2, t = lane & 3;
    int arrivals = counterBase;
    if (tid == 0) failed = 0;
    auto bsize = [&](int b) { return min(IB, n - b * IB); };
    // ONE CTA runs every leaf (and the solve / update product in front of it): the leaf is long straight-line code
    // that runs several times faster from a warm instruction cache than on an SM that sees it for the first time
    auto leaf_cta = [&](int j) { (void)j; return 0; };  // == leafCta below

    // factor the diagonal block held in S (lower triangle, identity padding), publish L(j,j) and X_j
    auto leaf_and_publish = [&](int j) {
        const int nj = bsize(j);
        leaf64(S, X, nj, &failed, dInfo, infoBase + j * IB, true);
        double* Ajj = A + (size_t)j * IB + (size_t)j * IB * lda;
        double* W = winv + (size_t)j * IB * IB;
        for (int e = tid; e < IB * IB; e += 256) {
            const int r = e % IB, c = e / IB;
            if (r < nj && c < nj && r >= c) Ajj[r + (size_t)c * lda] = S[r][c];
            W[e] = (r >= c) ? X[r][c] : 0.0;
        }
    };

    // ---- step 0: the first diagonal block comes straight from memory ----
    if (me == leaf_cta(0)) {
        const int nj = bsize(0);
        for (int e = tid; e < IB * IB; e += 256) {
            const int r = e % IB, c = e / IB;
            double v = (r == c) ? 1.0 : 0.0;
            if (r < nj && c < nj) v = (r >= c) ? __ldcg(A + r + (size_t)c * lda) : 0.0;
            S[r][c] = v;
            X[r][c] = 0.0;
        }
        __syncthreads();
        leaf_and_publish(0);
    }
    const int leafCta = 0;
    for (int j = 0; j < nblk; j++) {
        arrivals += G;
        grid_barrier(counter, arrivals);  // leaf(j) published, update(j-1) complete
        if (j == nblk - 1) break;
        const int nj = bsize(j);
        if (me == leafCta) TILE_CLK(j, 0);
        if (me == leafCta) {
            // ---- the chain of the tile, on one CTA, without waiting for anybody in between:
            //      A(j+1,j) <- A(j+1,j) X_j^T ;  A(j+1,j+1) -= A(j+1,j) A(j+1,j)^T ;  leaf(j+1) ----
            const int b = j + 1, nr = bsize(b);
            // opB[k][c] = X_j(c,k), straight from the inverse this CTA has just computed (X is zero above its diagonal
            // only where the leaf wrote it: mask)
            for (int e = tid; e < IB * IB; e += 256) {
                const int c = e % IB, k2 = e / IB;
                opB[k2 * PT_LD + c] = (c >= k2) ? X[c][k2] : 0.0;
            }
            stage_block(opA, A, lda, b * IB, j * IB, nr, nj, fast, tid);
            stage_commit_wait();
            double o[4][2][2];
            block_product(opA, opB, o, warp, lane);
            double* Abj = A + (size_t)b * IB + (size_t)j * IB * lda;
            __syncthreads();  // everybody has read opA: it now receives the solved block, as the operand of the update
#pragma unroll
            for (int jj = 0; jj < 2; jj++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int col = pn0 + frag_col(2 * t + h, jj);
#pragma unroll
                    for (int ii = 0; ii < 4; ii++) {
                        const int row = pm0 + frag_row<32>(g, ii);
                        const bool ok = row < nr && col < nj;
                        if (ok) Abj[row + (size_t)col * lda] = o[ii][jj][h];
                        opA[col * PT_LD + row] = ok ? o[ii][jj][h] : 0.0;
                    }
                }
            TILE_CLK(j, 1);
            arrivals += G;
            grid_arrive(counter);  // solve(j): my block is out; the update below needs nobody else's
            TILE_CLK(j, 2);
            block_product(opA, opA, o, warp, lane);
            for (int e = tid; e < IB * IB; e += 256) {
                const int r = e % IB, q = e / IB;
                S[r][q] = (r == q) ? 1.0 : 0.0;
                X[r][q] = 0.0;
            }
            __syncthreads();
            double* Abb = A + (size_t)b * IB + (size_t)b * IB * lda;
#pragma unroll
            for (int jj = 0; jj < 2; jj++)
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    const int col = pn0 + frag_col(2 * t + h, jj);
#pragma unroll
                    for (int ii = 0; ii < 4; ii++) {
                        const int row = pm0 + frag_row<32>(g, ii);
                        if (row >= nr || col >= nr || row < col) continue;  // lower triangle of the diagonal block
                        S[row][col] = __ldcg(Abb + row + (size_t)col * lda) - o[ii][jj][h];
                    }
                }
            __syncthreads();
            TILE_CLK(j, 3);
            leaf_and_publish(j + 1);
            TILE_CLK(j, 4);
            if (G > 1) continue;  // with one CTA the rest of the step is mine as well
        }
        // ---- solve(j): row blocks b = j+2 .. (item 0, b = j+1, was the leaf CTA's), spread over the other CTAs ----
        const int others = (G > 1) ? G - 1 : 1, myIdx = (G > 1) ? me - 1 : 0;
        {
            const int items = nblk - j - 2;
            bool haveX = false;
            for (int i = myIdx; i < items; i += others) {
                const int b = j + 2 + i, nr = bsize(b);
                if (!haveX) {  // opB[k][c] = X_j(c,k): the inverse as published (column-major 64 x 64, zeros above the diagonal)
                    stage_block(opB, winv + (size_t)j * IB * IB, IB, 0, 0, IB, IB, true, tid);
                    haveX = true;
                }
                stage_block(opA, A, lda, b * IB, j * IB, nr, nj, fast, tid);
                stage_commit_wait();
                double o[4][2][2];
                block_product(opA, opB, o, warp, lane);
                double* Abj = A + (size_t)b * IB + (size_t)j * IB * lda;
#pragma unroll
                for (int jj = 0; jj < 2; jj++)
#pragma unroll
                    for (int h = 0; h < 2; h++) {
                        const int col = pn0 + frag_col(2 * t + h, jj);
#pragma unroll
                        for (int ii = 0; ii < 4; ii++) {
                            const int row = pm0 + frag_row<32>(g, ii);
                            if (row < nr && col < nj) Abj[row + (size_t)col * lda] = o[ii][jj][h];
                        }
                    }
                __syncthreads();  // operands free for the next item
            }
        }
        if (G > 1) {
            arrivals += G;
            grid_barrier(counter, arrivals);  // solve(j) complete
        } else {
            __syncthreads();
        }
        // ---- update(j): blocks (b,c), j < c <= b, except (j+1,j+1) (the leaf CTA's), column-major enumeration ----
        {
            const int m = nblk - j - 1;  // block rows / columns of the trailing part
            int item = -1;
            for (int cc = 0; cc < m; cc++) {
                for (int bb2 = cc; bb2 < m; bb2++) {
                    if (cc == 0 && bb2 == 0) continue;
                    item++;
                    if (item % others != myIdx) continue;
                    const int b = j + 1 + bb2, c = j + 1 + cc, nr = bsize(b), nc = bsize(c);
                    stage_block(opA, A, lda, b * IB, j * IB, nr, nj, fast, tid);
                    stage_block(opB, A, lda, c * IB, j * IB, nc, nj, fast, tid);
                    stage_commit_wait();
                    double o[4][2][2];
                    block_product(opA, opB, o, warp, lane);
                    double* Abc = A + (size_t)b * IB + (size_t)c * IB * lda;
                    const bool diag = (b == c);
#pragma unroll
                    for (int jj = 0; jj < 2; jj++)
#pragma unroll
                        for (int h = 0; h < 2; h++) {
                            const int col = pn0 + frag_col(2 * t + h, jj);
#pragma unroll
                            for (int ii = 0; ii < 4; ii++) {
                                const int row = pm0 + frag_row<32>(g, ii);
                                if (row >= nr || col >= nc) continue;
                                if (diag && row < col) continue;  // the strict upper triangle of the tile is not ours
                                double* p = Abc + row + (size_t)col * lda;
                                *p = __ldcg(p) - o[ii][jj][h];
                            }
                        }
                    __syncthreads();
                }
            }
        }
    }
}

}  // namespace

#ifdef DPB_DEBUG_CLOCKS
// phase clocks of the last leaf of the last tile: [7] entry, [0] after chol(A11), [1] after inv(L11) | L21,
// [2] after the A22 update, [3] after chol(A22), [4] after inv(L22) | T
extern "C" int dplasma_b200_debug_read_tile_leaf(long long* leaf) {
    return (int)cudaMemcpyFromSymbol(leaf, g_leaf_clk, sizeof(long long) * 8);
}
// per step j (rows of 8), clocks of the chain CTA: 0 after barrier 1, 1 after its solve product, 2 after its arrival at
// barrier 2, 3 after its update product, 4 after leaf(j+1) is published
extern "C" int dplasma_b200_debug_read_tile_steps(long long* steps) {
    return (int)cudaMemcpyFromSymbol(steps, g_tile_clk, sizeof(long long) * 32 * 8);
}
#endif

// One launch for the whole tile.  `counter`: a device int owned by the context, only ever incremented -- the launch is
// told where it stands (counterBase) so that no reset (and no memset node) is needed between tiles.  Consequence: the
// tile factorizations of one context must run one after the other (they do: one panel stream per context).
int launch_potrf_tile(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                      int infoBase) {
    if (n <= 0) return 0;
    DPB_SMEM_OPT_IN(potrf_tile_kernel, PT_SMEM);
    const int nblk = (n + IB - 1) / IB;
    // enough CTAs that the first update phases (nblk (nblk - 1) / 2 block products) are two or three products deep
    int G = nblk <= 1 ? 1 : (nblk <= 8 ? 2 * nblk : std::min(4 * nblk, 64));
    if (const char* e = getenv("DPB_TILE_CTAS")) G = std::max(1, atoi(e));
    if (!ctx->tileSync) {
        DPB_CUDA(cudaMalloc(&ctx->tileSync, 64));
        DPB_CUDA(cudaMemset(ctx->tileSync, 0, 64));
        ctx->tileSyncCount = 0;
    }
    if (ctx->tileSyncCount > 0x70000000) {  // wrap: start again from zero behind everything enqueued so far
        DPB_CUDA(cudaMemsetAsync(ctx->tileSync, 0, 64, s));
        ctx->tileSyncCount = 0;
    }
    potrf_tile_kernel<<<G, 256, PT_SMEM, s>>>(A, lda, n, winv, ctx->tileSync, ctx->tileSyncCount, dInfo, infoBase);
    ctx->tileSyncCount += (2 * nblk - 1) * G;  // arrivals of this launch: two barriers per step, one in the last
    ctx->launches++;
    return (int)cudaGetLastError();
}

}  // namespace dpb
