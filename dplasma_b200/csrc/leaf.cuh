// Block-level building blocks of the 64 x 64 leaf: Cholesky of a diagonal block in shared memory and the inverse
// of its factor (shared by kernels.cu and potrf_tile.cu).
#pragma once
#include "internal.h"

namespace dpb {

// X <- inverse of the lower-triangular S (both IB x IB in shared memory, X zeroed on entry):
// 16x16 diagonal blocks by substitution, then two doubling levels
// inv([L11 0; L21 L22]) = [X11 0; -X22 L21 X11, X22].
__device__ __forceinline__ void invert_lower_64(double (*S)[IB + 1], double (*X)[IB + 1], int tid) {
    if (tid < IB) {
        const int r0 = (tid >> 4) << 4, c = tid & 15;
        X[r0 + c][r0 + c] = 1.0 / S[r0 + c][r0 + c];
        for (int i = c + 1; i < 16; i++) {
            double sum = 0.0;
            for (int l = c; l < i; l++) sum += S[r0 + i][r0 + l] * X[r0 + l][r0 + c];
            X[r0 + i][r0 + c] = -sum / S[r0 + i][r0 + i];
        }
    }
    __syncthreads();
    for (int b = 16; b < IB; b <<= 1) {
        const int E = (IB / (2 * b)) * b * b;
        // T = L21 * X11, parked (transposed) in the unused upper triangle of X
        for (int e = tid; e < E; e += 256) {
            const int p = e / (b * b), r = e - p * b * b, i = r % b, c = r / b, r0 = p * 2 * b;
            double t = 0.0;
            for (int l = c; l < b; l++) t += S[r0 + b + i][r0 + l] * X[r0 + l][r0 + c];
            X[r0 + c][r0 + b + i] = t;
        }
        __syncthreads();
        // X21 = -X22 * T
        for (int e = tid; e < E; e += 256) {
            const int p = e / (b * b), r = e - p * b * b, i = r % b, c = r / b, r0 = p * 2 * b;
            double x = 0.0;
            for (int l = 0; l <= i; l++) x += X[r0 + b + i][r0 + b + l] * X[r0 + c][r0 + b + l];
            X[r0 + b + i][r0 + c] = -x;
        }
        __syncthreads();
    }
}

// In-place lower Cholesky of the n x n (n <= IB) leading block of S (shared memory), 256 threads.
// Right-looking, unblocked.  report: write info for the first non-positive pivot.
__device__ __forceinline__ void leaf_cholesky_64(double (*S)[IB + 1], int n, int tid, int* failed, int* dInfo,
                                                 int infoBase, bool report) {
    for (int j = 0; j < n; j++) {
        const double d = S[j][j];
        if (report && tid == 0 && !(d > 0.0) && !*failed) {
            *failed = 1;
            if (*dInfo == 0) *dInfo = infoBase + j + 1;
        }
        const double s = sqrt(d), inv = 1.0 / s;
        __syncthreads();
        if (tid < IB) {
            if (tid == j) S[j][j] = s;
            else if (tid > j && tid < n) S[tid][j] *= inv;
        }
        __syncthreads();
        const int i = tid & (IB - 1);
        const double lij = S[i][j];
        for (int c = j + 1 + (tid >> 6); c < n; c += 4)
            if (i >= c && i < n) S[i][c] -= lij * S[c][j];
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------
// Fast 64x64 leaf: 2x2 blocks of 32, the sequential parts (32x32 Cholesky, 32x32 triangular
// inverse) done by ONE warp in registers with shuffles -- no block-wide barrier per column.
//   S (in/out): lower triangle of the block, identity beyond n.   X (out): inverse of L.
// 256 threads.  info rule as leaf_cholesky_64.
// ---------------------------------------------------------------------------------------
#ifdef DPB_DEBUG_CLOCKS
static __device__ long long g_leaf_clk[8];  // one copy per translation unit (no relocatable device code)
#endif
#ifdef DPB_DEBUG_CLOCKS
#define LEAF_CLK(i) do { if (tid == 0 && report) g_leaf_clk[i] = clock64(); } while (0)
#else
#define LEAF_CLK(i) do { } while (0)
#endif
// warp-level: lane r owns row r of the 32x32 block at S[o..o+32)[o..o+32)
// warp-level: lane r owns row r of the 32x32 block at S[o..o+32)[o..o+32), the row lives in registers.
// The sequential chain of a column is  broadcast of the pivot -> rsqrt -> scale -> update of the next pivot ;
// each lane keeps the running diagonal entry of ITS row in a register of its own (dg), so the pivot update needs
// no shuffle, and rsqrt(double) -- accurate to an ulp or two, far inside the 1e-11 the factor is held to -- gives
// the square root and the reciprocal with one multiplication each.  All the shuffles of a column are issued before
// the updates that use them (a warp issues in order).  The code is straight-line and long: it runs fast only from
// a warm instruction cache, which is why one CTA runs every leaf of a tile (potrf_tile.cu).
// Variants measured on B200 and rejected (us per 32x32 block, warm): block in shared memory with a rolled column
// loop 19.6 (store-to-load ordering serialises the column); register row with a rolled column loop and predicated
// static indices 29.7.
__device__ __forceinline__ void warp_chol32(double (*S)[IB + 1], double* invd, double* col, int o, int lane, int nvalid,
                                            int* failed, int* dInfo, int infoBase, bool report) {
    (void)col;
    double a[32];
#pragma unroll
    for (int c = 0; c < 32; c++) a[c] = S[o + lane][o + c];
    double dg = S[o + lane][o + lane];
    int firstBad = -1;
#pragma unroll
    for (int c = 0; c < 32; c++) {
        const double d = __shfl_sync(0xffffffffu, dg, c);
        if (firstBad < 0 && c < nvalid && !(d > 0.0)) firstBad = c;
        const double inv = rsqrt(d);
        // branch-free on purpose: every lane computes, selects keep what is its own (a lane-dependent branch here
        // becomes a divergent branch with a reconvergence point in front of every shuffle)
        const double l = a[c] * inv;   // lanes > c: the column entry
        const bool below = lane > c;
        a[c] = (lane == c) ? d * inv : (below ? l : a[c]);
        dg = below ? fma(-l, l, dg) : dg;
        if (lane == c) invd[o + c] = inv;
        double lc[32];
#pragma unroll
        for (int c2 = c + 1; c2 < 32; c2++) lc[c2] = __shfl_sync(0xffffffffu, l, c2);
#pragma unroll
        for (int c2 = c + 1; c2 < 32; c2++) {
            const double u = fma(-l, lc[c2], a[c2]);
            a[c2] = (lane > c2) ? u : a[c2];
        }
    }
    if (report && lane == 0 && firstBad >= 0 && !*failed) {
        *failed = 1;
        if (*dInfo == 0) *dInfo = infoBase + o + firstBad + 1;
    }
#pragma unroll
    for (int c = 0; c < 32; c++)
        if (c <= lane) S[o + lane][o + c] = a[c];
}
// warp-level: X block (o,o) = inverse of the lower-triangular S block (o,o); lane c owns column c.
// Column-oriented substitution: as soon as x[k] is final it is folded into every later entry, so the 31 updates of
// a step are independent of each other and only one multiply-add per step sits on the dependent chain.
__device__ __forceinline__ void warp_trinv32(double (*S)[IB + 1], double (*X)[IB + 1], const double* invd, int o,
                                             int lane) {
    double x[32];
#pragma unroll
    for (int i = 0; i < 32; i++) x[i] = (i == lane) ? 1.0 : 0.0;
#pragma unroll
    for (int k = 0; k < 32; k++) {
        x[k] *= invd[o + k];
#pragma unroll
        for (int i = k + 1; i < 32; i++) x[i] = fma(-S[o + i][o + k], x[k], x[i]);
    }
#pragma unroll
    for (int i = 0; i < 32; i++) X[o + i][o + lane] = (i >= lane) ? x[i] : 0.0;
}

// 32x32 Cholesky by the whole CTA (256 threads), left-looking by panels of 8 columns, block in shared memory:
//   (1) all threads: panel(rows >= c0, 8 columns) -= B(rows, 0:c0) B(c0:c0+8, 0:c0)^T     one output per thread
//   (2) warp 0, lane = row c0 + lane, its 8 panel entries in registers: the 8 columns in sequence
//       (pivot shuffle -> rsqrt -> scale -> own-diagonal update on the chain; 7 shuffles + multiply-adds beside it)
// The panel loop is rolled and its body is a few hundred instructions: the sequential part is 8 columns long,
// everything else is parallel multiply-adds from shared memory.
__device__ __forceinline__ void block_chol32(double (*S)[IB + 1], double* invd, int o, int tid, int nvalid, int* failed,
                                             int* dInfo, int infoBase, bool report) {
    const int lane = tid & 31, warp = tid >> 5;
    int firstBad = -1;
#pragma unroll 1
    for (int c0 = 0; c0 < 32; c0 += 8) {
        if (c0 > 0) {
            const int c = c0 + (tid & 7), r = c0 + (tid >> 3);
            if (r < 32 && r >= c) {
                double acc0 = 0.0, acc1 = 0.0;
#pragma unroll 4
                for (int k = 0; k < c0; k += 2) {
                    acc0 = fma(S[o + r][o + k], S[o + c][o + k], acc0);
                    acc1 = fma(S[o + r][o + k + 1], S[o + c][o + k + 1], acc1);
                }
                S[o + r][o + c] -= acc0 + acc1;
            }
            __syncthreads();
        }
        if (warp == 0) {
            const int r = c0 + lane;          // my row; rows beyond the block idle with zeros
            const bool live = r < 32;
            double row[8];
#pragma unroll
            for (int j = 0; j < 8; j++) row[j] = (live && r >= c0 + j) ? S[o + r][o + c0 + j] : 0.0;
            double dg = live ? S[o + r][o + r] : 1.0;   // running diagonal entry of my row (meaningful for lanes < 8 here)
#pragma unroll
            for (int j = 0; j < 8; j++) {
                const double d = __shfl_sync(0xffffffffu, (lane == j) ? row[j] : dg, j);   // pivot of column c0+j lives in lane j
                if (firstBad < 0 && c0 + j < nvalid && !(d > 0.0)) firstBad = c0 + j;
                const double inv = rsqrt(d);
                const double l = row[j] * inv;
                const bool below = lane > j;
                row[j] = (lane == j) ? d * inv : (below ? l : row[j]);
                if (lane == j) invd[o + c0 + j] = inv;
#pragma unroll
                for (int j2 = j + 1; j2 < 8; j2++) {
                    const double lj2 = __shfl_sync(0xffffffffu, l, j2);
                    const double u = fma(-l, lj2, row[j2]);
                    row[j2] = (lane >= j2) ? u : row[j2];
                }
            }
#pragma unroll
            for (int j = 0; j < 8; j++)
                if (live && r >= c0 + j) S[o + r][o + c0 + j] = row[j];
        }
        __syncthreads();
    }
    if (report && tid == 0 && firstBad >= 0 && !*failed) {
        *failed = 1;
        if (*dInfo == 0) *dInfo = infoBase + o + firstBad + 1;
    }
}

__device__ __forceinline__ void leaf_fast_64(double (*S)[IB + 1], double (*X)[IB + 1], int n, int tid, int* failed,
                                             int* dInfo, int infoBase, bool report) {
    const int warp = tid >> 5, lane = tid & 31;
    __shared__ double invd[IB];  // reciprocals of the diagonal of L
    LEAF_CLK(7);
    // 1. L11 = chol(A11)
    block_chol32(S, invd, 0, tid, min(n, 32), failed, dInfo, infoBase, report);
    LEAF_CLK(0);
    // 2. warp 0: X11 = inv(L11) ; warp 1: L21 = A21 L11^-T (one row per lane, column-oriented forward substitution)
    if (warp == 0) warp_trinv32(S, X, invd, 0, lane);
    else if (warp == 1) {
        double r[32];
#pragma unroll
        for (int c = 0; c < 32; c++) r[c] = S[32 + lane][c];
#pragma unroll
        for (int k = 0; k < 32; k++) {
            r[k] *= invd[k];
#pragma unroll
            for (int c = k + 1; c < 32; c++) r[c] = fma(-r[k], S[c][k], r[c]);
        }
#pragma unroll
        for (int c = 0; c < 32; c++) S[32 + lane][c] = r[c];
    }
    __syncthreads();
    LEAF_CLK(1);
    // 3. A22 -= L21 L21^T (lower part), all threads: 4 outputs each, accumulated side by side
    {
        const int i = tid & 31, c0 = tid >> 5;  // outputs (i, c0 + 8 q), q = 0..3
        double sum[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
        for (int k = 0; k < 32; k++) {
            const double li = S[32 + i][k];
#pragma unroll
            for (int q = 0; q < 4; q++) sum[q] = fma(li, S[32 + c0 + 8 * q][k], sum[q]);
        }
#pragma unroll
        for (int q = 0; q < 4; q++)
            if (i >= c0 + 8 * q) S[32 + i][32 + c0 + 8 * q] -= sum[q];
    }
    __syncthreads();
    LEAF_CLK(2);
    // 4. L22 = chol(A22)
    block_chol32(S, invd, 32, tid, max(0, min(n - 32, 32)), failed, dInfo, infoBase, report);
    LEAF_CLK(3);
    // 5. warp 0: X22 = inv(L22) ; warps 1-4: T = L21 X11 parked in X's upper-right block (transposed), 8 outputs
    //    per thread side by side
    if (warp == 0) warp_trinv32(S, X, invd, 32, lane);
    else if (warp <= 4) {
        const int i = lane, c0 = warp - 1;  // outputs (i, c0 + 4 q), q = 0..7
        double tq[8];
#pragma unroll
        for (int q = 0; q < 8; q++) tq[q] = 0.0;
#pragma unroll 4
        for (int k = 0; k < 32; k++) {
            const double li = S[32 + i][k];
#pragma unroll
            for (int q = 0; q < 8; q++) tq[q] = fma(li, X[k][c0 + 4 * q], tq[q]);  // X11 is zero above its diagonal
        }
#pragma unroll
        for (int q = 0; q < 8; q++) X[c0 + 4 * q][32 + i] = tq[q];
    }
    __syncthreads();
    LEAF_CLK(4);
    // 6. X21 = -X22 T, 4 outputs per thread side by side
    {
        const int i = tid & 31, c0 = tid >> 5;
        double xs[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll 8
        for (int k = 0; k < 32; k++) {
            const double xi = X[32 + i][32 + k];  // X22 is zero above its diagonal
#pragma unroll
            for (int q = 0; q < 4; q++) xs[q] = fma(xi, X[c0 + 8 * q][32 + k], xs[q]);
        }
#pragma unroll
        for (int q = 0; q < 4; q++) X[32 + i][c0 + 8 * q] = -xs[q];
    }
    __syncthreads();
}

// The leaf keeps a 32-entry row of the block per lane in registers.  Inlined into a kernel body with block products
// and staging loops around it, the register allocator ran out and moved that array to local memory -- a memory
// round trip on every step of the dependent chain.  As a function of its own the leaf gets a clean register file,
// for the price of one call.
static __device__ __noinline__ void leaf64(double (*S)[IB + 1], double (*X)[IB + 1], int n, int* failed, int* dInfo,
                                           int infoBase, bool report) {
    leaf_fast_64(S, X, n, (int)threadIdx.x, failed, dInfo, infoBase, report);
}

}  // namespace dpb
