// Out-of-core tiled Cholesky: a host matrix that does not fit the device memory it may use.
//
// The reference runs this case through PaRSEC's device tile cache with LRU eviction -- its `lowmem` GPU test limits
// the cache to 21 tiles (tests/Testings.cmake:147, --mca device_cuda_memory_number_of_blocks) and the stage-in / stage-out
// hooks of src/dplasmajdf_lapack_dtt.h:131-275 move tiles back and forth.  Here the schedule itself is written for
// the limit: LEFT-looking by blocks of W tile columns.
//
//   for each block of tile columns [c0, c1):
//       H2D   the block (rows >= c0 of its columns: lower trapezoid)            -> Cbuf
//       for the finished tile columns k < c0, in batches of Kb, streamed from the host (double-buffered):
//           Cbuf(m, n) -= L(m, k) L(n, k)^T     one grouped launch per batch, contraction spans the batch (K = Kb * nb)
//       factor the block in place (tile Cholesky, panel solve, update of the block's remaining columns)
//       D2H   the block
//
// Device memory: (nt - c0) x W tiles for the block plus 2 x (nt - c0) x Kb tiles for the streamed panels; host traffic
// is the matrix once each way plus ~ nt^3 / (6 W) tile reads of finished columns, all on the copy stream underneath
// the updates.  One process, lower triangle, real double; TILE, sym-packed and LAPACK storage.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include "internal.h"

namespace dpb {

// device bytes the factorization may use for the matrix: DPB_DEVICE_BUDGET_MB (tests, and the analogue of the
// reference's device_cuda_memory_number_of_blocks), else 85 % of what is free right now
static size_t device_budget() {
    if (const char* e = getenv("DPB_DEVICE_BUDGET_MB")) return (size_t)(atof(e) * 1024.0 * 1024.0);
    size_t freeB = 0, total = 0;
    if (cudaMemGetInfo(&freeB, &total) != cudaSuccess) return 0;
    return (size_t)(0.85 * (double)freeB);
}

bool ooc_wanted(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, int uplo) {
    (void)ctx;
    if (A->super.nodes != 1 || uplo != dplasmaLower || A->mtype != PARSEC_MATRIX_DOUBLE || A->mb != A->nb) return false;
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    cudaPointerAttributes at;
    const bool onDevice = cudaPointerGetAttributes(&at, bc->mat) == cudaSuccess &&
                          (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
    cudaGetLastError();
    if (onDevice) return false;
    const size_t need = (size_t)A->mt * A->nt * A->mb * A->nb * sizeof(double);  // the staged tile grid
    return need > device_budget();
}

int potrf_out_of_core(parsec_context_s* ctx, parsec_taskpool_s* tp) {
    const parsec_tiled_matrix_t* A = tp->A;
    DevMatrix full;
    int rc = describe_matrix(A, &full);
    if (rc) return -100 + rc;
    const int nt = full.nt, nb = full.nb;
    const size_t te = (size_t)full.tileElems;
    const size_t tileBytes = te * sizeof(double);
    const size_t budget = device_budget();
    // block width W and panel batch Kb: what the two panel buffers leave of the budget goes to the block
    int Kb = 4;
    long long W = ((long long)(budget / tileBytes) - 2LL * Kb * nt) / nt;
    while (W < 1 && Kb > 1) {
        Kb /= 2;
        W = ((long long)(budget / tileBytes) - 2LL * Kb * nt) / nt;
    }
    if (W < 1) {
        fprintf(stderr, "dplasma_b200: out-of-core dpotrf needs at least %d tiles of device memory (3 tile columns), budget is %zu\n",
                3 * nt, budget / tileBytes);
        return -130;
    }
    W = std::min<long long>(W, nt);
    Kb = (int)std::min<long long>(Kb, std::max<long long>(1, nt - W));
    double* Cbuf = nullptr;
    double* Pbuf = nullptr;
    DPB_CUDA(cudaMalloc(&Cbuf, (size_t)nt * W * tileBytes));
    if (cudaMalloc(&Pbuf, 2 * (size_t)nt * Kb * tileBytes) != cudaSuccess) {
        cudaFree(Cbuf);
        return -131;
    }
    rc = ensure_winv(ctx, nb);
    cudaStream_t sp = ctx->sPanel, sc = ctx->sCopy;
    cudaEvent_t evBlockIn, evBlockDone, evBlockOut, evP[2], evPfree[2];
    for (cudaEvent_t* e : {&evBlockIn, &evBlockDone, &evBlockOut, &evP[0], &evP[1], &evPfree[0], &evPfree[1]})
        cudaEventCreateWithFlags(e, cudaEventDisableTiming);
    bool outPending = false;
    int batches = 0;
    for (int c0 = 0; c0 < nt && !rc; c0 += (int)W) {
        const int c1 = (int)std::min<long long>(c0 + W, nt), w = c1 - c0, h = nt - c0;
        // the block as a matrix of its own: local tile (il, jl) <-> global tile (c0 + il, c0 + jl)
        DevMatrix V = full;
        V.base = Cbuf;
        V.lmt = h; V.lnt = w;
        V.P = V.Q = 1; V.prow = V.pcol = c0;
        V.owned = false;
        // ---- H2D: lower trapezoid of the block.  The copy stream is in order: the write-back of the previous block
        //      (which waited for its factorization) is ahead of these copies, so the buffer is free when they run ----
        for (int j = c0; j < c1 && !rc; j++) rc = stage_column_rows(A, &full, j, j, nt, V.tile(j - c0, j - c0), 0, sc);
        if (rc) break;
        DPB_CUDA(cudaEventRecord(evBlockIn, sc));
        DPB_CUDA(cudaStreamWaitEvent(sp, evBlockIn, 0));
        // ---- left-looking updates with the finished columns, streamed in batches of Kb ----
        for (int k0 = 0; k0 < c0 && !rc; k0 += Kb, batches++) {
            const int kb = std::min(Kb, c0 - k0), par = batches & 1;
            double* buf = Pbuf + (size_t)par * nt * Kb * te;
            if (batches >= 2) DPB_CUDA(cudaStreamWaitEvent(sc, evPfree[par], 0));
            for (int q = 0; q < kb && !rc; q++)  // rows >= c0 of finished column k0 + q, one k-tile each
                rc = stage_column_rows(A, &full, k0 + q, c0, nt, buf + (size_t)q * h * te, 0, sc);
            if (rc) break;
            DPB_CUDA(cudaEventRecord(evP[par], sc));
            DPB_CUDA(cudaStreamWaitEvent(sp, evP[par], 0));
            rc = trailing_update(ctx, sp, V, buf, (long long)te, buf, (long long)te, c0, c1, -1.0, kb * nb, nb, (long long)h * te);
            DPB_CUDA(cudaEventRecord(evPfree[par], sp));
        }
        if (rc) break;
        // ---- factor the block: right-looking inside it ----
        for (int kk = 0; kk < w && !rc; kk++) {
            const int k = c0 + kk, tk = full.rowsOfTile(k);
            double* Akk = V.tile(kk, kk);
            rc = potrf_tile_fused(ctx, sp, Akk, V.mb, tk, ctx->wInv, ctx->dInfo, k * nb);
            const int below = nt - k - 1;
            if (!rc && below > 0)
                rc = trsm_panel_lower(ctx, sp, Akk, V.mb, tk, ctx->wInv, V.tile(kk + 1, kk), V.mb, V.mb, below, (long long)te,
                                      full.rowsOfTile(nt - 1));
            if (!rc && kk + 1 < w) {
                const double* panel = V.tile(0, kk);  // tile of block row il at panel + il * te
                rc = trailing_update(ctx, sp, V, panel, (long long)te, panel, (long long)te, k + 1, c1, -1.0, nb);
            }
        }
        if (rc) break;
        DPB_CUDA(cudaEventRecord(evBlockDone, sp));
        // ---- D2H: the finished block ----
        DPB_CUDA(cudaStreamWaitEvent(sc, evBlockDone, 0));
        for (int j = c0; j < c1 && !rc; j++) rc = stage_column_rows(A, &full, j, j, nt, V.tile(j - c0, j - c0), 1, sc);
        DPB_CUDA(cudaEventRecord(evBlockOut, sc));
        outPending = true;
    }
    // the panel stream is the one wait() synchronises: it has to see the last write-back
    if (outPending) cudaStreamWaitEvent(sp, evBlockOut, 0);
    cudaStreamSynchronize(sp);
    cudaStreamSynchronize(sc);
    for (cudaEvent_t e : {evBlockIn, evBlockDone, evBlockOut, evP[0], evP[1], evPfree[0], evPfree[1]}) cudaEventDestroy(e);
    cudaFree(Cbuf);
    cudaFree(Pbuf);
    return rc;
}

}  // namespace dpb
