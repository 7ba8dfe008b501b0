// Tiled Cholesky driver: the static, look-ahead schedule that replaces the PaRSEC task graph
// of src/zpotrf_L.jdf (task classes potrf_zpotrf / potrf_ztrsm / potrf_zherk / potrf_zgemm,
// flows :109-114,209-217,321-325,423-428) and the public wrappers of src/zpotrf_wrapper.c.
//
// Schedule per step k (lower):
//   panel stream (high priority):  POTRF(k,k) -> TRSM(k+1..,k) -> update of tile column k+1
//   update stream:                 update of tile columns k+2.. with panel k  (one grouped launch)
// so that the panel of step k+1 is factored underneath the bulk update of step k, the same
// critical-path-first order the reference expresses with task priorities (:116,219,327,430).
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include "internal.h"

namespace dpb {

static cudaEvent_t get_event(parsec_context_s* ctx, size_t idx) {
    while (ctx->evPool.size() <= idx) {
        cudaEvent_t e;
        cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        ctx->evPool.push_back(e);
    }
    return ctx->evPool[idx];
}

// C(gm,gn) += alpha * rowPanel[il] * colPanel[jl]^T for the local tiles with
// gnFirst <= gn < gnEnd and gm >= gn; diagonal tiles only in their lower part (SYRK).
int trailing_update(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& C, const double* rowPanel,
                    long long rowStride, const double* colPanel, long long colStride, int gnFirst, int gnEnd,
                    double alpha, int K, int kTile, long long kTileStride, int rowFirstFixed, int jlMin, int jlMax) {
    // [jlMin, jlMax): additional clamp on the LOCAL tile columns (columns of a host matrix that have landed)
    const int jl0 = std::max(C.firstLocalColGE(gnFirst), jlMin), jl1 = std::min(std::min(C.firstLocalColGE(gnEnd), C.lnt), jlMax);
    const int rowEnd = std::min(C.firstLocalRowGE(C.mt), C.lmt);
    for (int c0 = jl0; c0 < jl1; c0 += kMaxCols) {
        const int ncols = std::min(kMaxCols, jl1 - c0);
        GemmDesc d;
        d.mode = GEMM_TRAILING;
        d.alpha = alpha; d.beta = 1.0; d.K = K;
        d.a_kmajor = d.b_kmajor = 0;
        d.preload = 1;
        d.kTile = 1 << 30; d.aKTileStride = d.bKTileStride = 0;
        if (kTile > 0) {  // the contraction runs over several tiles of a tile row (left-looking update)
            d.kTile = kTile;
            d.aKTileStride = d.bKTileStride = kTileStride;
        }
        d.Cbase = C.base; d.rowPanel = rowPanel; d.colPanel = colPanel;
        d.rowPanelStride = rowStride; d.colPanelStride = colStride; d.tileElems = C.tileElems;
        d.tileLdT = C.lmt; d.nb = C.nb;
        d.ncols = ncols; d.colFirst = c0;
        d.P = C.P; d.Q = C.Q; d.prow = C.prow; d.pcol = C.pcol;
        d.lastM = C.rowsOfTile(C.mt - 1); d.lastN = C.colsOfTile(C.nt - 1);
        d.mtGlobal = C.mt; d.ntGlobal = C.nt;
        d.rowEnd = rowEnd;
        d.full = ctx->gemmFull;
        if (ctx->cplx) d.fast = (K % 8 == 0) ? 1 : 0;
        else d.fast = ((C.nb % 2 == 0) && (kTile % 16 == 0) && (kTileStride % 2 == 0) && ((reinterpret_cast<uintptr_t>(rowPanel) & 15) == 0) &&
                  ((reinterpret_cast<uintptr_t>(colPanel) & 15) == 0) && (rowStride % 2 == 0) &&
                  (colStride % 2 == 0))
                     ? 1 : 0;
        long long tiles = 0;
        for (int c = 0; c < ncols; c++) {
            const int gn = (c0 + c) * C.Q + C.pcol;
            // rowFirstFixed >= 0: a rectangle of tiles (every column starts at that local tile row) instead of
            // the trapezoid below the diagonal (potri.cu)
            d.rowFirst[c] = std::min(rowFirstFixed >= 0 ? rowFirstFixed : C.firstLocalRowGE(gn), rowEnd);
            tiles += rowEnd - d.rowFirst[c];
        }
        if (tiles == 0) continue;
        const int bptBig = (C.nb + 127) / 128;
        const int cfg = gemm_pick_cfg(tiles * bptBig * bptBig);
        int bm, bn;
        gemm_cfg_dims(cfg, &bm, &bn);
        if (ctx->cplx) bm = bn = 64;
        d.bptM = (C.nb + bm - 1) / bm;
        d.bptN = (C.nb + bn - 1) / bn;
        long long items = 0;
        for (int c = 0; c < ncols; c++) {
            d.itemStart[c] = (int)items;
            items += (long long)(rowEnd - d.rowFirst[c]) * d.bptM * d.bptN;
        }
        d.itemStart[ncols] = (int)items;
        ctx->launches++;
        cudaEvent_t pe0 = nullptr, pe1 = nullptr;
        if (ctx->profile) {
            cudaEventCreate(&pe0);
            cudaEventCreate(&pe1);
            cudaEventRecord(pe0, s);
        }
        int rc = ctx->cplx ? zgemm_launch(d, (int)items, s) : gemm_launch(d, (int)items, cfg, s);
        if (ctx->profile) {
            cudaEventRecord(pe1, s);
            // algorithmic flops of this launch: 2 nb^2 K per GEMM tile, nb (nb+1) K per SYRK tile
            double fl = 0.0;
            for (int c = 0; c < ncols; c++) {
                const int gn = (c0 + c) * C.Q + C.pcol;
                for (int il = d.rowFirst[c]; il < rowEnd; il++) {
                    const int gm = il * C.P + C.prow;
                    const double tm = C.rowsOfTile(gm), tn = C.colsOfTile(gn);
                    fl += ((gm == gn) ? tm * (tm + 1.0) * K : 2.0 * tm * tn * K) * (ctx->cplx ? 4.0 : 1.0);
                }
            }
            ctx->profEv.push_back(pe0);
            ctx->profEv.push_back(pe1);
            ctx->profFlops.push_back(fl);
            ctx->profStream.push_back(s == ctx->sUpdate ? 1 : 0);
        }
        if (rc) return rc;
    }
    return 0;
}

// DPB_TRACE=1: timestamp the end of every phase of every step (debugging aid; adds event records)
struct TraceRec { int k; int phase; cudaEvent_t ev; };
static std::vector<TraceRec> g_trace;
static cudaEvent_t g_trace0 = nullptr;
bool trace_on() {
    static int on = -1;
    if (on < 0) on = getenv("DPB_TRACE") ? 1 : 0;
    return on == 1;
}
void trace_mark(cudaStream_t s, int k, int phase) {
    if (!trace_on()) return;
    cudaEvent_t e;
    cudaEventCreate(&e);
    cudaEventRecord(e, s);
    g_trace.push_back({k, phase, e});
}
void trace_begin(cudaStream_t s) {
    if (!trace_on()) return;
    if (!g_trace0) cudaEventCreate(&g_trace0);
    cudaEventRecord(g_trace0, s);
}
void trace_dump() {
    if (!trace_on() || !g_trace0) return;
    static const char* names[] = {"potrf", "trsm", "lookahead", "bulk", "bcast_diag", "bcast_panel", "catchup"};
    for (auto& r : g_trace) {
        float ms = 0;
        cudaEventElapsedTime(&ms, g_trace0, r.ev);
        fprintf(stderr, "TRACE k=%d %s end=%.3f ms\n", r.k, names[r.phase], ms);
        cudaEventDestroy(r.ev);
    }
    g_trace.clear();
}

// ---- single-process schedule (P = Q = 1): panels are read in place from the matrix ----
static int potrf_lower_1gpu(parsec_context_s* ctx, const DevMatrix& A) {
    const int nt = A.nt, nb = A.nb;
    trace_begin(ctx->sPanel);
    int rc = ensure_winv(ctx, nb);
    if (rc) return rc;
    cudaStream_t sp = ctx->sPanel, su = ctx->sUpdate;
    // the update stream must see everything enqueued on the panel stream so far (staging copies)
    cudaEvent_t ev0 = get_event(ctx, 0);
    DPB_CUDA(cudaEventRecord(ev0, sp));
    DPB_CUDA(cudaStreamWaitEvent(su, ev0, 0));
    size_t evi = 1;
    cudaEvent_t evUpdPrev = nullptr;
    // A host matrix that is still arriving tile column by tile column (pipelined stage-in): step k updates only the
    // columns [.., avail) that have landed; later arrivals are caught up with ALL panels factored so far in one
    // grouped launch (K spans the finished tile columns, read in place) and join the regular update from then on.
    const bool arriving = !ctx->evIn.empty() && (int)ctx->evIn.size() == nt && !ctx->cplx;
    if (!ctx->evIn.empty() && !arriving) {
        DPB_CUDA(cudaStreamWaitEvent(sp, ctx->evIn.back(), 0));
        DPB_CUDA(cudaStreamWaitEvent(su, ctx->evIn.back(), 0));
    }
    ArrivalModel model;
    if (arriving) model.init(A, ctx->h2dBw);
    int avail = arriving ? 0 : nt;
    const long long kts = (long long)A.lmt * A.tileElems;
    // columns [avail, upto) <- updates of panels 0 .. panels-1, on stream s, behind the copy of column upto-1
    auto catch_up = [&](cudaStream_t s, int upto, int panels) -> int {
        upto = std::min(upto, nt);
        if (upto <= avail) return 0;
        DPB_CUDA(cudaStreamWaitEvent(s, ctx->evIn[upto - 1], 0));  // copies are issued left to right
        if (panels <= 0) return 0;
        return trailing_update(ctx, s, A, A.tile(0, 0), A.tileElems, A.tile(0, 0), A.tileElems, panels, nt, -1.0, panels * nb,
                               nb, kts, -1, avail, upto);
    };
    for (int k = 0; k < nt; k++) {
        const int tk = A.rowsOfTile(k);
        double* winv = ctx->wInv + (size_t)(k & 1) * ctx->wInvSlot;
        double* Akk = A.tile(k, k);
        if (k >= avail) {  // first step, or a column that landed late: bring it up to date before factoring it
            rc = catch_up(sp, k + 1, k);
            if (rc) return rc;
            if (arriving) { model.waited(k); model.work(A, avail, k + 1, k); }
            avail = k + 1;
        }
        rc = potrf_tile_fused(ctx, sp, Akk, A.mb, tk, winv, ctx->dInfo, k * nb);
        if (rc) return rc;
        trace_mark(sp, k, 0);
        const int below = nt - k - 1;
        if (below <= 0) break;
        rc = trsm_panel_lower(ctx, sp, Akk, A.mb, tk, winv, A.tile(k + 1, k), A.mb, A.mb, below, A.tileElems,
                              A.rowsOfTile(nt - 1));
        if (rc) return rc;
        trace_mark(sp, k, 1);
        cudaEvent_t evPanel = get_event(ctx, evi++);
        DPB_CUDA(cudaEventRecord(evPanel, sp));
        // tile column k is final: a staged (host) matrix gets it back now, underneath the remaining steps
        if (ctx->curTp && ctx->earlyOut >= 0) {
            rc = stage_out_column(ctx, ctx->curTp->A, &ctx->curTp->dA, k, dplasmaLower, evPanel);
            if (rc) return rc;
            ctx->earlyOut = k + 1;
        }
        const double* panel = A.tile(0, k);  // tile (i,k) at panel + i*tileElems
        const int availOld = avail;
        // look-ahead: tile column k+1 on the panel stream (after the previous bulk update wrote it)
        if (evUpdPrev) DPB_CUDA(cudaStreamWaitEvent(sp, evUpdPrev, 0));
        if (k + 1 >= avail) {
            rc = catch_up(sp, k + 2, k + 1);  // not touched yet: panels 0 .. k at once
            // the two streams share the GPU: what the panel stream does also lets time pass for the copies
            if (arriving) { model.waited(k + 1); model.work(A, avail, k + 2, k + 1); }
            avail = k + 2;
        } else {
            rc = trailing_update(ctx, sp, A, panel, A.tileElems, panel, A.tileElems, k + 1, k + 2, -1.0, nb);
        }
        if (rc) return rc;
        trace_mark(sp, k, 2);
        // bulk: the tile columns k+2.. that had landed before this step, on the update stream
        bool bulk = false;
        if (k + 2 < std::min(nt, availOld)) {
            DPB_CUDA(cudaStreamWaitEvent(su, evPanel, 0));
            rc = trailing_update(ctx, su, A, panel, A.tileElems, panel, A.tileElems, k + 2, nt, -1.0, nb, 0, 0, -1, 0, availOld);
            if (rc) return rc;
            trace_mark(su, k, 3);
            bulk = true;
            if (arriving) model.work(A, k + 2, availOld, 1);
        }
        if (avail < nt) {
            // the columns expected to have landed by the time the GPU gets here: panels 0 .. k in one launch,
            // behind their copy event
            model.overhead(0.6e-3);  // tile factorization + panel solve of the step
            int target = std::max(avail, model.arrived());
            if (k + 3 >= nt) target = nt;
            if (target > avail) {
                DPB_CUDA(cudaStreamWaitEvent(su, evPanel, 0));
                rc = catch_up(su, target, k + 1);
                if (rc) return rc;
                trace_mark(su, k, 6);
                model.work(A, avail, target, k + 1);
                avail = target;
                bulk = true;
            }
        }
        if (bulk) {
            cudaEvent_t evUpd = get_event(ctx, evi++);
            DPB_CUDA(cudaEventRecord(evUpd, su));
            evUpdPrev = evUpd;
        } else {
            evUpdPrev = nullptr;
        }
    }
    // join: the panel stream is the one wait() synchronises / times
    cudaEvent_t evEnd = get_event(ctx, evi++);
    DPB_CUDA(cudaEventRecord(evEnd, su));
    DPB_CUDA(cudaStreamWaitEvent(sp, evEnd, 0));
    return 0;
}

int potrf_multi(parsec_context_s* ctx, parsec_taskpool_s* tp);  // potrf_dist.cu
void trace_dump();

int potrf_enqueue(parsec_context_s* ctx, parsec_taskpool_s* tp) {
    DevMatrix& A = tp->dA;
    if (A.mb != A.nb) return -3;
    ctx->cplx = (A.esz == 2);
    if (tp->uplo == dplasmaLower) {
        const int rc = (A.P * A.Q == 1) ? potrf_lower_1gpu(ctx, A) : potrf_multi(ctx, tp);
        ctx->cplx = 0;
        return rc;
    }
    // dplasmaUpper (src/zpotrf_U.jdf): U(k,n) = L(n,k)^H and every tile operation of the upper
    // DAG is the (conjugate) transpose of its lower counterpart, so the factorization runs as the lower
    // schedule on the transposed tile grid and is transposed back (two HBM-bound passes).
    if (A.P * A.Q != 1) {
        // grid: the process that holds tile (m, n) of the P x Q distribution holds tile (n, m) of the transpose
        // distributed over the Q x P grid with swapped coordinates -- each rank transposes its own tiles into a
        // second buffer, the lower schedule runs on that view with the row and column communicators exchanged,
        // and the factor is transposed back in place.  No tile crosses NVLink for the transposition.
        if (A.P * A.Q != ctx->nranks) { ctx->cplx = 0; return -303; }
        int rc = ensure_grid_comms(ctx, A.P, A.Q, A.ip, A.jq);
        DevMatrix W;
        W.base = nullptr;
        if (!rc) rc = transposed_copy(ctx, ctx->sPanel, A, dplasmaUpper, &W);
        if (!rc) {
            swap_grid_comms(ctx);
            parsec_taskpool_s low = *tp;
            low.uplo = dplasmaLower;
            low.dA = W;
            low.dA.owned = false;
            rc = potrf_multi(ctx, &low);
            swap_grid_comms(ctx);
        }
        if (!rc) rc = launch_transpose_local(ctx, ctx->sPanel, W, A, dplasmaLower);
        cudaStreamSynchronize(ctx->sPanel);
        if (W.base) cudaFree(W.base);
        ctx->cplx = 0;
        return rc;
    }
    int rc = launch_transpose_grid(ctx, ctx->sPanel, A);  // conjugate transpose for complex tiles
    if (!rc) rc = potrf_lower_1gpu(ctx, A);
    if (!rc) rc = launch_transpose_grid(ctx, ctx->sPanel, A);
    ctx->cplx = 0;
    return rc;
}

}  // namespace dpb

using namespace dpb;

extern "C" {

parsec_taskpool_t* dplasma_dpotrf_New(dplasma_enum_t uplo, parsec_tiled_matrix_t* A, int* info) {
    if (uplo != dplasmaUpper && uplo != dplasmaLower) {
        // dplasma_error("dplasma_zpotrf_New", "illegal value of uplo")  (src/zpotrf_wrapper.c:187-190)
        fprintf(stderr, "dplasma_dpotrf_New: illegal value of uplo\n");
        return nullptr;
    }
    if (info) *info = 0;  // src/zpotrf_wrapper.c:192
    parsec_taskpool_s* tp = new parsec_taskpool_s();
    tp->kind = TP_POTRF;
    tp->uplo = uplo;
    tp->A = A;
    tp->info = info;
    tp->hmb = 0;
    tp->staged = false;
    return tp;
}

void dplasma_dpotrf_setrecursive(parsec_taskpool_t* tp, int hmb) {
    if (tp) tp->hmb = hmb;  // accepted; the in-tile recursion block is fixed (IB)
}

void dplasma_dpotrf_Destruct(parsec_taskpool_t* tp) {
    if (!tp) return;
    stage_free(&tp->dA);
    delete tp;
}
void dplasma_dpotrf_Destroy(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }

int parsec_context_add_taskpool(parsec_context_t* ctx, parsec_taskpool_t* tp) {
    if (!ctx || !tp) return -1;
    ctx->queue.push_back(tp);
    return 0;
}

// Enqueue all GPU work of the queued taskpools; returns without waiting for the GPU.
int parsec_context_start(parsec_context_t* ctx) {
    if (!ctx) return -1;
    cudaSetDevice(ctx->device);
    // a queue that holds anything but factorizations runs operation by operation, in order (taskpools.cu)
    for (parsec_taskpool_s* tp : ctx->queue)
        if (tp->kind != TP_POTRF) return run_queue_in_order(ctx);
    DPB_CUDA(cudaMemsetAsync(ctx->dInfo, 0, sizeof(int), ctx->sPanel));
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    ctx->evIn.clear();
    for (parsec_taskpool_s* tp : ctx->queue) {
        if (tp->kind != TP_POTRF) continue;
        // A single queued lower factorization of a host matrix (tile or LAPACK storage, one process or a grid):
        // the stage-in is pipelined, the schedule follows the arrival of the tile columns.  Everything else:
        // plain stage-in.
        const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)tp->A;
        cudaPointerAttributes at;
        bool onDevice = cudaPointerGetAttributes(&at, bc->mat) == cudaSuccess &&
                        (at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged);
        cudaGetLastError();
        const bool pipe = ctx->queue.size() == 1 && tp->uplo == dplasmaLower && tp->A->mtype == PARSEC_MATRIX_DOUBLE &&
                          !onDevice && tp->A->mb == tp->A->nb && tp->A->nb % 16 == 0 && tp->A->nt >= 4 &&
                          !getenv("DPB_NO_PIPELINE");
        tp->ooc = ctx->queue.size() == 1 && ooc_wanted(ctx, tp->A, tp->uplo);
        if (tp->ooc) continue;  // never resident as a whole: the out-of-core schedule moves its own blocks
        int rc = pipe ? stage_in_pipelined(ctx, tp->A, &tp->dA, tp->uplo) : stage_in(ctx, tp->A, &tp->dA, true, tp->uplo);
        if (rc) return rc;
        tp->staged = true;
    }
    DPB_CUDA(cudaEventRecord(ctx->evStart, ctx->sPanel));
    std::vector<int> early(ctx->queue.size(), 0);
    for (size_t i = 0; i < ctx->queue.size(); i++) {
        parsec_taskpool_s* tp = ctx->queue[i];
        // early column write-back: single-process lower factorizations of staged matrices
        const bool canEarly = tp->dA.owned && tp->uplo == dplasmaLower;
        ctx->curTp = canEarly ? tp : nullptr;
        ctx->earlyOut = canEarly ? 0 : -1;
        int rc = tp->ooc ? potrf_out_of_core(ctx, tp) : potrf_enqueue(ctx, tp);
        early[i] = canEarly ? ctx->earlyOut : 0;
        ctx->curTp = nullptr;
        if (rc) return rc;
    }
    DPB_CUDA(cudaEventRecord(ctx->evStop, ctx->sPanel));
    for (size_t i = 0; i < ctx->queue.size(); i++) {
        parsec_taskpool_s* tp = ctx->queue[i];
        if (early[i] > 0) {
            // columns [0, early) already went out on the copy stream; the rest follow on the panel stream
            for (int jl = early[i]; jl < tp->dA.lnt; jl++) {
                cudaEvent_t e;
                cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
                cudaEventRecord(e, ctx->sPanel);
                int rc = stage_out_column(ctx, tp->A, &tp->dA, jl, tp->uplo, e);
                cudaEventDestroy(e);
                if (rc) return rc;
            }
            continue;
        }
        int rc = stage_out(ctx, tp->A, &tp->dA, tp->uplo);
        if (rc) return rc;
    }
    DPB_CUDA(cudaMemcpyAsync(ctx->hInfo, ctx->dInfo, sizeof(int), cudaMemcpyDeviceToHost, ctx->sPanel));
    return 0;
}

int parsec_context_wait(parsec_context_t* ctx) {
    if (!ctx) return -1;
    DPB_CUDA(cudaStreamSynchronize(ctx->sPanel));
    DPB_CUDA(cudaStreamSynchronize(ctx->sUpdate));
    DPB_CUDA(cudaStreamSynchronize(ctx->sCopy));
    float ms = 0.f;
    if (ctx->seqElapsed >= 0.0) ctx->lastElapsed = ctx->seqElapsed;
    else if (cudaEventElapsedTime(&ms, ctx->evStart, ctx->evStop) == cudaSuccess) ctx->lastElapsed = ms * 1e-3;
    ctx->seqElapsed = -1.0;
    if (ctx->h2dBytes > 0.0 && ctx->evH2d[0] && ctx->evH2d[1]) {
        float hms = 0.f;
        if (cudaEventElapsedTime(&hms, ctx->evH2d[0], ctx->evH2d[1]) == cudaSuccess && hms > 0.f) {
            const double bw = ctx->h2dBytes / (hms * 1e-3);
            ctx->h2dBw = ctx->h2dBw > 0.0 ? 0.5 * (ctx->h2dBw + bw) : bw;
        }
        ctx->h2dBytes = 0.0;
    }
    dpb::trace_dump();
    // the arrival events belong to this run only (other drivers reuse the schedule on resident matrices)
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    ctx->evIn.clear();
    for (parsec_taskpool_s* tp : ctx->queue) {
        if (!tp->done && tp->info && *tp->info == 0) *tp->info = *ctx->hInfo;  // rank-local until the wrapper reduces it
        stage_free(&tp->dA);
    }
    ctx->queue.clear();
    return 0;
}

int info_allreduce_max(parsec_context_s* ctx, int info);  // comm.cu

int dplasma_dpotrf(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    if (!parsec) return -1;
    int info = 0;
    parsec_taskpool_t* tp = dplasma_dpotrf_New(uplo, A, &info);
    // illegal uplo: New has printed the message and returned NULL; the reference's blocking wrapper then
    // skips the run and returns info == 0 (src/zpotrf_wrapper.c:318-335)
    if (!tp) return 0;
    int rc = parsec_context_add_taskpool(parsec, tp);
    if (!rc) rc = parsec_context_start(parsec);
    int rc2 = parsec_context_wait(parsec);
    dplasma_dpotrf_Destruct(tp);
    const int err = rc ? (rc < 0 ? rc : -rc - 1000) : (rc2 ? (rc2 < 0 ? rc2 : -rc2 - 1000) : 0);
    if (A->super.nodes > 1) {
        // src/zpotrf_wrapper.c:331-332.  A rank whose run failed still takes part, with INT_MAX, so that
        // the others neither wait for it forever nor report success.
        const int g = info_allreduce_max(parsec, err ? 0x7fffffff : info);
        if (err) return err;
        return g == 0x7fffffff ? -102 : g;
    }
    return err ? err : info;
}

// profile of the grouped trailing-update launches (the dominant kernel): mode 1 starts
// collecting, dplasma_b200_profile_read sums event-timed seconds / algorithmic flops / launches
void dplasma_b200_profile(parsec_context_t* ctx, int on) {
    if (!ctx) return;
    for (cudaEvent_t e : ctx->profEv) cudaEventDestroy(e);
    ctx->profEv.clear();
    ctx->profFlops.clear();
    ctx->profStream.clear();
    ctx->profile = on;
}
// which: -1 every launch; 1 the bulk updates on the update stream (launches on one stream never overlap each
// other, so their summed duration is a share of the step <= 1); 0 the look-ahead / left-looking launches on
// the panel stream, which run concurrently with the bulk update of the previous step
int dplasma_b200_profile_read_stream(parsec_context_t* ctx, int which, double* seconds, double* flops, int* launches,
                                     double* best_tflops) {
    if (!ctx) return -1;
    cudaDeviceSynchronize();
    double t = 0.0, f = 0.0, best = 0.0;
    int cnt = 0;
    const size_t n = ctx->profFlops.size();
    for (size_t i = 0; i < n; i++) {
        if (which >= 0 && ctx->profStream[i] != which) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, ctx->profEv[2 * i], ctx->profEv[2 * i + 1]) != cudaSuccess) continue;
        t += ms * 1e-3;
        f += ctx->profFlops[i];
        cnt++;
        if (ms > 0.5f) best = std::max(best, ctx->profFlops[i] / (ms * 1e-3) / 1e12);
    }
    if (seconds) *seconds = t;
    if (flops) *flops = f;
    if (launches) *launches = cnt;
    if (best_tflops) *best_tflops = best;
    return 0;
}
int dplasma_b200_profile_read(parsec_context_t* ctx, double* seconds, double* flops, int* launches,
                              double* best_tflops) {
    return dplasma_b200_profile_read_stream(ctx, -1, seconds, flops, launches, best_tflops);
}

// complex double: same driver, interleaved complex tiles (src/zpotrf_wrapper.c is the source all
// four precisions are generated from)
parsec_taskpool_t* dplasma_zpotrf_New(dplasma_enum_t uplo, parsec_tiled_matrix_t* A, int* info) {
    return dplasma_dpotrf_New(uplo, A, info);
}
void dplasma_zpotrf_Destruct(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
void dplasma_zpotrf_Destroy(parsec_taskpool_t* tp) { dplasma_dpotrf_Destruct(tp); }
int dplasma_zpotrf(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A) {
    if (A && A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return -3;
    return dplasma_dpotrf(parsec, uplo, A);
}

int dplasma_dpotrf_rec(parsec_context_t* parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t* A, int hmb) {
    (void)hmb;
    return dplasma_dpotrf(parsec, uplo, A);
}

}  // extern "C"
