// Multi-GPU tiled Cholesky: one process per GPU on a P x Q grid, owner-computes over the 2D
// block-cyclic layout.  The three tile broadcasts that PaRSEC derives from the flows of
// src/zpotrf_L.jdf are issued explicitly, once per step, as NCCL collectives over NVLink:
//   (1) L(k,k)           -> the ranks of grid column k%Q        (:112  T -> potrf_ztrsm(k+1.., k))
//   (2) L(m,k), m > k    -> the ranks of grid row m%P           (:213-214  -> herk / gemm operand A)
//   (3) L(n,k), n > k    -> the ranks of grid column n%Q        (:215  -> gemm operand B)
// (2) moves a whole contiguous slab of local tiles per rank in one broadcast; (3) is a grouped
// set of per-tile broadcasts whose root alternates with n%P.  The panel of step k+1 (factor,
// solve, broadcasts) runs on the high-priority panel stream underneath the bulk update of
// step k; panel buffers are double-buffered.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <vector>
#include "internal.h"

namespace dpb {

// ---------------------------------------------------------------------------------------
// The communication plan of one step: pure host arithmetic, exported for the CPU (gloo) tests.
// ---------------------------------------------------------------------------------------
struct PanelPlan {
    int kq, kp;        // grid column / row owning tile column k / the diagonal tile
    int il0, nrows;    // my local tile rows [il0, il0+nrows) have global index >= gmFirst
    int rowRoot;       // rank index inside my grid row (= grid column) that owns tile column k
    int jl0, ncols;    // my local tile cols [jl0, jl0+ncols) have global index >= gmFirst
    // per local column jl0+c: root (grid row) inside my grid column and its local row index there
    std::vector<int> colRoot, colSrcIl;
};

static void compute_panel_plan(const DevMatrix& A, int k, int gmFirst, PanelPlan* p) {
    p->kq = k % A.Q;
    p->kp = k % A.P;
    const int rowEnd = std::min(A.firstLocalRowGE(A.mt), A.lmt);
    const int colEnd = std::min(A.firstLocalColGE(A.nt), A.lnt);
    p->il0 = std::min(A.firstLocalRowGE(gmFirst), rowEnd);
    p->nrows = rowEnd - p->il0;
    p->rowRoot = p->kq;
    p->jl0 = std::min(A.firstLocalColGE(gmFirst), colEnd);
    p->ncols = colEnd - p->jl0;
    p->colRoot.resize(p->ncols);
    p->colSrcIl.resize(p->ncols);
    for (int c = 0; c < p->ncols; c++) {
        const int gn = (p->jl0 + c) * A.Q + A.pcol;
        p->colRoot[c] = gn % A.P;
        p->colSrcIl[c] = gn / A.P;  // gn = il*P + root
    }
}

static int ensure_panel_bufs(parsec_context_s* ctx, const DevMatrix& A, int nsets = 2) {
    const size_t diag = (size_t)A.tileElems + (size_t)((A.nb + IB - 1) / IB) * IB * IB * A.esz;
    const size_t need = (size_t)(A.lmt + A.lnt) * A.tileElems + diag;
    if (ctx->panelBase && ctx->panelSetElems >= need && ctx->panelSets >= nsets) return 0;
    if (ctx->panelBase) {
        // the panels of an earlier operation may still be in use by work in flight on the two streams
        cudaStreamSynchronize(ctx->sPanel);
        cudaStreamSynchronize(ctx->sUpdate);
        cudaFree(ctx->panelBase);
    }
    ctx->panelBase = nullptr;
    const size_t setElems = std::max(need, ctx->panelSetElems);
    const int sets = std::max(nsets, ctx->panelSets);
    DPB_CUDA(cudaMalloc(&ctx->panelBase, setElems * sets * sizeof(double)));
    ctx->panelSetElems = setElems;
    ctx->panelSets = sets;
    return 0;
}
struct PanelBufs {
    double* row;   // lmt tiles, indexed by local tile row
    double* col;   // lnt tiles, indexed by local tile column
    double* diag;  // one tile followed by the inverted diagonal blocks
};
static PanelBufs panel_bufs(parsec_context_s* ctx, const DevMatrix& A, int set) {
    PanelBufs b;
    b.row = ctx->panelBase + (size_t)set * ctx->panelSetElems;
    b.col = b.row + (size_t)A.lmt * A.tileElems;
    b.diag = b.col + (size_t)A.lnt * A.tileElems;
    return b;
}

struct PanelView {
    const double* rowPtr; long long rowStride;  // tile for local row il at rowPtr + il*rowStride
    const double* colPtr; long long colStride;  // tile for local col jl at colPtr + jl*colStride
};

// Make the tiles of global tile column k with global row >= gmFirst available as row / column
// panels on every rank.  src: on ranks of grid column k%Q, tile for local row il at src + il*srcStride.
// inPlaceRoot: the owning ranks read their own tiles where they lie instead of from the buffer.
static int panel_exchange(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int k, int gmFirst,
                          const double* src, long long srcStride, int set, bool inPlaceRoot, PanelView* out) {
    PanelPlan pl;
    compute_panel_plan(A, k, gmFirst, &pl);
    PanelBufs b = panel_bufs(ctx, A, set);
    const bool owner = (A.pcol == pl.kq);
    const size_t te = (size_t)A.tileElems;
    // ---- (2) row broadcast ----
    const double* rowPtr = b.row;
    long long rowStride = A.tileElems;
    if (A.Q == 1) {
        rowPtr = src; rowStride = srcStride;
    } else if (pl.nrows > 0) {
        const double* send = owner ? src + (size_t)pl.il0 * srcStride : b.row + (size_t)pl.il0 * te;
        double* recv = b.row + (size_t)pl.il0 * te;
        if (owner && inPlaceRoot && srcStride == A.tileElems) {
            recv = const_cast<double*>(send);  // in place on the root: no local copy
            rowPtr = src; rowStride = srcStride;
        } else if (owner && srcStride != A.tileElems) {
            return -310;  // non-contiguous source slabs are not used by any caller
        }
        int rc = bcast(ctx, ctx->commRow, send, recv, (size_t)pl.nrows * te, pl.rowRoot, s);
        if (rc) return rc;
    }
    // ---- (3) column ("transposed") broadcast ----
    const double* colPtr = b.col;
    long long colStride = A.tileElems;
    if (A.P == 1) {
        // all tile rows are local: the tile for local column jl is row tile gn = jl*Q + pcol
        colPtr = rowPtr + (size_t)A.pcol * rowStride;
        colStride = (long long)A.Q * rowStride;
    } else if (pl.ncols > 0) {
        int rc = group_start(ctx);
        if (rc) return -311;
        for (int c = 0; c < pl.ncols; c++) {
            const int jl = pl.jl0 + c;
            const double* send = rowPtr + (size_t)pl.colSrcIl[c] * rowStride;  // only dereferenced on the root
            rc = bcast(ctx, ctx->commCol, send, b.col + (size_t)jl * te, te, pl.colRoot[c], s);
            if (rc) { group_end(ctx); return rc; }
        }
        if (group_end(ctx)) return -312;
    }
    out->rowPtr = rowPtr; out->rowStride = rowStride;
    out->colPtr = colPtr; out->colStride = colStride;
    return 0;
}

static cudaEvent_t dist_event(parsec_context_s* ctx, size_t idx) {
    while (ctx->evPool.size() <= idx) {
        cudaEvent_t e;
        cudaEventCreateWithFlags(&e, cudaEventDisableTiming);
        ctx->evPool.push_back(e);
    }
    return ctx->evPool[idx];
}

// A host matrix that is still arriving (pipelined stage-in, ctx->evIn holds one event per local tile column):
// the right-looking step k updates only the local columns [.., avail) that have landed; columns that land later
// are caught up with every panel factored so far -- which therefore stay alive (one panel set each) during the
// first `window` steps -- and join the regular update from then on.  Purely local decisions: the collectives of a
// step do not depend on them.
int potrf_multi(parsec_context_s* ctx, parsec_taskpool_s* tp) {
    const DevMatrix& A = tp->dA;
    if (A.P * A.Q != ctx->nranks) {
        fprintf(stderr, "dplasma_b200: descriptor grid %dx%d does not match the %d initialised ranks\n", A.P, A.Q,
                ctx->nranks);
        return -303;
    }
    const bool arriving = !ctx->evIn.empty() && (int)ctx->evIn.size() == A.lnt;
    ArrivalModel model;
    int window = 0;
    if (arriving) {
        // the panels of the first `window` steps stay alive (one set each, ~ (lmt + lnt) tiles) so that columns
        // landing during those steps can be caught up
        model.init(A, ctx->h2dBw);
        window = std::min(A.nt, 8);
    }
    const int nsets = std::max(2, window);
    int rc = ensure_grid_comms(ctx, A.P, A.Q, A.ip, A.jq);
    if (!rc) rc = ensure_winv(ctx, A.nb);
    if (!rc) rc = ensure_panel_bufs(ctx, A, nsets);
    if (rc) return rc;
    cudaStream_t sp = ctx->sPanel, su = ctx->sUpdate;
    const int nt = A.nt, nb = A.nb;
    const size_t te = (size_t)A.tileElems;
    const size_t winvElems = (size_t)((nb + IB - 1) / IB) * IB * IB * A.esz;
    trace_begin(sp);
    cudaEvent_t ev0 = dist_event(ctx, 0);
    DPB_CUDA(cudaEventRecord(ev0, sp));
    DPB_CUDA(cudaStreamWaitEvent(su, ev0, 0));
    size_t evi = 1;
    cudaEvent_t evUpdPrev = nullptr;
    const int rowEnd = std::min(A.firstLocalRowGE(A.mt), A.lmt);
    const int colEnd = std::min(A.firstLocalColGE(A.nt), A.lnt);
    int avail = arriving ? 0 : A.lnt;   // local columns [0, avail) have landed and carry every update issued so far
    std::vector<PanelView> kept;         // panels 0 .. k-1 while columns are still missing
    // bring local columns [avail, upto) up to date with the panels kept so far, on stream s
    auto catch_up = [&](cudaStream_t s, int upto, int panels) -> int {
        upto = std::min(upto, colEnd);
        if (upto <= avail) return 0;
        DPB_CUDA(cudaStreamWaitEvent(s, ctx->evIn[upto - 1], 0));  // copies are issued left to right
        for (int q = 0; q < panels; q++) {
            const PanelView& v = kept[q];
            int r = trailing_update(ctx, s, A, v.rowPtr, v.rowStride, v.colPtr, v.colStride, q + 1, nt, -1.0, nb, 0, 0, -1,
                                    avail, upto);
            if (r) return r;
        }
        return 0;
    };
    for (int k = 0; k < nt; k++) {
        const int set = k % nsets;
        const int kp = k % A.P, kq = k % A.Q;
        const int tk = A.rowsOfTile(k);
        PanelBufs b = panel_bufs(ctx, A, set);
        const bool inCol = (A.pcol == kq);
        const int jlk = k / A.Q;  // local column index of tile column k on the ranks of grid column kq
        if (arriving && k == window) {
            // the panels kept so far are about to be overwritten: their last readers (the catch-up launches of the
            // previous step on the update stream) must be done, and everything still missing has to land now
            if (evUpdPrev) DPB_CUDA(cudaStreamWaitEvent(sp, evUpdPrev, 0));
            rc = catch_up(sp, colEnd, (int)kept.size());
            if (rc) return rc;
            avail = A.lnt;
            cudaEvent_t evAll = dist_event(ctx, evi++);
            DPB_CUDA(cudaEventRecord(evAll, sp));
            DPB_CUDA(cudaStreamWaitEvent(su, evAll, 0));
        }
        if (inCol && jlk >= avail) {
            // my part of the panel itself has not been touched yet: land it and apply panels 0 .. k-1 (look-ahead
            // normally did that one step ago; this is the first step, or a column that arrived late)
            rc = catch_up(sp, jlk + 1, (int)kept.size());
            if (rc) return rc;
            model.waited(jlk);
            model.work(A, avail, jlk + 1, (int)kept.size());
            cudaEvent_t evC = dist_event(ctx, evi++);
            DPB_CUDA(cudaEventRecord(evC, sp));
            DPB_CUDA(cudaStreamWaitEvent(su, evC, 0));
            avail = jlk + 1;
        }
        // ---- (1) diagonal tile: factor on the owner, broadcast down grid column kq ----
        if (inCol) {
            double* winv = b.diag + te;
            if (A.prow == kp) {
                // factor in place with the context workspace, then stage the inverses (and, when the
                // tile has to travel, the tile itself) in the panel buffer
                double* Akk = A.tile(k / A.P, jlk);
                rc = potrf_tile_fused(ctx, sp, Akk, A.mb, tk, ctx->wInv, ctx->dInfo, k * nb);
                if (rc) return rc;
                DPB_CUDA(cudaMemcpyAsync(winv, ctx->wInv, winvElems * sizeof(double), cudaMemcpyDeviceToDevice, sp));
                if (A.P > 1)
                    DPB_CUDA(cudaMemcpyAsync(b.diag, Akk, te * sizeof(double), cudaMemcpyDeviceToDevice, sp));
                trace_mark(sp, k, 0);
            }
            const double* Lkk = b.diag;
            if (A.P == 1) Lkk = A.tile(k / A.P, jlk);
            else {
                rc = bcast(ctx, ctx->commCol, b.diag, b.diag, te + winvElems, kp, sp);
                if (rc) return rc;
                trace_mark(sp, k, 4);
            }
            // ---- panel solve on my tiles of column k ----
            const int il0 = std::min(A.firstLocalRowGE(k + 1), rowEnd), cnt = rowEnd - il0;
            if (cnt > 0) {
                const int lastGm = (rowEnd - 1) * A.P + A.prow;
                rc = trsm_panel_lower(ctx, sp, Lkk, A.mb, tk, winv, A.tile(il0, jlk), A.mb, A.mb, cnt, A.tileElems,
                                      A.rowsOfTile(lastGm));
                if (rc) return rc;
                trace_mark(sp, k, 1);
            }
            // my tiles of tile column k are final: a staged (host) matrix gets them back now
            if (ctx->curTp && ctx->earlyOut >= 0) {
                cudaEvent_t evCol = dist_event(ctx, evi++);
                DPB_CUDA(cudaEventRecord(evCol, sp));
                rc = stage_out_column(ctx, ctx->curTp->A, &ctx->curTp->dA, jlk, dplasmaLower, evCol);
                if (rc) return rc;
                ctx->earlyOut = jlk + 1;
            }
        }
        if (k + 1 >= nt) break;
        // ---- (2) + (3): panel broadcasts ----
        PanelView pv;
        rc = panel_exchange(ctx, sp, A, k, k + 1, inCol ? A.tile(0, jlk) : nullptr, A.tileElems, set, true, &pv);
        if (rc) return rc;
        trace_mark(sp, k, 5);
        cudaEvent_t evPanel = dist_event(ctx, evi++);
        DPB_CUDA(cudaEventRecord(evPanel, sp));
        const int availOld = avail;
        if (avail < A.lnt) {
            kept.push_back(pv);
            // the look-ahead column must be complete before the next step factors it
            const int jl1 = (k + 1) / A.Q;
            if (A.pcol == (k + 1) % A.Q && jl1 >= avail) {
                if (evUpdPrev) DPB_CUDA(cudaStreamWaitEvent(sp, evUpdPrev, 0));
                rc = catch_up(sp, jl1 + 1, (int)kept.size());   // panels 0 .. k, this step's included
                if (rc) return rc;
                model.waited(jl1);
                model.work(A, avail, jl1 + 1, (int)kept.size());
                avail = jl1 + 1;
            }
        }
        // ---- look-ahead update of tile column k+1 on the panel stream (columns caught up just now already have it) ----
        if (evUpdPrev) DPB_CUDA(cudaStreamWaitEvent(sp, evUpdPrev, 0));
        rc = trailing_update(ctx, sp, A, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, k + 1, k + 2, -1.0, nb, 0, 0, -1, 0,
                             availOld);
        if (rc) return rc;
        trace_mark(sp, k, 2);
        // ---- bulk update on the update stream: the columns that had landed before this step ----
        bool bulk = false;
        if (k + 2 < nt) {
            DPB_CUDA(cudaStreamWaitEvent(su, evPanel, 0));
            rc = trailing_update(ctx, su, A, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, k + 2, nt, -1.0, nb, 0, 0, -1, 0,
                                 availOld);
            if (rc) return rc;
            trace_mark(su, k, 3);
            bulk = true;
            if (arriving) model.work(A, A.firstLocalColGE(k + 2), availOld, 1);
        }
        if (avail < A.lnt) {
            // columns expected to have landed by the time the GPU gets here: all panels so far, behind their copy
            // event, after the bulk
            model.overhead(1.5e-3);  // tile factorization, panel solve and panel broadcasts of the step
            int target = std::max(avail, model.arrived());
            // steps that find little to do are short, and the window is counted in steps: every step asks for its
            // share of the columns at least, so that the waiting is spread over steps that have work to do instead
            // of falling on the last step of the window
            target = std::max(target, (int)(((long long)A.lnt * (k + 1) + window - 1) / window));
            if (k + 1 >= window || k + 2 >= nt) target = A.lnt;
            if (target > avail) {
                if (avail != availOld) {  // the panel stream has just written columns [availOld, avail)
                    cudaEvent_t evL = dist_event(ctx, evi++);
                    DPB_CUDA(cudaEventRecord(evL, sp));
                    DPB_CUDA(cudaStreamWaitEvent(su, evL, 0));
                }
                DPB_CUDA(cudaStreamWaitEvent(su, evPanel, 0));
                rc = catch_up(su, target, (int)kept.size());
                if (rc) return rc;
                trace_mark(su, k, 6);
                model.waited(target - 1);
                model.work(A, avail, target, (int)kept.size());
                avail = target;
                bulk = true;
            }
        }
        if (bulk) {
            cudaEvent_t evUpd = dist_event(ctx, evi++);
            DPB_CUDA(cudaEventRecord(evUpd, su));
            evUpdPrev = evUpd;
        } else {
            evUpdPrev = nullptr;
        }
    }
    cudaEvent_t evEnd = dist_event(ctx, evi++);
    DPB_CUDA(cudaEventRecord(evEnd, su));
    DPB_CUDA(cudaStreamWaitEvent(sp, evEnd, 0));
    return 0;
}

// R <- R - L L^T on the lower tiles (the residual of check_dpotrf, src/dplasma_zcheck.c:89-99) for
// a matrix distributed over the grid: the factorization's own panel exchange, minus the solves.
int check_update_multi(parsec_context_s* ctx, const DevMatrix& L, const DevMatrix& R) {
    int rc = ensure_grid_comms(ctx, L.P, L.Q, L.ip, L.jq);
    if (!rc) rc = ensure_panel_bufs(ctx, L);
    if (rc) return rc;
    cudaStream_t s = ctx->sPanel;
    const size_t te = (size_t)L.tileElems;
    const int rowEnd = std::min(L.firstLocalRowGE(L.mt), L.lmt);
    for (int k = 0; k < L.nt; k++) {
        const int kp = k % L.P, kq = k % L.Q, jlk = k / L.Q;
        const int tk = L.rowsOfTile(k);
        PanelBufs b = panel_bufs(ctx, L, 0);
        const bool inCol = (L.pcol == kq);
        if (inCol) {
            // stage my tiles of column k (rows >= k) in the row buffer; the diagonal tile with its
            // strict upper part zeroed (dplasma_zlacpy(uplo), src/dplasma_zcheck.c:90)
            const int il0 = std::min(L.firstLocalRowGE(k), rowEnd), cnt = rowEnd - il0;
            if (cnt > 0)
                DPB_CUDA(cudaMemcpyAsync(b.row + (size_t)il0 * te, L.tile(il0, jlk), (size_t)cnt * te * sizeof(double),
                                         cudaMemcpyDeviceToDevice, s));
            if (L.prow == kp) {
                rc = launch_copy_tril(ctx, s, L.tile(k / L.P, jlk), L.mb, b.row + (size_t)(k / L.P) * te, L.mb, tk,
                                      dplasmaLower, 0);
                if (rc) return rc;
            }
        }
        PanelView pv;
        rc = panel_exchange(ctx, s, L, k, k, b.row, L.tileElems, 0, false, &pv);
        if (rc) return rc;
        rc = trailing_update(ctx, s, R, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, k, L.nt, -1.0, tk);
        if (rc) return rc;
    }
    return 0;
}

// ---------------------------------------------------------------------------------------
// Level-3 operations on a process grid (src/zsyrk_LN.jdf, src/zgemm_NT.jdf distributed by PaRSEC's remote dependencies):
// per tile column k of the operand, the factorization's panel exchange with gmFirst = 0 -- A(m,k) along the grid row
// of m, and the operand of the other side down the grid column of n -- then ONE grouped update launch of the local
// tiles of C.  Operands and C share the grid and the (square) tile size.  C has been scaled by beta already.
//   syrk:  C(m,n) += alpha A(m,k) A(n,k)^T   m >= n, diagonal tiles in their lower part
//   gemm:  C(m,n) += alpha A(m,k) B(n,k)^T   all local tiles
// ---------------------------------------------------------------------------------------
int syrk_lower_multi(parsec_context_s* ctx, cudaStream_t s, double alpha, const DevMatrix& A, const DevMatrix& C) {
    if (A.P != C.P || A.Q != C.Q || A.ip != C.ip || A.jq != C.jq || A.mb != C.mb || A.nb != C.nb || C.mb != C.nb) return -102;
    if (C.P * C.Q != ctx->nranks) return -303;
    int rc = ensure_grid_comms(ctx, C.P, C.Q, C.ip, C.jq);
    if (!rc) rc = ensure_panel_bufs(ctx, C);
    for (int k = 0; !rc && k < A.nt; k++) {
        const bool inCol = (A.pcol == k % A.Q);
        PanelView pv;
        rc = panel_exchange(ctx, s, C, k, 0, inCol ? A.tile(0, k / A.Q) : nullptr, A.tileElems, k & 1, true, &pv);
        if (!rc) rc = trailing_update(ctx, s, C, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, 0, C.nt, alpha, A.colsOfTile(k));
    }
    return rc;
}

int gemm_nt_multi(parsec_context_s* ctx, cudaStream_t s, double alpha, const DevMatrix& A, const DevMatrix& B,
                  const DevMatrix& C) {
    // square C only: both exchanges then run on the plan of one nt x nt tile grid
    if (C.M != C.N || C.mb != C.nb || A.mb != C.mb || A.nb != C.nb || B.mb != C.mb || B.nb != C.nb) return -102;
    if (A.P != C.P || A.Q != C.Q || B.P != C.P || B.Q != C.Q || A.ip != C.ip || A.jq != C.jq || B.ip != C.ip || B.jq != C.jq) return -102;
    if (C.P * C.Q != ctx->nranks) return -303;
    int rc = ensure_grid_comms(ctx, C.P, C.Q, C.ip, C.jq);
    if (!rc) rc = ensure_panel_bufs(ctx, C);
    ctx->gemmFull = 1;
    for (int k = 0; !rc && k < A.nt; k++) {
        const bool inCol = (C.pcol == k % C.Q);
        PanelView pa, pb;
        // rows of the product come from A's tile column k, columns from B's: two exchanges, one panel set each
        rc = panel_exchange(ctx, s, C, k, 0, inCol ? A.tile(0, k / A.Q) : nullptr, A.tileElems, 0, true, &pa);
        if (!rc) rc = panel_exchange(ctx, s, C, k, 0, inCol ? B.tile(0, k / B.Q) : nullptr, B.tileElems, 1, true, &pb);
        if (!rc) rc = trailing_update(ctx, s, C, pa.rowPtr, pa.rowStride, pb.colPtr, pb.colStride, 0, C.nt, alpha, A.colsOfTile(k), 0,
                                      0, 0);
    }
    ctx->gemmFull = 0;
    return rc;
}

// ---------------------------------------------------------------------------------------
// Distributed transpose of a triangle:  Tt(n,m) = T(m,n)^T for every stored tile (m,n) of T.
// Tile (m,n) lives on rank (m%P)*Q + n%Q, its image on rank (n%P)*Q + m%Q: point-to-point
// NCCL send/recv, one group per global tile column, then a local transposing copy.
// Used where the reference's DAGs read the triangular factor "the other way round"
// (ztrsm_LLT.jdf:132-137, zpotrf_U.jdf, zhemm in the axmb check).
// ---------------------------------------------------------------------------------------
int alloc_like(const DevMatrix& A, DevMatrix* out) {
    *out = A;
    out->base = nullptr;
    out->owned = true;
    size_t elems = (size_t)A.lmt * A.lnt * A.tileElems;
    if (elems == 0) elems = 1;
    DPB_CUDA(cudaMalloc(&out->base, elems * sizeof(double)));
    return 0;
}

// The transpose of a P x Q block-cyclic matrix, seen on the Q x P grid with swapped coordinates: the process that
// holds tile (m, n) holds tile (n, m) of the transpose, so the copy is local to each GPU (one HBM-bound launch) and
// what changes is the description -- P <-> Q, prow <-> pcol, RSRC <-> CSRC, local tile rows <-> columns.  The row
// communicator of the swapped grid is the column communicator of the original and vice versa (swap_grid_comms);
// an operation that works on ONE matrix (factorization, residual) can run on the swapped grid as it stands.
int transposed_copy(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int uplo, DevMatrix* W) {
    *W = A;
    W->base = nullptr;
    W->owned = true;
    std::swap(W->P, W->Q); std::swap(W->prow, W->pcol); std::swap(W->ip, W->jq);
    std::swap(W->lmt, W->lnt); std::swap(W->mt, W->nt); std::swap(W->M, W->N);
    std::swap(W->mb, W->nb);
    W->offR = W->offC = 0; W->userLmt = 0;
    const size_t elems = std::max<size_t>((size_t)A.lmt * A.lnt * A.tileElems, 1);
    if (cudaMalloc(&W->base, elems * sizeof(double)) != cudaSuccess) {
        cudaGetLastError();
        W->owned = false;
        return -131;
    }
    return launch_transpose_local(ctx, s, A, *W, uplo);
}
void swap_grid_comms(parsec_context_s* ctx) {
    std::swap(ctx->commRow, ctx->commCol);
    std::swap(ctx->gridP, ctx->gridQ);
    std::swap(ctx->gridIp, ctx->gridJq);
}

// ---- plan of the distributed transposition: host arithmetic, shared by dist_transpose_tri below and by
//      dplasma_b200_transpose_plan (exercised over gloo on the CPU, tests/test_dist_gloo.py) ----
struct XposeTile { int m, n, peer; };  // tile (m, n) of T; peer: logical rank it goes to (send list) / comes from (receive list)
struct XposePlan {
    int P, Q, mt;
    bool lower;
    size_t cap;  // per round: at most `cap` outgoing and `cap` incoming tiles per rank
    void rows_of(int n, int* m0, int* m1) const {
        *m0 = lower ? n : 0;
        *m1 = lower ? mt : n + 1;
    }
    static size_t cap_for(int mt, size_t tileBytes) {  // 1 GB each way, at least one full tile column, at most the triangle
        return std::max<size_t>((size_t)mt + 1, std::min<size_t>(((size_t)1 << 30) / std::max<size_t>(tileBytes, 1),
                                                                 (size_t)mt * (mt + 1) / 2));
    }
    // The columns [nA, nB) of the round that starts at nA: the same on every rank (bounds on what any rank sends or
    // receives: column n is sent by grid column n % Q, at most ceil(len / P) tiles per rank, and received by grid row
    // n % P, at most ceil(len / Q) tiles per rank).
    int round_end(int nA) const {
        std::vector<size_t> outQ(Q, 0), inP(P, 0);
        int nB = nA;
        while (nB < mt) {
            int m0, m1;
            rows_of(nB, &m0, &m1);
            const size_t o = (size_t)(m1 - m0 + P - 1) / P, i = (size_t)(m1 - m0 + Q - 1) / Q;
            if (nB > nA && (outQ[nB % Q] + o > cap || inP[nB % P] + i > cap)) break;
            outQ[nB % Q] += o; inP[nB % P] += i; nB++;
        }
        return nB;
    }
    // Both ends of a pair walk the tiles in the same order: column n ascending, row m ascending.
    // my tiles of the round, with where they go (peer == me: the tile stays)
    void send_list(int me, int nA, int nB, std::vector<XposeTile>* out) const {
        out->clear();
        const int prow = me / Q, pcol = me % Q;
        for (int n = nA; n < nB; n++) {
            if (n % Q != pcol) continue;
            int m0, m1;
            rows_of(n, &m0, &m1);
            for (int m = m0; m < m1; m++)
                if (m % P == prow) out->push_back({m, n, (n % P) * Q + (m % Q)});
        }
    }
    // the tiles that reach me from another rank
    void recv_list(int me, int nA, int nB, std::vector<XposeTile>* out) const {
        out->clear();
        const int prow = me / Q, pcol = me % Q;
        for (int n = nA; n < nB; n++) {
            if (n % P != prow) continue;
            int m0, m1;
            rows_of(n, &m0, &m1);
            for (int m = m0; m < m1; m++) {
                if (m % Q != pcol) continue;
                const int src = (m % P) * Q + (n % Q);
                if (src != me) out->push_back({m, n, src});
            }
        }
    }
};

// Tt <- T^T on the same grid (the `uplo` triangle of T; tile (m, n) goes, transposed, to tile (n, m) on process
// (n % P, m % Q)).  An all-to-all in rounds over blocks of tile columns: every rank transposes its outgoing tiles
// straight into a send staging area, packed by destination, ONE send and ONE receive per peer move the round's tiles
// (all pairs of the grid at the same time, large messages), and the received tiles are copied into place.  The staging
// is context-owned.
int dist_transpose_tri(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& T, int uplo, const DevMatrix& Tt) {
    if (T.mb != T.nb || T.mt != T.nt) return -320;
    if (T.esz != 1) return -2;  // real tiles only
    const size_t te = (size_t)T.tileElems;
    const int P = T.P, Q = T.Q, np = P * Q, me = T.prow * Q + T.pcol;
    XposePlan plan{P, Q, T.mt, uplo == dplasmaLower, XposePlan::cap_for(T.mt, te * sizeof(double))};
    const size_t cap = plan.cap;
    if (ctx->wXposeLen < 2 * cap * te) {
        if (ctx->wXpose) { cudaStreamSynchronize(s); cudaFree(ctx->wXpose); }
        ctx->wXpose = nullptr;
        ctx->wXposeLen = 0;
        DPB_CUDA(cudaMalloc(&ctx->wXpose, 2 * cap * te * sizeof(double)));
        ctx->wXposeLen = 2 * cap * te;
    }
    double* sendBuf = ctx->wXpose;
    double* recvBuf = ctx->wXpose + cap * te;
    std::vector<size_t> sendCnt(np), recvCnt(np), sendOff(np), recvOff(np), sendIdx(np), recvIdx(np);
    std::vector<XposeTile> sends, recvs;
    int rc = 0;
    for (int nA = 0; nA < T.nt && !rc;) {
        const int nB = plan.round_end(nA);
        plan.send_list(me, nA, nB, &sends);
        plan.recv_list(me, nA, nB, &recvs);
        std::fill(sendCnt.begin(), sendCnt.end(), 0);
        std::fill(recvCnt.begin(), recvCnt.end(), 0);
        for (const XposeTile& t : sends)
            if (t.peer != me) sendCnt[t.peer]++;
        for (const XposeTile& t : recvs) recvCnt[t.peer]++;
        size_t so = 0, ro = 0;
        for (int r = 0; r < np; r++) {
            sendOff[r] = so; so += sendCnt[r];
            recvOff[r] = ro; ro += recvCnt[r];
            sendIdx[r] = recvIdx[r] = 0;
        }
        if (so > cap || ro > cap) return -323;  // the bounds of round_end do not hold: a bug, not a size limit
        // ---- pack: my tiles, transposed, by destination; tiles that stay go straight into place ----
        for (size_t q = 0; q < sends.size() && !rc; q++) {
            const XposeTile& t = sends[q];
            double* to = (t.peer == me) ? Tt.tile(t.n / P, t.m / Q) : sendBuf + (sendOff[t.peer] + sendIdx[t.peer]++) * te;
            rc = launch_transpose(ctx, s, T.tile(t.m / P, t.n / Q), T.mb, to, T.mb, T.mb, T.nb);
        }
        if (rc) break;
        // ---- exchange: one message per peer and direction ----
        if (so || ro) {
            group_start(ctx);
            for (int r = 0; r < np && !rc; r++) {
                if (r == me) continue;
                const int peer = world_rank_of(T, r / Q, r % Q);
                if (sendCnt[r]) rc = p2p_send(ctx, sendBuf + sendOff[r] * te, sendCnt[r] * te, peer, s);
                if (!rc && recvCnt[r]) rc = p2p_recv(ctx, recvBuf + recvOff[r] * te, recvCnt[r] * te, peer, s);
            }
            if (group_end(ctx) && !rc) rc = -321;
        }
        // ---- unpack: received tiles into place, in the order their sender packed them ----
        for (size_t q = 0; q < recvs.size() && !rc; q++) {
            const XposeTile& t = recvs[q];
            if (cudaMemcpyAsync(Tt.tile(t.n / P, t.m / Q), recvBuf + (recvOff[t.peer] + recvIdx[t.peer]++) * te, te * sizeof(double),
                                cudaMemcpyDeviceToDevice, s) != cudaSuccess)
                rc = -322;
        }
        nA = nB;
    }
    cudaStreamSynchronize(s);  // the staging area may be reused by a later call on another stream
    return rc;
}

// ---------------------------------------------------------------------------------------
// Right-looking tiled triangular solve on the grid, "column panel" form: at step k the tiles of
// tile column k of the triangular matrix Tm (rows still to be updated, plus the diagonal tile)
// travel along the grid rows, the freshly solved RHS tile row k travels down the grid columns:
//   forward  (Tm lower): B(k) = Tm(k,k)^-1 B(k); B(m) -= Tm(m,k) B(k), m > k   (ztrsm_LLN.jdf)
//   backward (Tm upper): B(k) = Tm(k,k)^-1 B(k); B(m) -= Tm(m,k) B(k), m < k   (ztrsm_LUN.jdf)
// The transposed variants (LLT / LUT) run on a distributed transpose of the factor.
// ---------------------------------------------------------------------------------------
// The tile columns of Tm are final data: their broadcast along the grid rows for step k+1 is posted on the
// communication stream while step k computes (two panel sets alternate); the solved RHS row travels down the grid
// columns on the compute stream, where the next update needs it.
// Look-ahead: the chain of a step -- solve tile row k, send it down the columns, update the ONE tile row the next
// step solves -- stays on the compute stream; the update of all the other remaining rows goes to the update stream
// and runs underneath the chain of the next step (ztrsm_LLN.jdf gets the same overlap from its dataflow: the GEMM
// tasks of step k and the TRSM task of step k+1 are independent once row k+1 has its update).
static int solve_panels_multi(parsec_context_s* ctx, cudaStream_t s, bool lower, int diag, double alpha,
                              const DevMatrix& Tm, const DevMatrix& B) {
    int rc = ensure_grid_comms(ctx, Tm.P, Tm.Q, Tm.ip, Tm.jq);
    if (!rc) rc = ensure_winv(ctx, Tm.nb);
    if (!rc) rc = ensure_panel_bufs(ctx, Tm);
    if (!rc) rc = ensure_solve_ws(ctx, Tm.mb, B.nb);
    if (rc) return rc;
    const size_t te = (size_t)Tm.tileElems, teB = (size_t)B.tileElems;
    const int mt = Tm.mt;
    const int rowEnd = std::min(Tm.firstLocalRowGE(Tm.mt), Tm.lmt);
    const int colEndB = std::min(B.firstLocalColGE(B.nt), B.lnt);
    // solved tile row k of the RHS, one tile per local RHS tile column (context-owned, grown on demand)
    const size_t xSet = (size_t)std::max(colEndB, 1) * teB, xNeed = 2 * xSet;  // two sets: the bulk update of step k
    if (ctx->wSolveLen < xNeed) {                                               // reads its row while step k+1 receives
        if (ctx->wSolve) { cudaStreamSynchronize(s); cudaFree(ctx->wSolve); }
        ctx->wSolve = nullptr;
        DPB_CUDA(cudaMalloc(&ctx->wSolve, xNeed * sizeof(double)));
        ctx->wSolveLen = xNeed;
    }
    cudaStream_t sc = ctx->sComm;
    cudaStream_t su = (ctx->sUpdate != s) ? ctx->sUpdate : ctx->sPanel;  // bulk updates
    const size_t evBase = 900;  // events of this routine (a factorization's own count up from 0)
    cudaEvent_t evRecv[2] = {dist_event(ctx, evBase), dist_event(ctx, evBase + 1)};
    cudaEvent_t evFree[2] = {dist_event(ctx, evBase + 2), dist_event(ctx, evBase + 3)};
    cudaEvent_t evStart = dist_event(ctx, evBase + 4);
    cudaEvent_t evBulk[2] = {dist_event(ctx, evBase + 6), dist_event(ctx, evBase + 7)};   // bulk update of step kk done
    cudaEvent_t evX[2] = {dist_event(ctx, evBase + 8), dist_event(ctx, evBase + 9)};      // solved row kk is in place
    cudaEvent_t evStep[2] = {dist_event(ctx, evBase + 10), dist_event(ctx, evBase + 11)};  // chain of step kk done
    DPB_CUDA(cudaEventRecord(evStart, s));
    DPB_CUDA(cudaStreamWaitEvent(sc, evStart, 0));  // Tm may have been produced on s just before
    DPB_CUDA(cudaStreamWaitEvent(su, evStart, 0));
    auto rows_of = [&](int k, int* ilA, int* cnt) {
        // local tile rows that take part in step k: the diagonal tile's row and the rows still to update
        const int a = lower ? std::min(Tm.firstLocalRowGE(k), rowEnd) : 0;
        const int b = lower ? rowEnd : std::min(Tm.firstLocalRowGE(k + 1), rowEnd);
        *ilA = a;
        *cnt = b - a;
    };
    // (a) tile column k of Tm along the grid rows, into panel set kk & 1 (in place on the owners)
    auto post_panel = [&](int kk) -> int {
        const int k = lower ? kk : mt - 1 - kk, set = kk & 1;
        const int kq = k % Tm.Q, jlk = k / Tm.Q;
        int ilA, cnt;
        rows_of(k, &ilA, &cnt);
        if (kk >= 2) DPB_CUDA(cudaStreamWaitEvent(sc, evFree[set], 0));
        if (Tm.Q > 1 && cnt > 0) {
            const bool owner = (Tm.pcol == kq);
            PanelBufs pb = panel_bufs(ctx, Tm, set);
            double* buf = owner ? Tm.tile(ilA, jlk) : pb.row + (size_t)ilA * te;
            int r = bcast(ctx, ctx->commRow, buf, buf, (size_t)cnt * te, kq, sc);
            if (r) return r;
        }
        // the inverse blocks of the diagonal tile do not depend on the right-hand side: built here, one step ahead of
        // the chain that uses them (solve.cu, build_inverse_blocks)
        if (Tm.prow == k % Tm.P) {
            const int tk = Tm.rowsOfTile(k), ibs = solve_ibs_for(tk);
            const double* rowPtr = (Tm.Q == 1 || Tm.pcol == kq) ? Tm.tile(0, jlk) : panel_bufs(ctx, Tm, set).row;
            const double* Tkk = rowPtr + (size_t)(k / Tm.P) * te;
            double* winv64 = ctx->wInv + (size_t)set * ctx->wInvSlot;
            double* wbig = solve_ws_big(ctx, set);
            int r = build_inverse_blocks(ctx, sc, lower ? dplasmaLower : dplasmaUpper, diag, tk, Tkk, Tm.mb, winv64, wbig, Tm.mb,
                                         wbig + (size_t)Tm.mb * Tm.mb, ibs);
            if (r) return r;
        }
        DPB_CUDA(cudaEventRecord(evRecv[set], sc));
        return 0;
    };
    // DPB_TRACE_SOLVE: timing events per step -- chain: [0] start (panel there), [1] row solved, [2] row sent,
    // [3] look-ahead update done; bulk: [4], [5] around the bulk update on the update stream
    static const bool traceSolve = getenv("DPB_TRACE_SOLVE") != nullptr;
    std::vector<cudaEvent_t> tev;
    auto tmark = [&](cudaStream_t st) {
        if (!traceSolve) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, st);
        tev.push_back(e);
    };
    rc = post_panel(0);
    for (int kk = 0; kk < mt && !rc; kk++) {
        const int k = lower ? kk : mt - 1 - kk, set = kk & 1;
        const int kp = k % Tm.P, kq = k % Tm.Q, jlk = k / Tm.Q;
        const int tk = Tm.rowsOfTile(k);
        const double lalpha = (kk == 0) ? alpha : 1.0;
        if (kk + 1 < mt) {
            rc = post_panel(kk + 1);
            if (rc) break;
        }
        DPB_CUDA(cudaStreamWaitEvent(s, evRecv[set], 0));
        tmark(s);
        const double* rowPtr = panel_bufs(ctx, Tm, set).row;
        if (Tm.Q == 1 || Tm.pcol == kq) rowPtr = Tm.tile(0, jlk);
        // (b) solve the RHS tiles of tile row k (ranks of grid row kp, each for its own RHS columns)
        const int ilk = k / Tm.P;
        if (Tm.prow == kp) {
            const double* Tkk = rowPtr + (size_t)ilk * te;
            const int ibs = solve_ibs_for(tk);
            const int ul = lower ? dplasmaLower : dplasmaUpper;
            for (int jl = 0; jl < colEndB && !rc; jl++) {
                const int nc = B.colsOfTile(jl * B.Q + B.pcol);
                if (ibs > IB)
                    rc = trsm_left_big_entry(ctx, s, ul, false, tk, nc, lalpha, Tkk, Tm.mb, solve_ws_big(ctx, set), Tm.mb, ibs,
                                             solve_ws_xt(ctx), B.tile(ilk, jl), B.mb);
                else
                    rc = trsm_left_rec_entry(ctx, s, ul, false, tk, nc, lalpha, Tkk, Tm.mb, ctx->wInv + (size_t)set * ctx->wInvSlot,
                                             B.tile(ilk, jl), B.mb);
            }
            if (rc) break;
        }
        tmark(s);
        // (c) solved tile row k, TRANSPOSED (nc x tk, ld = B.nb), down the grid columns: as the second operand of the
        //     updates C(m,:) -= T(m,k) X(k,:) the transposed tile is an M-major operand (tensor-map TMA path of the
        //     engine) where the tile itself would be a K-major one (128-byte bulk rows, measured 19 TFLOP/s).  The buffer
        //     of this parity was last read by the bulk update of step kk-2.
        double* xrow = ctx->wSolve + (size_t)set * xSet;
        if (kk >= 2) DPB_CUDA(cudaStreamWaitEvent(s, evBulk[set], 0));
        if (Tm.prow == kp)
            for (int jl = 0; jl < colEndB && !rc; jl++)
                rc = launch_transpose(ctx, s, B.tile(ilk, jl), B.mb, xrow + (size_t)jl * teB, B.nb, tk, B.colsOfTile(jl * B.Q + B.pcol));
        if (rc) break;
        if (Tm.P > 1 && colEndB > 0) {
            group_start(ctx);
            for (int jl = 0; jl < colEndB && !rc; jl++)
                rc = bcast(ctx, ctx->commCol, xrow + (size_t)jl * teB, xrow + (size_t)jl * teB, teB, kp, s);
            if (group_end(ctx)) rc = -313;
            if (rc) break;
        }
        DPB_CUDA(cudaEventRecord(evX[set], s));
        tmark(s);
        // (d) update my remaining RHS tile rows [u0, u1): the row the next step solves here on the chain, the others
        //     on the update stream
        const int u0 = lower ? std::min(Tm.firstLocalRowGE(k + 1), rowEnd) : 0;
        const int u1 = lower ? rowEnd : std::min(Tm.firstLocalRowGE(k), rowEnd);
        const int kn = lower ? k + 1 : k - 1;  // global tile row of the next step
        const bool nextHere = kk + 1 < mt && kn % Tm.P == Tm.prow && u1 > u0;
        int la0 = u0, la1 = u0, b0 = u0, b1 = u1;  // look-ahead rows [la0, la1), bulk rows [b0, b1)
        if (nextHere) {
            if (lower) { la0 = u0; la1 = u0 + 1; b0 = u0 + 1; }
            else { la0 = u1 - 1; la1 = u1; b1 = u1 - 1; }
        }
        auto update_rows = [&](cudaStream_t st, int a, int b) -> int {
            if (b <= a) return 0;
            const int lastGm = (b - 1) * Tm.P + Tm.prow;
            int r = 0;
            for (int jl = 0; jl < colEndB && !r; jl++)
                r = gemm_single(ctx, st, B.mb, B.colsOfTile(jl * B.Q + B.pcol), tk, -1.0, rowPtr + (size_t)a * te, Tm.mb,
                                false, xrow + (size_t)jl * teB, B.nb, false, lalpha, B.tile(a, jl), B.mb, 0, b - a,
                                (long long)te, 0, (long long)teB, B.rowsOfTile(lastGm));
            return r;
        };
        if (la1 > la0) {
            // the bulk update of the previous step also writes this row: it has to be through first
            if (kk >= 1) DPB_CUDA(cudaStreamWaitEvent(s, evBulk[set ^ 1], 0));
            rc = update_rows(s, la0, la1);
            if (rc) break;
        }
        DPB_CUDA(cudaEventRecord(evStep[set], s));
        tmark(s);
        DPB_CUDA(cudaStreamWaitEvent(su, evX[set], 0));
        tmark(su);
        rc = update_rows(su, b0, b1);
        if (rc) break;
        tmark(su);
        DPB_CUDA(cudaEventRecord(evBulk[set], su));
        // panel set `set` is free again when both the chain and the bulk update of this step have read it
        DPB_CUDA(cudaStreamWaitEvent(su, evStep[set], 0));
        DPB_CUDA(cudaEventRecord(evFree[set], su));
    }
    // everything back on the caller's stream
    for (int e = 0; e < 2; e++) DPB_CUDA(cudaStreamWaitEvent(s, evBulk[e], 0));
    if (traceSolve && !tev.empty()) {
        cudaStreamSynchronize(s);
        cudaStreamSynchronize(su);
        double sum[5] = {0, 0, 0, 0, 0};  // solve, send, look-ahead, bulk, start-to-start
        const size_t steps = tev.size() / 6;
        for (size_t q = 0; q < steps; q++) {
            float ms;
            cudaEventElapsedTime(&ms, tev[6 * q], tev[6 * q + 1]); sum[0] += ms;
            cudaEventElapsedTime(&ms, tev[6 * q + 1], tev[6 * q + 2]); sum[1] += ms;
            cudaEventElapsedTime(&ms, tev[6 * q + 2], tev[6 * q + 3]); sum[2] += ms;
            cudaEventElapsedTime(&ms, tev[6 * q + 4], tev[6 * q + 5]); sum[3] += ms;
            if (q + 1 < steps) { cudaEventElapsedTime(&ms, tev[6 * q], tev[6 * q + 6]); sum[4] += ms; }
        }
        if (ctx->rank == 0)
            fprintf(stderr, "[solve]   %zu steps (%s): row solve %.2f ms, row send %.2f ms, look-ahead update (incl. wait for the "
                            "previous bulk) %.2f ms, bulk updates %.2f ms, step starts span %.2f ms\n",
                    steps, lower ? "forward" : "backward", sum[0], sum[1], sum[2], sum[3], sum[4]);
        for (cudaEvent_t e : tev) cudaEventDestroy(e);
    }
    // the communication stream must be idle before the caller reuses or frees the operands
    cudaEvent_t evDone = dist_event(ctx, evBase + 5);
    DPB_CUDA(cudaEventRecord(evDone, sc));
    DPB_CUDA(cudaStreamWaitEvent(s, evDone, 0));
    return rc;
}

int trsm_left_multi(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, double alpha,
                    const DevMatrix& A, const DevMatrix& B) {
    if (A.P != B.P || A.Q != B.Q || A.mb != B.mb || A.ip != B.ip || A.jq != B.jq) {
        fprintf(stderr, "dplasma_dtrsm: A and B must share the process grid and the row tiling\n");
        return -102;
    }
    if (A.P * A.Q != ctx->nranks) return -303;
    int rc = ensure_grid_comms(ctx, A.P, A.Q, A.ip, A.jq);
    if (rc) return rc;
    // DPB_TRACE_SOLVE=1: wall-clock phases of the call on stderr (synchronises the stream: measurement runs only)
    static const bool traceSolve = getenv("DPB_TRACE_SOLVE") != nullptr;
    auto now = [&]() {
        cudaStreamSynchronize(s);
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    };
    double t0 = traceSolve ? now() : 0.0;
    if (trans == dplasmaNoTrans) {
        rc = solve_panels_multi(ctx, s, uplo == dplasmaLower, diag, alpha, A, B);
        if (traceSolve && ctx->rank == 0)
            fprintf(stderr, "[solve] N=%d nrhs=%d %dx%d NoTrans: panels %.2f ms\n", A.M, B.N, A.P, A.Q, (now() - t0) * 1e3);
        return rc;
    }
    // op(A) = A^T: solve with the transposed factor, which is triangular the other way round
    DevMatrix At;
    rc = alloc_like(A, &At);
    double t1 = traceSolve ? now() : 0.0;
    if (!rc) rc = dist_transpose_tri(ctx, s, A, uplo, At);
    double t2 = traceSolve ? now() : 0.0;
    if (!rc) rc = solve_panels_multi(ctx, s, uplo != dplasmaLower, diag, alpha, At, B);
    cudaStreamSynchronize(s);
    double t3 = traceSolve ? now() : 0.0;
    stage_free(&At);
    if (traceSolve && ctx->rank == 0)
        fprintf(stderr, "[solve] N=%d nrhs=%d %dx%d Trans: alloc %.2f ms, transposition %.2f ms, panels %.2f ms, free %.2f ms\n", A.M,
                B.N, A.P, A.Q, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (now() - t3) * 1e3);
    return rc;
}

// B <- B - A X with A symmetric, stored in `uplo`, everything distributed on the grid
// (dplasma_zhemm(Left, uplo, -1, A, X, 1, B), src/dplasma_zcheck.c:199).
int symm_residual_multi(parsec_context_s* ctx, cudaStream_t s, int uplo, const DevMatrix& A, const DevMatrix& X,
                        const DevMatrix& B) {
    if (A.P * A.Q != ctx->nranks) return -303;
    int rc = ensure_grid_comms(ctx, A.P, A.Q, A.ip, A.jq);
    if (!rc) rc = ensure_panel_bufs(ctx, A);
    if (rc) return rc;
    DevMatrix At;
    rc = alloc_like(A, &At);
    if (!rc) rc = dist_transpose_tri(ctx, s, A, uplo, At);
    const size_t te = (size_t)A.tileElems, teB = (size_t)B.tileElems;
    const int rowEnd = std::min(A.firstLocalRowGE(A.mt), A.lmt);
    const int colEndB = std::min(B.firstLocalColGE(B.nt), B.lnt);
    double* xrow = nullptr;
    if (!rc && cudaMalloc(&xrow, (size_t)std::max(colEndB, 1) * teB * sizeof(double)) != cudaSuccess) rc = -200;
    PanelBufs pb = panel_bufs(ctx, A, 0);
    for (int n = 0; n < A.nt && !rc; n++) {
        const int np = n % A.P, nq = n % A.Q, jln = n / A.Q;
        const int tn = A.colsOfTile(n);
        // the full tile column n of the symmetric matrix, staged on the ranks of grid column nq
        if (A.pcol == nq) {
            // stored part comes from A, the mirrored part from the transposed copy
            const int split = std::min(A.firstLocalRowGE(n), rowEnd);  // local rows [0,split) have gm < n
            const DevMatrix& lowSrc = (uplo == dplasmaLower) ? At : A;   // rows above the diagonal tile
            const DevMatrix& highSrc = (uplo == dplasmaLower) ? A : At;  // rows from the diagonal tile down
            if (split > 0)
                DPB_CUDA(cudaMemcpyAsync(pb.row, lowSrc.tile(0, jln), (size_t)split * te * sizeof(double),
                                         cudaMemcpyDeviceToDevice, s));
            if (rowEnd > split)
                DPB_CUDA(cudaMemcpyAsync(pb.row + (size_t)split * te, highSrc.tile(split, jln),
                                         (size_t)(rowEnd - split) * te * sizeof(double), cudaMemcpyDeviceToDevice, s));
            if (A.prow == np)  // diagonal tile: mirror its stored triangle
                rc = launch_copy_tril(ctx, s, A.tile(n / A.P, jln), A.mb, pb.row + (size_t)(n / A.P) * te, A.mb, tn, uplo, 1);
        }
        if (!rc && A.Q > 1 && rowEnd > 0) rc = bcast(ctx, ctx->commRow, pb.row, pb.row, (size_t)rowEnd * te, nq, s);
        if (rc) break;
        const int iln = n / A.P;
        const double* xPtr = xrow;
        long long xStride = (long long)teB;
        if (A.P == 1) { xPtr = X.tile(iln, 0); xStride = (long long)X.lmt * teB; }
        else if (colEndB > 0) {
            group_start(ctx);
            for (int jl = 0; jl < colEndB && !rc; jl++)
                rc = bcast(ctx, ctx->commCol, X.tile(std::min(iln, std::max(X.lmt - 1, 0)), jl), xrow + (size_t)jl * teB, teB, np, s);
            if (group_end(ctx)) rc = -313;
        }
        if (rowEnd > 0) {
            const int lastGm = (rowEnd - 1) * A.P + A.prow;
            for (int jl = 0; jl < colEndB && !rc; jl++)
                rc = gemm_single(ctx, s, B.mb, B.colsOfTile(jl * B.Q + B.pcol), tn, -1.0, pb.row, A.mb, false,
                                 xPtr + (size_t)jl * xStride, B.mb, true, 1.0, B.tile(0, jl), B.mb, 0, rowEnd, (long long)te,
                                 0, (long long)teB, B.rowsOfTile(lastGm));
        }
    }
    cudaStreamSynchronize(s);
    if (xrow) cudaFree(xrow);
    stage_free(&At);
    return rc;
}

}  // namespace dpb

// Host-only view of the plan, for tests of the multi-rank logic without GPUs.
// out[0..5] = kq, kp, il0, nrows, jl0, ncols ; colRoot / colSrcIl must hold lnt ints.
// The plan of one round of the distributed transposition (dist_transpose_tri) for logical rank `me` of a P x Q grid:
// *nB = end of the round that starts at tile column nA; sendTiles / recvTiles = (m, n, peer) triples in the order the
// tiles are packed / unpacked (send list: peer == me for tiles that stay).  Returns 0, or -1 when a list is too short.
extern "C" long long dplasma_b200_transpose_cap(int mt, long long tileBytes) {
    return (long long)dpb::XposePlan::cap_for(mt, (size_t)tileBytes);
}
extern "C" int dplasma_b200_transpose_plan(int P, int Q, int mt, int uplo, long long cap, int me, int nA, int* nB,
                                           int* sendTiles, int maxSend, int* nSend, int* recvTiles, int maxRecv, int* nRecv) {
    if (P < 1 || Q < 1 || mt < 0 || me < 0 || me >= P * Q || nA < 0 || nA > mt || cap < 1 || !nB || !nSend || !nRecv) return -2;
    dpb::XposePlan plan{P, Q, mt, uplo == dplasmaLower, (size_t)cap};
    *nB = plan.round_end(nA);
    std::vector<dpb::XposeTile> sends, recvs;
    plan.send_list(me, nA, *nB, &sends);
    plan.recv_list(me, nA, *nB, &recvs);
    *nSend = (int)sends.size();
    *nRecv = (int)recvs.size();
    if ((int)sends.size() > maxSend || (int)recvs.size() > maxRecv) return -1;
    for (size_t q = 0; q < sends.size(); q++) { sendTiles[3 * q] = sends[q].m; sendTiles[3 * q + 1] = sends[q].n; sendTiles[3 * q + 2] = sends[q].peer; }
    for (size_t q = 0; q < recvs.size(); q++) { recvTiles[3 * q] = recvs[q].m; recvTiles[3 * q + 1] = recvs[q].n; recvTiles[3 * q + 2] = recvs[q].peer; }
    return 0;
}

extern "C" int dplasma_b200_panel_plan(const parsec_tiled_matrix_t* A, int k, int gmFirst, int* out, int* colRoot,
                                       int* colSrcIl) {
    using namespace dpb;
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    DevMatrix dm;
    dm.mb = A->mb; dm.nb = A->nb; dm.tileElems = (long long)A->mb * A->nb;
    dm.M = A->m; dm.N = A->n; dm.mt = A->mt; dm.nt = A->nt;
    dm.P = bc->grid.rows; dm.Q = bc->grid.cols;
    dm.prow = (bc->grid.rrank - bc->grid.ip + dm.P) % dm.P;
    dm.pcol = (bc->grid.crank - bc->grid.jq + dm.Q) % dm.Q;
    dm.lmt = bc->nb_elem_r; dm.lnt = bc->nb_elem_c;
    PanelPlan p;
    compute_panel_plan(dm, k, gmFirst, &p);
    out[0] = p.kq; out[1] = p.kp; out[2] = p.il0; out[3] = p.nrows; out[4] = p.jl0; out[5] = p.ncols;
    for (int c = 0; c < p.ncols; c++) {
        colRoot[c] = p.colRoot[c];
        colSrcIl[c] = p.colSrcIl[c];
    }
    return 0;
}

// Host-only plan of step k of the distributed inversion DAGs (src/ztrtri_L.jdf, src/zlauum_L.jdf): the operands of
// a step are tile ROW k to the left of the diagonal (n < k), which the owners in grid row kp send down their grid
// columns, and -- for the lauum products A(m,n) += A(k,m)^T A(k,n) -- the same tiles re-sent along the grid row of
// their transposed position.  out[0..3] = kp, kq, ncolsBefore (local tile columns with global n < k),
// nrowsBefore (local tile rows with global m < k); for each such local tile row il (global m): rowRootCol[il] =
// the grid column m % Q that holds A(k,m) after the column broadcast, rowSrcJl[il] = m / Q, its local column there.
extern "C" int dplasma_b200_rowpanel_plan(const parsec_tiled_matrix_t* A, int k, int* out, int* rowRootCol, int* rowSrcJl) {
    using namespace dpb;
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    DevMatrix dm;
    dm.mb = A->mb; dm.nb = A->nb;
    dm.M = A->m; dm.N = A->n; dm.mt = A->mt; dm.nt = A->nt;
    dm.P = bc->grid.rows; dm.Q = bc->grid.cols;
    dm.prow = (bc->grid.rrank - bc->grid.ip + dm.P) % dm.P;
    dm.pcol = (bc->grid.crank - bc->grid.jq + dm.Q) % dm.Q;
    dm.lmt = bc->nb_elem_r; dm.lnt = bc->nb_elem_c;
    out[0] = k % dm.P;
    out[1] = k % dm.Q;
    out[2] = std::min(dm.firstLocalColGE(k), dm.lnt);
    out[3] = std::min(dm.firstLocalRowGE(k), dm.lmt);
    for (int il = 0; il < out[3]; il++) {
        const int gm = il * dm.P + dm.prow;
        rowRootCol[il] = gm % dm.Q;
        rowSrcJl[il] = gm / dm.Q;
    }
    return 0;
}
