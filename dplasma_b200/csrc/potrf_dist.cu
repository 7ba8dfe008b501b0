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

static int ensure_panel_bufs(parsec_context_s* ctx, const DevMatrix& A) {
    const size_t diag = (size_t)A.tileElems + (size_t)((A.nb + IB - 1) / IB) * IB * IB;
    const size_t need = (size_t)(A.lmt + A.lnt) * A.tileElems + diag;
    if (ctx->panelBuf[0] && ctx->panelBufElems >= need) return 0;
    for (int i = 0; i < 2; i++) {
        if (ctx->panelBuf[i]) cudaFree(ctx->panelBuf[i]);
        ctx->panelBuf[i] = nullptr;
        DPB_CUDA(cudaMalloc(&ctx->panelBuf[i], need * sizeof(double)));
    }
    ctx->panelBufElems = need;
    return 0;
}
struct PanelBufs {
    double* row;   // lmt tiles, indexed by local tile row
    double* col;   // lnt tiles, indexed by local tile column
    double* diag;  // one tile followed by the inverted diagonal blocks
};
static PanelBufs panel_bufs(parsec_context_s* ctx, const DevMatrix& A, int set) {
    PanelBufs b;
    b.row = ctx->panelBuf[set];
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

int potrf_multi(parsec_context_s* ctx, parsec_taskpool_s* tp) {
    const DevMatrix& A = tp->dA;
    if (A.P * A.Q != ctx->nranks) {
        fprintf(stderr, "dplasma_b200: descriptor grid %dx%d does not match the %d initialised ranks\n", A.P, A.Q,
                ctx->nranks);
        return -303;
    }
    int rc = ensure_grid_comms(ctx, A.P, A.Q);
    if (!rc) rc = ensure_winv(ctx, A.nb);
    if (!rc) rc = ensure_panel_bufs(ctx, A);
    if (rc) return rc;
    cudaStream_t sp = ctx->sPanel, su = ctx->sUpdate;
    const int nt = A.nt, nb = A.nb;
    const size_t te = (size_t)A.tileElems;
    const size_t winvElems = (size_t)((nb + IB - 1) / IB) * IB * IB;
    cudaEvent_t ev0 = dist_event(ctx, 0);
    DPB_CUDA(cudaEventRecord(ev0, sp));
    DPB_CUDA(cudaStreamWaitEvent(su, ev0, 0));
    size_t evi = 1;
    cudaEvent_t evUpdPrev = nullptr;
    const int rowEnd = std::min(A.firstLocalRowGE(A.mt), A.lmt);
    for (int k = 0; k < nt; k++) {
        const int set = k & 1;
        const int kp = k % A.P, kq = k % A.Q;
        const int tk = A.rowsOfTile(k);
        PanelBufs b = panel_bufs(ctx, A, set);
        const bool inCol = (A.pcol == kq);
        const int jlk = k / A.Q;  // local column index of tile column k on the ranks of grid column kq
        // ---- (1) diagonal tile: factor on the owner, broadcast down grid column kq ----
        if (inCol) {
            double* winv = b.diag + te;
            if (A.prow == kp) {
                double* Akk = A.tile(k / A.P, jlk);
                if (A.P == 1) {
                    // no broadcast needed: factor in place, inverses straight into the buffer
                    rc = potrf_tile_lower(ctx, sp, Akk, A.mb, tk, winv, ctx->dInfo, k * nb);
                    if (rc) return rc;
                } else {
                    rc = potrf_tile_lower(ctx, sp, Akk, A.mb, tk, winv, ctx->dInfo, k * nb);
                    if (rc) return rc;
                    DPB_CUDA(cudaMemcpyAsync(b.diag, Akk, te * sizeof(double), cudaMemcpyDeviceToDevice, sp));
                }
            }
            const double* Lkk = b.diag;
            if (A.P == 1) Lkk = A.tile(k / A.P, jlk);
            else {
                rc = bcast(ctx, ctx->commCol, b.diag, b.diag, te + winvElems, kp, sp);
                if (rc) return rc;
            }
            // ---- panel solve on my tiles of column k ----
            const int il0 = std::min(A.firstLocalRowGE(k + 1), rowEnd), cnt = rowEnd - il0;
            if (cnt > 0) {
                const int lastGm = (rowEnd - 1) * A.P + A.prow;
                rc = trsm_panel_lower(ctx, sp, Lkk, A.mb, tk, winv, A.tile(il0, jlk), A.mb, A.mb, cnt, A.tileElems,
                                      A.rowsOfTile(lastGm));
                if (rc) return rc;
            }
        }
        if (k + 1 >= nt) break;
        // ---- (2) + (3): panel broadcasts ----
        PanelView pv;
        rc = panel_exchange(ctx, sp, A, k, k + 1, inCol ? A.tile(0, jlk) : nullptr, A.tileElems, set, true, &pv);
        if (rc) return rc;
        cudaEvent_t evPanel = dist_event(ctx, evi++);
        DPB_CUDA(cudaEventRecord(evPanel, sp));
        // ---- look-ahead update of tile column k+1 on the panel stream ----
        if (evUpdPrev) DPB_CUDA(cudaStreamWaitEvent(sp, evUpdPrev, 0));
        rc = trailing_update(ctx, sp, A, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, k + 1, k + 2, -1.0, nb);
        if (rc) return rc;
        // ---- bulk update on the update stream ----
        if (k + 2 < nt) {
            DPB_CUDA(cudaStreamWaitEvent(su, evPanel, 0));
            rc = trailing_update(ctx, su, A, pv.rowPtr, pv.rowStride, pv.colPtr, pv.colStride, k + 2, nt, -1.0, nb);
            if (rc) return rc;
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
    int rc = ensure_grid_comms(ctx, L.P, L.Q);
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

}  // namespace dpb

// Host-only view of the plan, for tests of the multi-rank logic without GPUs.
// out[0..5] = kq, kp, il0, nrows, jl0, ncols ; colRoot / colSrcIl must hold lnt ints.
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
