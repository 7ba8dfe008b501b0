// Host-side runtime of libdplasma_b200: the per-process context that takes PaRSEC's place on
// this path (streams, workspaces, taskpool queue), the 2D block-cyclic descriptors, and the
// staging of caller matrices (host / sym-packed / LAPACK storage) into the regular tile grid
// the kernels work on (the role of stage_in/stage_out in src/dplasmajdf_lapack_dtt.h:131-275).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include "internal.h"

using namespace dpb;

extern "C" {

const char* dplasma_b200_version(void) { return "dplasma_b200 0.1 (sm_100a)"; }

parsec_context_t* parsec_init(int nb_cores, int* pargc, char** pargv[]) {
    (void)nb_cores; (void)pargc; (void)pargv;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        fprintf(stderr, "dplasma_b200: no CUDA device -- this library has no CPU fallback\n");
        return nullptr;
    }
    parsec_context_s* ctx = new parsec_context_s();
    cudaGetDevice(&ctx->device);
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);  // hi is the numerically lowest = highest priority
    bool ok = cudaStreamCreateWithPriority(&ctx->sPanel, cudaStreamNonBlocking, hi) == cudaSuccess &&
              cudaStreamCreateWithPriority(&ctx->sUpdate, cudaStreamNonBlocking, lo) == cudaSuccess &&
              cudaStreamCreateWithPriority(&ctx->sComm, cudaStreamNonBlocking, hi) == cudaSuccess &&
              cudaStreamCreateWithPriority(&ctx->sCopy, cudaStreamNonBlocking, lo) == cudaSuccess &&
              cudaEventCreate(&ctx->evStart) == cudaSuccess && cudaEventCreate(&ctx->evStop) == cudaSuccess &&
              cudaMalloc(&ctx->dInfo, 64) == cudaSuccess &&
              cudaHostAlloc(&ctx->hInfo, 64, cudaHostAllocDefault) == cudaSuccess &&
              gemm_init_attributes() == 0 && zgemm_init_attributes() == 0;
    if (!ok) {
        fprintf(stderr, "dplasma_b200: context creation failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return nullptr;
    }
    cudaMemset(ctx->dInfo, 0, 64);
    return ctx;
}

int parsec_fini(parsec_context_t** pctx) {
    if (!pctx || !*pctx) return -1;
    parsec_context_s* ctx = *pctx;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    comm_destroy(ctx);
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    for (cudaEvent_t e : ctx->evPool) cudaEventDestroy(e);
    if (ctx->wInv) cudaFree(ctx->wInv);
    if (ctx->wNorm) cudaFree(ctx->wNorm);
    if (ctx->wSolve) cudaFree(ctx->wSolve);
    if (ctx->wXpose) cudaFree(ctx->wXpose);
    if (ctx->wSolveInv) cudaFree(ctx->wSolveInv);
    if (ctx->panelBase) cudaFree(ctx->panelBase);
    if (ctx->tileSync) cudaFree(ctx->tileSync);
    for (int i = 0; i < 2; i++)
        if (ctx->evH2d[i]) cudaEventDestroy(ctx->evH2d[i]);
    cudaFree(ctx->dInfo);
    cudaFreeHost(ctx->hInfo);
    cudaEventDestroy(ctx->evStart);
    cudaEventDestroy(ctx->evStop);
    cudaStreamDestroy(ctx->sPanel);
    cudaStreamDestroy(ctx->sUpdate);
    cudaStreamDestroy(ctx->sComm);
    cudaStreamDestroy(ctx->sCopy);
    delete ctx;
    *pctx = nullptr;
    return 0;
}

void* parsec_data_allocate(size_t bytes) {
    void* p = nullptr;
    if (bytes == 0) bytes = 8;
    if (cudaHostAlloc(&p, bytes, cudaHostAllocDefault) != cudaSuccess) {
        fprintf(stderr, "dplasma_b200: pinned allocation of %zu bytes failed\n", bytes);
        return nullptr;
    }
    return p;
}
void parsec_data_free(void* p) {
    if (p) cudaFreeHost(p);
}
void* dplasma_b200_device_allocate(size_t bytes) {
    void* p = nullptr;
    if (bytes == 0) bytes = 8;
    if (cudaMalloc(&p, bytes) != cudaSuccess) {
        fprintf(stderr, "dplasma_b200: device allocation of %zu bytes failed\n", bytes);
        return nullptr;
    }
    return p;
}
void dplasma_b200_device_free(void* p) {
    if (p) cudaFree(p);
}

double dplasma_b200_last_elapsed(parsec_context_t* ctx) { return ctx ? ctx->lastElapsed : 0.0; }
double dplasma_b200_last_check_result(parsec_context_t* ctx) { return ctx ? ctx->lastCheck : 0.0; }
long long dplasma_b200_kernel_launches(parsec_context_t* ctx) { return ctx ? ctx->launches : 0; }

double dplasma_b200_flops_dpotrf(double n) { return n * n * n / 3.0 + n * n / 2.0 + n / 6.0; }
double dplasma_b200_flops_dpotrs(double n, double nrhs) { return 2.0 * nrhs * n * n; }

// ---------------------------------------------------------------------------------------
// descriptors
// ---------------------------------------------------------------------------------------
static int local_tiles(int nt, int P, int p) { return nt > p ? (nt - p + P - 1) / P : 0; }

static void init_common(parsec_matrix_block_cyclic_t* dc, int mtype, int storage, int myrank, int mb, int nb,
                        int lm, int ln, int i, int j, int m, int n, int P, int Q, int kp, int kq, int ip,
                        int jq) {
    memset(dc, 0, sizeof(*dc));
    parsec_tiled_matrix_t* t = &dc->super;
    if (P < 1) P = 1;
    if (Q < 1) Q = 1;
    t->super.nodes = (uint32_t)(P * Q);
    t->super.myrank = (uint32_t)myrank;
    t->mtype = mtype;
    t->storage = storage;
    t->mb = mb; t->nb = nb; t->bsiz = mb * nb;
    t->lm = lm; t->ln = ln;
    t->lmt = (lm + mb - 1) / mb; t->lnt = (ln + nb - 1) / nb;
    t->i = i; t->j = j; t->m = m; t->n = n;
    t->mt = (m + mb - 1) / mb; t->nt = (n + nb - 1) / nb;
    dc->grid.rank = myrank; dc->grid.rows = P; dc->grid.cols = Q;
    dc->grid.rrank = myrank / Q; dc->grid.crank = myrank % Q;  // row-major grid (Cblacs_gridinit "Row")
    dc->grid.krows = kp; dc->grid.kcols = kq; dc->grid.ip = ip; dc->grid.jq = jq;
    const int prow = (dc->grid.rrank - ip + P) % P, pcol = (dc->grid.crank - jq + Q) % Q;
    dc->nb_elem_r = local_tiles(t->lmt, P, prow);
    dc->nb_elem_c = local_tiles(t->lnt, Q, pcol);
    // local element counts (numroc)
    int lr = 0, lc = 0;
    for (int g = prow; g < t->lmt; g += P) lr += (g == t->lmt - 1) ? lm - g * mb : mb;
    for (int g = pcol; g < t->lnt; g += Q) lc += (g == t->lnt - 1) ? ln - g * nb : nb;
    t->llm = lr; t->lln = lc; t->slm = lr; t->sln = lc;
    t->nb_local_tiles = dc->nb_elem_r * dc->nb_elem_c;
    dc->uplo = dplasmaUpperLower;
    dc->mat = nullptr;
}

void parsec_matrix_block_cyclic_init(parsec_matrix_block_cyclic_t* dc, int mtype, int storage, int myrank,
                                     int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P, int Q,
                                     int kp, int kq, int ip, int jq) {
    init_common(dc, mtype, storage, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq);
}

void parsec_matrix_sym_block_cyclic_init(parsec_matrix_sym_block_cyclic_t* dc, int mtype, int myrank, int mb,
                                         int nb, int lm, int ln, int i, int j, int m, int n, int P, int Q,
                                         int uplo) {
    init_common(dc, mtype, PARSEC_MATRIX_TILE, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, 1, 1, 0, 0);
    dc->uplo = uplo;
    // count the tiles of the stored triangle that live here
    int cnt = 0;
    const int prow = dc->grid.rrank, pcol = dc->grid.crank;
    for (int gn = pcol; gn < dc->super.lnt; gn += Q)
        for (int gm = prow; gm < dc->super.lmt; gm += P)
            if (uplo == dplasmaLower ? gm >= gn : gm <= gn) cnt++;
    dc->super.nb_local_tiles = cnt;
}

void parsec_matrix_block_cyclic_lapack_init(parsec_matrix_block_cyclic_t* dc, int mtype, int storage, int myrank,
                                            int mb, int nb, int lm, int ln, int i, int j, int m, int n, int P,
                                            int Q, int kp, int kq, int ip, int jq, int mloc, int nloc) {
    init_common(dc, mtype, storage, myrank, mb, nb, lm, ln, i, j, m, n, P, Q, kp, kq, ip, jq);
    dc->super.llm = mloc;  // leading dimension of the caller's local array (LLD)
    dc->super.lln = nloc;
}

void parsec_tiled_matrix_destroy(parsec_tiled_matrix_t* dc) { (void)dc; }

int dplasma_b200_rank_of(const parsec_tiled_matrix_t* A, int m, int n) {
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    const int pr = (m + bc->grid.ip) % bc->grid.rows, pc = (n + bc->grid.jq) % bc->grid.cols;
    return pr * bc->grid.cols + pc;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------
// staging
// ---------------------------------------------------------------------------------------
namespace dpb {

static bool is_device_ptr(const void* p) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

static int describe(const parsec_tiled_matrix_t* A, DevMatrix* dm) {
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    if (A->mtype != PARSEC_MATRIX_DOUBLE && A->mtype != PARSEC_MATRIX_COMPLEX_DOUBLE) return -1;
    dm->esz = (A->mtype == PARSEC_MATRIX_COMPLEX_DOUBLE) ? 2 : 1;
    if (bc->grid.krows != 1 || bc->grid.kcols != 1) return -2;
    dm->mb = A->mb; dm->nb = A->nb; dm->tileElems = (long long)A->mb * A->nb * dm->esz;
    dm->M = A->m; dm->N = A->n; dm->mt = A->mt; dm->nt = A->nt;
    dm->P = bc->grid.rows; dm->Q = bc->grid.cols;
    const bool whole = A->i == 0 && A->j == 0 && A->m == A->lm && A->n == A->ln;
    if (whole) {
        dm->prow = (bc->grid.rrank - bc->grid.ip + dm->P) % dm->P;
        dm->pcol = (bc->grid.crank - bc->grid.jq + dm->Q) % dm->Q;
        dm->ip = bc->grid.ip % dm->P; dm->jq = bc->grid.jq % dm->Q;
        dm->lmt = bc->nb_elem_r; dm->lnt = bc->nb_elem_c;
        dm->offR = dm->offC = 0; dm->userLmt = bc->nb_elem_r;
        return 0;
    }
    // Submatrix A(i:i+m, j:j+n): tile-aligned offsets only, like the reference's pdpotrf wrapper
    // (src/scalapack_wrappers/dplasma_wrapper_pdpotrf.c:209-211); not for sym-packed storage.
    if (A->i < 0 || A->j < 0 || A->i % A->mb != 0 || A->j % A->nb != 0 || A->i + A->m > A->lm || A->j + A->n > A->ln) return -3;
    if (bc->uplo != dplasmaUpperLower) return -3;
    if (A->i + A->m < A->lm && A->m % A->mb != 0) return -3;  // an interior submatrix must end on a tile boundary
    if (A->j + A->n < A->ln && A->n % A->nb != 0) return -3;
    const int i0 = A->i / A->mb, j0 = A->j / A->nb;
    dm->ip = (bc->grid.ip + i0) % dm->P; dm->jq = (bc->grid.jq + j0) % dm->Q;
    dm->prow = (bc->grid.rrank - dm->ip + dm->P) % dm->P;
    dm->pcol = (bc->grid.crank - dm->jq + dm->Q) % dm->Q;
    dm->lmt = local_tiles(dm->mt, dm->P, dm->prow);
    dm->lnt = local_tiles(dm->nt, dm->Q, dm->pcol);
    dm->offR = (dm->prow + i0) / dm->P; dm->offC = (dm->pcol + j0) / dm->Q;
    dm->userLmt = bc->nb_elem_r;
    return 0;
}

// Copies local tile rows [i0, i1) of local tile column jl between the caller's layout and device tiles laid out
// one after the other from devBase (tile of row i at devBase + (i - i0) * tileElems).  dir 0: in, 1: out.
// For sym-packed storage the rows must lie inside the stored triangle.
static int copy_column_rows(const parsec_tiled_matrix_t* A, const DevMatrix* dm, int jl, int i0, int i1, double* devBase,
                            int dir, cudaStream_t s) {
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    double* user = (double*)bc->mat;
    const bool sym = bc->uplo != dplasmaUpperLower;
    const int gn = jl * dm->Q + dm->pcol;
    if (i1 > dm->lmt) i1 = dm->lmt;
    if (i0 >= i1) return 0;
    if (A->storage == PARSEC_MATRIX_TILE) {
        size_t off;  // tile offset of (i0, jl) in the caller's array
        if (sym) {
            size_t packed = 0;
            for (int j = 0; j < jl; j++) {
                const int g = j * dm->Q + dm->pcol;
                int a = 0, b = dm->lmt;
                if (bc->uplo == dplasmaLower) a = dm->firstLocalRowGE(g);
                else b = std::min(dm->firstLocalRowGE(g + 1), dm->lmt);
                if (b > a) packed += (size_t)(b - a);
            }
            const int first = (bc->uplo == dplasmaLower) ? dm->firstLocalRowGE(gn) : 0;
            if (i0 < first) return -120;
            off = packed + (size_t)(i0 - first);
        } else {
            off = (size_t)(i0 + dm->offR) + (size_t)(jl + dm->offC) * dm->userLmt;
        }
        const size_t cnt = (size_t)(i1 - i0) * dm->tileElems;
        double* usr = user + off * dm->tileElems;
        if (dir == 0) DPB_CUDA(cudaMemcpyAsync(devBase, usr, cnt * sizeof(double), cudaMemcpyDefault, s));
        else DPB_CUDA(cudaMemcpyAsync(usr, devBase, cnt * sizeof(double), cudaMemcpyDefault, s));
        return 0;
    }
    // LAPACK storage: pitched copies, one per tile (src/dplasmajdf_lapack_dtt.h:161-178,238-255)
    const size_t lld = (size_t)A->llm, eb = 8 * (size_t)dm->esz;
    const int tn = dm->colsOfTile(gn);
    for (int il = i0; il < i1; il++) {
        const int gm = il * dm->P + dm->prow;
        const int tm = dm->rowsOfTile(gm);
        double* dev = devBase + (size_t)(il - i0) * dm->tileElems;
        double* usr = user + ((size_t)(il + dm->offR) * dm->mb + (size_t)(jl + dm->offC) * dm->nb * lld) * dm->esz;
        if (dir == 0)
            DPB_CUDA(cudaMemcpy2DAsync(dev, (size_t)dm->mb * eb, usr, lld * eb, (size_t)tm * eb, tn, cudaMemcpyDefault, s));
        else
            DPB_CUDA(cudaMemcpy2DAsync(usr, lld * eb, dev, (size_t)dm->mb * eb, (size_t)tm * eb, tn, cudaMemcpyDefault, s));
    }
    return 0;
}
int stage_column_rows(const parsec_tiled_matrix_t* A, const DevMatrix* dm, int jl, int i0, int i1, double* devBase, int dir,
                      cudaStream_t s) {
    return copy_column_rows(A, dm, jl, i0, i1, devBase, dir, s);
}
int describe_matrix(const parsec_tiled_matrix_t* A, DevMatrix* dm) { return describe(A, dm); }

// Copies one local tile column between the caller's layout and the regular device tile grid.
// dir 0: in, 1: out.  Only the `part` triangle moves (dplasmaUpperLower: everything).
static int copy_column(const parsec_tiled_matrix_t* A, DevMatrix* dm, int jl, int dir, int part, cudaStream_t s) {
    const int gn = jl * dm->Q + dm->pcol;
    int i0 = 0, i1 = dm->lmt;
    if (part == dplasmaLower) i0 = dm->firstLocalRowGE(gn);
    else if (part == dplasmaUpper) i1 = dm->firstLocalRowGE(gn + 1);
    if (i1 > dm->lmt) i1 = dm->lmt;
    if (i0 >= i1) return 0;
    return copy_column_rows(A, dm, jl, i0, i1, dm->tile(i0, jl), dir, s);
}

static int copy_layout(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int dir, int uploOnly) {
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    const bool sym = bc->uplo != dplasmaUpperLower;
    const int part = sym ? bc->uplo : uploOnly;  // which tiles move
    for (int jl = 0; jl < dm->lnt; jl++) {
        int rc = copy_column(A, dm, jl, dir, part, ctx->sPanel);
        if (rc) return rc;
    }
    return 0;
}

// Early write-back of one finished tile column on the copy stream (overlaps the rest of the factorization).
int stage_out_column(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int jl, int uploOnly,
                     cudaEvent_t ready) {
    if (!dm->owned) return 0;
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    const int part = (bc->uplo != dplasmaUpperLower) ? bc->uplo : uploOnly;
    DPB_CUDA(cudaStreamWaitEvent(ctx->sCopy, ready, 0));
    return copy_column(A, dm, jl, 1, part, ctx->sCopy);
}

int stage_in(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, bool copyData, int uploOnly) {
    int rc = describe(A, dm);
    if (rc) {
        fprintf(stderr, "dplasma_b200: unsupported descriptor (code %d)\n", rc);
        return -100 + rc;
    }
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    if (!bc->mat) return -110;
    // in place only for whole matrices: a submatrix is staged (its local tiles are not a dense tile grid)
    const bool regular = A->storage == PARSEC_MATRIX_TILE && bc->uplo == dplasmaUpperLower && dm->offR == 0 &&
                         dm->offC == 0 && dm->userLmt == dm->lmt;
    if (regular && is_device_ptr(bc->mat)) {
        dm->base = (double*)bc->mat;
        dm->owned = false;
        return 0;
    }
    size_t elems = (size_t)dm->lmt * dm->lnt * dm->tileElems;
    if (elems == 0) elems = 1;
    DPB_CUDA(cudaMalloc(&dm->base, elems * sizeof(double)));
    dm->owned = true;
    if (copyData) return copy_layout(ctx, A, dm, 0, uploOnly);
    return 0;
}

// Stage-in that does not wait: tile columns travel left to right on the copy stream, one event each,
// so that a left-looking start of the factorization can consume them as they arrive.
int stage_in_pipelined(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int uploOnly) {
    int rc = describe(A, dm);
    if (rc) return -100 + rc;
    const parsec_matrix_block_cyclic_t* bc = (const parsec_matrix_block_cyclic_t*)A;
    if (!bc->mat) return -110;
    size_t elems = (size_t)dm->lmt * dm->lnt * dm->tileElems;
    if (elems == 0) elems = 1;
    DPB_CUDA(cudaMalloc(&dm->base, elems * sizeof(double)));
    dm->owned = true;
    const int part = (bc->uplo != dplasmaUpperLower) ? bc->uplo : uploOnly;
    for (cudaEvent_t e : ctx->evIn) cudaEventDestroy(e);
    ctx->evIn.clear();
    // the copy rate of this stage-in feeds the arrival model of the next one (ArrivalModel, internal.h)
    for (int i = 0; i < 2; i++)
        if (!ctx->evH2d[i]) DPB_CUDA(cudaEventCreate(&ctx->evH2d[i]));
    const int rowEnd = std::min(dm->firstLocalRowGE(dm->mt), dm->lmt);
    double bytes = 0.0;
    DPB_CUDA(cudaEventRecord(ctx->evH2d[0], ctx->sCopy));
    for (int jl = 0; jl < dm->lnt; jl++) {
        rc = copy_column(A, dm, jl, 0, part, ctx->sCopy);
        if (rc) return rc;
        cudaEvent_t e;
        DPB_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        DPB_CUDA(cudaEventRecord(e, ctx->sCopy));
        ctx->evIn.push_back(e);
        const int gn = jl * dm->Q + dm->pcol;
        int i0 = 0, i1 = rowEnd;
        if (part == dplasmaLower) i0 = std::min(dm->firstLocalRowGE(gn), rowEnd);
        else if (part == dplasmaUpper) i1 = std::min(dm->firstLocalRowGE(gn + 1), rowEnd);
        bytes += (double)std::max(0, i1 - i0) * dm->tileElems * 8.0;
    }
    DPB_CUDA(cudaEventRecord(ctx->evH2d[1], ctx->sCopy));
    ctx->h2dBytes = bytes;
    return 0;
}

int stage_out(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int uploOnly) {
    if (!dm->owned) return 0;
    return copy_layout(ctx, A, dm, 1, uploOnly);
}

void stage_free(DevMatrix* dm) {
    if (dm->owned && dm->base) cudaFree(dm->base);
    dm->base = nullptr;
    dm->owned = false;
}

}  // namespace dpb
