// Internal structures of libdplasma_b200 (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <vector>
#include "../../include/dplasma_b200.h"
#include "gemm_dmma.cuh"

namespace dpb {

constexpr int IB = 64;  // leaf block of the in-tile recursion (diagonal blocks are inverted at this size)

#define DPB_CUDA(call)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            fprintf(stderr, "dplasma_b200: CUDA error %s at %s:%d: %s\n", #call, __FILE__,   \
                    __LINE__, cudaGetErrorString(e__));                                      \
            return (int)e__;                                                                 \
        }                                                                                    \
    } while (0)

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is a per-DEVICE setting: a process that drives several
// GPUs (one context per device, or the allocation-free tile entry called on each) needs it once per device.
#define DPB_SMEM_OPT_IN(kernel, bytes)                                                                        \
    do {                                                                                                      \
        static bool done__[64] = {false};                                                                     \
        int dev__ = 0;                                                                                        \
        DPB_CUDA(cudaGetDevice(&dev__));                                                                      \
        if (!done__[dev__ & 63]) {                                                                            \
            DPB_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes))); \
            done__[dev__ & 63] = true;                                                                        \
        }                                                                                                     \
    } while (0)

// Working layout in HBM: the regular local tile grid of a 2D block-cyclic matrix.
// Local tile (il, jl) <-> global tile (il*P + prow, jl*Q + pcol); tile at
// base + (il + jl*lmt) * tileElems, column-major, leading dimension mb.
struct DevMatrix {
    double* base = nullptr;
    int mb = 0, nb = 0;
    long long tileElems = 0;  // DOUBLES per tile (mb*nb*esz): all tile pointer arithmetic is in doubles
    int esz = 1;              // doubles per element: 1 real, 2 complex (interleaved re,im)
    int lmt = 0, lnt = 0;  // local tile rows / cols
    int M = 0, N = 0;      // global size
    int mt = 0, nt = 0;    // global tile counts
    int P = 1, Q = 1, prow = 0, pcol = 0;  // prow / pcol are LOGICAL grid coordinates: tile (0,0) lives on (0,0)
    int ip = 0, jq = 0;    // physical grid coordinates owning tile (0,0) (ScaLAPACK RSRC / CSRC)
    bool owned = false;    // staged copy allocated by the library
    // host-side only: where the (sub)matrix sits inside the caller's local array.  A tile-aligned submatrix
    // A(i:i+m, j:j+n) of a block-cyclic parent is itself block-cyclic with the source process shifted by
    // (i/mb, j/nb); its local tile (il, jl) is the parent's local tile (il + offR, jl + offC).
    int offR = 0, offC = 0, userLmt = 0;
    __host__ __device__ double* tile(int il, int jl) const { return base + ((size_t)il + (size_t)jl * lmt) * tileElems; }
    __host__ __device__ int rowsOfTile(int gm) const { return gm == mt - 1 ? M - gm * mb : mb; }
    __host__ __device__ int colsOfTile(int gn) const { return gn == nt - 1 ? N - gn * nb : nb; }
    // first local tile row whose global index is >= gm
    __host__ __device__ int firstLocalRowGE(int gm) const { return gm <= prow ? 0 : (gm - prow + P - 1) / P; }
    __host__ __device__ int firstLocalColGE(int gn) const { return gn <= pcol ? 0 : (gn - pcol + Q - 1) / Q; }
};

struct NcclApi;  // comm.cu

}  // namespace dpb

enum TaskKind { TP_POTRF = 1, TP_TRSM, TP_TRTRI, TP_LAUUM, TP_POINV, TP_GEMM, TP_SYRK };

struct parsec_taskpool_s {
    int kind;
    int uplo;
    parsec_tiled_matrix_t* A;
    int* info;
    int hmb;
    dpb::DevMatrix dA;
    bool staged;
    // operands of the other operations (taskpools.cu); A doubles as the first (const) operand
    int side = 0, trans = 0, trans2 = 0, diag = 0;
    double alpha = 1.0, beta = 0.0;
    const parsec_tiled_matrix_t* B = nullptr;
    parsec_tiled_matrix_t* C = nullptr;
    bool ooc = false;   // host matrix larger than the device memory it may use: out-of-core schedule (potrf_ooc.cu)
    bool done = false;  // ran to completion inside parsec_context_start (mixed queues run in order)
};

struct parsec_context_s {
    int device = 0;
    int rank = 0, nranks = 1;
    cudaStream_t sPanel = nullptr, sUpdate = nullptr, sComm = nullptr, sCopy = nullptr;
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    std::vector<cudaEvent_t> evPool;
    int* dInfo = nullptr;       // device int
    int* hInfo = nullptr;       // pinned host int
    double* wInv = nullptr;     // 2 slots, each: inverted diagonal blocks + factored diagonal blocks (2*(maxNB/IB)*IB*IB doubles)
    size_t wInvSlot = 0;
    double* wNorm = nullptr;    // row-sum scratch for norms
    double* wSolve = nullptr;   // solved RHS tile row of the grid solves
    size_t wSolveLen = 0;
    double* wSolveInv = nullptr;  // large inverse blocks of the tile solves: 2 x [n x n | scratch] + solved-tile buffer
    size_t wSolveInvLen = 0;
    int wSolveInvNb = 0;
    double* wXpose = nullptr;   // send + receive staging of the distributed transposition (grown on demand)
    size_t wXposeLen = 0;
    size_t wNormLen = 0;
    // multi-GPU panel workspaces: panelSets sets of panelSetElems doubles in one allocation (two alternate in
    // steady state; more are kept alive while a host matrix is still arriving, see potrf_dist.cu)
    double* panelBase = nullptr;
    size_t panelSetElems = 0;
    int panelSets = 0;
    // host -> device staging rate seen by the last pipelined stage-in (bytes/s, 0: not measured yet): the
    // schedule uses it to guess which tile columns will have landed by the time a step runs
    double h2dBw = 0.0;
    double h2dBytes = 0.0;
    int* tileSync = nullptr;     // arrival counter of the single-launch tile factorization (only ever incremented)
    int tileSyncCount = 0;       // arrivals accounted for by the launches enqueued so far
    cudaEvent_t evH2d[2] = {nullptr, nullptr};
    double lastElapsed = 0.0;
    double lastCheck = 0.0;
    long long launches = 0;
    std::vector<parsec_taskpool_s*> queue;
    dpb::NcclApi* nccl = nullptr;
    void* commWorld = nullptr;
    void* commRow = nullptr;
    void* commCol = nullptr;
    int gridP = 0, gridQ = 0, gridIp = 0, gridJq = 0;
    parsec_taskpool_s* curTp = nullptr;  // taskpool being enqueued (for early write-back of finished columns)
    int earlyOut = 0;                    // columns of curTp already written back by the copy stream
    std::vector<cudaEvent_t> evIn;       // per local tile column: its host-to-device copy has landed (pipelined stage-in)
    double seqElapsed = -1.0;  // >= 0: GPU seconds summed over a queue that ran operation by operation
    int cplx = 0;         // set by the drivers while they work on a complex (esz = 2) matrix
    int gemmFull = 0;     // grouped update of a general (not symmetric) matrix: no SYRK mask on diagonal tiles
    int gemmPreload = 1;  // internal products preload C into the accumulators; the tile seam API does not
    // optional per-kernel profile of the trailing-update launches (bench.py roofline leg)
    int profile = 0;
    std::vector<cudaEvent_t> profEv;   // pairs (start, stop)
    std::vector<double> profFlops;
    std::vector<int> profStream;       // 1: launched on the update stream (bulk), 0: panel stream
};

namespace dpb {
// ---------------------------------------------------------------------------------------
// Arrival model of a pipelined stage-in (host matrix still on its way, tile column by tile column, while the
// factorization runs).  All of it is a GUESS used to decide how much work to enqueue behind which copy event;
// correctness rests on the events alone (every stream waits for the copy of a column before touching it).
// A pessimistic guess only postpones the update of late columns (they are caught up later in one launch);
// an optimistic one makes a stream wait for data -- so the rate is discounted.
// ---------------------------------------------------------------------------------------
struct ArrivalModel {
    std::vector<double> landed;  // landed[jl]: modelled time at which local tile column jl is in HBM
    double now = 0.0;            // modelled time at which the GPU gets through what was enqueued so far
    double rate = 34e12;         // flop/s of the update kernel (on the fast side: `now` early = fewer columns expected)
    void init(const DevMatrix& A, double bytesPerSecond) {
        if (const char* e = getenv("DPB_H2D_GBS")) bytesPerSecond = atof(e) * 1e9 / 0.85;  // tests: force a rate
        const double bw = 0.85 * (bytesPerSecond > 1e6 ? bytesPerSecond : 20e9);
        landed.assign(A.lnt, 0.0);
        double t = 0.0;
        for (int jl = 0; jl < A.lnt; jl++) {
            t += tiles(A, jl, jl + 1) * A.tileElems * 8.0 / bw;
            landed[jl] = t;
        }
    }
    // local tiles on or below the diagonal in local tile columns [jl0, jl1)
    static double tiles(const DevMatrix& A, int jl0, int jl1) {
        const int rowEnd = std::min(A.firstLocalRowGE(A.mt), A.lmt);
        double n = 0.0;
        for (int jl = std::max(jl0, 0); jl < std::min(jl1, A.lnt); jl++)
            n += std::max(0, rowEnd - std::min(A.firstLocalRowGE(jl * A.Q + A.pcol), rowEnd));
        return n;
    }
    // `panels` rank-nb updates of the local columns [jl0, jl1) have been enqueued (either stream: they share the GPU)
    void work(const DevMatrix& A, int jl0, int jl1, int panels) {
        now += 2.0 * tiles(A, jl0, jl1) * panels * A.nb * (double)A.nb * A.nb / rate;
    }
    void overhead(double seconds) { now += seconds; }
    // a stream was made to wait for the copy of local column jl
    void waited(int jl) { if (jl >= 0 && jl < (int)landed.size()) now = std::max(now, landed[jl]); }
    // number of local columns modelled as landed by `now` (a prefix: copies are issued left to right)
    int arrived() const {
        int n = 0;
        while (n < (int)landed.size() && landed[n] <= now) n++;
        return n;
    }
    // steps a grid factorization should keep its panels for: until everything has landed even if the GPU had
    // all the work of every step available (the real window may be shorter; it is capped by memory)
    int window(const DevMatrix& A, int cap) const {
        ArrivalModel m = *this;
        int w = 0;
        while (w < cap && w < A.nt && m.arrived() < (int)landed.size()) {
            m.work(A, A.firstLocalColGE(w + 1), A.lnt, 1);
            m.overhead(1.5e-3);
            w++;
        }
        return w;
    }
};

// kernels.cu
int launch_potrf_leaf(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv,
                      int* dInfo, int infoBase);
int launch_trsm_panel_fused(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, int n, const double* winv,
                            double* B, int ldb, long long tileStride, int rowsPerTile, int lastTileRows, int ntiles);
int launch_ztranspose(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd, int m,
                      int n, int conj);
int launch_zconj(parsec_context_s* ctx, cudaStream_t s, double* p, long long ncomplex);
int launch_ztrtri_blocks(parsec_context_s* ctx, cudaStream_t s, const double* T, int ldt, int n, double* winv,
                         double* winvT);
int trsm_rec_entry(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, const double* winv, int n, double* B,
                   int ldb, int M, int batch, long long strideB, int lastBatchM);
int launch_zpotrf_leaf(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                       int infoBase);
int launch_potrf_colblock(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, int j0, double* winv,
                          double* wdiag, int* dInfo, int infoBase);
int launch_potrf_tile(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv, int* dInfo,
                      int infoBase);
int launch_potrf_diag_writeback(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, const double* wdiag);
int launch_generate(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int kind /*0 plgsy,1 plrnt*/,
                    double bump, int uplo, unsigned long long seed);
int launch_rowsum_sym(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int uplo, double* rowsum);
int launch_rowsum_ge(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, double* rowsum);
int launch_absmax(parsec_context_s* ctx, cudaStream_t s, const double* v, long long n, double* out);
int launch_copy_tril(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd,
                     int n, int uplo, int symmetrize);
int launch_fill(parsec_context_s* ctx, cudaStream_t s, double* p, long long n, double v);
int launch_scale_tile(parsec_context_s* ctx, cudaStream_t s, double* C, int ldc, int m, int n, double beta, int part);
int launch_trtri_leaf(parsec_context_s* ctx, cudaStream_t s, const double* A, int lda, int n, int uplo, int unit,
                      double* winv);
int launch_trtri_blocks(parsec_context_s* ctx, cudaStream_t s, const double* T, int ldt, int n, int uplo, int unit,
                        double* winv);
int trsm_left_rec_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* winv, double* B, int ldb);
int launch_transpose_grid(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A);
int transposed_copy(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& A, int uplo, DevMatrix* W);  // potrf_dist.cu
void swap_grid_comms(parsec_context_s* ctx);
int launch_transpose_local(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& S, const DevMatrix& D, int uplo);
int launch_transpose_upper_into(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst,
                                int ldd, int n);
int launch_transpose(parsec_context_s* ctx, cudaStream_t s, const double* src, int lds, double* dst, int ldd,
                     int m, int n);
// zgemm_dmma.cu
int zgemm_init_attributes();
int zgemm_launch(const GemmDesc& d, int nItems, cudaStream_t stream);
// tile_ops.cu
int gemm_single(parsec_context_s* ctx, cudaStream_t s, int M, int N, int K, double alpha, const double* A,
                int lda, bool aK, const double* B, int ldb, bool bK, double beta, double* C, int ldc,
                int lower, int batch = 1, long long sA = 0, long long sB = 0, long long sC = 0,
                int lastBatchM = -1);
int potrf_tile_lower(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv,
                     int* dInfo, int infoBase);
// fused variant: winv must be followed by a second region of the same size (see ensure_winv)
int potrf_tile_fused(parsec_context_s* ctx, cudaStream_t s, double* A, int lda, int n, double* winv,
                     int* dInfo, int infoBase);
int trsm_panel_lower(parsec_context_s* ctx, cudaStream_t s, const double* L, int ldl, int n, const double* winv,
                     double* B, int ldb, int M, int batch, long long strideB, int lastBatchM);
int trsm_left(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, int m, int n, double alpha,
              const double* T, int ldt, double* B, int ldb, double* winv);
int ensure_winv(parsec_context_s* ctx, int nb);
int solve_ibs_for(int n);
int ensure_solve_ws(parsec_context_s* ctx, int nb, int ncols);
double* solve_ws_big(parsec_context_s* ctx, int set);
double* solve_ws_xt(parsec_context_s* ctx);
int build_inverse_blocks(parsec_context_s* ctx, cudaStream_t s, int uplo, int diag, int n, const double* T, int ldt,
                         double* winv64, double* wbig, int ldw, double* y, int ibs);
int trsm_left_big_entry(parsec_context_s* ctx, cudaStream_t s, int uplo, bool trans, int n, int ncols, double alpha,
                        const double* T, int ldt, const double* wbig, int ldw, int ibs, double* xt, double* B, int ldb);
int ensure_wnorm(parsec_context_s* ctx, size_t n);
// staging.cu
int stage_in(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* out, bool copyData, int uploOnly);
int stage_out(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int uploOnly);
void stage_free(DevMatrix* dm);
int describe_matrix(const parsec_tiled_matrix_t* A, DevMatrix* dm);
// local tile rows [i0, i1) of local tile column jl <-> consecutive device tiles from devBase (dir 0: in, 1: out)
int stage_column_rows(const parsec_tiled_matrix_t* A, const DevMatrix* dm, int jl, int i0, int i1, double* devBase, int dir,
                      cudaStream_t s);
// potrf_ooc.cu
bool ooc_wanted(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, int uplo);
int potrf_out_of_core(parsec_context_s* ctx, parsec_taskpool_s* tp);
int stage_in_pipelined(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* out, int uploOnly);
int stage_out_column(parsec_context_s* ctx, const parsec_tiled_matrix_t* A, DevMatrix* dm, int jl, int uploOnly,
                     cudaEvent_t ready);
// comm.cu
int sum_allreduce(parsec_context_s* ctx, double* dbuf, size_t n, cudaStream_t s);
int bcast(parsec_context_s* ctx, void* comm, const double* send, double* recv, size_t count, int root, cudaStream_t s);
int p2p_send(parsec_context_s* ctx, const double* buf, size_t count, int peer, cudaStream_t s);
int p2p_recv(parsec_context_s* ctx, double* buf, size_t count, int peer, cudaStream_t s);
int group_start(parsec_context_s* ctx);
int group_end(parsec_context_s* ctx);
int ensure_grid_comms(parsec_context_s* ctx, int P, int Q, int ip = 0, int jq = 0);
// world rank of the process at LOGICAL grid coordinates (lr, lc) of A's distribution
inline int world_rank_of(const DevMatrix& A, int lr, int lc) { return ((lr + A.ip) % A.P) * A.Q + (lc + A.jq) % A.Q; }
void comm_destroy(parsec_context_s* ctx);
// potrf_dist.cu
int dist_transpose_tri(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& T, int uplo, const DevMatrix& Tt);
int alloc_like(const DevMatrix& A, DevMatrix* out);
int trsm_left_multi(parsec_context_s* ctx, cudaStream_t s, int uplo, int trans, int diag, double alpha,
                    const DevMatrix& A, const DevMatrix& B);
int symm_residual_multi(parsec_context_s* ctx, cudaStream_t s, int uplo, const DevMatrix& A, const DevMatrix& X,
                        const DevMatrix& B);
int check_update_multi(parsec_context_s* ctx, const DevMatrix& L, const DevMatrix& R);
// grid forms of the level-3 operations (potrf_dist.cu): the factorization's panel exchange with gmFirst = 0
int syrk_lower_multi(parsec_context_s* ctx, cudaStream_t s, double alpha, const DevMatrix& A, const DevMatrix& C);
int gemm_nt_multi(parsec_context_s* ctx, cudaStream_t s, double alpha, const DevMatrix& A, const DevMatrix& B, const DevMatrix& C);
// potri.cu
enum { OP_TRTRI = 1, OP_LAUUM = 2, OP_POTRF = 4 };
int run_inverse_chain(parsec_context_s* ctx, const char* name, int ops, int uplo, int diag, parsec_tiled_matrix_t* A,
                      int* info_out);
// taskpools.cu
int run_queue_in_order(parsec_context_s* ctx);
int launch_transpose_tiles(parsec_context_s* ctx, cudaStream_t s, const double* src, long long srcStride, int lds,
                           double* dst, long long dstStride, int ldd, int rows, int cols, int cnt);
// check.cu
int symm_residual_1gpu(parsec_context_s* ctx, cudaStream_t s, int uplo, const DevMatrix& A, const DevMatrix& X,
                       const DevMatrix& B, double* diagCopy);
int inf_norm_public(parsec_context_s* ctx, const DevMatrix& A, bool sym, int uplo, double* out);
// potrf.cu
int potrf_enqueue(parsec_context_s* ctx, parsec_taskpool_s* tp);
// DPB_TRACE=1: event timestamps at the end of every phase of every step, printed at parsec_context_wait
// (phases: 0 potrf tile, 1 panel solve, 2 look-ahead update, 3 bulk update, 4 diagonal-tile broadcast,
//  5 panel broadcasts, 6 catch-up of late columns)
bool trace_on();
void trace_begin(cudaStream_t s);
void trace_mark(cudaStream_t s, int k, int phase);
int trailing_update(parsec_context_s* ctx, cudaStream_t s, const DevMatrix& C, const double* rowPanel,
                    long long rowStride, const double* colPanel, long long colStride, int gnFirst,
                    int gnEnd, double alpha, int K, int kTile = 0, long long kTileStride = 0, int rowFirstFixed = -1,
                    int jlMin = 0, int jlMax = 1 << 30);
}  // namespace dpb
