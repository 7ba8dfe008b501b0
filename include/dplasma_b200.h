/*
 * dplasma_b200.h -- C ABI of libdplasma_b200.so: a B200-native (sm_100a) implementation of
 * DPLASMA's tiled Cholesky path (dpotrf / dpotrs / dposv / dtrsm-left / generators / checks /
 * pdpotrf_) behind the reference's own entry points.
 *
 * Every declaration cites the reference interface it replaces (paths relative to the
 * therault/dplasma tree).  Plain C: pointers and sizes only, no torch / C++ types.
 *
 * PaRSEC (the reference's runtime) is an external, un-vendored dependency of the reference;
 * the handful of PaRSEC names this path touches (parsec_context_t, parsec_taskpool_t,
 * parsec_tiled_matrix_t, parsec_matrix_block_cyclic_t, the *_init functions,
 * parsec_data_allocate/free, parsec_context_add_taskpool/start/wait) are provided here with
 * the field names the reference reads (SURVEY.md 8(b) "Descriptor surface to mirror"), so
 * that code written like tests/testing_zpotrf.c compiles and runs against this library.
 */
#ifndef DPLASMA_B200_H
#define DPLASMA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- enums: src/include/dplasma/constants.h:47,78-96 ---- */
typedef int dplasma_enum_t;
#define dplasmaNoTrans 111
#define dplasmaTrans 112
#define dplasmaConjTrans 113
#define dplasmaUpper 121
#define dplasmaLower 122
#define dplasmaUpperLower 123
#define dplasmaNonUnit 131
#define dplasmaUnit 132
#define dplasmaLeft 141
#define dplasmaRight 142
#define dplasmaInfNorm 175

/* ---- PaRSEC-shaped types (names as used by tests/common.h:182-190, src/dplasma_zcheck.c:80-83,
 *      src/scalapack_wrappers/common.c:147-152) ---- */
typedef struct parsec_context_s parsec_context_t;   /* opaque: one per process == one GPU */
typedef struct parsec_taskpool_s parsec_taskpool_t; /* opaque: a built, not yet run, operation */

enum parsec_matrix_type_e { PARSEC_MATRIX_DOUBLE = 4, PARSEC_MATRIX_COMPLEX_DOUBLE = 6 };
enum parsec_matrix_storage_e { PARSEC_MATRIX_LAPACK = 0, PARSEC_MATRIX_TILE = 1 };

typedef struct parsec_data_collection_s {
    uint32_t nodes;  /* number of ranks (processes, one GPU each) */
    uint32_t myrank;
} parsec_data_collection_t;

typedef struct parsec_tiled_matrix_s {
    parsec_data_collection_t super;
    int mtype;   /* parsec_matrix_type_e */
    int storage; /* parsec_matrix_storage_e */
    int mb, nb, bsiz; /* tile rows, cols, elements (mb*nb) */
    int lm, ln;       /* rows / cols of the whole matrix */
    int lmt, lnt;     /* tile rows / cols of the whole matrix */
    int llm, lln;     /* local rows / cols (LAPACK storage: llm is the leading dimension) */
    int i, j;         /* first row / col of the submatrix */
    int m, n;         /* submatrix rows / cols */
    int mt, nt;       /* submatrix tile rows / cols */
    int slm, sln;     /* local rows / cols of the submatrix */
    int nb_local_tiles;
} parsec_tiled_matrix_t;

typedef struct parsec_grid_2Dcyclic_s {
    int rank, rows, cols; /* my rank; P; Q  (row-major grid: rank = prow*Q + pcol) */
    int rrank, crank;     /* my grid row, my grid column */
    int krows, kcols;     /* k-cyclic factors (only 1 supported on this path: testing_zpotrf.c:37-38) */
    int ip, jq;           /* grid coordinates owning tile (0,0) */
} parsec_grid_2Dcyclic_t;

/* One struct serves parsec_matrix_block_cyclic_t and parsec_matrix_sym_block_cyclic_t.
 * `mat` may be HOST memory (staged through the GPU by the blocking calls) or DEVICE memory
 * (used in place).  Layouts:
 *   TILE storage: local tile (il,jl) at mat + (il + jl*nb_elem_r)*bsiz elements, each tile
 *       column-major with leading dimension mb (src/scalapack_wrappers/common.c:151-152).
 *   sym TILE storage: only tiles of the uplo triangle are stored, packed in the same order
 *       (local tile columns left to right, rows top to bottom inside a column).
 *   LAPACK storage: one column-major local array, leading dimension llm
 *       (src/scalapack_wrappers/dplasma_wrapper_pdpotrf.c:177-188). */
typedef struct parsec_matrix_block_cyclic_s {
    parsec_tiled_matrix_t super;
    parsec_grid_2Dcyclic_t grid;
    void *mat;
    int nb_elem_r, nb_elem_c; /* local tile rows / cols */
    int uplo;                 /* sym descriptors: dplasmaLower / dplasmaUpper; else UpperLower */
} parsec_matrix_block_cyclic_t;
typedef parsec_matrix_block_cyclic_t parsec_matrix_sym_block_cyclic_t;

/* ---- runtime shims ---- */
/* parsec_init (tests/common.c:680): creates the per-process context on the current CUDA
 * device; single rank until dplasma_b200_comm_init is called. */
parsec_context_t *parsec_init(int nb_cores, int *pargc, char **pargv[]);
/* parsec_fini (tests/common.c:737) */
int parsec_fini(parsec_context_t **pctx);
/* Multi-GPU: one process per GPU.  Rank 0 obtains an id, the application broadcasts the
 * 128 bytes to every rank by its own means (MPI_Bcast in the reference's world,
 * tests/common.c:648-659; torch.distributed here), then every rank calls comm_init.
 * Replaces parsec_remote_dep_set_ctx (scalapack_wrappers/dplasma_wrapper_parsec_init.c:112-114). */
int dplasma_b200_get_unique_id(void *id128);
int dplasma_b200_comm_init(parsec_context_t *ctx, const void *id128, int rank, int nranks);
/* parsec_context_add_taskpool / start / wait (tests/common.h:253-263): start enqueues all GPU
 * work of the added taskpools without blocking; wait synchronises and publishes info. */
int parsec_context_add_taskpool(parsec_context_t *ctx, parsec_taskpool_t *tp);
int parsec_context_start(parsec_context_t *ctx);
int parsec_context_wait(parsec_context_t *ctx);
/* parsec_data_allocate / parsec_data_free (tests/common.h:186-188): pinned host memory */
void *parsec_data_allocate(size_t bytes);
void parsec_data_free(void *p);
/* Device-resident variant for callers that keep the matrix in HBM (no reference counterpart:
 * the reference's matrices always live in host memory and are cached on the GPU by PaRSEC). */
void *dplasma_b200_device_allocate(size_t bytes);
void dplasma_b200_device_free(void *p);

/* descriptor initialisers: tests/testing_zposv.c:39-42, tests/testing_zpotrf.c:42-45,
 * src/scalapack_wrappers/dplasma_wrapper_pdpotrf.c:177-188 */
void parsec_matrix_block_cyclic_init(parsec_matrix_block_cyclic_t *dc, int mtype, int storage, int myrank,
                                     int mb, int nb, int lm, int ln, int i, int j, int m, int n,
                                     int P, int Q, int kp, int kq, int ip, int jq);
void parsec_matrix_sym_block_cyclic_init(parsec_matrix_sym_block_cyclic_t *dc, int mtype, int myrank,
                                         int mb, int nb, int lm, int ln, int i, int j, int m, int n,
                                         int P, int Q, int uplo);
void parsec_matrix_block_cyclic_lapack_init(parsec_matrix_block_cyclic_t *dc, int mtype, int storage,
                                            int myrank, int mb, int nb, int lm, int ln, int i, int j,
                                            int m, int n, int P, int Q, int kp, int kq, int ip,
                                            int jq, int mloc, int nloc);
void parsec_tiled_matrix_destroy(parsec_tiled_matrix_t *dc);
/* The per-step communication plan of the calling rank for tile column k (rows >= gmFirst):
 * which local tile rows / columns take part in the row and column panel broadcasts and, per
 * local column, the broadcast root (a grid row) and the root's local row index.  This is the
 * explicit form of the ranges in the JDF flows src/zpotrf_L.jdf:112,213-215 that PaRSEC turns
 * into messages.  Host arithmetic only.  out[6] = {kq, kp, il0, nrows, jl0, ncols}. */
int dplasma_b200_panel_plan(const parsec_tiled_matrix_t *A, int k, int gmFirst, int *out, int *colRoot,
                            int *colSrcIl);
/* The plan of step k of the distributed inversion DAGs (src/ztrtri_L.jdf:60-142, src/zlauum_L.jdf:38-130), whose
 * operands are tile ROW k left of the diagonal: out[4] = {kp, kq, local tile columns with n < k, local tile rows
 * with m < k}; per such local tile row (global m): the grid column m % Q that holds A(k,m) once the row has gone
 * down the grid columns, and m / Q, its local column index there (the root and source of the broadcast along the
 * grid row that the products A(m,n) += A(k,m)^T A(k,n) need).  Host arithmetic only. */
int dplasma_b200_rowpanel_plan(const parsec_tiled_matrix_t *A, int k, int *out, int *rowRootCol, int *rowSrcJl);
/* One round of the distributed transposition behind the transposed solves on a grid (the explicit form of the
 * A(n,k)^T dependencies of src/ztrsm_LLT.jdf:132-137): for logical rank `me` of a P x Q grid and the round that starts
 * at tile column nA of an mt x mt tile matrix (triangle `uplo`), *nB = first column after the round, and the (m, n, peer)
 * triples of the tiles it packs for each destination / unpacks from each source, in order (send list: peer == me for a
 * tile that stays on its rank).  cap = tiles a rank may send or receive per round (dplasma_b200_transpose_cap: 1 GB each
 * way).  Host arithmetic only; returns 0, -1 when a list is too short (counts are still set), -2 on bad arguments. */
long long dplasma_b200_transpose_cap(int mt, long long tileBytes);
int dplasma_b200_transpose_plan(int P, int Q, int mt, int uplo, long long cap, int me, int nA, int *nB,
                                int *sendTiles, int maxSend, int *nSend, int *recvTiles, int maxRecv, int *nRecv);
/* rank owning global tile (m,n): the data collection's rank_of (src/zpotrf_L.jdf:105) */
int dplasma_b200_rank_of(const parsec_tiled_matrix_t *A, int m, int n);

/* ---- public operations: src/include/dplasma/dplasma_z.h (d precision) ---- */
/* :172, src/zpotrf_wrapper.c:175-226.  Builds, does not run.  NULL + message on bad uplo. */
parsec_taskpool_t *dplasma_dpotrf_New(dplasma_enum_t uplo, parsec_tiled_matrix_t *A, int *info);
/* :239, src/zpotrf_wrapper.c:248-262 (BASELINE.json spells it _Destroy; both are exported) */
void dplasma_dpotrf_Destruct(parsec_taskpool_t *tp);
void dplasma_dpotrf_Destroy(parsec_taskpool_t *tp);
/* :82, src/zpotrf_wrapper.c:310-336.  0, >0 first bad leading minor (global, MAX over ranks) */
int dplasma_dpotrf(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
/* :83,:273, src/zpotrf_wrapper.c:390-416,52-59: hmb sets the in-tile recursion block */
int dplasma_dpotrf_rec(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A, int hmb);
void dplasma_dpotrf_setrecursive(parsec_taskpool_t *tp, int hmb);
/* :81, src/zposv_wrapper.c:64-114 */
int dplasma_dposv(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                  parsec_tiled_matrix_t *B);
/* :84, src/zpotrs_wrapper.c:58-98 */
int dplasma_dpotrs(parsec_context_t *parsec, dplasma_enum_t uplo, const parsec_tiled_matrix_t *A,
                   parsec_tiled_matrix_t *B);
/* :51, src/ztrsm_wrapper.c:263-310: all eight side/uplo/trans variants (ztrsm_{L,R}{L,U}{N,T}.jdf) on one
 * process; the Left variants also on a process grid */
int dplasma_dtrsm(parsec_context_t *parsec, dplasma_enum_t side, dplasma_enum_t uplo,
                  dplasma_enum_t trans, dplasma_enum_t diag, double alpha,
                  const parsec_tiled_matrix_t *A, parsec_tiled_matrix_t *B);
/* :108, src/zplgsy_wrapper.c:23-45,97-150 -> src/cores/core_zplgsy.c:72-150 (bit-exact) */
int dplasma_dplgsy(parsec_context_t *parsec, double bump, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                   unsigned long long int seed);
/* :109, src/zplrnt_wrapper.c -> src/cores/core_zplrnt.c:68-91 (bit-exact; diagdom must be 0) */
int dplasma_dplrnt(parsec_context_t *parsec, int diagdom, parsec_tiled_matrix_t *A,
                   unsigned long long int seed);
/* :280, src/dplasma_zcheck.c:64-134.  returns 0 correct / 1 suspicious */
int check_dpotrf(parsec_context_t *parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                 parsec_tiled_matrix_t *A0);
/* :279, src/dplasma_zcheck.c:180-225 (b is overwritten with b - A x like the reference) */
int check_daxmb(parsec_context_t *parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                parsec_tiled_matrix_t *b, parsec_tiled_matrix_t *x);
/* ---- inversion after the factorization; real double, one process.  The reference runs all
 *      tile bodies of these DAGs on the CPU (no CUDA body in ztrtri_*.jdf / zlauum_*.jdf / zpoinv_*.jdf). ---- */
/* dplasma_z.h:86, src/ztrtri_wrapper.c:194-230 (DAGs src/ztrtri_L.jdf, ztrtri_U.jdf).  Returns 0, -1 (uplo),
 * -2 (diag), -6 (A not square), -101 (could not run), or i > 0: A(i,i) is exactly zero (INFO = k*mb + info,
 * src/ztrtri_L.jdf:180-186) */
int dplasma_dtrtri(parsec_context_t *parsec, dplasma_enum_t uplo, dplasma_enum_t diag, parsec_tiled_matrix_t *A);
/* dplasma_z.h:78, src/zlauum_wrapper.c:184-214 (DAGs src/zlauum_L.jdf, zlauum_U.jdf): Lower A <- L^T L, Upper A <- U U^T */
int dplasma_dlauum(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
/* dplasma_z.h:85, src/zpotri_wrapper.c:55-86: inverse of an SPD matrix from its Cholesky factor */
int dplasma_dpotri(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
/* dplasma_z.h:79-80, src/zpoinv_wrapper.c:184-207,255-272: factor + invert in one call */
int dplasma_dpoinv(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
int dplasma_dpoinv_sync(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
/* dplasma_z.h:281, src/dplasma_zcheck.c:265-315: ||Id - Ainv A||_1 / (||A||_1 ||Ainv||_1 N eps) > 10 -> 1.
 * A is read as a general (full) matrix like the reference's dplasma_zhemm operand, Ainv from its uplo triangle. */
int check_dpoinv(parsec_context_t *parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                 parsec_tiled_matrix_t *Ainv);
/* ---- level-3 operations on the same engine; real double, one process ---- */
/* dplasma_z.h:28-30, src/zgemm_wrapper.c:645-733: C = alpha op(A) op(B) + beta C.  Returns 0, -1 (transA), -2 (transB),
 * -101 (tile sizes / matrix sizes do not match, or the call could not run) */
int dplasma_dgemm(parsec_context_t *parsec, dplasma_enum_t transA, dplasma_enum_t transB, double alpha,
                  const parsec_tiled_matrix_t *A, const parsec_tiled_matrix_t *B, double beta,
                  parsec_tiled_matrix_t *C);
/* dplasma_z.h:43, src/zsyrk_wrapper.c:219-258: C = alpha op(A) op(A)^T + beta C on the uplo triangle of C.
 * Returns 0, -1 (uplo), -2 (trans), -6 (C not square), -4 (A does not match C), -101 (could not run) */
int dplasma_dsyrk(parsec_context_t *parsec, dplasma_enum_t uplo, dplasma_enum_t trans, double alpha,
                  const parsec_tiled_matrix_t *A, double beta, parsec_tiled_matrix_t *C);
/* ---- taskpool forms of the operations above (dplasma_z.h:125-127,144-146,153-154,170-173,213,219,221,237-240): _New
 *      records the operation and returns NULL on an illegal argument; a context whose queue holds any of them
 *      runs the queue operation by operation in the order it was filled (what a composition such as
 *      src/zpotri_wrapper.c:68-80 needs); _Destruct frees the object ---- */
parsec_taskpool_t *dplasma_dtrsm_New(dplasma_enum_t side, dplasma_enum_t uplo, dplasma_enum_t trans,
                                     dplasma_enum_t diag, double alpha, const parsec_tiled_matrix_t *A,
                                     parsec_tiled_matrix_t *B);
parsec_taskpool_t *dplasma_dtrtri_New(dplasma_enum_t uplo, dplasma_enum_t diag, parsec_tiled_matrix_t *A, int *info);
parsec_taskpool_t *dplasma_dlauum_New(dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
parsec_taskpool_t *dplasma_dpoinv_New(dplasma_enum_t uplo, parsec_tiled_matrix_t *A, int *info);
parsec_taskpool_t *dplasma_dgemm_New(dplasma_enum_t transA, dplasma_enum_t transB, double alpha,
                                     const parsec_tiled_matrix_t *A, const parsec_tiled_matrix_t *B, double beta,
                                     parsec_tiled_matrix_t *C);
parsec_taskpool_t *dplasma_dsyrk_New(dplasma_enum_t uplo, dplasma_enum_t trans, double alpha,
                                     const parsec_tiled_matrix_t *A, double beta, parsec_tiled_matrix_t *C);
void dplasma_dtrsm_Destruct(parsec_taskpool_t *tp);
void dplasma_dtrtri_Destruct(parsec_taskpool_t *tp);
void dplasma_dlauum_Destruct(parsec_taskpool_t *tp);
void dplasma_dpoinv_Destruct(parsec_taskpool_t *tp);
void dplasma_dgemm_Destruct(parsec_taskpool_t *tp);
void dplasma_dsyrk_Destruct(parsec_taskpool_t *tp);
/* ---- complex double (zpotrf, BASELINE configs[4]); A->mtype must be PARSEC_MATRIX_COMPLEX_DOUBLE,
 *      tiles hold interleaved (re,im) dplasma_complex64_t.  Both triangles on one GPU, dplasmaLower on a grid.
 *      src/include/dplasma/dplasma_z.h:82,107,109,172,239,280 ---- */
parsec_taskpool_t *dplasma_zpotrf_New(dplasma_enum_t uplo, parsec_tiled_matrix_t *A, int *info);
void dplasma_zpotrf_Destruct(parsec_taskpool_t *tp);
void dplasma_zpotrf_Destroy(parsec_taskpool_t *tp);
int dplasma_zpotrf(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A);
/* src/zplghe_wrapper.c -> src/cores/core_zplghe.c:72-154 (bit-exact) */
int dplasma_zplghe(parsec_context_t *parsec, double bump, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                   unsigned long long int seed);
int dplasma_zplrnt(parsec_context_t *parsec, int diagdom, parsec_tiled_matrix_t *A, unsigned long long int seed);
int check_zpotrf(parsec_context_t *parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                 parsec_tiled_matrix_t *A0);
double dplasma_zlanhe(parsec_context_t *parsec, dplasma_enum_t ntype, dplasma_enum_t uplo,
                      const parsec_tiled_matrix_t *A);
double dplasma_zlange(parsec_context_t *parsec, dplasma_enum_t ntype, const parsec_tiled_matrix_t *A);
/* dplasma_z.h:81,84,279: complex solves and their check; dplasmaLower on one process in this build */
int dplasma_zposv(parsec_context_t *parsec, dplasma_enum_t uplo, parsec_tiled_matrix_t *A, parsec_tiled_matrix_t *B);
int dplasma_zpotrs(parsec_context_t *parsec, dplasma_enum_t uplo, const parsec_tiled_matrix_t *A,
                   parsec_tiled_matrix_t *B);
int check_zaxmb(parsec_context_t *parsec, int loud, dplasma_enum_t uplo, parsec_tiled_matrix_t *A,
                parsec_tiled_matrix_t *b, parsec_tiled_matrix_t *x);
/* the numbers the two checks print (src/dplasma_zcheck.c:105,203), for callers that want them */
double dplasma_b200_last_check_result(parsec_context_t *parsec);
/* dplasma_zlanhe / dplasma_zlange, InfNorm only (src/dplasma_zcheck.c:101-103,194-196) */
double dplasma_dlansy(parsec_context_t *parsec, dplasma_enum_t ntype, dplasma_enum_t uplo,
                      const parsec_tiled_matrix_t *A);
double dplasma_dlange(parsec_context_t *parsec, dplasma_enum_t ntype, const parsec_tiled_matrix_t *A);

/* seconds the GPU spent between parsec_context_start and the end of the last taskpool
 * (CUDA events on the library's streams); the reference times the same bracket on the host
 * (tests/common.h:258-263). */
double dplasma_b200_last_elapsed(parsec_context_t *parsec);
/* Event-timed profile of the grouped trailing-update launches (the kernel that replaces the
 * per-tile cublasDgemm/cublasDsyrk calls of src/zpotrf_L.jdf:342-346,463-468): on=1 starts
 * collecting; read returns summed seconds, algorithmic flops, launch count and the best
 * single-launch TFLOP/s.  No reference counterpart (PaRSEC's --enable-prof-trace plays this role). */
void dplasma_b200_profile(parsec_context_t *parsec, int on);
int dplasma_b200_profile_read(parsec_context_t *parsec, double *seconds, double *flops, int *launches,
                              double *best_tflops);
/* the same sums restricted to one stream: which = 1 the bulk updates (update stream; launches on one stream do
 * not overlap each other), 0 the look-ahead launches on the panel stream, -1 all */
int dplasma_b200_profile_read_stream(parsec_context_t *parsec, int which, double *seconds, double *flops,
                                     int *launches, double *best_tflops);
/* number of kernels this library launched since the context was created */
long long dplasma_b200_kernel_launches(parsec_context_t *parsec);

/* ---- kernel seam: stream-ordered, non-blocking, column-major tile ops
 *      with the argument shapes of the library calls they replace.  `stream` is a
 *      cudaStream_t passed as void*.  All pointers are device pointers.  Return 0 or a
 *      cudaError_t / negative argument index. ---- */
/* cusolverDnDpotrf, src/zpotrf_L.jdf:143 (K1).  *d_info (device int) receives
 * info_offset + j + 1 for the first non-positive pivot j if it was 0 before. */
int dplasma_b200_dpotrf_tile(parsec_context_t *ctx, void *stream, dplasma_enum_t uplo, int n, double *A,
                             int lda, int *d_info, int info_offset);
/* No counterpart in the reference.  Same computation as dplasma_b200_dpotrf_tile(Lower) for
 * callers that are not allowed to allocate device memory (a task body under a runtime that
 * owns the whole HBM, a no-sync inner loop, a fixed memory budget): the workspace lives in a
 * 512 KiB array inside the library image, so n <= 1024, nothing is allocated or freed, and no
 * context is needed.  d_info is required.  The first call on a device performs a one-time
 * set-up and must happen outside stream capture; calls that overlap on the GPU race on the
 * shared workspace.  Returns 0, a cudaError_t, or a negative argument index (-3: n too large). */
int dplasma_b200_dpotrf_tile_noalloc(void *stream, dplasma_enum_t uplo, int n, double *A, int lda,
                                     int *d_info, int info_offset);
/* cublasDtrsm, src/zpotrf_L.jdf:236-240 (K2).  Right/Lower/Trans/NonUnit and
 * Left/Upper/Trans/NonUnit as used by the factorization, and Left/{Lower,Upper}/{N,T} as
 * used by the solves (CORE_ztrsm at src/ztrsm_LLN.jdf:48). */
int dplasma_b200_dtrsm_tile(parsec_context_t *ctx, void *stream, dplasma_enum_t side, dplasma_enum_t uplo,
                            dplasma_enum_t trans, dplasma_enum_t diag, int m, int n, double alpha,
                            const double *T, int ldt, double *C, int ldc);
/* cublasDsyrk, src/zpotrf_L.jdf:342-346 (K3) */
int dplasma_b200_dsyrk_tile(parsec_context_t *ctx, void *stream, dplasma_enum_t uplo, dplasma_enum_t trans,
                            int n, int k, double alpha, const double *A, int lda, double beta, double *C,
                            int ldc);
/* cublasDgemm, src/zpotrf_L.jdf:463-468 (K4) */
int dplasma_b200_dgemm_tile(parsec_context_t *ctx, void *stream, dplasma_enum_t transA,
                            dplasma_enum_t transB, int m, int n, int k, double alpha, const double *A,
                            int lda, const double *B, int ldb, double beta, double *C, int ldc);

/* ---- ScaLAPACK-compatible entry: src/scalapack_wrappers/dplasma_wrapper_pdpotrf.c:133-297 ---- */
void pdpotrf_(char *UPLO, int *N, double *A, int *IA, int *JA, int *DESCA, int *info);
/* src/scalapack_wrappers/dplasma_wrapper_pdtrsm.c:203-327: like the reference's wrapper the operands are
 * always redistributed into tiles (:303-307); square equal blocks, IA=JA=IB=JB=1, whole matrices.
 * ScaLAPACK's pdtrsm has no info argument: argument errors are printed and the call returns. */
/* src/scalapack_wrappers/dplasma_wrapper_pdgemm.c:215-221: same restrictions as pdtrsm_ */
void pdgemm_(char *TRANSA, char *TRANSB, int *M, int *N, int *K, double *ALPHA, double *A, int *IA, int *JA,
             int *DESCA, double *B, int *IB, int *JB, int *DESCB, double *BETA, double *C, int *IC, int *JC,
             int *DESCC);
void pdtrsm_(char *SIDE, char *UPLO, char *TRANS, char *DIAG, int *M, int *N, double *ALPHA, double *A, int *IA,
             int *JA, int *DESCA, double *B, int *IB, int *JB, int *DESCB);
/* dplasma_wrapper_parsec_init.c:98-124 */
void parsec_init_wrapper_(void);
void parsec_fini_wrapper_(void);
/* The reference obtains the process grid from BLACS (Cblacs_gridinfo, common.h:298-301);
 * there is no BLACS here, so the grid of a context number is registered explicitly. */
void dplasma_b200_blacs_gridinit(int ictxt, int nprow, int npcol, int myrow, int mycol);
/* The reference's wrapper initialises PaRSEC lazily on the BLACS communicator
 * (dplasma_wrapper_parsec_init.c:72-96); here an application that already owns a context with
 * its inter-GPU communicators lends it to the wrapper. */
void dplasma_b200_wrapper_set_context(parsec_context_t *ctx);

/* src/flops.h:86-87,92-93,282 */
double dplasma_b200_flops_dpotrf(double n);
double dplasma_b200_flops_dpotrs(double n, double nrhs);

/* FP64 tensor-pipe ceiling: register-resident DMMA loop, returns TFLOP/s (no reference
 * counterpart; the reference's only peak probe is the SGEMM tool tools/gemmpeak/sgemmN.cu). */
double dplasma_b200_dmma_peak_tflops(int ctas_per_sm, int iters, double *seconds_out);

const char *dplasma_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DPLASMA_B200_H */
