// FP64 tensor-core (DMMA) GEMM engine for sm_100a -- the dense contraction behind every
// Level-3 step of the tiled Cholesky path.
//
// Replaces, inside the reference's task bodies, the per-tile library calls
//   cublasDgemm  (src/zpotrf_L.jdf:463-468)   C(m,n) -= L(m,k) L(n,k)^T
//   cublasDsyrk  (src/zpotrf_L.jdf:342-346)   C(m,m) -= L(m,k) L(m,k)^T   (lower part only)
// and, through inverted diagonal blocks, cublasDtrsm (src/zpotrf_L.jdf:236-240).
//
// FP64 has no tcgen05 / TMEM kind on sm_100a: the FP64 tensor path is the warp-level
// mma.sync.m8n8k4.f64 (SASS DMMA.8x8x4) with register accumulators.  The kernel is therefore
// an Ampere-style register-tiled GEMM fed by a Blackwell-style pipeline: one producer warp
// stages operand panels into shared memory with the TMA engine -- tensor-map tile copies
// (cp.async.bulk.tensor.4d, SASS UTMALDG) for operands stored M-major, 1-D bulk copies
// (cp.async.bulk, SASS UBLKCP) for operands stored K-major -- completing on mbarriers; consumer
// warps wait on the full barrier, run DMMA from conflict-free shared-memory layouts with
// register-double-buffered fragments (the loads of the next k-step, and of the next stage, are
// in flight while the current DMMAs issue), and release stages through empty barriers.
//
// Operand form is "NT": C[i][j] (+)= sum_k A[i][k] * B[j][k], A and B column-major with the
// contraction index as the slow dimension (a column of A is contiguous) -- exactly the
// layout of a column-major tile L(m,k).  "Kmajor" operands (rows of op(A) contiguous along k,
// i.e. A^T stored column-major) are supported by a transposing producer for the upper /
// solve variants.
#pragma once
#include <cuda.h>  // CUtensorMap (types only: the encoder is resolved through cudaGetDriverEntryPoint)
#include <cuda_runtime.h>
#include <stdint.h>

namespace dpb {

// ---------------------------------------------------------------------------------------
// Work description.  One launch covers either a single GEMM/SYRK, or the whole trailing
// update of one factorization step over tile storage ("grouped" form of SURVEY 8(b) B2).
// ---------------------------------------------------------------------------------------
enum GemmMode : int { GEMM_SINGLE = 0, GEMM_TRAILING = 1 };
static constexpr int kMaxCols = 160;  // max local tile columns in one trailing update

struct GemmDesc {
    int mode;
    // common
    double alpha, beta;  // C = beta*C + alpha*A*B^T
    int K;               // contraction length
    int fast;            // 1: TMA producer (operands 16 B aligned, even leading dimensions / strides; K-major
                         //    operands additionally K % 16 == 0); 0: element-wise producer
    int preload;         // 1: accumulators start as (beta/alpha)*C (hides the C read; sums into |C|-sized
                         //    accumulators, so only for tiles whose entries are of the size of the products);
                         // 0: BLAS order: product accumulated on its own, C = beta*C + alpha*acc in the epilogue
    int a_kmajor, b_kmajor;  // 0: operand stored M-major (NT form); 1: stored K-major
    // K may span several tiles of a tile-row: element (i,kk) of A is at
    //   A + (kk / kTile)*aKTileStride + (kk % kTile)*lda + i      (M-major)
    int kTile;
    long long aKTileStride, bKTileStride;
    // ---- GEMM_SINGLE ----
    const double* A;
    const double* B;
    double* C;
    int M, N, lda, ldb, ldc;
    int lower;   // 1: SYRK-like, touch only elements with (rowOff+i) >= (colOff+j); 2: only <=
    int rowOff, colOff;
    int gridM, gridN;
    int batch;                // independent problems, strided; the last one may have fewer rows
    long long batchStrideA, batchStrideB, batchStrideC;
    int lastBatchM;
    // ---- GEMM_TRAILING ----  (tile storage: tile (i,j) of the local array at
    //      Cbase + (i + j*tileLdT)*tileElems, each tile column-major with ld = nb)
    double* Cbase;
    const double* rowPanel;   // tile for local tile-row i at rowPanel + i*rowPanelStride
    const double* colPanel;   // tile for local tile-col j at colPanel + j*colPanelStride
    long long rowPanelStride, colPanelStride, tileElems;
    int tileLdT;              // local tile rows (stride between tile columns, in tiles)
    int nb;                   // tile size
    int bptM, bptN;           // CTA blocks per tile edge = ceil(nb / BM), ceil(nb / BN)
    int ncols;                // number of local tile columns in this update
    int colFirst;             // first local tile column j0
    int P, Q, prow, pcol;     // process grid and my coordinates: global m = i*P+prow, n = j*Q+pcol
    int lastM, lastN;         // rows in the last global tile-row / cols in last tile-col (ragged edge)
    int mtGlobal, ntGlobal;   // global tile counts
    int rowFirst[kMaxCols];   // per column: first local tile row with global m >= global n
    int rowEnd;               // local tile rows end (exclusive)
    int full;                 // 1: general product, diagonal tiles are updated in full (no SYRK mask)
    int itemStart[kMaxCols + 1];  // prefix sum of work items (CTA tiles) per column
    // ---- work order (both modes): which CTA tiles are resident together decides what the L2 has to hold ----
    int colGroup;  // TRAILING: tile columns walked together, tile row by tile row (0: one 64-wide strip at a time)
    int blkGroup;  // SINGLE: CTA block columns walked together, block row by block row (0: one block column at a time)
    // ---- tensor maps of the M-major operands (fast mode): 4-D views {rows, k, tile | batch, k-tile}
    alignas(64) CUtensorMap tmA;
    alignas(64) CUtensorMap tmB;
};

// ---------------------------------------------------------------------------------------
// PTX helpers
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE;\n"
        "bra WAIT_LOOP;\n"
        "DONE:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine, SASS UBLKCP), completes on an mbarrier
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// 4-D tensor-map tile copy global -> shared (SASS UTMALDG); out-of-bounds parts of the box arrive as zeros
__device__ __forceinline__ void tma_g2s_4d(void* dst, const CUtensorMap* tm, int c0, int c1, int c2, int c3,
                                           uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(
            smem_u32(dst)),
        "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
// 16-byte shared-memory load.  volatile: the consumer loop is software-pipelined by hand and the order of
// loads, DMMAs and barrier operations written in the source is the order that must reach the SASS.
__device__ __forceinline__ void lds128(double& x, double& y, uint32_t addr) {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(addr));
}
__device__ __forceinline__ void lds64(double& x, uint32_t addr) {
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(x) : "r"(addr));
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    // volatile: keeps every DMMA ordered against the (volatile) mbarrier waits/arrives, see the
    // stage-release note in the consumer loop
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1)
                 : "d"(a), "d"(b));
}

// ---------------------------------------------------------------------------------------
// Tile configuration
// ---------------------------------------------------------------------------------------
template <int BM_, int BN_, int STAGES_, int MINB_>
struct GemmCfg {
    static constexpr int BM = BM_, BN = BN_, BK = 16, WM = 32, WN = 32, STAGES = STAGES_, MINB = MINB_;
    static constexpr int MF = WM / 8, NF = WN / 8;  // 8x8 fragments per warp tile
    static constexpr int WARPS_M = BM / WM, WARPS_N = BN / WN;
    static constexpr int CONSUMER_WARPS = WARPS_M * WARPS_N;
    static constexpr int THREADS = (CONSUMER_WARPS + 1) * 32;
    // M-major operands are staged in blocks of 16 rows: [rows / 16][BK][16 doubles], one TMA box {16 rows, BK}
    // per block, written with the 128-byte swizzle (the 16-byte chunk j of k-row r lands at chunk j ^ (r & 7)).
    // A fragment load is one LDS.128 per lane -- two adjacent rows at one k -- and with the row-pair
    // permutation pair_of() below the 8 lanes of every quarter-warp hit 8 different 16-byte bank groups.
    static constexpr int MBLK = 16;
    static constexpr int MBLK_BYTES = MBLK * BK * 8;  // 2048: a multiple of the 1024-byte swizzle period
    // K-major operands are staged as [row][LDK]: one 128 B bulk copy per operand row, natural
    // fragment rows (8f + g), LDS.64 conflict-free for LDK = 4 (mod 16)
    static constexpr int LDK = BK + 4;
    __host__ __device__ static constexpr int a_stage(bool ak) { return ak ? BM * LDK : BM * BK; }  // doubles
    __host__ __device__ static constexpr int b_stage(bool bk) { return bk ? BN * LDK : BN * BK; }
    // + 1024: the kernel rounds the dynamic shared-memory base up to the swizzle period
    __host__ __device__ static constexpr size_t smem(bool ak, bool bk) {
        return (size_t)STAGES * (a_stage(ak) + b_stage(bk)) * 8 + 2 * STAGES * 8 + 16 + 1024;
    }
    static constexpr size_t SMEM = smem(false, false);
    static_assert((BM * BK * 8) % 1024 == 0 && (BM * LDK * 8) % 1024 == 0 && (BN * BK * 8) % 1024 == 0 &&
                      (BN * LDK * 8) % 1024 == 0, "stage bases must keep the 1024-byte swizzle alignment");
};
// 4 consumer warps, 32x32 warp tiles, 66 KB smem, 3 CTAs/SM: co-resident CTAs hide each other's prologue and
// epilogue (measured in round 1 against 128x128 and 128x64 CTA tiles at 1-2 CTAs/SM: 34.9 vs 32.6 / 32.3 TFLOP/s)
using CfgSmall = GemmCfg<64, 64, 4, 3>;
enum GemmCfgId : int { CFG_SMALL = 2 };

// Row pair (rows 2p, 2p+1 of a 16-row block) that fragment-row g of an m8n8k4 operand stands for.  Any
// permutation is legal as long as A, B and C agree; this one sends the two fragment rows that share a
// quarter-warp (g = 2q, 2q+1) to pairs 4 apart, which the swizzle then keeps on different bank groups
// for all four k-rows of a k-step.
__host__ __device__ __forceinline__ constexpr int pair_of(int g) { return (g >> 1) | ((g & 1) << 2); }
// row (within the warp tile) of M-major fragment f for fragment-row g: 2 adjacent rows, 16 apart per pair
__device__ __forceinline__ int mrow(int g, int f) { return 2 * pair_of(g) + (f & 1) + 16 * (f >> 1); }
// fragment maps of the kernels that keep their own padded staging (kernels.cu)
template <int WM>
__device__ __forceinline__ int frag_row(int g, int f) {
    if (WM == 64) return 8 * g + f;               // thread owns 8 contiguous rows
    return 2 * g + (f & 1) + 16 * (f >> 1);       // thread owns 2+2 rows, 16 apart
}
__device__ __forceinline__ int frag_col(int g, int f) { return 2 * g + (f & 1) + 16 * (f >> 1); }

struct BlockWork {
    const double* A;
    const double* B;
    double* C;
    int M, N;        // valid rows/cols in this CTA tile (<= BM/BN)
    int lda, ldb, ldc;
    int lower;       // mask: write only gi >= gj
    int gi0, gj0;    // global (within-problem) offsets of this CTA tile for the mask
    int aRow, aTile, bRow, bTile;  // tensor-map coordinates of the operand blocks: first row, tile | batch index
    bool valid;
};

template <class Cfg>
__host__ __device__ __forceinline__ BlockWork decode_work(const GemmDesc& d, int bid) {
    BlockWork w;
    w.valid = true;
    if (d.mode == GEMM_SINGLE) {
        // CTA tiles enumerated column-block major so that consecutive CTAs share the B panel
        const int per = d.gridM * d.gridN;
        const int bt = bid / per, r = bid - bt * per;
        int bj = r / d.gridM, bi = r - bj * d.gridM;
        if (d.blkGroup > 0) {
            // groups of blkGroup block columns, block row by block row inside a group: the CTAs resident together
            // (3 x 148) cover a ~28 x 16 patch of C and share 28 + 16 operand blocks instead of 444 + 2
            const int per_g = d.gridM * d.blkGroup;
            const int cg = r / per_g, rem = r - cg * per_g;
            const int gc = min(d.blkGroup, d.gridN - cg * d.blkGroup);
            bi = rem / gc;
            bj = cg * d.blkGroup + (rem - bi * gc);
        }
        int i0 = bi * Cfg::BM, j0 = bj * Cfg::BN;
        const int Mtot = (bt == d.batch - 1) ? d.lastBatchM : d.M;
        w.M = min(Cfg::BM, Mtot - i0);
        w.N = min(Cfg::BN, d.N - j0);
        if (w.M <= 0 || w.N <= 0) { w.valid = false; return w; }
        w.lda = d.lda; w.ldb = d.ldb; w.ldc = d.ldc;
        const double* Ab = d.A + (size_t)bt * d.batchStrideA;
        const double* Bb = d.B + (size_t)bt * d.batchStrideB;
        w.A = d.a_kmajor ? Ab + (size_t)i0 * d.lda : Ab + i0;
        w.B = d.b_kmajor ? Bb + (size_t)j0 * d.ldb : Bb + j0;
        w.C = d.C + (size_t)bt * d.batchStrideC + i0 + (size_t)j0 * d.ldc;
        w.lower = d.lower;
        w.gi0 = d.rowOff + i0; w.gj0 = d.colOff + j0;
        w.aRow = i0; w.aTile = (d.batchStrideA != 0) ? bt : 0;  // stride 0: one operand shared by the whole batch
        w.bRow = j0; w.bTile = (d.batchStrideB != 0) ? bt : 0;
        if (d.lower == 1 && (w.gi0 + w.M - 1 < w.gj0)) w.valid = false;  // entirely above diagonal
        if (d.lower == 2 && (w.gj0 + w.N - 1 < w.gi0)) w.valid = false;  // entirely below diagonal
        return w;
    }
    // GEMM_TRAILING: find the local tile column by binary search over the prefix array
    int lo = 0, hi = d.ncols;
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (d.itemStart[mid] <= bid) lo = mid; else hi = mid;
    }
    int c = lo;
    int i, j, bi, bj;
    if (d.colGroup > 0) {
        // Tile columns in groups of colGroup; inside a group tile row by tile row, the tiles of a row one after the
        // other, the CTA blocks of a tile together.  The CTAs resident at one time work on a few neighbouring tiles:
        // a row-panel tile is fetched once per group (not once per 64-wide strip) and the group's column-panel tiles
        // stay in the L2 while the rows stream by.  rowFirst is non-decreasing inside a group (the host checks):
        // rows [rowFirst[c0+l], rowFirst[c0+l+1]) hold the l+1 columns c0 .. c0+l, the last level runs to rowEnd.
        const int c0 = (c / d.colGroup) * d.colGroup, gc = min(d.colGroup, d.ncols - c0);
        const int bpt = d.bptM * d.bptN;
        const int r = bid - d.itemStart[c0];
        int q = r / bpt;
        const int e = r - q * bpt;
        i = d.rowEnd; c = c0;
        for (int lev = 0; lev < gc; lev++) {
            const int r0 = d.rowFirst[c0 + lev];
            const int r1 = (lev + 1 < gc) ? d.rowFirst[c0 + lev + 1] : d.rowEnd;
            const int cnt = (r1 - r0) * (lev + 1);
            if (q < cnt) {
                const int ri = q / (lev + 1);
                i = r0 + ri;
                c = c0 + (q - ri * (lev + 1));
                break;
            }
            q -= cnt;
        }
        j = d.colFirst + c;
        bj = e / d.bptM;
        bi = e - bj * d.bptM;
    } else {
        const int r = bid - d.itemStart[c];
        j = d.colFirst + c;                 // local tile column
        const int rowsInCol = (d.rowEnd - d.rowFirst[c]) * d.bptM;   // CTA block-rows in this tile column
        bj = r / rowsInCol;                 // CTA block column inside the tile
        const int rr = r - bj * rowsInCol;
        i = d.rowFirst[c] + rr / d.bptM;    // local tile row
        bi = rr - (rr / d.bptM) * d.bptM;
    }
    int gm = i * d.P + d.prow, gn = j * d.Q + d.pcol;   // global tile coordinates
    int tm = (gm == d.mtGlobal - 1) ? d.lastM : d.nb;
    int tn = (gn == d.ntGlobal - 1) ? d.lastN : d.nb;
    int i0 = bi * Cfg::BM, j0 = bj * Cfg::BN;
    w.M = min(Cfg::BM, tm - i0);
    w.N = min(Cfg::BN, tn - j0);
    if (w.M <= 0 || w.N <= 0) { w.valid = false; return w; }
    w.lda = w.ldb = w.ldc = d.nb;
    w.A = d.rowPanel + (size_t)i * d.rowPanelStride + i0;
    w.B = d.colPanel + (size_t)j * d.colPanelStride + j0;
    w.C = d.Cbase + ((size_t)i + (size_t)j * d.tileLdT) * d.tileElems + i0 + (size_t)j0 * d.nb;
    w.lower = (gm == gn) && !d.full;
    w.gi0 = i0; w.gj0 = j0;
    w.aRow = i0; w.aTile = i; w.bRow = j0; w.bTile = j;
    if (w.lower && (i0 + w.M - 1 < j0)) w.valid = false;
    return w;
}

// ---------------------------------------------------------------------------------------
// The kernel
// ---------------------------------------------------------------------------------------
template <class Cfg, bool AK, bool BKM>
__global__ void __launch_bounds__(Cfg::THREADS, Cfg::MINB) gemm_dmma_kernel(const __grid_constant__ GemmDesc d) {
    constexpr int BM = Cfg::BM, BN = Cfg::BN, BK = Cfg::BK, STAGES = Cfg::STAGES;
    constexpr int MF = Cfg::MF, NF = Cfg::NF, LDK = Cfg::LDK, MBB = Cfg::MBLK_BYTES;
    extern __shared__ unsigned char smem_dyn[];
    constexpr int A_STAGE = Cfg::a_stage(AK), B_STAGE = Cfg::b_stage(BKM);
    // the swizzle of the TMA engine is a function of the shared-memory address bits: 1024-byte aligned stages
    unsigned char* smem_raw = smem_dyn + ((1024u - (smem_u32(smem_dyn) & 1023u)) & 1023u);
    double* sA = reinterpret_cast<double*>(smem_raw);
    double* sB = sA + (size_t)STAGES * A_STAGE;
    uint64_t* full = reinterpret_cast<uint64_t*>(sB + (size_t)STAGES * B_STAGE);
    uint64_t* empty = full + STAGES;

    const BlockWork w = decode_work<Cfg>(d, blockIdx.x);
    if (!w.valid) return;  // uniform per CTA

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], Cfg::CONSUMER_WARPS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int nChunks = (d.K + BK - 1) / BK;

    if (warp == Cfg::CONSUMER_WARPS) {
        // ============================ producer warp ============================
        const bool fast = d.fast != 0;
        const int nBoxA = (w.M + Cfg::MBLK - 1) / Cfg::MBLK, nBoxB = (w.N + Cfg::MBLK - 1) / Cfg::MBLK;
        // what one stage brings in: whole boxes for the tensor-map copies (out-of-bounds parts count, they
        // arrive as zeros), 128-byte rows for the K-major bulk copies
        const uint32_t txBytes = (uint32_t)((AK ? w.M * BK * 8 : nBoxA * MBB) + (BKM ? w.N * BK * 8 : nBoxB * MBB));
        const int nA = AK ? w.M : nBoxA, nB = BKM ? w.N : nBoxB;  // copies per stage
        for (int c = 0; c < nChunks; c++) {
            const int s = c % STAGES;
            const uint32_t ph = (c / STAGES) & 1;
            if (c >= STAGES) mbar_wait(&empty[s], ph ^ 1);
            double* dA = sA + (size_t)s * A_STAGE;
            double* dB = sB + (size_t)s * B_STAGE;
            const int kk0 = c * BK;
            const int kt = kk0 / d.kTile, kin = kk0 - kt * d.kTile;
            if (fast) {
                if (lane == 0) mbar_expect_tx(&full[s], txBytes);
                __syncwarp();
                const double* Ab = w.A + (size_t)kt * d.aKTileStride;
                const double* Bb = w.B + (size_t)kt * d.bKTileStride;
                for (int q = lane; q < nA + nB; q += 32) {
                    if (q < nA) {
                        if (AK) bulk_g2s(dA + q * LDK, Ab + kin + (size_t)q * w.lda, BK * 8, &full[s]);
                        else tma_g2s_4d(reinterpret_cast<unsigned char*>(dA) + q * MBB, &d.tmA, w.aRow + q * Cfg::MBLK, kin,
                                        w.aTile, kt, &full[s]);
                    } else {
                        const int r = q - nA;
                        if (BKM) bulk_g2s(dB + r * LDK, Bb + kin + (size_t)r * w.ldb, BK * 8, &full[s]);
                        else tma_g2s_4d(reinterpret_cast<unsigned char*>(dB) + r * MBB, &d.tmB, w.bRow + r * Cfg::MBLK, kin,
                                        w.bTile, kt, &full[s]);
                    }
                }
            } else {
                // generic producer: element loads with zero fill (unaligned operands), same shared-memory layouts
                for (int r = 0; r < BK; r++) {
                    const int kk = kk0 + r;
                    const bool kin_range = kk < d.K;
                    const int kt2 = kk / d.kTile, kin2 = kk - kt2 * d.kTile;
                    for (int i = lane; i < BM; i += 32) {
                        double v = 0.0;
                        if (kin_range && i < w.M)
                            v = AK ? w.A[(size_t)kt2 * d.aKTileStride + kin2 + (size_t)i * w.lda]
                                   : w.A[(size_t)kt2 * d.aKTileStride + (size_t)kin2 * w.lda + i];
                        dA[AK ? i * LDK + r : ((i >> 4) * BK + r) * 16 + ((((i & 15) >> 1) ^ (r & 7)) << 1) + (i & 1)] = v;
                    }
                    for (int j = lane; j < BN; j += 32) {
                        double v = 0.0;
                        if (kin_range && j < w.N)
                            v = BKM ? w.B[(size_t)kt2 * d.bKTileStride + kin2 + (size_t)j * w.ldb]
                                    : w.B[(size_t)kt2 * d.bKTileStride + (size_t)kin2 * w.ldb + j];
                        dB[BKM ? j * LDK + r : ((j >> 4) * BK + r) * 16 + ((((j & 15) >> 1) ^ (r & 7)) << 1) + (j & 1)] = v;
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&full[s]);
            }
        }
        return;
    }

    // ================================ consumer warps ================================
    const int wm = warp % Cfg::WARPS_M, wn = warp / Cfg::WARPS_M;
    const int g = lane >> 2, t = lane & 3;
    const int wm0 = wm * Cfg::WM, wn0 = wn * Cfg::WN;

    // Accumulators start as (beta/alpha)*C: the read of C overlaps the pipeline fill instead of
    // sitting, un-overlapped, in the epilogue; the epilogue only stores alpha*acc.
    // thread holds, for fragment (i,j): row crow(i); columns ccol(2t, j) and ccol(2t+1, j)
    // (the C fragment's column index plays the role of B's fragment row g)
    const double alpha = d.alpha, beta = d.beta;
    // diagonal (SYRK) tiles carry the large diagonal entries of the matrix: never preload those
    const bool pre = d.preload && (w.lower == 0);
    const double cscale = (beta == 0.0 || !pre) ? 0.0 : beta / alpha;
    const double ebeta = pre ? 0.0 : beta;  // what the epilogue still has to add
    const bool vec = !AK && (w.M == BM) && (w.N == BN) && (w.lower == 0) && ((w.ldc & 1) == 0) &&
                     ((reinterpret_cast<uintptr_t>(w.C) & 15) == 0);
    // fragment -> tile coordinates (K-major operands use the natural 8f+g fragment rows)
    auto crow = [&](int i) { return wm0 + (AK ? 8 * i + g : mrow(g, i)); };
    auto ccol = [&](int c, int j) { return wn0 + (BKM ? 8 * j + c : mrow(c, j)); };
    double acc[MF][NF][2];
    if (cscale == 0.0) {
#pragma unroll
        for (int i = 0; i < MF; i++)
#pragma unroll
            for (int j = 0; j < NF; j++) acc[i][j][0] = acc[i][j][1] = 0.0;
    } else if (vec) {
#pragma unroll
        for (int j = 0; j < NF; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const double* cp = w.C + (size_t)ccol(2 * t + h, j) * w.ldc + wm0;
#pragma unroll
                for (int i = 0; i < MF; i += 2) {
                    const double2 x = __ldcs(reinterpret_cast<const double2*>(cp + mrow(g, i)));  // C streams: evict first
                    acc[i][j][h] = cscale * x.x;
                    acc[i + 1][j][h] = cscale * x.y;
                }
            }
    } else {
#pragma unroll
        for (int j = 0; j < NF; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int col = ccol(2 * t + h, j);
                const double* cp = w.C + (size_t)col * w.ldc;
                const int gj = w.gj0 + col;
#pragma unroll
                for (int i = 0; i < MF; i++) {
                    const int row = crow(i);
                    bool ok = (col < w.N) && (row < w.M);
                    if (w.lower == 1 && (w.gi0 + row < gj)) ok = false;
                    if (w.lower == 2 && (w.gi0 + row > gj)) ok = false;
                    acc[i][j][h] = ok ? cscale * cp[row] : 0.0;
                }
            }
    }

    // ---- fragment loads: per-thread byte offsets inside a stage, everything else is an immediate ----
    // M-major (swizzled blocks): k-row 4ks+t, chunk pair_of(g) ^ ((4ks+t) & 7); the chunk of odd ks differs from the
    // chunk of even ks in bit 2 only.  K-major: row 8v+g, k = 4ks+t.
    const uint32_t aE = AK ? (uint32_t)((wm0 + g) * LDK + t) * 8u
                           : (uint32_t)((wm0 / Cfg::MBLK) * MBB + t * 128 + ((pair_of(g) ^ t) << 4));
    const uint32_t bE = BKM ? (uint32_t)((wn0 + g) * LDK + t) * 8u
                            : (uint32_t)((wn0 / Cfg::MBLK) * MBB + t * 128 + ((pair_of(g) ^ t) << 4));
    const uint32_t aO = AK ? aE : (aE ^ 64u), bO = BKM ? bE : (bE ^ 64u);
    auto load_frags = [&](double (&a)[MF], double (&b)[NF], uint32_t stA, uint32_t stB, int ks) {
        const uint32_t pa = stA + ((ks & 1) ? aO : aE), pb = stB + ((ks & 1) ? bO : bE);
        if (AK) {
#pragma unroll
            for (int v = 0; v < MF; v++) lds64(a[v], pa + (uint32_t)(v * 8 * LDK + ks * 4) * 8u);
        } else {
#pragma unroll
            for (int v = 0; v < MF / 2; v++) lds128(a[2 * v], a[2 * v + 1], pa + (uint32_t)(v * MBB + ks * 512));
        }
        if (BKM) {
#pragma unroll
            for (int v = 0; v < NF; v++) lds64(b[v], pb + (uint32_t)(v * 8 * LDK + ks * 4) * 8u);
        } else {
#pragma unroll
            for (int v = 0; v < NF / 2; v++) lds128(b[2 * v], b[2 * v + 1], pb + (uint32_t)(v * MBB + ks * 512));
        }
    };
    auto mma_all = [&](const double (&a)[MF], const double (&b)[NF]) {
#pragma unroll
        for (int i = 0; i < MF; i++)
#pragma unroll
            for (int j = 0; j < NF; j++) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    };

    // ---- main loop, software-pipelined by hand: while the DMMAs of one k-step issue, the fragments of the next
    // k-step -- across the stage boundary too -- are already on their way from shared memory, so a warp never sits
    // at the top of a stage with an empty DMMA queue (with three warps per scheduler marching in step, that idle
    // gap was the 8 % the first version of this kernel lost) ----
    constexpr uint32_t A_BYTES = A_STAGE * 8, B_BYTES = B_STAGE * 8;
    const uint32_t sA0 = smem_u32(sA), sB0 = smem_u32(sB);
    uint32_t stA = sA0, stB = sB0;
    int s = 0;
    uint32_t ph = 0;
    double a0[MF], b0[NF], a1[MF], b1[NF];
    if (nChunks > 0) {
        mbar_wait(&full[0], 0);
        load_frags(a0, b0, stA, stB, 0);
    }
    for (int c = 0; c < nChunks; c++) {
        int ns = s + 1;
        uint32_t nA = stA + A_BYTES, nB = stB + B_BYTES, nph = ph;
        if (ns == STAGES) { ns = 0; nA = sA0; nB = sB0; nph ^= 1; }
#pragma unroll
        for (int ks = 0; ks < BK / 4; ks += 2) {
            load_frags(a1, b1, stA, stB, ks + 1);
            mma_all(a0, b0);
            if (ks + 2 < BK / 4) {
                load_frags(a0, b0, stA, stB, ks + 2);
            } else if (c + 1 < nChunks) {
                mbar_wait(&full[ns], nph);
                load_frags(a0, b0, nA, nB, 0);
            }
            mma_all(a1, b1);
        }
        // Every LDS of this stage has fed a DMMA that has been issued (in-order issue, and all of these are
        // volatile asm: the compiler keeps the order written here), so the stage can go back to the producer.
        // An arrive placed right after the last LDS instead could overtake that load (round-1 finding).
        if (lane == 0) mbar_arrive(&empty[s]);
        s = ns; stA = nA; stB = nB; ph = nph;
    }

    // ================================ epilogue ================================
    if (vec) {
#pragma unroll
        for (int j = 0; j < NF; j++)
#pragma unroll
            for (int h = 0; h < 2; h++) {
                double* cp = w.C + (size_t)ccol(2 * t + h, j) * w.ldc + wm0;
#pragma unroll
                for (int i = 0; i < MF; i += 2) {
                    double2 x;
                    x.x = alpha * acc[i][j][h];
                    x.y = alpha * acc[i + 1][j][h];
                    if (ebeta != 0.0) {
                        const double2 c0 = __ldcs(reinterpret_cast<const double2*>(cp + mrow(g, i)));
                        x.x += ebeta * c0.x;
                        x.y += ebeta * c0.y;
                    }
                    __stcs(reinterpret_cast<double2*>(cp + mrow(g, i)), x);
                }
            }
        return;
    }
#pragma unroll
    for (int j = 0; j < NF; j++) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int col = ccol(2 * t + h, j);
            if (col >= w.N) continue;
            double* cp = w.C + (size_t)col * w.ldc;
            const int gj = w.gj0 + col;
#pragma unroll
            for (int i = 0; i < MF; i++) {
                const int row = crow(i);
                if (row >= w.M) continue;
                if (w.lower == 1 && (w.gi0 + row < gj)) continue;
                if (w.lower == 2 && (w.gi0 + row > gj)) continue;
                double v = alpha * acc[i][j][h];
                if (ebeta != 0.0) v += ebeta * cp[row];
                cp[row] = v;
            }
        }
    }
}

// host side (gemm_dmma.cu)
int gemm_launch(const GemmDesc& d, int nItems, int cfg, cudaStream_t stream);
int gemm_pick_cfg(long long bigItems);
void gemm_cfg_dims(int cfg, int* bm, int* bn);
int gemm_init_attributes();

}  // namespace dpb
