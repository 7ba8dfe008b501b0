// dplasma_b200_dpotrf_tile_noalloc -- lower Cholesky of one tile for callers that must not
// allocate.
//
// Who needs this: code that runs where device allocation is off limits -- a task body inside a
// runtime that owns all GPU memory itself and hands the body nothing but its input/output tiles
// (PaRSEC's device module pre-reserves HBM in its zone allocator, so a cudaMalloc from a task can
// simply fail once the zone has taken the card), a latency-critical loop that cannot afford the
// implicit device synchronisation cudaMalloc/cudaFree bring, or a process running under a fixed
// memory budget where every byte was handed out at start-up.
//
// The regular entry point (dplasma_b200_dpotrf_tile) keeps the inverses of the 64x64 diagonal
// blocks of the factor in a workspace it cudaMallocs on first use and grows on demand.  This
// variant keeps that workspace in a fixed array that is part of the library image instead:
// 16 blocks x 64 x 64 doubles = 512 KiB, which bounds the tile size at n <= 1024 (the largest
// tile size of the BASELINE configurations).  Every GPU that loads the library carries those
// 512 KiB until the process exits, whether or not this function is ever called.
//
// Rules for the caller:
//  * The first call on each device looks up the device address of the array and raises the
//    shared-memory limits of the kernels it launches.  Make that first call outside any stream
//    capture, with the device you will use current.
//  * All calls on one device share the one array and it keeps whatever the last call left in
//    it.  Nothing in it has to start at zero: each 64x64 block is rewritten in full (zeros
//    included) by the leaf factorisation before any product reads it.  Two calls whose GPU work
//    overlaps in time (different streams, or different host threads) would overwrite each
//    other's blocks -- serialise them.
//  * Nothing is allocated or freed here: no cudaMalloc/cudaFree, no stream-ordered allocation,
//    no library handle creation.  Inputs and outputs are the caller's.
#include <cstdio>
#include "internal.h"

namespace {

constexpr int kNoAllocMaxN = 1024;
constexpr int kNoAllocBlocks = kNoAllocMaxN / dpb::IB;
constexpr int kMaxDevices = 64;

// The scratch itself.  In host code this name is only a handle for the runtime, never an
// address a kernel may be given: the device address comes from cudaGetSymbolAddress below.
static __device__ __align__(256) double g_noalloc_winv[kNoAllocBlocks * dpb::IB * dpb::IB];

static double* g_scratch_addr[kMaxDevices] = {nullptr};

// once per device, with that device current
static int noalloc_setup(int device, double** out) {
    if (g_scratch_addr[device]) {
        *out = g_scratch_addr[device];
        return 0;
    }
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_winv);
    if (e != cudaSuccess || !p) {
        fprintf(stderr, "dplasma_b200_dpotrf_tile_noalloc: cudaGetSymbolAddress failed on device %d: %s\n", device,
                cudaGetErrorString(e));
        return e != cudaSuccess ? (int)e : -1000;
    }
    int rc = dpb::gemm_init_attributes();  // dynamic shared-memory opt-in of the GEMM kernels; allocates nothing
    if (rc) {
        fprintf(stderr, "dplasma_b200_dpotrf_tile_noalloc: kernel attribute set-up failed (%d)\n", rc);
        return rc;
    }
    g_scratch_addr[device] = p;
    *out = p;
    return 0;
}

}  // namespace

extern "C" int dplasma_b200_dpotrf_tile_noalloc(void* stream, dplasma_enum_t uplo, int n, double* A, int lda,
                                                int* d_info, int info_offset) {
    if (uplo != dplasmaLower) {
        fprintf(stderr, "dplasma_b200_dpotrf_tile_noalloc: only dplasmaLower is provided (got %d)\n", uplo);
        return -2;
    }
    if (n < 0) return -3;
    if (n > kNoAllocMaxN) {
        fprintf(stderr,
                "dplasma_b200_dpotrf_tile_noalloc: n = %d exceeds the built-in workspace (n <= %d); "
                "use dplasma_b200_dpotrf_tile, which allocates its workspace\n", n, kNoAllocMaxN);
        return -3;
    }
    if (!A && n > 0) return -4;
    if (lda < (n > 1 ? n : 1)) return -5;
    if (!d_info) {
        fprintf(stderr, "dplasma_b200_dpotrf_tile_noalloc: d_info (a device int) is required\n");
        return -6;
    }
    if (n == 0) return 0;
    int device = -1;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess || device < 0 || device >= kMaxDevices) {
        fprintf(stderr, "dplasma_b200_dpotrf_tile_noalloc: no usable CUDA device (%s)\n", cudaGetErrorString(e));
        return e != cudaSuccess ? (int)e : -1001;
    }
    double* winv = nullptr;
    int rc = noalloc_setup(device, &winv);
    if (rc) return rc;
    // a context-shaped counter object on the host stack: the shared launch helpers only bump
    // its launch counter; it owns no CUDA resources
    parsec_context_s counter;
    counter.device = device;
    return dpb::potrf_tile_lower(&counter, (cudaStream_t)stream, A, lda, n, winv, d_info, info_offset);
}
