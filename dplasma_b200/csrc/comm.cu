// Inter-GPU communication: one process per GPU, NCCL over NVLink for the panel broadcasts
// that PaRSEC's remote-dependency engine performs over MPI in the reference
// (src/zpotrf_L.jdf:112,213-215; MPI_Allreduce of info at src/zpotrf_wrapper.c:331-332).
// NCCL is resolved at run time from the process (torch loads libnccl.so.2) or the loader path.
#include <cstdio>
#include <cstring>
#include <dlfcn.h>
#include "internal.h"

namespace dpb {

typedef struct { char internal[128]; } nccl_uid_t;
typedef void* nccl_comm_t;
enum { NCCL_INT = 2, NCCL_DOUBLE = 8, NCCL_SUM = 0, NCCL_MAX = 2 };

struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(nccl_uid_t*) = nullptr;
    int (*CommInitRank)(nccl_comm_t*, int, nccl_uid_t, int) = nullptr;
    int (*CommSplit)(nccl_comm_t, int, int, nccl_comm_t*, void*) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
};

static NcclApi* load_nccl() {
    static NcclApi api;
    if (api.h) return &api;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        fprintf(stderr, "dplasma_b200: cannot load libnccl.so.2: %s\n", dlerror());
        return nullptr;
    }
#define LOAD(name) *(void**)(&api.name) = dlsym(h, "nccl" #name); if (!api.name) { fprintf(stderr, "dplasma_b200: nccl" #name " missing\n"); return nullptr; }
    LOAD(GetUniqueId) LOAD(CommInitRank) LOAD(CommSplit) LOAD(CommDestroy) LOAD(Broadcast) LOAD(AllReduce)
    LOAD(GroupStart) LOAD(GroupEnd) LOAD(GetErrorString) LOAD(Send) LOAD(Recv)
#undef LOAD
    api.h = h;
    return &api;
}

int info_allreduce_max_impl(parsec_context_s* ctx, int info) {
    if (ctx->nranks <= 1 || !ctx->commWorld) return info;
    cudaStream_t s = ctx->sPanel;
    *ctx->hInfo = info;
    cudaMemcpyAsync(ctx->dInfo + 1, ctx->hInfo, sizeof(int), cudaMemcpyHostToDevice, s);
    int rc = ctx->nccl->AllReduce(ctx->dInfo + 1, ctx->dInfo + 1, 1, NCCL_INT, NCCL_MAX, ctx->commWorld, s);
    if (rc) fprintf(stderr, "dplasma_b200: ncclAllReduce failed: %s\n", ctx->nccl->GetErrorString(rc));
    cudaMemcpyAsync(ctx->hInfo, ctx->dInfo + 1, sizeof(int), cudaMemcpyDeviceToHost, s);
    cudaStreamSynchronize(s);
    return *ctx->hInfo;
}

int sum_allreduce(parsec_context_s* ctx, double* dbuf, size_t n, cudaStream_t s) {
    if (ctx->nranks <= 1 || !ctx->commWorld) return 0;
    int rc = ctx->nccl->AllReduce(dbuf, dbuf, n, NCCL_DOUBLE, NCCL_SUM, ctx->commWorld, s);
    if (rc) { fprintf(stderr, "dplasma_b200: ncclAllReduce failed: %s\n", ctx->nccl->GetErrorString(rc)); return -300; }
    return 0;
}

int bcast(parsec_context_s* ctx, void* comm, const double* send, double* recv, size_t count, int root, cudaStream_t s) {
    int rc = ctx->nccl->Broadcast(send, recv, count, NCCL_DOUBLE, root, comm, s);
    if (rc) { fprintf(stderr, "dplasma_b200: ncclBroadcast failed: %s\n", ctx->nccl->GetErrorString(rc)); return -301; }
    return 0;
}
int p2p_send(parsec_context_s* ctx, const double* buf, size_t count, int peer, cudaStream_t s) {
    int rc = ctx->nccl->Send(buf, count, NCCL_DOUBLE, peer, ctx->commWorld, s);
    if (rc) { fprintf(stderr, "dplasma_b200: ncclSend failed: %s\n", ctx->nccl->GetErrorString(rc)); return -305; }
    return 0;
}
int p2p_recv(parsec_context_s* ctx, double* buf, size_t count, int peer, cudaStream_t s) {
    int rc = ctx->nccl->Recv(buf, count, NCCL_DOUBLE, peer, ctx->commWorld, s);
    if (rc) { fprintf(stderr, "dplasma_b200: ncclRecv failed: %s\n", ctx->nccl->GetErrorString(rc)); return -306; }
    return 0;
}
int group_start(parsec_context_s* ctx) { return ctx->nccl->GroupStart(); }
int group_end(parsec_context_s* ctx) { return ctx->nccl->GroupEnd(); }

// row / column communicators of a P x Q grid are split lazily, once per grid shape.
// Ranks inside the row / column communicators are the LOGICAL grid coordinates of the distribution (the
// process that owns tile (0,0) is rank 0 of both), so that the broadcast roots k % Q, n % P of the step plan
// are communicator ranks as they stand, whatever RSRC / CSRC (grid.ip / grid.jq) the descriptor carries.
int ensure_grid_comms(parsec_context_s* ctx, int P, int Q, int ip, int jq) {
    if (ctx->gridP == P && ctx->gridQ == Q && ctx->gridIp == ip && ctx->gridJq == jq && ctx->commRow && ctx->commCol) return 0;
    if (!ctx->commWorld) { fprintf(stderr, "dplasma_b200: dplasma_b200_comm_init was not called\n"); return -302; }
    if (P * Q != ctx->nranks) { fprintf(stderr, "dplasma_b200: grid %dx%d does not match %d ranks\n", P, Q, ctx->nranks); return -303; }
    if (ctx->commRow) ctx->nccl->CommDestroy(ctx->commRow);
    if (ctx->commCol) ctx->nccl->CommDestroy(ctx->commCol);
    const int prow = ctx->rank / Q, pcol = ctx->rank % Q;
    const int lrow = (prow - ip % P + P) % P, lcol = (pcol - jq % Q + Q) % Q;
    int rc = ctx->nccl->CommSplit(ctx->commWorld, prow, lcol, &ctx->commRow, nullptr);  // ranks of my grid row, keyed by logical column
    if (!rc) rc = ctx->nccl->CommSplit(ctx->commWorld, pcol, lrow, &ctx->commCol, nullptr);  // my grid column, keyed by logical row
    if (rc) { fprintf(stderr, "dplasma_b200: ncclCommSplit failed: %s\n", ctx->nccl->GetErrorString(rc)); return -304; }
    ctx->gridP = P; ctx->gridQ = Q; ctx->gridIp = ip; ctx->gridJq = jq;
    return 0;
}

void comm_destroy(parsec_context_s* ctx) {
    if (!ctx->nccl) return;
    if (ctx->commRow) ctx->nccl->CommDestroy(ctx->commRow);
    if (ctx->commCol) ctx->nccl->CommDestroy(ctx->commCol);
    if (ctx->commWorld) ctx->nccl->CommDestroy(ctx->commWorld);
    ctx->commRow = ctx->commCol = ctx->commWorld = nullptr;
}

}  // namespace dpb

using namespace dpb;

extern "C" {

int info_allreduce_max(parsec_context_s* ctx, int info) { return info_allreduce_max_impl(ctx, info); }

int dplasma_b200_get_unique_id(void* id128) {
    NcclApi* api = load_nccl();
    if (!api || !id128) return -1;
    nccl_uid_t id;
    int rc = api->GetUniqueId(&id);
    if (rc) return -2;
    memcpy(id128, &id, sizeof(id));
    return 0;
}

int dplasma_b200_comm_init(parsec_context_t* ctx, const void* id128, int rank, int nranks) {
    if (!ctx || !id128 || nranks < 1 || rank < 0 || rank >= nranks) return -1;
    NcclApi* api = load_nccl();
    if (!api) return -2;
    cudaSetDevice(ctx->device);
    nccl_uid_t id;
    memcpy(&id, id128, sizeof(id));
    nccl_comm_t comm = nullptr;
    int rc = api->CommInitRank(&comm, nranks, id, rank);
    if (rc) {
        fprintf(stderr, "dplasma_b200: ncclCommInitRank failed: %s\n", api->GetErrorString(rc));
        return -3;
    }
    ctx->nccl = api;
    ctx->commWorld = comm;
    ctx->rank = rank;
    ctx->nranks = nranks;
    return 0;
}

}  // extern "C"
