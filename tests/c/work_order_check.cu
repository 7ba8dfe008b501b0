// Host-side check of the GEMM engine's work order (decode_work in dplasma_b200/csrc/gemm_dmma.cuh): whatever the
// order, the CTA tiles of one launch must cover every 64x64 block of every tile of the update exactly once.
// Compiled with nvcc, runs on the CPU (no kernel launch).  Exit code 0 = every case passed.
#include <cstdio>
#include <cstring>
#include <map>
#include <set>
#include <tuple>
#include <vector>
#include "../../dplasma_b200/csrc/gemm_dmma.cuh"
using namespace dpb;

static int fails = 0;

// trailing update of an nt x nt tile matrix after step k0 on rank (prow, pcol) of a P x Q grid
static void trailing_case(int nt, int nb, int N, int P, int Q, int prow, int pcol, int k0, int group, bool full) {
    GemmDesc d;
    memset(&d, 0, sizeof(d));
    d.mode = GEMM_TRAILING;
    d.nb = nb; d.K = nb;
    d.P = P; d.Q = Q; d.prow = prow; d.pcol = pcol;
    d.mtGlobal = d.ntGlobal = nt;
    d.lastM = d.lastN = N - (nt - 1) * nb;
    d.full = full;
    auto firstGE = [](int g, int me, int np) { return g <= me ? 0 : (g - me + np - 1) / np; };
    const int lmt = firstGE(nt, prow, P), lnt = firstGE(nt, pcol, Q);
    d.tileLdT = lmt; d.tileElems = (long long)nb * nb;
    d.rowEnd = lmt;
    const int jl0 = firstGE(k0 + 1, pcol, Q);
    d.colFirst = jl0; d.ncols = lnt - jl0;
    if (d.ncols <= 0 || d.ncols > kMaxCols) return;
    d.bptM = d.bptN = (nb + 63) / 64;
    long long items = 0;
    for (int c = 0; c < d.ncols; c++) {
        const int gn = (jl0 + c) * Q + pcol;
        d.rowFirst[c] = full ? 0 : std::min(firstGE(gn, prow, P), lmt);
        d.itemStart[c] = (int)items;
        items += (long long)(lmt - d.rowFirst[c]) * d.bptM * d.bptN;
    }
    d.itemStart[d.ncols] = (int)items;
    d.colGroup = group;
    d.Cbase = (double*)0x100000; d.rowPanel = (const double*)0x10; d.colPanel = (const double*)0x20;
    d.rowPanelStride = d.colPanelStride = d.tileElems;
    // expected: every (local tile i >= rowFirst[c], column c, bi, bj); blocks outside a ragged tile come back invalid
    std::set<std::tuple<long long, int, int>> seen;
    long long valid = 0, expectValid = 0;
    for (int c = 0; c < d.ncols; c++)
        for (int i = d.rowFirst[c]; i < lmt; i++) {
            const int gm = i * P + prow, gn = (jl0 + c) * Q + pcol;
            const int tm = gm == nt - 1 ? d.lastM : nb, tn = gn == nt - 1 ? d.lastN : nb;
            for (int bi = 0; bi * 64 < tm; bi++)
                for (int bj = 0; bj * 64 < tn; bj++) {
                    if (gm == gn && !full && bi * 64 + std::min(64, tm - bi * 64) - 1 < bj * 64) continue;
                    expectValid++;
                }
        }
    for (long long bid = 0; bid < items; bid++) {
        const BlockWork w = decode_work<CfgSmall>(d, (int)bid);
        if (!w.valid) continue;
        valid++;
        if (!seen.insert(std::make_tuple((long long)(w.C - d.Cbase), w.M, w.N)).second) {
            if (fails++ < 5) printf("trailing nt=%d nb=%d P=%d Q=%d k0=%d group=%d: block visited twice (bid %lld)\n", nt, nb, P, Q, k0, group, bid);
            return;
        }
        // operand pointers must belong to the tile row / column of the C block
        const long long off = w.C - d.Cbase, tile = off / d.tileElems, in = off % d.tileElems;
        const int i = (int)(tile % lmt), j = (int)(tile / lmt), i0 = (int)(in % nb), j0 = (int)(in / nb);
        if (w.A != d.rowPanel + (long long)i * d.rowPanelStride + i0 || w.B != d.colPanel + (long long)j * d.colPanelStride + j0 ||
            w.aTile != i || w.bTile != j || w.aRow != i0 || w.bRow != j0) {
            if (fails++ < 5) printf("trailing nt=%d nb=%d group=%d: operands of bid %lld do not match its C block\n", nt, nb, group, bid);
            return;
        }
    }
    if (valid != expectValid) {
        if (fails++ < 5) printf("trailing nt=%d nb=%d N=%d P=%d Q=%d (%d,%d) k0=%d group=%d full=%d: %lld blocks, expected %lld\n", nt, nb, N, P, Q,
                                prow, pcol, k0, group, (int)full, valid, expectValid);
    }
}

static void single_case(int M, int N, int batch, int lastM, int lower, int group) {
    GemmDesc d;
    memset(&d, 0, sizeof(d));
    d.mode = GEMM_SINGLE;
    d.M = M; d.N = N; d.lda = d.ldb = d.ldc = M + 2; d.K = 64;
    d.gridM = (M + 63) / 64; d.gridN = (N + 63) / 64;
    d.batch = batch; d.lastBatchM = lastM; d.lower = lower;
    d.batchStrideA = d.batchStrideC = (long long)(M + 2) * (N + 7);
    d.batchStrideB = 0;
    d.A = (const double*)0x10; d.B = (const double*)0x20; d.C = (double*)0x100000;
    d.blkGroup = group;
    std::set<long long> seen;
    long long valid = 0, expectValid = 0;
    for (int bt = 0; bt < batch; bt++) {
        const int Mt = bt == batch - 1 ? lastM : M;
        for (int bi = 0; bi * 64 < Mt; bi++)
            for (int bj = 0; bj * 64 < N; bj++) {
                const int m = std::min(64, Mt - bi * 64), n = std::min(64, N - bj * 64);
                if (lower == 1 && bi * 64 + m - 1 < bj * 64) continue;
                if (lower == 2 && bj * 64 + n - 1 < bi * 64) continue;
                expectValid++;
            }
    }
    const long long items = (long long)d.gridM * d.gridN * batch;
    for (long long bid = 0; bid < items; bid++) {
        const BlockWork w = decode_work<CfgSmall>(d, (int)bid);
        if (!w.valid) continue;
        valid++;
        if (!seen.insert((long long)(w.C - d.C)).second) {
            if (fails++ < 5) printf("single M=%d N=%d batch=%d group=%d: block visited twice\n", M, N, batch, group);
            return;
        }
        const long long off = (w.C - d.C) % d.batchStrideC;
        const int i0 = (int)(off % d.ldc), j0 = (int)(off / d.ldc);
        if (w.aRow != i0 || w.bRow != j0 || w.bTile != 0 || w.aTile != (int)((w.C - d.C) / d.batchStrideC)) {
            if (fails++ < 5) printf("single M=%d N=%d group=%d: operand coordinates do not match the C block\n", M, N, group);
            return;
        }
    }
    if (valid != expectValid)
        if (fails++ < 5) printf("single M=%d N=%d batch=%d lastM=%d lower=%d group=%d: %lld blocks, expected %lld\n", M, N, batch, lastM, lower, group, valid, expectValid);
}

int main() {
    int cases = 0;
    const int grids[][2] = {{1, 1}, {1, 2}, {2, 1}, {2, 2}, {2, 4}, {3, 2}};
    for (auto& pq : grids)
        for (int prow = 0; prow < pq[0]; prow++)
            for (int pcol = 0; pcol < pq[1]; pcol++)
                for (int group = 0; group <= 8; group++)
                    for (int k0 : {0, 1, 2, 5, 9})
                        for (int full = 0; full < 2; full++) {
                            trailing_case(13, 128, 13 * 128 - 37, pq[0], pq[1], prow, pcol, k0, group, full);
                            trailing_case(11, 512, 11 * 512, pq[0], pq[1], prow, pcol, k0, group, full);
                            trailing_case(9, 200, 9 * 200 - 150, pq[0], pq[1], prow, pcol, k0, group, full);
                            cases += 3;
                        }
    for (int group : {0, 1, 3, 16, 64})
        for (int lower = 0; lower < 3; lower++) {
            single_case(1000, 1000, 1, 1000, lower, group);
            single_case(64, 4096, 1, 64, 0, group);
            single_case(512, 200, 5, 130, 0, group);
            single_case(2048, 1100, 2, 2048, lower, group);
            cases += 4;
        }
    printf("%d cases, %d failures\n", cases, fails);
    return fails ? 1 : 0;
}
