"""Host-side check of the launch schedule of dplasma_b200/csrc/blas3.cu (dplasma_dgemm / dplasma_dsyrk): the
same transposes and batched products (offsets, strides, leading dimensions) replayed with numpy on flat
tile-storage buffers and compared with numpy.  Development aid; runs on the CPU."""
import numpy as np

from sim_inverse_schedule import Mem, gemm_single


class Grid:
    """tile storage of an M x N matrix with mb x nb tiles at offset `off` of mem"""

    def __init__(self, mem, off, M, N, mb, nb):
        self.mem, self.off, self.M, self.N, self.mb, self.nb = mem, off, M, N, mb, nb
        self.mt, self.nt = -(-M // mb), -(-N // nb)
        self.te = mb * nb
        self.size = self.mt * self.nt * self.te

    def tile(self, i, j):
        return self.off + (i + j * self.mt) * self.te

    def rows(self, i):
        return self.M - i * self.mb if i == self.mt - 1 else self.mb

    def cols(self, j):
        return self.N - j * self.nb if j == self.nt - 1 else self.nb

    def put(self, X):
        for j in range(self.nt):
            for i in range(self.mt):
                self.mem.mat(self.tile(i, j), self.mb, self.rows(i), self.cols(j))[:, :] = \
                    X[i * self.mb:i * self.mb + self.rows(i), j * self.nb:j * self.nb + self.cols(j)]

    def get(self):
        X = np.zeros((self.M, self.N))
        for j in range(self.nt):
            for i in range(self.mt):
                X[i * self.mb:i * self.mb + self.rows(i), j * self.nb:j * self.nb + self.cols(j)] = \
                    self.mem.mat(self.tile(i, j), self.mb, self.rows(i), self.cols(j))
        return X


def transpose_tiles(mem, src, sstride, lds, dst, dstride, ldd, rows, cols, cnt):
    for t in range(cnt):
        S = mem.mat(src + t * sstride, lds, rows, cols)
        S = np.where(np.isnan(S), 0.0, S)  # padding of ragged tiles is never used as data
        mem.mat(dst + t * dstride, ldd, cols, rows)[:, :] = S.T


def dgemm(tA, tB, M, N, K, mb, nb, kb, alpha, beta, rng):
    A = rng.standard_normal((K, M) if tA else (M, K))
    B = rng.standard_normal((N, K) if tB else (K, N))
    C = rng.standard_normal((M, N))
    mem = Mem(4 * (M + mb) * (N + nb) + 4 * (M + mb) * (K + kb) + 4 * (K + kb) * (N + nb) + 16 * mb * nb * kb)
    gA = Grid(mem, 0, *A.shape, kb if tA else mb, mb if tA else kb)
    gB = Grid(mem, gA.size, *B.shape, nb if tB else kb, kb if tB else nb)
    gC = Grid(mem, gA.size + gB.size, M, N, mb, nb)
    gA.put(A); gB.put(B); gC.put(C)
    mt, nt, kt = gC.mt, gC.nt, -(-K // kb)
    teAt, teBt = mb * kb, nb * kb
    At = gC.off + gC.size
    Bt = At + mt * teAt
    for k in range(kt):
        tk = K - k * kb if k == kt - 1 else kb
        if tA:
            transpose_tiles(mem, gA.tile(k, 0), gA.mt * gA.te, gA.mb, At, teAt, mb, tk, mb, mt)
        if not tB:
            transpose_tiles(mem, gB.tile(k, 0), gB.mt * gB.te, gB.mb, Bt, teBt, nb, tk, nb, nt)
        Aop, sA = (At, teAt) if tA else (gA.tile(0, k), gA.te)
        for n in range(nt):
            Bop = gB.tile(n, k) if tB else Bt + n * teBt
            gemm_single(mem, mb, gC.cols(n), tk, alpha, Aop, mb, False, Bop, nb, False, beta if k == 0 else 1.0,
                        gC.tile(0, n), mb, 0, mt, sA, 0, gC.te, gC.rows(mt - 1))
    want = alpha * (A.T if tA else A) @ (B.T if tB else B) + beta * C
    return np.abs(gC.get() - want).max()


if __name__ == "__main__":
    rng = np.random.default_rng(0)
    for tA in (False, True):
        for tB in (False, True):
            print("dgemm", tA, tB, dgemm(tA, tB, 150, 130, 170, 32, 48, 40, 0.7, -0.3, rng),
                  dgemm(tA, tB, 128, 128, 128, 64, 64, 64, 1.0, 0.0, rng))
    print("schedule ok")
