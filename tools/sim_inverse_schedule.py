"""Host-side check of the launch schedule of dplasma_b200/csrc/potri.cu: the same sequence of
batched products (operand offsets, strides, K-major flags, masks) replayed with numpy on a flat
tile-storage buffer and compared with numpy's inverse.  Development aid; runs on the CPU."""
import sys

import numpy as np


class Mem:
    def __init__(self, n):
        self.a = np.full(n, np.nan)

    def mat(self, off, ld, rows, cols):
        return np.lib.stride_tricks.as_strided(self.a[off:], shape=(rows, cols), strides=(8, 8 * ld))


def gemm_single(mem, M, N, K, alpha, A, lda, aK, B, ldb, bK, beta, Cc, ldc, lower, batch=1, sA=0, sB=0, sC=0, lastM=-1):
    if lastM < 0:
        lastM = M
    outs = []
    for b in range(batch):
        m = lastM if b == batch - 1 else M
        Am = mem.mat(A + b * sA, lda, K, m).T if aK else mem.mat(A + b * sA, lda, m, K)
        Bm = mem.mat(B + b * sB, ldb, K, N) if bK else mem.mat(B + b * sB, ldb, N, K).T
        Cm = mem.mat(Cc + b * sC, ldc, m, N)
        R = alpha * (Am @ Bm) + (beta * Cm if beta != 0 else 0)
        outs.append((Cm, R, m))
    for Cm, R, m in outs:  # all products of a launch read before any writes (no operand aliases its output)
        if lower == 1:
            mask = np.tril(np.ones((m, N), bool))
            Cm[mask] = R[mask]
        else:
            Cm[:, :] = R


def run(N, nb, diag_unit=False):
    nt = -(-N // nb)
    te = nb * nb
    lmt = nt
    cs = lmt * te
    rng = np.random.default_rng(0)
    G = rng.standard_normal((N, N))
    Afull = G @ G.T + N * np.eye(N)
    L = np.linalg.cholesky(Afull)
    if diag_unit:  # keep the unit-diagonal triangle well conditioned
        L = np.tril(L, -1) / N + np.diag(np.diag(L))
    Lin = L.copy()
    U0 = rng.standard_normal((N, N))
    Lin += np.triu(U0, 1)  # junk in the other triangle: must be ignored and preserved
    mem = Mem(nt * nt * te + te + nt * te + te)
    W = nt * nt * te
    Wp = W + te

    def tile(i, j):
        return (i + j * lmt) * te

    def rows(t):
        return N - t * nb if t == nt - 1 else nb

    for j in range(nt):
        for i in range(nt):
            blk = Lin[i * nb:(i + 1) * nb, j * nb:(j + 1) * nb]
            mem.mat(tile(i, j), nb, blk.shape[0], blk.shape[1])[:, :] = blk
    ld = nb
    W2 = Wp + nt * te

    def transpose_tiles(src, sstride, dst, dstride, rows_, cols_, cnt):
        for t in range(cnt):
            mem.mat(dst + t * dstride, ld, cols_, rows_)[:, :] = mem.mat(src + t * sstride, ld, rows_, cols_).T

    def trailing_update(mt_view, rowPanel, rowStride, colPanel, colStride, gn0, gn1, alpha, K, rowFirstFixed=-1, rowEnd=None):
        rowEnd = mt_view if rowEnd is None else rowEnd
        pend = []
        for n_ in range(gn0, gn1):
            first = rowFirstFixed if rowFirstFixed >= 0 else n_
            for m_ in range(first, rowEnd):
                tm = rows(m_) if mt_view == nt else nb
                Am = mem.mat(rowPanel + m_ * rowStride, nb, tm, K)
                Bm = mem.mat(colPanel + n_ * colStride, nb, nb, K)
                Cm = mem.mat(tile(m_, n_), nb, tm, nb)
                pend.append((Cm, Cm + alpha * (Am @ Bm.T), m_ == n_))
        for Cm, R, dg in pend:
            if dg:
                mask = np.tril(np.ones(Cm.shape, bool))
                Cm[mask] = R[mask]
            else:
                Cm[:, :] = R

    # ---- trtri
    for k in range(nt):
        tk, below = rows(k), nt - k - 1
        Akk = mem.mat(tile(k, k), ld, tk, tk)
        T = np.tril(Akk)
        if diag_unit:
            T = np.tril(Akk, -1) + np.eye(tk)
        mem.mat(W, ld, tk, tk)[:, :] = np.linalg.inv(T)  # identity + trsm_left
        if below > 0:
            mem.mat(W2, ld, tk, tk)[:, :] = mem.mat(W, ld, tk, tk).T.copy()
            n = below * te
            mem.a[Wp:Wp + n] = mem.a[tile(k + 1, k):tile(k + 1, k) + n]
            gemm_single(mem, ld, tk, tk, -1.0, Wp, ld, False, W2, ld, False, 0.0, tile(k + 1, k), ld, 0, below, te, 0, te, rows(nt - 1))
        if k > 0:
            transpose_tiles(tile(k, 0), cs, Wp, te, tk, nb, k)
            if below > 0:
                trailing_update(nt, tile(0, k), te, Wp, te, 0, k, 1.0, tk, rowFirstFixed=k + 1)
            gemm_single(mem, tk, nb, tk, 1.0, W, ld, False, Wp, ld, False, 0.0, tile(k, 0), ld, 0, k, 0, te, cs, tk)
        Wm = mem.mat(W, ld, tk, tk)
        mask = np.tril(np.ones((tk, tk), bool), -1 if diag_unit else 0)
        Akk[mask] = Wm[mask]
    got = np.zeros((N, N))
    for j in range(nt):
        for i in range(nt):
            r, c = rows(i), rows(j)
            got[i * nb:i * nb + r, j * nb:j * nb + c] = mem.mat(tile(i, j), nb, r, c)
    Lt = np.tril(L, -1) + np.eye(N) if diag_unit else np.tril(L)
    want = np.linalg.inv(Lt)
    sel = np.tril(np.ones((N, N), bool), -1 if diag_unit else 0)
    e1 = np.abs(got[sel] - want[sel]).max() / np.abs(want).max()
    assert np.array_equal(np.triu(got, 1), np.triu(Lin, 1))
    if diag_unit:
        return e1, 0.0
    # ---- lauum
    for k in range(nt):
        tk = rows(k)
        Akk = mem.mat(tile(k, k), ld, tk, tk)
        mem.mat(W2, ld, tk, tk)[:, :] = np.tril(Akk).T
        if k > 0:
            transpose_tiles(tile(k, 0), cs, Wp, te, tk, nb, k)
            trailing_update(k, Wp, te, Wp, te, 0, k, 1.0, tk)
            gemm_single(mem, tk, nb, tk, 1.0, W2, ld, False, Wp, ld, False, 0.0, tile(k, 0), ld, 0, k, 0, te, cs, tk)
        gemm_single(mem, tk, tk, tk, 1.0, W2, ld, False, W2, ld, False, 0.0, tile(k, k), ld, 1)
    got2 = np.zeros((N, N))
    for j in range(nt):
        for i in range(nt):
            r, c = rows(i), rows(j)
            got2[i * nb:i * nb + r, j * nb:j * nb + c] = mem.mat(tile(i, j), nb, r, c)
    want2 = np.linalg.inv(Afull)
    e2 = np.abs(np.tril(got2) - np.tril(want2)).max() / np.abs(want2).max()
    assert np.array_equal(np.triu(got2, 1), np.triu(Lin, 1))
    return e1, e2


if __name__ == "__main__":
    for N, nb in ((378, 93), (512, 64), (200, 256), (130, 64)):
        print(N, nb, run(N, nb), run(N, nb, True))
    print("schedule ok")
