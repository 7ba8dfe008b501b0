"""Host-side layout helpers shared by the tests: 2D block-cyclic tile storage <-> full matrix.
Mirrors the layout documented in include/dplasma_b200.h (local tile (il,jl) at
(il + jl*nb_elem_r)*bsiz, column-major inside a tile, ld = mb; sym: stored triangle packed)."""
import numpy as np

Upper, Lower, UpperLower = 121, 122, 123


def local_tile_rows(mt, P, p):
    return list(range(p, mt, P))


def full_to_local(A, mb, nb, P=1, Q=1, rank=0, sym_uplo=None):
    """Extract rank's local tile-storage array (flat float64) from the full matrix A."""
    M, N = A.shape
    mt, nt = -(-M // mb), -(-N // nb)
    prow, pcol = rank // Q, rank % Q
    rows, cols = local_tile_rows(mt, P, prow), local_tile_rows(nt, Q, pcol)
    out = []
    for gn in cols:
        for gm in rows:
            if sym_uplo == Lower and gm < gn:
                continue
            if sym_uplo == Upper and gm > gn:
                continue
            t = np.zeros((mb, nb), dtype=A.dtype, order="F")
            blk = A[gm * mb:(gm + 1) * mb, gn * nb:(gn + 1) * nb]
            t[:blk.shape[0], :blk.shape[1]] = blk
            out.append(t.ravel(order="F"))
    return np.concatenate(out) if out else np.zeros(0, dtype=A.dtype)


def local_to_full(flat, M, N, mb, nb, P=1, Q=1, rank=0, sym_uplo=None, out=None):
    """Scatter rank's local tile-storage array into the full matrix `out` (dtype follows `flat`)."""
    mt, nt = -(-M // mb), -(-N // nb)
    prow, pcol = rank // Q, rank % Q
    rows, cols = local_tile_rows(mt, P, prow), local_tile_rows(nt, Q, pcol)
    if out is None:
        out = np.zeros((M, N), dtype=flat.dtype, order="F")
    k = 0
    for gn in cols:
        for gm in rows:
            if sym_uplo == Lower and gm < gn:
                continue
            if sym_uplo == Upper and gm > gn:
                continue
            t = flat[k * mb * nb:(k + 1) * mb * nb].reshape((mb, nb), order="F")
            k += 1
            r, c = min(mb, M - gm * mb), min(nb, N - gn * nb)
            out[gm * mb:gm * mb + r, gn * nb:gn * nb + c] = t[:r, :c]
    return out


def lapack_local(A, mb, nb, P, Q, rank):
    """ScaLAPACK local array (column-major, LLD = local rows) of rank in a row-major P x Q grid."""
    M, N = A.shape
    mt, nt = -(-M // mb), -(-N // nb)
    prow, pcol = rank // Q, rank % Q
    ri = np.concatenate([np.arange(g * mb, min(M, (g + 1) * mb)) for g in range(prow, mt, P)] or [np.zeros(0, int)])
    ci = np.concatenate([np.arange(g * nb, min(N, (g + 1) * nb)) for g in range(pcol, nt, Q)] or [np.zeros(0, int)])
    return np.asfortranarray(A[np.ix_(ri.astype(int), ci.astype(int))]), ri.astype(int), ci.astype(int)


def load_fake_blacs():
    """compile tests/c/fake_blacs.c and load it RTLD_GLOBAL: the process now 'has a BLACS' (Cblacs_gridinfo,
    Cblacs_get, numroc_) for libdplasma_b200 to find, like an application linked with ScaLAPACK"""
    import ctypes
    import os
    import subprocess
    import tempfile
    here = os.path.dirname(os.path.abspath(__file__))
    so = os.path.join(tempfile.gettempdir(), f"libfakeblacs_{os.getuid()}.so")
    if not os.path.exists(so):
        tmp = so + f".{os.getpid()}"
        subprocess.check_call(["gcc", "-O1", "-shared", "-fPIC", "-o", tmp, os.path.join(here, "c", "fake_blacs.c")])
        os.replace(tmp, so)
    return ctypes.CDLL(so, mode=ctypes.RTLD_GLOBAL)
