"""CPU: the oracle's restatement of dplasma_zgemm / dplasma_zherk (d: dsyrk) as public tiled operations
 against numpy and, where it was built, against the reference's own CORE_dgemm / CORE_dsyrk
kernels driven in the same per-tile k order (oracle/ref_driver.c)."""
import numpy as np
import pytest

from oracle.oracle import Lower, Upper, NoTrans, Trans, ConjTrans


def op(X, t):
    return X if t == NoTrans else (X.T if t == Trans else X.conj().T)


@pytest.mark.parametrize("p", ["d", "z"])
def test_tiled_gemm_against_numpy(oracle, p):
    rng = np.random.default_rng(2)
    M, N, K, mb, nb, kb = 150, 130, 170, 32, 48, 40
    cplx = (lambda s: 1j * rng.standard_normal(s)) if p == "z" else (lambda s: 0)
    for ta in (NoTrans, Trans, ConjTrans):
        for tb in (NoTrans, Trans, ConjTrans):
            A = np.asfortranarray(rng.standard_normal((M, K) if ta == NoTrans else (K, M)) + cplx((M, K) if ta == NoTrans else (K, M)))
            B = np.asfortranarray(rng.standard_normal((K, N) if tb == NoTrans else (N, K)) + cplx((K, N) if tb == NoTrans else (N, K)))
            C0 = np.asfortranarray(rng.standard_normal((M, N)) + cplx((M, N)))
            C1 = C0.copy(order="F")
            assert oracle.gemm(ta, tb, 0.7, A, B, -0.3, C1, mb, nb, kb, p) == 0
            np.testing.assert_allclose(C1, 0.7 * op(A, ta) @ op(B, tb) - 0.3 * C0, rtol=1e-12, atol=1e-12)
    # argument codes (src/zgemm_wrapper.c:657-664) and the quick return (:716-719)
    assert oracle.gemm(7, NoTrans, 1.0, A, B, 0.0, C1, mb, nb, kb, p) == -1
    assert oracle.gemm(NoTrans, 7, 1.0, A, B, 0.0, C1, mb, nb, kb, p) == -2


@pytest.mark.parametrize("p", ["d", "z"])
def test_tiled_herk_against_numpy(oracle, p):
    rng = np.random.default_rng(4)
    N, K, nb, kb = 130, 170, 48, 40
    cplx = (lambda s: 1j * rng.standard_normal(s)) if p == "z" else (lambda s: 0)
    for uplo in (Lower, Upper):
        tri = np.tril if uplo == Lower else np.triu
        other = (lambda X: np.triu(X, 1)) if uplo == Lower else (lambda X: np.tril(X, -1))
        for tr in (NoTrans, ConjTrans):
            A = np.asfortranarray(rng.standard_normal((N, K) if tr == NoTrans else (K, N)) + cplx((N, K) if tr == NoTrans else (K, N)))
            G = rng.standard_normal((N, N)) + cplx((N, N))
            C0 = np.asfortranarray(G + G.conj().T)
            C1 = C0.copy(order="F")
            assert oracle.herk(uplo, tr, 0.7, A, -0.3, C1, nb, kb, p) == 0
            want = 0.7 * (A @ A.conj().T if tr == NoTrans else A.conj().T @ A) - 0.3 * C0
            np.testing.assert_allclose(tri(C1), tri(want), rtol=1e-12, atol=1e-12)
            assert np.array_equal(other(C1), other(C0))
    assert oracle.herk(7, NoTrans, 1.0, A, 0.0, C1, nb, kb, p) == -1
    assert oracle.herk(Lower, 7, 1.0, A, 0.0, C1, nb, kb, p) == -2


def test_gemm_syrk_against_reference_fixtures(oracle):
    """fixtures from the reference's CORE_dgemm / CORE_dsyrk (tests/golden/make_golden_blas3.py)"""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_r01_blas3.npz"))
    M, N, K, mb, nb, kb = 378, 300, 367, 93, 64, 80
    for ta, na in ((NoTrans, "N"), (Trans, "T")):
        for tb, nbn in ((NoTrans, "N"), (Trans, "T")):
            A = oracle.plrnt(*((M, K, mb, kb) if ta == NoTrans else (K, M, kb, mb)), 11)
            B = oracle.plrnt(*((K, N, kb, nb) if tb == NoTrans else (N, K, nb, kb)), 12)
            C = oracle.plrnt(M, N, mb, nb, 13)
            assert oracle.gemm(ta, tb, 0.75, A, B, -0.5, C, mb, nb, kb) == 0
            np.testing.assert_allclose(C.ravel(order="F")[::17], g[f"dgemm_{na}{nbn}_s17"], rtol=1e-12, atol=1e-13)
    for uplo, nu in ((Lower, "L"), (Upper, "U")):
        for tr, nt_ in ((NoTrans, "N"), (Trans, "T")):
            A = oracle.plrnt(*((N, K, nb, kb) if tr == NoTrans else (K, N, kb, nb)), 21)
            C = oracle.plrnt(N, N, nb, nb, 22)
            assert oracle.herk(uplo, tr, 0.5, A, -0.25, C, nb, kb) == 0
            np.testing.assert_allclose(C.ravel(order="F")[::17], g[f"dsyrk_{nu}{nt_}_s17"], rtol=1e-12, atol=1e-13)


def test_reference_gemm_syrk_agree_when_present(oracle):
    try:
        from oracle.oracle import Reference
        R = Reference()
    except (FileNotFoundError, OSError):
        pytest.skip("oracle/_ref not built on this machine")
    if not hasattr(R.lib, "ref_dgemm"):
        pytest.skip("oracle/_ref predates the level-3 drivers")
    M, N, K, mb, nb, kb = 150, 130, 170, 32, 48, 40
    for ta in (NoTrans, Trans):
        for tb in (NoTrans, Trans):
            A = oracle.plrnt(*((M, K, mb, kb) if ta == NoTrans else (K, M, kb, mb)), 11)
            B = oracle.plrnt(*((K, N, kb, nb) if tb == NoTrans else (N, K, nb, kb)), 12)
            C1 = oracle.plrnt(M, N, mb, nb, 13)
            C2 = C1.copy(order="F")
            oracle.gemm(ta, tb, 0.7, A, B, -0.3, C1, mb, nb, kb)
            R.gemm(ta, tb, 0.7, A, B, -0.3, C2, mb, nb, kb)
            np.testing.assert_allclose(C1, C2, rtol=1e-12, atol=1e-13)
    for uplo in (Lower, Upper):
        for tr in (NoTrans, Trans):
            A = oracle.plrnt(*((N, K, nb, kb) if tr == NoTrans else (K, N, kb, nb)), 21)
            C1 = oracle.plrnt(N, N, nb, nb, 22)
            C2 = C1.copy(order="F")
            oracle.herk(uplo, tr, 0.7, A, -0.3, C1, nb, kb)
            R.syrk(uplo, tr, 0.7, A, -0.3, C2, nb, kb)
            np.testing.assert_allclose(C1, C2, rtol=1e-12, atol=1e-13)
