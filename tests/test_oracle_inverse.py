"""CPU: the oracle's restatement of the inversion path (dtrtri / dlauum / dpotri / dpoinv and
check_dpoinv) against fixtures produced by the reference's own compiled tile
kernels (tests/golden/make_golden_inverse.py) and against numpy."""
import os

import numpy as np
import pytest

from oracle.oracle import Lower, Upper, UpperLower, NoTrans, Trans, ConjTrans, NonUnit, Unit, Left, Right

SEED_A = 3872
RTOL = 1e-12
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ginv():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_r01_inverse.npz"))


def tri_of(uplo):
    return np.tril if uplo == Lower else np.triu


@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
def test_inverse_chain_378_93_against_reference_fixtures(oracle, ginv, name, uplo):
    """tests/testing_zposv.c:217-254 on the ctest case -N 378 -t 93 (ragged last tile)"""
    N, nb = 378, 93
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    F = A0.copy(order="F")
    assert oracle.potrf(uplo, F, nb) == 0
    T = F.copy(order="F")
    assert oracle.trtri(uplo, NonUnit, T, nb) == 0
    np.testing.assert_allclose(T.ravel(order="F")[::31], ginv[f"d378_trtri_{name}_s31"], rtol=RTOL, atol=1e-15)
    Tu = F.copy(order="F")
    assert oracle.trtri(uplo, Unit, Tu, nb) == 0
    np.testing.assert_allclose(Tu.ravel(order="F")[::31], ginv[f"d378_trtri_unit_{name}_s31"], rtol=RTOL, atol=1e-15)
    assert np.array_equal(np.diag(Tu), np.diag(F))  # a Unit diagonal is not referenced
    U = F.copy(order="F")
    assert oracle.lauum(uplo, U, nb) == 0
    np.testing.assert_allclose(U.ravel(order="F")[::31], ginv[f"d378_lauum_{name}_s31"], rtol=RTOL, atol=1e-15)
    P = F.copy(order="F")
    assert oracle.potri(uplo, P, nb) == 0
    np.testing.assert_allclose(P.ravel(order="F")[::31], ginv[f"d378_potri_{name}_s31"], rtol=RTOL, atol=1e-16)
    np.testing.assert_allclose(np.diag(P), ginv[f"d378_potri_{name}_diag"], rtol=RTOL)
    # the other triangle is never touched by any of them
    other = (lambda M: np.triu(M, 1)) if uplo == Lower else (lambda M: np.tril(M, -1))
    for X in (T, Tu, U, P):
        assert np.array_equal(other(X), other(A0))
    # one call == factor + invert
    V = A0.copy(order="F")
    assert oracle.poinv(uplo, V, nb) == 0
    assert np.array_equal(V, P)
    # the reference's acceptance test (src/dplasma_zcheck.c:265-315)
    Afull = oracle.plgsy(float(N), UpperLower, N, nb, SEED_A)
    bad, res = oracle.check_poinv(uplo, Afull, P)
    assert bad == 0 and res < 10.0
    # ... and it does reject a wrong inverse
    Pbad = P.copy(order="F")
    Pbad[N // 2, N // 2] *= 1.001
    assert oracle.check_poinv(uplo, Afull, Pbad)[0] == 1
    # exactly singular triangle: INFO = k*mb + info (src/ztrtri_L.jdf:180-186)
    S = F.copy(order="F")
    S[200, 200] = 0.0
    assert oracle.trtri(uplo, NonUnit, S, nb) == int(ginv[f"d378_trtri_singular_info_{name}"][0]) == 201


@pytest.mark.parametrize("p", ["d", "z"])
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_inverse_against_numpy(oracle, p, uplo):
    rng = np.random.default_rng(11)
    N, nb = 150, 32
    G = rng.standard_normal((N, N)) + (1j * rng.standard_normal((N, N)) if p == "z" else 0)
    A = np.asfortranarray(G @ G.conj().T + N * np.eye(N))
    tri = tri_of(uplo)
    F = np.asfortranarray(tri(A))
    assert oracle.potrf(uplo, F, nb, p) == 0
    T = F.copy(order="F")
    assert oracle.trtri(uplo, NonUnit, T, nb, p) == 0
    np.testing.assert_allclose(tri(T) @ tri(F), np.eye(N), atol=1e-12)
    U = F.copy(order="F")
    oracle.lauum(uplo, U, nb, p)
    want = tri(F).conj().T @ tri(F) if uplo == Lower else tri(F) @ tri(F).conj().T
    np.testing.assert_allclose(tri(U), tri(want), rtol=1e-12, atol=1e-12)
    P = F.copy(order="F")
    assert oracle.potri(uplo, P, nb, p) == 0
    np.testing.assert_allclose(tri(P), tri(np.linalg.inv(A)), rtol=1e-9, atol=1e-12)
    bad, res = oracle.check_poinv(uplo, A, P, p)
    assert bad == 0


@pytest.mark.parametrize("p", ["d", "z"])
def test_trmm_tile_against_numpy(oracle, p):
    rng = np.random.default_rng(5)
    m, n = 40, 23
    for side in (Left, Right):
        na = m if side == Left else n
        G = np.asfortranarray(rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if p == "z" else 0))
        for uplo in (Lower, Upper):
            for trans in (NoTrans, Trans, ConjTrans):
                for diag in (NonUnit, Unit):
                    Tm = tri_of(uplo)(G).copy()
                    if diag == Unit:
                        np.fill_diagonal(Tm, 1.0)
                    op = {NoTrans: Tm, Trans: Tm.T, ConjTrans: Tm.conj().T}[trans]
                    B = np.asfortranarray(rng.standard_normal((m, n)) + (1j * rng.standard_normal((m, n)) if p == "z" else 0))
                    X = B.copy(order="F")
                    oracle.trmm_tile(side, uplo, trans, diag, 0.75, G, X, p)  # G's other triangle must be ignored
                    want = 0.75 * (op @ B if side == Left else B @ op)
                    np.testing.assert_allclose(X, want, rtol=1e-12, atol=1e-12)


def test_reference_inverse_agrees_when_present(oracle):
    try:
        from oracle.oracle import Reference
        R = Reference()
    except (FileNotFoundError, OSError):
        pytest.skip("oracle/_ref not built on this machine")
    if not hasattr(R.lib, "ref_dtrtri"):
        pytest.skip("oracle/_ref predates the inversion drivers")
    for uplo in (Lower, Upper):
        A0 = oracle.plgsy(131.0, uplo, 131, 32, 99)
        a, b = A0.copy(order="F"), A0.copy(order="F")
        assert oracle.poinv(uplo, a, 32) == R.poinv(uplo, b, 32) == 0
        np.testing.assert_allclose(a, b, rtol=RTOL, atol=1e-17)
