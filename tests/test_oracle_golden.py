"""CPU: the oracle (oracle/oracle.c) against fixtures produced by the reference's own
compiled generators and tile kernels (tests/golden/make_golden.py)."""
import numpy as np
import pytest

from oracle.oracle import Lower, Upper, UpperLower, NoTrans, Trans, ConjTrans, NonUnit, Left, Right

SEED_A, SEED_B, SEED_B_POSV = 3872, 3873, 2354
RTOL = 1e-12  # BLAS summation order differs; values agree to ~1e-15


def test_rnd64_jump_matches_definition(oracle):
    a, c, m = 6364136223846793005, 1, (1 << 64) - 1
    x = 3872
    for n in range(0, 70):
        assert oracle.rnd64_jump(n, 3872) == x
        x = (a * x + c) & m
    # jump-ahead composes
    assert oracle.rnd64_jump(1000003, 17) == oracle.rnd64_jump(3, oracle.rnd64_jump(1000000, 17))


def test_appendix_c_known_answers(oracle, golden):
    # SURVEY.md Appendix C, produced by CORE_dplgsy / CORE_dplrnt / CORE_dpotrf
    A = oracle.plgsy(8.0, UpperLower, 8, 4, SEED_A)
    assert A[0, 0] == 8.5
    assert A[1, 0] == A[0, 1] == -0.34199792146682739
    assert A[5, 2] == A[2, 5] == 0.1133708655834198
    assert A[7, 7] == 7.5208688974380493
    assert np.array_equal(A, golden["d8_A"])
    B = oracle.plrnt(8, 1, 4, 1, SEED_B)
    assert B[0, 0] == 0.5 and B[7, 0] == 0.37386718392372131
    assert np.array_equal(B, golden["d8_B"])
    L = oracle.plgsy(8.0, Lower, 8, 4, SEED_A)
    assert oracle.potrf(Lower, L, 4) == 0
    np.testing.assert_allclose(L[0, 0], 2.9154759474226504, rtol=1e-15)
    np.testing.assert_allclose(L[1, 0], -0.11730431930647948, rtol=1e-15)
    np.testing.assert_allclose(L[7, 7], 2.7222068734683491, rtol=1e-15)
    np.testing.assert_allclose(L, golden["d8_L"], rtol=RTOL, atol=1e-15)
    U = oracle.plgsy(8.0, Upper, 8, 4, SEED_A)
    assert oracle.potrf(Upper, U, 4) == 0
    np.testing.assert_allclose(U, golden["d8_U"], rtol=RTOL, atol=1e-15)


@pytest.mark.parametrize("p", ["d", "z"])
@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
def test_generator_and_factor_378_93(oracle, golden, p, name, uplo):
    N, nb = 378, 93
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A, p)
    assert np.array_equal(A0.ravel(order="F")[::31], golden[f"{p}378_A0_{name}_s31"])  # bit-exact
    F = A0.copy(order="F")
    assert oracle.potrf(uplo, F, nb, p) == 0
    np.testing.assert_allclose(F.ravel(order="F")[::31], golden[f"{p}378_F_{name}_s31"], rtol=RTOL, atol=1e-14)
    np.testing.assert_allclose(np.diag(F), golden[f"{p}378_F_{name}_diag"], rtol=RTOL)
    bad, res, _, _ = oracle.check_potrf(uplo, F, A0, p)
    assert bad == 0 and res < 60.0
    # the other triangle is never touched: diagonal tiles keep the generator's mirror image,
    # tiles across the diagonal are not even stored by the sym layout
    tri = (lambda M: np.triu(M, 1)) if uplo == Lower else (lambda M: np.tril(M, -1))
    assert np.array_equal(tri(F), tri(A0))


@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
def test_potrs_378_367(oracle, golden, name, uplo):
    N, nb, K = 378, 93, 367
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    F = A0.copy(order="F")
    assert oracle.potrf(uplo, F, nb) == 0
    B = oracle.plrnt(N, K, nb, nb, SEED_B_POSV)
    assert np.array_equal(B.ravel(order="F")[::41], golden["d378_B_s41"])
    X = B.copy(order="F")
    assert oracle.potrs(uplo, F, X, nb, nb) == 0
    np.testing.assert_allclose(X.ravel(order="F")[::41], golden[f"d378_X_{name}_s41"], rtol=1e-11, atol=1e-15)
    bad, res = oracle.check_axmb(uplo, A0, B.copy(order="F"), X)
    assert bad == 0 and res < 60.0
    # posv == potrf + potrs
    A2, X2 = A0.copy(order="F"), B.copy(order="F")
    assert oracle.posv(uplo, A2, X2, nb, nb) == 0
    assert np.array_equal(X2, X) and np.array_equal(A2, F)


@pytest.mark.parametrize("p", ["d", "z"])
def test_generator_independent_of_tiling(oracle, golden, p):
    a = oracle.plgsy(200.0, UpperLower, 200, 64, SEED_A, p)
    b = oracle.plgsy(200.0, UpperLower, 200, 25, SEED_A, p)
    assert np.array_equal(a, b)
    assert np.array_equal(a, a.conj().T)  # symmetric / hermitian by construction
    if p == "d":
        assert np.array_equal(a[::3, ::5], golden["d200_A_full"])
    r1 = oracle.plrnt(200, 30, 64, 7, SEED_B, p)
    r2 = oracle.plrnt(200, 30, 200, 30, SEED_B, p)
    assert np.array_equal(r1, r2)
    assert np.array_equal(oracle.plrnt(378, 5, 93, 5, SEED_B, p).ravel(order="F")[::41],
                          golden[f"{p}378_B3873_s41"])


def test_info_semantics(oracle, golden):
    # global 1-based index of the first failing leading minor: k*nb + iinfo (zpotrf_L.jdf:180-182)
    A = oracle.plgsy(0.0, Lower, 200, 64, SEED_A)
    assert oracle.potrf(Lower, A, 64) == int(golden["d200_nonspd_info_L"][0])
    A = oracle.plgsy(0.0, Upper, 200, 64, SEED_A)
    assert oracle.potrf(Upper, A, 64) == int(golden["d200_nonspd_info_U"][0])
    A = oracle.plgsy(200.0, Lower, 200, 64, SEED_A)
    A[150, 150] = -1.0
    assert oracle.potrf(Lower, A, 64) == int(golden["d200_pivot150_info_L"][0]) == 151
    bad, res, _, _ = oracle.check_potrf(Lower, A, oracle.plgsy(200.0, Lower, 200, 64, SEED_A))
    assert bad == 1


def test_tile_kernels_against_numpy(oracle):
    rng = np.random.default_rng(0)
    for p, dt in (("d", np.float64), ("z", np.complex128)):
        def rnd(*s):
            a = rng.standard_normal(s)
            if p == "z":
                a = a + 1j * rng.standard_normal(s)
            return np.asfortranarray(a.astype(dt))
        A, B, C = rnd(13, 7), rnd(11, 7), rnd(13, 11)
        C0 = C.copy()
        oracle.gemm_tile(NoTrans, ConjTrans, -1.0, A, B, 1.0, C, p)
        np.testing.assert_allclose(C, C0 - A @ B.conj().T, rtol=1e-13, atol=1e-13)
        S = rnd(13, 13)
        S0 = S.copy()
        oracle.herk_tile(Lower, NoTrans, -1.0, A, 1.0, S, p)
        ref = S0 - A @ A.conj().T
        np.testing.assert_allclose(np.tril(S, -1), np.tril(ref, -1), rtol=1e-13, atol=1e-13)
        np.testing.assert_allclose(np.diag(S), np.diag(ref).real, rtol=1e-13, atol=1e-13)
        assert np.array_equal(np.triu(S, 1), np.triu(S0, 1))
        T = np.asfortranarray(np.tril(rnd(7, 7)) + 7 * np.eye(7))
        X = A.copy(order="F")
        oracle.trsm_tile(Right, Lower, ConjTrans, NonUnit, 1.0, T, X, p)
        np.testing.assert_allclose(X @ T.conj().T, A, rtol=1e-12, atol=1e-12)
        Tu = np.asfortranarray(np.triu(rnd(13, 13)) + 13 * np.eye(13))
        Y = C0.copy(order="F")
        oracle.trsm_tile(Left, Upper, ConjTrans, NonUnit, 1.0, Tu, Y, p)
        np.testing.assert_allclose(Tu.conj().T @ Y, C0, rtol=1e-12, atol=1e-12)
        Y = C0.copy(order="F")
        oracle.trsm_tile(Left, Lower, NoTrans, NonUnit, 2.0, np.asfortranarray(Tu.conj().T), Y, p)
        np.testing.assert_allclose(Tu.conj().T @ Y, 2.0 * C0, rtol=1e-12, atol=1e-12)


def test_reference_library_agrees_when_present(oracle):
    """In the build container the reference's compiled kernels are available: compare live."""
    try:
        from oracle.oracle import Reference
        R = Reference()
    except (FileNotFoundError, OSError):
        pytest.skip("oracle/_ref not built on this machine")
    for p in "dz":
        a = oracle.plgsy(131.0, UpperLower, 131, 32, 99, p)
        assert np.array_equal(a, R.plgsy(131.0, UpperLower, 131, 32, 99, p))
        F1 = oracle.plgsy(131.0, Lower, 131, 32, 99, p)
        F2 = F1.copy(order="F")
        assert oracle.potrf(Lower, F1, 32, p) == R.potrf(Lower, F2, 32, p) == 0
        np.testing.assert_allclose(F1, F2, rtol=RTOL, atol=1e-15)


@pytest.mark.parametrize("p", ["d", "z"])
def test_tiled_trsm_all_eight_variants_against_numpy(oracle, p):
    """dplasma_ztrsm restated for Left and Right, both triangles, N / T(C): op(A) X = alpha B resp. X op(A) = alpha B"""
    rng = np.random.default_rng(3)
    M, R, nb = 150, 70, 32
    for side in (Left, Right):
        na = M if side == Left else R
        G = rng.standard_normal((na, na)) + (1j * rng.standard_normal((na, na)) if p == "z" else 0)
        for uplo in (Lower, Upper):
            Tm = np.asfortranarray((np.tril(G) if uplo == Lower else np.triu(G)) + na * np.eye(na)).astype(G.dtype)
            for trans in (NoTrans, ConjTrans):
                B = np.asfortranarray(rng.standard_normal((M, R)) + (1j * rng.standard_normal((M, R)) if p == "z" else 0)).astype(G.dtype)
                X = B.copy(order="F")
                mb, nbr = (nb, 24) if side == Left else (24, nb)
                assert oracle.trsm(side, uplo, trans, NonUnit, 0.5, Tm, X, mb, nbr, p) == 0
                op = Tm if trans == NoTrans else Tm.conj().T
                got = op @ X if side == Left else X @ op
                np.testing.assert_allclose(got, 0.5 * B, rtol=1e-11, atol=1e-11)
