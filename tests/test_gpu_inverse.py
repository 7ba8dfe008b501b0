"""GPU parity tests of the inversion path: dplasma_dtrtri / dlauum / dpotri /
dpoinv / check_dpoinv and the pdtrsm_ wrapper, called through the C ABI, against the CPU oracle
and fixtures from the reference's own tile kernels.  Tolerance: 1e-11 relative to max|.| for
values (fp64, different summation order); the reference's acceptance test is result <= 10
(src/dplasma_zcheck.c:303-309)."""
import ctypes as C
import os

import numpy as np
import pytest

import dplasma_b200 as dp
from oracle.oracle import NoTrans, Trans, NonUnit, Unit, Left, Right
from tests.util import full_to_local, local_to_full

pytestmark = pytest.mark.gpu
Lower, Upper, UpperLower = dp.Lower, dp.Upper, dp.UpperLower
SEED_A = 3872
VTOL = 1e-11
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ctx():
    c = dp.Context()
    yield c
    c.close()


@pytest.fixture(scope="module")
def ginv():
    return np.load(os.path.join(ROOT, "tests", "golden", "golden_r01_inverse.npz"))


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def make_host(ctx, A, mb, nb, kind="tile", uplo=UpperLower):
    M, N = A.shape
    d = dp.BlockCyclic(ctx.lib, M, N, mb, nb, kind=kind, uplo=uplo, where="host")
    d.host_array()[:d.elems] = full_to_local(A, mb, nb, sym_uplo=(uplo if kind == "sym" else None))
    return d


def read_host(d, M, N, mb, nb, kind="tile", uplo=UpperLower, out=None):
    return local_to_full(d.host_array()[:d.elems], M, N, mb, nb, sym_uplo=(uplo if kind == "sym" else None), out=out)


def tri_of(uplo):
    return np.tril if uplo == Lower else np.triu


def other_of(uplo):
    return (lambda M: np.triu(M, 1)) if uplo == Lower else (lambda M: np.tril(M, -1))


def factor(oracle, uplo, N, nb):
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    F = A0.copy(order="F")
    assert oracle.potrf(uplo, F, nb) == 0
    return A0, F


@pytest.mark.parametrize("N,nb", [(378, 93), (1024, 128), (640, 64), (200, 256), (1700, 320)])
@pytest.mark.parametrize("uplo", [Lower, Upper])
@pytest.mark.parametrize("diag", [NonUnit, Unit])
def test_dtrtri_vs_oracle(ctx, oracle, N, nb, uplo, diag):
    _, F = factor(oracle, uplo, N, nb)
    want = F.copy(order="F")
    assert oracle.trtri(uplo, diag, want, nb) == 0
    d = make_host(ctx, F, nb, nb)
    assert ctx.lib.dplasma_dtrtri(ctx.ptr, uplo, diag, d.ref) == 0
    got = read_host(d, N, N, nb, nb)
    assert relerr(tri_of(uplo)(got), tri_of(uplo)(want)) < VTOL
    assert np.array_equal(other_of(uplo)(got), other_of(uplo)(F))  # other triangle untouched
    if diag == Unit:
        assert np.array_equal(np.diag(got), np.diag(F))            # a Unit diagonal is not referenced
    d.free()


@pytest.mark.parametrize("N,nb", [(378, 93), (1024, 128), (640, 64), (200, 256)])
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_dlauum_vs_oracle(ctx, oracle, N, nb, uplo):
    _, F = factor(oracle, uplo, N, nb)
    want = F.copy(order="F")
    oracle.lauum(uplo, want, nb)
    d = make_host(ctx, F, nb, nb)
    assert ctx.lib.dplasma_dlauum(ctx.ptr, uplo, d.ref) == 0
    got = read_host(d, N, N, nb, nb)
    assert relerr(tri_of(uplo)(got), tri_of(uplo)(want)) < VTOL
    assert np.array_equal(other_of(uplo)(got), other_of(uplo)(F))
    d.free()


@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
@pytest.mark.parametrize("kind", ["tile", "sym", "lapack"])
def test_dpotri_378_93_reference_ctest_case(ctx, oracle, ginv, name, uplo, kind):
    """tests/testing_zposv.c:217-254 (ZPOTRF + ZPOTRI) on tests/Testings.cmake:146 -N 378 -t 93"""
    N, nb = 378, 93
    A0, F = factor(oracle, uplo, N, nb)
    if kind == "lapack":
        d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, kind="lapack", where="host")
        d.host_array()[:N * N] = F.ravel(order="F")
    else:
        d = make_host(ctx, F, nb, nb, kind, uplo)
    assert ctx.lib.dplasma_dpotri(ctx.ptr, uplo, d.ref) == 0
    if kind == "lapack":
        P = d.host_array()[:N * N].reshape((N, N), order="F").copy(order="F")
    else:
        P = read_host(d, N, N, nb, nb, kind, uplo, out=F.copy(order="F") if kind == "sym" else None)
    want = F.copy(order="F")
    assert oracle.potri(uplo, want, nb) == 0
    assert relerr(tri_of(uplo)(P), tri_of(uplo)(want)) < VTOL
    assert relerr(P.ravel(order="F")[::31], ginv[f"d378_potri_{name}_s31"]) < VTOL
    assert relerr(np.diag(P), ginv[f"d378_potri_{name}_diag"]) < VTOL
    assert np.array_equal(other_of(uplo)(P), other_of(uplo)(F))
    # the reference's acceptance test, by the oracle on the GPU result and by the GPU check itself
    Afull = oracle.plgsy(float(N), UpperLower, N, nb, SEED_A)
    bad, res = oracle.check_poinv(uplo, Afull, P)
    assert bad == 0 and res < 10.0
    dA = make_host(ctx, Afull, nb, nb)
    assert ctx.lib.check_dpoinv(ctx.ptr, 0, uplo, dA.ref, d.ref) == 0
    got = ctx.lib.dplasma_b200_last_check_result(ctx.ptr)
    assert got < 10.0 and 0.2 * res < got < 5.0 * res  # a rounding-level residual: same size, not same digits
    # ... and the GPU check rejects a wrong inverse
    if kind == "tile":
        Pbad = P.copy(order="F")
        Pbad[N // 2, N // 2] *= 1.001
        dbad = make_host(ctx, Pbad, nb, nb)
        assert ctx.lib.check_dpoinv(ctx.ptr, 0, uplo, dA.ref, dbad.ref) == 1
        dbad.free()
    dA.free()
    d.free()


@pytest.mark.parametrize("N,nb", [(378, 93), (1536, 256), (2048, 512), (65, 64), (7, 64)])
@pytest.mark.parametrize("uplo", [Lower, Upper])
@pytest.mark.parametrize("sync", [0, 1])
def test_dpoinv_vs_oracle_and_numpy(ctx, oracle, N, nb, uplo, sync):
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    want = A0.copy(order="F")
    assert oracle.poinv(uplo, want, nb) == 0
    d = make_host(ctx, A0, nb, nb)
    fn = ctx.lib.dplasma_dpoinv_sync if sync else ctx.lib.dplasma_dpoinv
    assert fn(ctx.ptr, uplo, d.ref) == 0
    got = read_host(d, N, N, nb, nb)
    assert relerr(tri_of(uplo)(got), tri_of(uplo)(want)) < VTOL
    assert np.array_equal(other_of(uplo)(got), other_of(uplo)(A0))
    Afull = oracle.plgsy(float(N), UpperLower, N, nb, SEED_A)
    assert relerr(tri_of(uplo)(got), tri_of(uplo)(np.linalg.inv(Afull))) < 1e-9
    d.free()


def test_inverse_info_and_argument_codes(ctx, oracle, ginv):
    N, nb = 378, 93
    for name, uplo in (("L", Lower), ("U", Upper)):
        _, F = factor(oracle, uplo, N, nb)
        S = F.copy(order="F")
        S[200, 200] = 0.0
        d = make_host(ctx, S, nb, nb)
        # exactly singular triangle: INFO = k*mb + info (src/ztrtri_L.jdf:180-186)
        assert ctx.lib.dplasma_dtrtri(ctx.ptr, uplo, NonUnit, d.ref) == int(ginv[f"d378_trtri_singular_info_{name}"][0])
        d.free()
        # not positive definite: dpoinv reports the factorization's info and inverts nothing
        A = oracle.plgsy(0.0, uplo, 200, 64, SEED_A)
        want = A.copy(order="F")
        winfo = oracle.potrf(uplo, want, 64)
        assert winfo > 0
        d = make_host(ctx, A, 64, 64)
        assert ctx.lib.dplasma_dpoinv(ctx.ptr, uplo, d.ref) == winfo
        d.free()
    s = dp.capi.BlockCyclicStruct()
    ctx.lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, 0, 4, 4, 8, 12, 0, 0, 8, 12, 1, 1, 1, 1, 0, 0)
    sq = dp.capi.BlockCyclicStruct()
    ctx.lib.parsec_matrix_block_cyclic_init(C.byref(sq), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, 0, 4, 4, 8, 8, 0, 0, 8, 8, 1, 1, 1, 1, 0, 0)
    # src/ztrtri_wrapper.c:203-216, zlauum_wrapper.c:192-200, zpotri_wrapper.c:62-65
    assert ctx.lib.dplasma_dtrtri(ctx.ptr, 7, NonUnit, C.byref(sq)) == -1
    assert ctx.lib.dplasma_dtrtri(ctx.ptr, Lower, 7, C.byref(sq)) == -2
    assert ctx.lib.dplasma_dtrtri(ctx.ptr, Lower, NonUnit, C.byref(s)) == -6
    assert ctx.lib.dplasma_dlauum(ctx.ptr, 7, C.byref(sq)) == -1
    assert ctx.lib.dplasma_dlauum(ctx.ptr, Upper, C.byref(s)) == -6
    assert ctx.lib.dplasma_dpotri(ctx.ptr, 7, C.byref(sq)) == -1
    assert ctx.lib.dplasma_dpoinv(ctx.ptr, 7, C.byref(sq)) == -1
    assert ctx.lib.dplasma_dpoinv_sync(ctx.ptr, 7, C.byref(sq)) == -1


@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_dpoinv_device_resident_8192_by_the_reference_check(ctx, uplo):
    """larger than the oracle finishes in seconds: generate, invert and check on the GPU only;
    A0 (full) x Ainv must pass check_dpoinv, and the round trip inv(inv(A)) must return A."""
    N, nb = 8192, 512
    lib = ctx.lib
    dA0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
    dA = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), UpperLower, dA0.ref, SEED_A) == 0
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), UpperLower, dA.ref, SEED_A) == 0
    assert lib.dplasma_dpoinv(ctx.ptr, uplo, dA.ref) == 0
    t = lib.dplasma_b200_last_elapsed(ctx.ptr)
    # dpotrf N^3/3 + dpotri 2N^3/3 (src/flops.h: FLOPS_DPOTRF + FLOPS_DPOTRI)
    print(f"\ndpoinv N={N} nb={nb} uplo={uplo}: {t * 1e3:.1f} ms, {N ** 3 / t / 1e12:.2f} TFLOP/s")
    assert lib.check_dpoinv(ctx.ptr, 0, uplo, dA0.ref, dA.ref) == 0
    assert lib.dplasma_b200_last_check_result(ctx.ptr) < 10.0
    # invert the inverse: back to A on the stored triangle (sampled through the symmetric norm of A0 - A)
    assert lib.dplasma_dpoinv(ctx.ptr, uplo, dA.ref) == 0
    assert lib.check_dpoinv(ctx.ptr, 0, uplo, dA0.ref, dA.ref) == 1  # A is not the inverse of A ...
    n1 = lib.dplasma_dlansy(ctx.ptr, dp.InfNorm, uplo, dA.ref)
    n0 = lib.dplasma_dlansy(ctx.ptr, dp.InfNorm, uplo, dA0.ref)
    assert n1 == pytest.approx(n0, rel=1e-10)                          # ... but it is A again
    dA0.free(); dA.free()


@pytest.mark.parametrize("side,uplo,trans,diag", [(b"L", b"L", b"N", b"N"), (b"L", b"U", b"T", b"N"), (b"R", b"L", b"T", b"N"),
                                                  (b"R", b"U", b"N", b"U"), (b"l", b"l", b"c", b"n")])
def test_pdtrsm_single_process(ctx, oracle, side, uplo, trans, diag):
    """src/scalapack_wrappers/dplasma_wrapper_pdtrsm.c:203-327 on a 1x1 grid, LAPACK-storage operands"""
    M, N, nb = 300, 170, 64
    e = {b"L": Left, b"R": Right, b"U": Upper, b"N": NoTrans, b"T": Trans, b"C": Trans}
    sd = e[side.upper()]
    ul = Lower if uplo.upper() == b"L" else Upper
    tr = e[trans.upper()]
    dg = Unit if diag.upper() == b"U" else NonUnit
    na = M if sd == Left else N
    T = oracle.plgsy(float(na), ul, na, nb, SEED_A)
    oracle.potrf(ul, T, nb)
    B = oracle.plrnt(M, N, nb, nb, 7)
    want = B.copy(order="F")
    assert oracle.trsm(sd, ul, tr, dg, 1.25, T, want, nb, nb) == 0
    ctx.lib.dplasma_b200_blacs_gridinit(0, 1, 1, 0, 0)
    ctx.lib.dplasma_b200_wrapper_set_context(ctx.ptr)
    descA = (C.c_int * 9)(1, 0, na, na, nb, nb, 0, 0, na)
    descB = (C.c_int * 9)(1, 0, M, N, nb, nb, 0, 0, M)
    m, n, one, alpha = C.c_int(M), C.c_int(N), C.c_int(1), C.c_double(1.25)
    Ta, Ba = np.asfortranarray(T.copy()), np.asfortranarray(B.copy())
    ctx.lib.pdtrsm_(side, uplo, trans, diag, C.byref(m), C.byref(n), C.byref(alpha), Ta.ctypes.data_as(C.c_void_p),
                    C.byref(one), C.byref(one), descA, Ba.ctypes.data_as(C.c_void_p), C.byref(one), C.byref(one), descB)
    assert relerr(Ba, want) < 1e-10
    assert np.array_equal(Ta, T)
    ctx.lib.dplasma_b200_wrapper_set_context(None)


def test_taskpool_composition_runs_in_queue_order(ctx, oracle):
    """src/zpotri_wrapper.c:68-80 (PARSEC_COMPOSITION): several _New taskpools, one start / wait.  Here
    potrf + trtri + lauum on one matrix, then trsm, gemm and syrk that consume each other's results."""
    N, nb, K = 378, 93, 200
    lib = ctx.lib
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A)
    want = A0.copy(order="F")
    assert oracle.poinv(Lower, want, nb) == 0
    d = make_host(ctx, A0, nb, nb)
    i1, i2 = C.c_int(-5), C.c_int(-5)
    tps = [lib.dplasma_dpotrf_New(Lower, d.ref, C.byref(i1)), lib.dplasma_dtrtri_New(Lower, NonUnit, d.ref, C.byref(i2)),
           lib.dplasma_dlauum_New(Lower, d.ref)]
    assert all(tps) and i1.value == 0 and i2.value == 0
    for tp in tps:
        assert lib.parsec_context_add_taskpool(ctx.ptr, tp) == 0
    assert lib.parsec_context_start(ctx.ptr) == 0
    assert lib.parsec_context_wait(ctx.ptr) == 0
    lib.dplasma_dpotrf_Destruct(tps[0]); lib.dplasma_dtrtri_Destruct(tps[1]); lib.dplasma_dlauum_Destruct(tps[2])
    assert i1.value == 0 and i2.value == 0
    assert lib.dplasma_b200_last_elapsed(ctx.ptr) > 0.0
    got = read_host(d, N, N, nb, nb)
    assert relerr(np.tril(got), np.tril(want)) < VTOL
    d.free()
    # one taskpool that does all three
    d = make_host(ctx, A0, nb, nb)
    tp = lib.dplasma_dpoinv_New(Lower, d.ref, C.byref(i1))
    lib.parsec_context_add_taskpool(ctx.ptr, tp)
    assert lib.parsec_context_start(ctx.ptr) == 0 and lib.parsec_context_wait(ctx.ptr) == 0
    lib.dplasma_dpoinv_Destruct(tp)
    assert i1.value == 0 and relerr(np.tril(read_host(d, N, N, nb, nb)), np.tril(want)) < VTOL
    d.free()
    # B <- L^-1 B ; G <- B^T B (gemm) ; S <- B^T B lower (syrk), in one queue
    _, F = factor(oracle, Lower, N, nb)
    B = oracle.plrnt(N, K, nb, nb, 7)
    wB = B.copy(order="F")
    assert oracle.trsm(Left, Lower, NoTrans, NonUnit, 1.0, F, wB, nb, nb) == 0
    wG = np.zeros((K, K), order="F")
    assert oracle.gemm(Trans, NoTrans, 1.0, wB, wB, 0.0, wG, nb, nb, nb) == 0
    dF, dB = make_host(ctx, F, nb, nb), make_host(ctx, B, nb, nb)
    dG, dS = make_host(ctx, np.full((K, K), np.nan, order="F"), nb, nb), make_host(ctx, np.zeros((K, K), order="F"), nb, nb)
    tps = [lib.dplasma_dtrsm_New(Left, Lower, NoTrans, NonUnit, 1.0, dF.ref, dB.ref),
           lib.dplasma_dgemm_New(Trans, NoTrans, 1.0, dB.ref, dB.ref, 0.0, dG.ref),
           lib.dplasma_dsyrk_New(Lower, Trans, 1.0, dB.ref, 0.0, dS.ref)]
    assert all(tps)
    for tp in tps:
        lib.parsec_context_add_taskpool(ctx.ptr, tp)
    assert lib.parsec_context_start(ctx.ptr) == 0 and lib.parsec_context_wait(ctx.ptr) == 0
    lib.dplasma_dtrsm_Destruct(tps[0]); lib.dplasma_dgemm_Destruct(tps[1]); lib.dplasma_dsyrk_Destruct(tps[2])
    assert relerr(read_host(dB, N, K, nb, nb), wB) < 1e-10
    assert relerr(read_host(dG, K, K, nb, nb), wG) < 1e-10
    assert relerr(np.tril(read_host(dS, K, K, nb, nb)), np.tril(wG)) < 1e-10
    for x in (dF, dB, dG, dS):
        x.free()
    # illegal arguments: NULL (src/ztrtri_wrapper.c:84-92, zgemm_wrapper.c:445-452)
    assert lib.dplasma_dtrtri_New(7, NonUnit, dF.ref, C.byref(i1)) is None
    assert lib.dplasma_dgemm_New(7, NoTrans, 1.0, dF.ref, dF.ref, 0.0, dF.ref) is None


def test_c_driver_with_the_four_checks_of_testing_zposv():
    """tests/c/testing_dposv.c: plain C against include/dplasma_b200.h, the four checks of tests/testing_zposv.c
    (posv; potrf + potrs; potrf + trsm + trsm; potrf + potri) for both triangles; exit code 0 iff all pass"""
    import subprocess
    exe = os.path.join(ROOT, "tests", "c", "testing_dposv")
    subprocess.check_call(["gcc", "-std=c11", "-O2", "-I" + os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "c", "testing_dposv.c"), "-o", exe,
                           "-L" + os.path.join(ROOT, "dplasma_b200", "lib"), "-ldplasma_b200",
                           "-Wl,-rpath," + os.path.join(ROOT, "dplasma_b200", "lib")])
    for args in (["378", "93", "367"], ["1700", "320", "400"]):   # tests/Testings.cmake:89-90,146-149
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stdout + out.stderr
        assert out.stdout.count("PASSED !") == 8 and "FAILED" not in out.stdout, out.stdout
        for what in ("DPOSV", "DPOTRF + DPOTRS", "DPOTRF + DTRSM + DTRSM", "DPOTRF + DPOTRI"):
            assert f"TESTING {what} (Upper) ....... PASSED" in out.stdout and f"TESTING {what} (Lower) ....... PASSED" in out.stdout
