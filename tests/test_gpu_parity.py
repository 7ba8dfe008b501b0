"""GPU parity tests: the CUDA path, called through the C ABI, against the CPU oracle and the
golden fixtures, on the same seeded inputs.  Tolerances:
  * generators: bit-exact;
  * factor / solve values: 1e-11 relative to max|.| (fp64, different summation order);
  * the reference's own acceptance test: residual <= 60 (src/dplasma_zcheck.c:119,215)."""
import ctypes as C

import numpy as np
import pytest

import dplasma_b200 as dp
from oracle.oracle import NoTrans, Trans, NonUnit, Unit, Left, Right
from tests.util import full_to_local, local_to_full, lapack_local

pytestmark = pytest.mark.gpu
Lower, Upper, UpperLower = dp.Lower, dp.Upper, dp.UpperLower
SEED_A, SEED_B, SEED_B_POSV = 3872, 3873, 2354
VTOL = 1e-11


@pytest.fixture(scope="module")
def ctx():
    c = dp.Context()
    yield c
    c.close()


def relerr(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-300)


def make_host(ctx, A, mb, nb, kind="tile", uplo=UpperLower):
    M, N = A.shape
    d = dp.BlockCyclic(ctx.lib, M, N, mb, nb, kind=kind, uplo=uplo, where="host")
    d.host_array()[:d.elems] = full_to_local(A, mb, nb, sym_uplo=(uplo if kind == "sym" else None))
    return d


def read_host(d, M, N, mb, nb, kind="tile", uplo=UpperLower, out=None):
    return local_to_full(d.host_array()[:d.elems], M, N, mb, nb, sym_uplo=(uplo if kind == "sym" else None), out=out)


# ------------------------------------------------------------------ generators (bit-exact)
@pytest.mark.parametrize("N,nb", [(8, 4), (378, 93), (1024, 128)])
@pytest.mark.parametrize("uplo", [Lower, Upper, UpperLower])
def test_dplgsy_bit_exact(ctx, oracle, golden, N, nb, uplo):
    d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, where="host")
    d.host_array()[:] = 0.0
    assert ctx.lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, d.ref, SEED_A) == 0
    got = read_host(d, N, N, nb, nb)
    want = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    assert np.array_equal(got, want)
    if N == 8 and uplo == UpperLower:
        assert np.array_equal(got, golden["d8_A"])
    if N == 378 and uplo == Lower:
        assert np.array_equal(got.ravel(order="F")[::31], golden["d378_A0_L_s31"])
    d.free()


def test_dplgsy_sym_descriptor_and_dplrnt(ctx, oracle, golden):
    N, nb = 378, 93
    d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, kind="sym", uplo=Lower, where="host")
    assert ctx.lib.dplasma_dplgsy(ctx.ptr, float(N), Lower, d.ref, SEED_A) == 0
    got = read_host(d, N, N, nb, nb, "sym", Lower)
    assert np.array_equal(got, oracle.plgsy(float(N), Lower, N, nb, SEED_A))
    d.free()
    b = dp.BlockCyclic(ctx.lib, N, 367, nb, nb, where="host")
    assert ctx.lib.dplasma_dplrnt(ctx.ptr, 0, b.ref, SEED_B_POSV) == 0
    got = read_host(b, N, 367, nb, nb)
    assert np.array_equal(got, oracle.plrnt(N, 367, nb, nb, SEED_B_POSV))
    assert np.array_equal(got.ravel(order="F")[::41], golden["d378_B_s41"])
    b.free()


# ------------------------------------------------------------------ kernel seam (B2)
def dev(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("m,n,k", [(128, 128, 128), (512, 512, 512), (93, 93, 93), (6, 93, 93), (200, 136, 64), (1024, 256, 512)])
def test_dgemm_tile_matches_oracle(ctx, oracle, m, n, k):
    import torch
    rng = np.random.default_rng(m * 7 + n)
    A = np.asfortranarray(rng.standard_normal((m, k)))
    B = np.asfortranarray(rng.standard_normal((n, k)))
    C0 = np.asfortranarray(rng.standard_normal((m, n)))
    want = C0.copy(order="F")
    oracle.gemm_tile(NoTrans, Trans, -1.0, A, B, 1.0, want)   # the K4 form: C -= A B^T
    tA, tB, tC = (torch.from_numpy(x.T.copy()).cuda() for x in (A, B, C0))  # row-major of X^T == col-major of X
    rc = ctx.lib.dplasma_b200_dgemm_tile(ctx.ptr, None, NoTrans, Trans, m, n, k, -1.0, dev(tA), m, dev(tB), n, 1.0, dev(tC), m)
    assert rc == 0
    torch.cuda.synchronize()
    got = tC.cpu().numpy().T
    assert relerr(got, want) < 1e-13
    # general transposes go through the element-wise producer: C = 0.5 A^T B + 2 C
    A2 = np.asfortranarray(rng.standard_normal((k, m)))
    B2 = np.asfortranarray(rng.standard_normal((k, n)))
    want = C0.copy(order="F")
    oracle.gemm_tile(Trans, NoTrans, 0.5, A2, B2, 2.0, want)
    tA, tB, tC = (torch.from_numpy(x.T.copy()).cuda() for x in (A2, B2, C0))
    assert ctx.lib.dplasma_b200_dgemm_tile(ctx.ptr, None, Trans, NoTrans, m, n, k, 0.5, dev(tA), k, dev(tB), k, 2.0, dev(tC), m) == 0
    torch.cuda.synchronize()
    assert relerr(tC.cpu().numpy().T, want) < 1e-13


@pytest.mark.parametrize("n,k", [(512, 512), (93, 93), (256, 64)])
def test_dsyrk_tile_lower_only(ctx, oracle, n, k):
    import torch
    rng = np.random.default_rng(n)
    A = np.asfortranarray(rng.standard_normal((n, k)))
    C0 = np.asfortranarray(rng.standard_normal((n, n)))
    want = C0.copy(order="F")
    oracle.herk_tile(Lower, NoTrans, -1.0, A, 1.0, want)      # K3
    tA, tC = (torch.from_numpy(x.T.copy()).cuda() for x in (A, C0))
    assert ctx.lib.dplasma_b200_dsyrk_tile(ctx.ptr, None, Lower, NoTrans, n, k, -1.0, dev(tA), n, 1.0, dev(tC), n) == 0
    torch.cuda.synchronize()
    got = tC.cpu().numpy().T
    assert relerr(np.tril(got), np.tril(want)) < 1e-13
    assert np.array_equal(np.triu(got, 1), np.triu(C0, 1))   # strict upper part untouched


@pytest.mark.parametrize("n", [64, 93, 512, 1024])
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_dpotrf_tile(ctx, oracle, n, uplo):
    import torch
    A0 = oracle.plgsy(float(n), UpperLower, n, n, 11)
    want = A0.copy(order="F")
    assert oracle.potrf_tile(uplo, want) == 0
    tA = torch.from_numpy(A0.T.copy()).cuda()
    info = torch.zeros(1, dtype=torch.int32, device="cuda")
    assert ctx.lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, uplo, n, dev(tA), n, dev(info), 0) == 0
    torch.cuda.synchronize()
    got = tA.cpu().numpy().T
    assert int(info.item()) == 0
    tri = np.tril if uplo == Lower else np.triu
    assert relerr(tri(got), tri(want)) < VTOL
    other = (lambda M: np.triu(M, 1)) if uplo == Lower else (lambda M: np.tril(M, -1))
    assert np.array_equal(other(got), other(A0))              # the other triangle is untouched
    # not positive definite -> info = offset + index of first bad pivot (src/zpotrf_L.jdf:180-182)
    Abad = A0.copy(order="F")
    j = n // 2
    Abad[j, j] = -1.0
    ref = Abad.copy(order="F")
    winfo = oracle.potrf_tile(uplo, ref)
    tA = torch.from_numpy(Abad.T.copy()).cuda()
    info.zero_()
    assert ctx.lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, uplo, n, dev(tA), n, dev(info), 1000) == 0
    torch.cuda.synchronize()
    assert int(info.item()) == 1000 + winfo == 1000 + j + 1


@pytest.mark.parametrize("m,n", [(512, 512), (93, 93), (6, 93), (1000, 256)])
def test_dtrsm_tile_right_lower_trans(ctx, oracle, m, n):
    import torch
    rng = np.random.default_rng(m + n)
    T = oracle.plgsy(float(n), UpperLower, n, n, 5)
    assert oracle.potrf_tile(Lower, T) == 0
    B0 = np.asfortranarray(rng.standard_normal((m, n)))
    want = B0.copy(order="F")
    oracle.trsm_tile(Right, Lower, Trans, NonUnit, 1.0, T, want)   # K2
    tT, tB = (torch.from_numpy(x.T.copy()).cuda() for x in (T, B0))
    assert ctx.lib.dplasma_b200_dtrsm_tile(ctx.ptr, None, Right, Lower, Trans, NonUnit, m, n, 1.0, dev(tT), n, dev(tB), m) == 0
    torch.cuda.synchronize()
    assert relerr(tB.cpu().numpy().T, want) < VTOL


@pytest.mark.parametrize("uplo,trans", [(Lower, NoTrans), (Lower, Trans), (Upper, NoTrans), (Upper, Trans)])
@pytest.mark.parametrize("m,n,diag", [(256, 100, NonUnit), (93, 37, NonUnit), (130, 64, Unit)])
def test_dtrsm_tile_left(ctx, oracle, uplo, trans, m, n, diag):
    import torch
    rng = np.random.default_rng(m * n)
    T = oracle.plgsy(float(m), UpperLower, m, m, 5)
    assert oracle.potrf_tile(uplo, T) == 0
    if diag == Unit:
        T = np.asfortranarray(T / 50.0)
    B0 = np.asfortranarray(rng.standard_normal((m, n)))
    want = B0.copy(order="F")
    oracle.trsm_tile(Left, uplo, trans, diag, 0.75, T, want)      # CORE_ztrsm at ztrsm_LLN.jdf:48
    tT, tB = (torch.from_numpy(x.T.copy()).cuda() for x in (T, B0))
    assert ctx.lib.dplasma_b200_dtrsm_tile(ctx.ptr, None, Left, uplo, trans, diag, m, n, 0.75, dev(tT), m, dev(tB), m) == 0
    torch.cuda.synchronize()
    assert relerr(tB.cpu().numpy().T, want) < VTOL


# ------------------------------------------------------------------ dpotrf: public API
@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
@pytest.mark.parametrize("kind", ["tile", "sym", "lapack"])
def test_dpotrf_378_93_reference_ctest_case(ctx, oracle, golden, name, uplo, kind):
    """tests/Testings.cmake:89  testing_dpotrf -N 378 -t 93 (ragged last tile, odd tile size)"""
    N, nb = 378, 93
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    if kind == "lapack":
        d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, kind="lapack", where="host")
        d.host_array()[:N * N] = A0.ravel(order="F")
    else:
        d = make_host(ctx, A0, nb, nb, kind, uplo)
    info = ctx.lib.dplasma_dpotrf(ctx.ptr, uplo, d.ref)
    assert info == 0
    if kind == "lapack":
        F = d.host_array()[:N * N].reshape((N, N), order="F").copy(order="F")
    else:
        F = read_host(d, N, N, nb, nb, kind, uplo, out=A0.copy(order="F") if kind == "sym" else None)
    want = A0.copy(order="F")
    assert oracle.potrf(uplo, want, nb) == 0
    tri = np.tril if uplo == Lower else np.triu
    assert relerr(tri(F), tri(want)) < VTOL
    assert relerr(F.ravel(order="F")[::31], golden[f"d378_F_{name}_s31"]) < VTOL
    # the reference's own acceptance test, evaluated by the oracle on the GPU result
    bad, res, _, _ = oracle.check_potrf(uplo, F, A0)
    assert bad == 0 and res < 60.0
    d.free()


@pytest.mark.parametrize("N,nb", [(1024, 128), (2048, 512), (1536, 256), (1700, 320)])
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_dpotrf_fast_path_vs_oracle(ctx, oracle, N, nb, uplo):
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    d = make_host(ctx, A0, nb, nb)
    assert ctx.lib.dplasma_dpotrf(ctx.ptr, uplo, d.ref) == 0
    F = read_host(d, N, N, nb, nb)
    want = A0.copy(order="F")
    assert oracle.potrf(uplo, want, nb) == 0
    tri = np.tril if uplo == Lower else np.triu
    assert relerr(tri(F), tri(want)) < VTOL
    other = (lambda M: np.triu(M, 1)) if uplo == Lower else (lambda M: np.tril(M, -1))
    assert np.array_equal(other(F), other(A0))
    d.free()


@pytest.mark.parametrize("kind", ["tile", "sym", "lapack"])
@pytest.mark.parametrize("gbs", [None, "0.02", "0.5", "1000"])
@pytest.mark.parametrize("N,nb", [(2048, 128), (1700, 112)])
def test_dpotrf_host_matrix_pipelined_stage_in(ctx, oracle, kind, gbs, N, nb):
    """host matrices take the pipelined path: the tile columns travel to the GPU while the factorization runs, each
    step updates the columns that have landed and later arrivals are caught up with all finished panels in one
    launch; finished columns are written back early.  The schedule follows a MODEL of the copy rate (events decide
    correctness): the result must be the same whatever the model believes -- from a link so slow that no column
    is ever expected (everything caught up at the last moment) to one so fast that everything is expected at once."""
    import os
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A)
    want = A0.copy(order="F")
    assert oracle.potrf(Lower, want, nb) == 0
    if gbs is None:
        os.environ.pop("DPB_H2D_GBS", None)
    else:
        os.environ["DPB_H2D_GBS"] = gbs
    try:
        if kind == "lapack":
            M = np.asfortranarray(np.tril(A0) + np.triu(np.full((N, N), 7.0), 1))   # junk above the diagonal must survive
            d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, kind="lapack", where="host")
            d.host_array()[:N * N] = M.ravel(order="F")
            assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 0
            F = d.host_array()[:N * N].reshape((N, N), order="F").copy()
            assert np.array_equal(np.triu(F, 1), np.triu(M, 1))
        else:
            d = make_host(ctx, A0, nb, nb, kind, Lower)
            assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 0
            F = read_host(d, N, N, nb, nb, kind, Lower, out=A0.copy(order="F") if kind == "sym" else None)
            assert np.array_equal(np.triu(F, 1), np.triu(A0, 1))
    finally:
        os.environ.pop("DPB_H2D_GBS", None)
    assert relerr(np.tril(F), np.tril(want)) < VTOL
    d.free()


def test_dpotrf_new_start_wait_destruct_device_resident(ctx, oracle):
    """the 3-call pattern of tests/common.h:253-267 on a matrix that lives in HBM"""
    import torch
    N, nb = 1024, 128
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A)
    t = torch.from_numpy(full_to_local(A0, nb, nb)).cuda()
    d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, where=t.data_ptr())
    info = C.c_int(-1)
    tp = ctx.lib.dplasma_dpotrf_New(Lower, d.ref, C.byref(info))
    assert tp and info.value == 0
    assert ctx.lib.parsec_context_add_taskpool(ctx.ptr, tp) == 0
    n0 = ctx.lib.dplasma_b200_kernel_launches(ctx.ptr)
    assert ctx.lib.parsec_context_start(ctx.ptr) == 0
    assert ctx.lib.parsec_context_wait(ctx.ptr) == 0
    ctx.lib.dplasma_dpotrf_Destruct(tp)
    assert info.value == 0
    assert ctx.lib.dplasma_b200_kernel_launches(ctx.ptr) > n0
    assert ctx.lib.dplasma_b200_last_elapsed(ctx.ptr) > 0.0
    F = local_to_full(t.cpu().numpy(), N, N, nb, nb)
    want = A0.copy(order="F")
    oracle.potrf(Lower, want, nb)
    assert relerr(np.tril(F), np.tril(want)) < VTOL


def test_info_semantics(ctx, oracle, golden):
    N, nb = 200, 64
    for uplo, key in ((Lower, "d200_nonspd_info_L"), (Upper, "d200_nonspd_info_U")):
        A = oracle.plgsy(0.0, uplo, N, nb, SEED_A)
        d = make_host(ctx, A, nb, nb)
        assert ctx.lib.dplasma_dpotrf(ctx.ptr, uplo, d.ref) == int(golden[key][0])
        d.free()
    A = oracle.plgsy(200.0, Lower, N, nb, SEED_A)
    A[150, 150] = -1.0
    d = make_host(ctx, A, nb, nb)
    assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 151 == int(golden["d200_pivot150_info_L"][0])
    d.free()
    # bad uplo: New prints and returns NULL, the blocking wrapper returns info == 0 (src/zpotrf_wrapper.c:318-335)
    d = make_host(ctx, A, nb, nb)
    assert ctx.lib.dplasma_dpotrf(ctx.ptr, 999, d.ref) == 0
    d.free()


# ------------------------------------------------------------------ checks and solves
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_checks_match_oracle(ctx, oracle, uplo):
    N, nb, K = 378, 93, 367
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    F = A0.copy(order="F")
    oracle.potrf(uplo, F, nb)
    dF, dA0 = make_host(ctx, F, nb, nb), make_host(ctx, A0, nb, nb)
    assert ctx.lib.check_dpotrf(ctx.ptr, 0, uplo, dF.ref, dA0.ref) == 0
    res = ctx.lib.dplasma_b200_last_check_result(ctx.ptr)
    _, want, _, anorm = oracle.check_potrf(uplo, F, A0)
    assert 0.0 <= res < 60.0 and abs(res - want) < 0.5 * max(want, 0.05)
    assert ctx.lib.dplasma_dlansy(ctx.ptr, dp.InfNorm, uplo, dA0.ref) == pytest.approx(anorm, rel=1e-13)
    # a corrupted factor must be flagged
    Fbad = F.copy(order="F")
    Fbad[200, 200] += 1e-3
    dFb = make_host(ctx, Fbad, nb, nb)
    assert ctx.lib.check_dpotrf(ctx.ptr, 0, uplo, dFb.ref, dA0.ref) == 1
    # solve residual
    B = oracle.plrnt(N, K, nb, nb, SEED_B_POSV)
    X = B.copy(order="F")
    oracle.potrs(uplo, F, X, nb, nb)
    dB, dX = make_host(ctx, B, nb, nb), make_host(ctx, X, nb, nb)
    assert ctx.lib.dplasma_dlange(ctx.ptr, dp.InfNorm, dB.ref) == pytest.approx(oracle.lange_inf(B), rel=1e-13)
    assert ctx.lib.check_daxmb(ctx.ptr, 0, uplo, dA0.ref, dB.ref, dX.ref) == 0
    res = ctx.lib.dplasma_b200_last_check_result(ctx.ptr)
    _, want = oracle.check_axmb(uplo, A0, B.copy(order="F"), X)
    assert 0.0 <= res < 60.0 and abs(res - want) < 0.5 * max(want, 0.05)
    for d in (dF, dA0, dFb, dB, dX):
        d.free()


@pytest.mark.parametrize("name,uplo", [("L", Lower), ("U", Upper)])
def test_dposv_378_93_367(ctx, oracle, golden, name, uplo):
    """tests/Testings.cmake:90  testing_dposv -N 378 -t 93 -K 367"""
    N, nb, K = 378, 93, 367
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    B = oracle.plrnt(N, K, nb, nb, SEED_B_POSV)
    dA, dB = make_host(ctx, A0, nb, nb), make_host(ctx, B, nb, nb)
    assert ctx.lib.dplasma_dposv(ctx.ptr, uplo, dA.ref, dB.ref) == 0
    X = read_host(dB, N, K, nb, nb)
    want = B.copy(order="F")
    Aw = A0.copy(order="F")
    assert oracle.posv(uplo, Aw, want, nb, nb) == 0
    assert relerr(X, want) < 1e-10
    assert relerr(X.ravel(order="F")[::41], golden[f"d378_X_{name}_s41"]) < 1e-10
    bad, res = oracle.check_axmb(uplo, A0, B.copy(order="F"), X)
    assert bad == 0 and res < 60.0
    dA.free(); dB.free()


@pytest.mark.parametrize("uplo,trans", [(Lower, NoTrans), (Lower, Trans), (Upper, NoTrans), (Upper, Trans)])
def test_dtrsm_tiled_left(ctx, oracle, uplo, trans):
    N, nb, K, nbr = 300, 64, 100, 48
    T = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    oracle.potrf(uplo, T, nb)
    B = oracle.plrnt(N, K, nb, nbr, 7)
    want = B.copy(order="F")
    assert oracle.trsm(Left, uplo, trans, NonUnit, 1.5, T, want, nb, nbr) == 0
    dT, dB = make_host(ctx, T, nb, nb), make_host(ctx, B, nb, nbr)
    assert ctx.lib.dplasma_dtrsm(ctx.ptr, Left, uplo, trans, NonUnit, 1.5, dT.ref, dB.ref) == 0
    assert relerr(read_host(dB, N, K, nb, nbr), want) < 1e-10
    dT.free(); dB.free()


def test_pdpotrf_single_process(ctx, oracle):
    """tools/cscalapack/pdpotrf.c:94: pdpotrf_("L", &n, A, &i1, &i1, descA, &info) on a 1x1 grid"""
    N, nb = 500, 64
    A0 = oracle.plgsy(float(N), UpperLower, N, nb, SEED_A)
    for u, uplo in ((b"L", Lower), (b"U", Upper)):
        A = np.asfortranarray(A0.copy())
        ctx.lib.dplasma_b200_blacs_gridinit(0, 1, 1, 0, 0)
        desc = (C.c_int * 9)(1, 0, N, N, nb, nb, 0, 0, N)
        n, one, info = C.c_int(N), C.c_int(1), C.c_int(-7)
        ctx.lib.pdpotrf_(u, C.byref(n), A.ctypes.data_as(C.c_void_p), C.byref(one), C.byref(one), desc, C.byref(info))
        assert info.value == 0
        want = A0.copy(order="F")
        oracle.potrf(uplo, want, nb)
        tri = np.tril if uplo == Lower else np.triu
        assert relerr(tri(A), tri(want)) < VTOL


@pytest.mark.parametrize("gN,ia,N", [(700, 129, 572), (700, 65, 256), (640, 1, 640)])
def test_pdpotrf_blacs_symbols_and_submatrix(ctx, oracle, gN, ia, N):
    """The reference's wrapper asks BLACS for the grid (Cblacs_gridinfo, common.h:250) and accepts sub( A ) =
    A(IA:IA+N-1, JA:JA+N-1) with block-aligned IA, JA (dplasma_wrapper_pdpotrf.c:209-211): a process that has those
    symbols is served by them, and only the uplo triangle of sub( A ) changes in the caller's array."""
    from tests.util import load_fake_blacs
    nb = 64
    fb = load_fake_blacs()
    fb.fake_blacs_set(42, 1, 1, 0, 0)
    calls0 = C.c_int.in_dll(fb, "g_fake_blacs_calls").value
    rng = np.random.default_rng(gN + ia)
    G0 = np.asfortranarray(rng.standard_normal((gN, gN)))
    S0 = oracle.plgsy(float(N), UpperLower, N, nb, SEED_A)
    o = ia - 1
    G0[o:o + N, o:o + N] = S0
    for u, uplo in ((b"L", Lower), (b"U", Upper)):
        G = G0.copy(order="F")
        desc = (C.c_int * 9)(1, 42, gN, gN, nb, nb, 0, 0, gN)
        n, i1, info = C.c_int(N), C.c_int(ia), C.c_int(-7)
        ctx.lib.pdpotrf_(u, C.byref(n), G.ctypes.data_as(C.c_void_p), C.byref(i1), C.byref(i1), desc, C.byref(info))
        assert info.value == 0
        want = S0.copy(order="F")
        assert oracle.potrf(uplo, want, nb) == 0
        tri = np.tril if uplo == Lower else np.triu
        sub = G[o:o + N, o:o + N]
        assert relerr(tri(sub), tri(want)) < VTOL
        changed = G != G0
        allowed = np.zeros_like(changed)
        allowed[o:o + N, o:o + N] = tri(np.ones((N, N), dtype=bool))
        assert not (changed & ~allowed).any()
    assert C.c_int.in_dll(fb, "g_fake_blacs_calls").value > calls0   # the grid came from "BLACS"
    # offsets that do not address the first row of a block are refused like the reference's assert
    n, bad, info = C.c_int(64), C.c_int(3), C.c_int(0)
    G = G0.copy(order="F")
    ctx.lib.pdpotrf_(b"L", C.byref(n), G.ctypes.data_as(C.c_void_p), C.byref(bad), C.byref(bad), desc, C.byref(info))
    assert info.value == -4 and np.array_equal(G, G0)


# ------------------------------------------------------------------ BASELINE sizes through properties
@pytest.mark.parametrize("N,nb", [(4096, 256), (32768, 512)])
def test_baseline_configs_residual_on_gpu(ctx, N, nb):
    """C1 / C2 of BASELINE.json: generate in HBM, factor in place, evaluate the reference's
    check (||L L^T - A|| / (||A|| N eps) <= 60) on the GPU."""
    lib = ctx.lib
    dA = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
    dA0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device")
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), Lower, dA.ref, SEED_A) == 0
    assert lib.dplasma_dplgsy(ctx.ptr, float(N), Lower, dA0.ref, SEED_A) == 0
    assert lib.dplasma_dpotrf(ctx.ptr, Lower, dA.ref) == 0
    assert lib.check_dpotrf(ctx.ptr, 0, Lower, dA.ref, dA0.ref) == 0
    res = lib.dplasma_b200_last_check_result(ctx.ptr)
    assert 0.0 < res < 60.0
    dA.free(); dA0.free()


# ------------------------------------------------------------------ complex double (zpotrf)
def zmake_host(ctx, A, nb):
    N = A.shape[0]
    d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, where="host", cplx=True)
    d.host_array()[:d.elems] = full_to_local(A, nb, nb).view(np.float64)
    return d


def zread_host(d, N, nb):
    return local_to_full(d.host_array()[:d.elems].view(np.complex128), N, N, nb, nb)


@pytest.mark.parametrize("N,nb", [(378, 93), (512, 64)])
@pytest.mark.parametrize("uplo", [Lower, Upper, UpperLower])
def test_zplghe_zplrnt_bit_exact(ctx, oracle, golden, N, nb, uplo):
    d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, where="host", cplx=True)
    d.host_array()[:] = 0.0
    assert ctx.lib.dplasma_zplghe(ctx.ptr, float(N), uplo, d.ref, SEED_A) == 0
    got = zread_host(d, N, nb)
    assert np.array_equal(got, oracle.plgsy(float(N), uplo, N, nb, SEED_A, "z"))
    if N == 378 and uplo == Lower:
        assert np.array_equal(got.ravel(order="F")[::31], golden["z378_A0_L_s31"])
    d.free()
    b = dp.BlockCyclic(ctx.lib, N, 5, nb, 5, where="host", cplx=True)
    assert ctx.lib.dplasma_zplrnt(ctx.ptr, 0, b.ref, SEED_B) == 0
    gotb = local_to_full(b.host_array()[:b.elems].view(np.complex128), N, 5, nb, 5)
    assert np.array_equal(gotb, oracle.plrnt(N, 5, nb, 5, SEED_B, "z"))
    b.free()


@pytest.mark.parametrize("N,nb", [(378, 93), (1024, 128), (2048, 512), (1700, 320)])
def test_zpotrf_lower_vs_oracle(ctx, oracle, golden, N, nb):
    """zpotrf (BASELINE configs[4] family): values against the oracle, the reference's residual test,
    and the frozen output of the reference's CORE_z* kernels for the 378/93 ctest size"""
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A, "z")
    d = zmake_host(ctx, A0, nb)
    assert ctx.lib.dplasma_zpotrf(ctx.ptr, Lower, d.ref) == 0
    F = zread_host(d, N, nb)
    want = A0.copy(order="F")
    assert oracle.potrf(Lower, want, nb, "z") == 0
    assert relerr(np.tril(F), np.tril(want)) < VTOL
    assert np.array_equal(np.triu(F, 1), np.triu(A0, 1))
    if N == 378:
        assert relerr(F.ravel(order="F")[::31], golden["z378_F_L_s31"]) < VTOL
    bad, res, _, _ = oracle.check_potrf(Lower, F, A0, "z")
    assert bad == 0 and res < 60.0
    # the GPU-side check agrees
    dA0 = zmake_host(ctx, A0, nb)
    assert ctx.lib.check_zpotrf(ctx.ptr, 0, Lower, d.ref, dA0.ref) == 0
    assert 0.0 <= ctx.lib.dplasma_b200_last_check_result(ctx.ptr) < 60.0
    assert ctx.lib.dplasma_zlanhe(ctx.ptr, dp.InfNorm, Lower, dA0.ref) == pytest.approx(oracle.lanhe_inf(Lower, A0, "z"), rel=1e-13)
    d.free(); dA0.free()


def test_zpotrf_info(ctx, oracle):
    N, nb = 200, 64
    A = oracle.plgsy(200.0, Lower, N, nb, SEED_A, "z")
    A[150, 150] = -1.0
    d = zmake_host(ctx, A, nb)
    assert ctx.lib.dplasma_zpotrf(ctx.ptr, Lower, d.ref) == 151
    d.free()


@pytest.mark.parametrize("N,nb", [(378, 93), (1024, 128)])
def test_zpotrf_upper_vs_oracle(ctx, oracle, golden, N, nb):
    """testing_zpotrf.c hard-codes dplasmaUpper (tests/testing_zpotrf.c:21): U^H U = A"""
    A0 = oracle.plgsy(float(N), Upper, N, nb, SEED_A, "z")
    d = zmake_host(ctx, A0, nb)
    assert ctx.lib.dplasma_zpotrf(ctx.ptr, Upper, d.ref) == 0
    F = zread_host(d, N, nb)
    want = A0.copy(order="F")
    assert oracle.potrf(Upper, want, nb, "z") == 0
    assert relerr(np.triu(F), np.triu(want)) < VTOL
    assert np.array_equal(np.tril(F, -1), np.tril(A0, -1))
    if N == 378:
        assert relerr(F.ravel(order="F")[::31], golden["z378_F_U_s31"]) < VTOL
    dA0 = zmake_host(ctx, A0, nb)
    assert ctx.lib.check_zpotrf(ctx.ptr, 0, Upper, d.ref, dA0.ref) == 0
    assert 0.0 <= ctx.lib.dplasma_b200_last_check_result(ctx.ptr) < 60.0
    d.free(); dA0.free()


def test_zpotrf_device_resident_residual(ctx):
    """complex factorization at a size with many tiles, generated and checked in HBM"""
    lib = ctx.lib
    N, nb = 8192, 512
    dA = dp.BlockCyclic(lib, N, N, nb, nb, where="device", cplx=True)
    dA0 = dp.BlockCyclic(lib, N, N, nb, nb, where="device", cplx=True)
    assert lib.dplasma_zplghe(ctx.ptr, float(N), Lower, dA.ref, SEED_A) == 0
    assert lib.dplasma_zplghe(ctx.ptr, float(N), Lower, dA0.ref, SEED_A) == 0
    assert lib.dplasma_zpotrf(ctx.ptr, Lower, dA.ref) == 0
    assert lib.check_zpotrf(ctx.ptr, 0, Lower, dA.ref, dA0.ref) == 0
    assert 0.0 < lib.dplasma_b200_last_check_result(ctx.ptr) < 60.0
    dA.free(); dA0.free()


# ------------------------------------------------------------------ edge cases and the C driver
@pytest.mark.parametrize("N,nb", [(1, 64), (7, 64), (64, 64), (65, 64), (129, 128)])
@pytest.mark.parametrize("uplo", [Lower, Upper])
def test_dpotrf_tiny_and_single_tile(ctx, oracle, N, nb, uplo):
    A0 = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    d = make_host(ctx, A0, nb, nb)
    assert ctx.lib.dplasma_dpotrf(ctx.ptr, uplo, d.ref) == 0
    F = read_host(d, N, N, nb, nb)
    want = A0.copy(order="F")
    assert oracle.potrf(uplo, want, nb) == 0
    tri = np.tril if uplo == Lower else np.triu
    assert relerr(tri(F), tri(want)) < VTOL
    B = oracle.plrnt(N, 1, nb, 1, SEED_B)            # one right-hand side
    dB = make_host(ctx, B, nb, 1)
    assert ctx.lib.dplasma_dpotrs(ctx.ptr, uplo, d.ref, dB.ref) == 0
    X = read_host(dB, N, 1, nb, 1)
    wx = B.copy(order="F")
    oracle.potrs(uplo, want, wx, nb, 1)
    assert relerr(X, wx) < 1e-10
    d.free(); dB.free()


def test_empty_matrix_and_alpha_zero(ctx, oracle):
    lib = ctx.lib
    d = dp.BlockCyclic(lib, 0, 0, 64, 64, where="host")
    assert lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 0          # nothing to do, info = 0
    d.free()
    n, one, info = C.c_int(0), C.c_int(1), C.c_int(-7)              # pdpotrf_ with N == 0 returns at once (:148-150)
    desc = (C.c_int * 9)(1, 0, 0, 0, 64, 64, 0, 0, 1)
    lib.pdpotrf_(b"L", C.byref(n), None, C.byref(one), C.byref(one), desc, C.byref(info))
    assert info.value == 0
    # dtrsm with alpha = 0 zeroes B (BLAS semantics)
    N, nb = 130, 64
    T = oracle.plgsy(float(N), Lower, N, nb, SEED_A)
    oracle.potrf(Lower, T, nb)
    B = oracle.plrnt(N, 9, nb, 9, 5)
    dT, dB = make_host(ctx, T, nb, nb), make_host(ctx, B, nb, 9)
    assert lib.dplasma_dtrsm(ctx.ptr, Left, Lower, NoTrans, NonUnit, 0.0, dT.ref, dB.ref) == 0
    assert not read_host(dB, N, 9, nb, 9).any()
    dT.free(); dB.free()


def test_c_driver_written_like_the_reference_test():
    """tests/c/testing_dpotrf.c: plain C against include/dplasma_b200.h, the call sequence of
    tests/testing_zpotrf.c; exit code 0 iff both of the reference's checks pass"""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "tests", "c", "testing_dpotrf")
    subprocess.check_call(["gcc", "-std=c11", "-O2", "-I" + os.path.join(root, "include"),
                           os.path.join(root, "tests", "c", "testing_dpotrf.c"), "-o", exe,
                           "-L" + os.path.join(root, "dplasma_b200", "lib"), "-ldplasma_b200",
                           "-Wl,-rpath," + os.path.join(root, "dplasma_b200", "lib")])
    for args in (["3200", "320", "L"], ["1700", "320", "U"]):      # the reference's GPU ctest sizes (Testings.cmake:146-148)
        out = subprocess.run([exe] + args, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stdout + out.stderr
        assert "Factorization is CORRECT" in out.stdout and "Solution is CORRECT" in out.stdout
        assert "[****] TIME(s)" in out.stdout


# ------------------------------------------------------------------ Right-side triangular solves
@pytest.mark.parametrize("uplo,trans", [(Lower, NoTrans), (Lower, Trans), (Upper, NoTrans), (Upper, Trans)])
@pytest.mark.parametrize("m,n,diag", [(200, 256, NonUnit), (37, 93, NonUnit), (64, 130, Unit)])
def test_dtrsm_tile_right_all_variants(ctx, oracle, uplo, trans, m, n, diag):
    import torch
    rng = np.random.default_rng(m + 3 * n)
    T = oracle.plgsy(float(n), UpperLower, n, n, 5)
    assert oracle.potrf_tile(uplo, T) == 0
    if diag == Unit:
        T = np.asfortranarray(T / 50.0)
    B0 = np.asfortranarray(rng.standard_normal((m, n)))
    want = B0.copy(order="F")
    oracle.trsm_tile(Right, uplo, trans, diag, 1.25, T, want)
    tT, tB = (torch.from_numpy(x.T.copy()).cuda() for x in (T, B0))
    assert ctx.lib.dplasma_b200_dtrsm_tile(ctx.ptr, None, Right, uplo, trans, diag, m, n, 1.25, dev(tT), n, dev(tB), m) == 0
    torch.cuda.synchronize()
    assert relerr(tB.cpu().numpy().T, want) < VTOL


@pytest.mark.parametrize("uplo,trans", [(Lower, NoTrans), (Lower, Trans), (Upper, NoTrans), (Upper, Trans)])
def test_dtrsm_tiled_right(ctx, oracle, uplo, trans):
    """dplasma_dtrsm(Right, ...): src/ztrsm_RLN.jdf / RLT / RUN / RUT"""
    M, N, mb, nb = 230, 300, 48, 64
    T = oracle.plgsy(float(N), uplo, N, nb, SEED_A)
    oracle.potrf(uplo, T, nb)
    B = oracle.plrnt(M, N, mb, nb, 7)
    want = B.copy(order="F")
    assert oracle.trsm(Right, uplo, trans, NonUnit, 1.5, T, want, mb, nb) == 0
    dT, dB = make_host(ctx, T, nb, nb), make_host(ctx, B, mb, nb)
    assert ctx.lib.dplasma_dtrsm(ctx.ptr, Right, uplo, trans, NonUnit, 1.5, dT.ref, dB.ref) == 0
    assert relerr(read_host(dB, M, N, mb, nb), want) < 1e-10
    dT.free(); dB.free()


# ------------------------------------------------------------------ complex solves
def zmake_rect(ctx, B, mb, nbr):
    M, R = B.shape
    d = dp.BlockCyclic(ctx.lib, M, R, mb, nbr, where="host", cplx=True)
    d.host_array()[:d.elems] = full_to_local(B, mb, nbr).view(np.float64)
    return d


@pytest.mark.parametrize("N,nb,K,nbr", [(378, 93, 67, 31), (1024, 128, 200, 96), (640, 64, 64, 64)])
def test_zposv_lower_vs_oracle(ctx, oracle, N, nb, K, nbr):
    """zposv = zpotrf + L then L^H solves (src/zposv_wrapper.c:102-110) against the oracle, plus check_zaxmb"""
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A, "z")
    B = oracle.plrnt(N, K, nb, nbr, SEED_B_POSV, "z")
    dA, dB = zmake_host(ctx, A0, nb), zmake_rect(ctx, B, nb, nbr)
    assert ctx.lib.dplasma_zposv(ctx.ptr, Lower, dA.ref, dB.ref) == 0
    X = local_to_full(dB.host_array()[:dB.elems].view(np.complex128), N, K, nb, nbr)
    Aw, want = A0.copy(order="F"), B.copy(order="F")
    assert oracle.posv(Lower, Aw, want, nb, nbr, "z") == 0
    assert relerr(X, want) < 1e-10
    bad, res = oracle.check_axmb(Lower, A0, B.copy(order="F"), X, "z")
    assert bad == 0 and res < 60.0
    dA0, dB0 = zmake_host(ctx, A0, nb), zmake_rect(ctx, B, nb, nbr)
    assert ctx.lib.check_zaxmb(ctx.ptr, 0, Lower, dA0.ref, dB0.ref, dB.ref) == 0
    got = ctx.lib.dplasma_b200_last_check_result(ctx.ptr)
    assert 0.0 <= got < 60.0 and abs(got - res) < 0.5 * max(res, 0.05)
    assert ctx.lib.dplasma_zlange(ctx.ptr, dp.InfNorm, dB.ref) == pytest.approx(oracle.lange_inf(np.asfortranarray(X), "z"), rel=1e-12)
    for d in (dA, dB, dA0, dB0):
        d.free()


def test_tile_seam_alpha_zero_scales_c(ctx):
    """cuBLAS / BLAS rule at the kernel seam: alpha == 0 (or k == 0) never reads A or B, C = beta C
    (beta == 0 stores zeros even over NaN); dsyrk touches only its triangle"""
    import torch
    m, n = 93, 70
    rng = np.random.default_rng(5)
    C0 = np.asfortranarray(rng.standard_normal((m, n)))
    junk = torch.full((8,), float("nan"), dtype=torch.float64, device="cuda")
    for beta in (2.5, 0.0):
        C1 = C0.copy()
        if beta == 0.0:
            C1[3, 4] = np.nan
        tC = torch.from_numpy(C1.T.copy()).cuda()
        assert ctx.lib.dplasma_b200_dgemm_tile(ctx.ptr, None, NoTrans, Trans, m, n, 64, 0.0, dev(junk), m, dev(junk), n, beta, dev(tC), m) == 0
        torch.cuda.synchronize()
        assert np.array_equal(tC.cpu().numpy().T, beta * C0 if beta else np.zeros_like(C0))
    S0 = np.asfortranarray(rng.standard_normal((n, n)))
    tS = torch.from_numpy(S0.T.copy()).cuda()
    assert ctx.lib.dplasma_b200_dsyrk_tile(ctx.ptr, None, Lower, NoTrans, n, 0, 1.0, dev(junk), n, 3.0, dev(tS), n) == 0
    torch.cuda.synchronize()
    got = tS.cpu().numpy().T
    assert np.array_equal(np.tril(got), np.tril(3.0 * S0)) and np.array_equal(np.triu(got, 1), np.triu(S0, 1))


def test_dpotrf_upper_more_than_65535_tile_pairs(ctx, oracle):
    """dplasmaUpper on one GPU transposes the tile grid in place: nt = 364 gives 66430 tile pairs, more than one
    grid dimension holds (the reference's own testing_zpotrf.c:21 is Upper-only)"""
    N, nb = 2906, 8
    A0 = oracle.plgsy(float(N), Upper, N, nb, SEED_A)
    d = make_host(ctx, A0, nb, nb)
    assert ctx.lib.dplasma_dpotrf(ctx.ptr, Upper, d.ref) == 0
    F = read_host(d, N, N, nb, nb)
    want = A0.copy(order="F")
    assert oracle.potrf(Upper, want, nb) == 0
    assert relerr(np.triu(F), np.triu(want)) < VTOL
    assert np.array_equal(np.tril(F, -1), np.tril(A0, -1))
    d.free()


@pytest.mark.parametrize("kind", ["tile", "sym", "lapack"])
@pytest.mark.parametrize("N,nb,budget_mb", [(4096, 256, 90), (1700, 112, 6.5), (2048, 128, 7)])
def test_dpotrf_out_of_core_host_matrix(ctx, oracle, kind, N, nb, budget_mb):
    """a host matrix larger than the device memory the factorization may use (the reference's `lowmem` GPU test limits
    PaRSEC's device tile cache the same way, tests/Testings.cmake:147): left-looking by blocks of tile columns, finished
    columns streamed back from the host as update operands.  Same result as the in-core path, whatever the storage."""
    import os
    A0 = oracle.plgsy(float(N), Lower, N, nb, SEED_A)
    want = A0.copy(order="F")
    assert oracle.potrf(Lower, want, nb) == 0
    os.environ["DPB_DEVICE_BUDGET_MB"] = str(budget_mb)
    try:
        if kind == "lapack":
            M = np.asfortranarray(np.tril(A0) + np.triu(np.full((N, N), 7.0), 1))
            d = dp.BlockCyclic(ctx.lib, N, N, nb, nb, kind="lapack", where="host")
            d.host_array()[:N * N] = M.ravel(order="F")
            assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 0
            F = d.host_array()[:N * N].reshape((N, N), order="F").copy()
            assert np.array_equal(np.triu(F, 1), np.triu(M, 1))
        else:
            d = make_host(ctx, A0, nb, nb, kind, Lower)
            assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d.ref) == 0
            F = read_host(d, N, N, nb, nb, kind, Lower, out=A0.copy(order="F") if kind == "sym" else None)
            assert np.array_equal(np.triu(F, 1), np.triu(A0, 1))
        # not positive definite: the global index of the first bad leading minor comes back like in core
        B0 = A0.copy(order="F")
        bad = 5 * nb + 3
        B0[bad, bad] = -1.0
        d2 = make_host(ctx, B0, nb, nb, "tile", Lower)
        assert ctx.lib.dplasma_dpotrf(ctx.ptr, Lower, d2.ref) == bad + 1
        d2.free()
    finally:
        os.environ.pop("DPB_DEVICE_BUDGET_MB", None)
    assert relerr(np.tril(F), np.tril(want)) < VTOL
    d.free()
