"""CPU: the C-ABI library loads and exports every symbol include/dplasma_b200.h declares;
descriptor initialisers (pure host logic) follow the 2D block-cyclic rules.  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import dplasma_b200 as dp
from tests.util import full_to_local, local_to_full

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "dplasma_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set()
    for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{}]*\)\s*;", hdr):
        n = m.group(1)
        if n not in ("defined",):
            names.add(n)
    return sorted(names)


def test_library_exports_every_declared_symbol():
    lib = dp.load()
    syms = declared_symbols()
    assert len(syms) >= 40
    missing = [s for s in syms if not hasattr(lib.dll, s)]
    assert not missing, missing
    assert b"sm_100a" in lib.dplasma_b200_version()


def test_no_device_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError):
        dp.Context()


def test_new_rejects_bad_uplo_and_sets_info():
    lib = dp.load()
    s = dp.capi.BlockCyclicStruct()
    lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, 0, 4, 4, 8, 8, 0, 0, 8, 8, 1, 1, 1, 1, 0, 0)
    info = C.c_int(77)
    assert lib.dplasma_dpotrf_New(999, C.byref(s), C.byref(info)) is None  # src/zpotrf_wrapper.c:187-190
    tp = lib.dplasma_dpotrf_New(dp.Lower, C.byref(s), C.byref(info))
    assert tp and info.value == 0  # :192
    lib.dplasma_dpotrf_Destruct(tp)
    # argument codes of dplasma_ztrsm (src/ztrsm_wrapper.c:274-296), checked before any GPU use
    assert lib.dplasma_dtrsm(None, 5, dp.Lower, dp.NoTrans, dp.NonUnit, 1.0, C.byref(s), C.byref(s)) == -1
    assert lib.dplasma_dtrsm(None, dp.Left, 5, dp.NoTrans, dp.NonUnit, 1.0, C.byref(s), C.byref(s)) == -2
    assert lib.dplasma_dtrsm(None, dp.Left, dp.Lower, 5, dp.NonUnit, 1.0, C.byref(s), C.byref(s)) == -3
    assert lib.dplasma_dtrsm(None, dp.Left, dp.Lower, dp.NoTrans, 5, 1.0, C.byref(s), C.byref(s)) == -4
    assert lib.dplasma_dposv(None, 7, C.byref(s), C.byref(s)) == -1   # src/zposv_wrapper.c:73-76
    assert lib.dplasma_dpotrs(None, 7, C.byref(s), C.byref(s)) == -1
    # inversion path: src/ztrtri_wrapper.c:203-216, zlauum_wrapper.c:192-200, zpotri_wrapper.c:62-65
    r = dp.capi.BlockCyclicStruct()
    lib.parsec_matrix_block_cyclic_init(C.byref(r), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, 0, 4, 4, 8, 12, 0, 0, 8, 12, 1, 1, 1, 1, 0, 0)
    assert lib.dplasma_dtrtri(None, 7, dp.NonUnit, C.byref(s)) == -1
    assert lib.dplasma_dtrtri(None, dp.Lower, 7, C.byref(s)) == -2
    assert lib.dplasma_dtrtri(None, dp.Lower, dp.NonUnit, C.byref(r)) == -6
    assert lib.dplasma_dlauum(None, 7, C.byref(s)) == -1
    assert lib.dplasma_dlauum(None, dp.Upper, C.byref(r)) == -6
    assert lib.dplasma_dpotri(None, 7, C.byref(s)) == -1
    assert lib.dplasma_dpoinv(None, 7, C.byref(s)) == -1
    assert lib.dplasma_dpoinv_sync(None, 7, C.byref(s)) == -1
    assert lib.dplasma_dpotri(None, dp.Lower, C.byref(s)) == -100  # no context: nothing runs
    # level-3 ops: src/zgemm_wrapper.c:657-710, src/zsyrk_wrapper.c:229-245 (s is 8x8, r is 8x12, tiles 4x4)
    assert lib.dplasma_dgemm(None, 7, dp.NoTrans, 1.0, C.byref(s), C.byref(r), 0.0, C.byref(r)) == -1
    assert lib.dplasma_dgemm(None, dp.NoTrans, 7, 1.0, C.byref(s), C.byref(r), 0.0, C.byref(r)) == -2
    assert lib.dplasma_dgemm(None, dp.NoTrans, dp.NoTrans, 1.0, C.byref(r), C.byref(r), 0.0, C.byref(r)) == -101
    assert lib.dplasma_dgemm(None, dp.NoTrans, dp.NoTrans, 0.0, C.byref(s), C.byref(r), 1.0, C.byref(r)) == 0
    assert lib.dplasma_dgemm(None, dp.NoTrans, dp.NoTrans, 1.0, C.byref(s), C.byref(r), 1.0, C.byref(r)) == -100
    assert lib.dplasma_dsyrk(None, 7, dp.NoTrans, 1.0, C.byref(r), 0.0, C.byref(s)) == -1
    assert lib.dplasma_dsyrk(None, dp.Lower, 7, 1.0, C.byref(r), 0.0, C.byref(s)) == -2
    assert lib.dplasma_dsyrk(None, dp.Lower, dp.NoTrans, 1.0, C.byref(r), 0.0, C.byref(r)) == -6
    assert lib.dplasma_dsyrk(None, dp.Lower, dp.Trans, 1.0, C.byref(r), 0.0, C.byref(s)) == -4
    # taskpool forms: NULL on an illegal argument, an object otherwise (nothing runs before parsec_context_start)
    assert lib.dplasma_dtrsm_New(5, dp.Lower, dp.NoTrans, dp.NonUnit, 1.0, C.byref(s), C.byref(s)) is None
    assert lib.dplasma_dtrtri_New(dp.Lower, 5, C.byref(s), C.byref(info)) is None
    assert lib.dplasma_dlauum_New(5, C.byref(s)) is None
    assert lib.dplasma_dpoinv_New(5, C.byref(s), C.byref(info)) is None
    assert lib.dplasma_dgemm_New(dp.NoTrans, 5, 1.0, C.byref(s), C.byref(s), 0.0, C.byref(s)) is None
    assert lib.dplasma_dsyrk_New(dp.Lower, 5, 1.0, C.byref(s), 0.0, C.byref(s)) is None
    info.value = 9
    for tp, destruct in ((lib.dplasma_dtrsm_New(dp.Left, dp.Lower, dp.NoTrans, dp.NonUnit, 1.0, C.byref(s), C.byref(s)), lib.dplasma_dtrsm_Destruct),
                         (lib.dplasma_dtrtri_New(dp.Lower, dp.NonUnit, C.byref(s), C.byref(info)), lib.dplasma_dtrtri_Destruct),
                         (lib.dplasma_dlauum_New(dp.Upper, C.byref(s)), lib.dplasma_dlauum_Destruct),
                         (lib.dplasma_dpoinv_New(dp.Upper, C.byref(s), C.byref(info)), lib.dplasma_dpoinv_Destruct),
                         (lib.dplasma_dgemm_New(dp.NoTrans, dp.Trans, 1.0, C.byref(s), C.byref(s), 0.0, C.byref(s)), lib.dplasma_dgemm_Destruct),
                         (lib.dplasma_dsyrk_New(dp.Lower, dp.Trans, 1.0, C.byref(s), 0.0, C.byref(s)), lib.dplasma_dsyrk_Destruct)):
        assert tp
        destruct(tp)
    assert info.value == 0


@pytest.mark.parametrize("M,N,mb,nb,P,Q", [(378, 378, 93, 93, 1, 1), (378, 367, 93, 93, 2, 2), (1000, 1000, 120, 120, 2, 4),
                                           (131072, 131072, 1024, 1024, 2, 4)])
def test_block_cyclic_descriptor(M, N, mb, nb, P, Q):
    lib = dp.load()
    mt, nt = -(-M // mb), -(-N // nb)
    total = 0
    for rank in range(P * Q):
        s = dp.capi.BlockCyclicStruct()
        lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, rank, mb, nb, M, N, 0, 0, M, N, P, Q, 1, 1, 0, 0)
        assert (s.super.mt, s.super.nt, s.super.bsiz) == (mt, nt, mb * nb)
        assert s.grid.rrank == rank // Q and s.grid.crank == rank % Q
        assert s.nb_elem_r == len(range(rank // Q, mt, P)) and s.nb_elem_c == len(range(rank % Q, nt, Q))
        assert s.super.nb_local_tiles == s.nb_elem_r * s.nb_elem_c
        total += s.super.nb_local_tiles
        # owner map == (m mod P)*Q + (n mod Q)
        for (m, n) in [(0, 0), (mt - 1, nt - 1), (mt // 2, nt // 3)]:
            assert lib.dplasma_b200_rank_of(C.byref(s), m, n) == (m % P) * Q + (n % Q)
    assert total == mt * nt
    # sym: only the stored triangle is counted
    if M == N:
        cnt = 0
        for rank in range(P * Q):
            s = dp.capi.BlockCyclicStruct()
            lib.parsec_matrix_sym_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, rank, mb, nb, M, N, 0, 0, M, N, P, Q, dp.Lower)
            cnt += s.super.nb_local_tiles
        assert cnt == mt * (mt + 1) // 2


def test_layout_helpers_roundtrip():
    rng = np.random.default_rng(1)
    A = np.asfortranarray(rng.standard_normal((378, 367)))
    out = np.zeros_like(A)
    for rank in range(6):
        flat = full_to_local(A, 93, 64, 2, 3, rank)
        local_to_full(flat, 378, 367, 93, 64, 2, 3, rank, out=out)
    assert np.array_equal(A, out)


def test_flops_model():
    lib = dp.load()
    assert lib.dplasma_b200_flops_dpotrf(4096.0) == pytest.approx(2.291488e10, rel=1e-6)
    assert lib.dplasma_b200_flops_dpotrf(32768.0) == pytest.approx(1.172866e13, rel=1e-6)  # C2
    assert lib.dplasma_b200_flops_dpotrs(65536.0, 1024.0) == pytest.approx(8.796093e12, rel=1e-6)
