"""GPU: the allocation-free tile Cholesky, dplasma_b200_dpotrf_tile_noalloc, through the C ABI.

Some callers may not allocate device memory at all (a task body under a runtime that has
reserved the whole GPU, a loop that cannot take cudaMalloc's implicit synchronisation, a
process with a fixed memory budget).  For them the library carries a 512 KiB workspace inside
its own image.  This test checks that entry point against the regular, allocating
dplasma_b200_dpotrf_tile on the same inputs, twice in a row (the workspace is shared between
calls and keeps the previous call's contents, so the second call runs on a dirty workspace),
and that sizes beyond the built-in workspace are refused.  It is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import dplasma_b200 as dp

pytestmark = pytest.mark.gpu
VTOL = 1e-11  # same tolerance as tests/test_gpu_parity.py


def dev(t):
    return C.c_void_p(t.data_ptr())


def test_noalloc_matches_allocating_entry_point_twice(oracle):
    import torch
    lib = dp.load()
    ctx = dp.Context(lib)
    try:
        # first a large tile (fills all 16 workspace blocks), then a small ragged one on the dirty workspace
        for call, (n, seed) in enumerate([(1024, 21), (448 + 37, 22)]):
            A0 = oracle.plgsy(float(n), dp.UpperLower, n, n, seed)
            tRef = torch.from_numpy(A0.T.copy()).cuda()
            tNew = torch.from_numpy(A0.T.copy()).cuda()
            iRef = torch.zeros(1, dtype=torch.int32, device="cuda")
            iNew = torch.zeros(1, dtype=torch.int32, device="cuda")
            assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, dev(tRef), n, dev(iRef), 0) == 0
            assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, n, dev(tNew), n, dev(iNew), 0) == 0
            torch.cuda.synchronize()
            ref, new = tRef.cpu().numpy().T, tNew.cpu().numpy().T
            assert int(iRef.item()) == 0 and int(iNew.item()) == 0
            err = np.abs(np.tril(new) - np.tril(ref)).max() / np.abs(ref).max()
            assert err < VTOL, (call, err)
            assert np.array_equal(np.triu(new, 1), np.triu(A0, 1))  # strict upper part untouched
            want = A0.copy(order="F")
            assert oracle.potrf_tile(dp.Lower, want) == 0
            assert np.abs(np.tril(new) - np.tril(want)).max() / np.abs(want).max() < VTOL
        # info is reported like the allocating entry point
        n = 200
        A0 = oracle.plgsy(float(n), dp.UpperLower, n, n, 5)
        A0[120, 120] = -3.0
        t = torch.from_numpy(A0.T.copy()).cuda()
        info = torch.zeros(1, dtype=torch.int32, device="cuda")
        assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, n, dev(t), n, dev(info), 7000) == 0
        torch.cuda.synchronize()
        assert int(info.item()) == 7000 + 121
        # refused: beyond the built-in workspace, unsupported triangle, missing info pointer
        big = torch.zeros(8, dtype=torch.float64, device="cuda")
        assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, 1025, dev(big), 1025, dev(info), 0) == -3
        assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Upper, 64, dev(big), 64, dev(info), 0) == -2
        assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, 2, dev(big), 2, None, 0) == -6
    finally:
        ctx.close()
