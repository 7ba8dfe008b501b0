"""GPU: the allocation-free tile Cholesky, dplasma_b200_dpotrf_tile_noalloc, through the C ABI.

Some callers may not allocate device memory at all (a task body under a runtime that has
reserved the whole GPU, a loop that cannot take cudaMalloc's implicit synchronisation, a
process with a fixed memory budget).  For them the library carries a 512 KiB workspace inside
its own image.  This test checks that entry point against the regular, allocating
dplasma_b200_dpotrf_tile on the same inputs, twice in a row (the workspace is shared between
calls and keeps the previous call's contents, so the second call runs on a dirty workspace),
and that sizes beyond the built-in workspace are refused.  The allocation-free entry is called
before any library context exists in this test (it needs none), the free device memory reported by
the driver must not move across its calls once the first call on the device has loaded the kernels,
and where two GPUs are visible the same process drives both (the array and the kernels' shared-memory
opt-in are per device).  It is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import dplasma_b200 as dp

pytestmark = pytest.mark.gpu
VTOL = 1e-11  # same tolerance as tests/test_gpu_parity.py


def dev(t):
    return C.c_void_p(t.data_ptr())


def free_bytes(torch):
    torch.cuda.synchronize()
    return torch.cuda.mem_get_info()[0]  # cudaMemGetInfo: what the driver says is free on the current device


def test_noalloc_matches_allocating_entry_point_twice(oracle):
    import time
    import torch
    lib = dp.load()
    ctx = None
    try:
        # first a large tile (fills all 16 workspace blocks), then a small ragged one on the dirty workspace
        for call, (n, seed) in enumerate([(1024, 21), (448 + 37, 22)]):
            A0 = oracle.plgsy(float(n), dp.UpperLower, n, n, seed)
            tRef = torch.from_numpy(A0.T.copy()).cuda()
            tNew = torch.from_numpy(A0.T.copy()).cuda()
            tAgain = torch.from_numpy(A0.T.copy()).cuda()
            iRef = torch.zeros(1, dtype=torch.int32, device="cuda")
            iNew = torch.zeros(2, dtype=torch.int32, device="cuda")
            # no context yet on the first round: the entry point needs none
            assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, n, dev(tNew), n, dev(iNew), 0) == 0
            # the kernels are loaded now: from here on the driver's free-memory figure must not move
            free0 = free_bytes(torch)
            t0 = time.perf_counter()
            assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, n, dev(tAgain), n, C.c_void_p(iNew.data_ptr() + 4), 0) == 0
            torch.cuda.synchronize()
            t_noalloc = time.perf_counter() - t0
            assert free_bytes(torch) == free0, "dplasma_b200_dpotrf_tile_noalloc changed the free device memory"
            assert torch.equal(tNew, tAgain)  # same input, same (dirty) workspace, same bits
            if ctx is None:
                ctx = dp.Context(lib)
                w = torch.from_numpy(A0.T.copy()).cuda()   # warm the allocating entry (workspace, kernels)
                assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, dev(w), n, dev(iRef), 0) == 0
                torch.cuda.synchronize()
                iRef.zero_()
            t0 = time.perf_counter()
            assert lib.dplasma_b200_dpotrf_tile(ctx.ptr, None, dp.Lower, n, dev(tRef), n, dev(iRef), 0) == 0
            torch.cuda.synchronize()
            t_alloc = time.perf_counter() - t0
            print(f"dpotrf tile n={n}: allocation-free entry {t_noalloc * 1e3:.3f} ms, allocating entry {t_alloc * 1e3:.3f} ms")
            ref, new = tRef.cpu().numpy().T, tNew.cpu().numpy().T
            assert int(iRef.item()) == 0 and iNew.cpu().tolist() == [0, 0]
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
        if ctx is not None:
            ctx.close()


def test_noalloc_on_a_second_device_of_the_same_process(oracle):
    """One process, two GPUs: the built-in array exists once per device and the kernels' shared-memory opt-in is a
    per-device setting, so the first call on the second device has to do its own set-up."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs in one process")
    lib = dp.load()
    n = 512 + 64 + 9
    A0 = oracle.plgsy(float(n), dp.UpperLower, n, n, 33)
    want = A0.copy(order="F")
    assert oracle.potrf_tile(dp.Lower, want) == 0
    for d in (1, 0, 1):
        with torch.cuda.device(d):
            t = torch.from_numpy(A0.T.copy()).cuda()
            info = torch.zeros(1, dtype=torch.int32, device="cuda")
            assert lib.dplasma_b200_dpotrf_tile_noalloc(None, dp.Lower, n, dev(t), n, dev(info), 0) == 0
            torch.cuda.synchronize()
            got = t.cpu().numpy().T
            assert int(info.item()) == 0
            assert np.abs(np.tril(got) - np.tril(want)).max() / np.abs(want).max() < VTOL, d
