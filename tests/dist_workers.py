"""Worker functions for the multi-process tests (spawned with torch.multiprocessing)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _init(rank, world, port, backend):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group(backend, rank=rank, world_size=world)
    return dist


def gloo_cholesky_worker(rank, world, port, P, Q, N, nb, q):
    """CPU: distributed right-looking tiled Cholesky with the ORACLE's tile kernels, driven by the
    library's own per-step communication plan (dplasma_b200_panel_plan) over gloo broadcasts."""
    import torch
    import dplasma_b200 as dp
    from oracle.oracle import Oracle, Lower, NoTrans, Trans, ConjTrans, NonUnit, Right
    from tests.util import full_to_local, local_to_full
    dist = _init(rank, world, port, "gloo")
    try:
        lib, orc = dp.load(), Oracle()
        prow, pcol = rank // Q, rank % Q
        row_groups = [dist.new_group([r * Q + c for c in range(Q)]) for r in range(P)]
        col_groups = [dist.new_group([r * Q + c for r in range(P)]) for c in range(Q)]
        A0 = orc.plgsy(float(N), Lower, N, nb, 3872)
        s = dp.capi.BlockCyclicStruct()
        lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, rank, nb, nb, N, N, 0, 0, N, N, P, Q, 1, 1, 0, 0)
        lmt, lnt, mt = s.nb_elem_r, s.nb_elem_c, s.super.mt
        flat = full_to_local(A0, nb, nb, P, Q, rank)
        tiles = flat.reshape(lnt, lmt, nb, nb) if flat.size else np.zeros((lnt, lmt, nb, nb))

        def tile(il, jl):  # column-major nb x nb view
            return tiles[jl, il].reshape((nb, nb)).T  # stored flat column-major -> Fortran view

        def tdim(g):
            return N - g * nb if g == mt - 1 else nb
        info = 0
        for k in range(mt):
            out = (C.c_int * 6)()
            cr, cs = (C.c_int * max(lnt, 1))(), (C.c_int * max(lnt, 1))()
            lib.dplasma_b200_panel_plan(C.byref(s), k, k + 1, out, cr, cs)
            kq, kp, il0, nrows, jl0, ncols = list(out)
            assert kq == k % Q and kp == k % P
            tk = tdim(k)
            jlk = k // Q
            if pcol == kq:
                Lkk = torch.zeros(nb, nb, dtype=torch.float64)
                if prow == kp:
                    T = np.asfortranarray(tile(k // P, jlk)[:tk, :tk])
                    ii = orc.potrf_tile(Lower, T)
                    if ii and not info:
                        info = k * nb + ii
                    tile(k // P, jlk)[:tk, :tk] = T
                    Lkk = torch.from_numpy(np.ascontiguousarray(tile(k // P, jlk)))
                dist.broadcast(Lkk, src=kp * Q + pcol, group=col_groups[pcol])
                Lk = np.asfortranarray(Lkk.numpy()[:tk, :tk])
                for il in range(il0, il0 + nrows):
                    gm = il * P + prow
                    B = np.asfortranarray(tile(il, jlk)[:tdim(gm), :tk])
                    orc.trsm_tile(Right, Lower, ConjTrans, NonUnit, 1.0, Lk, B)
                    tile(il, jlk)[:tdim(gm), :tk] = B
            if k + 1 >= mt:
                break
            # (2) row broadcast of the slab [il0, il0+nrows)
            slab = torch.zeros(max(nrows, 0), nb, nb, dtype=torch.float64)
            if pcol == kq and nrows > 0:
                slab = torch.from_numpy(np.stack([np.ascontiguousarray(tile(il, jlk)) for il in range(il0, il0 + nrows)]))
            if nrows > 0:
                dist.broadcast(slab, src=prow * Q + kq, group=row_groups[prow])
            rowp = {il0 + i: slab[i].numpy() for i in range(nrows)}
            # (3) column broadcasts, root alternating with n % P
            colp = {}
            for c in range(ncols):
                jl = jl0 + c
                t = torch.zeros(nb, nb, dtype=torch.float64)
                if cr[c] == prow:
                    t = torch.from_numpy(np.ascontiguousarray(rowp[cs[c]]))
                dist.broadcast(t, src=cr[c] * Q + pcol, group=col_groups[pcol])
                colp[jl] = t.numpy()
            # trailing update on my tiles
            for jl in range(jl0, jl0 + ncols):
                gn = jl * Q + pcol
                for il in range(il0, il0 + nrows):
                    gm = il * P + prow
                    if gm < gn:
                        continue
                    Am = np.asfortranarray(rowp[il][:tdim(gm), :nb])
                    Cmn = np.asfortranarray(tile(il, jl)[:tdim(gm), :tdim(gn)])
                    if gm == gn:
                        orc.herk_tile(Lower, NoTrans, -1.0, Am, 1.0, Cmn)
                    else:
                        Bn = np.asfortranarray(colp[jl][:tdim(gn), :nb])
                        orc.gemm_tile(NoTrans, ConjTrans, -1.0, Am, Bn, 1.0, Cmn)
                    tile(il, jl)[:tdim(gm), :tdim(gn)] = Cmn
        # compare my tiles with the single-process oracle
        want = A0.copy(order="F")
        winfo = orc.potrf(Lower, want, nb)
        got = local_to_full(tiles.reshape(-1), N, N, nb, nb, P, Q, rank)
        mask = np.zeros((N, N), dtype=bool)
        for jl in range(lnt):
            for il in range(lmt):
                gm, gn = il * P + prow, jl * Q + pcol
                if gm >= gn:
                    mask[gm * nb:gm * nb + tdim(gm), gn * nb:gn * nb + tdim(gn)] = True
        mask &= np.tril(np.ones((N, N), dtype=bool))
        err = float(np.abs(got - want)[mask].max()) if mask.any() else 0.0
        infos = [None] * world
        dist.all_gather_object(infos, info)
        q.put((rank, err, max(infos), winfo))
    finally:
        dist.destroy_process_group()


def nccl_potrf_worker(rank, world, port, P, Q, N, nb, q, check_values, cplx=False):
    """GPU: dplasma_dpotrf (or zpotrf) on a P x Q grid, one process per GPU, through the C ABI."""
    import torch
    import dplasma_b200 as dp
    from tests.util import full_to_local
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        dA = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where=1, cplx=cplx)   # descriptor first, buffer from torch
        tA = torch.zeros(max(dA.elems, 1), dtype=torch.float64, device="cuda")
        dA.s.mat = tA.data_ptr()
        dA0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device", cplx=cplx)
        gen = lib.dplasma_zplghe if cplx else lib.dplasma_dplgsy
        fac = lib.dplasma_zpotrf if cplx else lib.dplasma_dpotrf
        chk = lib.check_zpotrf if cplx else lib.check_dpotrf
        assert gen(ctx.ptr, float(N), dp.Lower, dA.ref, 3872) == 0
        assert gen(ctx.ptr, float(N), dp.Lower, dA0.ref, 3872) == 0
        info = fac(ctx.ptr, dp.Lower, dA.ref)
        t = lib.dplasma_b200_last_elapsed(ctx.ptr)
        bad = chk(ctx.ptr, 0, dp.Lower, dA.ref, dA0.ref)
        res = lib.dplasma_b200_last_check_result(ctx.ptr)
        err = -1.0
        if check_values:
            from oracle.oracle import Oracle
            orc = Oracle()
            pch = "z" if cplx else "d"
            want = orc.plgsy(float(N), dp.Lower, N, nb, 3872, pch)
            assert orc.potrf(dp.Lower, want, nb, pch) == 0
            wl = full_to_local(np.tril(want), nb, nb, P, Q, rank)
            from tests.util import local_to_full
            torch.cuda.synchronize()
            g = tA.cpu().numpy()[:dA.elems]
            if cplx:
                g = g.view(np.complex128)
            # compare only the lower-triangular entries (tiles above the diagonal are never touched)
            gl = full_to_local(np.tril(local_to_full(g, N, N, nb, nb, P, Q, rank)), nb, nb, P, Q, rank)
            err = float(np.abs(gl - wl).max() / np.abs(want).max())
        # not positive definite: info must come back as the MAX over ranks of the global index
        assert gen(ctx.ptr, 0.0, dp.Lower, dA.ref, 3872) == 0
        info_bad = fac(ctx.ptr, dp.Lower, dA.ref)
        q.put((rank, info, bad, res, err, t, info_bad))
        dA0.free()
        ctx.close()
    finally:
        dist.destroy_process_group()


def nccl_posv_worker(rank, world, port, P, Q, N, nb, K, nbr, uplo_name, q, check_values):
    """GPU: dplasma_dposv (+ check_dpotrf / check_daxmb) on a P x Q grid, either triangle."""
    import torch
    import dplasma_b200 as dp
    from tests.util import full_to_local, local_to_full
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        uplo = dp.Lower if uplo_name == "L" else dp.Upper

        def dev_desc(M_, N_, mb_, nb_):
            d = dp.BlockCyclic(lib, M_, N_, mb_, nb_, P, Q, rank, where=1)
            t = torch.zeros(max(d.elems, 1), dtype=torch.float64, device="cuda")
            d.s.mat = t.data_ptr()
            return d, t
        dA, tA = dev_desc(N, N, nb, nb)
        dA0, tA0 = dev_desc(N, N, nb, nb)
        dB, tB = dev_desc(N, K, nb, nbr)
        dX, tX = dev_desc(N, K, nb, nbr)
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, dA.ref, 3872) == 0
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, dA0.ref, 3872) == 0
        assert lib.dplasma_dplrnt(ctx.ptr, 0, dB.ref, 2354) == 0
        assert lib.dplasma_dplrnt(ctx.ptr, 0, dX.ref, 2354) == 0
        info = lib.dplasma_dposv(ctx.ptr, uplo, dA.ref, dX.ref)
        bad_f = lib.check_dpotrf(ctx.ptr, 0, uplo, dA.ref, dA0.ref)
        res_f = lib.dplasma_b200_last_check_result(ctx.ptr)
        bad_s = lib.check_daxmb(ctx.ptr, 0, uplo, dA0.ref, dB.ref, dX.ref)
        res_s = lib.dplasma_b200_last_check_result(ctx.ptr)
        errF = errX = -1.0
        if check_values:
            from oracle.oracle import Oracle
            orc = Oracle()
            A0 = orc.plgsy(float(N), uplo, N, nb, 3872)
            Bm = orc.plrnt(N, K, nb, nbr, 2354)
            F, Xw = A0.copy(order="F"), Bm.copy(order="F")
            assert orc.posv(uplo, F, Xw, nb, nbr) == 0
            torch.cuda.synchronize()
            tri = np.tril if uplo == dp.Lower else np.triu
            gotF = local_to_full(tA.cpu().numpy()[:dA.elems], N, N, nb, nb, P, Q, rank)
            wl = full_to_local(tri(F), nb, nb, P, Q, rank)
            gl = full_to_local(tri(gotF), nb, nb, P, Q, rank)
            errF = float(np.abs(gl - wl).max() / np.abs(F).max())
            gx = tX.cpu().numpy()[:dX.elems]
            wx = full_to_local(Xw, nb, nbr, P, Q, rank)
            errX = float(np.abs(gx - wx).max() / np.abs(Xw).max()) if gx.size else 0.0
        q.put((rank, info, bad_f, res_f, bad_s, res_s, errF, errX))
        ctx.close()
    finally:
        dist.destroy_process_group()


def nccl_pdpotrf_worker(rank, world, port, P, Q, N, nb, uplo_name, q):
    """GPU: the ScaLAPACK-compatible pdpotrf_ on a P x Q grid: each rank passes its own
    column-major local array (LLD = local rows) and the 9-int array descriptor, like
    tools/cscalapack/pdpotrf.c:94."""
    import torch
    import dplasma_b200 as dp
    from tests.util import lapack_local
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        lib.dplasma_b200_wrapper_set_context(C.c_void_p(ctx.ptr))
        from oracle.oracle import Oracle
        orc = Oracle()
        uplo = dp.Lower if uplo_name == "L" else dp.Upper
        A0 = orc.plgsy(float(N), dp.UpperLower, N, nb, 3872)
        loc, ri, ci = lapack_local(A0, nb, nb, P, Q, rank)
        lld = max(1, loc.shape[0])
        lib.dplasma_b200_blacs_gridinit(7, P, Q, rank // Q, rank % Q)
        desc = (C.c_int * 9)(1, 7, N, N, nb, nb, 0, 0, lld)
        n, one, info = C.c_int(N), C.c_int(1), C.c_int(-7)
        lib.pdpotrf_(uplo_name.encode(), C.byref(n), loc.ctypes.data_as(C.c_void_p), C.byref(one), C.byref(one), desc, C.byref(info))
        want = A0.copy(order="F")
        assert orc.potrf(uplo, want, nb) == 0
        tri = np.tril if uplo == dp.Lower else np.triu
        mask = tri(np.ones((N, N), dtype=bool))[np.ix_(ri, ci)]
        err = float(np.abs(loc - want[np.ix_(ri, ci)])[mask].max() / np.abs(want).max()) if mask.any() else 0.0
        # the other triangle of the caller's array is left alone
        untouched = bool(np.array_equal(loc[~mask], A0[np.ix_(ri, ci)][~mask]))
        q.put((rank, info.value, err, untouched))
        lib.dplasma_b200_wrapper_set_context(None)
        ctx.close()
    finally:
        dist.destroy_process_group()


def nccl_host_potrf_worker(rank, world, port, P, Q, N, nb, kind, q):
    """GPU: dplasma_dpotrf on a grid with the matrix in pinned HOST memory (tile or sym-packed storage):
    staged in, factored, finished tile columns written back early on the copy stream."""
    import torch
    import dplasma_b200 as dp
    from tests.util import full_to_local, local_to_full
    from oracle.oracle import Oracle
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        orc = Oracle()
        A0 = orc.plgsy(float(N), dp.Lower, N, nb, 3872)
        sym = dp.Lower if kind == "sym" else None
        d = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, kind=kind, uplo=dp.Lower if kind == "sym" else dp.UpperLower, where="host")
        loc0 = full_to_local(A0, nb, nb, P, Q, rank, sym_uplo=sym)
        d.host_array()[:d.elems] = loc0
        info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref)
        want = A0.copy(order="F")
        assert orc.potrf(dp.Lower, want, nb) == 0
        got = local_to_full(d.host_array()[:d.elems].copy(), N, N, nb, nb, P, Q, rank, sym_uplo=sym)
        wl = full_to_local(np.tril(want), nb, nb, P, Q, rank, sym_uplo=sym)
        gl = full_to_local(np.tril(got), nb, nb, P, Q, rank, sym_uplo=sym)
        err = float(np.abs(gl - wl).max() / np.abs(want).max()) if gl.size else 0.0
        q.put((rank, info, err))
        d.free()
        ctx.close()
    finally:
        dist.destroy_process_group()


def nccl_poinv_worker(rank, world, port, P, Q, N, nb, uplo_name, q):
    """GPU: dplasma_dpoinv, dplasma_dpotrf + dplasma_dpotri and dplasma_dtrtri on a P x Q grid against the oracle."""
    import torch
    import dplasma_b200 as dp
    from tests.util import full_to_local
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        uplo = dp.Lower if uplo_name == "L" else dp.Upper

        def dev_desc():
            d = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where=1)
            t = torch.zeros(max(d.elems, 1), dtype=torch.float64, device="cuda")
            d.s.mat = t.data_ptr()
            return d, t
        from oracle.oracle import Oracle, NonUnit
        orc = Oracle()
        tri = np.tril if uplo == dp.Lower else np.triu
        A0 = orc.plgsy(float(N), uplo, N, nb, 3872)
        F = A0.copy(order="F")
        assert orc.potrf(uplo, F, nb) == 0
        T = F.copy(order="F")
        assert orc.trtri(uplo, NonUnit, T, nb) == 0
        Pw = F.copy(order="F")
        assert orc.potri(uplo, Pw, nb) == 0

        def local_err(t, d, want):
            torch.cuda.synchronize()
            got = t.cpu().numpy()[:d.elems]
            w = full_to_local(want, nb, nb, P, Q, rank)
            # compare the stored triangle only: rebuild the mask in the local layout
            m = full_to_local(tri(np.ones((N, N))), nb, nb, P, Q, rank) != 0
            o = (full_to_local(np.ones((N, N)), nb, nb, P, Q, rank) != 0) & ~m   # inside the matrix, other triangle
            return float(np.abs(got[m] - w[m]).max() / np.abs(want).max()) if m.any() else 0.0, \
                bool(np.array_equal(got[o], full_to_local(A0, nb, nb, P, Q, rank)[o]))
        # (1) one call
        dA, tA = dev_desc()
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, dA.ref, 3872) == 0
        info1 = lib.dplasma_dpoinv(ctx.ptr, uplo, dA.ref)
        e1, other1 = local_err(tA, dA, Pw)
        # (2) factor, then potri
        dB, tB = dev_desc()
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, dB.ref, 3872) == 0
        info2 = lib.dplasma_dpotrf(ctx.ptr, uplo, dB.ref)
        info2 = info2 or lib.dplasma_dpotri(ctx.ptr, uplo, dB.ref)
        e2, other2 = local_err(tB, dB, Pw)
        # (3) factor, then trtri alone
        dC, tC = dev_desc()
        assert lib.dplasma_dplgsy(ctx.ptr, float(N), uplo, dC.ref, 3872) == 0
        info3 = lib.dplasma_dpotrf(ctx.ptr, uplo, dC.ref)
        info3 = info3 or lib.dplasma_dtrtri(ctx.ptr, uplo, NonUnit, dC.ref)
        e3, other3 = local_err(tC, dC, T)
        # (4) not positive definite: the factorization's info, MAX-reduced, nothing inverted
        dD, tD = dev_desc()
        assert lib.dplasma_dplgsy(ctx.ptr, 0.0, uplo, dD.ref, 3872) == 0
        info4 = lib.dplasma_dpoinv(ctx.ptr, uplo, dD.ref)
        q.put((rank, info1, e1, other1, info2, e2, other2, info3, e3, other3, info4))
        ctx.close()
    finally:
        dist.destroy_process_group()
