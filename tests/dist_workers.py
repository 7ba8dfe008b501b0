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


def nccl_potrf_worker(rank, world, port, P, Q, N, nb, q, check_values, cplx=False, uplo_name="L"):
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
        uplo = dp.Lower if uplo_name == "L" else dp.Upper
        tri = np.tril if uplo == dp.Lower else np.triu
        dA = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where=1, cplx=cplx)   # descriptor first, buffer from torch
        tA = torch.zeros(max(dA.elems, 1), dtype=torch.float64, device="cuda")
        dA.s.mat = tA.data_ptr()
        dA0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device", cplx=cplx)
        gen = lib.dplasma_zplghe if cplx else lib.dplasma_dplgsy
        fac = lib.dplasma_zpotrf if cplx else lib.dplasma_dpotrf
        chk = lib.check_zpotrf if cplx else lib.check_dpotrf
        assert gen(ctx.ptr, float(N), uplo, dA.ref, 3872) == 0
        assert gen(ctx.ptr, float(N), uplo, dA0.ref, 3872) == 0
        info = fac(ctx.ptr, uplo, dA.ref)
        t = lib.dplasma_b200_last_elapsed(ctx.ptr)
        bad = chk(ctx.ptr, 0, uplo, dA.ref, dA0.ref)
        res = lib.dplasma_b200_last_check_result(ctx.ptr)
        err = -1.0
        if check_values:
            from oracle.oracle import Oracle
            orc = Oracle()
            pch = "z" if cplx else "d"
            want = orc.plgsy(float(N), uplo, N, nb, 3872, pch)
            assert orc.potrf(uplo, want, nb, pch) == 0
            wl = full_to_local(tri(want), nb, nb, P, Q, rank)
            from tests.util import local_to_full
            torch.cuda.synchronize()
            g = tA.cpu().numpy()[:dA.elems]
            if cplx:
                g = g.view(np.complex128)
            # compare only the entries of the stored triangle (tiles of the other one are never touched)
            gl = full_to_local(tri(local_to_full(g, N, N, nb, nb, P, Q, rank)), nb, nb, P, Q, rank)
            err = float(np.abs(gl - wl).max() / np.abs(want).max())
        # not positive definite: info must come back as the MAX over ranks of the global index
        assert gen(ctx.ptr, 0.0, uplo, dA.ref, 3872) == 0
        info_bad = fac(ctx.ptr, uplo, dA.ref)
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


def nccl_pdpotrf_worker(rank, world, port, P, Q, N, nb, uplo_name, q, rsrc=0, csrc=0, ia=1, blacs=False):
    """GPU: the ScaLAPACK-compatible pdpotrf_ on a P x Q grid: each rank passes its own
    column-major local array (LLD = local rows) and the 9-int array descriptor, like
    tools/cscalapack/pdpotrf.c:94.  rsrc / csrc: the process row / column that owns block (0,0) (DESCA[6..7]);
    ia: sub( A ) starts at global row and column ia of a larger matrix; blacs: the grid comes from the
    process's "BLACS" (tests/c/fake_blacs.c) instead of dplasma_b200_blacs_gridinit."""
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
        S0 = orc.plgsy(float(N), dp.UpperLower, N, nb, 3872)
        o = ia - 1
        gN = o + N
        A0 = np.asfortranarray(np.random.default_rng(11).standard_normal((gN, gN)))
        A0[o:, o:] = S0
        # the process at physical (r, c) holds what the process at logical ((r - rsrc) % P, (c - csrc) % Q) would
        lrank = ((rank // Q - rsrc) % P) * Q + (rank % Q - csrc) % Q
        loc, ri, ci = lapack_local(A0, nb, nb, P, Q, lrank)
        loc0 = loc.copy()
        lld = max(1, loc.shape[0])
        if blacs:
            from tests.util import load_fake_blacs
            load_fake_blacs().fake_blacs_set(7, P, Q, rank // Q, rank % Q)
        else:
            lib.dplasma_b200_blacs_gridinit(7, P, Q, rank // Q, rank % Q)
        desc = (C.c_int * 9)(1, 7, gN, gN, nb, nb, rsrc, csrc, lld)
        n, one, info = C.c_int(N), C.c_int(ia), C.c_int(-7)
        lib.pdpotrf_(uplo_name.encode(), C.byref(n), loc.ctypes.data_as(C.c_void_p), C.byref(one), C.byref(one), desc, C.byref(info))
        want = A0.copy(order="F")
        sub = S0.copy(order="F")
        assert orc.potrf(uplo, sub, nb) == 0
        tri = np.tril if uplo == dp.Lower else np.triu
        full = np.zeros((gN, gN), dtype=bool)
        full[o:, o:] = tri(np.ones((N, N), dtype=bool))
        want[o:, o:] = sub
        mask = full[np.ix_(ri, ci)]
        err = float(np.abs(loc - want[np.ix_(ri, ci)])[mask].max() / np.abs(sub).max()) if mask.any() else 0.0
        # everything but the uplo triangle of sub( A ) is left alone in the caller's array
        untouched = bool(np.array_equal(loc[~mask], loc0[~mask]))
        q.put((rank, info.value, err, untouched))
        lib.dplasma_b200_wrapper_set_context(None)
        ctx.close()
    finally:
        dist.destroy_process_group()


def nccl_host_potrf_worker(rank, world, port, P, Q, N, nb, kind, q, gbs=None):
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
        if gbs is not None:
            # what the schedule BELIEVES about the host link, a different belief on every rank: each rank decides by
            # itself which of its columns a step updates, the collectives of a step must not depend on it
            os.environ["DPB_H2D_GBS"] = str(float(gbs) * (1 + 3 * rank))
        info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref)
        os.environ.pop("DPB_H2D_GBS", None)
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


def gloo_potri_worker(rank, world, port, P, Q, N, nb, q):
    """CPU: distributed tiled trtri + lauum (lower) with the ORACLE's tile kernels, driven by the library's own
    per-step plans (dplasma_b200_panel_plan for the tile column below the diagonal, dplasma_b200_rowpanel_plan for
    the tile row left of it) over gloo broadcasts.  Step bodies: src/ztrtri_L.jdf:21-187, src/zlauum_L.jdf:38-155."""
    import torch
    import dplasma_b200 as dp
    from oracle.oracle import Oracle, Lower, NoTrans, ConjTrans, NonUnit, Left, Right
    from tests.util import full_to_local, local_to_full
    dist = _init(rank, world, port, "gloo")
    try:
        lib, orc = dp.load(), Oracle()
        prow, pcol = rank // Q, rank % Q
        row_groups = [dist.new_group([r * Q + c for c in range(Q)]) for r in range(P)]
        col_groups = [dist.new_group([r * Q + c for r in range(P)]) for c in range(Q)]
        A0 = orc.plgsy(float(N), Lower, N, nb, 3872)
        F = A0.copy(order="F")
        assert orc.potrf(Lower, F, nb) == 0
        s = dp.capi.BlockCyclicStruct()
        lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, rank, nb, nb, N, N, 0, 0, N, N, P, Q, 1, 1, 0, 0)
        lmt, lnt, mt = s.nb_elem_r, s.nb_elem_c, s.super.mt
        flat = full_to_local(F, nb, nb, P, Q, rank)
        tiles = flat.reshape(lnt, lmt, nb, nb) if flat.size else np.zeros((lnt, lmt, nb, nb))

        def tile(il, jl):
            return tiles[jl, il].reshape((nb, nb)).T

        def tdim(g):
            return N - g * nb if g == mt - 1 else nb

        def bc(t, src, group):
            x = torch.from_numpy(np.ascontiguousarray(t))
            dist.broadcast(x, src=src, group=group)
            return x.numpy()

        def plans(k):
            out = (C.c_int * 6)()
            cr, cs = (C.c_int * max(lnt, 1))(), (C.c_int * max(lnt, 1))()
            lib.dplasma_b200_panel_plan(C.byref(s), k, k + 1, out, cr, cs)
            kq, kp, il0, nrows = out[0], out[1], out[2], out[3]
            ro = (C.c_int * 4)()
            rc, rj = (C.c_int * max(lmt, 1))(), (C.c_int * max(lmt, 1))()
            lib.dplasma_b200_rowpanel_plan(C.byref(s), k, ro, rc, rj)
            assert ro[0] == kp == k % P and ro[1] == kq == k % Q
            return kp, kq, il0, nrows, ro[2], ro[3], list(rc), list(rj)

        def diag_tile_everywhere_it_is_needed(k, kp, kq, down_column, along_row):
            """A(k,k) from its owner down grid column kq and / or along grid row kp (ztrsm_in_A, readA)"""
            t = np.zeros((nb, nb))
            if prow == kp and pcol == kq:
                t = np.array(tile(k // P, k // Q))
            if down_column and pcol == kq:
                t = bc(t, kp * Q + kq, col_groups[kq])
            if along_row and prow == kp:
                t = bc(t, kp * Q + kq, row_groups[kp])
            return t

        def row_k_down_the_columns(k, kp, ncb):
            """tile row k, n < k: every rank of grid row kp sends its tiles down its grid column"""
            return {jl: bc(np.array(tile(k // P, jl)) if prow == kp else np.zeros((nb, nb)), kp * Q + pcol, col_groups[pcol])
                    for jl in range(ncb)}

        # ---------------- trtri ----------------
        for k in range(mt):
            kp, kq, il0, nrows, ncb, nrb, rootc, srcj = plans(k)
            tk, ilk, jlk = tdim(k), k // P, k // Q
            Akk = np.asfortranarray(diag_tile_everywhere_it_is_needed(k, kp, kq, True, True)[:tk, :tk])
            # trtri_ztrsmR: A(m,k) = -A(m,k) inv(A(k,k)), my rows m > k of tile column k
            if pcol == kq:
                for il in range(il0, il0 + nrows):
                    gm = il * P + prow
                    B = np.asfortranarray(tile(il, jlk)[:tdim(gm), :tk])
                    orc.trsm_tile(Right, Lower, NoTrans, NonUnit, -1.0, Akk, B)
                    tile(il, jlk)[:tdim(gm), :tk] = B
            # the new column slab along the grid rows (root kq), tile row k down the grid columns (root kp)
            rowp = {}
            if nrows > 0:
                slab = np.stack([np.array(tile(il, jlk)) for il in range(il0, il0 + nrows)]) if pcol == kq else np.zeros((nrows, nb, nb))
                slab = bc(slab, prow * Q + kq, row_groups[prow])
                rowp = {il0 + i: slab[i] for i in range(nrows)}
            colp = row_k_down_the_columns(k, kp, ncb)
            # trtri_zgemm: A(m,n) += A(m,k) A(k,n), m > k > n
            for jl in range(ncb):
                for il in range(il0, il0 + nrows):
                    gm = il * P + prow
                    Am = np.asfortranarray(rowp[il][:tdim(gm), :tk])
                    Bn = np.asfortranarray(colp[jl][:tk, :nb])
                    Cmn = np.asfortranarray(tile(il, jl)[:tdim(gm), :nb])
                    orc.gemm_tile(NoTrans, NoTrans, 1.0, Am, Bn, 1.0, Cmn)
                    tile(il, jl)[:tdim(gm), :nb] = Cmn
            if prow == kp:
                # trtri_ztrsmL: A(k,n) = inv(A(k,k)) A(k,n), n < k
                for jl in range(ncb):
                    B = np.asfortranarray(tile(ilk, jl)[:tk, :nb])
                    orc.trsm_tile(Left, Lower, NoTrans, NonUnit, 1.0, Akk, B)
                    tile(ilk, jl)[:tk, :nb] = B
                if pcol == kq:
                    T = np.asfortranarray(tile(ilk, jlk)[:tk, :tk])
                    assert orc.trtri_tile(Lower, NonUnit, T) == 0
                    tile(ilk, jlk)[:tk, :tk] = T
        # ---------------- lauum ----------------
        for k in range(mt):
            kp, kq, il0, nrows, ncb, nrb, rootc, srcj = plans(k)
            tk, ilk, jlk = tdim(k), k // P, k // Q
            colp = row_k_down_the_columns(k, kp, ncb)
            # A(k,m) for my tile rows m < k: along my grid row from the grid column that received it
            rowt = {il: bc(colp[srcj[il]] if pcol == rootc[il] else np.zeros((nb, nb)), prow * Q + rootc[il], row_groups[prow])
                    for il in range(nrb)}
            for jl in range(ncb):
                gn = jl * Q + pcol
                An = np.asfortranarray(colp[jl][:tk, :nb])
                for il in range(nrb):
                    gm = il * P + prow
                    if gm < gn:
                        continue
                    Cmn = np.asfortranarray(tile(il, jl))
                    if gm == gn:  # lauum_zherk
                        orc.herk_tile(Lower, ConjTrans, 1.0, An, 1.0, Cmn)
                    else:         # lauum_zgemm
                        orc.gemm_tile(ConjTrans, NoTrans, 1.0, np.asfortranarray(rowt[il][:tk, :nb]), An, 1.0, Cmn)
                    tile(il, jl)[:, :] = Cmn
            Akk = np.asfortranarray(diag_tile_everywhere_it_is_needed(k, kp, kq, False, True)[:tk, :tk])
            if prow == kp:
                for jl in range(ncb):  # lauum_ztrmm
                    B = np.asfortranarray(tile(ilk, jl)[:tk, :nb])
                    orc.trmm_tile(Left, Lower, ConjTrans, NonUnit, 1.0, Akk, B)
                    tile(ilk, jl)[:tk, :nb] = B
                if pcol == kq:         # lauum_zlauum
                    T = np.asfortranarray(tile(ilk, jlk)[:tk, :tk])
                    orc.lauum_tile(Lower, T)
                    tile(ilk, jlk)[:tk, :tk] = T
        want = F.copy(order="F")
        assert orc.potri(Lower, want, nb) == 0
        got = local_to_full(tiles.reshape(-1), N, N, nb, nb, P, Q, rank)
        mask = np.zeros((N, N), dtype=bool)
        for jl in range(lnt):
            for il in range(lmt):
                gm, gn = il * P + prow, jl * Q + pcol
                if gm >= gn:
                    mask[gm * nb:gm * nb + tdim(gm), gn * nb:gn * nb + tdim(gn)] = True
        low = mask & np.tril(np.ones((N, N), dtype=bool))
        up = mask & ~np.tril(np.ones((N, N), dtype=bool))          # strict upper parts of my diagonal tiles
        err = float(np.abs(got - want)[low].max() / np.abs(want).max()) if low.any() else 0.0
        untouched = bool(np.array_equal(got[up], F[up]))
        q.put((rank, err, untouched))
    finally:
        dist.destroy_process_group()


def gloo_gemm_syrk_worker(rank, world, port, P, Q, M, N, K, nb, q):
    """CPU: distributed C = alpha A B^T + beta C (dgemm NT) and C = alpha A A^T + beta C (dsyrk Lower NoTrans) on a
    P x Q grid with the ORACLE's tile kernels, driven by dplasma_b200_panel_plan with gmFirst = 0: at step k tile
    column k of A goes along the grid rows; tile column k of B (of A for syrk) goes along the grid rows too and
    then, from the rank on the diagonal of the grid pattern, down the grid column of its transposed position --
    the three broadcasts of the factorization (src/zpotrf_L.jdf:112,213-215) without the solves."""
    import torch
    import dplasma_b200 as dp
    from oracle.oracle import Oracle, Lower, NoTrans, ConjTrans
    from tests.util import full_to_local, local_to_full
    dist = _init(rank, world, port, "gloo")
    try:
        lib, orc = dp.load(), Oracle()
        prow, pcol = rank // Q, rank % Q
        row_groups = [dist.new_group([r * Q + c for c in range(Q)]) for r in range(P)]
        col_groups = [dist.new_group([r * Q + c for r in range(P)]) for c in range(Q)]
        alpha, beta = 0.75, -0.5

        def desc(m, n):
            s = dp.capi.BlockCyclicStruct()
            lib.parsec_matrix_block_cyclic_init(C.byref(s), dp.MATRIX_DOUBLE, dp.MATRIX_TILE, rank, nb, nb, m, n, 0, 0, m, n, P, Q, 1, 1, 0, 0)
            return s

        class Local:
            def __init__(self, X):
                self.m, self.n = X.shape
                self.s = desc(self.m, self.n)
                self.lmt, self.lnt, self.mt, self.nt = self.s.nb_elem_r, self.s.nb_elem_c, self.s.super.mt, self.s.super.nt
                flat = full_to_local(X, nb, nb, P, Q, rank)
                self.t = flat.reshape(self.lnt, self.lmt, nb, nb) if flat.size else np.zeros((self.lnt, self.lmt, nb, nb))

            def tile(self, il, jl):
                return self.t[jl, il].reshape((nb, nb)).T

            def rows(self, g):
                return self.m - g * nb if g == self.mt - 1 else nb

            def cols(self, g):
                return self.n - g * nb if g == self.nt - 1 else nb

            def full(self):
                return local_to_full(self.t.reshape(-1), self.m, self.n, nb, nb, P, Q, rank)

        def bc(t, src, group):
            x = torch.from_numpy(np.ascontiguousarray(t))
            dist.broadcast(x, src=src, group=group)
            return x.numpy()

        def row_panel(X, k):
            """tile column k of X, all rows: my grid row's tiles of it, from the rank in grid column k % Q"""
            out = (C.c_int * 6)()
            cr, cs = (C.c_int * max(X.lnt, 1))(), (C.c_int * max(X.lnt, 1))()
            lib.dplasma_b200_panel_plan(C.byref(X.s), k, 0, out, cr, cs)
            kq, il0, nrows = out[0], out[2], out[3]
            assert il0 == 0 and nrows == X.lmt
            if nrows == 0:
                return {}
            slab = np.stack([np.array(X.tile(il, k // Q)) for il in range(nrows)]) if pcol == kq else np.zeros((nrows, nb, nb))
            slab = bc(slab, prow * Q + kq, row_groups[prow])
            return {il: slab[il] for il in range(nrows)}

        def col_panel(X, k, rowp, ncols_c, owner_desc):
            """X(n,k) for my local tile columns n of C: down my grid column from grid row n % P, which holds it in rowp"""
            out = (C.c_int * 6)()
            cr, cs = (C.c_int * max(ncols_c, 1))(), (C.c_int * max(ncols_c, 1))()
            lib.dplasma_b200_panel_plan(C.byref(owner_desc), k, 0, out, cr, cs)
            assert out[4] == 0 and out[5] == ncols_c
            return {jl: bc(rowp[cs[jl]] if cr[jl] == prow else np.zeros((nb, nb)), cr[jl] * Q + pcol, col_groups[pcol])
                    for jl in range(ncols_c)}

        # ---------------- dgemm NT: C (M x N) = alpha A (M x K) B (N x K)^T + beta C ----------------
        A = Local(orc.plrnt(M, K, nb, nb, 11))
        B = Local(orc.plrnt(N, K, nb, nb, 12))
        Cfull = orc.plrnt(M, N, nb, nb, 13)
        Cm = Local(Cfull)
        # the plan of the transposed broadcast is the one of an N x N matrix with C's column distribution
        sq = desc(N, N)
        for k in range(A.nt):
            tk = A.cols(k)
            ap = row_panel(A, k)
            bp = col_panel(B, k, row_panel(B, k), Cm.lnt, sq)
            for jl in range(Cm.lnt):
                gn = jl * Q + pcol
                Bn = np.asfortranarray(bp[jl][:Cm.cols(gn), :tk])
                for il in range(Cm.lmt):
                    gm = il * P + prow
                    T = np.asfortranarray(Cm.tile(il, jl)[:Cm.rows(gm), :Cm.cols(gn)])
                    orc.gemm_tile(NoTrans, ConjTrans, alpha, np.asfortranarray(ap[il][:Cm.rows(gm), :tk]), Bn, beta if k == 0 else 1.0, T)
                    Cm.tile(il, jl)[:Cm.rows(gm), :Cm.cols(gn)] = T
        want = Cfull.copy(order="F")
        assert orc.gemm(NoTrans, ConjTrans, alpha, orc.plrnt(M, K, nb, nb, 11), orc.plrnt(N, K, nb, nb, 12), beta, want, nb, nb, nb) == 0
        mask = local_to_full(full_to_local(np.ones((M, N)), nb, nb, P, Q, rank), M, N, nb, nb, P, Q, rank) != 0
        err_gemm = float(np.abs(Cm.full() - want)[mask].max() / np.abs(want).max()) if mask.any() else 0.0
        # ---------------- dsyrk Lower NoTrans: S (M x M) = alpha A A^T + beta S ----------------
        Sfull = orc.plrnt(M, M, nb, nb, 22)
        S = Local(Sfull)
        for k in range(A.nt):
            tk = A.cols(k)
            ap = row_panel(A, k)
            cp = col_panel(A, k, ap, S.lnt, S.s)
            for jl in range(S.lnt):
                gn = jl * Q + pcol
                An = np.asfortranarray(cp[jl][:S.cols(gn), :tk])
                for il in range(S.lmt):
                    gm = il * P + prow
                    if gm < gn:
                        continue
                    T = np.asfortranarray(S.tile(il, jl)[:S.rows(gm), :S.cols(gn)])
                    zb = beta if k == 0 else 1.0
                    if gm == gn:
                        orc.herk_tile(Lower, NoTrans, alpha, An, zb, T)
                    else:
                        orc.gemm_tile(NoTrans, ConjTrans, alpha, np.asfortranarray(ap[il][:S.rows(gm), :tk]), An, zb, T)
                    S.tile(il, jl)[:S.rows(gm), :S.cols(gn)] = T
        wantS = Sfull.copy(order="F")
        assert orc.herk(Lower, NoTrans, alpha, orc.plrnt(M, K, nb, nb, 11), beta, wantS, nb, nb) == 0
        maskS = local_to_full(full_to_local(np.ones((M, M)), nb, nb, P, Q, rank), M, M, nb, nb, P, Q, rank) != 0
        d = np.abs(S.full() - wantS)   # the upper triangle must still be the input: wantS keeps it too
        err_syrk = float(d[maskS].max() / np.abs(wantS).max()) if maskS.any() else 0.0
        q.put((rank, err_gemm, err_syrk))
    finally:
        dist.destroy_process_group()


def nccl_blas3_worker(rank, world, port, P, Q, N, K, nb, q):
    """GPU: dplasma_dsyrk (Lower / NoTrans) and dplasma_dgemm (NoTrans, Trans) on a P x Q grid, device-resident
    operands, against numpy on the full matrices (the single-process forms are checked against the oracle and the
    reference's CORE kernels in tests/test_gpu_blas3.py)."""
    import torch
    import dplasma_b200 as dp
    from oracle.oracle import Oracle
    from tests.util import full_to_local
    torch.cuda.set_device(rank)
    dist = _init(rank, world, port, "nccl")
    try:
        lib = dp.load()
        ctx = dp.Context(lib)
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
        orc = Oracle()
        A = orc.plrnt(N, K, nb, nb, 101)
        B = orc.plrnt(N, K, nb, nb, 202)
        C0 = orc.plrnt(N, N, nb, nb, 303)

        def dev(Mx):
            d = dp.BlockCyclic(lib, Mx.shape[0], Mx.shape[1], nb, nb, P, Q, rank, where=1)
            t = torch.from_numpy(full_to_local(Mx, nb, nb, P, Q, rank).copy()).cuda() if d.elems else torch.zeros(1, dtype=torch.float64, device="cuda")
            d.s.mat = t.data_ptr()
            return d, t
        dA, tA = dev(A)
        dB, tB = dev(B)
        alpha, beta = -0.75, 0.5
        # dsyrk: only the lower triangle of C changes
        dC, tC = dev(C0)
        rc1 = lib.dplasma_dsyrk(ctx.ptr, dp.Lower, dp.NoTrans, alpha, dA.ref, beta, dC.ref)
        torch.cuda.synchronize()
        want = np.where(np.tril(np.ones((N, N), dtype=bool)), alpha * (A @ A.T) + beta * C0, C0)
        got = tC.cpu().numpy()[:dC.elems]
        e1 = float(np.abs(got - full_to_local(want, nb, nb, P, Q, rank)).max() / np.abs(want).max()) if dC.elems else 0.0
        # dgemm NT: the whole of C
        dC2, tC2 = dev(C0)
        rc2 = lib.dplasma_dgemm(ctx.ptr, dp.NoTrans, dp.Trans, alpha, dA.ref, dB.ref, beta, dC2.ref)
        torch.cuda.synchronize()
        want2 = alpha * (A @ B.T) + beta * C0
        got2 = tC2.cpu().numpy()[:dC2.elems]
        e2 = float(np.abs(got2 - full_to_local(want2, nb, nb, P, Q, rank)).max() / np.abs(want2).max()) if dC2.elems else 0.0
        q.put((rank, rc1, e1, rc2, e2))
        ctx.close()
    finally:
        dist.destroy_process_group()


def gloo_transpose_worker(rank, world, port, P, Q, N, nb, uplo_name, cap, q):
    """CPU: the distributed transposition behind the transposed grid solves (dist_transpose_tri), driven by the
    library's own round plan (dplasma_b200_transpose_plan): tiles packed by destination in plan order, one message per
    peer and direction and round over gloo, unpacked in plan order; result == the transpose's own block-cyclic tiles."""
    import ctypes as C
    import torch
    import dplasma_b200 as dp
    dist = _init(rank, world, port, "gloo")
    try:
        lib = dp.load()
        uplo = dp.Lower if uplo_name == "L" else dp.Upper
        mt = (N + nb - 1) // nb
        rng = np.random.default_rng(1234)
        F = rng.standard_normal((N, N))
        prow, pcol = rank // Q, rank % Q

        def tile(Mx, m, n):
            t = np.zeros((nb, nb))
            blk = Mx[m * nb:(m + 1) * nb, n * nb:(n + 1) * nb]
            t[:blk.shape[0], :blk.shape[1]] = blk
            return t
        in_tri = (lambda m, n: m >= n) if uplo == dp.Lower else (lambda m, n: m <= n)
        mine = {(m, n): tile(F, m, n) for m in range(mt) for n in range(mt)
                if in_tri(m, n) and m % P == prow and n % Q == pcol}
        if cap <= 0:
            cap = lib.dplasma_b200_transpose_cap(mt, nb * nb * 8)
        got = {}
        nA, rounds, max_out, max_in = 0, 0, 0, 0
        IntArr = C.c_int * (3 * mt * mt + 3)
        while nA < mt:
            nB, ns, nr = C.c_int(0), C.c_int(0), C.c_int(0)
            sl, rl = IntArr(), IntArr()
            rc = lib.dplasma_b200_transpose_plan(P, Q, mt, uplo, cap, rank, nA, C.byref(nB), sl, mt * mt + 1, C.byref(ns),
                                                 rl, mt * mt + 1, C.byref(nr))
            assert rc == 0 and nB.value > nA, (rc, nA, nB.value)
            sends = [(sl[3 * i], sl[3 * i + 1], sl[3 * i + 2]) for i in range(ns.value)]
            recvs = [(rl[3 * i], rl[3 * i + 1], rl[3 * i + 2]) for i in range(nr.value)]
            packed = {}
            for m, n, dst in sends:
                assert nA <= n < nB.value and (m, n) in mine, (m, n)
                if dst == rank:
                    got[(n, m)] = mine[(m, n)].T.copy()
                else:
                    packed.setdefault(dst, []).append(mine[(m, n)].T.copy())
            incoming = {}
            for m, n, src in recvs:
                assert src != rank
                incoming[src] = incoming.get(src, 0) + 1
            max_out = max(max_out, sum(len(v) for v in packed.values()))
            max_in = max(max_in, sum(incoming.values()))
            reqs, bufs = [], {}
            for r in range(world):
                if r == rank:
                    continue
                if r in packed:
                    reqs.append(dist.isend(torch.from_numpy(np.stack(packed[r])), dst=r))
                if r in incoming:
                    bufs[r] = torch.empty((incoming[r], nb, nb), dtype=torch.float64)
                    reqs.append(dist.irecv(bufs[r], src=r))
            for rq in reqs:
                rq.wait()
            idx = {}
            for m, n, src in recvs:
                k = idx.get(src, 0)
                idx[src] = k + 1
                got[(n, m)] = bufs[src][k].numpy().copy()
            nA = nB.value
            rounds += 1
        # what I must hold now: the tiles of F^T in the opposite triangle that the distribution gives me
        Ft = F.T
        want = {(a, b) for a in range(mt) for b in range(mt) if in_tri(b, a) and a % P == prow and b % Q == pcol}
        err = 0.0
        ok = set(got) == want
        for (a, b) in want & set(got):
            err = max(err, float(np.abs(got[(a, b)] - tile(Ft, a, b)).max()))
        q.put((rank, ok, err, rounds, max_out, max_in, int(cap)))
    finally:
        dist.destroy_process_group()
