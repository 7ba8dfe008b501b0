"""GPU, one process per GPU over NCCL: dplasma_dpotrf / check_dpotrf on a P x Q grid.
Skipped where fewer GPUs are visible than the grid needs."""
import socket

import pytest

pytestmark = pytest.mark.gpu


def free_port():
    """a rendezvous port OUTSIDE the kernel's ephemeral range (32768-60999): a port handed out by bind(0) can be
    taken by any outgoing connection between the moment it is released here and the moment the workers listen on it"""
    import os
    import random
    rng = random.Random(os.getpid() * 7919 + int.from_bytes(os.urandom(4), "little"))
    for _ in range(200):
        p = rng.randrange(20000, 32000)
        s = socket.socket()
        try:
            s.bind(("127.0.0.1", p))
        except OSError:
            continue
        finally:
            s.close()
        return p
    raise RuntimeError("no free rendezvous port")


def run_grid(P, Q, N, nb, check_values, cplx=False, uplo="L"):
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_potrf_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_potrf_worker, args=(r, world, port, P, Q, N, nb, q, check_values, cplx, uplo)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("N,nb", [(1000, 128), (4096, 256)])
def test_dpotrf_grid_matches_oracle(P, Q, N, nb):
    res = run_grid(P, Q, N, nb, True)
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)
        assert err < 1e-11, (rank, err)
        assert info_bad == res[0][6] and info_bad > 0     # MAX-reduced, identical on every rank


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("cplx", [False, True])
def test_potrf_grid_upper_matches_oracle(P, Q, cplx):
    """dplasmaUpper (what testing_zpotrf.c:21 runs by default) on a grid: each rank transposes its own tiles and the
    lower schedule runs on the Q x P grid with swapped coordinates; real and complex, ragged last tile"""
    res = run_grid(P, Q, 1000, 128, True, cplx=cplx, uplo="U")
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)
        assert err < 1e-11, (rank, err)
        assert info_bad == res[0][6] and info_bad > 0


@pytest.mark.parametrize("P,Q,N,nb", [(1, 2, 32768, 1024), (2, 2, 49152, 1024), (2, 4, 65536, 1024), (2, 4, 131072, 1024)])
def test_dpotrf_grid_large_residual(P, Q, N, nb):
    """the last case is BASELINE.json configs[2] (the north-star run) through the reference's residual check"""
    res = run_grid(P, Q, N, nb, False)
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)


def run_posv(P, Q, N, nb, K, nbr, uplo, check_values):
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_posv_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_posv_worker, args=(r, world, port, P, Q, N, nb, K, nbr, uplo, q, check_values)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=420) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    return res


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("uplo", ["L", "U"])
def test_dposv_grid_matches_oracle(P, Q, uplo):
    """dposv on a grid, both triangles; ragged sizes like the reference's mpi ctest (-N 457 -t 19 -K 367)"""
    res = run_posv(P, Q, 1000, 128, 367, 96, uplo, True)
    for rank, info, bad_f, res_f, bad_s, res_s, errF, errX in res:
        assert info == 0 and bad_f == 0 and bad_s == 0, (rank, info, bad_f, res_f, bad_s, res_s)
        assert 0 <= res_f < 60 and 0 <= res_s < 60
        assert errF < 1e-11 and errX < 1e-10, (rank, errF, errX)


@pytest.mark.parametrize("P,Q,N,nb,K,nbr", [(1, 2, 16384, 1024, 1024, 512), (2, 4, 65536, 1024, 1024, 256)])
def test_dposv_grid_large_residual(P, Q, N, nb, K, nbr):
    """BASELINE configs[3]: dposv N=65536 NRHS=1024 on 8 GPUs (P=2, Q=4) through the reference's two checks"""
    res = run_posv(P, Q, N, nb, K, nbr, "L", False)
    for rank, info, bad_f, res_f, bad_s, res_s, errF, errX in res:
        assert info == 0 and bad_f == 0 and bad_s == 0, (rank, info, bad_f, res_f, bad_s, res_s)
        assert 0 <= res_f < 60 and 0 <= res_s < 60


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 2), (2, 4)])
@pytest.mark.parametrize("uplo", ["L", "U"])
def test_pdpotrf_grid(P, Q, uplo):
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_pdpotrf_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_pdpotrf_worker, args=(r, world, port, P, Q, 1100, 128, uplo, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, info, err, untouched in res:
        assert info == 0 and err < 1e-11 and untouched, (rank, info, err, untouched)


@pytest.mark.parametrize("P,Q,rsrc,csrc,ia", [(1, 2, 0, 1, 1), (2, 1, 1, 0, 129), (1, 2, 0, 1, 257), (2, 2, 1, 1, 129), (2, 4, 1, 3, 385)])
def test_pdpotrf_grid_rsrc_csrc_and_submatrix(P, Q, rsrc, csrc, ia):
    """block (0,0) on a process other than (0,0) (DESCA RSRC / CSRC) and a block-aligned sub( A ); the grid comes from
    the process's BLACS symbols like in an application linked with ScaLAPACK"""
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_pdpotrf_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    for uplo in ("L", "U"):
        port = free_port()
        procs = [ctx.Process(target=nccl_pdpotrf_worker, args=(r, world, port, P, Q, 1100, 128, uplo, q, rsrc, csrc, ia, True))
                 for r in range(world)]
        for p in procs:
            p.start()
        res = [q.get(timeout=240) for _ in range(world)]
        for p in procs:
            p.join(timeout=120)
            assert p.exitcode == 0
        for rank, info, err, untouched in res:
            assert info == 0 and err < 1e-11 and untouched, (rank, uplo, info, err, untouched)


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 2)])
def test_zpotrf_grid_matches_oracle(P, Q):
    res = run_grid(P, Q, 1000, 128, True, cplx=True)
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)
        assert err < 1e-11, (rank, err)
        assert info_bad == res[0][6] and info_bad > 0


def test_zpotrf_baseline_config5_residual():
    """BASELINE configs[4]: zpotrf complex double N=49152 NB=512 on 4 x B200 (P=2, Q=2)"""
    res = run_grid(2, 2, 49152, 512, False, cplx=True)
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("kind,gbs", [("tile", None), ("sym", None), ("tile", 0.002), ("sym", 0.05), ("tile", 500.0)])
def test_dpotrf_grid_host_matrix(P, Q, kind, gbs):
    """the reference's own situation: the distributed matrix lives in host memory on every rank.  The tile columns
    travel to the GPUs while the factorization runs (the panels of the first steps are kept until every column has
    landed and been caught up); gbs forces what each rank believes about its copy rate"""
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_host_potrf_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_host_potrf_worker, args=(r, world, port, P, Q, 3000, 128, kind, q, gbs)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, info, err in res:
        assert info == 0 and err < 1e-11, (rank, info, err)


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("uplo", ["L", "U"])
def test_dpoinv_dpotri_dtrtri_grid_match_oracle(P, Q, uplo):
    """the inversion path on a process grid (ragged last tile): one-call dpoinv, dpotrf + dpotri, dpotrf + dtrtri"""
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_poinv_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_poinv_worker, args=(r, world, port, P, Q, 1000, 128, uplo, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, info1, e1, o1, info2, e2, o2, info3, e3, o3, info4 in res:
        assert info1 == 0 and info2 == 0 and info3 == 0, (rank, info1, info2, info3)
        assert e1 < 1e-11 and e2 < 1e-11 and e3 < 1e-11, (rank, e1, e2, e3)
        assert o1 and o2 and o3, (rank, o1, o2, o3)      # the other triangle is never touched
        assert info4 == res[0][10] and info4 > 0           # MAX-reduced, identical on every rank


@pytest.mark.parametrize("P,Q", [(1, 2), (2, 1), (2, 2), (2, 4)])
@pytest.mark.parametrize("N,K,nb", [(1000, 700, 128), (2048, 1024, 256)])
def test_dgemm_dsyrk_grid(P, Q, N, K, nb):
    """level-3 operations on a process grid (src/zgemm_NT.jdf, src/zsyrk_LN.jdf): the panels of the operands travel
    like the factorization's, one grouped update launch per tile column of the operand; ragged last tiles included"""
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_blas3_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_blas3_worker, args=(r, world, port, P, Q, N, K, nb, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    for rank, rc1, e1, rc2, e2 in res:
        assert rc1 == 0 and rc2 == 0, (rank, rc1, rc2)
        assert e1 < 1e-12 and e2 < 1e-12, (rank, e1, e2)
