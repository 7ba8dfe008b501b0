"""CPU, world_size > 1 (gloo): the multi-rank logic of the tiled Cholesky -- 2D block-cyclic
ownership and the per-step broadcast plan the library computes (dplasma_b200_panel_plan) --
exercised end to end with the oracle's tile kernels standing in for the GPU kernels."""
import socket

import pytest


def free_port():
    """outside the kernel's ephemeral range, see tests/test_gpu_multi.py"""
    import os
    import random
    rng = random.Random(os.getpid() * 104729 + int.from_bytes(os.urandom(4), "little"))
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


@pytest.mark.parametrize("P,Q,N,nb", [(1, 2, 200, 32), (2, 1, 200, 32), (2, 2, 232, 24)])
def test_distributed_cholesky_plan_over_gloo(P, Q, N, nb):
    import torch.multiprocessing as mp
    from tests.dist_workers import gloo_cholesky_worker
    world = P * Q
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=gloo_cholesky_worker, args=(r, world, port, P, Q, N, nb, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, info, winfo in res:
        assert err < 1e-12, (rank, err)
        assert info == winfo == 0


@pytest.mark.parametrize("P,Q,N,nb", [(1, 2, 200, 32), (2, 1, 200, 32), (2, 2, 232, 24), (2, 3, 250, 24)])
def test_distributed_trtri_lauum_plan_over_gloo(P, Q, N, nb):
    """the communication pattern of the inversion DAGs on a grid (tile column below the diagonal along the grid
    rows, tile row left of it down the grid columns and, for lauum, on along the grid row of its transposed
    position) with the library's plans and the oracle's tile kernels; result == single-process oracle potri"""
    import torch.multiprocessing as mp
    from tests.dist_workers import gloo_potri_worker
    world = P * Q
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=gloo_potri_worker, args=(r, world, port, P, Q, N, nb, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err, untouched in res:
        assert err < 1e-12, (rank, err)
        assert untouched, rank


@pytest.mark.parametrize("P,Q,M,N,K,nb", [(1, 2, 150, 130, 170, 32), (2, 1, 150, 130, 170, 32), (2, 2, 140, 100, 90, 24), (2, 3, 160, 150, 100, 24)])
def test_distributed_gemm_syrk_plan_over_gloo(P, Q, M, N, K, nb):
    """dgemm (NT) and dsyrk (Lower, NoTrans) on a grid are the factorization's panel exchange with gmFirst = 0:
    the library's plan + the oracle's tile kernels over gloo reproduce the single-process oracle"""
    import torch.multiprocessing as mp
    from tests.dist_workers import gloo_gemm_syrk_worker
    world = P * Q
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=gloo_gemm_syrk_worker, args=(r, world, port, P, Q, M, N, K, nb, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err_gemm, err_syrk in res:
        assert err_gemm < 1e-12 and err_syrk < 1e-12, (rank, err_gemm, err_syrk)


@pytest.mark.parametrize("P,Q,N,nb,uplo,cap", [(1, 2, 200, 32, "L", 0), (2, 1, 200, 32, "U", 0), (2, 2, 232, 24, "L", 11),
                                               (2, 3, 250, 24, "U", 12), (2, 2, 232, 24, "U", 0)])
def test_distributed_transposition_plan_over_gloo(P, Q, N, nb, uplo, cap):
    """the all-to-all in rounds behind the transposed solves on a grid (csrc/potrf_dist.cu: dist_transpose_tri) with the
    library's own round plan; cap > 0 forces several rounds"""
    import torch.multiprocessing as mp
    from tests.dist_workers import gloo_transpose_worker
    world = P * Q
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=gloo_transpose_worker, args=(r, world, port, P, Q, N, nb, uplo, cap, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, err, rounds, max_out, max_in, used_cap in res:
        assert ok and err == 0.0, (rank, ok, err)
        assert max_out <= used_cap and max_in <= used_cap, (rank, max_out, max_in, used_cap)
        if cap:
            assert rounds > 1
