"""GPU, one process per GPU over NCCL: dplasma_dpotrf / check_dpotrf on a P x Q grid.
Skipped where fewer GPUs are visible than the grid needs."""
import socket

import pytest

pytestmark = pytest.mark.gpu


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def run_grid(P, Q, N, nb, check_values):
    import torch
    import torch.multiprocessing as mp
    from tests.dist_workers import nccl_potrf_worker
    world = P * Q
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = free_port()
    procs = [ctx.Process(target=nccl_potrf_worker, args=(r, world, port, P, Q, N, nb, q, check_values)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=600) for _ in range(world)]
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


@pytest.mark.parametrize("P,Q,N,nb", [(1, 2, 32768, 1024), (2, 2, 49152, 1024), (2, 4, 65536, 1024)])
def test_dpotrf_grid_large_residual(P, Q, N, nb):
    res = run_grid(P, Q, N, nb, False)
    for rank, info, bad, resid, err, t, info_bad in res:
        assert info == 0 and bad == 0 and 0.0 <= resid < 60.0, (rank, info, bad, resid)
