"""BASELINE configs[3]-shaped dposv on a process grid (bench.bench_posv), alone: times of dpotrf and dpotrs and the
two residuals.  python -m torch.distributed.run --nproc-per-node N ... tools/posv_probe.py [N nb nrhs nbr]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    import dplasma_b200 as dp
    import bench
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    if world > 1:
        dist.init_process_group("nccl")
    lib = dp.load()
    ctx = dp.Context(lib)
    if world > 1:
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
    a = [int(x) for x in sys.argv[1:5]]
    N, nb, nrhs, nbr = a if len(a) == 4 else (32768 if world < 8 else 65536, 1024, 1024, 256)
    P, Q = bench.grid_for(world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    out = bench.bench_posv(lib, dp, ctx, torch, dist if world > 1 else None, barrier, rank, world, P, Q, N, nb, nrhs, nbr)
    if rank == 0:
        print(json.dumps(out))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
