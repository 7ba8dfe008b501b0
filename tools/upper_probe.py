"""dplasmaUpper beside dplasmaLower on a process grid, at bench.py's workload for the same number of GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        tools/upper_probe.py [--cplx] > gpurun_out/upper.json

Prints one JSON line: seconds (max over ranks, the library's own CUDA events around each factorization), TFLOP/s and
the residual of src/dplasma_zcheck.c:89-119 for both triangles.  Not part of the product: a measurement script.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cplx", action="store_true")
    ap.add_argument("--N", type=int, default=0)
    ap.add_argument("--nb", type=int, default=0)
    args = ap.parse_args()
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
    P, Q = bench.grid_for(world)
    N, nb = bench.workload_for(world)
    if args.cplx:
        N, nb = (N * 3) // 4, 512
    N, nb = args.N or N, args.nb or nb
    gen = lib.dplasma_zplghe if args.cplx else lib.dplasma_dplgsy
    fac = lib.dplasma_zpotrf if args.cplx else lib.dplasma_dpotrf
    chk = lib.check_zpotrf if args.cplx else lib.check_dpotrf
    flops = bench.flops_zpotrf(N) if args.cplx else bench.flops_dpotrf(N)
    dA = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device", cplx=args.cplx)
    dA0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device", cplx=args.cplx)

    def allmax(*v):
        t = torch.tensor(v, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    out = {"n_gpus": world, "N": N, "NB": nb, "P": P, "Q": Q, "dtype": "z" if args.cplx else "d"}
    for name, uplo in (("lower", dp.Lower), ("upper", dp.Upper)):
        ts = []
        for it in range(3):
            assert gen(ctx.ptr, float(N), uplo, dA.ref, 3872) == 0
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            info = fac(ctx.ptr, uplo, dA.ref)
            assert info == 0, info
            ts.append(allmax(lib.dplasma_b200_last_elapsed(ctx.ptr))[0])
        assert gen(ctx.ptr, float(N), uplo, dA0.ref, 3872) == 0
        bad = chk(ctx.ptr, 0, uplo, dA.ref, dA0.ref)
        bad, res = allmax(float(bad), lib.dplasma_b200_last_check_result(ctx.ptr))
        out[name] = {"seconds": min(ts[1:]), "tflops": flops / min(ts[1:]) / 1e12, "all_seconds": ts, "residual": res,
                     "passed": bool(bad == 0 and 0 <= res < 60)}
    if rank == 0:
        print(json.dumps(out))
    dA.free()
    dA0.free()
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
