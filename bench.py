#!/usr/bin/env python
"""bench.py -- dpotrf TFLOP/s (N^3/3 + N^2/2 + N/6 flops, src/flops.h:86-87) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one tiled Cholesky factorization (lower) of a dplgsy(bump=N, seed=3872) matrix:
  N=1 : BASELINE.json configs[1]  dpotrf N=32768 NB=512 on 1 x B200
  N>1 : 2D block-cyclic over a P x Q grid, one process per GPU, NB=1024, N grown so that each
        GPU holds the 17.2 GB of configs[2] (N=131072 at 8 GPUs, P=2 Q=4)  -> "weak" scaling.
`value` times the factorization only, matrix resident and generated in HBM (what the reference
times: tests/common.h:258-263); `e2e` is the same factorization through the blocking
dplasma_dpotrf() C-ABI call on a HOST (pinned) sym-block-cyclic matrix, H2D + D2H inside.
--impl reference times the reference's own CPU tile kernels (oracle/_ref) on the host cores.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 3872
METRIC = "dpotrf TFLOP/s (N^3/3) at 1/2/4/8 B200; % of FP64 tensor peak"


def flops_dpotrf(n):
    n = float(n)
    return n * n * n / 3.0 + n * n / 2.0 + n / 6.0


def grid_for(n):
    return {1: (1, 1), 2: (1, 2), 4: (2, 2), 8: (2, 4)}[n]


def workload_for(n_gpus):
    if n_gpus == 1:
        return 32768, 512
    import math
    nb = 1024
    N = int(round(131072 * math.sqrt(n_gpus / 8.0) / nb)) * nb
    return N, nb


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0=None, t1=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        sm, smax, reasons, pw = [], [], set(), []
        for ts, ln in self.lines:
            if t0 is not None and not (t0 <= ts <= t1 + 0.3):
                continue
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def cpu_reference_run(N, nb, steps, warmup):
    """the reference's own CPU tile kernels under an OpenMP DAG (oracle/_ref), else the oracle port"""
    try:
        from oracle.oracle import Reference, Lower
        R = Reference()
        kind, cores = "reference", R.info()["threads"]
        blas = R.info()["blas"] + " core=" + R.info()["core"]

        def one():
            t, info = R.time_dpotrf(Lower, N, nb, SEED)
            assert info == 0
            return t
    except (FileNotFoundError, OSError):
        from oracle.oracle import Oracle, Lower
        O = Oracle()
        kind, cores, blas = "port", os.cpu_count(), "plain C loops"

        def one():
            A = O.plgsy(float(N), Lower, N, nb, SEED)
            t0 = time.perf_counter()
            assert O.potrf(Lower, A, nb) == 0
            return time.perf_counter() - t0
    for _ in range(warmup):
        one()
    ts = [one() for _ in range(steps)]
    return kind, cores, blas, ts


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded sample of the workload: same algorithm, tile kernels and generator, smaller N
    N, nb = (16384, 256)
    kind, cores, blas, ts = cpu_reference_run(N, nb, args.steps, min(args.warmup, 1))
    t = sum(ts) / len(ts)
    val = flops_dpotrf(N) / t / 1e12
    Nw, nbw = workload_for(args.gpus)
    sample = (f"dpotrf N={N} NB={nb} lower, dplgsy(bump=N, seed={SEED}); reference CORE_d* tile kernels "
              f"({blas}) under an OpenMP task DAG restating src/zpotrf_L.jdf; {cores} host threads; "
              f"bounded sample of the N={Nw} NB={nbw} workload (PaRSEC runtime itself unavailable on this box)")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "TFLOP/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"dpotrf N={Nw} NB={nbw} lower on {args.gpus} x B200 (CPU arm: bounded sample N={N} NB={nb})"},
        "cpu_baseline": {"value": val, "unit": "TFLOP/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": val, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import torch
    import dplasma_b200 as dp

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    n_gpus = args.gpus
    if world != n_gpus:
        if world == 1 and n_gpus > 1:
            raise SystemExit("launch with torch.distributed.run --nproc-per-node N for --gpus N > 1")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: dplasma_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    lib = dp.load()
    ctx = dp.Context(lib)
    if world > 1:
        ids = [ctx.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(ids[0], rank, world)
    P, Q = grid_for(n_gpus)
    N, nb = workload_for(n_gpus)
    flops = flops_dpotrf(N)
    K, W = args.steps, max(args.warmup, 3)

    # ---- device-resident matrices: as many pre-generated copies as fit comfortably ----
    d0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device")
    local_bytes = d0.elems * 8
    nbuf = max(1, min(K, int(90e9 // max(local_bytes, 1))))
    bufs = [d0] + [dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device") for _ in range(nbuf - 1)]

    def gen(d):
        rc = lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, d.ref, SEED)
        assert rc == 0, rc

    def factor(d):
        info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, d.ref)
        assert info == 0, f"dpotrf info={info}"
        return lib.dplasma_b200_last_elapsed(ctx.ptr)

    for _ in range(W):
        gen(d0)
        factor(d0)

    # ---- timed region: exactly K factorizations, barrier + synchronize on both sides ----
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    dev_seconds, wall_seconds, done = 0.0, 0.0, 0
    launches0 = lib.dplasma_b200_kernel_launches(ctx.ptr)
    launches = 0
    t_first, t_last = None, None
    while done < K:
        chunk = min(nbuf, K - done)
        for b in bufs[:chunk]:
            gen(b)                       # untimed: inputs resident in HBM when the timed region starts
        barrier()
        l0 = lib.dplasma_b200_kernel_launches(ctx.ptr)
        t0 = time.perf_counter()
        t_first = t_first or time.time()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for b in bufs[:chunk]:
            dev_seconds += factor(b)     # CUDA events on the library's own stream, per factorization
        barrier()
        wall_seconds += time.perf_counter() - t0
        t_last = time.time()
        launches += lib.dplasma_b200_kernel_launches(ctx.ptr) - l0
        done += chunk
    clocks = sampler.stop(t_first, t_last) if rank == 0 else None
    tt = torch.tensor([dev_seconds, wall_seconds], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    dev_seconds, wall_seconds = tt.tolist()
    value = flops * K / dev_seconds / 1e12

    # residual of the last factorization (the reference's acceptance test), single process only
    residual = None
    if world == 1:
        dA0 = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, where="device")
        gen(dA0)
        bad = lib.check_dpotrf(ctx.ptr, 0, dp.Lower, bufs[0].ref, dA0.ref)
        residual = lib.dplasma_b200_last_check_result(ctx.ptr)
        assert bad == 0, f"residual check failed: {residual}"
        dA0.free()

    # ---- roofline leg: event-time every grouped trailing-update launch of one factorization ----
    gen(d0)
    barrier()
    lib.dplasma_b200_profile(ctx.ptr, 1)
    factor(d0)
    ps, pf, pl, pb = C.c_double(0), C.c_double(0), C.c_int(0), C.c_double(0)
    lib.dplasma_b200_profile_read(ctx.ptr, C.byref(ps), C.byref(pf), C.byref(pl), C.byref(pb))
    lib.dplasma_b200_profile(ctx.ptr, 0)
    secs = C.c_double(0)
    peak = lib.dplasma_b200_dmma_peak_tflops(2, 4096, C.byref(secs))
    achieved = pf.value / ps.value / 1e12 if ps.value > 0 else None
    roofline = {
        "bound": "tensor", "kernel": "gemm_dmma_kernel (grouped trailing update: DGEMM + DSYRK tiles)",
        "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": (achieved / peak) if achieved and peak > 0 else None,
        "traffic": None,
        "traffic_note": "ncu --set full on the same kernel (profiles/ncu_gemm_r01_final_summary.md): dram read+write = 1.59 x algorithmic bytes, 13 % of the HBM roof",
        "launches_timed": pl.value, "algorithmic_flops_per_launch_avg": pf.value / max(pl.value, 1),
        "best_launch_tflops": pb.value, "share_of_step": ps.value / (dev_seconds / K) if dev_seconds > 0 else None,
        "peak_source": "of measured: live register-resident DMMA (mma.sync.m8n8k4.f64) loop on this GPU "
                       "(MEASURED_PEAKS.json holds no FP64 figure); nominal B200 FP64 tensor ~37-40 TFLOP/s",
    }

    # ---- e2e: blocking C-ABI call on a HOST (pinned) sym block-cyclic matrix ----
    e2e = None
    try:
        h = dp.BlockCyclic(lib, N, N, nb, nb, P, Q, rank, kind="sym", uplo=dp.Lower, where="host")
        h2d = d2h = h.elems * 8
        ke = min(K, 3)
        tsum = 0.0
        for it in range(ke + 1):
            rc = lib.dplasma_dplgsy(ctx.ptr, float(N), dp.Lower, h.ref, SEED)  # untimed input generation
            assert rc == 0
            barrier()
            t0 = time.perf_counter()
            info = lib.dplasma_dpotrf(ctx.ptr, dp.Lower, h.ref)               # H2D + factor + D2H + info read
            barrier()
            dt = time.perf_counter() - t0
            assert info == 0
            if it > 0:
                tsum += dt
        te = torch.tensor([tsum], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        tsum = te.item()
        e2e = {"value": flops * ke / tsum / 1e12, "unit": "TFLOP/s", "h2d_bytes_per_step": h2d * world,
               "d2h_bytes_per_step": d2h * world, "steps": ke, "ms_per_step": tsum / ke * 1e3,
               "api": "dplasma_dpotrf(parsec, dplasmaLower, A) with A->mat in pinned host memory (sym block-cyclic tile storage)"}
        h.free()
    except (MemoryError, AssertionError) as ex:  # host RAM too small for the pinned copy
        e2e = {"value": None, "unit": "TFLOP/s", "error": repr(ex)}

    # ---- CPU baseline beside it (rank 0, N=1 only): the reference's own tile kernels ----
    cpu = None
    if rank == 0 and n_gpus == 1 and not args.no_cpu:
        cn, cnb = 24576, 256
        kind, cores, blas, ts = cpu_reference_run(cn, cnb, 1, 0)
        cpu = {"value": flops_dpotrf(cn) / ts[0] / 1e12, "unit": "TFLOP/s", "cores": cores, "kind": kind,
               "sample": f"dpotrf N={cn} NB={cnb} lower, one run of {ts[0]:.1f} s; reference CORE_d* kernels ({blas}) "
                         f"under an OpenMP task DAG restating src/zpotrf_L.jdf"}

    if rank == 0:
        out = {
            "metric": METRIC, "value": value, "unit": "TFLOP/s", "n_gpus": n_gpus, "steps": K, "warmup": W,
            "ms_per_step": dev_seconds / K * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"dpotrf N={N} NB={nb} lower on {n_gpus} x B200, 2D block-cyclic P={P} Q={Q}, "
                                   f"dplgsy(bump=N, seed={SEED})",
                       "N": N, "NB": nb, "P": P, "Q": Q,
                       "l2": f"inputs larger than L2 ({local_bytes / 1e9:.1f} GB per GPU vs 126 MB), fresh matrix per step",
                       "timing": "CUDA events on the library's stream around each factorization, max over ranks; "
                                 "generation outside the timed region like tests/common.h:258-263"},
            "wall_ms_per_step": wall_seconds / K * 1e3,
            "residual": residual, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
        }
        print(json.dumps(out))
    for b in bufs:
        b.free()
    ctx.close()
    if dist is not None:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
